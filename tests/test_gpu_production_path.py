"""GPU parity on the path the benchmark times: tables of >= 2^18 rows, where the FP16-split kernel runs as a
cascade (coarse one-product pass over all rows -> three-product pass over the undecided rows -> full bookkeeping
over the near-tie rows -> FP64 re-check) with its DEFAULT options.  Every case is compared with the FP64 oracle:
centres, cost, iteration count (check_fit_parity) and the label of every row (kmb200_assign + assert_labels_match).

Shapes: the headline geometry (d = 256, k = 1000: BN = 128, 8 centroid tiles, serpentine walk with resident B
groups), C2's (d = 64, k = 100), and data on which the cascade's hand-offs happen -- heavily overlapping blobs,
uniform rows, small-integer rows with masses of exactly tied distances -- so that "coarse pass dropped" and
"near-tie copy overflowed" are exercised at scale and, with a tightened coarse-copy size, in the middle of a fit.
Tolerances are BASELINE.json's (helpers.py)."""
import ctypes as C

import numpy as np
import pytest

from helpers import assert_labels_match
from oracle import kmeans_oracle as ko
from test_gpu_parity import check_fit_parity, gpu_assign

pytestmark = pytest.mark.gpu
MIN_ROWS = 1 << 18      # kTwoLevelMinRows in csrc/kmb200.cu: the two-level scheme and the cascade start here


@pytest.fixture(scope="module")
def m(built_lib):
    import oap_mllib_b200 as mod
    assert mod.lib().kmb200_device_ok() == 1, "no sm_100 GPU visible: the CUDA path cannot run"
    return mod


def unique_row_init(X, k):
    """Rows floor(j*n/k) like row_init_centers, moved forward past rows already taken (integer data repeats rows;
    duplicated initial centres would leave clusters empty at once, which is a different test)."""
    n = X.shape[0]
    seen, out = set(), []
    for j in range(k):
        i = (j * n) // k
        while X[i].tobytes() in seen:
            i = (i + 1) % n
        seen.add(X[i].tobytes())
        out.append(X[i])
    return np.asarray(out, np.float64)


def fit_labels_profile(m, X, init, max_iter, tol=1e-4):
    """Fit parity, then the label of every row at the oracle's final centres; returns the fit's profile."""
    assert X.shape[0] >= MIN_ROWS
    got_c, _, _ = check_fit_parity(m, X, init, tol=tol, max_iter=max_iter)
    prof = m.kmeans.profile()
    X32 = np.asarray(X, np.float32)
    t = m.OneDAL.vectorsToMergedNumericTables(X32)
    try:
        labels, sizes, _ = gpu_assign(m, t, got_c)
    finally:
        t.release()
    flipped = assert_labels_match(labels, X32.astype(np.float64), got_c)
    assert sizes.sum() == X.shape[0] and np.array_equal(sizes, np.bincount(labels, minlength=init.shape[0]))
    print(f"  {X.shape[0]} x {X.shape[1]}, k = {init.shape[0]}: coarse-undecided rows / iteration "
          f"{prof['coarse_undecided_rows'] / max(prof['iterations'], 1):.0f}, re-checked {prof['rechecked_rows']}, "
          f"cascade dropped at {prof['cascade_dropped_at']}, two-level dropped at {prof['two_level_dropped_at']}, "
          f"near-tie labels that differ {flipped}")
    return prof


def test_headline_geometry_default_cascade(m):
    """d = 256, k = 1000 (C3's shape), 400 000 rows, row init: the cascade must have run in every iteration."""
    X, _, _ = ko.gaussian_blobs(400_000, 256, 1000, seed=31)
    prof = fit_labels_profile(m, X, ko.row_init_centers(X, 1000), max_iter=6)
    assert prof["coarse_undecided_rows"] > 0 and prof["cascade_dropped_at"] == -1 and prof["two_level_dropped_at"] == -1


def test_headline_geometry_near_duplicate_centres(m):
    """Same geometry, but 40 of the initial centres are copies of others displaced by ~1e-6: whole clusters of rows
    sit inside the error bound of both schemes and must be decided in FP64 exactly as the oracle decides them."""
    X, _, _ = ko.gaussian_blobs(300_000, 256, 1000, seed=32)
    init = ko.row_init_centers(X, 1000)
    rng = np.random.default_rng(5)
    init[25::25] = init[24::25][:len(init[25::25])] + 1e-6 * rng.standard_normal((len(init[25::25]), 256))
    fit_labels_profile(m, X, init, max_iter=5)


@pytest.mark.parametrize("force_cascade", [False, True])
def test_c2_geometry(m, force_cascade):
    """d = 64, k = 100 (C2's shape): one centroid tile and two chunks per row tile -- by default the fit skips the coarse
    pass there (three-product fast pass + near-tie pass); with "cascade" = 2 the coarse pass runs all the same."""
    X, _, _ = ko.gaussian_blobs(300_000, 64, 100, seed=33)
    try:
        if force_cascade:
            m._lib.check(m.lib().kmb200_set_option(b"cascade", 2.0))
        prof = fit_labels_profile(m, X, ko.row_init_centers(X, 100), max_iter=8)
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"cascade", 1.0))
    assert prof["assign_variant"] == (1 if force_cascade else 2)
    assert (prof["coarse_undecided_rows"] > 0) == force_cascade


@pytest.mark.parametrize("kind", ["overlapping", "uniform", "integers"])
def test_hard_data_hand_offs(m, kind):
    """Data the cascade is bad at: the coarse pass (and, on the integer rows, the near-tie copy too) gives up and the
    fit continues on the next scheme down; the result is the oracle's all the same."""
    rng = np.random.default_rng(41)
    if kind == "overlapping":      # blobs with sigma = 6 around centres ~ U(-10, 10): clusters overlap heavily
        n, d, k = 300_000, 64, 100
        cen = rng.uniform(-10, 10, (k, d))
        X = (cen[rng.integers(0, k, n)] + 6.0 * rng.standard_normal((n, d))).astype(np.float32)
        init = ko.row_init_centers(X, k)
    elif kind == "uniform":        # no structure at all: every row is about as far from every centre
        n, d, k = 300_000, 32, 64
        X = rng.uniform(0, 1, (n, d)).astype(np.float32)
        init = ko.row_init_centers(X, k)
    else:                          # small integers: distances to the initial centres are integers, exact ties abound
        n, d, k = 270_000, 16, 48
        X = rng.integers(0, 4, (n, d)).astype(np.float32)
        init = unique_row_init(X, k)
    try:
        m._lib.check(m.lib().kmb200_set_option(b"cascade", 2.0))   # (these shapes would skip the coarse pass by default)
        prof = fit_labels_profile(m, X, init, max_iter=6)
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"cascade", 1.0))
    if kind == "integers":
        assert prof["rechecked_rows"] > n // 16     # far more near-ties than the near-tie copy holds


@pytest.mark.parametrize("cand1_frac", [0.08, 0.16])
def test_coarse_pass_dropped_in_mid_fit(m, cand1_frac):
    """With the coarse pass's copy sized between the undecided fractions of the first and of later iterations, the
    hand-off to the two-level scheme happens inside the fit (the profile says in which iteration)."""
    X, _, _ = ko.gaussian_blobs(300_000, 128, 500, seed=35)
    try:
        m._lib.check(m.lib().kmb200_set_option(b"cand1_frac", cand1_frac))
        prof = fit_labels_profile(m, X, ko.row_init_centers(X, 500), max_iter=6)
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"cand1_frac", 0.3))
    assert prof["iterations"] >= 2


@pytest.mark.parametrize("d", [1030, 1536])
def test_wide_rows_beyond_one_accumulate_pass(m, d):
    """d > 1024: the accumulate kernel walks a row in several passes of 256 float4 lanes, the last one partial."""
    X, _, _ = ko.gaussian_blobs(3000, d, 12, seed=d)
    check_fit_parity(m, X, ko.row_init_centers(X, 12), tol=1e-4, max_iter=4)


def test_c4_geometry_wide_cascade(m):
    """d = 512 (C4's shape): the one-product pass of the FP16-split kernel over all rows, the streaming 3xTF32 kernel over
    the rows it leaves undecided."""
    X, _, _ = ko.gaussian_blobs(270_000, 512, 2000, seed=36)
    # (centres after one oracle step: from raw rows the first pass leaves more than cand1_frac of the rows undecided
    # and the fit drops the cascade at once -- that hand-off is test_hard_data_hand_offs' business)
    init, _, _, _ = ko.lloyd_step([X.astype(np.float64)], ko.row_init_centers(X, 2000))
    prof = fit_labels_profile(m, X, init, max_iter=3)
    assert prof["assign_variant"] == 8 and prof["coarse_undecided_rows"] > 0 and prof["cascade_dropped_at"] == -1


@pytest.mark.parametrize("n,d,k", [(9000, 384, 200), (6000, 512, 700), (5000, 288, 40), (4000, 480, 1100)])
def test_wide_cascade_forced_on_small_tables(m, n, d, k):
    """The same scheme forced onto small tables (10 to 16 chunks of 32 columns, every BN, ragged row tiles)."""
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + d)
    try:
        m._lib.check(m.lib().kmb200_set_option(b"two_level", 2.0))
        check_fit_parity(m, X, ko.row_init_centers(X, k), tol=1e-4, max_iter=5)
        assert m.kmeans.profile()["assign_variant"] == 8
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"two_level", 1.0))
