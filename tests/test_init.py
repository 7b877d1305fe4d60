"""Initialisation: the generators Spark's init draws from, the driver-side local k-means++,
and -- on the GPU -- the k-means|| row passes against the CPU restatement (oracle/spark_init.py).  java.util.Random is
pinned by its well-known first values; XORShiftRandom is restated from Spark's published source and unpinned (neither the
file nor vectors for it are in the reference tree): what is checked for it is that the three implementations here (C++
in the library, Python in the oracle, Python in the host mirror) are the same generator."""
import numpy as np
import pytest

from oracle import kmeans_oracle as ko
from oracle import spark_init as si


def test_java_util_random_known_values():
    assert si.JavaRandom(0).next_double() == 0.730967787376657
    assert si.JavaRandom(0).next_int() == -1155484576
    assert si.JavaRandom(42).next_int() == -1170105035
    assert si.JavaRandom(42).next_int(10) == 0
    from oap_mllib_b200.init import JavaRandom
    assert JavaRandom(0).nextDouble() == 0.730967787376657 and JavaRandom(42).nextInt(10) == 0


def test_xorshift_library_and_oracle_are_the_same_generator(built_lib):
    import oap_mllib_b200 as m
    lib = m.lib()
    for seed in (0, 1, -7, 42, 2 ** 40 + 3, -(2 ** 62)):
        r = si.XORShiftRandom(seed)
        assert lib.kmb200_xorshift_next_int(seed) == r.next_int()
        r = si.XORShiftRandom(seed)
        want = [r.next_double() for _ in range(5)]
        assert [lib.kmb200_xorshift_next_double(seed, i) for i in range(5)] == want
        assert all(0.0 <= u < 1.0 for u in want)
    assert si.XORShiftRandom(0).next_int() != 0          # Spark's XORShiftRandomSuite: "XORShift with zero seed"


def test_local_kmeans_pp_host_mirror_matches_oracle():
    from oap_mllib_b200.init import local_kmeans_pp
    rng = np.random.default_rng(3)
    pts = rng.normal(size=(400, 7)) + 6.0 * rng.integers(0, 5, size=(400, 1))
    w = rng.integers(0, 50, 400).astype(np.float64)
    a = local_kmeans_pp(0, pts, w, 9, 30)
    b = si.local_kmeans_pp(0, pts, w, 9, 30)
    np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("n,d,k,steps,parts", [(20000, 16, 10, 2, None), (60000, 64, 40, 2, [0, 25000, 25001, 60000]),
                                                (300_000, 32, 16, 3, [0, 100_000, 300_000])])
def test_kmeans_parallel_row_passes_match_the_oracle(built_lib, n, d, k, steps, parts):
    import oap_mllib_b200 as m
    from oap_mllib_b200 import init as gi
    assert m.lib().kmb200_device_ok() == 1
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + d)
    want_c, want_rows = si.init_kmeans_parallel(X.astype(np.float64), k, steps, seed=5, part_rows=parts, part_index0=3)
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        got_c, got_rows = gi.init_kmeans_parallel(t, k, steps, seed=5, part_rows=parts, part_index0=3)
    finally:
        t.release()
    assert got_rows == want_rows                      # the same rows kept in every step, in the same order
    assert got_c.shape == want_c.shape
    np.testing.assert_allclose(got_c, want_c, rtol=1e-9, atol=1e-9)


@pytest.mark.gpu
def test_random_init_and_fit_with_gpu_init(built_lib):
    import oap_mllib_b200 as m
    from oap_mllib_b200 import init as gi
    X, _, _ = ko.gaussian_blobs(30000, 24, 12, seed=9)
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        got = gi.init_random(t, 12, seed=11)
    finally:
        t.release()
    np.testing.assert_array_equal(got, si.init_random(X.astype(np.float64), 12, seed=11))
    for mode in ("k-means||", "random"):
        model = m.KMeans(k=12, initMode=mode, maxIter=30, seed=3).fit(X)
        init = si.init_kmeans_parallel(X.astype(np.float64), 12, 2, 3)[0] if mode == "k-means||" else si.init_random(X.astype(np.float64), 12, 3)
        want_c, want_cost, want_it = ko.lloyd(X.astype(np.float64), init, 1e-4, 30)
        assert model.numIter == want_it
        assert abs(model.trainingCost - want_cost) <= 1e-5 * want_cost


@pytest.mark.gpu
def test_init_centers_are_rows_and_deduplicated(built_lib):
    """The reference's 50 x 3 fixture (row i = [i % 5] * 3): only 5 distinct rows, so both modes return at most 5 centres,
    all of them rows of the data, none twice (KMeans.scala:357-358, :414-417)."""
    import oap_mllib_b200 as m
    X = np.array([[float(i % 5)] * 3 for i in range(1, 51)])
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        for mode in ("random", "k-means||"):
            c = m.KMeans(k=5, initMode=mode, seed=1)._init_centers(t)
            assert 1 <= c.shape[0] <= 5 and c.shape[1] == 3
            assert len(np.unique(c, axis=0)) == c.shape[0]
            assert all(any((r == x).all() for x in X) for r in c)
    finally:
        t.release()
