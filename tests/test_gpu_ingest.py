"""Input layouts either side of the hot path: columnar, sparse (CSR), tables created
with an upper bound, pinned staging.  The checker is the oracle's restatement of the reference's conversion
(every row densified to double[] and stored row-major)."""
import ctypes as C

import numpy as np
import pytest

from oracle import kmeans_oracle as ko

pytestmark = pytest.mark.gpu


@pytest.fixture()
def m(built_lib):
    import oap_mllib_b200 as mod
    assert mod.lib().kmb200_device_ok() == 1, "no sm_100 GPU visible: the CUDA path cannot run"
    yield mod
    mod._lib.check(mod.lib().kmb200_set_option(b"stage_mb", 256.0))


def gpu_fit(m, table, init, max_iter=5):
    impl = m.KMeansDALImpl(init.shape[0], max_iter, 1e-4, "euclidean", init, 1, 1)
    model = impl.train(table)
    return model.clusterCenters, model.trainingCost, model.numIter


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n,d,stage_mb", [(1000, 7, 256.0), (4097, 40, 0.25), (33, 256, 256.0), (70001, 64, 1.0)])
def test_columns_land_as_the_reference_rows(m, dtype, n, d, stage_mb):
    rng = np.random.default_rng(n + d)
    cols = tuple(rng.normal(size=n).astype(dtype) * (c + 1) for c in range(d))
    m._lib.check(m.lib().kmb200_set_option(b"stage_mb", stage_mb))   # small staging buffers: several chunks
    t = m.OneDAL.vectorsToMergedNumericTables(cols)
    try:
        got = t.read_rows(0, n)
    finally:
        t.release()
    np.testing.assert_array_equal(got, ko.dense_rows(cols).astype(np.float32))


def test_csr_rows_land_as_the_reference_rows_and_fit_the_same(m):
    sp = pytest.importorskip("scipy.sparse")
    n, d, k = 20000, 96, 24
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=3)
    X[np.abs(X) < 9.0] = 0.0                       # ~85 % zeros, some empty rows
    X[5] = 0.0
    csr = sp.csr_matrix(X.astype(np.float64))
    m._lib.check(m.lib().kmb200_set_option(b"stage_mb", 0.5))
    t = m.OneDAL.vectorsToMergedNumericTables(csr)
    try:
        np.testing.assert_array_equal(t.read_rows(0, n), ko.dense_rows(csr).astype(np.float32))
        init = ko.dense_rows(csr)[np.arange(k) * (n // k)]
        a = gpu_fit(m, t, init)
    finally:
        t.release()
    t2 = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        b = gpu_fit(m, t2, init)
    finally:
        t2.release()
    # (row order inside a cluster's bucket is not fixed run to run, so FP64 sums may differ in the last bits)
    assert a[2] == b[2] and abs(a[1] - b[1]) <= 1e-12 * b[1]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-11, atol=1e-13)
    want_c, want_cost, want_it = ko.lloyd(ko.dense_rows(csr).astype(np.float32).astype(np.float64), init, 1e-4, 5)
    assert a[2] == want_it and abs(a[1] - want_cost) <= 1e-5 * want_cost


def test_csr_rejects_bad_input(m):
    t = m.Table(4, 5)
    try:
        with pytest.raises(m.Kmb200Error):   # column index 5 is outside [0, 5)
            t.append_csr(0, [0, 1, 2, 2, 3], [0, 5, 1], [1.0, 2.0, 3.0])
        with pytest.raises(m.Kmb200Error):   # indptr decreases
            t.append_csr(0, [0, 2, 1, 2, 3], [0, 1, 2], [1.0, 2.0, 3.0])
        with pytest.raises(m.Kmb200Error):   # more rows than the table holds
            t.append_csr(2, [0, 0, 0, 0], [], [])
        t.append_csr(0, [0, 1, 2, 2, 3], [0, 4, 1], [1.0, 2.0, 3.0])
        want = np.zeros((4, 5), np.float32)
        want[0, 0], want[1, 4], want[3, 1] = 1.0, 2.0, 3.0
        np.testing.assert_array_equal(t.read_rows(0, 4), want)
    finally:
        t.release()


def test_table_with_an_upper_bound_is_cut_to_the_rows_appended(m):
    n, d, k = 6000, 48, 20
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=11)
    init = ko.row_init_centers(X, k)
    t = m.Table(10000, d)                         # upper bound, no counting pass
    try:
        t.append(0, X[:2500])
        t.append(2500, X[2500:])
        with pytest.raises(m.Kmb200Error):
            t.truncate(10001)
        with pytest.raises(m.Kmb200Error):
            t.truncate(0)
        t.truncate(n)
        assert m.lib().kmb200_table_rows(t.handle) == n
        a = gpu_fit(m, t, init)
        t.truncate(n // 2)                        # after a fit: the workspace is rebuilt for the new row count
        c = gpu_fit(m, t, init)
    finally:
        t.release()
    t2 = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        b = gpu_fit(m, t2, init)
    finally:
        t2.release()
    # (row order inside a cluster's bucket is not fixed run to run, so FP64 sums may differ in the last bits)
    assert a[2] == b[2] and abs(a[1] - b[1]) <= 1e-12 * b[1]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-11, atol=1e-13)
    want_c, want_cost, want_it = ko.lloyd(X[:n // 2].astype(np.float64), init, 1e-4, 5)
    assert c[2] == want_it and abs(c[1] - want_cost) <= 1e-5 * want_cost


def test_pinned_staging_buffer(m):
    n, d = 3000, 24
    p = C.c_void_p()
    m._lib.check(m.lib().kmb200_pinned_alloc(n * d * 8, C.byref(p)))
    try:
        buf = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(n, d))
        buf[:] = np.random.default_rng(0).normal(size=(n, d))
        t = m.Table(n, d)
        try:
            t.append(0, buf)
            np.testing.assert_array_equal(t.read_rows(0, n), buf.astype(np.float32))
        finally:
            t.release()
    finally:
        m._lib.check(m.lib().kmb200_pinned_free(p))
    assert m.lib().kmb200_pinned_alloc(0, C.byref(p)) != 0
