"""PCA path on the GPU against the FP64 oracle (oracle/pca_oracle.py), through the C ABI
(kmb200_pca_train / kmb200_pca_gram), the Python mirror of PCADALImpl and the JNI natives.

Tolerances.  The reference's own test compares with Spark's RowMatrix PCA at absTol 1e-5 on the explained variance and
on the absolute values of the principal components (IntelPCASuite.scala:70-84).  The CUDA path is held to a tenth of that
on the same quantities -- explained variance: 1e-6 absolute; principal axes with a relative eigenvalue gap >= 1e-2: 1e-6
absolute, signs included (both sides use the "largest entry positive" rule) -- and, beyond what the reference checks:
eigenvalues 2e-6 relative to the largest, every entry of the centred Gram matrix 2e-6 of sqrt(G_ii G_jj).  (Measured:
7e-7 on the Gram matrix -- the FP32 accumulation of the tensor core over 128 rows -- and 1e-8 on the components.)"""
import ctypes as C

import numpy as np
import pytest

from oracle import pca_oracle as po

pytestmark = pytest.mark.gpu
GRAM_TOL, EIG_TOL, EXPL_TOL, AXIS_TOL = 2e-6, 2e-6, 1e-6, 1e-6


@pytest.fixture(scope="module")
def m(built_lib):
    import oap_mllib_b200 as mod
    assert mod.lib().kmb200_device_ok() == 1, "no sm_100 GPU visible: the CUDA path cannot run"
    return mod


def decaying_table(n, d, seed, shift=3.0, dtype=np.float32):
    """Rows with a geometrically decaying spectrum in a random orthogonal basis, plus a column offset."""
    rng = np.random.default_rng(seed)
    basis = np.linalg.qr(rng.standard_normal((d, d)))[0]
    sig = 10.0 * 0.93 ** np.arange(d) + 0.05
    return ((rng.standard_normal((n, d)) * sig) @ basis + rng.uniform(-shift, shift, d)).astype(dtype)


def check_against_oracle(m, parts, center=True, k=None):
    """Single-GPU: the rows of all parts in one table.  Returns (eigenvalues, axes) of the CUDA path."""
    from oap_mllib_b200 import pca
    X = np.vstack(parts)
    n, d = X.shape
    table = m.OneDAL.vectorsToMergedNumericTables(list(parts))
    try:
        g, mean, cnt = pca.gram(table, center)
        ev = np.empty(d)
        axes = np.empty((d, d))
        m._lib.check(m.lib().kmb200_pca_train(table.handle, None, 1 if center else 0, ev.ctypes.data_as(C.c_void_p),
                                              axes.ctypes.data_as(C.c_void_p)))
    finally:
        table.release()
    X32 = [np.asarray(p, np.float32) for p in parts]     # the table holds FP32: the oracle sees the same values
    g_ref, mean_ref, n_ref = po.centred_gram(X32, center)
    assert cnt == n_ref == n
    assert np.allclose(mean, mean_ref, rtol=0, atol=1e-12 * max(1.0, np.abs(mean_ref).max()))
    scale = np.sqrt(np.outer(np.diag(g_ref), np.diag(g_ref)))
    scale[scale == 0] = 1.0
    assert (np.abs(g - g_ref) / scale).max() <= GRAM_TOL
    assert np.array_equal(g, g.T)
    ev_ref, axes_ref = po.pca_svd_dense(X32, center)      # the reference's route (QR per block + SVD of the factors)
    assert np.abs(ev - ev_ref).max() <= EIG_TOL * ev_ref[0]
    assert np.all(np.diff(ev) <= 0) and ev.min() >= 0
    assert np.abs(axes @ axes.T - np.eye(d)).max() <= 1e-12
    total = ev_ref.sum()
    assert np.abs(ev / ev.sum() - ev_ref / total).max() <= EXPL_TOL
    # axes: those whose eigenvalue is separated from its neighbours (an axis inside a cluster of eigenvalues is not defined)
    gaps = np.full(d, np.inf)
    gaps[:-1] = np.minimum(gaps[:-1], ev_ref[:-1] - ev_ref[1:])
    gaps[1:] = np.minimum(gaps[1:], ev_ref[:-1] - ev_ref[1:])
    stable = gaps >= 1e-2 * ev_ref[0]
    if k is not None:
        stable[k:] = False
    assert stable.any()
    assert np.abs(axes[stable] - axes_ref[stable]).max() <= AXIS_TOL
    return ev, axes


def test_reference_fixture_through_the_estimator(m):
    """IntelPCASuite.scala:40-85 / PCAExample.scala:37-50: the 3 x 5 matrix, k = 3, against Spark's published numbers."""
    from test_pca_oracle import FIXTURE, SPARK_EXPLAINED, SPARK_PROJECTION
    model = m.PCA(k=3).fit(FIXTURE)
    assert model.pc.shape == (5, 3) and model.explainedVariance.shape == (3,)
    assert np.allclose(model.explainedVariance[:2], SPARK_EXPLAINED, rtol=0, atol=1e-6)     # the reference's absTol is 1e-5
    assert abs(model.explainedVariance[2]) <= 1e-6
    assert np.allclose(np.abs(model.transform(FIXTURE)[:, :2]), np.abs(SPARK_PROJECTION), rtol=0, atol=1e-5)
    with pytest.raises(ValueError):
        m.PCA(k=6).fit(FIXTURE)                                                             # PCA.scala:93-94


@pytest.mark.parametrize("n,d", [(1, 4), (2, 3), (31, 17), (33, 32), (257, 64), (1000, 100), (5000, 256), (4097, 255)])
def test_small_tables(m, n, d):
    """One stage, ragged last stage, fewer stages than CTA pairs, d below / at one 256-column panel."""
    check_against_oracle(m, [decaying_table(n, d, seed=n + d)], center=n > 1)


@pytest.mark.parametrize("n,d", [(3000, 257), (4000, 300), (3000, 512), (2500, 700)])
def test_several_column_panels(m, n, d):
    """d > 256: panel pairs off the diagonal (two strips per stage, three products)."""
    check_against_oracle(m, [decaying_table(n, d, seed=d)])


@pytest.mark.parametrize("center", [True, False])
def test_long_table_accumulation(m, center):
    """300 000 rows: every CTA pair runs tens of level-1 segments and several promotions to FP64; partitions of uneven size."""
    X = decaying_table(300_000, 256, seed=11, shift=40.0)     # means ~ 4 sigma of the leading axis: centring matters
    check_against_oracle(m, [X[:70_001], X[70_001:70_002], X[70_002:]], center=center)


@pytest.mark.parametrize("n,d", [(100_000, 128), (70_001, 64), (50_000, 33), (90_003, 32), (60_000, 7)])
def test_narrow_tables_with_and_without_the_row_split(m, n, d):
    """d <= 128: the two CTAs of a pair take different stages; d <= 64: and two stages each, side by side in the strip
    (default, "pca_split" = 2); with 1 / 0 only the first / neither.  All against the oracle, ragged numbers of stages."""
    X = decaying_table(n, d, seed=n)
    try:
        for split in (2.0, 1.0, 0.0):
            m._lib.check(m.lib().kmb200_set_option(b"pca_split", split))
            check_against_oracle(m, [X])
    finally:
        m._lib.check(m.lib().kmb200_set_option(b"pca_split", 2.0))


def test_column_scales_spread_over_30_orders(m):
    """Every column has its own power-of-two scale: a column of 1e-20s keeps its relative accuracy next to one of 1e+10s
    (one table-wide scale would flush it to zero in FP16)."""
    rng = np.random.default_rng(5)
    n, d = 20_000, 48
    mag = 10.0 ** rng.uniform(-20, 10, d)                      # per-column magnitude; the offset is of the same order
    X = ((rng.standard_normal((n, d)) + 2.0 * rng.standard_normal(d)) * mag).astype(np.float32)
    from oap_mllib_b200 import pca
    table = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        g, _, _ = pca.gram(table, True)
    finally:
        table.release()
    g_ref, _, _ = po.centred_gram([X], True)
    assert (np.abs(g - g_ref) / np.sqrt(np.outer(np.diag(g_ref), np.diag(g_ref)))).max() <= GRAM_TOL


def test_constant_and_zero_columns(m):
    X = decaying_table(5000, 40, seed=9)
    X[:, 3] = 0.0
    X[:, 7] = 12.5
    ev, axes = check_against_oracle(m, [X])
    assert ev[-1] <= 1e-9 * ev[0] and ev[-2] <= 1e-9 * ev[0]


def test_results_are_bit_reproducible(m):
    from oap_mllib_b200 import pca
    X = decaying_table(200_000, 192, seed=21)
    table = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        first = pca.gram(table, True)[0]
        for _ in range(5):
            assert np.array_equal(pca.gram(table, True)[0], first)
    finally:
        table.release()


def test_column_statistics_follow_the_appends(m):
    """The column sums / maxima the centring and the column scales need are kept with the table by appends that continue
    where the last one ended; an overwrite in the middle makes the PCA call recompute them (three more kernels)."""
    from oap_mllib_b200 import pca

    def gram_matches(table, X):
        g, mean, n = pca.gram(table, True)
        g_ref, mean_ref, _ = po.centred_gram([X], True)
        assert n == X.shape[0] and np.allclose(mean, mean_ref, rtol=0, atol=1e-12 * np.abs(mean_ref).max())
        assert (np.abs(g - g_ref) / np.sqrt(np.outer(np.diag(g_ref), np.diag(g_ref)))).max() <= GRAM_TOL
        return m._lib.profile()["kernel_launches"]

    X = decaying_table(50_000, 40, seed=13, shift=20.0)
    t = m.Table(60_000, 40)                   # sized for more rows than arrive, truncated below (OneDAL.scala:150-160 does the same)
    try:
        t.append(0, X[:20_000])
        t.append(20_000, X[20_000:])
        t.truncate(50_000)
        kept = gram_matches(t, X)
        X2 = X.copy()
        X2[1000:2000] = X2[1000:2000] * 3.0 + 7.0
        t.append(1000, X2[1000:2000])          # out of order: the running statistics are dropped ...
        assert gram_matches(t, X2) == kept + 3     # ... and recomputed by the call
        assert gram_matches(t, X2) == kept         # (and then kept again)
        t.append(0, X[:30_000])                # a new upload from row 0 starts them over
        t.append(30_000, X[30_000:])
        assert gram_matches(t, X) == kept
    finally:
        t.release()


def test_pcadalimpl_mirror_and_model(m):
    """PCADALImpl.train -> MLlibPCAModel: pc is d x k with one component per column, explainedVariance sums below 1."""
    X = decaying_table(20_000, 30, seed=2)
    model = m.PCADALImpl(k=5, executorNum=1, executorCores=1).train([X[:9000], X[9000:]])
    ev_ref, axes_ref = po.pca_svd_dense([X])
    pc_ref, expl_ref = po.model(ev_ref, axes_ref, 5)
    assert model.k == 5 and model.pc.shape == (30, 5)
    assert np.abs(model.pc - pc_ref).max() <= AXIS_TOL and np.abs(model.explainedVariance - expl_ref).max() <= EXPL_TOL


def test_errors(m):
    t = m.Table(10, 4)
    with pytest.raises(m.Kmb200Error):
        m._lib.check(m.lib().kmb200_pca_train(t.handle, None, 1, None, None))
    t.release()


def test_jni_pca_natives(built_lib):
    """Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDAL (the reference's symbol: rows taken as they are) and
    ..._cPCATrainDALCentered, driven through the fake JNIEnv of test_jni_shim.py."""
    from test_jni_shim import P, FakeJvm, fn
    h, vm = C.CDLL(built_lib), FakeJvm()
    X = decaying_table(6000, 20, seed=4).astype(np.float64)
    n, d = X.shape
    new_table = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable", C.c_int64, [P, P, C.c_int64, C.c_int32])
    set_batch = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cSetDoubleBatch", None, [P, P, C.c_int64, C.c_int32, P, C.c_int32, C.c_int32])
    get_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cGetCenters", P, [P, P, C.c_int64])
    free_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeCenters", None, [P, P, C.c_int64])
    free_tab = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeDataMemory", None, [P, P, C.c_int64])
    c_init = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1init", C.c_int32, [P, P, C.c_int32, C.c_int32, P, P])
    cleanup = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1cleanup", None, [P, P])
    sig = [P, P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, P]
    train = fn(h, "Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDAL", C.c_int64, sig)
    train_centered = fn(h, "Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDALCentered", C.c_int64, sig)
    table = new_table(vm.env, None, n, d)
    assert table != 0 and not vm.thrown, vm.thrown
    blk = np.ascontiguousarray(X)
    set_batch(vm.env, None, table, 0, vm.new(blk), n, d)
    assert c_init(vm.env, None, 1, 0, vm.new(np.zeros(128, np.int8)), vm.new(object())) == 1
    X32 = X.astype(np.float32)
    for native, center in ((train, False), (train_centered, True)):
        result = vm.new(object())
        assert native(vm.env, None, table, 3, 1, 1, result) == 0 and not vm.thrown, vm.thrown
        f = vm.fields[result]
        axes = vm.objs[get_cent(vm.env, None, f["pcNumericTable"])].reshape(d, d)
        ev = vm.objs[get_cent(vm.env, None, f["explainedVarianceNumericTable"])]
        ev_ref, axes_ref = po.pca_svd_dense([X32], center)
        assert ev.shape == (d,) and np.abs(ev - ev_ref).max() <= EIG_TOL * ev_ref[0]
        assert np.abs(axes[:3] - axes_ref[:3]).max() <= AXIS_TOL
        free_cent(vm.env, None, f["pcNumericTable"])
        free_cent(vm.env, None, f["explainedVarianceNumericTable"])
    assert train(vm.env, None, 0, 3, 1, 1, vm.new(object())) == 0 and vm.thrown      # null handle -> Java exception
    cleanup(vm.env, None)
    free_tab(vm.env, None, table)
