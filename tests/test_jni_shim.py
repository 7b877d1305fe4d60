"""The JNI layer (csrc/jni_shim.cpp) driven through a hand-built JNIEnv function table: there is no
JVM in this image, so the slots the shim uses are filled with ctypes callbacks.  Slot numbers are
those of the JNI specification's interface function table (mirrored in csrc/jni_min.h)."""
import ctypes as C
import os

import numpy as np
import pytest

SLOTS = dict(FindClass=6, ThrowNew=14, GetObjectClass=31, GetFieldID=94, SetIntField=109, SetLongField=110,
             SetDoubleField=112, GetStringUTFChars=169, ReleaseStringUTFChars=170, GetArrayLength=171, GetObjectArrayElement=173,
             NewByteArray=176, NewLongArray=180, NewDoubleArray=182, GetByteArrayRegion=200, SetByteArrayRegion=208,
             SetIntArrayRegion=211, SetLongArrayRegion=212, SetDoubleArrayRegion=214, GetPrimitiveArrayCritical=222, ReleasePrimitiveArrayCritical=223)
P = C.c_void_p


class FakeJvm:
    """Objects are small integers; arrays map to numpy buffers; fields are recorded in dicts."""

    def __init__(self):
        self.objs, self.fields, self.field_names, self.thrown, self.keep = {}, {}, {}, [], []
        self.next = 100
        table = (P * 235)()
        def put(name, restype, argtypes, fn):
            cb = C.CFUNCTYPE(restype, *argtypes)(fn)
            self.keep.append(cb)
            table[SLOTS[name]] = C.cast(cb, P)
        put("FindClass", P, [P, C.c_char_p], lambda env, n: 7)
        put("ThrowNew", C.c_int, [P, P, C.c_char_p], lambda env, c, msg: self.thrown.append(msg.decode()) or 0)
        put("GetObjectClass", P, [P, P], lambda env, o: 8)
        put("GetFieldID", P, [P, P, C.c_char_p, C.c_char_p], self._field_id)
        put("SetIntField", None, [P, P, P, C.c_int32], self._set_field)
        put("SetLongField", None, [P, P, P, C.c_int64], self._set_field)
        put("SetDoubleField", None, [P, P, P, C.c_double], self._set_field)
        put("GetStringUTFChars", P, [P, P, P], lambda env, s, c: C.cast(self.objs[s], P).value)
        put("ReleaseStringUTFChars", None, [P, P, P], lambda env, s, c: None)
        put("NewByteArray", P, [P, C.c_int32], lambda env, n: self.new(np.zeros(n, np.int8)))
        put("NewDoubleArray", P, [P, C.c_int32], lambda env, n: self.new(np.zeros(n, np.float64)))
        put("NewLongArray", P, [P, C.c_int32], lambda env, n: self.new(np.zeros(n, np.int64)))
        put("SetIntArrayRegion", None, [P, P, C.c_int32, C.c_int32, P],
            lambda env, a, s, n, buf: C.memmove(self.objs[a].ctypes.data + 4 * s, buf, 4 * n) and None)
        put("SetLongArrayRegion", None, [P, P, C.c_int32, C.c_int32, P],
            lambda env, a, s, n, buf: C.memmove(self.objs[a].ctypes.data + 8 * s, buf, 8 * n) and None)
        put("GetByteArrayRegion", None, [P, P, C.c_int32, C.c_int32, P],
            lambda env, a, s, n, buf: C.memmove(buf, self.objs[a].ctypes.data + s, n) and None)
        put("SetByteArrayRegion", None, [P, P, C.c_int32, C.c_int32, P],
            lambda env, a, s, n, buf: C.memmove(self.objs[a].ctypes.data + s, buf, n) and None)
        put("SetDoubleArrayRegion", None, [P, P, C.c_int32, C.c_int32, P],
            lambda env, a, s, n, buf: C.memmove(self.objs[a].ctypes.data + 8 * s, buf, 8 * n) and None)
        put("GetObjectArrayElement", P, [P, P, C.c_int32], lambda env, a, i: self.objs[a][i])
        put("GetPrimitiveArrayCritical", P, [P, P, P], lambda env, a, c: self.objs[a].ctypes.data)
        put("ReleasePrimitiveArrayCritical", None, [P, P, P, C.c_int32], lambda env, a, p, m: None)
        self.table = table
        self.table_ptr = P(C.addressof(table))          # JNIEnv  = pointer to the table
        self.env = C.pointer(self.table_ptr)             # JNIEnv* = what natives receive

    def new(self, obj):
        self.next += 1
        self.objs[self.next] = obj
        return self.next

    def string(self, s):
        return self.new(C.create_string_buffer(s.encode()))

    def _field_id(self, env, cls, name, sig):
        fid = 1000 + len(self.field_names)
        self.field_names[fid] = name.decode()
        return fid

    def _set_field(self, env, obj, fid, value):
        self.fields.setdefault(obj, {})[self.field_names[fid]] = value


@pytest.fixture()
def jni(built_lib):
    h = C.CDLL(built_lib)
    return h, FakeJvm()


def fn(h, name, restype, argtypes):
    f = getattr(h, name)
    f.restype, f.argtypes = restype, argtypes
    return f


def test_platform_gate_and_setenv(jni):
    h, vm = jni
    ok = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cCheckPlatformCompatibility", C.c_uint8, [P, P])(vm.env, None)
    assert ok in (0, 1)
    setenv = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_setEnv", C.c_int32, [P, P, P, P, C.c_uint8])
    assert setenv(vm.env, None, vm.string("KMB200_JNI_TEST"), vm.string("ofi"), 1) == 0
    assert os.environ.get("KMB200_JNI_TEST") == "ofi" or True     # os.environ caches; the libc env is what counts
    libc = C.CDLL(None)
    libc.getenv.restype = C.c_char_p
    assert libc.getenv(b"KMB200_JNI_TEST") == b"ofi"
    assert fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_isRoot", C.c_uint8, [P, P])(vm.env, None) == 1
    assert fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_rankID", C.c_int32, [P, P])(vm.env, None) == 0


@pytest.mark.gpu
def test_jni_fit_matches_oracle(jni):
    from oracle import kmeans_oracle as ko
    h, vm = jni
    X, _, _ = ko.gaussian_blobs(5000, 24, 9, seed=3)
    X = X.astype(np.float64)
    init = ko.row_init_centers(X, 9)
    n, d = X.shape
    new_table = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable", C.c_int64, [P, P, C.c_int64, C.c_int32])
    set_batch = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cSetDoubleBatch", None,
                   [P, P, C.c_int64, C.c_int32, P, C.c_int32, C.c_int32])
    new_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewCenters", C.c_int64, [P, P, C.c_int32, C.c_int32])
    set_val = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_setNumericTableValue", None,
                 [P, P, C.c_int64, C.c_int32, C.c_int32, C.c_double])
    get_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cGetCenters", P, [P, P, C.c_int64])
    free_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeCenters", None, [P, P, C.c_int64])
    free_tab = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeDataMemory", None, [P, P, C.c_int64])
    c_init = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1init", C.c_int32, [P, P, C.c_int32, C.c_int32, P, P])
    cleanup = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1cleanup", None, [P, P])
    compute = fn(h, "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALComputeWithInitCenters", C.c_int64,
                 [P, P, C.c_int64, C.c_int64, C.c_int32, C.c_double, C.c_int32, C.c_int32, C.c_int32, P])

    table = new_table(vm.env, None, n, d)
    assert table != 0 and not vm.thrown, vm.thrown
    for r0 in range(0, n, 2048):                      # row batches, like the patched OneDAL.scala
        blk = np.ascontiguousarray(X[r0:r0 + 2048])
        set_batch(vm.env, None, table, r0, vm.new(blk), blk.shape[0], d)
    cent = new_cent(vm.env, None, 9, d)
    for i in range(9):
        for j in range(d):
            set_val(vm.env, None, cent, i, j, float(init[i, j]))
    param = vm.new(object())
    assert c_init(vm.env, None, 1, 0, vm.new(np.zeros(128, np.int8)), param) == 1
    assert vm.fields[param] == {"commSize": 1, "rankId": 0}
    result = vm.new(object())
    out = compute(vm.env, None, table, cent, 9, 1e-4, 15, 1, 1, result)
    assert out != 0 and not vm.thrown, vm.thrown
    got = vm.objs[get_cent(vm.env, None, out)].reshape(9, d)
    want_c, want_cost, want_it = ko.lloyd(X.astype(np.float32).astype(np.float64), init, 1e-4, 15)
    assert vm.fields[result]["iterationNum"] == want_it
    assert abs(vm.fields[result]["totalCost"] - want_cost) <= 1e-5 * want_cost
    np.testing.assert_allclose(got, want_c, rtol=1e-5, atol=1e-5)
    cleanup(vm.env, None)
    free_cent(vm.env, None, out)
    free_cent(vm.env, None, cent)
    free_tab(vm.env, None, table)
    # errors surface as a Java exception, not as a crash: a table with a wrong column count
    bad = new_cent(vm.env, None, 2, d + 1)
    t2 = new_table(vm.env, None, 10, d)
    assert compute(vm.env, None, t2, bad, 2, 1e-4, 1, 1, 1, result) == 0
    assert vm.thrown and "dimension" in vm.thrown[-1]
    free_cent(vm.env, None, bad)
    free_tab(vm.env, None, t2)


@pytest.mark.gpu
def test_jni_columnar_sparse_and_truncate(jni):
    """cSetColumnsBatch (double[][]), cSetSparseBatch (CSR arrays) and cTruncateTable land the same rows as the
    row-batch native; checked by reading the table back through the C ABI."""
    from oracle import kmeans_oracle as ko
    h, vm = jni
    n, d = 700, 9
    X = np.random.default_rng(2).normal(size=(n, d))
    X[np.abs(X) < 0.8] = 0.0
    new_table = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable", C.c_int64, [P, P, C.c_int64, C.c_int32])
    set_cols = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cSetColumnsBatch", None,
                  [P, P, C.c_int64, C.c_int32, P, C.c_int32, C.c_int32])
    set_sparse = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cSetSparseBatch", None,
                    [P, P, C.c_int64, C.c_int32, P, P, P, C.c_int32])
    truncate = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cTruncateTable", None, [P, P, C.c_int64, C.c_int64])
    free_tab = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeDataMemory", None, [P, P, C.c_int64])
    read_rows = fn(h, "kmb200_table_read_rows", C.c_int, [P, C.c_int64, C.c_int64, P])
    rows = fn(h, "kmb200_table_rows", C.c_int64, [P])

    table = new_table(vm.env, None, 2 * n + 100, d)            # upper bound: no counting pass
    cols = [np.ascontiguousarray(X[:, c]) for c in range(d)]
    set_cols(vm.env, None, table, 0, vm.new([vm.new(c) for c in cols]), n, d)
    indptr = np.zeros(n + 1, np.int64)
    indices, values = [], []
    for r in range(n):
        nz = np.flatnonzero(X[r])
        indices.extend(nz.tolist())
        values.extend(X[r, nz].tolist())
        indptr[r + 1] = len(indices)
    set_sparse(vm.env, None, table, n, vm.new(indptr), vm.new(np.array(indices, np.int32)),
               vm.new(np.array(values, np.float64)), n)
    truncate(vm.env, None, table, 2 * n)
    assert not vm.thrown, vm.thrown
    assert rows(table) == 2 * n
    got = np.empty((2 * n, d), np.float32)
    assert read_rows(table, 0, 2 * n, got.ctypes.data_as(P)) == 0
    want = ko.dense_rows(X).astype(np.float32)
    np.testing.assert_array_equal(got[:n], want)
    np.testing.assert_array_equal(got[n:], want)
    truncate(vm.env, None, table, 10 ** 9)                      # beyond the capacity: a Java exception, not a crash
    assert vm.thrown and "truncate" in vm.thrown[-1]
    free_tab(vm.env, None, table)


@pytest.mark.gpu
def test_jni_summary_and_init_natives(jni):
    """cKMeansDALAssign (the summary pass of KMeans.fit on the table that is still resident), cKMeansInitCosts and
    cKMeansInitSample (the row passes of initKMeansParallel) against the oracle."""
    from oracle import kmeans_oracle as ko
    from oracle import spark_init as si
    h, vm = jni
    n, d, k = 8000, 12, 7
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=5)
    X64 = X.astype(np.float64)
    new_table = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable", C.c_int64, [P, P, C.c_int64, C.c_int32])
    set_batch = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cSetDoubleBatch", None,
                   [P, P, C.c_int64, C.c_int32, P, C.c_int32, C.c_int32])
    new_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewCenters", C.c_int64, [P, P, C.c_int32, C.c_int32])
    set_val = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_setNumericTableValue", None,
                 [P, P, C.c_int64, C.c_int32, C.c_int32, C.c_double])
    free_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeCenters", None, [P, P, C.c_int64])
    free_tab = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeDataMemory", None, [P, P, C.c_int64])
    assign = fn(h, "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALAssign", C.c_double, [P, P, C.c_int64, C.c_int64, P, P])
    init_costs = fn(h, "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansInitCosts", C.c_double, [P, P, C.c_int64, C.c_int64, C.c_uint8])
    init_sample = fn(h, "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansInitSample", P,
                     [P, P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_double, P, C.c_int32, C.c_int32])

    def centres_handle(cen):
        hd = new_cent(vm.env, None, cen.shape[0], d)
        for i in range(cen.shape[0]):
            for j in range(d):
                set_val(vm.env, None, hd, i, j, float(cen[i, j]))
        return hd

    table = new_table(vm.env, None, n, d)
    set_batch(vm.env, None, table, 0, vm.new(np.ascontiguousarray(X64)), n, d)
    # summary pass
    cen = ko.row_init_centers(X, k)
    hc = centres_handle(cen)
    pred, sizes = vm.new(np.zeros(n, np.int32)), vm.new(np.zeros(k, np.int64))
    cost = assign(vm.env, None, table, hc, pred, sizes)
    assert not vm.thrown, vm.thrown
    want_labels, best, _ = ko.assign(X64, cen)
    assert np.array_equal(vm.objs[pred], want_labels)
    assert np.array_equal(vm.objs[sizes], np.bincount(want_labels, minlength=k))
    assert abs(cost - best.sum()) <= 1e-5 * best.sum()
    free_cent(vm.env, None, hc)
    # one k-means|| step: costs to the first centre, then the sampling filter over two partitions
    first = X64[si.first_center_row(123, n)][None, :]
    hf = centres_handle(first)
    total = init_costs(vm.env, None, table, hf, 1)
    want_costs = ((X64 - first[0]) ** 2).sum(1)
    assert abs(total - want_costs.sum()) <= 1e-12 * want_costs.sum()
    parts = np.array([0, 3000, n], np.int64)
    got = init_sample(vm.env, None, table, 123, 0, k, total, vm.new(parts), 2, 4)
    want = []
    for p in range(2):
        rand = si.XORShiftRandom(si._to_i32(123 ^ (0 << 16) ^ (4 + p)))
        for r in range(parts[p], parts[p + 1]):
            if rand.next_double() < 2.0 * want_costs[r] * k / total:
                want.append(r)
    assert not vm.thrown, vm.thrown
    assert vm.objs[got].tolist() == want and len(want) > 0
    free_cent(vm.env, None, hf)
    free_tab(vm.env, None, table)


def _jni_rank(rank, world, lib_path, q_id, q_out):
    """One executor process: device, table, rendezvous through the JNI natives, collective compute."""
    from oracle import kmeans_oracle as ko
    h, vm = C.CDLL(lib_path), FakeJvm()
    n, d, k = 40000, 32, 20
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=77)
    init = ko.row_init_centers(X, k)
    r0, r1 = (n * rank) // world, (n * (rank + 1)) // world
    set_dev = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1setDevice", C.c_int32, [P, P, C.c_int32])
    get_id = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1getUniqueId", P, [P, P])
    c_init = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1init", C.c_int32, [P, P, C.c_int32, C.c_int32, P, P])
    cleanup = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_c_1cleanup", None, [P, P])
    is_root = fn(h, "Java_org_apache_spark_ml_util_OneCCL_00024_isRoot", C.c_uint8, [P, P])
    new_table = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable", C.c_int64, [P, P, C.c_int64, C.c_int32])
    set_batch = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cSetDoubleBatch", None,
                   [P, P, C.c_int64, C.c_int32, P, C.c_int32, C.c_int32])
    new_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cNewCenters", C.c_int64, [P, P, C.c_int32, C.c_int32])
    set_val = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_setNumericTableValue", None,
                 [P, P, C.c_int64, C.c_int32, C.c_int32, C.c_double])
    get_cent = fn(h, "Java_org_apache_spark_ml_util_OneDAL_00024_cGetCenters", P, [P, P, C.c_int64])
    compute = fn(h, "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALComputeWithInitCenters", C.c_int64,
                 [P, P, C.c_int64, C.c_int64, C.c_int32, C.c_double, C.c_int32, C.c_int32, C.c_int32, P])
    assert set_dev(vm.env, None, rank) == rank
    if rank == 0:
        ident = bytes(vm.objs[get_id(vm.env, None)].view(np.uint8))
        for _ in range(world - 1):
            q_id.put(ident)
    else:
        ident = q_id.get(timeout=120)
    table = new_table(vm.env, None, r1 - r0, d)
    set_batch(vm.env, None, table, 0, vm.new(np.ascontiguousarray(X[r0:r1], np.float64)), r1 - r0, d)
    cent = new_cent(vm.env, None, k, d)
    for i in range(k):
        for j in range(d):
            set_val(vm.env, None, cent, i, j, float(init[i, j]))
    param, result = vm.new(object()), vm.new(object())
    assert c_init(vm.env, None, world, rank, vm.new(np.frombuffer(ident, np.int8).copy()), param) == 1
    assert vm.fields[param] == {"commSize": world, "rankId": rank}
    out = compute(vm.env, None, table, cent, k, 1e-4, 10, world, 1, result)
    ok = not vm.thrown
    if rank == 0:
        assert is_root(vm.env, None) == 1 and out != 0
        got = vm.objs[get_cent(vm.env, None, out)].reshape(k, d)
        want_c, want_cost, want_it = ko.lloyd(X.astype(np.float64), init, 1e-4, 10, n_ranks=world)
        ok = ok and vm.fields[result]["iterationNum"] == want_it and abs(vm.fields[result]["totalCost"] - want_cost) <= 1e-5 * want_cost \
            and np.allclose(got, want_c, rtol=1e-5, atol=1e-5)
    else:
        ok = ok and out == 0
    # the PCA native on the same tables and communicator (section 8 row f4): the result fields are set on rank 0 only
    from oracle import pca_oracle as po
    pca = fn(h, "Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDALCentered", C.c_int64,
             [P, P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, P])
    pres = vm.new(object())
    ok = ok and pca(vm.env, None, table, 3, world, 1, pres) == 0 and not vm.thrown
    if rank == 0:
        f = vm.fields[pres]
        axes = vm.objs[get_cent(vm.env, None, f["pcNumericTable"])].reshape(d, d)
        ev = vm.objs[get_cent(vm.env, None, f["explainedVarianceNumericTable"])]
        X32 = X.astype(np.float32)
        ev_ref, axes_ref = po.pca_svd_dense([X32[(n * r) // world:(n * (r + 1)) // world] for r in range(world)])
        ok = ok and np.abs(ev - ev_ref).max() <= 2e-6 * ev_ref[0] and np.abs(axes[:3] - axes_ref[:3]).max() <= 1e-6
    else:
        ok = ok and pres not in vm.fields
    cleanup(vm.env, None)
    q_out.put((rank, bool(ok), vm.thrown))


@pytest.mark.gpu
def test_jni_two_executors_rendezvous_and_fit(built_lib):
    """Two executor processes, one GPU each: c_setDevice, c_getUniqueId on rank 0 (the bytes travel to rank 1 the way
    the Spark closure would carry them), c_init(size = 2), the collective compute natives (K-means, then PCA), c_cleanup."""
    import multiprocessing as mp

    import oap_mllib_b200 as m
    if m.lib().kmb200_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q_id, q_out = ctx.Queue(), ctx.Queue()
    procs = [ctx.Process(target=_jni_rank, args=(r, 2, built_lib, q_id, q_out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
