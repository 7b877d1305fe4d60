// JNI layer of libkmeans_b200.so: the symbols the reference's Scala objects bind, re-implemented
// over the C ABI (include/kmb200.h).  Names follow the javah headers of the reference
// (mllib-dal/src/main/native/javah/*.h; "_00024" is the '$' of a Scala object):
//
//   kept (same name, same JNI signature)
//     KMeansDALImpl.cKMeansDALComputeWithInitCenters (JJIDIIILorg/.../KMeansResult;)J
//                                   javah/org_apache_spark_ml_clustering_KMeansDALImpl.h:10-16
//     OneDAL$.cSetDoubleBatch (JI[DII)V, setNumericTableValue (JIID)V, cFreeDataMemory (J)V,
//     cCheckPlatformCompatibility ()Z          javah/org_apache_spark_ml_util_OneDAL__.h
//     OneCCL$.c_cleanup ()V, isRoot ()Z, rankID ()I, setEnv (Ljava/lang/String;Ljava/lang/String;Z)I
//                                   javah/org_apache_spark_ml_util_OneCCL__.h
//   changed / new (the oneDAL Java classes DALMatrix/HomogenNumericTable are gone)
//     OneDAL$.cNewTable (JI)J, cNewCenters (II)J, cGetCenters (J)[D, cFreeCenters (J)V
//     OneCCL$.c_getUniqueId ()[B, c_init (II[BLorg/apache/spark/ml/util/CCLParam;)I, c_setDevice (I)I
//     KMeansDALImpl.cKMeansDALAssign (JJ[I[J)D   -- the summary pass (labels, cluster sizes, cost) on the GPU
//
// No JVM exists in this image: the layer is exercised through a hand-built JNIEnv function table
// (tests/test_jni_shim.py).  It cannot be linked against a real JVM here -- INTEGRATION.md says so.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <sys/file.h>
#include <unistd.h>
#include <mutex>
#include <cstdio>
#include <vector>

#include "../../include/kmb200.h"
#include "jni_min.h"

namespace {

template <typename Fn>
Fn slot(JNIEnvPtr env, int idx) {
    return reinterpret_cast<Fn>((*env)[idx]);
}

void throw_runtime(JNIEnvPtr env, const char *msg) {
    auto find = slot<jclass (*)(JNIEnvPtr, const char *)>(env, kJniFindClass);
    auto thrw = slot<jint (*)(JNIEnvPtr, jclass, const char *)>(env, kJniThrowNew);
    if (!find || !thrw) return;
    jclass cls = find(env, "java/lang/RuntimeException");
    if (cls) thrw(env, cls, msg);
}

struct Centers {  // host k x d FP64 buffer handed around as a jlong (replaces NumericTablePtr*)
    int k, d;
    std::vector<double> v;
};

kmb200_comm *g_comm = nullptr;  // process-global like the reference's g_comms (native/OneCCL.cpp:38-40)
int g_world = 1, g_rank = 0;
int g_device = -1;              // set by c_setDevice; -1 = not chosen yet (c_init picks one)

// Page-locked staging buffer of the calling thread.  Java arrays are copied into it inside a SHORT critical
// region (a memcpy, nothing else: no CUDA call, no allocation, no blocking while the garbage collector is held
// off), the region is released, and only then does the upload run -- from pinned memory, i.e. at the link rate.
struct Staging {
    void *p = nullptr;
    size_t bytes = 0;
    ~Staging() { if (p) kmb200_pinned_free(p); }
    void *get(size_t want) {
        if (bytes >= want) return p;
        if (p) kmb200_pinned_free(p);
        p = nullptr;
        bytes = 0;
        if (kmb200_pinned_alloc(want, &p) != 0) return nullptr;
        bytes = want;
        return p;
    }
};
thread_local Staging g_staging;

// Which GPU this executor process uses.  Order: the ordinal given to c_setDevice (the Scala side passes
// spark.oap.mllib.b200.device or the address Spark's GPU scheduling assigned to the task); the environment variable
// KMB200_DEVICE; else the first GPU of the node that no other executor process has claimed -- a claim is an
// exclusive flock on /tmp/kmb200-gpu-<i>.lock held until the process exits, so that executors which all see every
// GPU do not pile up on GPU 0.  Chosen once per process, before the first table is created.
std::mutex g_device_mu;
int resolve_device() {
    std::lock_guard<std::mutex> lk(g_device_mu);
    if (g_device >= 0) return g_device;
    const int count = kmb200_device_count();
    if (count <= 0) return 0;   // the compute calls will report the missing device
    if (const char *e = getenv("KMB200_DEVICE")) {
        g_device = atoi(e) % count;
        if (g_device < 0) g_device = 0;
        return g_device;
    }
    for (int i = 0; i < count && g_device < 0; ++i) {
        char path[64];
        snprintf(path, sizeof(path), "/tmp/kmb200-gpu-%d.lock", i);
        const int fd = open(path, O_CREAT | O_RDWR, 0666);
        if (fd < 0) continue;
        if (flock(fd, LOCK_EX | LOCK_NB) == 0) g_device = i;   // fd stays open: the claim lasts as long as the process
        else close(fd);
    }
    if (g_device < 0) g_device = (int)(getpid() % count);   // more executors than GPUs on this node
    return g_device;
}
bool use_device(JNIEnvPtr env) {
    if (kmb200_device_count() <= 0) return true;   // nothing to select; compute calls fail with their own message
    if (kmb200_set_device(resolve_device()) != 0) {
        throw_runtime(env, kmb200_last_error());
        return false;
    }
    return true;
}
constexpr size_t kStagingBytes = (size_t)64 << 20;

// Copies count elements of a primitive Java array, starting at element `first`, into dst.  false + a pending
// Java exception when the JVM cannot give access to the array.
template <typename T>
bool copy_from_array(JNIEnvPtr env, jarray arr, size_t first, size_t count, T *dst) {
    if (count == 0) return true;
    auto get = slot<void *(*)(JNIEnvPtr, jarray, jboolean *)>(env, kJniGetPrimitiveArrayCritical);
    auto rel = slot<void (*)(JNIEnvPtr, jarray, void *, jint)>(env, kJniReleasePrimitiveArrayCritical);
    T *src = arr ? static_cast<T *>(get(env, arr, nullptr)) : nullptr;
    if (!src) {
        throw_runtime(env, "cannot access the Java array (null array or GetPrimitiveArrayCritical failed)");
        return false;
    }
    memcpy(dst, src + first, count * sizeof(T));
    rel(env, arr, src, JNI_ABORT);
    return true;
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------- OneDAL$
JNIEXPORT jboolean JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cCheckPlatformCompatibility(JNIEnvPtr, jobject) {
    return kmb200_device_ok() ? 1 : 0;
}

JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable(JNIEnvPtr env, jobject, jlong rows,
                                                                             jint cols) {
    kmb200_table *t = nullptr;
    if (!use_device(env)) return 0;   // conversion tasks run on threads of their own: each selects the executor's GPU
    if (kmb200_table_create(rows, cols, &t) != 0) {
        throw_runtime(env, kmb200_last_error());
        return 0;
    }
    return reinterpret_cast<jlong>(t);
}

// Same shape as the reference's native (native/OneDAL.cpp:50-60) but meant to be called with row
// BATCHES (numRows >> 1), not once per row.
JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cSetDoubleBatch(JNIEnvPtr env, jobject,
                                                                                  jlong numTableAddr, jint curRows,
                                                                                  jdoubleArray batch, jint numRows,
                                                                                  jint numCols) {
    kmb200_table *t = reinterpret_cast<kmb200_table *>(numTableAddr);
    if (numRows <= 0 || numCols <= 0) return;
    const size_t row_bytes = (size_t)numCols * sizeof(double);
    const size_t rows_per = std::max<size_t>(1, kStagingBytes / row_bytes);
    double *stage = static_cast<double *>(g_staging.get(std::min((size_t)numRows, rows_per) * row_bytes));
    if (!stage) {
        throw_runtime(env, kmb200_last_error());
        return;
    }
    for (size_t r = 0; r < (size_t)numRows; r += rows_per) {
        const size_t h = std::min(rows_per, (size_t)numRows - r);
        if (!copy_from_array<double>(env, batch, r * numCols, h * numCols, stage)) return;
        if (kmb200_table_append_f64(t, (int64_t)curRows + (int64_t)r, stage, (int64_t)h, numCols) != 0) {
            throw_runtime(env, kmb200_last_error());
            return;
        }
    }
}

// ---- other input layouts (no counterpart in the reference, which densifies every row on the JVM with
// Vector.toArray, OneDAL.scala:135-139): see include/kmb200.h, "other input layouts"
// columns: double[numCols][numRows]
JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cSetColumnsBatch(JNIEnvPtr env, jobject,
                                                                                   jlong numTableAddr, jint curRows,
                                                                                   jobjectArray columns, jint numRows,
                                                                                   jint numCols) {
    auto elem = slot<jobject (*)(JNIEnvPtr, jobjectArray, jsize)>(env, kJniGetObjectArrayElement);
    kmb200_table *t = reinterpret_cast<kmb200_table *>(numTableAddr);
    if (numRows <= 0 || numCols <= 0) return;
    // chunks of rows: every column's slice is copied (one short critical region per column) into a column-major
    // block of the staging buffer, then the block is uploaded
    const size_t rows_per = std::max<size_t>(1, kStagingBytes / ((size_t)numCols * sizeof(double)));
    const size_t chunk = std::min((size_t)numRows, rows_per);
    double *stage = static_cast<double *>(g_staging.get(chunk * numCols * sizeof(double)));
    if (!stage) {
        throw_runtime(env, kmb200_last_error());
        return;
    }
    std::vector<jarray> arrays((size_t)numCols);
    std::vector<const double *> ptrs((size_t)numCols);
    for (int c = 0; c < numCols; ++c) {
        arrays[c] = columns ? elem(env, columns, c) : nullptr;
        ptrs[c] = stage + (size_t)c * chunk;
    }
    for (size_t r = 0; r < (size_t)numRows; r += chunk) {
        const size_t h = std::min(chunk, (size_t)numRows - r);
        for (int c = 0; c < numCols; ++c)
            if (!copy_from_array<double>(env, arrays[c], r, h, stage + (size_t)c * chunk)) return;
        if (kmb200_table_append_columns_f64(t, (int64_t)curRows + (int64_t)r, ptrs.data(), (int64_t)h, numCols) != 0) {
            throw_runtime(env, kmb200_last_error());
            return;
        }
    }
}

// sparse rows: indptr long[numRows + 1], indices int[nnz], values double[nnz]
JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cSetSparseBatch(JNIEnvPtr env, jobject,
                                                                                  jlong numTableAddr, jint curRows,
                                                                                  jlongArray indptr, jintArray indices,
                                                                                  jdoubleArray values, jint numRows) {
    if (numRows <= 0) return;
    // the three arrays are copied out one after the other (short critical regions), then scattered on the device
    std::vector<int64_t> ip((size_t)numRows + 1);
    if (!copy_from_array<int64_t>(env, indptr, 0, ip.size(), ip.data())) return;
    const int64_t base = ip[0], nnz = ip[numRows] - base;
    if (nnz < 0) {
        throw_runtime(env, "cSetSparseBatch: indptr decreases");
        return;
    }
    void *stage = g_staging.get(std::max<size_t>((size_t)nnz * 12, 16));
    if (!stage) {
        throw_runtime(env, kmb200_last_error());
        return;
    }
    double *va = static_cast<double *>(stage);
    int32_t *ix = reinterpret_cast<int32_t *>(va + nnz);
    if (nnz > 0 && (!copy_from_array<double>(env, values, (size_t)base, (size_t)nnz, va) ||
                    !copy_from_array<int32_t>(env, indices, (size_t)base, (size_t)nnz, ix)))
        return;
    for (auto &v : ip) v -= base;   // the staged copies start at the first nonzero of this batch
    if (kmb200_table_append_csr_f64(reinterpret_cast<kmb200_table *>(numTableAddr), curRows, numRows, ip.data(), ix, va) != 0)
        throw_runtime(env, kmb200_last_error());
}

JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cTruncateTable(JNIEnvPtr env, jobject, jlong addr,
                                                                                 jlong rows) {
    if (kmb200_table_truncate(reinterpret_cast<kmb200_table *>(addr), rows) != 0) throw_runtime(env, kmb200_last_error());
}

JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cTableRows(JNIEnvPtr, jobject, jlong addr) {
    return kmb200_table_rows(reinterpret_cast<kmb200_table *>(addr));
}

JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cFreeDataMemory(JNIEnvPtr, jobject, jlong addr) {
    kmb200_table_free(reinterpret_cast<kmb200_table *>(addr));
}

JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cNewCenters(JNIEnvPtr, jobject, jint k, jint d) {
    Centers *c = new Centers{k, d, std::vector<double>((size_t)k * d, 0.0)};
    return reinterpret_cast<jlong>(c);
}

// element store used by OneDAL.makeNumericTable(Array[OldVector]) (OneDAL.scala:62-79)
JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_setNumericTableValue(JNIEnvPtr, jobject, jlong addr,
                                                                                       jint row, jint column,
                                                                                       jdouble value) {
    Centers *c = reinterpret_cast<Centers *>(addr);
    if (c && row >= 0 && row < c->k && column >= 0 && column < c->d) c->v[(size_t)row * c->d + column] = value;
}

// one call instead of k*d getDoubleValue calls (OneDAL.scala:37-52)
JNIEXPORT jdoubleArray JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cGetCenters(JNIEnvPtr env, jobject, jlong addr) {
    Centers *c = reinterpret_cast<Centers *>(addr);
    auto mk = slot<jdoubleArray (*)(JNIEnvPtr, jsize)>(env, kJniNewDoubleArray);
    auto set = slot<void (*)(JNIEnvPtr, jdoubleArray, jsize, jsize, const jdouble *)>(env, kJniSetDoubleArrayRegion);
    jdoubleArray arr = mk(env, (jsize)c->v.size());
    if (arr) set(env, arr, 0, (jsize)c->v.size(), c->v.data());
    return arr;
}

JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cFreeCenters(JNIEnvPtr, jobject, jlong addr) {
    delete reinterpret_cast<Centers *>(addr);
}

// ---------------------------------------------------------------- OneCCL$
JNIEXPORT jbyteArray JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_c_1getUniqueId(JNIEnvPtr env, jobject) {
    uint8_t id[KMB200_UNIQUE_ID_BYTES];
    if (kmb200_comm_unique_id(id) != 0) {
        throw_runtime(env, kmb200_last_error());
        return nullptr;
    }
    auto mk = slot<jbyteArray (*)(JNIEnvPtr, jsize)>(env, kJniNewByteArray);
    auto set = slot<void (*)(JNIEnvPtr, jbyteArray, jsize, jsize, const jbyte *)>(env, kJniSetByteArrayRegion);
    jbyteArray arr = mk(env, KMB200_UNIQUE_ID_BYTES);
    if (arr) set(env, arr, 0, KMB200_UNIQUE_ID_BYTES, reinterpret_cast<const jbyte *>(id));
    return arr;
}

// replaces c_init(size, rank, "ip_port", param) (native/OneCCL.cpp:47-86): the unique id arrives as
// bytes; commSize / rankId are written back into the CCLParam object like the reference does.
JNIEXPORT jint JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_c_1init(JNIEnvPtr env, jobject, jint size, jint rank,
                                                                          jbyteArray id, jobject param) {
    uint8_t buf[KMB200_UNIQUE_ID_BYTES];
    auto getr = slot<void (*)(JNIEnvPtr, jbyteArray, jsize, jsize, jbyte *)>(env, kJniGetByteArrayRegion);
    getr(env, id, 0, KMB200_UNIQUE_ID_BYTES, reinterpret_cast<jbyte *>(buf));
    if (!use_device(env)) return 0;
    if (size > 1) {
        if (kmb200_comm_init(size, rank, buf, &g_comm) != 0) {
            throw_runtime(env, kmb200_last_error());
            return 0;
        }
    }
    g_world = size;
    g_rank = rank;
    auto cls_of = slot<jclass (*)(JNIEnvPtr, jobject)>(env, kJniGetObjectClass);
    auto fid = slot<jfieldID (*)(JNIEnvPtr, jclass, const char *, const char *)>(env, kJniGetFieldID);
    auto setl = slot<void (*)(JNIEnvPtr, jobject, jfieldID, jlong)>(env, kJniSetLongField);
    jclass cls = cls_of(env, param);
    setl(env, param, fid(env, cls, "commSize", "J"), (jlong)g_world);
    setl(env, param, fid(env, cls, "rankId", "J"), (jlong)g_rank);
    return 1;
}

// Device ordinal of this executor (no counterpart in the reference, which has no devices).  Call it before the
// first table is created; returns the ordinal in use (an earlier choice is kept once tables may exist on it).
JNIEXPORT jint JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_c_1setDevice(JNIEnvPtr env, jobject, jint ordinal) {
    {
        std::lock_guard<std::mutex> lk(g_device_mu);
        if (g_device < 0 && ordinal >= 0 && ordinal < kmb200_device_count()) g_device = ordinal;
    }
    if (ordinal < 0 || ordinal >= kmb200_device_count()) {
        throw_runtime(env, "c_setDevice: no such CUDA device");
        return -1;
    }
    return resolve_device();
}

JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_c_1cleanup(JNIEnvPtr, jobject) {
    kmb200_comm_destroy(g_comm);
    g_comm = nullptr;
    g_world = 1;
    g_rank = 0;
}

JNIEXPORT jboolean JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_isRoot(JNIEnvPtr, jobject) { return g_rank == 0; }
JNIEXPORT jint JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_rankID(JNIEnvPtr, jobject) { return g_rank; }

JNIEXPORT jint JNICALL Java_org_apache_spark_ml_util_OneCCL_00024_setEnv(JNIEnvPtr env, jobject, jstring key,
                                                                         jstring value, jboolean overwrite) {
    auto getc = slot<const char *(*)(JNIEnvPtr, jstring, jboolean *)>(env, kJniGetStringUTFChars);
    auto relc = slot<void (*)(JNIEnvPtr, jstring, const char *)>(env, kJniReleaseStringUTFChars);
    const char *k = getc(env, key, nullptr);
    const char *v = getc(env, value, nullptr);
    const int err = setenv(k, v, overwrite);
    relc(env, key, k);
    relc(env, value, v);
    return err;
}

// ---------------------------------------------------------------- KMeansDALImpl
// Same name and signature as the reference (native/KMeansDALImpl.cpp:175-179).  pNumTabData is the
// kmb200_table handle, pNumTabCenters a Centers handle; returns a NEW Centers handle on rank 0 and
// 0 elsewhere, fills result.iterationNum / result.totalCost on rank 0 (reference :225-248).
// cluster_num follows the centres table; executor_cores is unused on the GPU.
JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALComputeWithInitCenters(
    JNIEnvPtr env, jobject, jlong pNumTabData, jlong pNumTabCenters, jint cluster_num, jdouble tolerance,
    jint iteration_num, jint executor_num, jint executor_cores, jobject resultObj) {
    (void)cluster_num;
    (void)executor_num;
    (void)executor_cores;
    kmb200_table *table = reinterpret_cast<kmb200_table *>(pNumTabData);
    Centers *init = reinterpret_cast<Centers *>(pNumTabCenters);
    if (!table || !init) {
        throw_runtime(env, "null table or centres handle");
        return 0;
    }
    Centers *out = new Centers{init->k, init->d, std::vector<double>(init->v.size())};
    double cost = 0.0;
    int32_t iters = 0;
    const int rc = kmb200_fit(table, g_comm, init->v.data(), init->k, init->d, tolerance, iteration_num, out->v.data(),
                              &cost, &iters);
    if (rc != 0) {
        delete out;
        throw_runtime(env, kmb200_last_error());
        return 0;
    }
    if (g_rank != 0) {
        delete out;
        return 0;
    }
    auto cls_of = slot<jclass (*)(JNIEnvPtr, jobject)>(env, kJniGetObjectClass);
    auto fid = slot<jfieldID (*)(JNIEnvPtr, jclass, const char *, const char *)>(env, kJniGetFieldID);
    auto seti = slot<void (*)(JNIEnvPtr, jobject, jfieldID, jint)>(env, kJniSetIntField);
    auto setd = slot<void (*)(JNIEnvPtr, jobject, jfieldID, jdouble)>(env, kJniSetDoubleField);
    jclass clazz = cls_of(env, resultObj);
    seti(env, resultObj, fid(env, clazz, "iterationNum", "I"), iters);
    setd(env, resultObj, fid(env, clazz, "totalCost", "D"), cost);
    return reinterpret_cast<jlong>(out);
}

// The summary pass of KMeans.fit (spark-3.1.1/ml/clustering/KMeans.scala:359-368: model.transform + clusterSizes, one
// more JVM pass over every row in the reference) on the GPU, over the table that is still resident after the fit.
// predictions: int[local rows] or null; clusterSizes: long[k] (global sizes on every rank); returns the global cost
// of the assignment.  Collective like the compute call.
JNIEXPORT jdouble JNICALL Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALAssign(
    JNIEnvPtr env, jobject, jlong pNumTabData, jlong pNumTabCenters, jintArray predictions, jlongArray clusterSizes) {
    kmb200_table *table = reinterpret_cast<kmb200_table *>(pNumTabData);
    Centers *cen = reinterpret_cast<Centers *>(pNumTabCenters);
    if (!table || !cen) {
        throw_runtime(env, "null table or centres handle");
        return 0.0;
    }
    const int64_t n = kmb200_table_rows(table);
    std::vector<int32_t> labels(predictions ? (size_t)n : 0);
    std::vector<int64_t> sizes((size_t)cen->k);
    double cost = 0.0;
    if (kmb200_assign(table, g_comm, cen->v.data(), cen->k, cen->d, predictions ? labels.data() : nullptr, sizes.data(),
                      &cost) != 0) {
        throw_runtime(env, kmb200_last_error());
        return 0.0;
    }
    if (predictions) {
        auto seti = slot<void (*)(JNIEnvPtr, jintArray, jsize, jsize, const jint *)>(env, kJniSetIntArrayRegion);
        seti(env, predictions, 0, (jsize)n, labels.data());
    }
    if (clusterSizes) {
        auto setl = slot<void (*)(JNIEnvPtr, jlongArray, jsize, jsize, const jlong *)>(env, kJniSetLongArrayRegion);
        setl(env, clusterSizes, 0, (jsize)cen->k, reinterpret_cast<const jlong *>(sizes.data()));
    }
    return cost;
}

// ---- initialisation passes (no counterpart in the reference's natives: its initialisation
// runs as Spark jobs on the JVM, mllib/clustering/KMeans.scala:370-432).  The Scala loop of initKMeansParallel stays;
// its three row passes call these instead of `data.zip(costs).map`, `mapPartitionsWithIndex { filter }` and `countByValue`.
JNIEXPORT jdouble JNICALL Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansInitCosts(
    JNIEnvPtr env, jobject, jlong pNumTabData, jlong pNewCenters, jboolean firstStep) {
    kmb200_table *table = reinterpret_cast<kmb200_table *>(pNumTabData);
    Centers *cen = reinterpret_cast<Centers *>(pNewCenters);
    if (!table || !cen) {
        throw_runtime(env, "null table or centres handle");
        return 0.0;
    }
    double sum = 0.0;
    if (kmb200_init_costs(table, g_comm, cen->v.data(), cen->k, cen->d, firstStep ? 1 : 0, &sum) != 0) throw_runtime(env, kmb200_last_error());
    return sum;
}

// partRows: long[numPartitions + 1] row offsets of this executor's Spark partitions inside the table; returns the kept
// rows (table row numbers, ascending)
JNIEXPORT jlongArray JNICALL Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansInitSample(
    JNIEnvPtr env, jobject, jlong pNumTabData, jint seed, jint step, jint k, jdouble sumCosts, jlongArray partRows,
    jint numPartitions, jint partIndex0) {
    kmb200_table *table = reinterpret_cast<kmb200_table *>(pNumTabData);
    if (!table || numPartitions < 1) {
        throw_runtime(env, "cKMeansInitSample: bad argument");
        return nullptr;
    }
    std::vector<int64_t> parts((size_t)numPartitions + 1);
    if (!copy_from_array<int64_t>(env, partRows, 0, parts.size(), parts.data())) return nullptr;
    std::vector<int64_t> rows((size_t)std::max(16 * k, 1024));
    int64_t count = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (kmb200_init_sample(table, seed, step, k, sumCosts, parts.data(), numPartitions, partIndex0, rows.data(),
                               (int64_t)rows.size(), &count) != 0) {
            throw_runtime(env, kmb200_last_error());
            return nullptr;
        }
        if (count <= (int64_t)rows.size()) break;
        rows.resize((size_t)count);
    }
    auto mk = slot<jlongArray (*)(JNIEnvPtr, jsize)>(env, kJniNewLongArray);
    auto setl = slot<void (*)(JNIEnvPtr, jlongArray, jsize, jsize, const jlong *)>(env, kJniSetLongArrayRegion);
    jlongArray arr = mk(env, (jsize)count);
    if (arr && count > 0) setl(env, arr, 0, (jsize)count, reinterpret_cast<const jlong *>(rows.data()));
    return arr;
}

// ---------------------------------------------------------------- PCADALImpl
// Same name and signature as the reference (native/PCADALImpl.cpp:38-42; javah/org_apache_spark_ml_feature_PCADALImpl.h).
// pNumTabData is the kmb200_table handle.  On rank 0, result.pcNumericTable / result.explainedVarianceNumericTable
// receive handles of a d x d table (one principal axis per row) and a 1 x d table (eigenvalues), read with cGetCenters
// where the reference reads HomogenNumericTable.getDoubleArray (PCADALImpl.scala:108-136) and freed with cFreeCenters.
// Like the reference's native code it takes the rows as they are (PCADALImpl.scala:37 has centred them on the JVM);
// cPCATrainDALCentered subtracts the column means on the device instead, so that `normalizeData` can go.
static jlong pca_train_jni(JNIEnvPtr env, jlong pNumTabData, jobject resultObj, bool center) {
    kmb200_table *table = reinterpret_cast<kmb200_table *>(pNumTabData);
    if (!table || !resultObj) {
        throw_runtime(env, "null table handle or result object");
        return 0;
    }
    const int d = kmb200_table_cols(table);
    Centers *axes = new Centers{d, d, std::vector<double>((size_t)d * d)};
    Centers *ev = new Centers{1, d, std::vector<double>((size_t)d)};
    if (kmb200_pca_train(table, g_comm, center ? 1 : 0, ev->v.data(), axes->v.data()) != 0) {
        delete axes;
        delete ev;
        throw_runtime(env, kmb200_last_error());
        return 0;
    }
    if (g_rank != 0) {
        delete axes;
        delete ev;
        return 0;
    }
    auto cls_of = slot<jclass (*)(JNIEnvPtr, jobject)>(env, kJniGetObjectClass);
    auto fid = slot<jfieldID (*)(JNIEnvPtr, jclass, const char *, const char *)>(env, kJniGetFieldID);
    auto setl = slot<void (*)(JNIEnvPtr, jobject, jfieldID, jlong)>(env, kJniSetLongField);
    jclass clazz = cls_of(env, resultObj);
    setl(env, resultObj, fid(env, clazz, "pcNumericTable", "J"), reinterpret_cast<jlong>(axes));
    setl(env, resultObj, fid(env, clazz, "explainedVarianceNumericTable", "J"), reinterpret_cast<jlong>(ev));
    return 0;
}
JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDAL(JNIEnvPtr env, jobject, jlong pNumTabData, jint k,
                                                                                 jint executor_num, jint executor_cores,
                                                                                 jobject resultObj) {
    (void)k;
    (void)executor_num;
    (void)executor_cores;
    return pca_train_jni(env, pNumTabData, resultObj, false);
}
JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDALCentered(JNIEnvPtr env, jobject, jlong pNumTabData,
                                                                                         jint k, jint executor_num,
                                                                                         jint executor_cores, jobject resultObj) {
    (void)k;
    (void)executor_num;
    (void)executor_cores;
    return pca_train_jni(env, pNumTabData, resultObj, true);
}

}  // extern "C"
