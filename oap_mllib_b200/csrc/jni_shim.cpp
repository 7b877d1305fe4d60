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
//     OneCCL$.c_getUniqueId ()[B, c_init (II[BLorg/apache/spark/ml/util/CCLParam;)I
//
// No JVM exists in this image: the layer is exercised through a hand-built JNIEnv function table
// (tests/test_jni_shim.py).  It cannot be linked against a real JVM here -- INTEGRATION.md says so.
#include <cstdlib>
#include <cstring>
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

}  // namespace

extern "C" {

// ---------------------------------------------------------------- OneDAL$
JNIEXPORT jboolean JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cCheckPlatformCompatibility(JNIEnvPtr, jobject) {
    return kmb200_device_ok() ? 1 : 0;
}

JNIEXPORT jlong JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable(JNIEnvPtr env, jobject, jlong rows,
                                                                             jint cols) {
    kmb200_table *t = nullptr;
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
    auto get = slot<void *(*)(JNIEnvPtr, jarray, jboolean *)>(env, kJniGetPrimitiveArrayCritical);
    auto rel = slot<void (*)(JNIEnvPtr, jarray, void *, jint)>(env, kJniReleasePrimitiveArrayCritical);
    double *values = static_cast<double *>(get(env, batch, nullptr));
    const int rc = kmb200_table_append_f64(reinterpret_cast<kmb200_table *>(numTableAddr), curRows, values, numRows, numCols);
    rel(env, batch, values, JNI_ABORT);
    if (rc != 0) throw_runtime(env, kmb200_last_error());
}

// ---- other input layouts (no counterpart in the reference, which densifies every row on the JVM with
// Vector.toArray, OneDAL.scala:135-139): see include/kmb200.h, "other input layouts"
// columns: double[numCols][numRows]
JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cSetColumnsBatch(JNIEnvPtr env, jobject,
                                                                                   jlong numTableAddr, jint curRows,
                                                                                   jobjectArray columns, jint numRows,
                                                                                   jint numCols) {
    auto elem = slot<jobject (*)(JNIEnvPtr, jobjectArray, jsize)>(env, kJniGetObjectArrayElement);
    auto get = slot<void *(*)(JNIEnvPtr, jarray, jboolean *)>(env, kJniGetPrimitiveArrayCritical);
    auto rel = slot<void (*)(JNIEnvPtr, jarray, void *, jint)>(env, kJniReleasePrimitiveArrayCritical);
    std::vector<jarray> arrays((size_t)numCols);
    std::vector<const double *> ptrs((size_t)numCols);
    for (int c = 0; c < numCols; ++c) arrays[c] = elem(env, columns, c);
    for (int c = 0; c < numCols; ++c) ptrs[c] = static_cast<const double *>(get(env, arrays[c], nullptr));
    const int rc = kmb200_table_append_columns_f64(reinterpret_cast<kmb200_table *>(numTableAddr), curRows, ptrs.data(),
                                                   numRows, numCols);
    for (int c = numCols - 1; c >= 0; --c) rel(env, arrays[c], const_cast<double *>(ptrs[c]), JNI_ABORT);
    if (rc != 0) throw_runtime(env, kmb200_last_error());
}

// sparse rows: indptr long[numRows + 1], indices int[nnz], values double[nnz]
JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cSetSparseBatch(JNIEnvPtr env, jobject,
                                                                                  jlong numTableAddr, jint curRows,
                                                                                  jlongArray indptr, jintArray indices,
                                                                                  jdoubleArray values, jint numRows) {
    auto get = slot<void *(*)(JNIEnvPtr, jarray, jboolean *)>(env, kJniGetPrimitiveArrayCritical);
    auto rel = slot<void (*)(JNIEnvPtr, jarray, void *, jint)>(env, kJniReleasePrimitiveArrayCritical);
    int64_t *ip = static_cast<int64_t *>(get(env, indptr, nullptr));
    int32_t *ix = static_cast<int32_t *>(get(env, indices, nullptr));
    double *va = static_cast<double *>(get(env, values, nullptr));
    const int rc = kmb200_table_append_csr_f64(reinterpret_cast<kmb200_table *>(numTableAddr), curRows, numRows, ip, ix, va);
    rel(env, values, va, JNI_ABORT);
    rel(env, indices, ix, JNI_ABORT);
    rel(env, indptr, ip, JNI_ABORT);
    if (rc != 0) throw_runtime(env, kmb200_last_error());
}

JNIEXPORT void JNICALL Java_org_apache_spark_ml_util_OneDAL_00024_cTruncateTable(JNIEnvPtr env, jobject, jlong addr,
                                                                                 jlong rows) {
    if (kmb200_table_truncate(reinterpret_cast<kmb200_table *>(addr), rows) != 0) throw_runtime(env, kmb200_last_error());
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
    if (kmb200_set_device(0) != 0 && kmb200_device_count() > 0) { /* one executor = one visible GPU */ }
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

}  // extern "C"
