// libkmeans_b200.so -- host side of the C ABI declared in include/kmb200.h.
// Table ingest (host -> HBM), NCCL rendezvous, and the Lloyd driver that replaces the loop in the
// reference's mllib-dal/src/main/native/KMeansDALImpl.cpp:175-249.  There is no CPU fallback: every
// compute entry point fails with KMB200_ERR_DEVICE when no sm_100 GPU is usable.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>  // types only; the library is dlopen'ed so single-GPU use needs no NCCL at all

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/kmb200.h"
#include "assign_tc.cuh"
#include "assign_tc2.cuh"
#include "assign_pair_ss.cuh"
#include "assign_pair_h16.cuh"
#include "fused_smallk.cuh"
#include "gram_pair.cuh"
#include "kernels_misc.cuh"

using namespace kmb;

// ------------------------------------------------------------------------------------ errors
static thread_local std::string g_err;
static thread_local kmb200_profile g_prof;

static int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CU_TRY(expr)                                                                                   \
    do {                                                                                               \
        cudaError_t e__ = (expr);                                                                      \
        if (e__ != cudaSuccess)                                                                        \
            return fail(e__ == cudaErrorMemoryAllocation ? KMB200_ERR_MEMORY : KMB200_ERR_DEVICE,      \
                        "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__);  \
    } while (0)
#define KM_TRY(expr)            \
    do {                        \
        int rc__ = (expr);      \
        if (rc__ != 0) return rc__; \
    } while (0)

// After every kernel launch.  KMB200_SYNC_CHECK=1 in the environment (diagnostics) also waits for the kernel and
// names the launch site when it faulted -- asynchronous faults otherwise surface at some later, unrelated call.
static int g_sync_check = -1;
static int post_launch(int line) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) {
        if (g_sync_check < 0) {
            const char *v = getenv("KMB200_SYNC_CHECK");
            g_sync_check = v && atoi(v) != 0;
        }
        if (g_sync_check) e = cudaDeviceSynchronize();
    }
    if (e != cudaSuccess)
        return fail(e == cudaErrorMemoryAllocation ? KMB200_ERR_MEMORY : KMB200_ERR_DEVICE,
                    "kernel launched at kmb200.cu:%d failed: %s", line, cudaGetErrorString(e));
    return 0;
}

// ------------------------------------------------------------------------------------ options
static double g_eps_scale = 1.0;
static int g_passes = 3;
static int g_force_simt = 0;
static int g_knockout = 0;
static int g_kernel = 2;   // 2 = X-resident / multicast kernel where it applies (ld <= 256), 1 = first-generation kernel
static int g_fused = 1;    // 1 = one-pass CUDA-core kernel when k*ld is small (HBM-bound regime)
static int g_h16 = 1;      // 1 = FP16-split pair kernel (assign_pair_h16.cuh) where it applies, 0 = TF32 + BF16 corrections
static int g_two_level = 1;  // FP16-split kernel: 1 = fast bookkeeping over all rows + full bookkeeping over the near-tie rows
                             //   (tables of >= kTwoLevelMinRows rows), 2 = always, 0 = never
static int64_t g_cand_cap = 0;   // candidate-buffer rows (0 = rows / 16); tests shrink it to force the overflow path
static size_t g_stage_bytes = (size_t)256 << 20;   // device staging buffer of the columnar / sparse / FP64 uploads ("stage_mb")
static int g_dbg_products = 3;   // kmb200_debug_scores: 1 = dump the coarse one-product scores ("dbg_products")
static int g_cascade = 1;    // FP16-split kernel: 1 = coarse one-product pass first (needs the two-level scheme), 0 = off
static double g_cand1_frac = 0.3;   // rows of the coarse pass's copy, as a fraction of the table ("cand1_frac")
static int g_rb_cap = 4;   // experiments: cap on the B-ring group slots of the pair kernel
static int g_cluster = 2;  // 2 = CTA pairs (tcgen05 cta_group::2, centroid tiles split across the pair), 1 = single CTAs
static void env_overrides() {
    static bool done = false;
    if (done) return;
    done = true;
    if (const char *e = getenv("KMB200_KERNEL")) g_kernel = atoi(e) == 1 ? 1 : 2;
    if (const char *e = getenv("KMB200_CLUSTER")) g_cluster = atoi(e) == 1 ? 1 : 2;
    if (const char *e = getenv("KMB200_H16")) g_h16 = atoi(e) != 0;
}
static long long *g_role_prof = nullptr;  // device buffer, allocated by option "role_prof"

// ------------------------------------------------------------------------------------ NCCL (dlopen)
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static int nccl_load() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.handle) return 0;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *nm : names) {
        h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return fail(KMB200_ERR_COMM, "cannot dlopen libnccl.so.2: %s", dlerror());
#define KMB_SYM(field, name)                                                        \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));        \
    if (!g_nccl.field) return fail(KMB200_ERR_COMM, "NCCL symbol %s missing", name);
    KMB_SYM(GetUniqueId, "ncclGetUniqueId")
    KMB_SYM(CommInitRank, "ncclCommInitRank")
    KMB_SYM(CommDestroy, "ncclCommDestroy")
    KMB_SYM(CommAbort, "ncclCommAbort")
    KMB_SYM(AllReduce, "ncclAllReduce")
    KMB_SYM(AllGather, "ncclAllGather")
    KMB_SYM(GetErrorString, "ncclGetErrorString")
#undef KMB_SYM
    g_nccl.handle = h;
    return 0;
}
#define NCCL_TRY(expr)                                                                          \
    do {                                                                                        \
        ncclResult_t r__ = (expr);                                                              \
        if (r__ != ncclSuccess)                                                                 \
            return fail(KMB200_ERR_COMM, "%s failed: %s", #expr, g_nccl.GetErrorString(r__));   \
    } while (0)

struct kmb200_comm {
    ncclComm_t comm = nullptr;
    int world = 1, rank = 0;
    bool broken = false;   // a collective or a kernel failed on it: destroy with ncclCommAbort (peers may never arrive)
};

// ------------------------------------------------------------------------------------ table
struct Workspace {
    int k = 0, kpad = 0, bn = 0;
    double *cent = nullptr;       // [kpad x ld] FP64 master copy of the centres (padding zero)
    double *cent2 = nullptr;      // [kpad x ld] output of finalize; swapped with cent every iteration
    float *bhi = nullptr, *blo = nullptr, *cnorm = nullptr, *cstats = nullptr;
    double *cnorm64 = nullptr;        // [kpad + 4] FP64 ||c||^2, +inf beyond k (full re-check)
    double *rc_score = nullptr;       // partial winners of a full re-check split over several CTAs
    int32_t *rc_idx = nullptr;
    __half *b16 = nullptr;            // [kpad x 2*ld64] FP16 split of -2c*sw: g1 | g2 (split16_kernel)
    __nv_bfloat16 *bcorr = nullptr;   // [kpad x 2*ld] BF16 correction operand (see split_centers_kernel)
    int32_t *labels = nullptr, *order = nullptr, *flag_count = nullptr;
    int2 *flag_list = nullptr;    // near-tie rows: (row, runner-up centre or -1)
    int32_t *counts = nullptr, *offsets = nullptr, *cursor = nullptr, *chunk_off = nullptr;
    int32_t *hb = nullptr, *segsum = nullptr;   // stable bucketing: per-block label histograms [blocks x k], segment sums
    double *part = nullptr, *cpart = nullptr, *red = nullptr, *costj = nullptr;
    int32_t *flags = nullptr;      // [0] centres that moved > tol, [1] empty clusters
    double *fused_part = nullptr;  // per-CTA partials of the one-pass small-k kernel
    float *fused_stage = nullptr;  // FP32 centres (1024) + norms (32) on their way to constant memory
    int fused_ctas = 0;
    int64_t max_chunks = 0;
    CUtensorMap tm_x, tm_bhi, tm_blo;
    CUtensorMap tm_b16_half;               // FP16 [kpad x 2*ld64], box 64 values (128 bytes) x bn/2 rows
    CUtensorMap tm_bcorr_half;             // BF16 [kpad x 2*ld], box 64 values (128 bytes) x bn/2 rows
    CUtensorMap tm_bhi_half, tm_blo_half;  // boxes of bn/2 rows: each CTA of a pair loads half a centroid tile
    // two-level bookkeeping of the FP16-split kernel: the near-tie rows of the fast pass, gathered
    float *xc = nullptr;              // [cand_cap x ld] copies of the candidate rows
    float *xnc = nullptr;             // [cand_cap] their ||x||^2
    int32_t *cand_list = nullptr;     // [cand_cap] their row numbers
    int32_t *labels_c = nullptr;      // [cand_cap] their labels (scattered back after the re-check)
    int64_t cand_cap = 0;
    CUtensorMap tm_xc;
    // cascade: the rows the coarse one-product pass could not decide (a larger copy, one level above xc)
    float *xc1 = nullptr;             // [cand1_cap x ld]
    float *xnc1 = nullptr;
    int32_t *cand1_list = nullptr;
    int32_t *labels_c1 = nullptr;
    int64_t cand1_cap = 0;
    CUtensorMap tm_xc1;
    bool cascade_off = false;         // the coarse pass left too many rows undecided once in this fit: not used again
    int recheck_level = 0;            // set by launch_assign: the lists index 0 = the table, 1 = xc1, 2 = xc
    bool scatter1 = false, scatter2 = false;   // labels_c -> (labels_c1 | labels) via cand_list; labels_c1 -> labels via cand1_list
    bool two_level_active = false;    // set by launch_assign for the launch_recheck that follows
    bool two_level_off = false;       // the candidate buffer overflowed once in this fit: full bookkeeping from then on
    // per-call state (one fit / assign at a time per table)
    float *dbg_scores = nullptr;      // set only inside kmb200_debug_scores
    cudaEvent_t *main_ev = nullptr;   // [2] events bracketing the main assign kernel of the current iteration (or null)
    bool ready = false;               // every buffer and tensor map below exists (set last by ws_prepare)
    // pinned host mirror read once per iteration
    struct Host {
        int32_t flags[2];
        int32_t flag_count;
        int32_t full_count;
        int32_t cand_count;
        int32_t cand1_count;
        double cost;
    } *host = nullptr;
};

// CUDA events of one table (created on the table's device), reused by every fit
struct EventPool {
    std::vector<cudaEvent_t> ev;
    int per_iter = 9;   // 7 phase boundaries + 2 around the main assign kernel
    int cap_iters = 0;
    ~EventPool() {
        for (auto e : ev) cudaEventDestroy(e);
    }
    int ensure(int iters) {
        iters = std::min(iters, 1024);
        while ((int)ev.size() < iters * per_iter) {
            cudaEvent_t e;
            CU_TRY(cudaEventCreate(&e));
            ev.push_back(e);
        }
        cap_iters = iters;
        return 0;
    }
};

// Buffers of the PCA path (pca_host.inl), allocated on a table's first PCA call and kept with it: allocation and
// release per call cost more than the kernels on a table of a few million rows.
struct PcaWs {
    int pairs = 0;
    double *gsum = nullptr, *inv = nullptr, *mean_used = nullptr, *partial = nullptr, *G = nullptr;
    float *gmax = nullptr, *scale = nullptr, *offs = nullptr;
    long long *prof = nullptr;
    double *host = nullptr;   // pinned: [ld * ld] Gram matrix, [ld + 1] sums, [ld] means
    cudaEvent_t ev[5] = {};
    ~PcaWs() {
        void *ptrs[] = {gsum, inv, mean_used, partial, G, gmax, scale, offs, prof};
        for (void *q : ptrs)
            if (q) cudaFree(q);
        if (host) cudaFreeHost(host);
        for (auto e : ev)
            if (e) cudaEventDestroy(e);
    }
};

struct kmb200_table {
    PcaWs *pca = nullptr;
    int device = 0;
    int64_t rows = 0;       // logical rows (kmb200_table_truncate may lower it)
    int64_t capacity = 0;   // rows allocated at creation
    int d = 0, ld = 0;
    float *X = nullptr;
    float *xnorm = nullptr;
    float *xabs = nullptr;   // [1] max |x_ij| over everything ever appended (scale of the FP16 split)
    void *stage = nullptr;
    size_t stage_bytes = 0;
    // Running column statistics of rows [0, stats_rows) (PCA: sums for the centring, max |x| for the column scales),
    // kept up to date by every append that continues where the last one ended (launch_rownorm); anything else --
    // a gap, an overwrite, a truncation below stats_rows -- clears stats_ok and the PCA call recomputes them in one pass.
    double *colsum = nullptr;      // [ld + 1] column sums, then the row count
    float *colmax = nullptr;       // [ld]
    double *cs_psum = nullptr, *cs_tsum = nullptr;   // scratch: per-block partials [4 * num_sms][ld], one batch's sums [ld + 1]
    float *cs_pmax = nullptr, *cs_tmax = nullptr;
    int64_t stats_rows = 0;
    bool stats_ok = false;
    double *init_cost = nullptr;   // [rows] k-means|| initialisation: cost of every row to the nearest centre chosen so far
    double *init_tmp = nullptr;    // [rows + 1024] scratch of those passes
    cudaStream_t stream = nullptr;
    int num_sms = 0;
    Workspace ws;
    EventPool events;
};

static void ws_free(Workspace &w) {
    void *ptrs[] = {w.cent, w.cent2, w.bhi, w.blo, w.bcorr, w.b16, w.cnorm64, w.rc_score, w.rc_idx, w.cnorm, w.cstats, w.labels, w.order, w.flag_list, w.flag_count,
                    w.counts, w.offsets, w.cursor, w.chunk_off, w.hb, w.segsum, w.part, w.cpart, w.red, w.costj, w.flags, w.fused_part, w.fused_stage,
                    w.xc, w.xnc, w.cand_list, w.labels_c, w.xc1, w.xnc1, w.cand1_list, w.labels_c1};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (w.host) cudaFreeHost(w.host);
    w = Workspace();
}

// ------------------------------------------------------------------------------------ device gate
extern "C" int kmb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int kmb200_device_ok(void) {
    const int n = kmb200_device_count();
    for (int i = 0; i < n; ++i) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10)
            return 1;
    }
    return 0;
}

static int require_device() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(KMB200_ERR_DEVICE, "no usable CUDA device: %s (libkmeans_b200 has no CPU fallback)",
                    cudaGetErrorString(e));
    }
    int major = 0;
    CU_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10)
        return fail(KMB200_ERR_DEVICE, "device %d has compute capability %d.x; sm_100 (B200) is required", dev,
                    major);
    return 0;
}

extern "C" int kmb200_set_device(int32_t ordinal) {
    CU_TRY(cudaSetDevice(ordinal));
    return 0;
}

// ------------------------------------------------------------------------------------ tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
// 2-D BF16 row-major [rows x ld2] tensor, box = 64 values (128 bytes) x box_rows rows, 128-byte swizzle.
static int make_map_bf16(CUtensorMap *m, const void *base, uint64_t rows, uint64_t ld2, uint32_t box_rows, bool fp16 = false);
// 2-D FP32 row-major [rows x ld] tensor, box = 32 columns x box_rows rows, 128-byte swizzle.
static int make_map(CUtensorMap *m, const float *base, uint64_t rows, uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(KMB200_ERR_DEVICE, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {ld, rows};
    cuuint64_t strides[1] = {ld * sizeof(float)};
    cuuint32_t box[2] = {32, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(KMB200_ERR_DEVICE, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

static int make_map_bf16(CUtensorMap *m, const void *base, uint64_t rows, uint64_t ld2, uint32_t box_rows, bool fp16) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(KMB200_ERR_DEVICE, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {ld2, rows};
    cuuint64_t strides[1] = {ld2 * 2};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, fp16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(KMB200_ERR_DEVICE, "cuTensorMapEncodeTiled(bf16) failed with CUresult %d", (int)r);
    return 0;
}

// ------------------------------------------------------------------------------------ table API
extern "C" int kmb200_table_create(int64_t rows, int32_t cols, kmb200_table **out) {
    if (!out || rows <= 0 || cols <= 0) return fail(KMB200_ERR_ARG, "table_create: rows and cols must be positive");
    if (rows >= (int64_t)2147483647 - 1024)
        return fail(KMB200_ERR_LIMIT, "table_create: at most 2^31-1025 rows per GPU (row indices are int32)");
    KM_TRY(require_device());
    kmb200_table *t = new kmb200_table();
    cudaGetDevice(&t->device);
    t->rows = rows;
    t->capacity = rows;
    t->d = cols;
    t->ld = (cols + 31) / 32 * 32;
    cudaDeviceGetAttribute(&t->num_sms, cudaDevAttrMultiProcessorCount, t->device);
    cudaError_t e = cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&t->X, (size_t)rows * t->ld * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&t->xnorm, (size_t)rows * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&t->xabs, sizeof(float));
    if (e == cudaSuccess) e = cudaMemsetAsync(t->xabs, 0, sizeof(float), t->stream);
    if (e == cudaSuccess && t->ld != t->d) e = cudaMemsetAsync(t->X, 0, (size_t)rows * t->ld * sizeof(float), t->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(t->xnorm, 0, (size_t)rows * sizeof(float), t->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
    if (e != cudaSuccess) {
        const int code = e == cudaErrorMemoryAllocation ? KMB200_ERR_MEMORY : KMB200_ERR_DEVICE;
        fail(code, "table_create(%lld x %d): %s", (long long)rows, cols, cudaGetErrorString(e));
        cudaGetLastError();
        if (t->X) cudaFree(t->X);
        if (t->xnorm) cudaFree(t->xnorm);
        if (t->xabs) cudaFree(t->xabs);
        if (t->stream) cudaStreamDestroy(t->stream);
        delete t;
        return code;
    }
    *out = t;
    return 0;
}

extern "C" int kmb200_table_free(kmb200_table *t) {
    if (!t) return 0;
    cudaSetDevice(t->device);
    cudaStreamSynchronize(t->stream);
    ws_free(t->ws);
    delete t->pca;
    if (t->stage) cudaFree(t->stage);
    if (t->init_cost) cudaFree(t->init_cost);
    if (t->init_tmp) cudaFree(t->init_tmp);
    void *stats[] = {t->colsum, t->colmax, t->cs_psum, t->cs_tsum, t->cs_pmax, t->cs_tmax};
    for (void *q : stats)
        if (q) cudaFree(q);
    cudaFree(t->X);
    cudaFree(t->xnorm);
    cudaFree(t->xabs);
    cudaStreamDestroy(t->stream);
    delete t;
    return 0;
}

extern "C" int64_t kmb200_table_rows(const kmb200_table *t) { return t ? t->rows : -1; }
extern "C" int32_t kmb200_table_cols(const kmb200_table *t) { return t ? t->d : -1; }

// Column statistics of the rows just written (see kmb200_table::colsum).  row0 == 0 starts over.
static int table_stats_append(kmb200_table *t, int64_t row0, int64_t nrows) {
    if (nrows <= 0) return 0;
    const bool reset = row0 == 0;
    if (reset) {
        t->stats_rows = 0;
        t->stats_ok = true;
    }
    if (!t->stats_ok) return 0;
    const int ld = t->ld;
    // out of order -> recompute at the PCA call; so do appends of a few rows at a time (three launches per append
    // would cost more than the one pass they save)
    if (row0 != t->stats_rows || (!reset && (size_t)nrows * ld * sizeof(float) < (size_t)64 * 1024)) {
        t->stats_ok = false;
        return 0;
    }
    if (!t->colsum) {
        CU_TRY(cudaMalloc(&t->colsum, ((size_t)ld + 1) * sizeof(double)));
        CU_TRY(cudaMalloc(&t->colmax, (size_t)ld * sizeof(float)));
        CU_TRY(cudaMalloc(&t->cs_tsum, ((size_t)ld + 1) * sizeof(double)));
        CU_TRY(cudaMalloc(&t->cs_tmax, (size_t)ld * sizeof(float)));
        CU_TRY(cudaMalloc(&t->cs_psum, (size_t)t->num_sms * 4 * ld * sizeof(double)));
        CU_TRY(cudaMalloc(&t->cs_pmax, (size_t)t->num_sms * 4 * ld * sizeof(float)));
    }
    const int blocks = (int)std::min<int64_t>((int64_t)t->num_sms * 4, std::max<int64_t>(1, nrows / 64));
    colstats_kernel<<<blocks, 256, 0, t->stream>>>(t->X + (size_t)row0 * ld, nrows, ld, t->cs_psum, t->cs_pmax);
    KM_TRY(post_launch(__LINE__));
    colstats_reduce_kernel<<<(ld + 7) / 8, 256, 0, t->stream>>>(t->cs_psum, t->cs_pmax, blocks, ld, nrows, t->cs_tsum, t->cs_tmax);
    KM_TRY(post_launch(__LINE__));
    colstats_merge_kernel<<<(ld + 256) / 256, 256, 0, t->stream>>>(t->cs_tsum, t->cs_tmax, ld, reset ? 1 : 0, t->colsum, t->colmax);
    KM_TRY(post_launch(__LINE__));
    t->stats_rows = row0 + nrows;
    return 0;
}

static int launch_rownorm(kmb200_table *t, int64_t row0, int64_t nrows) {
    const int64_t warps = (nrows + 3) / 4;
    int blocks = (int)std::min<int64_t>((warps * 32 + 255) / 256, (int64_t)t->num_sms * 16);
    if (blocks < 1) blocks = 1;
    rownorm_kernel<<<blocks, 256, 0, t->stream>>>(t->X, t->xnorm, t->xabs, row0, nrows, t->ld);
    KM_TRY(post_launch(__LINE__));
    return table_stats_append(t, row0, nrows);
}

template <typename T>
static int append_rows(kmb200_table *t, int64_t row0, const T *rows, int64_t nrows, int32_t ncols) {
    if (!t || !rows) return fail(KMB200_ERR_ARG, "table_append: null argument");
    if (ncols != t->d) return fail(KMB200_ERR_ARG, "table_append: ncols %d != table cols %d", ncols, t->d);
    if (row0 < 0 || nrows < 0 || row0 + nrows > t->rows)
        return fail(KMB200_ERR_ARG, "table_append: rows [%lld, %lld) outside the table of %lld rows",
                    (long long)row0, (long long)(row0 + nrows), (long long)t->rows);
    if (nrows == 0) return 0;
    CU_TRY(cudaSetDevice(t->device));
    if (sizeof(T) == sizeof(float)) {
        // FP32 input lands directly in the table (strided copy when the row is padded)
        const int64_t max_h = 1 << 20;
        if (t->ld == t->d) {
            CU_TRY(cudaMemcpyAsync(t->X + row0 * t->ld, rows, (size_t)nrows * t->d * sizeof(float), cudaMemcpyHostToDevice,
                                   t->stream));
        } else
        for (int64_t r = 0; r < nrows; r += max_h) {
            const int64_t h = std::min(max_h, nrows - r);
            CU_TRY(cudaMemcpy2DAsync(t->X + (row0 + r) * t->ld, (size_t)t->ld * 4, rows + r * t->d, (size_t)t->d * 4,
                                     (size_t)t->d * 4, (size_t)h, cudaMemcpyHostToDevice, t->stream));
        }
    } else {
        // FP64 input: staged in device memory, rounded to FP32 by a kernel
        const size_t want = std::min<size_t>((size_t)nrows * t->d * sizeof(T), std::max<size_t>(g_stage_bytes, (size_t)t->d * sizeof(T)));
        if (t->stage_bytes < want) {
            if (t->stage) cudaFree(t->stage);
            t->stage = nullptr;
            t->stage_bytes = 0;
            CU_TRY(cudaMalloc(&t->stage, want));
            t->stage_bytes = want;
        }
        const int64_t rows_per = std::max<int64_t>(1, (int64_t)(t->stage_bytes / ((size_t)t->d * sizeof(T))));
        for (int64_t r = 0; r < nrows; r += rows_per) {
            const int64_t h = std::min(rows_per, nrows - r);
            CU_TRY(cudaMemcpyAsync(t->stage, rows + r * t->d, (size_t)h * t->d * sizeof(T), cudaMemcpyHostToDevice,
                                   t->stream));
            const int64_t total = h * t->d;
            const int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)t->num_sms * 32);
            convert_rows_kernel<T><<<blocks, 256, 0, t->stream>>>(reinterpret_cast<const T *>(t->stage),
                                                                  t->X + (row0 + r) * t->ld, h, t->d, t->ld);
            KM_TRY(post_launch(__LINE__));
        }
    }
    KM_TRY(launch_rownorm(t, row0, nrows));
    CU_TRY(cudaStreamSynchronize(t->stream));
    return 0;
}

extern "C" int kmb200_table_append_f64(kmb200_table *t, int64_t row0, const double *rows, int64_t nrows,
                                       int32_t ncols) {
    return append_rows<double>(t, row0, rows, nrows, ncols);
}
extern "C" int kmb200_table_append_f32(kmb200_table *t, int64_t row0, const float *rows, int64_t nrows,
                                       int32_t ncols) {
    return append_rows<float>(t, row0, rows, nrows, ncols);
}

static int ensure_stage(kmb200_table *t, size_t want) {
    if (t->stage_bytes >= want) return 0;
    if (t->stage) cudaFree(t->stage);
    t->stage = nullptr;
    t->stage_bytes = 0;
    CU_TRY(cudaMalloc(&t->stage, want));
    t->stage_bytes = want;
    return 0;
}

static int check_append_range(const kmb200_table *t, const void *p, int64_t row0, int64_t nrows, int32_t ncols,
                              const char *what) {
    if (!t || !p) return fail(KMB200_ERR_ARG, "%s: null argument", what);
    if (ncols != t->d) return fail(KMB200_ERR_ARG, "%s: ncols %d != table cols %d", what, ncols, t->d);
    if (row0 < 0 || nrows < 0 || row0 + nrows > t->rows)
        return fail(KMB200_ERR_ARG, "%s: rows [%lld, %lld) outside the table of %lld rows", what, (long long)row0,
                    (long long)(row0 + nrows), (long long)t->rows);
    return 0;
}

// Columnar ingest: one contiguous host buffer per column (an Arrow primitive array's data buffer, a Spark
// ColumnVector, a NumPy column).  Columns are staged in device memory in chunks of rows and transposed there.
template <typename T>
static int append_columns(kmb200_table *t, int64_t row0, const T *const *columns, int64_t nrows, int32_t ncols) {
    KM_TRY(check_append_range(t, columns, row0, nrows, ncols, "table_append_columns"));
    for (int c = 0; c < ncols; ++c)
        if (!columns[c]) return fail(KMB200_ERR_ARG, "table_append_columns: column %d is null", c);
    if (nrows == 0) return 0;
    CU_TRY(cudaSetDevice(t->device));
    const size_t row_bytes = (size_t)t->d * sizeof(T);
    const int64_t chunk = std::max<int64_t>(32, std::min<int64_t>(nrows, (int64_t)(g_stage_bytes / row_bytes)) / 32 * 32);
    KM_TRY(ensure_stage(t, (size_t)chunk * row_bytes));
    T *stage = reinterpret_cast<T *>(t->stage);
    for (int64_t r = 0; r < nrows; r += chunk) {
        const int64_t h = std::min(chunk, nrows - r);
        for (int c = 0; c < ncols; ++c)
            CU_TRY(cudaMemcpyAsync(stage + (size_t)c * chunk, columns[c] + r, (size_t)h * sizeof(T), cudaMemcpyHostToDevice,
                                   t->stream));
        const int64_t tiles = ((h + 31) / 32) * ((t->d + 31) / 32);
        const int blocks = (int)std::min<int64_t>(tiles, (int64_t)t->num_sms * 16);
        transpose_columns_kernel<T><<<blocks, dim3(32, 8), 0, t->stream>>>(stage, chunk, t->X + (row0 + r) * t->ld, h, t->d,
                                                                           t->ld);
        KM_TRY(post_launch(__LINE__));
    }
    KM_TRY(launch_rownorm(t, row0, nrows));
    CU_TRY(cudaStreamSynchronize(t->stream));
    return 0;
}
extern "C" int kmb200_table_append_columns_f64(kmb200_table *t, int64_t row0, const double *const *columns, int64_t nrows,
                                               int32_t ncols) {
    return append_columns<double>(t, row0, columns, nrows, ncols);
}
extern "C" int kmb200_table_append_columns_f32(kmb200_table *t, int64_t row0, const float *const *columns, int64_t nrows,
                                               int32_t ncols) {
    return append_columns<float>(t, row0, columns, nrows, ncols);
}

// Sparse rows in CSR form.  indptr has nrows + 1 entries (indptr[0] need not be 0), indices / values are addressed
// by indptr's values.
extern "C" int kmb200_table_append_csr_f64(kmb200_table *t, int64_t row0, int64_t nrows, const int64_t *indptr,
                                           const int32_t *indices, const double *values) {
    if (!t || !indptr) return fail(KMB200_ERR_ARG, "table_append_csr: null argument");
    KM_TRY(check_append_range(t, indptr, row0, nrows, t->d, "table_append_csr"));
    if (nrows == 0) return 0;
    for (int64_t r = 0; r < nrows; ++r)
        if (indptr[r + 1] < indptr[r]) return fail(KMB200_ERR_ARG, "table_append_csr: indptr decreases at row %lld", (long long)r);
    if (indptr[nrows] > indptr[0] && (!indices || !values)) return fail(KMB200_ERR_ARG, "table_append_csr: null indices / values");
    CU_TRY(cudaSetDevice(t->device));
    // chunks of rows whose nonzeros fit the staging buffer: [indptr (h+1) int64 | values nnz f64 | indices nnz i32]
    const int64_t max_nnz = std::max<int64_t>(1024, (int64_t)(g_stage_bytes / 32)), max_rows = std::max<int64_t>(32, (int64_t)(g_stage_bytes / 256));
    KM_TRY(ensure_stage(t, (size_t)(max_rows + 1) * 8 + (size_t)max_nnz * 12 + 64));
    uint8_t *sb = reinterpret_cast<uint8_t *>(t->stage);
    int64_t *d_indptr = reinterpret_cast<int64_t *>(sb);
    double *d_values = reinterpret_cast<double *>(sb + (size_t)(max_rows + 1) * 8);
    int32_t *d_indices = reinterpret_cast<int32_t *>(sb + (size_t)(max_rows + 1) * 8 + (size_t)max_nnz * 8);
    int32_t *d_bad = reinterpret_cast<int32_t *>(sb + (size_t)(max_rows + 1) * 8 + (size_t)max_nnz * 12);
    CU_TRY(cudaMemsetAsync(d_bad, 0, sizeof(int32_t), t->stream));
    int64_t r = 0;
    while (r < nrows) {
        int64_t h = 0;
        while (r + h < nrows && h < max_rows && indptr[r + h + 1] - indptr[r] <= max_nnz) ++h;
        if (h == 0) return fail(KMB200_ERR_LIMIT, "table_append_csr: row %lld has more than %lld nonzeros", (long long)(row0 + r), (long long)max_nnz);
        const int64_t base = indptr[r], nnz = indptr[r + h] - base;
        CU_TRY(cudaMemcpyAsync(d_indptr, indptr + r, (size_t)(h + 1) * 8, cudaMemcpyHostToDevice, t->stream));
        if (nnz > 0) {
            CU_TRY(cudaMemcpyAsync(d_values, values + base, (size_t)nnz * 8, cudaMemcpyHostToDevice, t->stream));
            CU_TRY(cudaMemcpyAsync(d_indices, indices + base, (size_t)nnz * 4, cudaMemcpyHostToDevice, t->stream));
        }
        CU_TRY(cudaMemsetAsync(t->X + (row0 + r) * t->ld, 0, (size_t)h * t->ld * sizeof(float), t->stream));
        const int blocks = (int)std::min<int64_t>((h * 32 + 255) / 256, (int64_t)t->num_sms * 16);
        csr_scatter_kernel<<<blocks, 256, 0, t->stream>>>(d_indptr, base, d_indices, d_values, t->X + (row0 + r) * t->ld, h,
                                                          t->d, t->ld, d_bad);
        KM_TRY(post_launch(__LINE__));
        CU_TRY(cudaStreamSynchronize(t->stream));   // the staging buffer is reused by the next chunk
        r += h;
    }
    int32_t bad = 0;
    CU_TRY(cudaMemcpy(&bad, d_bad, sizeof(bad), cudaMemcpyDeviceToHost));
    KM_TRY(launch_rownorm(t, row0, nrows));
    CU_TRY(cudaStreamSynchronize(t->stream));
    if (bad) return fail(KMB200_ERR_ARG, "table_append_csr: a column index is outside [0, %d)", t->d);
    return 0;
}

// Callers that do not know the row count in advance (the reference runs a whole Spark job just to count rows per
// partition, Utils.scala:28-49) create the table with an upper bound and cut it to the rows actually appended.
extern "C" int kmb200_table_truncate(kmb200_table *t, int64_t rows) {
    if (!t) return fail(KMB200_ERR_ARG, "table_truncate: null table");
    if (rows <= 0 || rows > t->capacity)
        return fail(KMB200_ERR_ARG, "table_truncate: rows must be in [1, %lld] (got %lld)", (long long)t->capacity, (long long)rows);
    if (rows == t->rows) return 0;
    CU_TRY(cudaSetDevice(t->device));
    ws_free(t->ws);   // tensor maps and per-row buffers depend on the row count
    if (t->init_cost) cudaFree(t->init_cost);
    if (t->init_tmp) cudaFree(t->init_tmp);
    t->init_cost = t->init_tmp = nullptr;
    t->rows = rows;
    if (rows != t->stats_rows) t->stats_ok = false;
    return 0;
}

// Page-locked host memory for the caller's staging buffers (a JVM wraps it in a direct ByteBuffer): uploads from
// it run at the link rate and do not pass through the driver's bounce buffers.
extern "C" int kmb200_pinned_alloc(size_t bytes, void **out) {
    if (!out || bytes == 0) return fail(KMB200_ERR_ARG, "pinned_alloc: bad argument");
    KM_TRY(require_device());
    CU_TRY(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return 0;
}
extern "C" int kmb200_pinned_free(void *p) {
    if (!p) return 0;
    CU_TRY(cudaFreeHost(p));
    return 0;
}

extern "C" int kmb200_table_read_rows(const kmb200_table *t, int64_t row0, int64_t nrows, float *out) {
    if (!t || !out) return fail(KMB200_ERR_ARG, "table_read_rows: null argument");
    if (row0 < 0 || nrows < 0 || row0 + nrows > t->rows) return fail(KMB200_ERR_ARG, "table_read_rows: range");
    if (nrows == 0) return 0;
    CU_TRY(cudaSetDevice(t->device));
    CU_TRY(cudaMemcpy2DAsync(out, (size_t)t->d * 4, t->X + row0 * t->ld, (size_t)t->ld * 4, (size_t)t->d * 4,
                             (size_t)nrows, cudaMemcpyDeviceToHost, t->stream));
    CU_TRY(cudaStreamSynchronize(t->stream));
    return 0;
}

extern "C" int kmb200_table_fill_blobs(kmb200_table *t, const float *centres, int32_t k_true, uint64_t seed,
                                       int64_t row_offset) {
    if (!t || !centres || k_true <= 0) return fail(KMB200_ERR_ARG, "table_fill_blobs: bad argument");
    CU_TRY(cudaSetDevice(t->device));
    float *dc = nullptr;
    CU_TRY(cudaMalloc(&dc, (size_t)k_true * t->d * sizeof(float)));
    cudaError_t e = cudaMemcpyAsync(dc, centres, (size_t)k_true * t->d * sizeof(float), cudaMemcpyHostToDevice, t->stream);
    if (e == cudaSuccess) {
        blobs_kernel<<<t->num_sms * 16, 256, 0, t->stream>>>(t->X, t->rows, t->d, t->ld, dc, k_true, seed, row_offset);
        e = cudaGetLastError();
    }
    int rc = 0;
    if (e == cudaSuccess) rc = launch_rownorm(t, 0, t->rows);
    if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
    cudaFree(dc);
    if (e != cudaSuccess) return fail(KMB200_ERR_DEVICE, "table_fill_blobs: %s", cudaGetErrorString(e));
    return rc;
}

// ------------------------------------------------------------------------------------ communicator
extern "C" int kmb200_comm_unique_id(uint8_t id[KMB200_UNIQUE_ID_BYTES]) {
    if (!id) return fail(KMB200_ERR_ARG, "comm_unique_id: null");
    KM_TRY(nccl_load());
    ncclUniqueId uid;
    static_assert(sizeof(ncclUniqueId) == KMB200_UNIQUE_ID_BYTES, "ncclUniqueId size");
    NCCL_TRY(g_nccl.GetUniqueId(&uid));
    memcpy(id, &uid, sizeof(uid));
    return 0;
}

extern "C" int kmb200_comm_init(int32_t world, int32_t rank, const uint8_t id[KMB200_UNIQUE_ID_BYTES],
                                kmb200_comm **out) {
    if (!out || !id || world < 1 || rank < 0 || rank >= world) return fail(KMB200_ERR_ARG, "comm_init: bad argument");
    KM_TRY(require_device());
    KM_TRY(nccl_load());
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    kmb200_comm *c = new kmb200_comm();
    c->world = world;
    c->rank = rank;
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, uid, rank);
    if (r != ncclSuccess) {
        delete c;
        return fail(KMB200_ERR_COMM, "ncclCommInitRank(world=%d, rank=%d) failed: %s", world, rank,
                    g_nccl.GetErrorString(r));
    }
    *out = c;
    return 0;
}
extern "C" int32_t kmb200_comm_rank(const kmb200_comm *c) { return c ? c->rank : 0; }
extern "C" int32_t kmb200_comm_size(const kmb200_comm *c) { return c ? c->world : 1; }
extern "C" int kmb200_comm_destroy(kmb200_comm *c) {
    if (!c) return 0;
    if (c->comm && c->broken && g_nccl.CommAbort) g_nccl.CommAbort(c->comm);
    else if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    delete c;
    return 0;
}

// ------------------------------------------------------------------------------------ workspace
static size_t ld64(int ld) { return (size_t)(ld + 63) / 64 * 64; }
// (the CTA-pair kernels need N >= 32: with the FP16-split kernel enabled, k <= 16 is padded to 32 columns rather than
// sent to the single-CTA kernel)
static int pick_bn(int k) { return k <= 16 ? (g_h16 ? 32 : 16) : k <= 32 ? 32 : k <= 64 ? 64 : 128; }

template <typename T>
static int dmalloc(T **p, size_t count) {
    CU_TRY(cudaMalloc(reinterpret_cast<void **>(p), std::max<size_t>(count, 1) * sizeof(T)));
    return 0;
}

static int ws_prepare_impl(kmb200_table *t, int k);
static int ws_prepare(kmb200_table *t, int k) {
    Workspace &w = t->ws;
    if (w.ready && w.k == k && w.bn == pick_bn(k)) return 0;
    const int rc = ws_prepare_impl(t, k);
    if (rc != 0) {   // e.g. out of memory half way: leave nothing behind that a later call could mistake for a workspace
        cudaGetLastError();
        ws_free(w);
    }
    return rc;
}
static int ws_prepare_impl(kmb200_table *t, int k) {
    Workspace &w = t->ws;
    ws_free(w);
    w.k = k;
    w.bn = pick_bn(k);
    w.kpad = (k + w.bn - 1) / w.bn * w.bn;
    const size_t ld = t->ld, n = (size_t)t->rows;
    w.max_chunks = (int64_t)((n + kSegChunk - 1) / kSegChunk) + k;
    KM_TRY(dmalloc(&w.cent, (size_t)w.kpad * ld));
    KM_TRY(dmalloc(&w.cent2, (size_t)w.kpad * ld));
    CU_TRY(cudaMemset(w.cent2, 0, (size_t)w.kpad * ld * sizeof(double)));
    KM_TRY(dmalloc(&w.bhi, (size_t)w.kpad * ld));
    KM_TRY(dmalloc(&w.blo, (size_t)w.kpad * ld));
    KM_TRY(dmalloc(&w.bcorr, (size_t)w.kpad * ld * 2));
    KM_TRY(dmalloc(&w.b16, (size_t)w.kpad * ld64(t->ld) * 2));
    KM_TRY(dmalloc(&w.cnorm, (size_t)w.kpad));
    KM_TRY(dmalloc(&w.cnorm64, (size_t)w.kpad + 4));
    KM_TRY(dmalloc(&w.rc_score, (size_t)t->num_sms * 8 * 8 * kRecheckMaxSplit));   // nsplit > 1 only below 8 rows per CTA
    KM_TRY(dmalloc(&w.rc_idx, (size_t)t->num_sms * 8 * 8 * kRecheckMaxSplit));
    {
        std::vector<double> inf((size_t)w.kpad + 4, INFINITY);   // the 4 slots past kpad are never rewritten
        CU_TRY(cudaMemcpy(w.cnorm64, inf.data(), inf.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    KM_TRY(dmalloc(&w.cstats, 8));
    KM_TRY(dmalloc(&w.labels, n));
    KM_TRY(dmalloc(&w.order, n));
    KM_TRY(dmalloc(&w.flag_list, n));
    KM_TRY(dmalloc(&w.flag_count, 4));
    KM_TRY(dmalloc(&w.counts, (size_t)k + 1));
    KM_TRY(dmalloc(&w.offsets, (size_t)k + 1));
    KM_TRY(dmalloc(&w.cursor, (size_t)k + 1));
    KM_TRY(dmalloc(&w.chunk_off, (size_t)k + 1));
    if (k <= kHistSmemMax) {
        const size_t blocks = (n + kBucketRows - 1) / kBucketRows, nseg = (blocks + kScanSeg - 1) / kScanSeg;
        KM_TRY(dmalloc(&w.hb, blocks * (size_t)k));
        KM_TRY(dmalloc(&w.segsum, nseg * (size_t)k));
    }
    KM_TRY(dmalloc(&w.part, (size_t)w.max_chunks * ld));
    KM_TRY(dmalloc(&w.cpart, (size_t)w.max_chunks));
    KM_TRY(dmalloc(&w.red, (size_t)k * ld + k + 1));
    KM_TRY(dmalloc(&w.costj, (size_t)k));
    KM_TRY(dmalloc(&w.flags, 4));
    w.fused_ctas = t->num_sms * 3;   // partials for up to 3 CTAs of 256 threads per SM (the MMA variant; the CUDA-core kernel runs 2)
    KM_TRY(dmalloc(&w.fused_part, (size_t)w.fused_ctas * ((size_t)k * ld + k + 1)));
    KM_TRY(dmalloc(&w.fused_stage, 1024 + 32 + 2048));
    CU_TRY(cudaMallocHost(reinterpret_cast<void **>(&w.host), sizeof(Workspace::Host)));
    KM_TRY(make_map(&w.tm_x, t->X, (uint64_t)t->rows, ld, kTileM));
    KM_TRY(make_map(&w.tm_bhi, w.bhi, (uint64_t)w.kpad, ld, (uint32_t)w.bn));
    KM_TRY(make_map(&w.tm_blo, w.blo, (uint64_t)w.kpad, ld, (uint32_t)w.bn));
    KM_TRY(make_map(&w.tm_bhi_half, w.bhi, (uint64_t)w.kpad, ld, (uint32_t)std::max(w.bn / 2, 8)));
    KM_TRY(make_map(&w.tm_blo_half, w.blo, (uint64_t)w.kpad, ld, (uint32_t)std::max(w.bn / 2, 8)));
    KM_TRY(make_map_bf16(&w.tm_bcorr_half, w.bcorr, (uint64_t)w.kpad, 2 * ld, (uint32_t)std::max(w.bn / 2, 8)));
    KM_TRY(make_map_bf16(&w.tm_b16_half, w.b16, (uint64_t)w.kpad, 2 * ld64(t->ld), (uint32_t)std::max(w.bn / 2, 8), true));
    w.ready = true;
    return 0;
}

// ------------------------------------------------------------------------------------ launches
// Relative bound on |tensor-core score - exact score| / (2|x||c|max + |c|max^2).  3xTF32: the FP32
// accumulator in TMEM is updated 3*ld/8 times per score and each update may lose up to one ulp
// (the tensor core truncates), plus the TF32 hi/lo split residuals (~3*2^-22) and the FP32 norm /
// final add; measured maxima sit at ~0.35 of this (tests/test_gpu_parity.py).
static float score_eps(int ld, bool bf16_corr = false) {
    // BF16 corrections: 2*ld/8 accumulator updates + split residuals (each correction product carries
    // two BF16 roundings on a term that is already 2^-11 of the main one: <= 2^-18 in total)
    if (g_passes == 3 && bf16_corr) return (float)(g_eps_scale * ldexp(2.0 * ld / 8.0 + 16.0 + 64.0, -24));
    if (g_passes == 3) return (float)(g_eps_scale * ldexp(3.0 * ld / 8.0 + 16.0, -24));
    return (float)(g_eps_scale * 1.25 * ldexp(1.0, -10));
}

// What one launch of an assign kernel reads and writes: the table itself or one of the gathered copies.
struct H16Src {
    const CUtensorMap *tm;    // rows to score
    const float *xnorm;
    int32_t *labels;
    int64_t rows;
    int32_t *cand_list;       // MODE 2: where the undecided rows are listed
    int32_t *cand_count;
    int64_t cand_cap;
    bool prof;                // role counters (diagnostics) for this launch
};

template <int BN>
static int launch_assign_bn(kmb200_table *t, const H16Src *src = nullptr) {
    Workspace &w = t->ws;
    using Cfg = AssignCfg<BN>;
    static bool attr_set[64] = {false};
    const int stages = Cfg::stages();
    const int smem = Cfg::smem_bytes(stages);
    if (!attr_set[t->device & 63]) {
        CU_TRY(cudaFuncSetAttribute(assign_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set[t->device & 63] = true;
    }
    const int64_t rows = src ? src->rows : t->rows;
    AssignParams p;
    p.cnorm = w.cnorm;
    p.xnorm = src ? src->xnorm : t->xnorm;
    p.cstats = w.cstats;
    p.labels = src ? src->labels : w.labels;
    p.flag_count = w.flag_count;
    p.flag_list = w.flag_list;
    p.flag_cap = (int32_t)t->rows;
    p.n = (int32_t)rows;
    p.num_row_tiles = (int32_t)((rows + kTileM - 1) / kTileM);
    p.num_col_tiles = w.kpad / BN;
    p.num_k_chunks = t->ld / kChunkK;
    p.stages = stages;
    p.passes = g_passes;
    p.eps = score_eps(t->ld);
    p.dbg_scores = w.dbg_scores;
    p.kpad = w.kpad;
    p.knockout = g_knockout;
    p.role_prof = g_role_prof;
    const int grid = std::min(p.num_row_tiles, t->num_sms);
    assign_tc_kernel<BN><<<grid, kAssignThreads, smem, t->stream>>>(src ? *src->tm : w.tm_x, w.tm_bhi, w.tm_blo, p);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    return 0;
}

template <int BN, int CL>
static int launch_assign2_bn(kmb200_table *t) {
    Workspace &w = t->ws;
    using Cfg = Assign2Cfg<BN, CL>;
    static bool attr_set[64] = {false};
    const int kc = t->ld / kChunkK;
    const int stages = Cfg::stages(kc);
    const int smem = Cfg::smem_bytes(kc, stages);
    if (!attr_set[t->device & 63]) {
        CU_TRY(cudaFuncSetAttribute(assign_tc2_kernel<BN, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[t->device & 63] = true;
    }
    AssignParams p;
    p.cnorm = w.cnorm;
    p.xnorm = t->xnorm;
    p.cstats = w.cstats;
    p.labels = w.labels;
    p.flag_count = w.flag_count;
    p.flag_list = w.flag_list;
    p.flag_cap = (int32_t)t->rows;
    p.n = (int32_t)t->rows;
    p.num_row_tiles = (int32_t)((t->rows + kTileM - 1) / kTileM);
    p.num_col_tiles = w.kpad / BN;
    p.num_k_chunks = kc;
    p.stages = stages;
    p.passes = 3;
    p.eps = score_eps(t->ld);
    p.dbg_scores = w.dbg_scores;
    p.kpad = w.kpad;
    p.knockout = g_knockout;
    p.role_prof = g_role_prof;
    const int clusters = std::min((p.num_row_tiles + CL - 1) / CL, t->num_sms / CL);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(clusters * CL));
    cfg.blockDim = dim3(kAssignThreads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = t->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CU_TRY(cudaLaunchKernelEx(&cfg, assign_tc2_kernel<BN, CL>, w.tm_x, CL == 2 ? w.tm_bhi_half : w.tm_bhi,
                              CL == 2 ? w.tm_blo_half : w.tm_blo, p));
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    return 0;
}

template <int BN, bool DUMP>   // DUMP: plus the score dump used by kmb200_debug_scores
static int launch_assign_pair_bn(kmb200_table *t, bool *applicable) {
    Workspace &w = t->ws;
    using Cfg = PairSsCfg<BN>;
    auto kernel = assign_pair_ss_kernel<BN, DUMP>;
    static bool attr_set[64] = {false};
    const int kc = t->ld / kChunkK;
    int rb = std::min(Cfg::group_slots(kc), g_rb_cap);
    if (rb == 3) rb = 2;   // the kernel's completion ring has period 4: 2 or 4 group slots
    *applicable = rb >= 2;
    if (rb < 2) return 0;
    const int smem = Cfg::smem_bytes(kc, rb);
    if (!attr_set[t->device & 63]) {
        CU_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[t->device & 63] = true;
    }
    AssignParams p;
    p.cnorm = w.cnorm;
    p.xnorm = t->xnorm;
    p.cstats = w.cstats;
    p.labels = w.labels;
    p.flag_count = w.flag_count;
    p.flag_list = w.flag_list;
    p.flag_cap = (int32_t)t->rows;
    p.n = (int32_t)t->rows;
    p.num_row_tiles = (int32_t)((t->rows + kTileM - 1) / kTileM);
    p.num_col_tiles = w.kpad / BN;
    p.num_k_chunks = kc;
    p.stages = rb;
    p.passes = 3;
    p.eps = score_eps(t->ld, true);
    p.dbg_scores = w.dbg_scores;
    p.kpad = w.kpad;
    p.knockout = g_knockout;
    p.role_prof = g_role_prof;
    const int pairs = std::min((p.num_row_tiles + 1) / 2, t->num_sms / 2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(pairs * 2));
    cfg.blockDim = dim3(kPairSsThreads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = t->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CU_TRY(cudaLaunchKernelEx(&cfg, kernel, w.tm_x, w.tm_bhi_half, w.tm_bcorr_half, p));
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    return 0;
}

// FP16-split pair kernel (assign_pair_h16.cuh).  The score error: 3*ld/16 accumulator updates (one unit each, as for 3xTF32;
// measured maximum at ld = 256: 20 x 2^-24 against 80), split residuals 3*2^-22 and dropped h2.g2 2^-22, FP32 norm, final FMA.
static float score_eps_h16(int ld) { return (float)(g_eps_scale * ldexp(3.0 * ld / 16.0 + 32.0, -24)); }

template <int BN>
static bool h16_fits(const kmb200_table *t) {
    const int kc = t->ld / kChunkK;
    return PairH16Cfg<BN>::x_slots(kc) >= 1 && PairH16Cfg<BN>::b_groups(kc) >= 2;
}
static bool h16_applies(const kmb200_table *t) {
    const Workspace &w = t->ws;
    if (!g_h16 || g_force_simt || g_kernel != 2 || g_passes != 3 || g_cluster != 2 || t->ld > 256 || w.bn < 32) return false;
    if ((t->rows + kTileM - 1) / kTileM < 2) return false;
    return w.bn == 32 ? h16_fits<32>(t) : w.bn == 64 ? h16_fits<64>(t) : h16_fits<128>(t);
}

constexpr int64_t kTwoLevelMinRows = 1 << 18;

// 256 < ld <= 512 (BASELINE config 4, d = 512): the three-product kernel does not fit (h1 and h2 of a row tile would
// take all 512 TMEM columns), the one-product pass does.  Large tables get the cascade with the streaming 3xTF32
// kernel as its second stage (launch_assign_h16_wide); everything else stays on the streaming kernel alone.
static bool h16_wide_applies(const kmb200_table *t) {
    const Workspace &w = t->ws;
    if (!g_h16 || !g_cascade || !g_two_level || g_force_simt || g_kernel != 2 || g_passes != 3 || g_cluster != 2) return false;
    if (t->ld <= 256 || t->ld > 512 || w.bn < 32 || w.dbg_scores || w.cascade_off) return false;
    if (g_two_level == 1 && t->rows < kTwoLevelMinRows) return false;
    if ((t->rows + kTileM - 1) / kTileM < 2) return false;
    return w.bn == 32 ? h16_fits<32>(t) : w.bn == 64 ? h16_fits<64>(t) : h16_fits<128>(t);
}

// (Re)allocates the candidate buffers when the two-level scheme applies to this table; false = use full bookkeeping.
static bool two_level_prepare(kmb200_table *t) {
    Workspace &w = t->ws;
    if (!g_two_level || w.two_level_off || w.dbg_scores) return false;
    if (g_two_level == 1 && t->rows < kTwoLevelMinRows) return false;
    int64_t cap = g_cand_cap > 0 ? g_cand_cap : t->rows / 16;
    cap = std::max<int64_t>((cap + kTileM - 1) / kTileM * kTileM, 2 * kTileM);
    if (w.cand_cap == cap && w.xc) return true;
    for (void *ptr : {(void *)w.xc, (void *)w.xnc, (void *)w.cand_list, (void *)w.labels_c})
        if (ptr) cudaFree(ptr);
    w.xc = nullptr; w.xnc = nullptr; w.cand_list = nullptr; w.labels_c = nullptr;
    w.cand_cap = 0;
    bool ok = cudaMalloc(reinterpret_cast<void **>(&w.xc), (size_t)cap * t->ld * sizeof(float)) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void **>(&w.xnc), (size_t)cap * sizeof(float)) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void **>(&w.cand_list), (size_t)cap * sizeof(int32_t)) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void **>(&w.labels_c), (size_t)cap * sizeof(int32_t)) == cudaSuccess;
    if (ok) ok = make_map(&w.tm_xc, w.xc, (uint64_t)cap, (uint64_t)t->ld, kTileM) == 0;
    if (!ok) {   // not enough memory left for the copies: the single-pass scheme needs none
        cudaGetLastError();
        for (void *ptr : {(void *)w.xc, (void *)w.xnc, (void *)w.cand_list, (void *)w.labels_c})
            if (ptr) cudaFree(ptr);
        w.xc = nullptr; w.xnc = nullptr; w.cand_list = nullptr; w.labels_c = nullptr;
        w.two_level_off = true;
        return false;
    }
    w.cand_cap = cap;
    return true;
}

// (Re)allocates the copy that holds the rows the coarse pass leaves undecided; false = no cascade.
static bool cascade_prepare(kmb200_table *t) {
    Workspace &w = t->ws;
    if (!g_cascade || w.cascade_off) return false;
    int64_t cap = (int64_t)(g_cand1_frac * (double)t->rows);
    cap = std::max<int64_t>((cap + kTileM - 1) / kTileM * kTileM, 2 * kTileM);
    if (w.cand1_cap == cap && w.xc1) return true;
    for (void *ptr : {(void *)w.xc1, (void *)w.xnc1, (void *)w.cand1_list, (void *)w.labels_c1})
        if (ptr) cudaFree(ptr);
    w.xc1 = nullptr; w.xnc1 = nullptr; w.cand1_list = nullptr; w.labels_c1 = nullptr;
    w.cand1_cap = 0;
    bool ok = cudaMalloc(reinterpret_cast<void **>(&w.xc1), (size_t)cap * t->ld * sizeof(float)) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void **>(&w.xnc1), (size_t)cap * sizeof(float)) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void **>(&w.cand1_list), (size_t)cap * sizeof(int32_t)) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void **>(&w.labels_c1), (size_t)cap * sizeof(int32_t)) == cudaSuccess;
    if (ok) ok = make_map(&w.tm_xc1, w.xc1, (uint64_t)cap, (uint64_t)t->ld, kTileM) == 0;
    if (!ok) {   // not enough memory left: the two-level scheme alone needs a fifth of it
        cudaGetLastError();
        for (void *ptr : {(void *)w.xc1, (void *)w.xnc1, (void *)w.cand1_list, (void *)w.labels_c1})
            if (ptr) cudaFree(ptr);
        w.xc1 = nullptr; w.xnc1 = nullptr; w.cand1_list = nullptr; w.labels_c1 = nullptr;
        w.cascade_off = true;
        return false;
    }
    w.cand1_cap = cap;
    return true;
}

// Score error of the one-product pass, the part relative to (2|x||c| + |c|^2): ld/16 accumulator updates, FP32 norm
// and final FMA ...
static float score_eps_h16_one(int ld) { return (float)(g_eps_scale * ldexp(ld / 16.0 + 32.0, -24)); }
// ... and its operand roundings: x sx -> h1 and -2c sw -> g1 are each within 2^-11 (relative, element by element), so every
// product is within 2^-10 + 2^-22 and the cross term -2 x.c within (2^-10 + 2^-22) * 2 |x| |c| (Cauchy-Schwarz).  The
// centre norm is added in FP32 and carries no such error, so this part of the bound scales with 2 |x| |c| alone.
static float score_eps_h16_one_x() { return (float)(g_eps_scale * (ldexp(1.0, -10) + ldexp(1.0, -22))); }

template <int BN, int MODE, int NP>
static int launch_assign_h16_bn(kmb200_table *t, const H16Src &src) {
    Workspace &w = t->ws;
    using Cfg = PairH16Cfg<BN>;
    auto kernel = assign_pair_h16_kernel<BN, MODE, NP>;
    static bool attr_set[64] = {false};
    const int kc = t->ld / kChunkK;
    const int smem = Cfg::smem_bytes(kc);
    if (!attr_set[t->device & 63]) {
        CU_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[t->device & 63] = true;
    }
    AssignParams p;
    p.cnorm = w.cnorm;
    p.xnorm = src.xnorm;
    p.cstats = w.cstats;
    p.labels = src.labels;
    p.flag_count = w.flag_count;
    p.flag_list = w.flag_list;
    p.flag_cap = (int32_t)t->rows;
    p.n = (int32_t)src.rows;
    p.num_row_tiles = (int32_t)((src.rows + kTileM - 1) / kTileM);
    p.num_col_tiles = w.kpad / BN;
    p.num_k_chunks = kc;
    p.stages = Cfg::x_slots(kc) | (Cfg::b_groups(kc) << 8);
    p.passes = 3;
    p.eps = NP == 1 ? score_eps_h16_one(t->ld) : score_eps_h16(t->ld);
    p.eps_x = NP == 1 ? score_eps_h16_one_x() : 0.f;
    p.dbg_scores = w.dbg_scores;
    p.kpad = w.kpad;
    p.knockout = g_knockout;
    p.role_prof = src.prof ? g_role_prof : nullptr;
    p.sqrt_ld = sqrtf((float)t->ld);
    p.n_dev = nullptr;
    p.cand_list = src.cand_list;
    p.cand_count = src.cand_count;
    p.cand_cap = (int32_t)src.cand_cap;
    const int pairs = std::min((p.num_row_tiles + 1) / 2, t->num_sms / 2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(pairs * 2));
    cfg.blockDim = dim3(kPairH16Threads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = t->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CU_TRY(cudaLaunchKernelEx(&cfg, kernel, *src.tm, w.tm_b16_half, p));
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    return 0;
}

template <int MODE, int NP>
static int launch_assign_h16_mode(kmb200_table *t, const H16Src &src) {
    switch (t->ws.bn) {
        case 32: return launch_assign_h16_bn<32, MODE, NP>(t, src);
        case 64: return launch_assign_h16_bn<64, MODE, NP>(t, src);
        default: return launch_assign_h16_bn<128, MODE, NP>(t, src);
    }
}

static int read_count(kmb200_table *t, int32_t *host_dst, const int32_t *dev_src) {
    CU_TRY(cudaMemcpyAsync(host_dst, dev_src, sizeof(int32_t), cudaMemcpyDeviceToHost, t->stream));
    CU_TRY(cudaStreamSynchronize(t->stream));
    return 0;
}

static int launch_gather(kmb200_table *t, const float *X, const float *xnorm, const int32_t *list, const int32_t *count,
                         int64_t cap, float *xc, float *xnc) {
    gather_rows_kernel<<<t->num_sms * 8, 256, 0, t->stream>>>(X, xnorm, list, count, (int32_t)cap, t->ld, xc, xnc);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    return 0;
}

// Near-tie bookkeeping over `src` (the table or the coarse pass's copy, level 0 / 1): either one full-bookkeeping
// pass, or -- two-level scheme -- a FAST pass (winner + runner-up value, 6 instead of 11 instructions per score) that
// lists the ~1 % of rows whose two best scores are inside the error bound, a gather of those rows into a dense copy
// (xc) and the full-bookkeeping pass over the copy only.  Leaves the (row, runner-up) / (row, all) lists for
// launch_recheck and records which copy they index.
static int launch_h16_fine(kmb200_table *t, const H16Src &src, const float *srcX, int level, bool two_level, bool bracket) {
    Workspace &w = t->ws;
    if (!two_level) {
        w.recheck_level = level;
        if (level == 0) g_prof.assign_variant = 3;
        return launch_assign_h16_mode<0, 3>(t, src);
    }
    if (level == 0) g_prof.assign_variant = 2;
    H16Src fast = src;
    fast.cand_list = w.cand_list;
    fast.cand_count = w.flag_count + 2;
    fast.cand_cap = w.cand_cap;
    if (bracket && w.main_ev) CU_TRY(cudaEventRecord(w.main_ev[0], t->stream));
    KM_TRY((launch_assign_h16_mode<2, 3>(t, fast)));
    if (bracket && w.main_ev) {
        CU_TRY(cudaEventRecord(w.main_ev[1], t->stream));
        w.main_ev = nullptr;   // recorded: the caller must not bracket the whole phase
    }
    // The candidate count decides what runs next, so it is read back here: one short host round trip.  Overflow of
    // the candidate buffer (a large share of near-ties) switches this fit to the single-pass scheme on the spot.
    KM_TRY(read_count(t, &w.host->cand_count, w.flag_count + 2));
    if ((int64_t)w.host->cand_count > w.cand_cap) {
        w.two_level_off = true;
        if (g_prof.two_level_dropped_at < 0) g_prof.two_level_dropped_at = g_prof.iterations;
        CU_TRY(cudaMemsetAsync(w.flag_count, 0, 3 * sizeof(int32_t), t->stream));   // [3] (the coarse count) stays
        w.recheck_level = level;
        return launch_assign_h16_mode<0, 3>(t, src);
    }
    w.recheck_level = 2;
    w.scatter2 = true;
    if (w.host->cand_count == 0) return 0;   // no near-ties at all: the lists stay empty
    KM_TRY(launch_gather(t, srcX, src.xnorm, w.cand_list, w.flag_count + 2, w.cand_cap, w.xc, w.xnc));
    H16Src fine = {&w.tm_xc, w.xnc, w.labels_c, (int64_t)w.host->cand_count, nullptr, w.flag_count + 2, 0, false};
    return launch_assign_h16_mode<0, 3>(t, fine);
}

// The assign phase on the FP16-split kernel.  Cascade (tables of >= 2^18 rows): a COARSE pass over all rows with
// the h1.g1 product only (a third of the tensor work; error bound 2^-10) decides every row whose two best scores
// are further apart than that bound -- 80 % of them on the benchmark's blob data -- and lists the rest; those rows
// are gathered into a dense copy and go through the three-product kernel (two-level bookkeeping as above), and
// what that cannot decide goes to the FP64 re-check.  Every row is still scored against every centre on the
// tensor cores, and the labels are the FP64 arg-min whatever path a row takes.  If the coarse pass leaves more
// than cand1_frac (30 %) of the rows undecided the fit drops it and continues with the two-level scheme.
static int launch_assign_h16(kmb200_table *t) {
    Workspace &w = t->ws;
    w.two_level_active = false;
    w.recheck_level = 0;
    w.scatter1 = w.scatter2 = false;
    H16Src table = {&w.tm_x, t->xnorm, w.labels, t->rows, nullptr, w.flag_count + 2, 0, true};
    if (w.dbg_scores) return g_dbg_products == 1 ? launch_assign_h16_mode<1, 1>(t, table) : launch_assign_h16_mode<1, 3>(t, table);
    const bool two_level = two_level_prepare(t);
    // The coarse pass pays when a row tile carries real tensor work.  With one centroid tile and one or two chunks (C2:
    // d = 64, k = 100) a pass costs the same ~3000 cycles of hand-offs per row tile whatever it multiplies, so a second
    // pass over 13 % of the rows only adds to the first: such shapes go straight to the three-product fast pass.
    const bool worth = (w.kpad / w.bn) * ((t->ld / kChunkK + 1) / 2) >= 4 || g_cascade == 2;
    if (two_level && worth && cascade_prepare(t)) {
        H16Src coarse = table;
        coarse.cand_list = w.cand1_list;
        coarse.cand_count = w.flag_count + 3;
        coarse.cand_cap = w.cand1_cap;
        cudaEvent_t *ev = w.main_ev;
        if (ev) CU_TRY(cudaEventRecord(ev[0], t->stream));
        KM_TRY((launch_assign_h16_mode<2, 1>(t, coarse)));
        if (ev) CU_TRY(cudaEventRecord(ev[1], t->stream));
        KM_TRY(read_count(t, &w.host->cand1_count, w.flag_count + 3));
        if ((int64_t)w.host->cand1_count <= w.cand1_cap) {
            g_prof.assign_variant = 1;
            g_prof.coarse_undecided_rows += w.host->cand1_count;
            w.main_ev = nullptr;   // bracketed above
            w.scatter1 = true;
            w.recheck_level = 1;
            if (w.host->cand1_count == 0) return 0;
            KM_TRY(launch_gather(t, t->X, t->xnorm, w.cand1_list, w.flag_count + 3, w.cand1_cap, w.xc1, w.xnc1));
            H16Src mid = {&w.tm_xc1, w.xnc1, w.labels_c1, (int64_t)w.host->cand1_count, nullptr, w.flag_count + 2, 0, false};
            return launch_h16_fine(t, mid, w.xc1, 1, !w.two_level_off, false);
        }
        w.cascade_off = true;   // too many undecided rows: the coarse pass is not worth its time on this data
        if (g_prof.cascade_dropped_at < 0) g_prof.cascade_dropped_at = g_prof.iterations;
        CU_TRY(cudaMemsetAsync(w.flag_count, 0, 4 * sizeof(int32_t), t->stream));
    }
    return launch_h16_fine(t, table, t->X, 0, two_level && !w.two_level_off, true);
}

static int launch_assign_stream(kmb200_table *t, const H16Src *src) {
    switch (t->ws.bn) {
        case 16: return launch_assign_bn<16>(t, src);
        case 32: return launch_assign_bn<32>(t, src);
        case 64: return launch_assign_bn<64>(t, src);
        default: return launch_assign_bn<128>(t, src);
    }
}

// 256 < ld <= 512: coarse one-product pass of the FP16-split kernel over all rows, then the streaming 3xTF32 kernel
// (full bookkeeping) over the gathered copy of the rows it left undecided, then the FP64 re-check as everywhere.
// *done = false: the cascade does not apply (any more) and nothing was launched.
static int launch_assign_h16_wide(kmb200_table *t, bool *done) {
    Workspace &w = t->ws;
    *done = false;
    w.two_level_active = false;
    w.recheck_level = 0;
    w.scatter1 = w.scatter2 = false;
    if (!cascade_prepare(t)) return 0;
    H16Src coarse = {&w.tm_x, t->xnorm, w.labels, t->rows, w.cand1_list, w.flag_count + 3, w.cand1_cap, true};
    cudaEvent_t *ev = w.main_ev;
    if (ev) CU_TRY(cudaEventRecord(ev[0], t->stream));
    KM_TRY((launch_assign_h16_mode<2, 1>(t, coarse)));
    if (ev) CU_TRY(cudaEventRecord(ev[1], t->stream));
    KM_TRY(read_count(t, &w.host->cand1_count, w.flag_count + 3));
    if ((int64_t)w.host->cand1_count > w.cand1_cap) {
        w.cascade_off = true;   // too many undecided rows: the streaming kernel alone from here on
        if (g_prof.cascade_dropped_at < 0) g_prof.cascade_dropped_at = g_prof.iterations;
        CU_TRY(cudaMemsetAsync(w.flag_count, 0, 4 * sizeof(int32_t), t->stream));
        return 0;
    }
    *done = true;
    g_prof.assign_variant = 8;
    g_prof.coarse_undecided_rows += w.host->cand1_count;
    w.main_ev = nullptr;   // bracketed above
    w.scatter1 = true;
    w.recheck_level = 1;
    if (w.host->cand1_count == 0) return 0;
    KM_TRY(launch_gather(t, t->X, t->xnorm, w.cand1_list, w.flag_count + 3, w.cand1_cap, w.xc1, w.xnc1));
    H16Src mid = {&w.tm_xc1, w.xnc1, w.labels_c1, (int64_t)w.host->cand1_count, nullptr, nullptr, 0, false};
    return launch_assign_stream(t, &mid);
}

static int launch_assign_pair(kmb200_table *t, bool *applicable) {
    switch (t->ws.bn) {
#define KMB_PAIR(BN_) \
    return t->ws.dbg_scores ? launch_assign_pair_bn<BN_, true>(t, applicable) : launch_assign_pair_bn<BN_, false>(t, applicable)
        case 32: KMB_PAIR(32);
        case 64: KMB_PAIR(64);
        case 128: KMB_PAIR(128);
#undef KMB_PAIR
        default: *applicable = false; return 0;   // pair MMAs need N >= 32
    }
}

template <int CL>
static int launch_assign2(kmb200_table *t) {
    switch (t->ws.bn) {
        case 16: return launch_assign2_bn<16, 1>(t);   // pair MMAs need N >= 32
        case 32: return launch_assign2_bn<32, CL>(t);
        case 64: return launch_assign2_bn<64, CL>(t);
        default: return launch_assign2_bn<128, CL>(t);
    }
}

static int fused_kp(int k) { return k <= 8 ? 8 : k <= 16 ? 16 : 32; }
static bool fused_applies(const kmb200_table *t) {
    const Workspace &w = t->ws;
    return g_fused && !g_force_simt && (t->ld == 32 || t->ld == 64) && w.k <= 32 &&
           (int64_t)fused_kp(w.k) * t->ld <= kFusedMaxKLd;
}

// FP32 copy of the centres padded to KP rows (+inf norm for padding), staged in device memory and then
// copied device-to-device into the kernel's __constant__ tables.
__global__ void fused_prepare_kernel(const double *__restrict__ cent, int k, int kp, int ld, float *__restrict__ c32,
                                     float *__restrict__ n32) {
    for (int i = threadIdx.x; i < kp * ld; i += blockDim.x) c32[i] = i < k * ld ? (float)cent[i] : 0.f;
    for (int j = threadIdx.x; j < 32; j += blockDim.x) {
        double s = 0.0;
        if (j < k)
            for (int c = 0; c < ld; ++c) s += cent[(size_t)j * ld + c] * cent[(size_t)j * ld + c];
        n32[j] = j < k ? (float)s : __int_as_float(0x7f800000);
    }
}

template <int LD, int KP>
static int launch_fused_inst(kmb200_table *t, float eps) {
    Workspace &w = t->ws;
    static bool attr_set[64] = {false};
    if (!attr_set[t->device & 63]) {
        CU_TRY(cudaFuncSetAttribute(fused_smallk_kernel<LD, KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set[t->device & 63] = true;
    }
    auto *counter = reinterpret_cast<unsigned long long *>(w.flag_count);
    fused_smallk_kernel<LD, KP><<<t->num_sms * 2, kFusedWarps * 32, fused_smem_bytes(w.k, LD), t->stream>>>(
        t->X, w.cent, w.cstats, t->rows, w.k, eps, w.labels, w.fused_part, counter);
    KM_TRY(post_launch(__LINE__));
    return 0;
}

// One read of X: labels, [sums | counts | cost] in w.red.  flag_count[0..1] holds (as one 64-bit counter) the
// number of rows re-scored in FP64.
static int launch_fused(kmb200_table *t) {
    Workspace &w = t->ws;
    g_prof.assign_variant = 0;
    const int k = w.k, ld = t->ld, kp = fused_kp(k);
    CU_TRY(cudaMemsetAsync(w.flag_count, 0, 4 * sizeof(int32_t), t->stream));
    if (ld == 32 && k <= 16 && g_fused == 1) {
        // k <= 16, ld = 32 (C5): the distances as warp-level FP16-split MMAs (fused_mma_kernel)
        static bool attr_set[64] = {false};
        if (!attr_set[t->device & 63]) {
            CU_TRY(cudaFuncSetAttribute(fused_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
            attr_set[t->device & 63] = true;
        }
        static_assert(sizeof(FusedMmaPrep) <= (1024 + 32 + 2048) * sizeof(float), "staging buffer");
        FusedMmaPrep *prep = reinterpret_cast<FusedMmaPrep *>(w.fused_stage);
        fused_mma_prepare_kernel<<<1, 256, 0, t->stream>>>(w.cent, k, t->xabs, prep);
        KM_TRY(post_launch(__LINE__));
        const float eps_mma = (float)(g_eps_scale * ldexp(3.0 * 32 / 16.0 + 32.0, -24));   // as the tcgen05 FP16-split kernel
        auto *counter = reinterpret_cast<unsigned long long *>(w.flag_count);
        fused_mma_kernel<<<t->num_sms * 3, kFusedWarps * 32, fused_mma_smem_bytes(k), t->stream>>>(
            t->X, t->xnorm, w.cent, prep, t->rows, k, eps_mma, w.labels, w.fused_part, counter);
        KM_TRY(post_launch(__LINE__));
        const int psize1 = k * ld + k + 1;
        fused_reduce_kernel<<<(psize1 + 255) / 256, 256, 0, t->stream>>>(w.fused_part, t->num_sms * 3, psize1, w.red);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches += 3;
        return 0;
    }
    const float eps = (float)(g_eps_scale * ldexp((double)ld + 4.0, -24));   // FP32 FMA chain of ld terms
    float *stage_c = w.fused_stage, *stage_n = w.fused_stage + 1024;
    fused_prepare_kernel<<<1, 256, 0, t->stream>>>(w.cent, k, kp, ld, stage_c, stage_n);
    KM_TRY(post_launch(__LINE__));
    CU_TRY(cudaMemcpyToSymbolAsync(kc_cent, stage_c, (size_t)kp * ld * sizeof(float), 0, cudaMemcpyDeviceToDevice, t->stream));
    CU_TRY(cudaMemcpyToSymbolAsync(kc_norm, stage_n, 32 * sizeof(float), 0, cudaMemcpyDeviceToDevice, t->stream));
    int rc = 0;
    if (ld == 32) rc = kp == 8 ? launch_fused_inst<32, 8>(t, eps) : kp == 16 ? launch_fused_inst<32, 16>(t, eps) : launch_fused_inst<32, 32>(t, eps);
    else rc = kp == 8 ? launch_fused_inst<64, 8>(t, eps) : launch_fused_inst<64, 16>(t, eps);
    KM_TRY(rc);
    const int psize1 = k * ld + k + 1;
    fused_reduce_kernel<<<(psize1 + 255) / 256, 256, 0, t->stream>>>(w.fused_part, t->num_sms * 2, psize1, w.red);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches += 3;
    return 0;
}

static int launch_split(kmb200_table *t) {
    Workspace &w = t->ws;
    CU_TRY(cudaMemsetAsync(w.cstats, 0, 8 * sizeof(float), t->stream));
    env_overrides();
    const bool narrow = h16_applies(t);
    if (narrow || h16_wide_applies(t)) {
        const int64_t count = (int64_t)w.k * t->ld;
        cent_absmax_kernel<<<(int)std::min<int64_t>((count + 255) / 256, 1024), 256, 0, t->stream>>>(w.cent, count, w.cstats);
        KM_TRY(post_launch(__LINE__));
        split16_kernel<<<w.kpad, 128, 0, t->stream>>>(w.cent, w.b16, w.cnorm, w.cnorm64, w.cstats, t->xabs, w.k, t->ld,
                                                     (int)ld64(t->ld));
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches += 2;
        if (narrow) return 0;   // (the wide cascade also needs the TF32 operands of its second stage)
    }
    split_centers_kernel<<<w.kpad, 128, 0, t->stream>>>(w.cent, w.bhi, w.blo, w.bcorr, w.cnorm, w.cnorm64, w.cstats,
                                                        w.k, t->ld);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    return 0;
}

static int launch_assign(kmb200_table *t) {
    Workspace &w = t->ws;
    CU_TRY(cudaMemsetAsync(w.flag_count, 0, 4 * sizeof(int32_t), t->stream));
    if (g_force_simt) {
        recheck_kernel<<<t->num_sms * 8, 256, 0, t->stream>>>(t->X, w.cent, nullptr, nullptr, 0, t->rows, w.k, t->ld,
                                                             w.labels);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches++;
        g_prof.assign_variant = 7;
        return 0;
    }
    env_overrides();
    if (h16_applies(t)) return launch_assign_h16(t);
    if (g_kernel == 2 && g_passes == 3 && t->ld <= 256) {
        const int row_tiles = (int)((t->rows + kTileM - 1) / kTileM);
        if (g_cluster == 2 && row_tiles >= 2) {
            bool applicable = false;
            const int rc = launch_assign_pair(t, &applicable);
            if (applicable) g_prof.assign_variant = 4;
            if (rc != 0 || applicable) return rc;
        }
        g_prof.assign_variant = 5;
        return launch_assign2<1>(t);
    }
    if (h16_wide_applies(t)) {
        bool done = false;
        KM_TRY(launch_assign_h16_wide(t, &done));
        if (done) return 0;
    }
    g_prof.assign_variant = 6;
    return launch_assign_stream(t, nullptr);
}

static int launch_recheck(kmb200_table *t) {
    Workspace &w = t->ws;
    if (g_force_simt) return 0;
    // the lists index the table or one of the gathered copies (launch_assign_h16)
    const float *X = w.recheck_level == 2 ? w.xc : w.recheck_level == 1 ? w.xc1 : t->X;
    int32_t *labels = w.recheck_level == 2 ? w.labels_c : w.recheck_level == 1 ? w.labels_c1 : w.labels;
    recheck_kernel<<<t->num_sms * 4, 256, 0, t->stream>>>(X, w.cent, w.flag_list, w.flag_count, (int32_t)t->rows,
                                                         t->rows, w.k, t->ld, labels);
    KM_TRY(post_launch(__LINE__));
    const int rgrid = t->num_sms * 8;
    recheck_full_kernel<<<rgrid, 256, 0, t->stream>>>(X, w.cent, w.cnorm64, w.flag_list, w.flag_count + 1,
                                                     (int32_t)t->rows, w.k, t->ld, labels, w.rc_score, w.rc_idx);
    KM_TRY(post_launch(__LINE__));
    recheck_merge_kernel<<<8, 256, 0, t->stream>>>(w.flag_list, w.flag_count + 1, (int32_t)t->rows, rgrid, w.rc_score,
                                                  w.rc_idx, labels);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches += 3;
    if (w.scatter2 && w.recheck_level == 2) {   // decided labels of the near-tie copy -> the level they were gathered from
        scatter_labels_kernel<<<t->num_sms * 2, 256, 0, t->stream>>>(w.cand_list, w.flag_count + 2, (int32_t)w.cand_cap,
                                                                    w.labels_c, w.scatter1 ? w.labels_c1 : w.labels);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches++;
    }
    if (w.scatter1) {   // labels of the coarse pass's undecided rows -> the table's labels
        scatter_labels_kernel<<<t->num_sms * 2, 256, 0, t->stream>>>(w.cand1_list, w.flag_count + 3, (int32_t)w.cand1_cap,
                                                                    w.labels_c1, w.labels);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches++;
    }
    return 0;
}

static int launch_bucket(kmb200_table *t) {
    Workspace &w = t->ws;
    const int k = w.k;
    const int blocks = (int)((t->rows + kBucketRows - 1) / kBucketRows);
    const size_t sh = k <= kHistSmemMax ? (size_t)k * sizeof(int32_t) : 0;
    if (k <= kHistSmemMax) {
        // stable: the rows of a cluster keep their row order, so the FP64 sums are reproducible run to run
        const int nseg = (blocks + kScanSeg - 1) / kScanSeg;
        const dim3 sgrid((unsigned)((k + 255) / 256), (unsigned)nseg);
        block_hist_kernel<<<blocks, 256, sh, t->stream>>>(w.labels, t->rows, k, w.hb);
        KM_TRY(post_launch(__LINE__));
        colsum_seg_kernel<<<sgrid, 256, 0, t->stream>>>(w.hb, blocks, k, w.segsum);
        KM_TRY(post_launch(__LINE__));
        seg_scan_kernel<<<(k + 255) / 256, 256, 0, t->stream>>>(w.segsum, nseg, k, w.counts);
        KM_TRY(post_launch(__LINE__));
        colscan_seg_kernel<<<sgrid, 256, 0, t->stream>>>(w.hb, blocks, k, w.segsum);
        KM_TRY(post_launch(__LINE__));
        scan_kernel<<<1, 1024, 0, t->stream>>>(w.counts, k, w.offsets, w.cursor, w.chunk_off);
        KM_TRY(post_launch(__LINE__));
        scatter_stable_kernel<<<blocks, 256, sh, t->stream>>>(w.labels, t->rows, k, w.offsets, w.hb, w.order);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches += 6;
        return 0;
    }
    CU_TRY(cudaMemsetAsync(w.counts, 0, ((size_t)k + 1) * sizeof(int32_t), t->stream));
    label_hist_kernel<<<blocks, 256, sh, t->stream>>>(w.labels, t->rows, k, w.counts);
    KM_TRY(post_launch(__LINE__));
    scan_kernel<<<1, 1024, 0, t->stream>>>(w.counts, k, w.offsets, w.cursor, w.chunk_off);
    KM_TRY(post_launch(__LINE__));
    scatter_kernel<<<blocks, 256, sh, t->stream>>>(w.labels, t->rows, k, w.cursor, w.order);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches += 3;
    return 0;
}

static int launch_accumulate(kmb200_table *t) {
    Workspace &w = t->ws;
    const int ld = t->ld;
    segment_sums_kernel<<<(unsigned)w.max_chunks, 256, 0, t->stream>>>(t->X, w.cent, w.order, w.offsets, w.chunk_off,
                                                                        w.k, ld, w.part, w.cpart);
    KM_TRY(post_launch(__LINE__));
    segment_reduce_kernel<<<w.k, 128, 0, t->stream>>>(w.part, w.cpart, w.offsets, w.chunk_off, w.k, ld, w.red, w.costj);
    KM_TRY(post_launch(__LINE__));
    cost_sum_kernel<<<1, 1024, 0, t->stream>>>(w.costj, w.k, ld, w.red);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches += 3;
    return 0;
}

static int do_allreduce(kmb200_table *t, kmb200_comm *comm) {
    if (!comm || comm->world == 1) return 0;
    Workspace &w = t->ws;
    const ncclResult_t r = g_nccl.AllReduce(w.red, w.red, (size_t)w.k * t->ld + w.k + 1, ncclDouble, ncclSum, comm->comm, t->stream);
    if (r != ncclSuccess) {
        comm->broken = true;
        return fail(KMB200_ERR_COMM, "ncclAllReduce failed: %s", g_nccl.GetErrorString(r));
    }
    return 0;
}

static int upload_centers(kmb200_table *t, const double *centers, int k, int d) {
    Workspace &w = t->ws;
    CU_TRY(cudaMemsetAsync(w.cent, 0, (size_t)w.kpad * t->ld * sizeof(double), t->stream));
    CU_TRY(cudaMemcpy2DAsync(w.cent, (size_t)t->ld * 8, centers, (size_t)d * 8, (size_t)d * 8, (size_t)k,
                             cudaMemcpyHostToDevice, t->stream));
    return 0;
}
static int download_centers(kmb200_table *t, double *out, int k, int d) {
    Workspace &w = t->ws;
    CU_TRY(cudaMemcpy2DAsync(out, (size_t)d * 8, w.cent, (size_t)t->ld * 8, (size_t)d * 8, (size_t)k,
                             cudaMemcpyDeviceToHost, t->stream));
    return 0;
}

// ------------------------------------------------------------------------------------ empty clusters
// oneDAL's finalizeCompute gives an empty cluster the farthest remaining candidate row and takes that
// row's distance out of the objective (recalled from the published source, pinned by no vector
// of the reference's tests).  m = number of empty clusters; every rank finds its m farthest rows ON THE DEVICE -- FP64
// distance of every row to its own centre, then a radix select over the bits of those distances (12 bits per round,
// one small read-back per round, typically two rounds) and a short list of the rows at or above the threshold --
// the lists are all-gathered and merged in (distance desc, rank asc, row asc) order identically on every rank.
struct DevBuf {   // frees what it holds on every exit path
    std::vector<void *> ptrs;
    ~DevBuf() { for (void *p : ptrs) if (p) cudaFree(p); }
    template <typename T> int alloc(T **p, size_t count) {
        *p = nullptr;
        CU_TRY(cudaMalloc(reinterpret_cast<void **>(p), std::max<size_t>(count, 1) * sizeof(T)));
        ptrs.push_back(*p);
        return 0;
    }
};

// The m rows with the largest dist (ties: lower row first), as (distance, row) in that order; fewer than m when the
// table is shorter.
static int select_farthest(kmb200_table *t, const double *dist, int m, std::vector<std::pair<double, int32_t>> *out) {
    const int64_t n = t->rows;
    cudaStream_t st = t->stream;
    const unsigned want = (unsigned)std::min<int64_t>(m, n);
    const unsigned cap = want + 16384;   // the threshold bin may hold this many rows beyond the ones wanted
    DevBuf buf;
    unsigned int *hist = nullptr, *sel = nullptr, *count = nullptr;
    unsigned long long *keys = nullptr;
    int32_t *rows = nullptr;
    KM_TRY(buf.alloc(&hist, 1 << kDigitBits));
    KM_TRY(buf.alloc(&sel, 4));
    KM_TRY(buf.alloc(&count, 1));
    KM_TRY(buf.alloc(&keys, cap));
    KM_TRY(buf.alloc(&rows, cap));
    out->clear();
    if (want == 0) return 0;
    const int grid = std::min<int64_t>((n + 255) / 256, (int64_t)t->num_sms * 8);
    unsigned long long prefix = 0;
    unsigned need = want;
    // digits at bits 52..63 (the sign bit of a distance is 0), 40..51, 28..39, 16..27, 4..15; what is left below bit 4
    // is 16 ulps of an FP64 distance: rows that close count as tied and are listed together
    int shift = 64 - kDigitBits;
    for (int round = 0;; ++round) {
        CU_TRY(cudaMemsetAsync(hist, 0, sizeof(unsigned int) << kDigitBits, st));
        dist_digit_hist_kernel<<<grid, 256, 0, st>>>(dist, n, prefix, shift, round == 0, hist);
        KM_TRY(post_launch(__LINE__));
        digit_select_kernel<<<1, 1024, 0, st>>>(hist, need, sel);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches += 2;
        unsigned h[3];
        CU_TRY(cudaMemcpyAsync(h, sel, sizeof(h), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        prefix = (prefix << kDigitBits) | h[0];
        need -= h[1];                      // still wanted from the threshold bin, 1 <= need <= h[2]
        if (h[2] - need <= 16384 || shift < kDigitBits) break;   // the bin is small enough to be listed whole
        shift -= kDigitBits;
    }
    CU_TRY(cudaMemsetAsync(count, 0, sizeof(unsigned int), st));
    dist_collect_kernel<<<grid, 256, 0, st>>>(dist, n, prefix, shift, keys, rows, count, cap);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    unsigned got = 0;
    CU_TRY(cudaMemcpyAsync(&got, count, sizeof(got), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (got > cap)
        return fail(KMB200_ERR_LIMIT, "empty-cluster candidates: %u rows tie for the farthest %u (exactly equal distances); "
                                      "more than the %u the candidate list holds", got, want, cap);
    std::vector<unsigned long long> hk(got);
    std::vector<int32_t> hr(got);
    CU_TRY(cudaMemcpyAsync(hk.data(), keys, (size_t)got * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(hr.data(), rows, (size_t)got * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    std::vector<int> idx(got);
    for (unsigned i = 0; i < got; ++i) idx[i] = (int)i;
    std::sort(idx.begin(), idx.end(), [&](int a, int b) { return hk[a] != hk[b] ? hk[a] > hk[b] : hr[a] < hr[b]; });
    for (unsigned i = 0; i < std::min(got, want); ++i) {
        double v;
        memcpy(&v, &hk[idx[i]], sizeof(v));
        out->emplace_back(v, hr[idx[i]]);
    }
    return 0;
}

static int handle_empty_clusters(kmb200_table *t, kmb200_comm *comm, int m, double tol, double *cost,
                                 int *moved_extra) {
    Workspace &w = t->ws;
    const int k = w.k, ld = t->ld, d = t->d;
    const int world = comm ? comm->world : 1;
    const int64_t n = t->rows;
    cudaStream_t st = t->stream;
    g_prof.empty_cluster_events++;
    const int rec = 2 + d;  // distance, local row, row values
    std::vector<double> mine((size_t)m * rec, 0.0), all((size_t)m * rec * world, 0.0);
    for (int e = 0; e < m; ++e) mine[(size_t)e * rec] = mine[(size_t)e * rec + 1] = -1.0;   // "no candidate"
    DevBuf buf;
    double *dist = nullptr, *dgather = nullptr;
    int32_t *drows = nullptr;
    float *dvals = nullptr;
    KM_TRY(buf.alloc(&dist, (size_t)n));
    // distances to the centres this iteration's assignment used: w.cent still holds them (finalize
    // wrote the new centres to w.cent2)
    rowdist_kernel<<<t->num_sms * 8, 256, 0, st>>>(t->X, w.cent, w.labels, n, ld, dist);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    std::vector<std::pair<double, int32_t>> far;
    KM_TRY(select_farthest(t, dist, m, &far));
    if (!far.empty()) {
        const int got = (int)far.size();
        std::vector<int32_t> hrows(got);
        for (int e = 0; e < got; ++e) hrows[e] = far[e].second;
        KM_TRY(buf.alloc(&drows, (size_t)got));
        KM_TRY(buf.alloc(&dvals, (size_t)got * d));
        CU_TRY(cudaMemcpyAsync(drows, hrows.data(), (size_t)got * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        gather_list_kernel<<<std::min(got, 1024), 256, 0, st>>>(t->X, drows, got, d, ld, dvals);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches++;
        std::vector<float> hvals((size_t)got * d);
        CU_TRY(cudaMemcpyAsync(hvals.data(), dvals, hvals.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        for (int e = 0; e < got; ++e) {
            double *r = &mine[(size_t)e * rec];
            r[0] = far[e].first;
            r[1] = (double)far[e].second;
            for (int c = 0; c < d; ++c) r[2 + c] = (double)hvals[(size_t)e * d + c];
        }
    }
    if (world > 1) {
        const size_t cnt = (size_t)m * rec;
        KM_TRY(buf.alloc(&dgather, cnt * (world + 1)));
        CU_TRY(cudaMemcpyAsync(dgather, mine.data(), cnt * sizeof(double), cudaMemcpyHostToDevice, st));
        ncclResult_t r = g_nccl.AllGather(dgather, dgather + cnt, cnt, ncclDouble, comm->comm, st);
        if (r != ncclSuccess) {
            comm->broken = true;
            return fail(KMB200_ERR_COMM, "ncclAllGather failed: %s", g_nccl.GetErrorString(r));
        }
        CU_TRY(cudaMemcpyAsync(all.data(), dgather + cnt, cnt * world * sizeof(double), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
    } else {
        all = mine;
    }
    // merge: (distance desc, rank asc, row asc); records are already rank-major, row order inside a
    // rank is distance desc with row-asc ties, so a stable sort by distance keeps the rule.
    std::vector<int> idx((size_t)m * world);
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return all[(size_t)a * rec] > all[(size_t)b * rec]; });
    // counts and the centres of this iteration, one read-back each; then substitute in cluster-index order
    std::vector<double> red_counts(k), oldc((size_t)k * ld);
    CU_TRY(cudaMemcpyAsync(red_counts.data(), w.red + (size_t)k * ld, (size_t)k * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(oldc.data(), w.cent, (size_t)k * ld * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    int used = 0;
    for (int j = 0; j < k; ++j) {
        if (red_counts[j] > 0.0) continue;
        if (used >= (int)idx.size() || !(all[(size_t)idx[used] * rec] > 0.0))
            return fail(KMB200_ERR_EMPTY, "cluster %d is empty and no positive-distance candidate row is left", j);
        const double *r = &all[(size_t)idx[used] * rec];   // (`all` outlives the copies below: synchronised before return)
        ++used;
        double s = 0.0;
        for (int c = 0; c < d; ++c) {
            const double df = r[2 + c] - oldc[(size_t)j * ld + c];
            s += df * df;
        }
        if (!(s <= tol * tol)) (*moved_extra)++;
        CU_TRY(cudaMemcpyAsync(w.cent2 + (size_t)j * ld, r + 2, (size_t)d * sizeof(double), cudaMemcpyHostToDevice, st));
        *cost -= r[0];
    }
    CU_TRY(cudaStreamSynchronize(st));
    return 0;
}

// ------------------------------------------------------------------------------------ profile
static void prof_reset() {
    memset(&g_prof, 0, sizeof(g_prof));
    g_prof.cascade_dropped_at = g_prof.two_level_dropped_at = -1;
}

extern "C" int kmb200_get_profile(kmb200_profile *out) {
    if (!out) return fail(KMB200_ERR_ARG, "get_profile: null");
    *out = g_prof;
    return 0;
}

// ------------------------------------------------------------------------------------ the driver
// Compute calls on one device run one at a time: the one-pass small-k kernel keeps its centres in __constant__
// memory, which every table on a device shares, and the tunables of kmb200_set_option are process-wide.  Calls
// on different devices run concurrently.
static std::mutex g_device_mu[64];
struct DeviceLock {
    std::unique_lock<std::mutex> lk;
    explicit DeviceLock(int dev) : lk(g_device_mu[dev & 63]) {}
};
static int check_fit_args(kmb200_table *t, const double *centers, int k, int d) {
    if (!t || !centers) return fail(KMB200_ERR_ARG, "null table or centres");
    if (k <= 0) return fail(KMB200_ERR_ARG, "k must be positive (got %d)", k);
    if (d != t->d) return fail(KMB200_ERR_ARG, "centre dimension %d != table cols %d", d, t->d);
    if (k > 65536) return fail(KMB200_ERR_LIMIT, "k = %d is beyond the supported 65536", k);
    return 0;
}

static int fit_impl(kmb200_table *t, kmb200_comm *comm, const double *init_centers, int32_t k, int32_t d, double tol,
                    int32_t max_iter, double *out_centers, double *out_cost, int32_t *out_iters);
extern "C" int kmb200_fit(kmb200_table *t, kmb200_comm *comm, const double *init_centers, int32_t k, int32_t d,
                          double tol, int32_t max_iter, double *out_centers, double *out_cost, int32_t *out_iters) {
    const int rc = fit_impl(t, comm, init_centers, k, d, tol, max_iter, out_centers, out_cost, out_iters);
    // a rank that fails inside the collective loop leaves its peers waiting: the communicator is torn down with
    // ncclCommAbort rather than ncclCommDestroy
    if (rc != 0 && rc != KMB200_ERR_ARG && rc != KMB200_ERR_LIMIT && comm) comm->broken = true;
    return rc;
}
static int fit_impl(kmb200_table *t, kmb200_comm *comm, const double *init_centers, int32_t k, int32_t d, double tol,
                    int32_t max_iter, double *out_centers, double *out_cost, int32_t *out_iters) {
    KM_TRY(check_fit_args(t, init_centers, k, d));
    if (!out_centers || !out_cost || !out_iters) return fail(KMB200_ERR_ARG, "fit: null output");
    if (max_iter < 0) return fail(KMB200_ERR_ARG, "fit: max_iter must be >= 0");
    if (comm && comm->world > 1) KM_TRY(nccl_load());
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    KM_TRY(require_device());
    prof_reset();
    KM_TRY(ws_prepare(t, k));
    Workspace &w = t->ws;
    cudaStream_t st = t->stream;
    KM_TRY(upload_centers(t, init_centers, k, d));
    w.two_level_off = false;
    w.cascade_off = false;
    KM_TRY(launch_split(t));
    KM_TRY(t->events.ensure(max_iter));

    double cost = 0.0;
    int it = 0;
    bool converged = false;
    const bool fused = fused_applies(t);
    const int P = t->events.per_iter;
    std::vector<char> main_timed((size_t)std::max(t->events.cap_iters, 1), 0);
    for (it = 0; it < max_iter && !converged; ++it) {
        g_prof.iterations = it;   // (the hand-off records of launch_assign_h16 read it)
        const bool timed = it < t->events.cap_iters;
        cudaEvent_t *ev = timed ? &t->events.ev[(size_t)it * P] : nullptr;
        if (timed) CU_TRY(cudaEventRecord(ev[0], st));
        if (fused) {
            KM_TRY(launch_fused(t));   // distances + argmin + sums + cost in one read of X
            if (timed) {
                CU_TRY(cudaEventRecord(ev[1], st));
                CU_TRY(cudaEventRecord(ev[2], st));
                CU_TRY(cudaEventRecord(ev[3], st));
            }
        } else {
            w.main_ev = timed ? ev + 7 : nullptr;
            KM_TRY(launch_assign(t));
            if (timed) CU_TRY(cudaEventRecord(ev[1], st));
            // the two-level launcher brackets its main kernel itself (and clears w.main_ev); otherwise the phase IS the kernel
            const bool main_separate = timed && w.main_ev == nullptr;
            w.main_ev = nullptr;
            if (timed) main_timed[it] = main_separate;
            KM_TRY(launch_recheck(t));
            if (timed) CU_TRY(cudaEventRecord(ev[2], st));
            KM_TRY(launch_bucket(t));
            if (timed) CU_TRY(cudaEventRecord(ev[3], st));
            KM_TRY(launch_accumulate(t));
        }
        if (timed) CU_TRY(cudaEventRecord(ev[4], st));
        KM_TRY(do_allreduce(t, comm));
        if (timed) CU_TRY(cudaEventRecord(ev[5], st));
        CU_TRY(cudaMemsetAsync(w.flags, 0, 4 * sizeof(int32_t), st));
        finalize_kernel<<<k, 128, 0, st>>>(w.red, w.cent, w.cent2, k, t->ld, d, tol, w.flags);
        KM_TRY(post_launch(__LINE__));
        g_prof.kernel_launches++;
        // one host round trip per iteration: the reference's iteration count must be reproduced
        // exactly, so the converged flag (and the rare empty-cluster count) is read every time
        CU_TRY(cudaMemcpyAsync(&w.host->cost, w.red + (size_t)k * t->ld + k, sizeof(double), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaMemcpyAsync(w.host->flags, w.flags, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaMemcpyAsync(&w.host->flag_count, w.flag_count, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        cost = w.host->cost;
        int moved = w.host->flags[0];
        if (w.host->flags[1] > 0) KM_TRY(handle_empty_clusters(t, comm, w.host->flags[1], tol, &cost, &moved));
        std::swap(w.cent, w.cent2);
        KM_TRY(launch_split(t));  // TF32 split + norms of the new centres for the next assignment
        if (timed) CU_TRY(cudaEventRecord(ev[6], st));
        if (fused) {
            g_prof.rechecked_rows += w.host->flag_count;   // low word of the 64-bit counter (exact below 2^31)
        } else {
            g_prof.rechecked_rows += g_force_simt ? 0 : std::min<int64_t>((int64_t)w.host->flag_count + w.host->full_count, t->rows);
            g_prof.full_rechecks += g_force_simt ? 0 : w.host->full_count;
        }
        converged = tol >= 0.0 && moved == 0;
    }
    KM_TRY(download_centers(t, out_centers, k, d));
    CU_TRY(cudaStreamSynchronize(st));
    *out_cost = cost;
    *out_iters = it;
    g_prof.iterations = it;
    const int timed_iters = std::min(it, t->events.cap_iters);
    for (int i = 0; i < timed_iters; ++i) {
        cudaEvent_t *ev = &t->events.ev[(size_t)i * P];
        float ms[6] = {0, 0, 0, 0, 0, 0};
        for (int s = 0; s < 6; ++s) cudaEventElapsedTime(&ms[s], ev[s], ev[s + 1]);
        g_prof.assign_ms += ms[0];
        g_prof.recheck_ms += ms[1];
        g_prof.bucket_ms += ms[2];
        g_prof.accumulate_ms += ms[3];
        g_prof.allreduce_ms += ms[4];
        g_prof.finalize_ms += ms[5];
        float main_ms = ms[0];
        if (!fused && main_timed[i]) cudaEventElapsedTime(&main_ms, ev[7], ev[8]);
        g_prof.assign_main_ms += main_ms;
    }
    if (timed_iters > 0) {
        float tot = 0.f;
        cudaEventElapsedTime(&tot, t->events.ev[0], t->events.ev[(size_t)(timed_iters - 1) * P + 6]);
        g_prof.total_ms = tot;
    }
    return 0;
}

extern "C" int kmb200_assign(kmb200_table *t, kmb200_comm *comm, const double *centers, int32_t k, int32_t d,
                             int32_t *labels_host, int64_t *sizes_host, double *cost_host) {
    KM_TRY(check_fit_args(t, centers, k, d));
    if (comm && comm->world > 1) KM_TRY(nccl_load());
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    KM_TRY(require_device());
    prof_reset();
    KM_TRY(ws_prepare(t, k));
    Workspace &w = t->ws;
    cudaStream_t st = t->stream;
    KM_TRY(t->events.ensure(1));
    cudaEvent_t *ev = &t->events.ev[0];
    KM_TRY(upload_centers(t, centers, k, d));
    w.two_level_off = false;
    w.cascade_off = false;
    KM_TRY(launch_split(t));
    const bool fused = fused_applies(t);
    CU_TRY(cudaEventRecord(ev[0], st));
    if (fused) {
        KM_TRY(launch_fused(t));
        CU_TRY(cudaEventRecord(ev[1], st));
        CU_TRY(cudaEventRecord(ev[2], st));
        CU_TRY(cudaEventRecord(ev[3], st));
    } else {
        KM_TRY(launch_assign(t));
        CU_TRY(cudaEventRecord(ev[1], st));
        KM_TRY(launch_recheck(t));
        CU_TRY(cudaEventRecord(ev[2], st));
        KM_TRY(launch_bucket(t));
        CU_TRY(cudaEventRecord(ev[3], st));
        KM_TRY(launch_accumulate(t));
    }
    CU_TRY(cudaEventRecord(ev[4], st));
    KM_TRY(do_allreduce(t, comm));
    CU_TRY(cudaEventRecord(ev[5], st));
    if (labels_host)
        CU_TRY(cudaMemcpyAsync(labels_host, w.labels, (size_t)t->rows * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    std::vector<double> tail((size_t)k + 1);
    CU_TRY(cudaMemcpyAsync(tail.data(), w.red + (size_t)k * t->ld, ((size_t)k + 1) * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(&w.host->flag_count, w.flag_count, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (sizes_host)
        for (int j = 0; j < k; ++j) sizes_host[j] = (int64_t)llround(tail[j]);
    if (cost_host) *cost_host = tail[k];
    g_prof.iterations = 1;
    if (fused) {
        g_prof.rechecked_rows = w.host->flag_count;
    } else {
        g_prof.rechecked_rows = g_force_simt ? 0 : std::min<int64_t>((int64_t)w.host->flag_count + w.host->full_count, t->rows);
        g_prof.full_rechecks = g_force_simt ? 0 : w.host->full_count;
    }
    cudaEventElapsedTime(&g_prof.assign_ms, ev[0], ev[1]);
    cudaEventElapsedTime(&g_prof.recheck_ms, ev[1], ev[2]);
    cudaEventElapsedTime(&g_prof.bucket_ms, ev[2], ev[3]);
    cudaEventElapsedTime(&g_prof.accumulate_ms, ev[3], ev[4]);
    cudaEventElapsedTime(&g_prof.allreduce_ms, ev[4], ev[5]);
    cudaEventElapsedTime(&g_prof.total_ms, ev[0], ev[5]);
    return 0;
}

// ------------------------------------------------------------------------------------ k-means|| initialisation (f3)
// Spark's java.util.Random with XORShiftRandom's next(bits) (org.apache.spark.util.random.XORShiftRandom: the seed goes
// through two MurmurHash3.bytesHash rounds of its big-endian bytes, then x ^= x << 21; x ^= x >>> 35; x ^= x << 4).
// Restated from the published Spark source; the reference tree does not hold that file or test vectors for it.
namespace {
inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
uint32_t murmur3_bytes_hash(const uint8_t *data, int len, uint32_t seed) {   // scala.util.hashing.MurmurHash3.bytesHash
    uint32_t h = seed;
    int i = 0, left = len;
    auto mix_last = [](uint32_t hash, uint32_t k) { k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u; return hash ^ k; };
    while (left >= 4) {
        const uint32_t k = data[i] | (data[i + 1] << 8) | (data[i + 2] << 16) | ((uint32_t)data[i + 3] << 24);
        h = mix_last(h, k);
        h = rotl32(h, 13);
        h = h * 5 + 0xe6546b64u;
        i += 4;
        left -= 4;
    }
    uint32_t k = 0;
    if (left == 3) k ^= data[i + 2] << 16;
    if (left >= 2) k ^= data[i + 1] << 8;
    if (left >= 1) { k ^= data[i]; h = mix_last(h, k); }
    h ^= (uint32_t)len;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}
struct XorShiftRandom {
    uint64_t seed;
    explicit XorShiftRandom(int64_t init) {
        uint8_t bytes[8];
        for (int i = 0; i < 8; ++i) bytes[i] = (uint8_t)((uint64_t)init >> (56 - 8 * i));   // ByteBuffer.putLong: big-endian
        const uint32_t low = murmur3_bytes_hash(bytes, 8, 0x3c074a61u);   // MurmurHash3.arraySeed
        const uint32_t high = murmur3_bytes_hash(bytes, 8, low);
        seed = ((uint64_t)high << 32) | low;
    }
    int32_t next(int bits) {
        uint64_t x = seed ^ (seed << 21);
        x ^= x >> 35;
        x ^= x << 4;
        seed = x;
        return (int32_t)(x & ((1ull << bits) - 1));
    }
    int32_t next_int() { return next(32); }
    double next_double() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }   // java.util.Random.nextDouble
};
}  // namespace

extern "C" int32_t kmb200_xorshift_next_int(int64_t seed) { return XorShiftRandom(seed).next_int(); }
extern "C" double kmb200_xorshift_next_double(int64_t seed, int32_t skip) {
    XorShiftRandom r(seed);
    for (int i = 0; i < skip; ++i) r.next_double();
    return r.next_double();
}

// One distance pass of initKMeansParallel (spark-3.1.1/mllib/clustering/KMeans.scala:395-401): cost_i = min(cost_i,
// min_j ||x_i - c_j||^2) over the m centres added in the last step, and the sum of the new costs over all ranks.  The
// nearest of the m centres comes from the assign path (FP64-faithful labels), the distance to it is FP64 direct-form.
extern "C" int kmb200_init_costs(kmb200_table *t, kmb200_comm *comm, const double *new_centers, int32_t m, int32_t d,
                                 int32_t first_step, double *sum_costs) {
    KM_TRY(check_fit_args(t, new_centers, m, d));
    if (!sum_costs) return fail(KMB200_ERR_ARG, "init_costs: null output");
    if (comm && comm->world > 1) KM_TRY(nccl_load());
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    KM_TRY(require_device());
    prof_reset();
    if (!t->init_cost) {
        CU_TRY(cudaMalloc(&t->init_cost, (size_t)t->rows * sizeof(double)));
        CU_TRY(cudaMalloc(&t->init_tmp, ((size_t)t->rows + 1024 + 8) * sizeof(double)));
        first_step = 1;
    }
    KM_TRY(ws_prepare(t, m));
    Workspace &w = t->ws;
    cudaStream_t st = t->stream;
    KM_TRY(upload_centers(t, new_centers, m, d));
    w.two_level_off = false;
    w.cascade_off = false;
    KM_TRY(launch_split(t));
    if (fused_applies(t)) {
        KM_TRY(launch_fused(t));
    } else {
        KM_TRY(launch_assign(t));
        KM_TRY(launch_recheck(t));
    }
    rowdist_kernel<<<t->num_sms * 8, 256, 0, st>>>(t->X, w.cent, w.labels, t->rows, t->ld, t->init_tmp);
    KM_TRY(post_launch(__LINE__));
    const int blocks = (int)std::min<int64_t>(1024, (t->rows + 255) / 256);
    double *part = t->init_tmp + t->rows, *total = part + 1024;
    init_min_kernel<<<blocks, 256, 0, st>>>(t->init_cost, t->init_tmp, t->rows, first_step != 0, part);
    KM_TRY(post_launch(__LINE__));
    sum_parts_kernel<<<1, 32, 0, st>>>(part, blocks, total);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches += 3;
    if (comm && comm->world > 1) {
        const ncclResult_t r = g_nccl.AllReduce(total, total, 1, ncclDouble, ncclSum, comm->comm, st);
        if (r != ncclSuccess) {
            comm->broken = true;
            return fail(KMB200_ERR_COMM, "ncclAllReduce failed: %s", g_nccl.GetErrorString(r));
        }
    }
    CU_TRY(cudaMemcpyAsync(sum_costs, total, sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    return 0;
}

// The sampling step of initKMeansParallel (KMeans.scala:406-410): partition p of this table (rows part_rows[p] ..
// part_rows[p + 1], Spark partition index part_index0 + p) draws one nextDouble per row, in row order, from
// XORShiftRandom(seed ^ (step << 16) ^ index), and keeps the rows with u < 2.0 * cost * k / sumCosts.  The costs come
// back from the device in pinned chunks; the generator is sequential by definition and runs on the host.
extern "C" int kmb200_init_sample(kmb200_table *t, int32_t seed, int32_t step, int32_t k, double sum_costs,
                                  const int64_t *part_rows, int32_t nparts, int32_t part_index0, int64_t *rows_out,
                                  int64_t cap, int64_t *count_out) {
    if (!t || !part_rows || !count_out || nparts < 1 || (cap > 0 && !rows_out))
        return fail(KMB200_ERR_ARG, "init_sample: bad argument");
    if (!t->init_cost) return fail(KMB200_ERR_ARG, "init_sample: kmb200_init_costs has not run on this table");
    if (part_rows[0] != 0 || part_rows[nparts] != t->rows)
        return fail(KMB200_ERR_ARG, "init_sample: the partitions must cover the table's rows");
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    const int64_t chunk = 1 << 22;
    double *host = nullptr;
    CU_TRY(cudaMallocHost(reinterpret_cast<void **>(&host), (size_t)chunk * sizeof(double)));
    int64_t count = 0;
    int rc = 0;
    for (int p = 0; p < nparts && rc == 0; ++p) {
        if (part_rows[p + 1] < part_rows[p]) { rc = fail(KMB200_ERR_ARG, "init_sample: partition offsets decrease"); break; }
        XorShiftRandom rand((int64_t)(seed ^ (step << 16) ^ (part_index0 + p)));   // Int arithmetic, widened like Scala does
        for (int64_t r0 = part_rows[p]; r0 < part_rows[p + 1]; r0 += chunk) {
            const int64_t h = std::min(chunk, part_rows[p + 1] - r0);
            cudaError_t e = cudaMemcpyAsync(host, t->init_cost + r0, (size_t)h * sizeof(double), cudaMemcpyDeviceToHost, t->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
            if (e != cudaSuccess) { rc = fail(KMB200_ERR_DEVICE, "init_sample: %s", cudaGetErrorString(e)); break; }
            for (int64_t i = 0; i < h; ++i) {
                if (rand.next_double() < 2.0 * host[i] * k / sum_costs) {
                    if (count < cap) rows_out[count] = r0 + i;
                    ++count;
                }
            }
        }
    }
    cudaFreeHost(host);
    *count_out = count;
    return rc;
}

#include "pca_host.inl"

// Test hook: the raw tensor-core scores ||c_j||^2 - 2 x.c_j (FP32) of every row against every padded
// centre column, so tests can measure the 3xTF32 error against FP64 and validate the near-tie bound.
extern "C" int kmb200_debug_scores(kmb200_table *t, const double *centers, int32_t k, int32_t d, float *scores_host,
                                   int32_t *kpad_out) {
    KM_TRY(check_fit_args(t, centers, k, d));
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    KM_TRY(require_device());
    KM_TRY(ws_prepare(t, k));
    Workspace &w = t->ws;
    if (kpad_out) *kpad_out = w.kpad;
    if (!scores_host) return 0;
    float *dev = nullptr;
    CU_TRY(cudaMalloc(&dev, (size_t)t->rows * w.kpad * sizeof(float)));
    int rc = upload_centers(t, centers, k, d);
    if (!rc) rc = launch_split(t);
    const int simt = g_force_simt;
    g_force_simt = 0;
    w.dbg_scores = dev;
    if (!rc) rc = launch_assign(t);
    w.dbg_scores = nullptr;
    g_force_simt = simt;
    cudaError_t e = cudaMemcpyAsync(scores_host, dev, (size_t)t->rows * w.kpad * sizeof(float), cudaMemcpyDeviceToHost, t->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
    cudaFree(dev);
    if (rc) return rc;
    if (e != cudaSuccess) return fail(KMB200_ERR_DEVICE, "debug_scores: %s", cudaGetErrorString(e));
    return 0;
}

// Diagnostics: copies one of the workspace arrays of the last fit / assign to the host (tools/repro_check.py).
extern "C" int kmb200_debug_dump(kmb200_table *t, const char *name, size_t offset, void *out, size_t bytes) {
    if (!t || !name || !out) return fail(KMB200_ERR_ARG, "debug_dump: null argument");
    Workspace &w = t->ws;
    const void *src = nullptr;
    if (!strcmp(name, "cand_list")) src = w.cand_list;
    else if (!strcmp(name, "labels_c")) src = w.labels_c;
    else if (!strcmp(name, "cand1_list")) src = w.cand1_list;
    else if (!strcmp(name, "labels_c1")) src = w.labels_c1;
    else if (!strcmp(name, "flag_list")) src = w.flag_list;
    else if (!strcmp(name, "flag_count")) src = w.flag_count;
    else if (!strcmp(name, "xc")) src = w.xc;
    else if (!strcmp(name, "xc1")) src = w.xc1;
    else if (!strcmp(name, "labels")) src = w.labels;
    if (!src) return fail(KMB200_ERR_ARG, "debug_dump: no array '%s'", name);
    CU_TRY(cudaSetDevice(t->device));
    CU_TRY(cudaMemcpy(out, static_cast<const uint8_t *>(src) + offset, bytes, cudaMemcpyDeviceToHost));
    return 0;
}

// ------------------------------------------------------------------------------------ no-allocation entry
// Cluster sizes from a label vector without touching the device allocator.  The label histogram
// needs k int32 counters of scratch; the regular path (launch_bucket) keeps them in the workspace
// that ws_prepare cudaMallocs.  Some hosts of this library must not allocate on the device at call
// time: an executor whose GPU memory is handed, whole, to a pool allocator (e.g. an RMM pool shared
// with other Spark plugins) gets out-of-memory from cudaMalloc although the pool has room, and a
// latency-bound serving loop cannot afford the device-wide synchronisation cudaMalloc / cudaFree
// imply.  For them the counters live in the module itself: a fixed array that every GPU loading the
// library carries until the process exits (256 KiB per GPU).
constexpr int kNoAllocMaxClusters = 65536;
static __device__ __align__(256) int32_t g_noalloc_counters[kNoAllocMaxClusters];
// g_noalloc_counters is only a host-side handle here: the device address is looked up once per
// device (with that device current) and remembered.
static int32_t *g_noalloc_addr[64] = {nullptr};
static std::mutex g_noalloc_mu;

extern "C" int kmb200_cluster_sizes_noalloc(const int32_t *labels_dev, int64_t n, int32_t k, int64_t *sizes_dev,
                                            void *cuda_stream) {
    if (!labels_dev && n > 0) return fail(KMB200_ERR_ARG, "cluster_sizes_noalloc: labels_dev is null");
    if (!sizes_dev) return fail(KMB200_ERR_ARG, "cluster_sizes_noalloc: sizes_dev is null");
    if (n < 0 || k <= 0) return fail(KMB200_ERR_ARG, "cluster_sizes_noalloc: n must be >= 0 and k > 0");
    if (n > 2147483647LL)   // the built-in counters are 32-bit: one cluster could hold every row
        return fail(KMB200_ERR_LIMIT, "cluster_sizes_noalloc: n = %lld is beyond the 2^31-1 rows the 32-bit scratch counters can hold", (long long)n);
    if (k > kNoAllocMaxClusters)
        return fail(KMB200_ERR_LIMIT,
                    "cluster_sizes_noalloc: k = %d is larger than the built-in scratch of %d counters; use "
                    "kmb200_assign, which allocates its own",
                    k, kNoAllocMaxClusters);
    KM_TRY(require_device());
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(KMB200_ERR_LIMIT, "cluster_sizes_noalloc: device ordinal %d not supported", dev);
    int32_t *counters = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_addr[dev]) {
            void *addr = nullptr;
            CU_TRY(cudaGetSymbolAddress(&addr, g_noalloc_counters));  // this device's copy of the array
            g_noalloc_addr[dev] = static_cast<int32_t *>(addr);
        }
        counters = g_noalloc_addr[dev];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // the array still holds the previous call's histogram: the k counters used now must start at zero
    CU_TRY(cudaMemsetAsync(counters, 0, (size_t)k * sizeof(int32_t), st));
    if (n > 0) {
        const int64_t blocks = (n + kBucketRows - 1) / kBucketRows;
        if (blocks > 2147483647LL) return fail(KMB200_ERR_LIMIT, "cluster_sizes_noalloc: too many rows");
        const size_t sh = k <= kHistSmemMax ? (size_t)k * sizeof(int32_t) : 0;
        label_hist_kernel<<<(unsigned)blocks, 256, sh, st>>>(labels_dev, n, k, counters);
        KM_TRY(post_launch(__LINE__));
    }
    counts_to_i64_kernel<<<(k + 255) / 256, 256, 0, st>>>(counters, k, reinterpret_cast<unsigned long long *>(sizes_dev));
    KM_TRY(post_launch(__LINE__));
    return 0;
}

// ------------------------------------------------------------------------------------ misc
extern "C" int kmb200_set_option(const char *name, double value) {
    if (!name) return fail(KMB200_ERR_ARG, "set_option: null name");
    if (!strcmp(name, "eps_scale")) {
        if (!(value > 0)) return fail(KMB200_ERR_ARG, "eps_scale must be > 0");
        g_eps_scale = value;
    } else if (!strcmp(name, "passes")) {
        if (value != 1 && value != 3) return fail(KMB200_ERR_ARG, "passes must be 1 or 3");
        g_passes = (int)value;
    } else if (!strcmp(name, "force_simt")) {
        g_force_simt = value != 0;
    } else if (!strcmp(name, "role_prof")) {
        // diagnostics: value != 0 allocates the counter buffer; value == 0 prints and frees it
        if (value != 0 && !g_role_prof) {
            CU_TRY(cudaMalloc(&g_role_prof, 16 * sizeof(long long)));
            CU_TRY(cudaMemset(g_role_prof, 0, 16 * sizeof(long long)));
        } else if (value == 0 && g_role_prof) {
            long long h[16];
            CU_TRY(cudaDeviceSynchronize());
            CU_TRY(cudaMemcpy(h, g_role_prof, sizeof(h), cudaMemcpyDeviceToHost));
            fprintf(stderr, "[kmb200 role_prof] last assign launch, CTA 0 MMA warp: units=%lld cycles/unit: wait_acc=%.1f wait_full=%.1f wait_a=%.1f issue=%.1f\n",
                    h[4], (double)h[0] / h[4], (double)h[1] / h[4], (double)h[2] / h[4], (double)h[3] / h[4]);
            if (h[13] + h[14] + h[15] > 0)
                fprintf(stderr, "[kmb200 role_prof] a_full wait at the first centroid tile of a row tile, total cycles for units 1, 2, 3: %lld %lld %lld (tiles %lld)\n",
                        h[13], h[14], h[15], h[4]);
            for (int c = 0; c < 2; ++c)
                if (h[5 + 4 * c + 3] > 0)
                    fprintf(stderr, "[kmb200 role_prof] CTA %d transform warp: chunks=%lld cycles/chunk: load+split=%.1f wait_a_empty=%.1f store+arrive=%.1f\n",
                            c, h[8 + 4 * c], (double)h[5 + 4 * c] / h[8 + 4 * c], (double)h[6 + 4 * c] / h[8 + 4 * c],
                            (double)h[7 + 4 * c] / h[8 + 4 * c]);
            cudaFree(g_role_prof);
            g_role_prof = nullptr;
        }
    } else if (!strcmp(name, "kernel")) {
        if (value != 1 && value != 2) return fail(KMB200_ERR_ARG, "kernel must be 1 or 2");
        env_overrides();
        g_kernel = (int)value;
    } else if (!strcmp(name, "fused")) {
        g_fused = value == 2 ? 2 : value != 0;   // 2 = the CUDA-core kernel also where the MMA variant applies (tests)
    } else if (!strcmp(name, "stage_mb")) {
        if (value < 1.0 / 1024 || value > 4096) return fail(KMB200_ERR_ARG, "stage_mb must be in [1/1024, 4096]");
        g_stage_bytes = (size_t)(value * 1048576.0);
    } else if (!strcmp(name, "dbg_products")) {
        g_dbg_products = value == 1 ? 1 : 3;
    } else if (!strcmp(name, "cascade")) {
        g_cascade = value == 2 ? 2 : value != 0;   // 2 = also where the shape makes it unprofitable (tests)
    } else if (!strcmp(name, "cand1_frac")) {
        if (!(value > 0.0 && value <= 1.0)) return fail(KMB200_ERR_ARG, "cand1_frac must be in (0, 1]");
        g_cand1_frac = value;
    } else if (!strcmp(name, "two_level")) {
        g_two_level = (int)value;
    } else if (!strcmp(name, "cand_cap")) {
        g_cand_cap = (int64_t)value;
    } else if (!strcmp(name, "h16")) {
        g_h16 = value != 0;
    } else if (!strcmp(name, "rb_cap")) {
        g_rb_cap = (int)value;
    } else if (!strcmp(name, "cluster")) {
        if (value != 1 && value != 2) return fail(KMB200_ERR_ARG, "cluster must be 1 or 2");
        env_overrides();
        g_cluster = (int)value;
    } else if (!strcmp(name, "knockout")) {
        g_knockout = (int)value;  // timing experiments only: results are invalid when non-zero
    } else if (!strcmp(name, "pca_flush_rows")) {
        if (!(value >= 32 && value <= 1048576)) return fail(KMB200_ERR_ARG, "pca_flush_rows must be in [32, 2^20]");
        g_pca_flush_rows = (int)value;
    } else if (!strcmp(name, "pca_promote")) {
        if (!(value >= 1 && value <= 4096)) return fail(KMB200_ERR_ARG, "pca_promote must be in [1, 4096]");
        g_pca_promote = (int)value;
    } else if (!strcmp(name, "pca_ls") || !strcmp(name, "pca_hs")) {
        const int v = (int)value;
        if (v < 0 || v > kGramMaxRing || (v & 1)) return fail(KMB200_ERR_ARG, "%s must be even and at most %d", name, kGramMaxRing);
        (name[4] == 'l' ? g_pca_ls : g_pca_hs) = v;
    } else if (!strcmp(name, "pca_role_prof")) {
        g_pca_role_prof = value != 0;
    } else if (!strcmp(name, "pca_split")) {
        g_pca_split = value < 0 ? 0 : value > 2 ? 2 : (int)value;
    } else if (!strcmp(name, "pca_rows")) {
        g_pca_rows = value == 32 ? 32 : 64;
    } else if (!strcmp(name, "pca_knockout")) {
        g_pca_knockout = (int)value;
    } else if (!strcmp(name, "pca_desc_swap")) {
        g_pca_desc_swap = value != 0;
    } else {
        return fail(KMB200_ERR_ARG, "unknown option '%s'", name);
    }
    return 0;
}

extern "C" const char *kmb200_last_error(void) { return g_err.c_str(); }
extern "C" const char *kmb200_version(void) { return "kmeans_b200 0.2 (sm_100a; tcgen05 cta_group::2 FP16-split assign with FP64 re-check, FP64 accumulate)"; }
