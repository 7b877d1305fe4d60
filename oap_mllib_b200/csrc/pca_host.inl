// PCA entry points -- included by kmb200.cu, which provides fail / CU_TRY / KM_TRY,
// DevBuf, DeviceLock, make_map, the NCCL table and the profile.  Replaces, behind the same boundary, the body of
// Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDAL (mllib-dal/src/main/native/PCADALImpl.cpp:38-190):
//   step1Local on every rank   -> column statistics pass + gram_pair_kernel over the resident FP32 table
//   all-gather of the partials -> one NCCL all-reduce of the d x d FP64 Gram matrix (+ column sums, row count)
//   step2Master on rank 0      -> sym_eig_desc on the host (every rank gets the same answer; the JNI shim hands it to
//                                 the JVM on rank 0 only, like the reference)
#include "gram_pair.cuh"
#include "sym_eig.h"
#include <chrono>

static int g_pca_flush_rows = 128;   // rows a level-1 (tensor core, FP32) accumulator holds ("pca_flush_rows")
static int g_pca_promote = 16;        // level-1 segments per promotion of the level-2 sums to FP64 ("pca_promote")
static int g_pca_desc_swap = 0;       // debugging ("pca_desc_swap")
static int g_pca_knockout = 0;        // timing experiments ("pca_knockout"): results are invalid when non-zero

// Centred (center != 0) or raw Gram matrix of the whole (distributed) table: G [d x d] row-major on the host, the column
// means used (0 when center == 0) and the global row count.  Collective.
static int pca_gram(kmb200_table *t, kmb200_comm *comm, int center, std::vector<double> *G_out, std::vector<double> *mean_out,
                    int64_t *n_out) {
    const int ld = t->ld, d = t->d;
    const int ld2 = (ld + 255) / 256 * 256;
    const int64_t n = t->rows;
    const int world = comm ? comm->world : 1;
    cudaStream_t st = t->stream;
    DevBuf buf;
    double *psum = nullptr, *gsum = nullptr, *inv = nullptr, *mean_used = nullptr, *partial = nullptr, *G = nullptr;
    float *pmax = nullptr, *gmax = nullptr, *scale = nullptr, *offs = nullptr;
    const int sblocks = (int)std::min<int64_t>((int64_t)t->num_sms * 4, std::max<int64_t>(1, n / 64));
    KM_TRY(buf.alloc(&psum, (size_t)sblocks * ld));
    KM_TRY(buf.alloc(&pmax, (size_t)sblocks * ld));
    KM_TRY(buf.alloc(&gsum, (size_t)ld + 1));
    KM_TRY(buf.alloc(&gmax, (size_t)ld));
    KM_TRY(buf.alloc(&scale, (size_t)ld2));
    KM_TRY(buf.alloc(&offs, (size_t)ld2));
    KM_TRY(buf.alloc(&inv, (size_t)ld));
    KM_TRY(buf.alloc(&mean_used, (size_t)ld));
    KM_TRY(buf.alloc(&G, (size_t)ld * ld));
    struct Events {   // device-side phase times for the profile
        cudaEvent_t e[5] = {};
        ~Events() { for (auto x : e) if (x) cudaEventDestroy(x); }
    } ev;
    for (auto &x : ev.e) CU_TRY(cudaEventCreate(&x));
    CU_TRY(cudaEventRecord(ev.e[0], st));
    // ---- pass 1: column sums and maxima
    colstats_kernel<<<sblocks, 256, 0, st>>>(t->X, n, ld, psum, pmax);
    KM_TRY(post_launch(__LINE__));
    colstats_reduce_kernel<<<(ld + 7) / 8, 256, 0, st>>>(psum, pmax, sblocks, ld, n, gsum, gmax);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches += 2;
    if (world > 1) {
        ncclResult_t r = g_nccl.AllReduce(gsum, gsum, (size_t)ld + 1, ncclDouble, ncclSum, comm->comm, st);
        if (r == ncclSuccess) r = g_nccl.AllReduce(gmax, gmax, (size_t)ld, ncclFloat, ncclMax, comm->comm, st);
        if (r != ncclSuccess) {
            comm->broken = true;
            return fail(KMB200_ERR_COMM, "ncclAllReduce failed: %s", g_nccl.GetErrorString(r));
        }
    }
    pca_prep_kernel<<<(ld2 + 255) / 256, 256, 0, st>>>(gsum, gmax, ld, ld2, center, scale, offs, inv, mean_used);
    KM_TRY(post_launch(__LINE__));
    g_prof.kernel_launches++;
    CU_TRY(cudaEventRecord(ev.e[1], st));
    // ---- pass 2: the Gram matrix, one launch per 256-column panel pair on or above the diagonal
    const int stages_total = (int)((n + kGramRows - 1) / kGramRows);
    const int pairs = std::max(1, std::min(stages_total, t->num_sms / 2));
    KM_TRY(buf.alloc(&partial, (size_t)pairs * 65536));
    CUtensorMap tm;
    KM_TRY(make_map(&tm, t->X, (uint64_t)t->rows, (uint64_t)ld, kGramRows));
    static bool attr_set[64][2] = {};
    for (int same = 0; same < 2; ++same) {
        if (attr_set[t->device & 63][same]) continue;
        if (same) CU_TRY(cudaFuncSetAttribute(gram_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, gram_smem_bytes(true)));
        else CU_TRY(cudaFuncSetAttribute(gram_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, gram_smem_bytes(false)));
        attr_set[t->device & 63][same] = true;
    }
    GramParams p;
    p.scale = scale;
    p.offs = offs;
    p.partial = partial;
    p.n_rows = (int32_t)n;
    p.stages_total = stages_total;
    p.flush_stages = std::max(1, g_pca_flush_rows / kGramRows);
    p.promote_segs = std::max(1, g_pca_promote);
    p.desc_swap = g_pca_desc_swap;
    p.knockout = g_pca_knockout;
    const int panels = ld2 / 256;
    for (int I = 0; I < panels; ++I) {
        for (int J = I; J < panels; ++J) {
            CU_TRY(cudaMemsetAsync(partial, 0, (size_t)pairs * 65536 * sizeof(double), st));
            p.colA0 = I * 256;
            p.colB0 = J * 256;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)(pairs * 2));
            cfg.blockDim = dim3(kGramThreads);
            cfg.dynamicSmemBytes = (size_t)gram_smem_bytes(I == J);
            cfg.stream = st;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            if (I == J) CU_TRY(cudaLaunchKernelEx(&cfg, gram_pair_kernel<true>, tm, p));
            else CU_TRY(cudaLaunchKernelEx(&cfg, gram_pair_kernel<false>, tm, p));
            KM_TRY(post_launch(__LINE__));
            gram_reduce_kernel<<<dim3(1, 256), 256, 0, st>>>(partial, pairs, inv, ld, p.colA0, p.colB0, G);
            KM_TRY(post_launch(__LINE__));
            g_prof.kernel_launches += 2;
        }
    }
    CU_TRY(cudaEventRecord(ev.e[2], st));
    if (world > 1) {
        const ncclResult_t r = g_nccl.AllReduce(G, G, (size_t)ld * ld, ncclDouble, ncclSum, comm->comm, st);
        if (r != ncclSuccess) {
            comm->broken = true;
            return fail(KMB200_ERR_COMM, "ncclAllReduce failed: %s", g_nccl.GetErrorString(r));
        }
    }
    CU_TRY(cudaEventRecord(ev.e[3], st));
    std::vector<double> hG((size_t)ld * ld), hsum((size_t)ld + 1), hmean((size_t)ld);
    CU_TRY(cudaMemcpyAsync(hG.data(), G, hG.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(hsum.data(), gsum, hsum.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(hmean.data(), mean_used, hmean.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaEventRecord(ev.e[4], st));
    CU_TRY(cudaStreamSynchronize(st));
    CU_TRY(cudaEventElapsedTime(&g_prof.accumulate_ms, ev.e[0], ev.e[1]));   // column statistics (+ their all-reduce)
    CU_TRY(cudaEventElapsedTime(&g_prof.assign_ms, ev.e[1], ev.e[2]));       // Gram kernels + the reduction of their partials
    g_prof.assign_main_ms = g_prof.assign_ms;
    CU_TRY(cudaEventElapsedTime(&g_prof.allreduce_ms, ev.e[2], ev.e[3]));
    CU_TRY(cudaEventElapsedTime(&g_prof.total_ms, ev.e[0], ev.e[4]));
    g_prof.iterations = 1;
    const double ntot = hsum[ld];
    G_out->assign((size_t)d * d, 0.0);
    mean_out->assign((size_t)d, 0.0);
    // The kernel centred with the FP32 rounding of the mean; the exact mean of what it summed is
    // ybar = sum / N - mu_f, and sum (y - ybar)(y - ybar)^T = sum y y^T - N ybar ybar^T.
    std::vector<double> ybar((size_t)d, 0.0);
    for (int c = 0; c < d; ++c) {
        if (center) {
            ybar[c] = hsum[c] / ntot - hmean[c];
            (*mean_out)[c] = hsum[c] / ntot;
        }
    }
    for (int i = 0; i < d; ++i)
        for (int j = 0; j < d; ++j) (*G_out)[(size_t)i * d + j] = hG[(size_t)i * ld + j] - ntot * (ybar[i] * ybar[j]);
    *n_out = (int64_t)ntot;
    return 0;
}

static int pca_check(kmb200_table *t, kmb200_comm *comm, const char *what) {
    if (!t) return fail(KMB200_ERR_ARG, "%s: null table", what);
    if (t->rows < 1) return fail(KMB200_ERR_ARG, "%s: empty table", what);
    if (comm && comm->world > 1) KM_TRY(nccl_load());
    return 0;
}

extern "C" int kmb200_pca_gram(kmb200_table *t, kmb200_comm *comm, int32_t center, double *gram, double *mean, int64_t *n_total) {
    KM_TRY(pca_check(t, comm, "pca_gram"));
    if (!gram) return fail(KMB200_ERR_ARG, "pca_gram: null output");
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    KM_TRY(require_device());
    prof_reset();
    std::vector<double> G, mu;
    int64_t n = 0;
    KM_TRY(pca_gram(t, comm, center, &G, &mu, &n));
    memcpy(gram, G.data(), G.size() * sizeof(double));
    if (mean) memcpy(mean, mu.data(), mu.size() * sizeof(double));
    if (n_total) *n_total = n;
    return 0;
}

// Eigenvalues of G / (N - 1), descending, and the principal axes as ROWS (entry of largest magnitude positive): the two
// result tables of the reference (PCADALImpl.cpp:150-186: pca::eigenvalues 1 x d, pca::eigenvectors d x d).
extern "C" int kmb200_pca_train(kmb200_table *t, kmb200_comm *comm, int32_t center, double *eigenvalues, double *eigenvectors) {
    KM_TRY(pca_check(t, comm, "pca_train"));
    if (!eigenvalues || !eigenvectors) return fail(KMB200_ERR_ARG, "pca_train: null output");
    DeviceLock device_lock(t->device);
    CU_TRY(cudaSetDevice(t->device));
    KM_TRY(require_device());
    prof_reset();
    std::vector<double> G, mu;
    int64_t n = 0;
    KM_TRY(pca_gram(t, comm, center, &G, &mu, &n));
    const int d = t->d;
    const auto h0 = std::chrono::steady_clock::now();
    if (!sym_eig_desc(G.data(), d, eigenvalues, eigenvectors)) return fail(KMB200_ERR_DEVICE, "pca_train: the QL iteration did not converge");
    g_prof.finalize_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - h0).count();   // host eigen-solver
    const double denom = n > 1 ? (double)(n - 1) : 1.0;
    for (int i = 0; i < d; ++i) {
        eigenvalues[i] = std::max(eigenvalues[i], 0.0) / denom;
        double *v = eigenvectors + (size_t)i * d;
        int big = 0;
        for (int c = 1; c < d; ++c)
            if (std::fabs(v[c]) > std::fabs(v[big])) big = c;
        if (v[big] < 0.0)
            for (int c = 0; c < d; ++c) v[c] = -v[c];
    }
    return 0;
}

// Host only (no device needed): the eigen-solver by itself, for the CPU tests.
extern "C" int kmb200_sym_eig(const double *a, int32_t n, double *eigenvalues, double *eigenvectors) {
    if (!a || !eigenvalues || !eigenvectors || n < 1) return fail(KMB200_ERR_ARG, "sym_eig: bad argument");
    if (!sym_eig_desc(a, n, eigenvalues, eigenvectors)) return fail(KMB200_ERR_DEVICE, "sym_eig: the QL iteration did not converge");
    return 0;
}
