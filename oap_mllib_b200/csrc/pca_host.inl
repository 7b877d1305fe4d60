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

static int g_pca_flush_rows = 256;   // rows a level-1 (tensor core, FP32) accumulator holds ("pca_flush_rows")
static int g_pca_promote = 16;        // level-1 segments per promotion of the level-2 sums to FP64 ("pca_promote")
static int g_pca_desc_swap = 0;       // debugging ("pca_desc_swap")
static int g_pca_split = 2;           // narrow tables: 0 = off, 1 = the CTAs of a pair take different stages (d <= 128), 2 = and two
                                      // stages per CTA for d <= 64 ("pca_split")
static int g_pca_rows = 64;           // rows per stage on the diagonal panels, 64 or 32 ("pca_rows")
static int g_pca_ls = 0, g_pca_hs = 0;   // ring depths ("pca_ls", "pca_hs"; 0 = default, even, <= 8)
static int g_pca_role_prof = 0;       // diagnostics ("pca_role_prof"): print pair 0's role cycle counters to stderr
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
    const int Rd = g_pca_rows == 32 ? 32 : 64;   // rows per stage on the diagonal panels
    const int split = !g_pca_split ? 0 : ld <= 64 ? std::min(g_pca_split, 2) : ld <= 128 ? 1 : 0;   // stages per pair and step: 1 << split
    const int64_t min_stages = (n + (Rd << split) - 1) / (Rd << split);
    const int pairs = (int)std::max<int64_t>(1, std::min<int64_t>(min_stages, t->num_sms / 2));
    if (t->pca && t->pca->pairs < pairs) {   // (the table was longer when they were sized)
        delete t->pca;
        t->pca = nullptr;
    }
    if (!t->pca) {
        PcaWs *w = new PcaWs();
        w->pairs = pairs;
        auto allocate = [&]() -> int {
            CU_TRY(cudaMalloc(&w->gsum, ((size_t)ld + 1) * sizeof(double)));
            CU_TRY(cudaMalloc(&w->gmax, (size_t)ld * sizeof(float)));
            CU_TRY(cudaMalloc(&w->scale, (size_t)ld2 * sizeof(float)));
            CU_TRY(cudaMalloc(&w->offs, (size_t)ld2 * sizeof(float)));
            CU_TRY(cudaMalloc(&w->inv, (size_t)ld * sizeof(double)));
            CU_TRY(cudaMalloc(&w->mean_used, (size_t)ld * sizeof(double)));
            CU_TRY(cudaMalloc(&w->G, (size_t)ld * ld * sizeof(double)));
            CU_TRY(cudaMalloc(&w->partial, (size_t)pairs * 65536 * sizeof(double)));
            CU_TRY(cudaMalloc(&w->prof, 64 * sizeof(long long)));
            CU_TRY(cudaMallocHost(reinterpret_cast<void **>(&w->host), ((size_t)ld * ld + 2 * (size_t)ld + 1) * sizeof(double)));
            for (auto &x : w->ev) CU_TRY(cudaEventCreate(&x));
            return 0;
        };
        const int rc = allocate();
        if (rc != 0) {   // all or nothing
            delete w;
            return rc;
        }
        t->pca = w;
    }
    PcaWs &ws = *t->pca;
    double *gsum = ws.gsum, *inv = ws.inv, *mean_used = ws.mean_used, *partial = ws.partial, *G = ws.G;
    float *gmax = ws.gmax, *scale = ws.scale, *offs = ws.offs;
    struct { cudaEvent_t *e; } ev{ws.ev};
    CU_TRY(cudaEventRecord(ev.e[0], st));
    // ---- column sums and maxima: kept with the table by its appends; one pass over the rows when they are not current
    if (!(t->stats_ok && t->stats_rows == n)) {
        KM_TRY(table_stats_append(t, 0, n));
        g_prof.kernel_launches += 3;
    }
    CU_TRY(cudaMemcpyAsync(gsum, t->colsum, ((size_t)ld + 1) * sizeof(double), cudaMemcpyDeviceToDevice, st));
    CU_TRY(cudaMemcpyAsync(gmax, t->colmax, (size_t)ld * sizeof(float), cudaMemcpyDeviceToDevice, st));
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
    CUtensorMap tm32, tm64;
    KM_TRY(make_map(&tm32, t->X, (uint64_t)t->rows, (uint64_t)ld, 32));
    KM_TRY(make_map(&tm64, t->X, (uint64_t)t->rows, (uint64_t)ld, 64));
    static bool attr_set[64] = {};
    if (!attr_set[t->device & 63]) {
        CU_TRY(cudaFuncSetAttribute(gram_pair_kernel<true, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        CU_TRY(cudaFuncSetAttribute(gram_pair_kernel<true, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        CU_TRY(cudaFuncSetAttribute(gram_pair_kernel<false, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[t->device & 63] = true;
    }
    GramParams p;
    p.scale = scale;
    p.offs = offs;
    p.partial = partial;
    p.n_rows = (int32_t)n;
    p.promote_segs = std::max(1, g_pca_promote);
    p.desc_swap = g_pca_desc_swap;
    p.knockout = g_pca_knockout;
    p.prof = nullptr;
    long long *dprof = nullptr;
    if (g_pca_role_prof) {
        dprof = ws.prof;
        CU_TRY(cudaMemsetAsync(dprof, 0, 64 * sizeof(long long), st));
        p.prof = dprof;
    }
    const int panels = ld2 / 256;
    for (int I = 0; I < panels; ++I) {
        for (int J = I; J < panels; ++J) {
            CU_TRY(cudaMemsetAsync(partial, 0, (size_t)pairs * 65536 * sizeof(double), st));
            const int R = I == J ? Rd : 32;
            size_t cfg_smem = 0;
            p.colA0 = I * 256;
            p.colB0 = J * 256;
            p.split = split;
            p.stages_total = (int32_t)((n + ((int64_t)R << split) - 1) / ((int64_t)R << split));
            // (three products off the diagonal panels: 3/2 of the adds per row, so 2/3 of the rows per segment)
            p.flush_stages = std::max(1, (I == J ? g_pca_flush_rows : g_pca_flush_rows * 2 / 3) / R);
            int ls = 0, hs = 0;
            gram_default_rings(I == J, R, &ls, &hs);
            if (g_pca_ls > 0) ls = g_pca_ls;
            if (g_pca_hs > 0) hs = g_pca_hs;
            const int smem = gram_smem_bytes(I == J, R, ls, hs);
            if (smem > 227 * 1024) return fail(KMB200_ERR_ARG, "pca: landing / operand rings of %d / %d stages need %d bytes of shared memory", ls, hs, smem);
            p.ls = ls;
            p.hs = hs;
            cfg_smem = (size_t)smem;
            cudaLaunchConfig_t cfg = {};
            cfg.dynamicSmemBytes = cfg_smem;
            cfg.gridDim = dim3((unsigned)(pairs * 2));
            cfg.blockDim = dim3(kGramThreads);
            cfg.stream = st;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            if (I != J) {
                CU_TRY(cudaLaunchKernelEx(&cfg, gram_pair_kernel<false, 32>, tm32, p));
            } else if (R == 64) {
                CU_TRY(cudaLaunchKernelEx(&cfg, gram_pair_kernel<true, 64>, tm64, p));
            } else {
                CU_TRY(cudaLaunchKernelEx(&cfg, gram_pair_kernel<true, 32>, tm32, p));
            }
            KM_TRY(post_launch(__LINE__));
            if (dprof && I == 0 && J == 0) {
                long long h[64];
                CU_TRY(cudaMemcpyAsync(h, dprof, sizeof(h), cudaMemcpyDeviceToHost, st));
                CU_TRY(cudaStreamSynchronize(st));
                fprintf(stderr, "[gram roles, pair 0, cycles] producer total %lld wait-free-slot %lld | issuer0 total %lld wait-acc %lld wait-operands %lld | issuer1 total %lld wait-acc %lld wait-operands %lld | flush0 total %lld wait %lld | flush1 total %lld wait %lld\n",
                        h[0], h[1], h[4], h[5], h[6], h[8], h[9], h[10], h[12], h[13], h[14], h[15]);
                for (int c = 0; c < 2; ++c)
                    for (int g = 0; g < 2; ++g)
                        fprintf(stderr, "[gram roles] transform cta %d group %d: total %lld wait-landing %lld wait-operand-slot %lld fence+arrive %lld\n", c, g,
                                h[16 + 8 * c + 4 * g], h[17 + 8 * c + 4 * g], h[18 + 8 * c + 4 * g], h[19 + 8 * c + 4 * g]);
            }
            gram_reduce_kernel<<<dim3(1, 256), 256, 0, st>>>(partial, pairs, inv, ld, p.colA0, p.colB0, split, G);
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
    double *hG = ws.host, *hsum = hG + (size_t)ld * ld, *hmean = hsum + ld + 1;   // pinned
    CU_TRY(cudaMemcpyAsync(hG, G, (size_t)ld * ld * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(hsum, gsum, ((size_t)ld + 1) * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(hmean, mean_used, (size_t)ld * sizeof(double), cudaMemcpyDeviceToHost, st));
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
    const auto w0 = std::chrono::steady_clock::now();
    KM_TRY(pca_gram(t, comm, center, &G, &mu, &n));
    const int d = t->d;
    const auto h0 = std::chrono::steady_clock::now();
    if (getenv("KMB200_PCA_TRACE"))
        fprintf(stderr, "[pca trace] pca_gram wall %.2f ms (device %.2f ms)\n", std::chrono::duration<float, std::milli>(h0 - w0).count(), g_prof.total_ms);
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
