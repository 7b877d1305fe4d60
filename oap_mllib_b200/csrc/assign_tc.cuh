// Assignment step on the 5th-gen tensor cores: for every row x of the table,
//     label(x) = argmin_j ( ||c_j||^2 - 2 x . c_j ),   lowest j among equal minima,
// the step oneDAL's kmeans::Distributed<step1Local,double> performs for the reference
// (mllib-dal/src/main/native/KMeansDALImpl.cpp:70-77).
//
// Shape of the kernel (one persistent CTA per SM, 384 threads, warp-specialised):
//   warp 0      TMA producer   X chunk [128 rows x 32 floats] + B_hi/B_lo chunk [BN x 32] per stage,
//                              128-byte swizzled, multi-stage mbarrier ring
//   warp 1      MMA issuer     one lane issues tcgen05.mma.kind::tf32 (M=128, N=BN, K=8), A from
//                              TMEM, B from shared memory, FP32 accumulators in TMEM (2 buffers)
//   warp 2      TMEM allocator
//   warps 4-7   transform      thread r owns row r: reads its 32 floats of the X chunk from shared
//                              memory, splits x = hi + lo (round-to-nearest TF32 each) and writes both
//                              halves to TMEM as the A operand (2 buffers), so the 3xTF32 product
//                              x.c ~= lo.c_hi + hi.c_lo + hi.c_hi keeps FP32-level accuracy
//   warps 8-11  epilogue       thread r owns row r: tcgen05.ld of the accumulator tile, + ||c_j||^2,
//                              running (best, second best, argmin) across centroid tiles, label
//                              store, and a near-tie flag when second-best - best is inside the
//                              error bound (those rows are re-evaluated in FP64 by recheck_kernel)
// B holds -2*c split into TF32 hi/lo, precomputed per iteration (finalize/split kernel).
#pragma once
#include "ptx_sm100.cuh"

namespace kmb {

constexpr int kTileM = 128;         // rows per CTA tile == TMEM lanes
constexpr int kChunkK = 32;         // floats per K chunk == one 128-byte swizzle row
constexpr int kXStageBytes = kTileM * kChunkK * 4;
constexpr int kMaxStages = 8;
constexpr int kAssignThreads = 384;

template <int BN>
struct AssignCfg {
    static constexpr int kBStageBytes = BN * kChunkK * 4;
    static constexpr int kStageBytes = kXStageBytes + 2 * kBStageBytes;
    static constexpr int kACols = 128;  // 2 buffers x (32 hi + 32 lo)
    static constexpr int kAccCol0 = kACols;
    static constexpr int kTmemColsUsed = kACols + 2 * BN;
    static constexpr int kTmemCols = kTmemColsUsed <= 256 ? 256 : 512;
    static constexpr int kTailBytes = 2 * BN * 4 + (2 * kMaxStages + 8) * 8 + 16;
    static int stages() {
        int s = (227 * 1024 - 1024 - kTailBytes) / kStageBytes;
        return s > kMaxStages ? kMaxStages : s;
    }
    static int smem_bytes(int stages) { return 1024 + stages * kStageBytes + kTailBytes; }
};

struct AssignParams {
    const float *cnorm;    // [kpad]  ||c_j||^2 (FP32 of the FP64 norm), +inf for padding columns
    const float *xnorm;    // [n]     ||x_i||^2
    const float *cstats;   // [0] = max_j ||c_j||
    int32_t *labels;       // [n]
    int32_t *flag_count;   // [0] (row, runner-up) entries at the front of flag_list, [1] full-scan entries at its back
    int2 *flag_list;       // [flag_cap] (row, runner-up centre) or (row, -1) when a third centre is inside the bound too
    int32_t flag_cap;
    int32_t n;             // rows in the table
    int32_t num_row_tiles; // ceil(n / 128)
    int32_t num_col_tiles; // kpad / BN
    int32_t num_k_chunks;  // ld / 32
    int32_t stages;
    int32_t passes;        // 3 = 3xTF32, 1 = single TF32 pass
    float eps;             // relative error bound of one score
    uint32_t keep_mask = 0xffffff00u;   // (see the one-product epilogue of assign_pair_h16.cuh)
    float eps_x = 0.f;     // FP16-split kernel, one-product pass: error relative to 2 |x| |c| alone (operand roundings)
    float *dbg_scores;     // optional [n x kpad] dump of the tensor-core scores (tests only), else nullptr
    int32_t kpad;
    long long *role_prof;  // optional cycle counters of the MMA-issuer warp of CTA 0 (diagnostics), else nullptr
    int32_t knockout;      // debugging only: 1 = no TMA, 2 = no transform work, 4 = no epilogue work (results invalid)
    float sqrt_ld;         // sqrt(ld) (FP16-split kernel: absolute part of the near-tie bound)
    // FP16-split kernel, two-level bookkeeping (assign_pair_h16.cuh)
    const int32_t *n_dev = nullptr;   // optional: row count in device memory (min with n)
    int32_t *cand_list = nullptr;     // FAST mode: rows whose best / runner-up gap is inside the bound
    int32_t *cand_count = nullptr;
    int32_t cand_cap = 0;
};

// Near-tie bookkeeping shared by the assign kernels: rows whose winner must be decided in FP64.
// (row, runner-up) entries grow from the front of flag_list (flag_count[0]); rows with a third centre
// inside the bound need all k centres re-scored and are listed from the BACK (flag_count[1]), so the
// two kinds can be processed by differently shaped kernels.
__device__ __forceinline__ void flag_row(const AssignParams &p, int row, int winner, int runner, bool need_full) {
    if (need_full) {
        const int pos = atomicAdd(p.flag_count + 1, 1);
        if (pos < p.flag_cap) p.flag_list[p.flag_cap - 1 - pos] = make_int2(row, -1);
    } else {
        const int pos = atomicAdd(p.flag_count, 1);
        // both centre indices fit 16 bits (k <= 65536): the re-check then needs no dependent label load
        if (pos < p.flag_cap) p.flag_list[pos] = make_int2(row, (winner << 16) | runner);
    }
}

template <int BN>
__global__ void __launch_bounds__(kAssignThreads, 1)
assign_tc_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_bhi,
                 const __grid_constant__ CUtensorMap tm_blo, const AssignParams p) {
    using Cfg = AssignCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);  // swizzle-128B needs 1024-byte alignment

    const int S = p.stages;
    uint8_t *tail = smem + S * Cfg::kStageBytes;
    float *cn_s = reinterpret_cast<float *>(tail);  // [2][BN]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail + 2 * BN * 4);
    uint64_t *full = bars;
    uint64_t *empty = bars + kMaxStages;
    uint64_t *a_full = bars + 2 * kMaxStages;
    uint64_t *a_empty = a_full + 2;
    uint64_t *acc_full = a_full + 4;
    uint64_t *acc_empty = a_full + 6;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kMaxStages + 8);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_x);
        tma_prefetch_desc(&tm_bhi);
        tma_prefetch_desc(&tm_blo);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&a_full[b], 128);
            mbar_init(&a_empty[b], 1);
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int NT = p.num_col_tiles;
    const int KC = p.num_k_chunks;

    if (warp == 0) {
        // ================================ TMA producer ================================
        // The whole warp walks the loop (converged), one elected lane issues: this keeps the
        // uniform-datapath instructions (UTMALDG) out of divergent code.
        uint32_t s = 0, ph = 0;
        const uint32_t stage_tx = kXStageBytes + (p.passes == 3 ? 2 : 1) * Cfg::kBStageBytes;
        for (int rt = blockIdx.x; rt < p.num_row_tiles; rt += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                for (int kc = 0; kc < KC; ++kc) {
                    mbar_wait(&empty[s], ph ^ 1);
                    if (elect_one()) {
                        uint8_t *st = smem + s * Cfg::kStageBytes;
                        if (p.knockout & 1) {
                            mbar_arrive(&full[s]);
                        } else {
                            mbar_arrive_expect_tx(&full[s], stage_tx);
                            tma_load_2d(&tm_x, &full[s], st, kc * kChunkK, rt * kTileM);
                            tma_load_2d(&tm_bhi, &full[s], st + kXStageBytes, kc * kChunkK, nt * BN);
                            if (p.passes == 3)
                                tma_load_2d(&tm_blo, &full[s], st + kXStageBytes + Cfg::kBStageBytes, kc * kChunkK,
                                            nt * BN);
                        }
                    }
                    __syncwarp();
                    if (++s == (uint32_t)S) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        constexpr uint32_t idesc = make_idesc_tf32(kTileM, BN);
        uint32_t it = 0, tile = 0, s = 0, ph = 0;
        long long t_wait_acc = 0, t_wait_full = 0, t_wait_a = 0, t_issue = 0, t_mark = 0;
        const bool prof = p.role_prof != nullptr && blockIdx.x == 0;
        for (int rt = blockIdx.x; rt < p.num_row_tiles; rt += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt, ++tile) {
                const uint32_t ab = tile & 1, aph = (tile >> 1) & 1;
                if (prof) t_mark = clock64();
                mbar_wait(&acc_empty[ab], aph ^ 1);  // epilogue has drained this accumulator
                if (prof) { const long long t = clock64(); t_wait_acc += t - t_mark; t_mark = t; }
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + Cfg::kAccCol0 + ab * BN;
                for (int kc = 0; kc < KC; ++kc, ++it) {
                    const uint32_t xb = it & 1, xph = (it >> 1) & 1;
                    mbar_wait(&full[s], ph);     // B chunk landed in shared memory
                    if (prof) { const long long t = clock64(); t_wait_full += t - t_mark; t_mark = t; }
                    mbar_wait(&a_full[xb], xph); // transform wrote the A chunk to TMEM
                    if (prof) { const long long t = clock64(); t_wait_a += t - t_mark; t_mark = t; }
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t a_hi = tmem_base + xb * 64;
                        const uint32_t a_lo = a_hi + 32;
                        const uint32_t bsm = smem_u32(smem + s * Cfg::kStageBytes + kXStageBytes);
                        const uint64_t dhi = make_kmajor_sw128_desc(bsm);
                        const uint64_t dlo = make_kmajor_sw128_desc(bsm + Cfg::kBStageBytes);
#pragma unroll
                        for (int ks = 0; ks < kChunkK / 8; ++ks) {
                            // advancing 8 TF32 values (32 bytes) inside the swizzle atom: +2 in the
                            // 16-byte-granular start-address field
                            const uint64_t koff = static_cast<uint64_t>(ks * 2);
                            const uint32_t first = (kc | ks) == 0 ? 0u : 1u;
                            if (p.passes == 3) {
                                mma_tf32_ts(d_tmem, a_lo + ks * 8, dhi + koff, idesc, first);  // small terms first
                                mma_tf32_ts(d_tmem, a_hi + ks * 8, dlo + koff, idesc, 1u);
                                mma_tf32_ts(d_tmem, a_hi + ks * 8, dhi + koff, idesc, 1u);
                            } else {
                                mma_tf32_ts(d_tmem, a_hi + ks * 8, dhi + koff, idesc, first);
                            }
                        }
                        tc_commit(&empty[s]);     // shared-memory stage may be refilled
                        tc_commit(&a_empty[xb]);  // TMEM A buffer may be rewritten
                        if (kc == KC - 1) tc_commit(&acc_full[ab]);  // accumulator tile complete -> epilogue
                    }
                    __syncwarp();
                    if (prof) { const long long t = clock64(); t_issue += t - t_mark; t_mark = t; }
                    if (++s == (uint32_t)S) { s = 0; ph ^= 1; }
                }
            }
        }
        if (prof && lane == 0) {
            p.role_prof[0] = t_wait_acc;
            p.role_prof[1] = t_wait_full;
            p.role_prof[2] = t_wait_a;
            p.role_prof[3] = t_issue;
            p.role_prof[4] = it;
        }
    } else if (warp >= 4 && warp < 8) {
        // ================================ transform ===================================
        const int r = threadIdx.x - 128;  // row inside the tile == TMEM lane
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        uint32_t it = 0, s = 0, ph = 0;
        for (int rt = blockIdx.x; rt < p.num_row_tiles; rt += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                for (int kc = 0; kc < KC; ++kc, ++it) {
                    const uint32_t xb = it & 1, xph = (it >> 1) & 1;
                    mbar_wait(&full[s], ph);
                    const uint8_t *xs = smem + s * Cfg::kStageBytes + r * 128;
                    if (++s == (uint32_t)S) { s = 0; ph ^= 1; }
                    if (p.knockout & 2) {
                        mbar_wait(&a_empty[xb], xph ^ 1);
                        mbar_arrive(&a_full[xb]);
                        continue;
                    }
                    float x[32];
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        const float4 v = *reinterpret_cast<const float4 *>(xs + ((g ^ (r & 7)) << 4));
                        x[4 * g + 0] = v.x;
                        x[4 * g + 1] = v.y;
                        x[4 * g + 2] = v.z;
                        x[4 * g + 3] = v.w;
                    }
                    uint32_t hi[32], lo[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float h = to_tf32_rna(x[i]);
                        hi[i] = __float_as_uint(h);
                        lo[i] = __float_as_uint(to_tf32_rna(x[i] - h));
                    }
                    mbar_wait(&a_empty[xb], xph ^ 1);  // MMAs that read this A buffer have completed
                    tc_fence_after();
                    const uint32_t a_addr = tmem_base + lane_addr + xb * 64;
                    tmem_st32(a_addr, hi);
                    if (p.passes == 3) tmem_st32(a_addr + 32, lo);
                    tmem_wait_st();
                    tc_fence_before();
                    mbar_arrive(&a_full[xb]);
                }
            }
        }
    } else if (warp >= 8) {
        // ================================ epilogue ====================================
        const int r = threadIdx.x - 256;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const float cmax = p.cstats[0];
        uint32_t tile = 0;
        for (int rt = blockIdx.x; rt < p.num_row_tiles; rt += gridDim.x) {
            float best = __int_as_float(0x7f800000), second = best, third = best;
            int bidx = 0, sidx = 0;
            for (int nt = 0; nt < NT; ++nt, ++tile) {
                const uint32_t ab = tile & 1, aph = (tile >> 1) & 1;
                float *cn = cn_s + ab * BN;
                if (r < BN) cn[r] = p.cnorm[nt * BN + r];
                asm volatile("bar.sync 1, 128;" ::: "memory");  // epilogue warps only
                mbar_wait(&acc_full[ab], aph);
                tc_fence_after();
                if (p.knockout & 4) {
                    mbar_arrive(&acc_empty[ab]);
                    continue;
                }
                const uint32_t acc = tmem_base + lane_addr + Cfg::kAccCol0 + ab * BN;
                constexpr int G = BN >= 32 ? 32 : 16;
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += G) {
                    uint32_t v[G];
                    if constexpr (G == 32) tmem_ld32(acc + c0, v);
                    else tmem_ld16(acc + c0, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < G; ++i) {
                        const float sc = __uint_as_float(v[i]) + cn[c0 + i];
                        if (p.dbg_scores != nullptr && rt * kTileM + r < p.n)
                            p.dbg_scores[(size_t)(rt * kTileM + r) * p.kpad + nt * BN + c0 + i] = sc;
                        // running top-3 values with the indices of the best two; strict '<' keeps
                        // the lowest index among equal scores
                        const int j = nt * BN + c0 + i;
                        const bool lt1 = sc < best, lt2 = sc < second;
                        third = fminf(third, fmaxf(sc, second));
                        sidx = lt1 ? bidx : (lt2 ? j : sidx);
                        second = fminf(second, fmaxf(sc, best));
                        bidx = lt1 ? j : bidx;
                        best = fminf(best, sc);
                    }
                }
                tc_fence_before();
                mbar_arrive(&acc_empty[ab]);
            }
            const int row = rt * kTileM + r;
            if (row < p.n) {
                p.labels[row] = bidx;
                const float xn = sqrtf(p.xnorm[row]);
                const float bound = 2.0f * p.eps * (2.0f * xn * cmax + cmax * cmax);
                if (second - best <= bound) flag_row(p, row, bidx, sidx, third - best <= bound);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

}  // namespace kmb
