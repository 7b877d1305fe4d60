// CTA-pair assign kernel, "H16" variant (ld <= 256, BN >= 32): two-term FP16 splits of both operands.
//
//   x * sx = h1 + h2 (+ <= 2^-22 |x sx|),   -2c * sw = g1 + g2 (+ <= 2^-22 |2c sw|)      (sx, sw powers of two)
//   -2 x.c  ~=  (h2.g1 + h1.g2 + h1.g1) / (sx sw)        three kind::f16 products, FP32 accumulation in TMEM
//
// FP16 products are exact in FP32, the dropped h2.g2 term is 2^-22 of the main one, so the score error is
// the same size as the TF32 + BF16-correction scheme's (assign_pair_ss.cuh) at 3/4 of its tensor-pipe time:
// 3 x (ld/16) MMA instructions per score tile instead of 2 x (ld/8).  Two further differences:
//  * both A operands (h1, h2) of a whole 128 x ld row tile live in TMEM (2 x ld/2 columns) for ALL the
//    centroid tiles, so the transform warps touch each X element once per row tile, not once per centroid
//    tile; the X tile only passes through shared memory;
//  * the B ring is a plain full/empty ring of [g1 | g2] chunks of 64 K-values (128-byte rows).
// The power-of-two scales keep |h|, |g| < 2^15; values below the FP16 normal range (2^-29 of the largest
// element) lose relative precision, which the near-tie bound accounts for with an absolute term
// (cstats[4], cstats[5]); such rows are re-scored in FP64 like every other flagged row.
#pragma once
#include <cuda_fp16.h>

#include "assign_pair_ss.cuh"

namespace kmb {

constexpr int kPairH16Threads = 640;

template <int BN>
struct PairH16Cfg {
    static constexpr int kBRows = BN / 2;                // centroid rows held by one CTA
    static constexpr int kBHalfBytes = kBRows * 128;     // g1 (or g2) for one chunk of 64 K-values
    static constexpr int kBStageBytes = 2 * kBHalfBytes; // [g1 | g2]
    static constexpr int kA2Col0 = 128;                  // h1 at columns [0, 128), h2 at [128, 256)
    static constexpr int kAccCol0 = 256;
    static constexpr int kTmemCols = 512;
    static constexpr int kNumBars = 8 /*bfull*/ + 8 /*bempty*/ + 8 /*xfull*/ + 8 /*xempty*/ + 8 /*a_full*/ + 4 /*a_free*/ + 2 + 2;
    static constexpr int kXchgBytes = 2 * 128 * 5 * 4;
    static constexpr int kTailBytes = 2 * BN * 4 + kNumBars * 8 + 16 + kXchgBytes;
    static int b_stages(int kc) {
        int s = (227 * 1024 - 1024 - kTailBytes - kc * kXStageBytes) / kBStageBytes;
        return s > 8 ? 8 : s;
    }
    static int smem_bytes(int kc, int bs) { return 1024 + kc * kXStageBytes + bs * kBStageBytes + kTailBytes; }
};

// tm_b: FP16 [kpad x 2*ld64] (g1 in columns [0, ld64), g2 in [ld64, 2*ld64)), boxes of 64 values x BN/2 rows.
// p.stages = B ring stages (2..8).  p.cstats: [0] max ||c||, [2] 1/(sx sw), [3] sx, [4]/[5] absolute error units.
template <int BN, bool DBG>
__global__ void __launch_bounds__(kPairH16Threads, 1)
assign_pair_h16_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_b,
                       const AssignParams p) {
    using Cfg = PairH16Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

    const int BS = p.stages;
    const int NT = p.num_col_tiles;
    const int KC = p.num_k_chunks;       // 32-value chunks of X, <= 8
    const int C64 = (KC + 1) >> 1;       // 64-value chunks of B
    const int ld64 = C64 * 64;
    uint8_t *xs = smem;                                  // [KC][128 rows x 128 B], swizzled
    uint8_t *bs = smem + KC * kXStageBytes;              // [BS][g1 chunk | g2 chunk] (this CTA's rows)
    uint8_t *tail = bs + BS * Cfg::kBStageBytes;
    float *cn_s = reinterpret_cast<float *>(tail);       // [2][BN]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail + 2 * BN * 4);
    uint64_t *bfull = bars;           // [8]  (leader) B stage landed in BOTH CTAs
    uint64_t *bempty = bars + 8;      // [8]  the MMAs that read the stage completed (multicast commit)
    uint64_t *xfull = bars + 16;      // [8]  X chunk landed
    uint64_t *xempty = bars + 24;     // [8]  X chunk read by its transform warpgroup
    uint64_t *a_full = bars + 32;     // [8]  h1/h2 of a 32-value chunk written (leader: its 4 warps + the peer's 4)
    uint64_t *a_free = bars + 40;     // [4]  last MMAs of the row tile that read a 64-value chunk completed (multicast)
    uint64_t *acc_full = bars + 44;   // [2]
    uint64_t *acc_empty = bars + 46;  // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + Cfg::kNumBars);
    uint32_t *xchg = tmem_slot + 4;   // [2][5][128]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    // row tiles are dealt to pairs; both CTAs always run the same number of tiles (a tile index beyond
    // the table is a dummy: TMA zero-fills it and nothing is stored)
    const int tile0 = blockIdx.x;
    const int base0 = tile0 - (int)rank;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_x);
        tma_prefetch_desc(&tm_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < 8; ++s) {
            mbar_init(&bfull[s], 1);
            mbar_init(&bempty[s], 1);
            mbar_init(&xfull[s], 1);
            mbar_init(&xempty[s], 4);
            mbar_init(&a_full[s], leader ? 8 : 4);
        }
        for (int s = 0; s < 4; ++s) mbar_init(&a_free[s], 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], leader ? 16 : 8);   // two epilogue warpgroups per CTA
        }
        fence_mbar_init();
    }
    cluster_sync_all();
    if (warp == 2) tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();               // the peer's barriers and TMEM exist before anything touches them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================ B producer ==================================
        uint32_t it = 0;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                const int brow = nt * BN + (int)rank * Cfg::kBRows;
                for (int c = 0; c < C64; ++c, ++it) {
                    const uint32_t s = it % BS, ph = (it / BS) & 1;
                    mbar_wait(&bempty[s], ph ^ 1);
                    if (elect_one()) {
                        uint8_t *st = bs + s * Cfg::kBStageBytes;
                        // both CTAs' bytes are accounted on the leader's barrier; the peer cannot run ahead of
                        // the ring (bempty comes from the leader's MMAs), and bytes that land before the leader
                        // arms the phase only make its transaction count transiently negative
                        if (leader) mbar_arrive_expect_tx(&bfull[s], 2 * Cfg::kBStageBytes);
                        tma_load_2d_pair(&tm_b, &bfull[s], st, c * 64, brow);
                        tma_load_2d_pair(&tm_b, &bfull[s], st + Cfg::kBHalfBytes, ld64 + c * 64, brow);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == 3) {
        // ================================ X producer ==================================
        uint32_t xph = 0;
        for (int rt = tile0; rt - (int)rank < p.num_row_tiles; rt += gridDim.x) {
            for (int kc = 0; kc < KC; ++kc) {
                mbar_wait(&xempty[kc], xph ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&xfull[kc], kXStageBytes);
                    tma_load_2d(&tm_x, &xfull[kc], xs + kc * kXStageBytes, kc * kChunkK, rt * kTileM);
                }
                __syncwarp();
            }
            xph ^= 1;
        }
    } else if (warp == 1) {
        if (leader) {
            // ================================ MMA issuer ==============================
            constexpr uint32_t idesc = make_idesc_f16(2 * kTileM, BN);
            uint32_t it = 0, tile = 0, rtn = 0;
            long long t_wait_acc = 0, t_wait = 0, t_issue = 0, t_mark = 0;
            const bool prof = p.role_prof != nullptr && blockIdx.x == 0;
            for (int base = base0; base < p.num_row_tiles; base += gridDim.x, ++rtn) {
                for (int nt = 0; nt < NT; ++nt, ++tile) {
                    const uint32_t ab = tile & 1, aph = (tile >> 1) & 1;
                    if (prof) t_mark = clock64();
                    mbar_wait(&acc_empty[ab], aph ^ 1);
                    if (prof) { const long long t = clock64(); t_wait_acc += t - t_mark; t_mark = t; }
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + Cfg::kAccCol0 + ab * BN;
                    for (int c = 0; c < C64; ++c, ++it) {
                        const uint32_t s = it % BS, ph = (it / BS) & 1;
                        const int nk = (2 * c + 1 < KC) ? 4 : 2;   // 16-value steps in this chunk
                        if (nt == 0) {
                            // first centroid tile of a row tile: the chunk's h1/h2 must have been written
                            const uint32_t k0 = 2 * c, k1 = (2 * c + 1 < KC) ? 2 * c + 1 : 2 * c;
                            mbar_wait3(&a_full[k0], rtn & 1, &a_full[k1], rtn & 1, &bfull[s], ph);
                        } else {
                            mbar_wait(&bfull[s], ph);
                        }
                        if (prof) { const long long t = clock64(); t_wait += t - t_mark; t_mark = t; }
                        tc_fence_after();
                        if (elect_one()) {
                            const uint32_t bsm = smem_u32(bs + s * Cfg::kBStageBytes);
                            const uint64_t d1 = make_kmajor_sw128_desc(bsm);
                            const uint64_t d2 = make_kmajor_sw128_desc(bsm + Cfg::kBHalfBytes);
                            const uint32_t a1 = tmem_base + c * 32, a2 = a1 + Cfg::kA2Col0;
                            // small terms first
                            for (int ks = 0; ks < nk; ++ks)
                                mma_bf16_ts_pair(d_tmem, a2 + ks * 8, d1 + static_cast<uint64_t>(ks * 2), idesc, (c | ks) == 0 ? 0u : 1u);
                            for (int ks = 0; ks < nk; ++ks)
                                mma_bf16_ts_pair(d_tmem, a1 + ks * 8, d2 + static_cast<uint64_t>(ks * 2), idesc, 1u);
                            for (int ks = 0; ks < nk; ++ks)
                                mma_bf16_ts_pair(d_tmem, a1 + ks * 8, d1 + static_cast<uint64_t>(ks * 2), idesc, 1u);
                            tc_commit_pair(&bempty[s]);
                            if (nt == NT - 1) tc_commit_pair(&a_free[c]);   // the chunk's h1/h2 may be overwritten
                            if (c == C64 - 1) tc_commit_pair(&acc_full[ab]);
                        }
                        __syncwarp();
                        if (prof) { const long long t = clock64(); t_issue += t - t_mark; t_mark = t; }
                    }
                }
            }
            if (prof && lane == 0) {
                p.role_prof[0] = t_wait_acc;
                p.role_prof[1] = t_wait;
                p.role_prof[2] = 0;
                p.role_prof[3] = t_issue;
                p.role_prof[4] = it;
            }
        }
    } else if ((warp >= 4 && warp < 8) || (warp >= 12 && warp < 16)) {
        // ================================ transform ===================================
        // once per row tile: two warpgroups, warps 4-7 take the even 32-value chunks, warps 12-15 the odd ones
        const int wg = warp >= 12 ? 1 : 0;
        const int r = threadIdx.x & 127;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const float sx = p.cstats[3];
        uint32_t rtn = 0;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x, ++rtn) {
            for (int kc = wg; kc < KC; kc += 2) {
                mbar_wait(&xfull[kc], rtn & 1);
                const uint8_t *xrow = xs + kc * kXStageBytes + r * 128;
                float x[32];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 v = *reinterpret_cast<const float4 *>(xrow + ((q ^ (r & 7)) << 4));
                    x[4 * q + 0] = v.x * sx;
                    x[4 * q + 1] = v.y * sx;
                    x[4 * q + 2] = v.z * sx;
                    x[4 * q + 3] = v.w * sx;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&xempty[kc]);   // the chunk is in registers: the next row tile's may land
                uint32_t h1[16], h2[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const __half2 a = __floats2half2_rn(x[2 * i], x[2 * i + 1]);      // element 2i in the low half
                    const float2 af = __half22float2(a);
                    const __half2 b = __floats2half2_rn(x[2 * i] - af.x, x[2 * i + 1] - af.y);   // residuals are exact
                    h1[i] = *reinterpret_cast<const uint32_t *>(&a);
                    h2[i] = *reinterpret_cast<const uint32_t *>(&b);
                }
                // the previous row tile's last MMAs on this chunk have completed
                if (rtn > 0) mbar_wait(&a_free[kc >> 1], (rtn - 1) & 1);
                tc_fence_after();
                if (!(p.knockout & 2)) {
                    tmem_st16(tmem_base + lane_addr + kc * 16, h1);
                    tmem_st16(tmem_base + lane_addr + Cfg::kA2Col0 + kc * 16, h2);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(&a_full[kc]);
                    else mbar_arrive_remote_relaxed(&a_full[kc], 0);
                }
            }
        }
    } else if ((warp >= 8 && warp < 12) || warp >= 16) {
        // ================================ epilogue ====================================
        // as in assign_pair_ss_kernel: two warpgroups, each drains half of the columns of every accumulator tile
        const int eg = warp >= 16 ? 1 : 0;
        const int r = threadIdx.x & 127;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        constexpr int HALF = BN >= 64 ? BN / 2 : BN;
        const int col0 = eg * HALF;
        const bool has_cols = col0 < BN;
        const float cmax = p.cstats[0];
        const float unscale = p.cstats[2];
        uint32_t tile = 0, rtn = 0;
        for (int rt = tile0; rt - (int)rank < p.num_row_tiles; rt += gridDim.x, ++rtn) {
            constexpr int NS = 4;
            const float inf = __int_as_float(0x7f800000);
            float best[NS], second[NS], third[NS];
            int bidx[NS], sidx[NS];
#pragma unroll
            for (int q = 0; q < NS; ++q) { best[q] = second[q] = third[q] = inf; bidx[q] = sidx[q] = 0; }
            for (int nt = 0; nt < NT; ++nt, ++tile) {
                const uint32_t ab = tile & 1, aph = (tile >> 1) & 1;
                float *cn = cn_s + ab * BN;
                if (eg == 0 && r < BN) cn[r] = p.cnorm[nt * BN + r];
                asm volatile("bar.sync 1, 256;" ::: "memory");   // both epilogue warpgroups
                mbar_wait(&acc_full[ab], aph);
                tc_fence_after();
                if (has_cols && !(p.knockout & 4)) {
                    const uint32_t acc = tmem_base + lane_addr + Cfg::kAccCol0 + ab * BN;
#pragma unroll 1
                    for (int c0 = col0; c0 < col0 + HALF; c0 += 32) {
                        uint32_t v[32];
                        tmem_ld32(acc + c0, v);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const int q = i & (NS - 1);
                            const float sc = fmaf(__uint_as_float(v[i]), unscale, cn[c0 + i]);
                            if constexpr (DBG) {
                                if (rt * kTileM + r < p.n)
                                    p.dbg_scores[(size_t)(rt * kTileM + r) * p.kpad + nt * BN + c0 + i] = sc;
                            }
                            const int j = nt * BN + c0 + i;
                            const bool lt1 = sc < best[q], lt2 = sc < second[q];
                            third[q] = fminf(third[q], fmaxf(sc, second[q]));
                            sidx[q] = lt1 ? bidx[q] : (lt2 ? j : sidx[q]);
                            second[q] = fminf(second[q], fmaxf(sc, best[q]));
                            bidx[q] = lt1 ? j : bidx[q];
                            best[q] = fminf(best[q], sc);
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(&acc_empty[ab]);
                    else mbar_arrive_remote_relaxed(&acc_empty[ab], 0);
                }
            }
            // merge sets 1..3 into set 0: lower value first, equal values -> lower index first
            float mb = best[0], ms = second[0], mt = third[0];
            int mbi = bidx[0], msi = sidx[0];
            auto insert = [&](float v, int j, bool has_index) {
                if (has_index && (v < mb || (v == mb && j < mbi))) { mt = ms; ms = mb; msi = mbi; mb = v; mbi = j; }
                else if (has_index && (v < ms || (v == ms && j < msi))) { mt = ms; ms = v; msi = j; }
                else if (v < mt) mt = v;
            };
#pragma unroll
            for (int q = 1; q < NS; ++q) {
                insert(best[q], bidx[q], true);
                insert(second[q], sidx[q], true);
                insert(third[q], 0, false);
            }
            uint32_t *xb = xchg + (rtn & 1) * 5 * 128;
            if (eg == 1) {
                xb[0 * 128 + r] = __float_as_uint(mb);
                xb[1 * 128 + r] = __float_as_uint(ms);
                xb[2 * 128 + r] = __float_as_uint(mt);
                xb[3 * 128 + r] = (uint32_t)mbi;
                xb[4 * 128 + r] = (uint32_t)msi;
            }
            asm volatile("bar.sync 2, 256;" ::: "memory");
            if (eg == 0) {
                insert(__uint_as_float(xb[0 * 128 + r]), (int)xb[3 * 128 + r], true);
                insert(__uint_as_float(xb[1 * 128 + r]), (int)xb[4 * 128 + r], true);
                insert(__uint_as_float(xb[2 * 128 + r]), 0, false);
                const int row = rt * kTileM + r;
                if (row < p.n && !(p.knockout & 4)) {
                    p.labels[row] = mbi;
                    const float xn = sqrtf(p.xnorm[row]);
                    // relative part (accumulation, split residuals) + absolute part (FP16 subnormal range)
                    const float bound = 2.0f * (p.eps * (2.0f * xn * cmax + cmax * cmax) +
                                                p.sqrt_ld * (p.cstats[4] * 2.0f * cmax + p.cstats[5] * xn));
                    if (ms - mb <= bound) flag_row(p, row, mbi, msi, mt - mb <= bound);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();               // no CTA leaves while its peer can still arrive on its barriers
    if (warp == 2) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
}

}  // namespace kmb
