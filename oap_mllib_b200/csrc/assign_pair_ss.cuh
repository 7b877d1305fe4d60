// CTA-pair assign kernel, "SS" variant (ld <= 256, BN >= 32): the fallback of the FP16-split kernel ("h16" = 0).
// TF32 main product + BF16 correction products; the
// main product reads its A operand STRAIGHT FROM THE RESIDENT X TILE in shared memory
// (tcgen05.mma ... kind::tf32 with a shared-memory A descriptor: the tensor core drops the low 13
// mantissa bits of x itself), so the transform warps only produce the correction operand
//     [bf16(x - trunc_tf32(x)) x 32 | bf16(x) x 32]  = 32 TMEM columns per chunk.
// That halves the transform's work and TMEM traffic and, above all, leaves room for EIGHT chunk
// buffers of the A operand instead of four: the commit -> transform -> arrive loop crosses the CTA
// pair twice (~2500 cycles) and with 4 buffers it, not the tensor pipe, set the pace.
#pragma once
// (derived from the round-1 kernel that kept the whole A operand in TMEM, since removed)
#include <cuda_bf16.h>

#include "assign_tc.cuh"

namespace kmb {

constexpr int kPairSsThreads = 640;   // + warps 16-19: second epilogue warpgroup

template <int BN>
struct PairSsCfg {
    static constexpr int kBRows = BN / 2;                     // centroid rows held by one CTA
    static constexpr int kBChunkBytes = kBRows * kChunkK * 4; // one of B_hi / B_lo for one K chunk
    static constexpr int kBStageBytes = 2 * kBChunkBytes;     // hi + lo
    static constexpr int kAccCol0 = 256;                      // A corrections: 8 chunk buffers x 32 columns
    static constexpr int kTmemCols = 512;
    static constexpr int kNumBars = 8 /*full*/ + 8 /*gdone*/ + 8 /*xfull*/ + 8 /*xempty*/ + 8 /*a_full*/ + 8 /*cdone*/ + 2 + 2;
    static constexpr int kXchgBytes = 2 * 128 * 5 * 4;         // top-3 sets handed from epilogue warpgroup 1 to 0
    static constexpr int kTailBytes = 2 * BN * 4 + kNumBars * 8 + 16 + kXchgBytes;
    // B ring: RB group slots of two stages each (RB <= 4)
    static int group_slots(int kc) {
        int s = (227 * 1024 - 1024 - kTailBytes - kc * kXStageBytes) / (2 * kBStageBytes);
        return s > 4 ? 4 : s;
    }
    static int smem_bytes(int kc, int rb) { return 1024 + kc * kXStageBytes + rb * 2 * kBStageBytes + kTailBytes; }
};

// tm_bhi: FP32 [kpad x ld], tm_blo: the BF16 correction operand [kpad x 2*ld]; boxes of BN/2 rows.
// p.stages = RB (group slots of the B ring).
// DBG: also dump every score to p.dbg_scores (tests only; costs epilogue instructions).
template <int BN, bool DBG>
__global__ void __launch_bounds__(kPairSsThreads, 1)
assign_pair_ss_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_bhi,
                   const __grid_constant__ CUtensorMap tm_blo, const AssignParams p) {
    using Cfg = PairSsCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

    const int RB = p.stages;             // 2 or 4 group slots in the B ring
    constexpr int P = 4;                 // period of the gdone ring: covers the B ring (RB) and the 4 A group slots
    const int NT = p.num_col_tiles;
    const int KC = p.num_k_chunks;       // <= 8
    const int G = (KC + 1) >> 1;         // groups (chunk pairs) per centroid tile
    uint8_t *xs = smem;                                  // [KC][128 rows x 128 B], swizzled
    uint8_t *bs = smem + KC * kXStageBytes;              // [RB][2][B_hi chunk | B_lo chunk] (this CTA's rows)
    uint8_t *tail = bs + RB * 2 * Cfg::kBStageBytes;
    float *cn_s = reinterpret_cast<float *>(tail);       // [2][BN]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail + 2 * BN * 4);
    uint64_t *full = bars;            // [8]  (leader) B stage landed in BOTH CTAs
    uint64_t *gdone = bars + 8;       // [8]  all MMAs of a group completed (multicast commit): its B stages
                                      //      and its A slot may be overwritten
    uint64_t *xfull = bars + 16;      // [8]  X chunk landed
    uint64_t *xempty = bars + 24;     // [8]  X chunk read for the last time by the transform
    uint64_t *a_full = bars + 32;     // [8]  correction chunk buffer written (leader: its 4 warps + the peer's 4 warps)
    // (bars + 40 .. 47 unused)
    uint64_t *acc_full = bars + 48;   // [2]  accumulator tile complete (multicast commit)
    uint64_t *acc_empty = bars + 50;  // [2]  accumulator drained (leader: its 4 warps + the peer's 4 warps)
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
        tma_prefetch_desc(&tm_bhi);
        tma_prefetch_desc(&tm_blo);
    }
    if (warp == 1 && lane == 0) {
        // leader: its 4 warps + the peer's 4 warps, which arrive remotely (relaxed: their TMEM stores / loads
        // were already ordered by tcgen05.wait + fence); no relay warps in this variant
        const uint32_t group_count = leader ? 8 : 4;
        const uint32_t drain_count = leader ? 16 : 8;   // two epilogue warpgroups per CTA
        for (int s = 0; s < 8; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&gdone[s], 1);
            mbar_init(&xfull[s], 1);
            mbar_init(&xempty[s], 5);   // the 4 transform warps that read the chunk + the commit of the MMAs that read it
        }
        for (int b = 0; b < 8; ++b) mbar_init(&a_full[b], group_count);   // one transform warpgroup per chunk
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], drain_count);
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

    // "group g - D has completed" on the gdone ring (D = 4 for the correction buffers, RB for the B slot)
    auto wait_group_done = [&](uint32_t g, uint32_t D) {
        if (g >= D) {
            const uint32_t q = g - D;
            mbar_wait(&gdone[q % P], (q / P) & 1);
        }
    };

    if (warp == 0) {
        // ================================ B producer ==================================
        uint32_t grp = 0;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                for (int g = 0; g < G; ++g, ++grp) {
                    const int nch = (2 * g + 1 < KC) ? 2 : 1;
                    const uint32_t slot = grp % RB;
                    wait_group_done(grp, RB);
                    if (elect_one()) {
                        const int brow = nt * BN + (int)rank * Cfg::kBRows;
                        for (int c = 0; c < 2; ++c) {
                            const uint32_t s = slot * 2 + c;
                            uint8_t *st = bs + s * Cfg::kBStageBytes;
                            const int kc = 2 * g + c;
                            if (c >= nch || (p.knockout & 1)) {
                                // an odd chunk count leaves the group's second stage unused: keep its
                                // barrier phase in step with the slot
                                if (leader) mbar_arrive(&full[s]);
                            } else {
                                // both CTAs' bytes are accounted on the leader's barrier; the peer cannot
                                // run ahead of the ring (gdone comes from the leader's MMAs), so its bytes
                                // always land in the phase the leader is arming
                                if (leader) mbar_arrive_expect_tx(&full[s], 2 * Cfg::kBStageBytes);
                                tma_load_2d_pair(&tm_bhi, &full[s], st, kc * kChunkK, brow);
                                tma_load_2d_pair(&tm_blo, &full[s], st + Cfg::kBChunkBytes, kc * 2 * kChunkK, brow);
                            }
                        }
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
            constexpr uint32_t idesc = make_idesc_tf32(2 * kTileM, BN);
            constexpr uint32_t idesc16 = make_idesc_bf16(2 * kTileM, BN);
            uint32_t grp = 0, tile = 0;
            long long t_wait_acc = 0, t_wait = 0, t_issue = 0, t_mark = 0;
            const bool prof = p.role_prof != nullptr && blockIdx.x == 0;
            for (int base = base0; base < p.num_row_tiles; base += gridDim.x) {
                for (int nt = 0; nt < NT; ++nt, ++tile) {
                    const uint32_t ab = tile & 1, aph = (tile >> 1) & 1;
                    if (prof) t_mark = clock64();
                    mbar_wait(&acc_empty[ab], aph ^ 1);
                    if (prof) { const long long t = clock64(); t_wait_acc += t - t_mark; t_mark = t; }
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + Cfg::kAccCol0 + ab * BN;
                    for (int g = 0; g < G; ++g, ++grp) {
                        const int nch = (2 * g + 1 < KC) ? 2 : 1;
                        // running chunk index (two per group, the second may be a phantom): buffer n % 8
                        const uint32_t n0 = grp * 2, b0 = n0 & 7, b1 = b0 + 1, bp = (n0 >> 3) & 1;
                        const uint32_t bslot = grp % RB, bph = (grp / RB) & 1;
                        const uint32_t s0 = bslot * 2, s1 = s0 + 1;
                        // ONE poll for the whole group: the MMA queue is only a few instructions deep, so
                        // every cycle this warp spends between two issue bursts is a cycle the tensor pipe idles
                        mbar_wait4(&a_full[b0], bp, &a_full[b1], bp, &full[s0], bph, &full[s1], bph);
                        if (prof) { const long long t = clock64(); t_wait += t - t_mark; t_mark = t; }
                        tc_fence_after();
                        if (elect_one()) {
#pragma unroll
                            for (int c = 0; c < 2; ++c) {
                                if (c < nch) {
                                    const int kc = 2 * g + c;
                                    const uint32_t a_corr = tmem_base + (b0 + c) * 32;
                                    const uint32_t bsm = smem_u32(bs + (s0 + c) * Cfg::kBStageBytes);
                                    const uint64_t dhi = make_kmajor_sw128_desc(bsm);
                                    const uint64_t dco = make_kmajor_sw128_desc(bsm + Cfg::kBChunkBytes);
                                    const uint64_t dx = make_kmajor_sw128_desc(smem_u32(xs + kc * kXStageBytes));
                                    // corrections first (small terms): K = 64 BF16 values = 4 instructions, A from TMEM
#pragma unroll
                                    for (int ks = 0; ks < 4; ++ks)
                                        mma_bf16_ts_pair(d_tmem, a_corr + ks * 8, dco + static_cast<uint64_t>(ks * 2), idesc16,
                                                         (g | c | ks) == 0 ? 0u : 1u);
                                    // main product: A = the resident X chunk in shared memory (raw FP32; the tensor
                                    // core uses the upper 19 bits), B = TF32-rounded -2c
#pragma unroll
                                    for (int ks = 0; ks < kChunkK / 8; ++ks)
                                        mma_tf32_ss_pair(d_tmem, dx + static_cast<uint64_t>(ks * 2), dhi + static_cast<uint64_t>(ks * 2), idesc, 1u);
                                    if (nt == NT - 1) tc_commit_pair(&xempty[kc]);   // last MMAs that read this X chunk
                                }
                            }
                            tc_commit_pair(&gdone[grp & 3]);   // the group's B stages and correction buffers are free
                            if (g == G - 1) tc_commit_pair(&acc_full[ab]);
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
                p.role_prof[4] = grp;
            }
        }
    } else if ((warp >= 4 && warp < 8) || (warp >= 12 && warp < 16)) {
        // ================================ transform ===================================
        // two warpgroups: warps 4-7 convert the first chunk of every group, warps 12-15 the second
        const int mychunk = warp >= 12 ? 1 : 0;
        const int r = threadIdx.x & 127;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        uint32_t grp = 0, xph = 0;
        long long t_load = 0, t_wait_a = 0, t_store = 0, t_mark = 0;
        const bool prof = p.role_prof != nullptr && blockIdx.x < 2 && warp == 4;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                for (int g = 0; g < G; ++g, ++grp) {
                    const int nch = (2 * g + 1 < KC) ? 2 : 1;
                    const uint32_t n = grp * 2 + mychunk, bb = n & 7;   // running chunk index
                    uint32_t corr[32];
                    const int kc = 2 * g + mychunk;
                    if (prof) t_mark = clock64();
                    if (mychunk < nch) {
                        if (nt == 0) mbar_wait(&xfull[kc], xph);   // X chunk of this row tile has landed
                        const uint8_t *xrow = xs + kc * kXStageBytes + r * 128;
                        float x[32];
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float4 v = *reinterpret_cast<const float4 *>(xrow + ((q ^ (r & 7)) << 4));
                            x[4 * q + 0] = v.x;
                            x[4 * q + 1] = v.y;
                            x[4 * q + 2] = v.z;
                            x[4 * q + 3] = v.w;
                        }
                        // what the tensor core will NOT see of x: lo = x - (x with its low 13 mantissa bits cleared),
                        // exact in FP32.  corr = [bf16(lo) x 32 | bf16(x) x 32], element 2i in the low half of a word.
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float l0 = __fsub_rn(x[2 * i], __uint_as_float(__float_as_uint(x[2 * i]) & 0xffffe000u));
                            const float l1 = __fsub_rn(x[2 * i + 1], __uint_as_float(__float_as_uint(x[2 * i + 1]) & 0xffffe000u));
                            const __nv_bfloat162 a2 = __floats2bfloat162_rn(l0, l1);
                            const __nv_bfloat162 b2 = __floats2bfloat162_rn(x[2 * i], x[2 * i + 1]);
                            corr[i] = *reinterpret_cast<const uint32_t *>(&a2);
                            corr[16 + i] = *reinterpret_cast<const uint32_t *>(&b2);
                        }
                        if (nt == NT - 1) {   // this warp's last read of the chunk (the MMAs' last read is committed by the issuer)
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&xempty[kc]);
                        }
                    }
                    if (prof) { const long long t = clock64(); t_load += t - t_mark; t_mark = t; }
                    // the MMAs of the group that used this buffer 4 groups ago have completed (also keeps an idle
                    // warpgroup -- odd chunk count -- in step with the MMAs, or its arrive below would overrun)
                    wait_group_done(grp, 4);
                    tc_fence_after();
                    if (prof) { const long long t = clock64(); t_wait_a += t - t_mark; t_mark = t; }
                    if (mychunk < nch && !(p.knockout & 2)) tmem_st32(tmem_base + lane_addr + bb * 32, corr);
                    tmem_wait_st();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (leader) mbar_arrive(&a_full[bb]);
                        else mbar_arrive_remote_relaxed(&a_full[bb], 0);
                    }
                    if (prof) { const long long t = clock64(); t_store += t - t_mark; t_mark = t; }
                }
            }
            xph ^= 1;
        }
        if (prof && lane == 0) {
            long long *o = p.role_prof + 5 + 4 * blockIdx.x;
            o[0] = t_load;
            o[1] = t_wait_a;
            o[2] = t_store;
            o[3] = grp;
        }
    } else if ((warp >= 8 && warp < 12) || warp >= 16) {
        // ================================ epilogue ====================================
        // Two warpgroups: each drains half of the columns of every accumulator tile (the running top-3 update is
        // a latency-bound chain; one warpgroup alone drained a tile more slowly than the tensor pipe filled it).
        const int eg = warp >= 16 ? 1 : 0;
        const int r = threadIdx.x & 127;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        constexpr int HALF = BN >= 64 ? BN / 2 : BN;      // columns per warpgroup (BN = 32: warpgroup 1 only syncs)
        const int col0 = eg * HALF;
        const bool has_cols = col0 < BN;
        const float cmax = p.cstats[0];
        uint32_t tile = 0, rtn = 0;
        for (int rt = tile0; rt - (int)rank < p.num_row_tiles; rt += gridDim.x, ++rtn) {
            // Four independent running top-3 sets (columns i mod 4): the update of one set is a serial
            // chain of min/max/select, four chains give the ALU pipe independent work.  Inside a set the
            // columns arrive in ascending order, so strict '<' keeps the lowest index; the sets are
            // merged once per row tile with an explicit (value, index) tie rule.
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
                            const float sc = __uint_as_float(v[i]) + cn[c0 + i];
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
                insert(third[q], 0, false);   // can only ever become the merged third value
            }
            // warpgroup 1 hands its set to warpgroup 0 (double-buffered by row-tile parity)
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
                    const float bound = 2.0f * p.eps * (2.0f * xn * cmax + cmax * cmax);
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
