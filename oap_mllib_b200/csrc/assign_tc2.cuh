// Second-generation assign kernel for rows of at most 256 floats (ld <= 256): same arithmetic as
// assign_tc_kernel (3xTF32, FP32 accumulators in TMEM, top-3 epilogue), restructured to cut the
// L2 -> SM traffic and the per-chunk handshakes that limited the first kernel:
//   * the X tile (128 rows x ld) stays RESIDENT in shared memory for all centroid tiles of a row
//     tile (loaded once by its own producer warp; the first kernel re-read it k/BN times from L2);
//   * CL == 2: the kernel runs as CTA PAIRS (clusters of two CTAs on one TPC) and the MMAs are
//     tcgen05.mma.cta_group::2 with M = 256: each CTA owns 128 rows (its own X tile, its own A
//     operand and accumulators in its own TMEM) while every centroid tile is SPLIT across the pair
//     -- each CTA loads and holds only BN/2 of its rows -- so centroid bytes per SM, per L2 read
//     and per shared-memory read are halved and twice as many stages fit.  Only the leader CTA
//     issues MMAs; tcgen05.commit multicasts the completion to both CTAs' barriers;
//   * the TMEM A operand has FOUR single-chunk buffers (4 x 64 columns), so the transform runs up to
//     three chunks ahead of the MMAs (the commit -> transform -> arrive chain crosses CTAs in pair
//     mode and is slow), and the MMA warp polls all barriers of a chunk pair together, so one
//     handshake covers 24 MMAs.
// Roles (per CTA): warp 0 B producer, warp 1 MMA issuer (leader CTA only), warp 2 TMEM allocator,
// warp 3 X producer, warps 4-7 transform (thread r = row r), warps 8-11 epilogue (thread r = row r).
#pragma once
#include "assign_tc.cuh"

namespace kmb {

template <int BN, int CL>
struct Assign2Cfg {
    static constexpr int kBRows = BN / CL;                    // centroid rows held by one CTA
    static constexpr int kBChunkBytes = kBRows * kChunkK * 4; // one of B_hi / B_lo for one K chunk
    static constexpr int kBStageBytes = 2 * kBChunkBytes;     // hi + lo
    static constexpr int kACols = 256;                        // 4 buffers x (32 hi + 32 lo)
    static constexpr int kAccCol0 = kACols;
    static constexpr int kTmemCols = 512;
    static constexpr int kTailBytes = 2 * BN * 4 + (4 * kMaxStages + 12) * 8 + 16;
    static int stages(int kc) {
        int s = (227 * 1024 - 1024 - kTailBytes - kc * kXStageBytes) / kBStageBytes;
        return s > kMaxStages ? kMaxStages : s;
    }
    static int smem_bytes(int kc, int stages) { return 1024 + kc * kXStageBytes + stages * kBStageBytes + kTailBytes; }
};

// tm_bhi / tm_blo must have a box of BN / CL rows.
template <int BN, int CL>
__global__ void __launch_bounds__(kAssignThreads, 1)
assign_tc2_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_bhi,
                  const __grid_constant__ CUtensorMap tm_blo, const AssignParams p) {
    using Cfg = Assign2Cfg<BN, CL>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

    const int S = p.stages;
    const int NT = p.num_col_tiles;
    const int KC = p.num_k_chunks;       // <= 8
    const int G = (KC + 1) >> 1;         // chunk pairs per centroid tile
    uint8_t *xs = smem;                                  // [KC][128 rows x 128 B], swizzled
    uint8_t *bs = smem + KC * kXStageBytes;              // [S][B_hi chunk | B_lo chunk] (this CTA's rows)
    uint8_t *tail = bs + S * Cfg::kBStageBytes;
    float *cn_s = reinterpret_cast<float *>(tail);       // [2][BN]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail + 2 * BN * 4);
    uint64_t *full = bars;                   // B stage landed (pair mode: both halves, on the leader)
    uint64_t *empty = bars + kMaxStages;     // B stage consumed by the MMAs
    uint64_t *xfull = bars + 2 * kMaxStages; // X chunk landed
    uint64_t *xempty = bars + 3 * kMaxStages;// X chunk read for the last time by the transform
    uint64_t *a_full = bars + 4 * kMaxStages;   // [4] (leader) A chunk written by every transform warp of the pair
    uint64_t *a_empty = a_full + 4;             // [4] MMAs reading the A chunk have completed
    uint64_t *acc_full = a_full + 8;            // [2]
    uint64_t *acc_empty = a_full + 10;          // [2] (leader) accumulator drained by every epilogue warp of the pair
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 4 * kMaxStages + 12);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = CL > 1 ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    // row tiles are dealt to clusters; both CTAs of a pair always run the same number of tiles
    // (a tile index beyond the table is a dummy: TMA zero-fills it and nothing is stored)
    const int tile0 = blockIdx.x;
    const int base0 = tile0 - (int)rank;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_x);
        tma_prefetch_desc(&tm_bhi);
        tma_prefetch_desc(&tm_blo);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(&full[s], 1);        // the (leader's) arrive.expect_tx; both CTAs' bytes are counted on it
            mbar_init(&empty[s], 1);
            mbar_init(&xfull[s], 1);
            mbar_init(&xempty[s], 4);          // one arrive per transform warp
        }
        // leader: 4 local warps (+ 1 relayed arrive standing for the peer's 4 warps); peer: 4 local warps
        const uint32_t group_count = (CL > 1 && leader) ? 5 : 4;
        for (int b = 0; b < 4; ++b) {
            mbar_init(&a_full[b], group_count);
            mbar_init(&a_empty[b], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], group_count);
        }
        fence_mbar_init();
    }
    if (CL > 1) cluster_sync_all();
    if (warp == 2) {
        if (CL == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
        else tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();   // the peer's barriers and TMEM exist before anything touches them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================ B producer ==================================
        uint32_t s = 0, ph = 0;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                for (int kc = 0; kc < KC; ++kc) {
                    mbar_wait(&empty[s], ph ^ 1);
                    if (elect_one()) {
                        uint8_t *st = bs + s * Cfg::kBStageBytes;
                        const int brow = nt * BN + (int)rank * Cfg::kBRows;
                        if (p.knockout & 1) {
                            if (leader) mbar_arrive(&full[s]);
                        } else if (CL == 1) {
                            mbar_arrive_expect_tx(&full[s], Cfg::kBStageBytes);
                            tma_load_2d(&tm_bhi, &full[s], st, kc * kChunkK, brow);
                            tma_load_2d(&tm_blo, &full[s], st + Cfg::kBChunkBytes, kc * kChunkK, brow);
                        } else {
                            // both CTAs' bytes are accounted on the leader's barrier
                            if (leader) mbar_arrive_expect_tx(&full[s], 2 * Cfg::kBStageBytes);
                            tma_load_2d_pair(&tm_bhi, &full[s], st, kc * kChunkK, brow);
                            tma_load_2d_pair(&tm_blo, &full[s], st + Cfg::kBChunkBytes, kc * kChunkK, brow);
                            // the peer cannot run ahead of the ring (its empty[] comes from the leader's
                            // MMAs), so its bytes always land in the phase the leader is arming
                        }
                    }
                    __syncwarp();
                    if (++s == (uint32_t)S) { s = 0; ph ^= 1; }
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
        // ================================ MMA issuer (leader CTA) ======================
        if (leader) {
            constexpr uint32_t idesc = make_idesc_tf32(kTileM * CL, BN);
            uint32_t grp = 0, ch = 0, tile = 0, s = 0, ph = 0;
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
                        const uint32_t b0 = ch & 3, bp0 = (ch >> 2) & 1;
                        const uint32_t b1 = (ch + 1) & 3, bp1 = ((ch + 1) >> 2) & 1;
                        uint32_t s1 = s + 1, ph1 = ph;
                        if (s1 == (uint32_t)S) { s1 = 0; ph1 ^= 1; }
                        if (nch == 2) mbar_wait4(&a_full[b0], bp0, &a_full[b1], bp1, &full[s], ph, &full[s1], ph1);
                        else mbar_wait2(&a_full[b0], bp0, &full[s], ph);
                        if (prof) { const long long t = clock64(); t_wait += t - t_mark; t_mark = t; }
                        tc_fence_after();
                        if (elect_one()) {
#pragma unroll
                            for (int c = 0; c < 2; ++c) {
                                if (c < nch) {
                                    const uint32_t sc = c == 0 ? s : s1;
                                    const uint32_t bb = c == 0 ? b0 : b1;
                                    const uint32_t a_hi = tmem_base + bb * 64;
                                    const uint32_t a_lo = a_hi + 32;
                                    const uint32_t bsm = smem_u32(bs + sc * Cfg::kBStageBytes);
                                    const uint64_t dhi = make_kmajor_sw128_desc(bsm);
                                    const uint64_t dlo = make_kmajor_sw128_desc(bsm + Cfg::kBChunkBytes);
#pragma unroll
                                    for (int ks = 0; ks < kChunkK / 8; ++ks) {
                                        const uint64_t koff = static_cast<uint64_t>(ks * 2);
                                        const uint32_t first = (g | c | ks) == 0 ? 0u : 1u;
                                        if (CL == 1) {
                                            mma_tf32_ts(d_tmem, a_lo + ks * 8, dhi + koff, idesc, first);  // small terms first
                                            mma_tf32_ts(d_tmem, a_hi + ks * 8, dlo + koff, idesc, 1u);
                                            mma_tf32_ts(d_tmem, a_hi + ks * 8, dhi + koff, idesc, 1u);
                                        } else {
                                            mma_tf32_ts_pair(d_tmem, a_lo + ks * 8, dhi + koff, idesc, first);
                                            mma_tf32_ts_pair(d_tmem, a_hi + ks * 8, dlo + koff, idesc, 1u);
                                            mma_tf32_ts_pair(d_tmem, a_hi + ks * 8, dhi + koff, idesc, 1u);
                                        }
                                    }
                                    if (CL == 1) {
                                        tc_commit(&empty[sc]);
                                        tc_commit(&a_empty[bb]);
                                    } else {          // both CTAs' producers / transforms may proceed
                                        tc_commit_pair(&empty[sc]);
                                        tc_commit_pair(&a_empty[bb]);
                                    }
                                }
                            }
                            if (g == G - 1) {
                                if (CL == 1) tc_commit(&acc_full[ab]);
                                else tc_commit_pair(&acc_full[ab]);
                            }
                        }
                        __syncwarp();
                        if (prof) { const long long t = clock64(); t_issue += t - t_mark; t_mark = t; }
                        s = nch == 2 ? s1 + 1 : s1;
                        ph = ph1;
                        if (nch == 2 && s == (uint32_t)S) { s = 0; ph ^= 1; }
                        ch += nch;
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
        } else {
            // ---- peer CTA: relay "A chunk written" to the leader.  A remote mbarrier arrive costs
            // several hundred cycles; doing it here keeps that latency off the transform warps.
            uint32_t ch = 0;
            for (int base = base0; base < p.num_row_tiles; base += gridDim.x)
                for (int nt = 0; nt < NT; ++nt)
                    for (int kc = 0; kc < KC; ++kc, ++ch) {
                        mbar_wait(&a_full[ch & 3], (ch >> 2) & 1);
                        if (lane == 0) mbar_arrive_remote(&a_full[ch & 3], 0);
                        __syncwarp();
                    }
        }
    } else if (warp == 2 && CL > 1 && !leader) {
        // ---- peer CTA: relay "accumulator drained" to the leader
        uint32_t tile = 0;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x)
            for (int nt = 0; nt < NT; ++nt, ++tile) {
                mbar_wait(&acc_empty[tile & 1], (tile >> 1) & 1);
                if (lane == 0) mbar_arrive_remote(&acc_empty[tile & 1], 0);
                __syncwarp();
            }
    } else if (warp >= 4 && warp < 8) {
        // ================================ transform ===================================
        const int r = threadIdx.x - 128;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        uint32_t ch = 0, xph = 0;
        long long t_load = 0, t_wait_a = 0, t_store = 0, t_mark = 0;
        const bool prof = p.role_prof != nullptr && blockIdx.x < 2 && warp == 4;
        for (int base = base0; base < p.num_row_tiles; base += gridDim.x) {
            for (int nt = 0; nt < NT; ++nt) {
                for (int kc = 0; kc < KC; ++kc, ++ch) {
                    const uint32_t bb = ch & 3, bph = (ch >> 2) & 1;
                    if (prof) t_mark = clock64();
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
                    uint32_t hi[32], lo[32];
                    // Veltkamp split with 2^13 + 1: hi = x rounded to 11 significant bits (exactly a TF32
                    // value), lo = x - hi exactly; the tensor core drops lo's low 13 bits (|error| <=
                    // 2^-21 |x|).  All four operations run on the FMA pipe, leaving the ALU pipe (min /
                    // max / select) to the epilogue warps that share these schedulers.
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float t = __fmul_rn(x[i], 8193.0f);
                        const float h = __fsub_rn(t, __fsub_rn(t, x[i]));
                        hi[i] = __float_as_uint(h);
                        lo[i] = __float_as_uint(__fsub_rn(x[i], h));
                    }
                    if (nt == NT - 1) {   // last read of this chunk: the X producer may reload it
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&xempty[kc]);
                    }
                    if (prof) { const long long t = clock64(); t_load += t - t_mark; t_mark = t; }
                    mbar_wait(&a_empty[bb], bph ^ 1);   // MMAs that read this A buffer have completed
                    if (prof) { const long long t = clock64(); t_wait_a += t - t_mark; t_mark = t; }
                    tc_fence_after();
                    if (!(p.knockout & 2)) {
                        const uint32_t a_addr = tmem_base + lane_addr + bb * 64;
                        tmem_st32(a_addr, hi);
                        tmem_st32(a_addr + 32, lo);
                    }
                    tmem_wait_st();
                    tc_fence_before();
                    __syncwarp();          // one (possibly remote) arrive per warp
                    if (lane == 0) mbar_arrive(&a_full[bb]);   // local; the peer's relay warp forwards it
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
            o[3] = ch;
        }
    } else if (warp >= 8) {
        // ================================ epilogue ====================================
        const int r = threadIdx.x - 256;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const float cmax = p.cstats[0];
        uint32_t tile = 0;
        for (int rt = tile0; rt - (int)rank < p.num_row_tiles; rt += gridDim.x) {
            float best = __int_as_float(0x7f800000), second = best, third = best;
            int bidx = 0, sidx = 0;
            for (int nt = 0; nt < NT; ++nt, ++tile) {
                const uint32_t ab = tile & 1, aph = (tile >> 1) & 1;
                float *cn = cn_s + ab * BN;
                if (r < BN) cn[r] = p.cnorm[nt * BN + r];
                asm volatile("bar.sync 1, 128;" ::: "memory");
                mbar_wait(&acc_full[ab], aph);
                tc_fence_after();
                if (!(p.knockout & 4)) {
                    const uint32_t acc = tmem_base + lane_addr + Cfg::kAccCol0 + ab * BN;
                    constexpr int GW = BN >= 32 ? 32 : 16;
#pragma unroll 1
                    for (int c0 = 0; c0 < BN; c0 += GW) {
                        uint32_t v[GW];
                        if constexpr (GW == 32) tmem_ld32(acc + c0, v);
                        else tmem_ld16(acc + c0, v);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < GW; ++i) {
                            const float sc = __uint_as_float(v[i]) + cn[c0 + i];
                            if (p.dbg_scores != nullptr && rt * kTileM + r < p.n)
                                p.dbg_scores[(size_t)(rt * kTileM + r) * p.kpad + nt * BN + c0 + i] = sc;
                            const int j = nt * BN + c0 + i;
                            const bool lt1 = sc < best, lt2 = sc < second;
                            third = fminf(third, fmaxf(sc, second));
                            sidx = lt1 ? bidx : (lt2 ? j : sidx);
                            second = fminf(second, fmaxf(sc, best));
                            bidx = lt1 ? j : bidx;
                            best = fminf(best, sc);
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[ab]);    // local; the peer's relay warp forwards it
            }
            const int row = rt * kTileM + r;
            if (row < p.n && !(p.knockout & 4)) {
                p.labels[row] = bidx;
                const float xn = sqrtf(p.xnorm[row]);
                const float bound = 2.0f * p.eps * (2.0f * xn * cmax + cmax * cmax);
                if (second - best <= bound) flag_row(p, row, bidx, sidx, third - best <= bound);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();   // no CTA leaves while its peer can still arrive on its barriers
    if (warp == 2) {
        if (CL == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
        else tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    }
}

}  // namespace kmb
