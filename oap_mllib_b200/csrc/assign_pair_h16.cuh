// CTA-pair assign kernel, "H16" variant (ld <= 256, BN >= 32; the one-product pass alone up to ld = 512): two-term FP16
// splits of both operands.
//
//   x * sx = h1 + h2 (+ <= 2^-22 |x sx|),   -2c * sw = g1 + g2 (+ <= 2^-22 |2c sw|)      (sx, sw powers of two)
//   -2 x.c  ~=  (h2.g1 + h1.g2 + h1.g1) / (sx sw)        three kind::f16 products, FP32 accumulation in TMEM
//
// FP16 products are exact in FP32, the dropped h2.g2 term is 2^-22 of the main one, so the score error is
// smaller than the TF32 + BF16-correction scheme's (assign_pair_ss.cuh) at 3/4 of its tensor-pipe time:
// 3 x (ld/16) MMA instructions per score tile instead of 2 x (ld/8).  Further differences:
//  * both A operands (h1, h2) of a whole 128 x ld row tile live in TMEM (2 x ld/2 columns) for ALL the
//    centroid tiles, so the transform warps touch each X element once per row tile, not once per centroid
//    tile; the X tile only passes through a small shared-memory ring (ld <= 128: two A buffers);
//  * the B operand [g1 | g2] of a whole centroid tile is ONE ring entry (2..4 "groups"), synchronised once per
//    tile; centroid tile nt always lives in group nt % NG and odd row tiles walk the tiles backwards
//    (serpentine), so the tiles around a row-tile switch are not reloaded;
//  * two bookkeeping modes (MODE below): the full running (best, runner-up, third) update, or a fast one that
//    only lists the rows whose two best scores are inside the error bound for a second, full pass.
// The power-of-two scales keep |h|, |g| < 2^15; values below the FP16 normal range (2^-29 of the largest
// element) lose relative precision, which the near-tie bound accounts for with an absolute term
// (cstats[4], cstats[5]); such rows are re-scored in FP64 like every other flagged row.
#pragma once
#include <cuda_fp16.h>

#include "assign_pair_ss.cuh"

namespace kmb {

constexpr int kPairH16Threads = 640;
constexpr uint32_t kTransformSleepNs = 128;   // between polls of the transform warps' wait for the A columns (a tile of slack)

template <int BN>
struct PairH16Cfg {
    static constexpr int kBRows = BN / 2;                // centroid rows held by one CTA
    static constexpr int kBHalfBytes = kBRows * 128;     // g1 (or g2) for one chunk of 64 K-values
    static constexpr int kBUnitBytes = 2 * kBHalfBytes;  // [g1 | g2]
    static constexpr int kA2Col0 = 128;                  // h1 at columns [0, 128), h2 at [128, 256)
    static constexpr int kAccCol0 = 256;
    static constexpr int kTmemCols = 512;
    static constexpr int kNumBars = 4 /*bfull*/ + 8 /*tile_done*/ + 8 /*xfull*/ + 8 /*xempty*/ + 16 /*a_full*/ + 4 /*a_free*/ + 3 /*acc_empty*/;
    static constexpr int kXchgBytes = 2 * 128 * 5 * 4;
    static constexpr int kTailBytes = 2 * BN * 4 + kNumBars * 8 + 16 + kXchgBytes;
    static constexpr int kBudget = 227 * 1024 - 1024 - kTailBytes;
    // ld > 256 (more than 8 chunks; one-product pass only, see the kernel): the units hold g1 alone
    static int unit_bytes(int kc) { return kc > 8 ? kBHalfBytes : kBUnitBytes; }
    static int group_bytes(int kc) { return ((kc + 1) / 2) * unit_bytes(kc); }   // B operand of one centroid tile
    // X ring slots (32-value chunks): as many as fit next to two B groups, at most one row tile, at most 8; an EVEN
    // number >= 2, because each transform warpgroup owns half of them (see "xfull / xempty" in the barrier notes)
    static int x_slots(int kc) {
        int s = (kBudget - 2 * group_bytes(kc)) / kXStageBytes;
        if (s > kc) s = kc;
        s &= ~1;
        return s < 2 ? 2 : s;
    }
    static int b_groups(int kc) {
        int g = (kBudget - x_slots(kc) * kXStageBytes) / group_bytes(kc);
        return g >= 4 ? 4 : g == 3 ? 2 : g;   // 2 or 4: see "bfull" in the barrier notes of the kernel
    }
    static int smem_bytes(int kc) { return 1024 + x_slots(kc) * kXStageBytes + b_groups(kc) * group_bytes(kc) + kTailBytes; }
};

// tm_b: FP16 [kpad x 2*ld64] (g1 in columns [0, ld64), g2 in [ld64, 2*ld64)), boxes of 64 values x BN/2 rows.
// p.stages = X ring slots | B groups << 8.  p.cstats: [0] max ||c||, [2] 1/(sx sw), [3] sx, [4]/[5] absolute error
// units.
//
// Synchronisation is per CENTROID TILE, not per K chunk: every tcgen05.commit / mbarrier wait pair costs the
// issuing warp several hundred cycles during which the (shallow) MMA queue drains, so the B operand of a whole
// centroid tile is one ring entry ("group", 2..4 of them), and ONE commit per tile (tile_done) tells the
// epilogue that the accumulator is complete and the B producer that the group may be refilled.
// MODE: 0 = full bookkeeping (best, runner-up with indices, third value -> p.flag_list), 1 = the same plus the score
// dump of kmb200_debug_scores, 2 = FAST: best with index and the runner-up VALUE only; rows whose gap is inside the
// bound are appended to p.cand_list and go through a MODE 0 launch over a gathered copy (see launch_assign_h16).
// p.n_dev (optional): the row count is read from device memory (min with p.n).
// NP: 3 = the three products above; 1 = h1.g1 ONLY (a third of the tensor work, score error ~2^-10 instead of
// ~2^-20): the coarse first pass of the cascade in launch_assign_h16, always with MODE 2.  It needs neither h2 nor
// g2, so h1 is double-buffered in TMEM for every ld <= 256 and only the g1 halves of the B units are loaded.
// NP = 1 also runs with 256 < ld <= 512 ("wide": up to 16 chunks): h1 of a row tile then fills all 256 A columns (one
// buffer, released in two halves like the three-product kernel's), and the B units are packed g1 only.  The three-
// product kernel cannot (h1 and h2 of 512 values would need all 512 TMEM columns): the rows this pass leaves undecided
// go through the streaming 3xTF32 kernel (launch_assign_h16_wide).
//
// ---- Barrier rings: why no parity wait can alias ------------------------------------------------------------------
// An mbarrier parity wait for "completion number u" of a barrier is only right while the barrier has completed u or
// u + 1 phases when the waiter looks: with u - 1 it passes too early, with u + 2 it blocks for good.  So for every
// ring: (lo) completion u - 1 has happened before any waiter asks for u, and (hi) completion u + 1 cannot happen before
// every waiter of u has passed.  T = running centroid-tile index (tile), owner(T) = issuer T & 1, NG in {2, 4} groups.
// Two facts used throughout: (F1) an issuer handles its tiles in order; (F2) issuing tile T needs acc_empty of T - A
// (A = 2 or 3 accumulator buffers), which the epilogue gives after draining T - A, and the epilogue drains tiles in order
// after waiting for each, so "T issued" implies "every tile <= T - A is complete".
//  * tile_done[8], slot T & 7, completion number T >> 3, one arrival (owner's multicast commit).
//      - epilogue waiter of T: (lo) it passed T - 8 itself.  (hi) T + 8 uses the accumulator of T + 8 - A >= T + 5: it
//        cannot even be issued before the epilogue drained that tile, which comes after T.
//      - B producer, before the LOAD for tile T, waits for q = T - NG, the previous user of that group (in forward and
//        backward row tiles alike: a tile that needs a load has walk index i >= NG, and walk index i - NG holds the same
//        group).  (lo) needs q - 8 complete: the producer's previous load was for T' >= T - RES - 1 >= T - 5 (at most RES
//        <= 4 resident tiles, which need no load, lie between two loads) and it waited for T' - NG then; by F2 everything
//        <= T' - NG - 3 is complete as well, and q - 8 = T - NG - 8 <= T' - NG - 3.  (hi) q + 8 >= T + 4 is not complete:
//        T itself needs this load, the owner of T is stuck at T (F1), and the other issuer gets no further than T + 1
//        (T + 3 needs the accumulator of T + 1 or of T, i.e. the epilogue past T).
//  * bfull[4] (leader), one completion per load into the group.  Waiter: owner(T) for the load of T, or -- resident
//    tile -- for the load that filled the group.  (lo) the previous load into the group served tile T - NG, same owner
//    because NG is even (b_groups() never returns 3): that issuer waited for it itself.  (hi) the next load into the
//    group is issued only after tile_done of its previous user, i.e. of T or of a later tile.
//    (The peer CTA's bytes may land before the leader arms the phase: its producer also waited for tile_done(T - NG),
//    so the previous phase is complete and the transaction count is just negative for a moment.)
//  * xfull / xempty[XS]: TWO full / empty rings of XS / 2 slots, one per transform warpgroup (even / odd chunks), each
//    with its own running index: a slot is only ever consumed by one warpgroup, which therefore waits for every use of
//    it in order.  (One ring shared by both warpgroups is NOT safe: the uses of a slot then alternate between them, a
//    warpgroup asks for use u without ever having waited for use u - 1, and when the loads land out of order -- a mix of
//    L2 hits and misses, e.g. on a freshly gathered copy -- it passes on the wrong phase and converts a chunk that has
//    not landed yet.  Measured before the split: about one row tile in 10^5 with wrong low-order scores, and now and
//    then a CTA that exited with a load still in flight -- "unspecified launch failure".)
//  * a_full[buffer][chunk] (leader; 4 local + 4 remote transform warps), one completion per use of the A buffer.
//    Waiters: BOTH issuers, in every row tile (their first own tile; with one tile per row tile the issuer without a tile
//    waits too) -> (lo).  (hi) the next use is written only after a_free of this use, which needs both issuers' commits,
//    which they make after their waits.
//  * a_free[4] (both CTAs, 2 arrivals = both issuers' commits after their last tile of the row tile).  Waiters: the
//    transform warps before rewriting the buffer.  (lo) they waited for the previous release before the previous write.
//    (hi) the next release needs MMAs on the data they are about to write.
//  * acc_empty[A] (leader; 8 local + 8 remote epilogue warps): owner(T) waits for the drain of T - A (buffer T mod A,
//    completion number T div A); the epilogue drains T only after tile_done(T).  One phase per use on both sides; the
//    two owners alternate on a buffer when A = 3, and each has seen the other's previous use complete through F2.
// Shapes that reach every branch: tools/stress.py (NT = 1, NT <= NG, NT > NG, odd / even row tiles per pair with a dummy
// last tile, one / two A buffers, every BN), tests/test_gpu_production_path.py and tests/test_gpu_parity.py.
template <int BN, int MODE, int NP = 3>
__global__ void __launch_bounds__(kPairH16Threads, 1)
assign_pair_h16_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_b,
                       const AssignParams p) {
    using Cfg = PairH16Cfg<BN>;
    constexpr bool DBG = MODE == 1, FAST = MODE == 2, ONE = NP == 1;
    // One-product FAST pass: the column index rides in the low mantissa bits of the (quantised) score, see the epilogue.
    constexpr bool PACK = FAST && ONE;
    int n_rows = p.n;
    if (p.n_dev != nullptr) {
        const int c = *p.n_dev;
        n_rows = c < p.n ? c : p.n;
    }
    const int num_row_tiles = (n_rows + kTileM - 1) / kTileM;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

    const int XS = p.stages & 255, NG = p.stages >> 8;
    const uint32_t XSW = XS >> 1;        // X ring slots per transform warpgroup
    const int NT = p.num_col_tiles;
    const int KC = p.num_k_chunks;       // 32-value chunks of X, <= 8 (one-product pass: <= 16)
    const bool WIDE = KC > 8;
    const int unit_bytes = WIDE ? Cfg::kBHalfBytes : Cfg::kBUnitBytes;
    const int C64 = (KC + 1) >> 1;       // 64-value units of B
    const int C64a = C64 >> 1;           // units [0, C64a) of A are released first in a row tile's last centroid tile
    const int ld64 = C64 * 64;
    const int group_bytes = C64 * unit_bytes;
    const int RES = NG < NT ? NG : NT;   // centroid tiles still resident in the B groups at a row-tile switch
    // ld <= 128: h1 / h2 of a row tile take at most 64 columns each, so there are TWO A buffers (row tile parity)
    // and the transform of the next row tile overlaps this one's MMAs completely -- with few centroid tiles per
    // row tile (k <= 128: one) the a_free -> transform -> a_full handshake would otherwise be the critical path.
    // (knockout bit 32, experiments: the one-product pass with two A buffers and two accumulators instead of one and three)
    const bool ALT = ONE && !WIDE && (p.knockout & 32);
    const bool ADB = KC <= 4 || ALT;
    const uint32_t abuf_cols = (ALT && KC > 4) ? 128 : 64;       // columns between the two A buffers
    // TMEM columns: three-product kernel: h1 [0, 128), h2 [128, 256), two accumulators [256, 512).  One-product pass,
    // ld <= 256: h1 alone in [0, 128) and THREE accumulators [128, 512) -- with two, the MMAs of tile T + 2 wait for the
    // epilogue of tile T, and its latency chain (commit -> wait -> tcgen05.ld -> 64 scores -> arrive -> wait) is about
    // two tiles of tensor time (round-2 ncu: tensor pipe 50 % active, the kernel ran at 41 ms without any epilogue
    // work against 59 ms with it, and 25 % fewer epilogue instructions bought 4 %); a third buffer gives the chain a
    // whole extra tile.  One-product pass, ld > 256: h1 [0, 256), two accumulators.
    const uint32_t NACC = (ONE && !WIDE && !ALT) ? 3u : 2u;
    const uint32_t acc_col0 = NACC == 3 ? 128u : 256u;
    uint8_t *xs = smem;                                  // [XS][128 rows x 128 B], swizzled
    uint8_t *bs = smem + XS * kXStageBytes;              // [NG][C64][g1 chunk | g2 chunk] (this CTA's rows)
    uint8_t *tail = bs + NG * group_bytes;
    float *cn_s = reinterpret_cast<float *>(tail);       // [2][BN]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail + 2 * BN * 4);
    uint64_t *bfull = bars;           // [4]  (leader) B group landed in BOTH CTAs
    uint64_t *tile_done = bars + 4;   // [8]  all MMAs of a centroid tile completed (multicast commit), ring over tiles
    uint64_t *xfull = bars + 12;      // [8]  X chunk landed
    uint64_t *xempty = bars + 20;     // [8]  X chunk read by its transform warpgroup
    uint64_t *a_full = bars + 28;     // [2][8]  h1/h2 of a 32-value chunk written (leader: its 4 warps + the peer's 4)
    uint64_t *a_free = bars + 44;     // [4]  units [0, C64a) / [C64a, C64) of A were read for the last time in this row
                                      //      tile (multicast); with two A buffers: [2 * buffer + 1], once per row tile
    uint64_t *acc_empty = bars + 48;  // [3]  accumulator drained (leader: 8 local + 8 remote warps)
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + Cfg::kNumBars);
    uint32_t *xchg = tmem_slot + 4;   // [2][5][128]

    // Roles by PHYSICAL warp: the scheduler prefers the highest warp id, so the warps whose latency matters (MMA
    // issuer, producers: 16-19; transform: 8-15) sit above the ALU-bound epilogue warps (0-7).  `warp` below is the
    // logical index the role tests use; warp & 3 (the TMEM lane quarter) is the same for both.
    const int pw = threadIdx.x >> 5;
    const int warp = pw >= 16 ? pw - 16 : pw >= 12 ? pw : pw >= 8 ? pw - 4 : pw >= 4 ? pw + 12 : pw + 8;
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
        for (int s = 0; s < 4; ++s) mbar_init(&bfull[s], 1);
        for (int s = 0; s < 8; ++s) mbar_init(&tile_done[s], 1);
        for (int s = 0; s < 8; ++s) {
            mbar_init(&xfull[s], 1);
            mbar_init(&xempty[s], 4);
            mbar_init(&a_full[s], leader ? 8 : 4);
            mbar_init(&a_full[8 + s], leader ? 8 : 4);
        }
        for (int s = 0; s < 4; ++s) mbar_init(&a_free[s], 2);   // both MMA issuers commit
        for (int b = 0; b < 3; ++b) mbar_init(&acc_empty[b], leader ? 16 : 8);   // two epilogue warpgroups per CTA
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
        uint32_t tile = 0, rtn = 0;
        for (int base = base0; base < num_row_tiles; base += gridDim.x, ++rtn) {
            for (int i = 0; i < NT; ++i, ++tile) {
                const int nt = (rtn & 1) ? NT - 1 - i : i;   // serpentine order, see the MMA issuer
                const int brow = nt * BN + (int)rank * Cfg::kBRows;
                const uint32_t g = nt % NG;
                // Before a LOAD: the MMAs of the tile that last used this group have completed.  A parity wait is only
                // meaningful while its barrier is within one phase of the awaited one, hence the ring of 8 and no wait
                // at resident tiles: the MMAs cannot pass a tile that needs this load, so tile q + 8 is never done
                // (NG <= 4 < 8), and at most RES <= 4 resident tiles lie between two loads, so tile q - 8 always is.
                // (Resident tiles do not depend on this warp at all -- with k <= NG tiles per row tile the MMAs may
                // run arbitrarily far ahead of it.)
                const bool resident = rtn > 0 && i < RES;
                if (!resident && tile >= (uint32_t)NG) {
                    const uint32_t q = tile - NG;
                    mbar_wait_backoff(&tile_done[q & 7], (q >> 3) & 1, kTransformSleepNs);   // (the ring gives this warp NG tiles of slack)
                }
                if (!resident && elect_one()) {
                    uint8_t *st = bs + g * group_bytes;
                    // both CTAs' bytes are accounted on the leader's barrier; the peer cannot run ahead of
                    // the ring (tile_done comes from the leader's MMAs), and bytes that land before the leader
                    // arms the phase only make its transaction count transiently negative
                    if (p.knockout & 1) {   // timing experiments only: no B traffic
                        if (leader) mbar_arrive(&bfull[g]);
                    } else {
                        if (leader) mbar_arrive_expect_tx(&bfull[g], ONE ? 2 * C64 * Cfg::kBHalfBytes : 2 * group_bytes);
                        for (int c = 0; c < C64; ++c) {
                            tma_load_2d_pair(&tm_b, &bfull[g], st + c * unit_bytes, c * 64, brow);
                            if constexpr (!ONE)
                                tma_load_2d_pair(&tm_b, &bfull[g], st + c * unit_bytes + Cfg::kBHalfBytes, ld64 + c * 64, brow);
                        }
                    }
                }
                __syncwarp();
            }
        }
    } else if (warp == 3) {
        // ================================ X producer ==================================
        uint32_t xj[2] = {0, 0};   // running chunk index of each transform warpgroup's ring: slot wg * XSW + xj % XSW
        for (int rt = tile0; rt - (int)rank < num_row_tiles; rt += gridDim.x) {
            for (int kc = 0; kc < KC; ++kc) {
                const uint32_t wg = kc & 1, j = xj[wg]++;
                const uint32_t s = wg * XSW + j % XSW, ph = (j / XSW) & 1;
                mbar_wait_backoff(&xempty[s], ph ^ 1, kTransformSleepNs);   // (XSW slots of slack)
                if (elect_one()) {
                    mbar_arrive_expect_tx(&xfull[s], kXStageBytes);
                    tma_load_2d(&tm_x, &xfull[s], xs + s * kXStageBytes, kc * kChunkK, rt * kTileM);
                }
                __syncwarp();
            }
        }
    } else if (warp == 1 || warp == 2) {
        if (leader) {
            // ================================ MMA issuers =============================
            // TWO issuing warps (logical warps 1 and 2), one per accumulator buffer: issuer q takes the centroid
            // tiles with tile % 2 == q.  Between two tiles an issuer spends > 1000 cycles in the tile's commit and
            // the next tile's waits while the (shallow) MMA queue drains; with two of them one is issuing while the
            // other waits.  tcgen05.commit only tracks the MMAs of the thread that executes it, hence: tile_done is
            // committed by the tile's owner, and everything that releases the A operand (a_free) is committed by
            // BOTH issuers (barrier count 2), each after its own last tile of the row tile.
            const uint32_t issuer = warp - 1;
            constexpr uint32_t idesc = make_idesc_f16(2 * kTileM, BN);
            uint32_t tile = 0, rtn = 0, bpar = 0;   // bpar: parity of the next load into each B group
            long long t_wait_acc = 0, t_wait = 0, t_issue = 0, t_mark = 0;
            const bool prof = p.role_prof != nullptr && blockIdx.x == 0 && issuer == 0;
            for (int base = base0; base < num_row_tiles; base += gridDim.x, ++rtn) {
                const uint32_t abuf = ADB ? (rtn & 1) : 0, apar = ADB ? (rtn >> 1) & 1 : rtn & 1;
                bool a_checked = false;   // this issuer has waited for the row tile's A operand
                for (int i = 0; i < NT; ++i, ++tile) {
                    // Serpentine order over the centroid tiles: odd row tiles walk them backwards, and centroid
                    // tile nt always lives in B group nt % NG -- so the last RES = min(NG, NT) tiles of one row tile
                    // are exactly the first RES of the next and are still in shared memory (no reload; with
                    // NT <= NG the whole B operand stays resident).  Which of two exactly tied columns is met
                    // first now depends on the direction; tied rows are re-checked in FP64 with the index rule.
                    const int nt = (rtn & 1) ? NT - 1 - i : i;
                    const uint32_t own = tile & 1, ab = tile % NACC, aph = (tile / NACC) & 1;
                    const uint32_t g = nt % NG;
                    const bool resident = rtn > 0 && i < RES;
                    const uint32_t bfull_par = resident ? ((bpar >> g) & 1) ^ 1 : (bpar >> g) & 1;   // resident: the load that filled it
                    if (!resident) bpar ^= 1u << g;
                    if (own != issuer) {
                        // not this issuer's tile; a row tile of ONE tile leaves this issuer without any MMA on its
                        // A operand, but the release barriers still expect its arrival
                        if (NT == 1) {
                            // (it also waits for the A operand like the owner does: an issuer that skipped a phase of
                            // the a_full barriers could pass its next parity wait on them too early)
                            for (int kc = 0; kc < KC; ++kc) mbar_wait(&a_full[abuf * 8 + kc], apar);
                            if (elect_one()) {
                                if (!ADB && C64a > 0) tc_commit_pair(&a_free[0]);
                                tc_commit_pair(&a_free[abuf * 2 + 1]);
                            }
                        }
                        __syncwarp();
                        continue;
                    }
                    if (prof) t_mark = clock64();
                    mbar_wait2(&acc_empty[ab], aph ^ 1, &bfull[g], bfull_par);
                    if (prof) { const long long t = clock64(); t_wait_acc += t - t_mark; t_mark = t; }
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc_col0 + ab * BN;
                    const uint32_t bsm0 = smem_u32(bs + g * group_bytes);
                    const bool last_tile = i == NT - 1;           // of the row tile
                    const bool my_last = last_tile || i == NT - 2;   // this issuer's last tile of the row tile
                    // one unit (64 K-values) of this tile: its MMAs and whatever is committed after them
                    auto issue_unit = [&](int c) {
                        const int nk = (2 * c + 1 < KC) ? 4 : 2;   // 16-value steps in this unit
                        const uint32_t bsm = bsm0 + c * unit_bytes;
                        const uint64_t d1 = make_kmajor_sw128_desc(bsm);
                        const uint64_t d2 = make_kmajor_sw128_desc(bsm + Cfg::kBHalfBytes);
                        const uint32_t a1 = tmem_base + abuf * abuf_cols + c * 32, a2 = a1 + Cfg::kA2Col0;
                        if constexpr (ONE) {
                            for (int ks = 0; ks < nk; ++ks)
                                mma_bf16_ts_pair(d_tmem, a1 + ks * 8, d1 + static_cast<uint64_t>(ks * 2), idesc, (c | ks) == 0 ? 0u : 1u);
                        } else {
                            // small terms first (knockout 8 / 16: timing experiments only, MMAs left out)
                            if (!(p.knockout & 8)) {
                                for (int ks = 0; ks < nk; ++ks)
                                    mma_bf16_ts_pair(d_tmem, a2 + ks * 8, d1 + static_cast<uint64_t>(ks * 2), idesc, (c | ks) == 0 ? 0u : 1u);
                                for (int ks = 0; ks < nk; ++ks)
                                    mma_bf16_ts_pair(d_tmem, a1 + ks * 8, d2 + static_cast<uint64_t>(ks * 2), idesc, 1u);
                            }
                            if (!(p.knockout & 16))
                                for (int ks = 0; ks < nk; ++ks)
                                    mma_bf16_ts_pair(d_tmem, a1 + ks * 8, d1 + static_cast<uint64_t>(ks * 2), idesc, 1u);
                        }
                        // Release of the A operand after this issuer's last tile of the row tile.  One A buffer: in
                        // two halves by the owner of the row tile's last tile (a release per unit measured 5 %
                        // slower: every commit stalls the issuer for a few hundred cycles), both halves at once by
                        // the other issuer.  Two A buffers: one release, the buffer has a row tile of slack.
                        if (my_last) {
                            if (!ADB && C64a > 0 && (last_tile ? c == C64a - 1 : c == C64 - 1)) tc_commit_pair(&a_free[0]);
                            if (c == C64 - 1) tc_commit_pair(&a_free[abuf * 2 + 1]);
                        }
                        if (c == C64 - 1) tc_commit_pair(&tile_done[tile & 7]);
                    };
                    if (a_checked) {
                        // the row tile's A operand is known to be complete: the whole tile in one go
                        if (prof) { const long long t = clock64(); t_wait += t - t_mark; t_mark = t; }
                        if (elect_one())
                            for (int c = 0; c < C64; ++c) issue_unit(c);
                        __syncwarp();
                        if (prof) { const long long t = clock64(); t_issue += t - t_mark; t_mark = t; }
                    } else {
                        for (int c = 0; c < C64; ++c) {
                            // this issuer's first tile of the row tile: the unit's h1/h2 must have been written
                            const uint32_t k0 = 2 * c, k1 = (2 * c + 1 < KC) ? 2 * c + 1 : 2 * c;
                            mbar_wait2(&a_full[abuf * 8 + k0], apar, &a_full[abuf * 8 + k1], apar);
                            tc_fence_after();
                            if (prof) {
                                const long long t = clock64();
                                t_wait += t - t_mark;
                                // (the issuer runs ahead of the tensor pipe by the depth of the MMA queue: time spent
                                // here is not necessarily time the pipe idles)
                                if (i == 0 && lane == 0 && c >= 1 && c <= 3) p.role_prof[12 + c] += t - t_mark;   // a_full wait per unit
                                t_mark = t;
                            }
                            if (elect_one()) issue_unit(c);
                            __syncwarp();
                            if (prof) { const long long t = clock64(); t_issue += t - t_mark; t_mark = t; }
                        }
                    }
                    a_checked = true;
                }
            }
            if (prof && lane == 0) {
                p.role_prof[0] = t_wait_acc;
                p.role_prof[1] = t_wait;
                p.role_prof[2] = 0;
                p.role_prof[3] = t_issue;
                p.role_prof[4] = tile;
            }
        }
    } else if ((warp >= 4 && warp < 8) || (warp >= 12 && warp < 16)) {
        // ================================ transform ===================================
        // once per row tile: two warpgroups, warps 4-7 take the even 32-value chunks, warps 12-15 the odd ones
        const int wg = warp >= 12 ? 1 : 0;
        const int r = threadIdx.x & 127;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const float sx = p.cstats[3];
        uint32_t rtn = 0, xj = 0;   // xj: running index of this warpgroup's X ring (same sequence as the producer's)
        long long t_x = 0, t_free = 0, t_store = 0, t_mark = 0, n_chunks = 0;
        const bool prof = p.role_prof != nullptr && blockIdx.x < 2 && warp == 4;
        for (int base = base0; base < num_row_tiles; base += gridDim.x, ++rtn) {
            for (int kc = wg; kc < KC; kc += 2, ++xj) {
                const uint32_t xslot = wg * XSW + xj % XSW;
                if (prof) t_mark = clock64();
                mbar_wait_parked(&xfull[xslot], (xj / XSW) & 1);
                const uint8_t *xrow = xs + xslot * kXStageBytes + r * 128;
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
                if (lane == 0) mbar_arrive(&xempty[xslot]);   // the chunk is in registers: the slot may be refilled
                uint32_t h1[16], h2[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const __half2 a = __floats2half2_rn(x[2 * i], x[2 * i + 1]);      // element 2i in the low half
                    const float2 af = __half22float2(a);
                    const __half2 b = __floats2half2_rn(x[2 * i] - af.x, x[2 * i + 1] - af.y);   // residuals are exact
                    h1[i] = *reinterpret_cast<const uint32_t *>(&a);
                    h2[i] = *reinterpret_cast<const uint32_t *>(&b);
                }
                // the last MMAs that read this chunk's TMEM columns have completed
                // (barriers of their own, one phase per use of the buffer: this warp gets here a whole row tile early,
                // and a parity wait is only meaningful within one phase of its barrier -- tile_done would alias)
                const uint32_t abuf = ADB ? (rtn & 1) : 0;
                if (prof) { const long long t = clock64(); t_x += t - t_mark; t_mark = t; }
                if (ADB) {
                    if (rtn >= 2) mbar_wait_backoff(&a_free[abuf * 2 + 1], ((rtn >> 1) - 1) & 1, kTransformSleepNs);
                } else if (rtn > 0) {
                    mbar_wait_backoff(&a_free[(kc >> 1) < C64a ? 0 : 1], (rtn - 1) & 1, kTransformSleepNs);
                }
                tc_fence_after();
                if (prof) {
                    const long long t = clock64();
                    t_free += t - t_mark;
                    t_mark = t;
                }
                if (p.knockout & 128) {   // timing experiments: the stores as usual, but of zeros
#pragma unroll
                    for (int i = 0; i < 16; ++i) h1[i] = 0u;
                }
                if (!(p.knockout & 2)) {
                    tmem_st16(tmem_base + lane_addr + abuf * abuf_cols + kc * 16, h1);
                    if constexpr (!ONE) tmem_st16(tmem_base + lane_addr + Cfg::kA2Col0 + abuf * 64 + kc * 16, h2);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(&a_full[abuf * 8 + kc]);
                    else mbar_arrive_remote_relaxed(&a_full[abuf * 8 + kc], 0);
                }
                if (prof) { const long long t = clock64(); t_store += t - t_mark; ++n_chunks; }
            }
        }
        if (prof && lane == 0) {
            long long *o = p.role_prof + 5 + 4 * blockIdx.x;
            o[0] = t_x;       // X chunk landed + read + converted
            o[1] = t_free;    // waiting for the TMEM columns
            o[2] = t_store;   // tcgen05.st + wait + arrive
            o[3] = n_chunks;
        }
    } else if ((warp >= 8 && warp < 12) || warp >= 16) {
        // ================================ epilogue ====================================
        // Two warpgroups, each drains half of the columns of every accumulator tile.  The TMEM loads are
        // double-buffered in batches of 16 columns and the centre norms of the NEXT tile are fetched while this one
        // is processed.  (The running top-3 update is 10 ALU + 1 FMA instructions per score; a form balanced 6 + 6
        // over the two pipes with predicated FMA-pipe moves ran the tile faster in cycles but, the kernel being
        // power-limited at this size, 1.5 % slower in time -- left out.)
        const int eg = warp >= 16 ? 1 : 0;
        const int r = threadIdx.x & 127;
        const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
        constexpr int HALF = BN >= 64 ? BN / 2 : BN;
        const int col0 = eg * HALF;
        const bool has_cols = col0 < BN;
        const float cmax = p.cstats[0];
        const float unscale = p.cstats[2];
        const bool stages_cn = eg == 0 && r < BN;
        // (PACK: the +inf norm of a padding column becomes a huge finite one -- a column byte inserted into the bits of
        //  +inf would make a NaN, and two NaNs in one pair would pull `second` down to `best`)
        constexpr float kPadNorm = 3.0e38f;
        if (stages_cn) cn_s[r] = PACK ? fminf(p.cnorm[r], kPadNorm) : p.cnorm[r];   // row tile 0 starts with centroid tile 0
        asm volatile("bar.sync 1, 256;" ::: "memory");   // both epilogue warpgroups
        uint32_t tile = 0, rtn = 0;
        for (int rt = tile0; rt - (int)rank < num_row_tiles; rt += gridDim.x, ++rtn) {
            // Four independent running top-3 sets (columns i mod 4).  Inside a set the columns arrive in ascending
            // order, so strict '<' keeps the lowest index; the sets are merged once per row tile with an explicit
            // (value, index) tie rule.  Indices are carried as floats (exact below 2^24).
            constexpr int NS = 4;
            const float inf = __int_as_float(0x7f800000);
            float best[NS], second[NS], third[NS], bidxf[NS], sidxf[NS];
#pragma unroll
            for (int q = 0; q < NS; ++q) { best[q] = second[q] = third[q] = inf; bidxf[q] = sidxf[q] = 0.f; }
            // ||x||^2 of this thread's row (near-tie bound at the end of the row tile): fetched now, off the critical path
            float xnsq = 0.f;
            if ((eg == 0 || PACK) && rt * kTileM + r < n_rows) xnsq = __ldg(p.xnorm + rt * kTileM + r);
            // PACK: the column's position in this warpgroup's half of the tile (0..63) replaces the low mantissa byte of
            // the score (one LOP3; the value moves by < 2^-16 of itself, which the near-tie bound below allows for).  The
            // minimum of such values carries its own arg-min, so the running update needs no compare and no select: per
            // PAIR of scores min, max, max, 3-input min, min -- 4.5 instructions per score instead of 6, 3.5 instead of 5
            // of them on the ALU pipe, which is the busiest unit of this kernel (68 % in the round-2 ncu capture).
            // Equal packed values of one row tile differ in the column byte; which of them wins does not matter, the
            // row is inside the bound and is decided later in FP64.
            float bprev[NS], btile[NS];
            const uint32_t keep = p.keep_mask;   // 0xffffff00, from a kernel parameter so that it lives in a register and
                                                 // "mask, then insert the column" is ONE 3-input LOP3 with the column as immediate
            if constexpr (PACK) {
#pragma unroll
                for (int q = 0; q < NS; ++q) { bprev[q] = inf; btile[q] = 0.f; }
            }
            for (int i = 0; i < NT; ++i, ++tile) {
                const int nt = (rtn & 1) ? NT - 1 - i : i;   // serpentine order, see the MMA issuer
                const uint32_t ab = tile % NACC, cb = tile & 1;   // accumulator buffer, buffer of the staged norms
                const float *cn = cn_s + cb * BN;
                float cn_next = 0.f;
                if (stages_cn) {
                    // the next tile in the walk (the next row tile starts where this one ends)
                    const int nn = i + 1 == NT ? nt : ((rtn & 1) ? nt - 1 : nt + 1);
                    cn_next = __ldg(p.cnorm + nn * BN + r);
                    if constexpr (PACK) cn_next = fminf(cn_next, kPadNorm);
                }
                mbar_wait_parked(&tile_done[tile & 7], (tile >> 3) & 1);
                tc_fence_after();
                if (has_cols && !(p.knockout & 4)) {
                    const uint32_t acc = tmem_base + lane_addr + acc_col0 + ab * BN + col0;
                    uint32_t va[16], vb[16];
                    auto consume = [&](const uint32_t (&v)[16], int c0) {
                        const float basef = (float)(nt * BN + c0);
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const int q = i & (NS - 1);
                            const float sc = fmaf(__uint_as_float(v[i]), unscale, cn[c0 + i]);
                            if constexpr (DBG) {
                                if (rt * kTileM + r < n_rows)
                                    p.dbg_scores[(size_t)(rt * kTileM + r) * p.kpad + nt * BN + c0 + i] = sc;
                            }
                            const float jf = basef + (float)i;
                            if constexpr (FAST) {   // 6 instructions per score instead of 11
                                const bool lt1 = sc < best[q];
                                second[q] = fminf(second[q], fmaxf(sc, best[q]));
                                bidxf[q] = lt1 ? jf : bidxf[q];
                                best[q] = fminf(best[q], sc);
                            } else {
                                const bool lt1 = sc < best[q], lt2 = sc < second[q];
                                third[q] = fminf(third[q], fmaxf(sc, second[q]));
                                sidxf[q] = lt1 ? bidxf[q] : (lt2 ? jf : sidxf[q]);
                                second[q] = fminf(second[q], fmaxf(sc, best[q]));
                                bidxf[q] = lt1 ? jf : bidxf[q];
                                best[q] = fminf(best[q], sc);
                            }
                        }
                    };
                    // PACK (see above): B = first column of the batch inside this warpgroup's half, a literal at every
                    // call so that the column numbers below are immediates
                    auto consume_pack = [&](const uint32_t (&v)[16], const int B) {
#pragma unroll
                        for (int i = 0; i < 16; i += 2) {
                            const int q = (i >> 1) & (NS - 1);
                            const uint32_t ua = __float_as_uint(fmaf(__uint_as_float(v[i]), unscale, cn[col0 + B + i]));
                            const uint32_t ub = __float_as_uint(fmaf(__uint_as_float(v[i + 1]), unscale, cn[col0 + B + i + 1]));
                            const float sa = __uint_as_float((ua & keep) | (uint32_t)(B + i));
                            const float sb = __uint_as_float((ub & keep) | (uint32_t)(B + i + 1));
                            const float lo = fminf(sa, sb), hi = fmaxf(sa, sb);
                            second[q] = fmin3(second[q], fmaxf(best[q], lo), hi);
                            best[q] = fminf(best[q], lo);
                        }
                    };
                    if constexpr (FAST) {
                        // The fast update is short enough that the drain is bound by the LATENCY of tcgen05.ld (one
                        // load per 16 columns, each issued after the previous wait): up to three loads in flight.
                        constexpr int NB = HALF / 16;   // 2 or 4 batches
                        uint32_t vc[16];
                        tmem_ld16(acc, va);
                        tmem_ld16(acc + 16, vb);
                        if constexpr (NB > 2) tmem_ld16(acc + 32, vc);
                        tmem_wait_ld();
                        if (p.knockout & 64) {   // timing experiments: the TMEM loads without the bookkeeping
                            if constexpr (NB > 3) { tmem_ld16(acc + 48, va); tmem_wait_ld(); }
                            best[0] = fminf(best[0], __uint_as_float(va[0] ^ vb[1] ^ vc[2]));
                        } else {
                        if constexpr (PACK) consume_pack(va, 0); else consume(va, col0);
                        if constexpr (NB > 3) tmem_ld16(acc + 48, va);
                        if constexpr (PACK) consume_pack(vb, 16); else consume(vb, col0 + 16);
                        if constexpr (NB > 2) { if constexpr (PACK) consume_pack(vc, 32); else consume(vc, col0 + 32); }
                        if constexpr (NB > 3) {
                            tmem_wait_ld();
                            if constexpr (PACK) consume_pack(va, 48); else consume(va, col0 + 48);
                        }
                        }
                        if constexpr (PACK) {   // which centroid tile the running minimum of every set comes from
#pragma unroll
                            for (int q = 0; q < NS; ++q) {
                                btile[q] = best[q] < bprev[q] ? (float)nt : btile[q];
                                bprev[q] = best[q];
                            }
                        }
                    } else {
                        tmem_ld16(acc, va);
#pragma unroll 1
                        for (int c0 = 0; c0 < HALF; c0 += 32) {
                            tmem_wait_ld();
                            tmem_ld16(acc + c0 + 16, vb);
                            consume(va, col0 + c0);
                            tmem_wait_ld();
                            if (c0 + 32 < HALF) tmem_ld16(acc + c0 + 32, va);
                            consume(vb, col0 + c0 + 16);
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(&acc_empty[ab]);
                    else mbar_arrive_remote_relaxed(&acc_empty[ab], 0);
                }
                if (stages_cn) cn_s[(cb ^ 1) * BN + r] = cn_next;   // last read one tile ago, before the previous bar.sync
                asm volatile("bar.sync 1, 256;" ::: "memory");
            }
            int bidx[NS], sidx[NS];
#pragma unroll
            for (int q = 0; q < NS; ++q) { bidx[q] = (int)bidxf[q]; sidx[q] = (int)sidxf[q]; }
            if constexpr (PACK) {
#pragma unroll
                for (int q = 0; q < NS; ++q)
                    bidx[q] = best[q] < inf ? (int)btile[q] * BN + col0 + (int)(__float_as_uint(best[q]) & 0xffu) : 0;
            }
            // merge sets 1..3 into set 0: lower value first, equal values -> lower index first
            float mb = best[0], ms = second[0], mt = third[0];
            int mbi = bidx[0], msi = sidx[0];
            auto insert = [&](float v, int j, bool has_index) {
                if (has_index && (v < mb || (v == mb && j < mbi))) { mt = ms; ms = mb; msi = mbi; mb = v; mbi = j; }
                else if (has_index && (v < ms || (v == ms && j < msi))) { mt = ms; ms = v; msi = j; }
                else if (v < mt) mt = v;
            };
            // FAST mode: (best, index, runner-up value) only
            auto merge2 = [&](float bq, float sq, int jq) {
                const bool win = bq < mb || (bq == mb && jq < mbi);
                ms = win ? fminf(mb, sq) : fminf(ms, bq);
                mbi = win ? jq : mbi;
                mb = win ? bq : mb;
            };
#pragma unroll
            for (int q = 1; q < NS; ++q) {
                if constexpr (FAST) {
                    merge2(best[q], second[q], bidx[q]);
                } else {
                    insert(best[q], bidx[q], true);
                    insert(second[q], sidx[q], true);
                    insert(third[q], 0, false);
                }
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
                if constexpr (FAST) {
                    merge2(__uint_as_float(xb[0 * 128 + r]), __uint_as_float(xb[1 * 128 + r]), (int)xb[3 * 128 + r]);
                } else {
                    insert(__uint_as_float(xb[0 * 128 + r]), (int)xb[3 * 128 + r], true);
                    insert(__uint_as_float(xb[1 * 128 + r]), (int)xb[4 * 128 + r], true);
                    insert(__uint_as_float(xb[2 * 128 + r]), 0, false);
                }
                const int row = rt * kTileM + r;
                if (row < n_rows && !(p.knockout & 4)) {
                    p.labels[row] = mbi;
                    const float xn = sqrtf(xnsq);
                    // relative part (accumulation, split residuals) + absolute part (FP16 subnormal range)
                    // (eps_x: the part of the error that is proportional to the cross term alone -- the operand
                    //  roundings of the one-product pass -- and not to the centre norm, which is added exactly)
                    float bound = 2.0f * (p.eps_x * 2.0f * xn * cmax + p.eps * (2.0f * xn * cmax + cmax * cmax) +
                                          p.sqrt_ld * (p.cstats[4] * 2.0f * cmax + p.cstats[5] * xn));
                    if constexpr (PACK) bound += 3.1e-5f * (2.0f * xn * cmax + cmax * cmax);   // 2 x 2^-16: the packed byte
                    if constexpr (FAST) {
                        // (in FAST mode `ms` is exact -- the merge below only needs values for it -- msi / mt are not)
                        if (ms - mb <= bound) {
                            const int pos = atomicAdd(p.cand_count, 1);   // may run past cand_cap: the host checks
                            if (pos < p.cand_cap) p.cand_list[pos] = row;
                        }
                    } else {
                        if (ms - mb <= bound) flag_row(p, row, mbi, msi, mt - mb <= bound);
                    }
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
