// Gram matrix of the (centred, column-scaled) table on the tensor cores: the contraction of the PCA path
// (the reference's pca::Distributed<step1Local, double, svdDense> of
// mllib-dal/src/main/native/PCADALImpl.cpp:64-71 reduces its block to a d x d factor on the host cores).
//
//   y_rc = (x_rc - mu_c) * s_c          mu_c: column mean (or 0), s_c: power of two with max_r |y_rc| in [2^13, 2^15)
//   y = h + l (+ <= 2^-22 |y|)          two-term FP16 split, as in assign_pair_h16.cuh
//   G^ = sum_r y_r y_r^T  ~=  H^T L + L^T H + H^T H        three kind::f16 products per 16 rows, FP32 accumulation
//   (diagonal panels: A = H^T (2 L) + H^T H and G^ = (A + A^T) / 2 -- two products)
//
// One CTA pair (tcgen05 cta_group::2) owns a 256-column panel pair (A columns [colA0, colA0 + 256) x B columns
// [colB0, colB0 + 256)) and a contiguous range of R-row stages (R = 64 on the diagonal, 32 off it); M = 256 output rows
// are split over the pair (CTA r:
// A columns colA0 + 128 r ...), and so are the N = 256 output columns of the B operand (CTA r supplies B columns
// colB0 + 128 r ... from its own shared memory).  On the diagonal (colA0 == colB0) A and B are the SAME shared-memory
// image, so every CTA reads only its own 128-column strip of X and the table is read from HBM once.
//
// Both operands are "MN-major": the contraction index is the table ROW, and a row-major [rows][columns] tile has the
// M / N index contiguous.  That is exactly the image a transform warp gets by converting the TMA-landed FP32 tile in
// place -- no transpose anywhere:
//   landing tile   [R rows][32 floats], 128-byte swizzle (TMA)           4 per strip and stage
//   operand tiles  H, L: [2 sub-tiles][R rows][64 halves], 128-byte swizzle:  canonical MN-major SW128 layout
//                  ((8,8,2),(8,k)) : ((1,8,LBO),(64,SBO)) halves with LBO = 128 R bytes (next 64 columns), SBO = 1024 B
//                  (next 8 rows); a K = 16 step advances the start address by 2048 B.
//
// Accumulation in three levels.  The tensor core adds into FP32 and TRUNCATES: measured on this kernel, the diagonal of G
// comes out low by 6e-9 of itself per accumulated row (1.5e-6 after 256 rows, 1.5e-5 after 2048), so
//   level 1  a TMEM accumulator only ever holds `flush_stages * R` rows (default 256; 160 with three products);
//   level 2  the flush warps then add it (FP32, round to nearest: no drift) into a second TMEM array;
//   level 3  every `promote_segs` segments that one is added into this pair's FP64 partial in global memory
//            (P[n][m], 256 x 256, lanes contiguous; no other writer: plain read-modify-write) and cleared.
// TMEM holds 512 columns = level 1 (256) + level 2 (256), so level 1 cannot be double-buffered in time.  It is split by
// COLUMNS instead: the N = 256 output columns are two halves of 128 (half h = sub-tile h of both CTAs' strips), each
// with its own MMA-issuing warp, its own four flush warps and segment boundaries half a segment apart -- while one half
// is being drained (TMEM reads run at 64 B / cycle: about 2000 cycles per drain) the tensor pipe works on the other.
//
// Warp roles (608 threads): 0-7 flush (warp & 3 = TMEM lane quarter, warp >> 2 = column half), 8-15 transform (two
// groups of four; sixteen warps ran slower), 16 TMA producer, 17 / 18 MMA issuers of half 0 / 1 (leader CTA).  Barriers:
// every ring is a plain full / free pair whose waiters wait for EVERY use in order (see the parity rule in assign_pair_h16.cuh):
//   land_full[LS] (tx) / land_free[LS] (the 4 transform warps of the stage's group)           per CTA
//   hl_full[HS] (leader: 4 local + 4 remote transform warps; both issuers wait) / hl_free[HS] (both issuers' multicast commits)
//   acc_full[half] (multicast commit) / acc_empty[half] (leader: 4 local + 4 remote flush warps)
#pragma once
#include <cuda_fp16.h>

#include "ptx_sm100.cuh"

namespace kmb {

constexpr int kGramThreads = 608;   // 19 warps: 8 flush, 8 transform, producer, 2 issuers
constexpr int kGramStrip = 128;                     // columns per CTA
// Rows per stage R (R / 16 K-steps): 64 on the diagonal panels (one strip per stage), 32 off the diagonal (two strips).
// Every stage costs a fixed ~1000 cycles of hand-offs in the transform warps (proxy fence, cluster-scope arrive) and a
// wait + commit in each issuer, so the rows per stage are what amortises them (measured with the knock-outs of
// tools/pca_check.py: at R = 32 the barrier skeleton alone, without any load, conversion or MMA, ran 40 % as long as the
// whole kernel).  Landing / operand ring depths are EVEN: the two transform groups take the even / the odd stages, and
// a ring slot must always be consumed by the same group (parity rule); they are launch parameters (p.ls, p.hs;
// defaults: gram_default_rings).
constexpr int kGramMaxRing = 8;
constexpr int kGramMaxBars = 4 * kGramMaxRing + 4;

struct GramParams {
    const float *scale;     // [ld2] s_c     (ld2 = ld rounded up to 256: entries beyond ld are 1 / 0)
    const float *offs;      // [ld2] -mu_c s_c
    double *partial;        // [pairs][256 n][256 m]
    int32_t n_rows;         // rows of the table
    int32_t stages_total;   // steps over all pairs: ceil(n_rows / (R << split))
    int32_t ls, hs;         // landing / operand ring depths (even, <= 8)
    int32_t flush_stages;   // stages per level-1 segment
    int32_t promote_segs;   // level-1 segments per level-3 promotion
    int32_t colA0, colB0;   // first column of the A / B panel (multiples of 256)
    int32_t split;          // diagonal panel of a narrow table: 1 (d <= 128) = the two CTAs of a pair take DIFFERENT stages,
                            // 2 (d <= 64) = and each CTA two of them, side by side in its strip; see the kernel
    int32_t desc_swap;      // debugging: exchange LBO and SBO in the operand descriptors
    long long *prof;        // optional cycle counters of pair 0's roles (diagnostics), else nullptr
    int32_t knockout;       // timing experiments (results invalid): 1 = no MMAs, 2 = no conversion, 4 = no level-2 / level-3 work, 8 = no TMA
};

static inline int gram_smem_bytes(bool same, int rows, int ls, int hs) {
    const int strips = same ? 1 : 2;
    return 1024 + strips * (ls + hs) * 4 * rows * 128 + kGramMaxBars * 8 + 16 + 2 * 2 * kGramStrip * 4;
}
// Default ring depths.  The landing ring needs few slots (the producer waits for a free slot 80 % of the time); the
// operand ring is what lets the issuer of one column half run ahead while the other waits for its accumulator.
static inline void gram_default_rings(bool same, int rows, int *ls, int *hs) {
    if (!same) { *ls = 2; *hs = 4; }
    else if (rows == 64) { *ls = 2; *hs = 4; }
    else { *ls = 4; *hs = 8; }
}

// MN-major operand, 128-byte swizzle (see the header): start address, LBO, SBO in 16-byte units
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
    d |= static_cast<uint64_t>((lbo >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>((sbo >> 4) & 0x3FFF) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}
__device__ __forceinline__ void mma_f16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// Eight consecutive columns of one row: FP32 (TMA-landed, swizzled) -> y = x s + o -> FP16 pair (h, l).  CHECK: the row
// may lie past the end of the table (TMA zero-filled it; with centring y would be the offset, not 0).
template <bool SAME, bool CHECK>
__device__ __forceinline__ void gram_convert8(const uint8_t *src, int q, int rr, const float (&sc)[8], const float (&of)[8], bool valid,
                                              uint4 &hv, uint4 &lv) {
    const float4 v0 = *reinterpret_cast<const float4 *>(src + (((2 * q) ^ rr) << 4));
    const float4 v1 = *reinterpret_cast<const float4 *>(src + (((2 * q + 1) ^ rr) << 4));
    const float x[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    uint32_t hw[4], lw[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float y0 = fmaf(x[2 * j], sc[2 * j], of[2 * j]), y1 = fmaf(x[2 * j + 1], sc[2 * j + 1], of[2 * j + 1]);
        if (CHECK) {
            y0 = valid ? y0 : 0.f;
            y1 = valid ? y1 : 0.f;
        }
        const __half2 a = __floats2half2_rn(y0, y1);
        const float2 af = __half22float2(a);
        // the residuals are exact; diagonal panels store 2 l (exact too): H^T H + H^T (2 L) is half of G^ + G^^T's
        // asymmetric part, and gram_reduce_kernel averages entry (i, j) with (j, i) -- two products instead of three
        const float l2 = SAME ? 2.f : 1.f;
        const __half2 b = __floats2half2_rn(l2 * (y0 - af.x), l2 * (y1 - af.y));
        hw[j] = *reinterpret_cast<const uint32_t *>(&a);
        lw[j] = *reinterpret_cast<const uint32_t *>(&b);
    }
    hv = make_uint4(hw[0], hw[1], hw[2], hw[3]);
    lv = make_uint4(lw[0], lw[1], lw[2], lw[3]);
}

// Level-1 segments of column half h: the first one of half 1 is half as long, so the two halves are never drained at
// the same time.  Stage g of nst ends a segment when it is the last one, or the f0 + i fs -th.
__device__ __forceinline__ int gram_first_segment(int h, int fs) { return (h == 1 && fs >= 2) ? fs / 2 : fs; }
__device__ __forceinline__ bool gram_segment_ends(int g, int nst, int f0, int fs) {
    return g + 1 == nst || (g + 1 >= f0 && (g + 1 - f0) % fs == 0);
}

template <bool SAME, int R>
__global__ void __launch_bounds__(kGramThreads, 1)
gram_pair_kernel(const __grid_constant__ CUtensorMap tm_x, const GramParams p) {
    constexpr int STRIPS = SAME ? 1 : 2;
    const int kGramLS = p.ls, kGramHS = p.hs;
    constexpr int kGramRows = R;
    constexpr int kGramLandBytes = R * 128;            // one [R x 32 floats] box
    constexpr int kGramSubBytes = R * 128;             // one [R x 64 halves] sub-tile
    constexpr int kGramOpBytes = 2 * kGramSubBytes;    // H (or L) of one strip and stage
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
    uint8_t *land = smem;                                                   // [LS][STRIPS][4 boxes]
    uint8_t *ops = land + kGramLS * STRIPS * 4 * kGramLandBytes;            // [HS][STRIPS][H | L][2 sub-tiles]
    uint8_t *tail = ops + kGramHS * STRIPS * 2 * kGramOpBytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail);
    uint64_t *land_full = bars, *land_free = bars + kGramLS;
    uint64_t *hl_full = bars + 2 * kGramLS, *hl_free = hl_full + kGramHS;
    uint64_t *acc_full = hl_free + kGramHS, *acc_empty = acc_full + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + kGramMaxBars);
    float *col_s = reinterpret_cast<float *>(tmem_slot + 4);               // [STRIPS][128] scale, then [STRIPS][128] offset

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair = blockIdx.x >> 1, pairs = gridDim.x >> 1;
    // this pair's stages: [st0, st1)
    const int st0 = (int)((long long)p.stages_total * pair / pairs);
    const int st1 = (int)((long long)p.stages_total * (pair + 1) / pairs);
    const int nst = st1 - st0;
    const int FS = p.flush_stages;
    // p.split (tables of at most 128 columns): with one strip the second CTA of a pair would only see padding.  Instead
    // both CTAs take columns [0, 128) of DIFFERENT stages (2 g and 2 g + 1).  The pair's MMA then produces
    // [A0^T B0 | A0^T B1] in CTA 0's lanes and [A1^T B0 | A1^T B1] in CTA 1's: the blocks on the diagonal are the two
    // stages' Gram matrices, the cross blocks are never read -- twice the rows per unit of tensor work.  split == 2
    // (at most 64 columns): each CTA also puts TWO stages side by side in its strip (sub-tile s = columns [0, 64) of its
    // stage s), four stages per pair; the useful blocks are then the four 64 x 64 ones on the diagonal.
    const int split = SAME ? p.split : 0;
    const int PS = 1 << split, CS = split == 2 ? 2 : 1;   // stages per pair and step, stages per CTA and step
    auto stage_of = [&](int g, int sub) { return (st0 + g) * PS + (split ? (int)rank * CS : 0) + sub; };
    const int colA = p.colA0 + (split ? 0 : (int)rank * kGramStrip), colB = p.colB0 + (int)rank * kGramStrip;

    if (threadIdx.x < STRIPS * kGramStrip) {
        const int s = threadIdx.x / kGramStrip, c = threadIdx.x % kGramStrip;
        const int col = (s == 0 ? colA : colB) + c;
        col_s[s * kGramStrip + c] = p.scale[col];
        col_s[STRIPS * kGramStrip + s * kGramStrip + c] = p.offs[col];
    }
    if (warp == 16 && lane == 0) tma_prefetch_desc(&tm_x);
    if (warp == 17 && lane == 0) {
        for (int s = 0; s < kGramLS; ++s) {
            mbar_init(&land_full[s], 1);
            mbar_init(&land_free[s], 4);
        }
        for (int s = 0; s < kGramHS; ++s) {
            mbar_init(&hl_full[s], leader ? 8 : 4);
            mbar_init(&hl_free[s], 2);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], leader ? 8 : 4);
        }
        fence_mbar_init();
    }
    cluster_sync_all();
    if (warp == 16) tmem_alloc_pair<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 16) {
        // ================================ TMA producer ================================
        const uint64_t evict_first = 0x12F0000000000000ull;
        const bool prof = p.prof != nullptr && blockIdx.x == 0;
        long long t_wait = 0, t0 = clock64();
        for (int g = 0; g < nst; ++g) {
            const uint32_t s = g % kGramLS, ph = (g / kGramLS) & 1;
            const long long tm = prof ? clock64() : 0;
            mbar_wait_parked(&land_free[s], ph ^ 1);
            if (prof) t_wait += clock64() - tm;
            if (elect_one()) {
                if (p.knockout & 8) {
                    mbar_arrive(&land_full[s]);
                } else {
                    mbar_arrive_expect_tx(&land_full[s], STRIPS * 4 * kGramLandBytes);
                    uint8_t *dst = land + s * STRIPS * 4 * kGramLandBytes;
                    const int row = stage_of(g, 0) * kGramRows;
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        if (split == 2) tma_load_2d_hint(&tm_x, &land_full[s], dst + b * kGramLandBytes, colA + 32 * (b & 1), stage_of(g, b >> 1) * kGramRows, evict_first);
                        else tma_load_2d_hint(&tm_x, &land_full[s], dst + b * kGramLandBytes, colA + 32 * b, row, evict_first);
                    }
                    if constexpr (!SAME) {
#pragma unroll
                        for (int b = 0; b < 4; ++b)
                            tma_load_2d_hint(&tm_x, &land_full[s], dst + (4 + b) * kGramLandBytes, colB + 32 * b, row, evict_first);
                    }
                }
            }
            __syncwarp();
        }
        if (prof && lane == 0) {
            p.prof[0] = clock64() - t0;   // producer: total, waiting for a free landing slot
            p.prof[1] = t_wait;
        }
    } else if (warp >= 17) {
        // ================================ MMA issuers =================================
        // warp 17 + h: column half h (N = 128: sub-tile h of both CTAs' B strips), level-1 accumulator at columns 128 h
        if (leader) {
            const int h = warp - 17;
            constexpr uint32_t idesc = make_idesc_f16(256, 128) | (1u << 15) | (1u << 16);   // A and B MN-major
            const uint32_t lbo = p.desc_swap ? 1024u : (uint32_t)kGramSubBytes, sbo = p.desc_swap ? (uint32_t)kGramSubBytes : 1024u;
            const int F0 = gram_first_segment(h, FS);
            const uint32_t d_tmem = tmem_base + h * 128;
            uint32_t seg = 0;
            bool seg_start = true;
            const bool prof = p.prof != nullptr && blockIdx.x == 0;
            long long t_acc = 0, t_full = 0, t0 = clock64();
            for (int g = 0; g < nst; ++g) {
                const uint32_t s = g % kGramHS, ph = (g / kGramHS) & 1;
                long long tm = prof ? clock64() : 0;
                if (seg_start && seg >= 1) mbar_wait(&acc_empty[h], (seg - 1) & 1);
                if (prof) { const long long t = clock64(); t_acc += t - tm; tm = t; }
                mbar_wait(&hl_full[s], ph);
                if (prof) t_full += clock64() - tm;
                tc_fence_after();
                const bool seg_end = gram_segment_ends(g, nst, F0, FS);
                if (elect_one()) {
                    const uint32_t a_h = smem_u32(ops + s * STRIPS * 2 * kGramOpBytes), a_l = a_h + kGramOpBytes;
                    const uint32_t b_h = (SAME ? a_h : a_h + 2 * kGramOpBytes) + h * kGramSubBytes, b_l = b_h + kGramOpBytes;
#pragma unroll
                    for (int ks = 0; ks < ((p.knockout & 1) ? 0 : kGramRows / 16); ++ks) {
                        const uint32_t off = ks * 16 * 128;
                        const uint64_t dah = make_mnmajor_sw128_desc(a_h + off, lbo, sbo), dal = make_mnmajor_sw128_desc(a_l + off, lbo, sbo);
                        const uint64_t dbh = make_mnmajor_sw128_desc(b_h + off, lbo, sbo), dbl = make_mnmajor_sw128_desc(b_l + off, lbo, sbo);
                        mma_f16_ss_pair(d_tmem, dah, dbl, idesc, (seg_start && ks == 0) ? 0u : 1u);
                        if constexpr (!SAME) mma_f16_ss_pair(d_tmem, dal, dbh, idesc, 1u);   // (diagonal panels: see L2 below)
                        mma_f16_ss_pair(d_tmem, dah, dbh, idesc, 1u);
                    }
                    tc_commit_pair(&hl_free[s]);
                    if (seg_end) tc_commit_pair(&acc_full[h]);
                }
                __syncwarp();
                seg_start = seg_end;
                if (seg_end) ++seg;
            }
            if (prof && lane == 0) {
                p.prof[4 + 4 * h] = clock64() - t0;   // issuer h: total, waiting for its accumulator, waiting for operands
                p.prof[5 + 4 * h] = t_acc;
                p.prof[6 + 4 * h] = t_full;
            }
        }
    } else if (warp >= 8) {
        // ================================ transform ===================================
        // Two groups of four warps: group (tw >> 2) converts the stages of that parity, so a warp has two stage times
        // for the latency chain wait -> LDS -> convert -> wait -> STS -> fence -> (remote) arrive.  Warp: box tw & 3 (32
        // columns of the strip), all R rows in steps of eight; lane: row lane & 7, 8-column group lane >> 3.  A
        // quarter warp (one column group, eight rows) reads 16-byte chunks (2q) ^ row and writes chunk
        // ((box & 1) * 4 + q) ^ row of eight different rows: eight different bank groups both ways.
        const int tw = warp - 8;
        const int box = tw & 3, grp = tw >> 2;
        const int rr = lane & 7, q = lane >> 3;
        float sc[STRIPS][8], of[STRIPS][8];
#pragma unroll
        for (int s = 0; s < STRIPS; ++s)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int cpos = (split == 2 ? (box & 1) : box) * 32 + q * 8 + j;   // column of the strip this value belongs to
                sc[s][j] = col_s[s * kGramStrip + cpos];
                of[s][j] = col_s[STRIPS * kGramStrip + s * kGramStrip + cpos];
            }
        const uint32_t sub = box >> 1, chunk = (box & 1) * 4 + q;
        const bool prof = p.prof != nullptr && blockIdx.x < 2 && box == 0;
        long long t_land = 0, t_free = 0, t_fence = 0, t0 = clock64();
        for (int g = grp; g < nst; g += 2) {
            const uint32_t ls = g % kGramLS, lph = (g / kGramLS) & 1, hs = g % kGramHS;
            const int rows_left = p.n_rows - stage_of(g, split == 2 ? box >> 1 : 0) * kGramRows;   // rows of this warp's stage inside the table
            const bool full = rows_left >= kGramRows;
            long long tm = prof ? clock64() : 0;
            mbar_wait_parked(&land_full[ls], lph);
            if (prof) t_land += clock64() - tm;
#pragma unroll
            for (int s = 0; s < STRIPS; ++s) {
#pragma unroll
                for (int pass = 0; pass < R / 32; ++pass) {
                    uint4 hv[4], lv[4];
#pragma unroll
                    for (int it = 0; it < ((p.knockout & 2) ? 0 : 4); ++it) {
                        const int row = pass * 32 + it * 8 + rr;
                        const uint8_t *src = land + ((ls * STRIPS + s) * 4 + box) * kGramLandBytes + row * 128;
                        if (full) gram_convert8<SAME, false>(src, q, rr, sc[s], of[s], true, hv[it], lv[it]);
                        else gram_convert8<SAME, true>(src, q, rr, sc[s], of[s], row < rows_left, hv[it], lv[it]);
                    }
                    if (s == STRIPS - 1 && pass == R / 32 - 1) {
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&land_free[ls]);   // the stage's tiles are in registers
                    }
                    // the MMAs that read the previous contents of this operand slot have completed
                    if (s == 0 && pass == 0 && g >= kGramHS) {
                        tm = prof ? clock64() : 0;
                        mbar_wait_parked(&hl_free[hs], ((g / kGramHS) - 1) & 1);
                        if (prof) t_free += clock64() - tm;
                    }
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int row = pass * 32 + it * 8 + rr;
                        uint8_t *dst = ops + (hs * STRIPS + s) * 2 * kGramOpBytes + sub * kGramSubBytes + row * 128 + ((chunk ^ rr) << 4);
                        *reinterpret_cast<uint4 *>(dst) = hv[it];
                        *reinterpret_cast<uint4 *>(dst + kGramOpBytes) = lv[it];
                    }
                }
            }
            tm = prof ? clock64() : 0;
            fence_proxy_async();   // generic-proxy stores -> visible to the tensor core's reads
            __syncwarp();
            if (lane == 0) {
                if (leader) mbar_arrive(&hl_full[hs]);
                else mbar_arrive_remote_cta(&hl_full[hs], 0);
            }
            if (prof) t_fence += clock64() - tm;
        }
        if (prof && lane == 0) {
            long long *o = p.prof + 16 + 8 * blockIdx.x + 4 * grp;   // transform (CTA, group): total, landing wait, operand-slot wait, fence + arrive
            o[0] = clock64() - t0;
            o[1] = t_land;
            o[2] = t_free;
            o[3] = t_fence;
        }
    } else {
        // ================================ flush =======================================
        // warp & 3: TMEM lane quarter (output rows m), warp >> 2: column half.  TMEM column c of half h is output column
        // (c >> 6) * 128 + 64 h + (c & 63) of the panel (columns 0-63 come from CTA 0's sub-tile h, 64-127 from CTA 1's).
        // (Tried and slower: all eight warps draining whichever half completes next, level 1 read first and released at
        // once -- the drains of the two halves then queue behind each other; 16-column pieces for a 72-register budget.)
        const int qd = warp & 3, h = warp >> 2;
        const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
        const int F0 = gram_first_segment(h, FS);
        const int nseg = nst <= F0 ? 1 : 1 + (nst - F0 + FS - 1) / FS;
        double *P = p.partial + (size_t)pair * 65536 + (int)rank * 128 + qd * 32 + lane;
        const uint32_t l1 = tmem_base + lane_addr + h * 128, l2 = l1 + 256;
        const bool prof = p.prof != nullptr && blockIdx.x == 0 && qd == 0;
        long long t_wait = 0, t0 = clock64();
        for (int seg = 0; seg < nseg; ++seg) {
            const long long tm = prof ? clock64() : 0;
            mbar_wait_parked(&acc_full[h], seg & 1);
            if (prof) t_wait += clock64() - tm;
            tc_fence_after();
            const bool promote = (seg + 1) % p.promote_segs == 0 || seg + 1 == nseg;
            const bool fresh = seg % p.promote_segs == 0;   // level 2 was cleared (or never written)
            // split: only the diagonal blocks are read (columns from this CTA's own B; split == 2: and only by the lane
            // half whose rows belong to column half h)
            const bool useful = split != 2 || (qd >> 1) == h;
            const int ch0 = (p.knockout & 4) ? 3 : split ? 2 * (int)rank : 0, ch1 = !useful ? 0 : split ? 2 * (int)rank + 2 : 4;
            if (ch0 >= ch1) {   // nothing to read: release only
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(&acc_empty[h]);
                    else mbar_arrive_remote_relaxed(&acc_empty[h], 0);
                }
            }
#pragma unroll 1
            for (int ch = ch0; ch < ch1; ++ch) {
                uint32_t a[32], b[32];
                tmem_ld32(l1 + ch * 32, a);
                if (!fresh) tmem_ld32(l2 + ch * 32, b);
                tmem_wait_ld();
                if (ch == ch1 - 1) {   // level 1 has been read: the MMAs of the next segment may overwrite it
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (leader) mbar_arrive(&acc_empty[h]);
                        else mbar_arrive_remote_relaxed(&acc_empty[h], 0);   // (TMEM reads: ordered by the tcgen05 fence, no memory to publish)
                    }
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float v = __uint_as_float(a[j]) + (fresh ? 0.f : __uint_as_float(b[j]));
                    b[j] = __float_as_uint(v);
                }
                if (promote) {
                    const int c0 = ch * 32;
                    double *dst = P + (size_t)((c0 >> 6) * 128 + 64 * h + (c0 & 63)) * 256;
#pragma unroll
                    for (int j0 = 0; j0 < 32; j0 += 8) {
                        double acc[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[j] = dst[(size_t)(j0 + j) * 256];
#pragma unroll
                        for (int j = 0; j < 8; ++j) dst[(size_t)(j0 + j) * 256] = acc[j] + (double)__uint_as_float(b[j0 + j]);
                    }
                } else {
                    tmem_st32(l2 + ch * 32, b);
                }
            }
            tmem_wait_st();
        }
        if (prof && lane == 0) {
            p.prof[12 + 2 * h] = clock64() - t0;   // flush warps of half h: total, waiting for a complete accumulator
            p.prof[13 + 2 * h] = t_wait;
        }
    }
    // teardown: every MMA has completed (the flush warps waited for the last acc_full) before TMEM goes away
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 16) tmem_dealloc_pair<512>(tmem_base);
}

// ------------------------------------------------------------------ the passes around it
// Column sums (FP64) and column max |x| of a row range per block: psum [grid][ld], pmax [grid][ld].  Thread t handles the
// float4 column group t % G of every RL-th row; the row lanes are combined in a fixed order (bit-reproducible sums).
__global__ void __launch_bounds__(256) colstats_kernel(const float *__restrict__ X, int64_t n, int ld, double *__restrict__ psum,
                                                       float *__restrict__ pmax) {
    __shared__ double ssum[256 * 4];
    __shared__ float smax[256 * 4];
    const int groups = ld >> 2;
    const int64_t rows_per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per, r1 = r0 + rows_per < n ? r0 + rows_per : n;
    for (int g0 = 0; g0 < groups; g0 += 256) {
        const int G = groups - g0 < 256 ? groups - g0 : 256, RL = 256 / G;
        const int g = threadIdx.x % G, rl = threadIdx.x / G;
        double s[4] = {0.0, 0.0, 0.0, 0.0};
        float m[4] = {0.f, 0.f, 0.f, 0.f};
        if (rl < RL) {
            auto add = [&](const float4 v) {
                s[0] += v.x; s[1] += v.y; s[2] += v.z; s[3] += v.w;
                m[0] = fmaxf(m[0], fabsf(v.x)); m[1] = fmaxf(m[1], fabsf(v.y));
                m[2] = fmaxf(m[2], fabsf(v.z)); m[3] = fmaxf(m[3], fabsf(v.w));
            };
            int64_t r = r0 + rl;
            for (; r + 7 * (int64_t)RL < r1; r += 8 * (int64_t)RL) {   // eight independent loads in flight per thread
                float4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __ldcs(reinterpret_cast<const float4 *>(X + (r + u * (int64_t)RL) * ld) + g0 + g);
#pragma unroll
                for (int u = 0; u < 8; ++u) add(v[u]);
            }
            for (; r < r1; r += RL) add(__ldcs(reinterpret_cast<const float4 *>(X + r * ld) + g0 + g));
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            ssum[threadIdx.x * 4 + j] = s[j];
            smax[threadIdx.x * 4 + j] = m[j];
        }
        __syncthreads();
        if (rl == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                double a = 0.0;
                float b = 0.f;
                for (int l = 0; l < RL; ++l) {
                    a += ssum[(l * G + g) * 4 + j];
                    b = fmaxf(b, smax[(l * G + g) * 4 + j]);
                }
                psum[(size_t)blockIdx.x * ld + (g0 + g) * 4 + j] = a;
                pmax[(size_t)blockIdx.x * ld + (g0 + g) * 4 + j] = b;
            }
        }
        __syncthreads();
    }
}
// out: sum[ld] then (double) row count of this rank; omax[ld].  One block per 8 columns, 32 lanes over the partial rows,
// combined in a fixed order.
__global__ void __launch_bounds__(256) colstats_reduce_kernel(const double *__restrict__ psum, const float *__restrict__ pmax, int blocks,
                                                              int ld, int64_t n, double *__restrict__ osum, float *__restrict__ omax) {
    __shared__ double ss[32][8];
    __shared__ float sm[32][8];
    const int c = blockIdx.x * 8 + (threadIdx.x & 7), l = threadIdx.x >> 3;
    double a = 0.0;
    float b = 0.f;
    if (c < ld) {
        for (int i = l; i < blocks; i += 32) {
            a += psum[(size_t)i * ld + c];
            b = fmaxf(b, pmax[(size_t)i * ld + c]);
        }
    }
    ss[l][threadIdx.x & 7] = a;
    sm[l][threadIdx.x & 7] = b;
    __syncthreads();
    if (l == 0 && c < ld) {
        a = 0.0;
        b = 0.f;
        for (int i = 0; i < 32; ++i) {
            a += ss[i][threadIdx.x & 7];
            b = fmaxf(b, sm[i][threadIdx.x & 7]);
        }
        osum[c] = a;
        omax[c] = b;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) osum[ld] = (double)n;
}
// running statistics (+)= one batch's (kmb200_table::colsum / colmax); sums [ld + 1] with the row count last
__global__ void colstats_merge_kernel(const double *__restrict__ tsum, const float *__restrict__ tmax, int ld, int reset,
                                      double *__restrict__ colsum, float *__restrict__ colmax) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c <= ld) colsum[c] = (reset ? 0.0 : colsum[c]) + tsum[c];
    if (c < ld) colmax[c] = fmaxf(reset ? 0.f : colmax[c], tmax[c]);
}
// Per column, from the GLOBAL sums / maxima: mu = centre ? float(sum / N) : 0; s = 2^(13 - floor(log2(max|x| + |mu|)));
// scale = s, offs = -mu s, inv = 1 / s (double), mean_used = mu (double).  Columns [ld, ld2) get scale 1, offset 0.
__global__ void pca_prep_kernel(const double *__restrict__ gsum, const float *__restrict__ gmax, int ld, int ld2, int center,
                                float *__restrict__ scale, float *__restrict__ offs, double *__restrict__ inv, double *__restrict__ mean_used) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ld2) return;
    if (c >= ld) {
        scale[c] = 1.f;
        offs[c] = 0.f;
        return;
    }
    const double nrows = gsum[ld];
    const float mu = center ? (float)(gsum[c] / nrows) : 0.f;
    const float a = gmax[c] + fabsf(mu);
    int e = 0;
    if (a > 0.f && a < 3.0e38f) {
        e = ilogbf(a);
        e = e < -100 ? -100 : e > 100 ? 100 : e;
    }
    const float s = (a > 0.f) ? ldexpf(1.f, 13 - e) : 1.f;
    scale[c] = s;
    offs[c] = -mu * s;
    inv[c] = 1.0 / (double)s;
    mean_used[c] = (double)mu;
}
// G[i][j] (i in the A panel, j in the B panel) = inv_i inv_j * sum over pairs of P[pair][j][i]; written into the full
// [ld x ld] matrix (and mirrored, so that the panel pairs above the diagonal fill the whole matrix).  Fixed order.
// split: the pair's two (four) stages left their sums in the 128 x 128 (64 x 64) blocks on the diagonal of P (see the kernel).
__global__ void gram_reduce_kernel(const double *__restrict__ partial, int pairs, const double *__restrict__ inv, int ld, int colA0,
                                   int colB0, int split, double *__restrict__ G) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;   // 0..255: lanes contiguous in P
    const int nn = blockIdx.y;                             // 0..255
    const int i = colA0 + m, j = colB0 + nn;
    if (m >= 256 || i >= ld || j >= ld) return;
    auto entry = [&](int r, int c) {   // sum over the pairs of P[c][r] (and of the second stage's block)
        double a = 0.0;
        for (int q = 0; q < pairs; ++q) {
            a += partial[(size_t)q * 65536 + (size_t)c * 256 + r];
            if (split == 1) a += partial[(size_t)q * 65536 + (size_t)(128 + c) * 256 + 128 + r];
            if (split == 2)
                for (int b = 1; b < 4; ++b) a += partial[(size_t)q * 65536 + (size_t)(64 * b + c) * 256 + 64 * b + r];
        }
        return a * (inv[i] * inv[j]);
    };
    const double a = entry(m, nn);
    if (colA0 == colB0) {
        // diagonal panel: entry (i, j) and (j, i) of the product differ only in the order of the FP32 additions (and, with
        // two products, A = H^T (2 L) + H^T H is only half of the sum): average
        G[(size_t)i * ld + j] = 0.5 * (a + entry(nn, m));
    } else {
        G[(size_t)i * ld + j] = a;
        G[(size_t)j * ld + i] = a;
    }
}

}  // namespace kmb
