// One-pass Lloyd step for the HBM-bound low-k regime (BASELINE config 5: 400M x 32, k = 16): when k*ld is
// small the distance work is a few hundred FMAs per row and the whole iteration -- distances, argmin,
// per-cluster FP64 sums, counts and the direct-form cost -- fits in ONE read of X on the CUDA cores.  The
// tensor-core path needs two reads (assign, then accumulate) plus bucketing; here the bound is N*d*4 bytes at
// HBM bandwidth.  Same semantics as the other path: first minimum wins, near-ties (top-2 gap inside the FP32
// error bound) are re-scored in FP64 before the row is accumulated.
//
// Per warp, batches of 32 rows: coalesced float4 loads -> padded shared-memory tile -> thread t owns row t for
// the distance phase (x in registers, centres broadcast from shared memory) -> lanes own columns for the
// accumulation phase (FP64 partial sums per warp in shared memory, no atomics).  Per-CTA partials go to global
// memory and are reduced in a fixed order by fused_reduce_kernel into the [k*ld sums | k counts | cost] buffer.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels_misc.cuh"

namespace kmb {

constexpr int kFusedWarps = 8;
// Centres of the current iteration in constant memory (FP32, padded to KP rows; norm = +inf for padding):
// with the centre loop fully unrolled they become immediate constant-bank operands of the FMAs, so the
// distance phase touches neither shared memory nor the load/store unit.  Refreshed (device-to-device) by
// the host wrapper before every launch.
__constant__ float kc_cent[1024];
__constant__ float kc_norm[32];
constexpr int kFusedMaxKLd = 1024;   // k * ld <= 1024: 8 KB of FP64 partials per warp

__host__ __device__ inline size_t fused_smem_bytes(int k, int ld) {
    const size_t cen = (size_t)(k + 1) * ld * 4 + (size_t)(k + 4) * 4;  // FP32 centres + norms (padded to even k)
    const size_t tile = (size_t)kFusedWarps * 32 * (ld + 4) * 4;        // per-warp X tile, padded rows
    const size_t part = (size_t)kFusedWarps * ((size_t)k * ld + k) * 8; // per-warp FP64 sums + counts
    return cen + tile + part + 64;
}

// Accumulation of one 32-row batch (rows in `tile`, label of row `lane` in bidx; rows >= nrows do not exist): lanes own
// columns.  The rows are first ordered by label (a counting sort inside the warp: __match_any_sync gives a row's rank among
// the rows of its label, a 32-entry shared-memory histogram and one warp scan give the labels' offsets), then walked in
// that order by a fully unrolled loop: the running FP64 sum of a column stays in a register while the label does not
// change and goes to the shared-memory partial once per (label, batch).  No data-dependent loop control: the only branch
// is the warp-uniform "label changed" test.  (Round 1 looped over the distinct labels with ballot / ffs loops: about 1000
// of the 1860 instructions of a batch.)  scratch: 72 ints of this warp.
template <int LD, int STRIDE>
__device__ __forceinline__ void accumulate_sorted(const float *tile, const float *c32, double *mypart, int k, int nrows,
                                                  int bidx, int lane, int *scratch, double &cost) {
    int *sc = scratch, *sp = scratch + 40;
    const int key = lane < nrows ? bidx : 32;           // rows beyond the table sort last
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int lower = __popc(peers & ((1u << lane) - 1u));
    sc[lane] = 0;
    __syncwarp();
    if (lower == 0 && key < 32) sc[key] = __popc(peers);   // one lane per distinct label
    __syncwarp();
    const int cnt = sc[lane];                             // rows of label `lane` in this batch
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    __syncwarp();
    sc[lane] = inc - cnt;                                 // first sorted position of label `lane`
    if (lane == 0) sc[32] = nrows;
    __syncwarp();
    sp[sc[key] + lower] = lane | (key << 8);
    if (lane < k && cnt > 0) mypart[(size_t)k * LD + lane] += (double)cnt;
    __syncwarp();
    const int mine = sp[lane];                            // the row at sorted position `lane`, and its label
    int lprev = -1;
    double s0 = 0.0, s1 = 0.0;
    float q0 = 0.f, q1 = 0.f, c0 = 0.f, c1 = 0.f;
#pragma unroll
    for (int r = 0; r < 32; ++r) {
        const int e = __shfl_sync(0xffffffffu, mine, r);
        const int pr = e & 0xff, lr = e >> 8;
        if (lr != lprev) {                                // warp-uniform
            if (lprev >= 0) {
                double *dst = mypart + (size_t)lprev * LD;
                dst[lane] += s0;
                if (LD > 32) dst[32 + lane] += s1;
            }
            lprev = lr;
            if (lr >= 32) { lprev = -1; break; }          // only rows beyond the table follow
            s0 = s1 = 0.0;
            c0 = c32[(size_t)lr * LD + lane];
            if (LD > 32) c1 = c32[(size_t)lr * LD + 32 + lane];
        }
        const float v0 = tile[pr * STRIDE + lane];
        s0 += (double)v0;
        q0 = fmaf(v0 - c0, v0 - c0, q0);
        if (LD > 32) {
            const float v1 = tile[pr * STRIDE + 32 + lane];
            s1 += (double)v1;
            q1 = fmaf(v1 - c1, v1 - c1, q1);
        }
    }
    if (lprev >= 0) {
        double *dst = mypart + (size_t)lprev * LD;
        dst[lane] += s0;
        if (LD > 32) dst[32 + lane] += s1;
    }
    cost += (double)q0 + (double)q1;
}

template <int LD, int KP>
__global__ void __launch_bounds__(kFusedWarps * 32)
fused_smallk_kernel(const float *__restrict__ X, const double *__restrict__ cent, const float *__restrict__ cstats,
                    int64_t n, int k, float eps, int32_t *__restrict__ labels, double *__restrict__ cta_part,
                    unsigned long long *__restrict__ recheck_counter) {
    extern __shared__ __align__(16) uint8_t fsm[];
    __shared__ int s_sort[kFusedWarps][72];   // per warp: [0..32] label counts / offsets, [40..71] rows in label order
    constexpr int STRIDE = LD + 4;   // floats per tile row (16-byte aligned, spreads the banks)
    const int k2 = k;
    float *c32 = reinterpret_cast<float *>(fsm);                 // [k2][LD]
    float *cn32 = c32 + (size_t)k2 * LD;                         // [k2]
    float *tiles = cn32 + ((k2 + 3) & ~3);                       // [warps][32][STRIDE]
    double *parts = reinterpret_cast<double *>(tiles + (size_t)kFusedWarps * 32 * STRIDE);   // [warps][k*LD + k]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int psize = k * LD + k;

    for (int i = threadIdx.x; i < k2 * LD; i += blockDim.x) c32[i] = i < k * LD ? (float)cent[i] : 0.f;
    for (int j = threadIdx.x; j < k2; j += blockDim.x) {
        double s = 0.0;
        if (j < k)
            for (int c = 0; c < LD; ++c) s += cent[(size_t)j * LD + c] * cent[(size_t)j * LD + c];
        cn32[j] = j < k ? (float)s : __int_as_float(0x7f800000);
    }
    double *mypart = parts + (size_t)w * psize;
    for (int i = lane; i < psize; i += 32) mypart[i] = 0.0;
    __syncthreads();
    const float cmax = cstats[0];
    float *tile = tiles + (size_t)w * 32 * STRIDE;
    double cost = 0.0;
    unsigned long long rechecked = 0;

    const int64_t nbatches = (n + 31) / 32;
    const int64_t gw = (int64_t)blockIdx.x * kFusedWarps + w, gstride = (int64_t)gridDim.x * kFusedWarps;
    for (int64_t b = gw; b < nbatches; b += gstride) {
        const int64_t row0 = b * 32;
        const int nrows = (int)min((int64_t)32, n - row0);
        // ---- 1. coalesced load of 32 rows into the padded tile
        constexpr int VPR = LD / 4;            // float4 per row
        constexpr int RPI = 32 / VPR;          // rows per warp-wide load instruction
#pragma unroll
        for (int it = 0; it < 32 / RPI; ++it) {
            const int r = it * RPI + lane / VPR, q = lane % VPR;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < nrows) v = __ldg(reinterpret_cast<const float4 *>(X + (row0 + r) * LD) + q);
            *reinterpret_cast<float4 *>(tile + r * STRIDE + q * 4) = v;
        }
        __syncwarp();
        // ---- 2. thread t scores row t against every centre (FP32 FMA chain), top-2 tracking
        float x[LD];
#pragma unroll
        for (int q = 0; q < VPR; ++q) {
            const float4 v = *reinterpret_cast<const float4 *>(tile + lane * STRIDE + q * 4);
            x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
        }
        float xn = 0.f;
#pragma unroll
        for (int c = 0; c < LD; ++c) xn = fmaf(x[c], x[c], xn);
        float best = __int_as_float(0x7f800000), second = best;
        int bidx = 0;
        // fully unrolled over the KP padded centres: every kc_cent[...] below is an immediate constant operand
#pragma unroll
        for (int j = 0; j < KP; ++j) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int c = 0; c < LD; c += 4) {
                a0 = fmaf(x[c], kc_cent[j * LD + c], a0);
                a1 = fmaf(x[c + 1], kc_cent[j * LD + c + 1], a1);
                a2 = fmaf(x[c + 2], kc_cent[j * LD + c + 2], a2);
                a3 = fmaf(x[c + 3], kc_cent[j * LD + c + 3], a3);
            }
            const float sc = fmaf(-2.0f, (a0 + a1) + (a2 + a3), kc_norm[j]);   // +inf norm for a padding centre
            second = fminf(second, fmaxf(sc, best));
            if (sc < best) bidx = j;       // strict: lowest index wins ties
            best = fminf(best, sc);
        }
        // ---- 3. near-tie: decide in FP64 over all centres (rare)
        const float bound = 2.0f * eps * (2.0f * sqrtf(xn) * cmax + cmax * cmax);
        if (lane < nrows && second - best <= bound) {
            double dbest = INFINITY;
            int di = 0;
            for (int j = 0; j < k; ++j) {
                const double *cj = cent + (size_t)j * LD;
                double dot = 0.0, cn = 0.0;
#pragma unroll
                for (int c = 0; c < LD; ++c) {
                    const double cv = cj[c];
                    dot = fma((double)x[c], cv, dot);
                    cn = fma(cv, cv, cn);
                }
                const double sc = cn - 2.0 * dot;
                if (sc < dbest) { dbest = sc; di = j; }
            }
            bidx = di;
            ++rechecked;
        }
        // ---- 4. label store (coalesced)
        if (lane < nrows) labels[row0 + lane] = bidx;
        // ---- 5. accumulate (lanes own columns; see accumulate_sorted)
        accumulate_sorted<LD, STRIDE>(tile, c32, mypart, k, nrows, bidx, lane, s_sort[w], cost);
        __syncwarp();
    }
    // ---- per-CTA reduction of the 8 warp partials (fixed order), then one record per CTA in global memory
    cost = warp_sum(cost);
    __shared__ double s_cost[kFusedWarps];
    if (lane == 0) s_cost[w] = cost;
    __syncthreads();
    double *out = cta_part + (size_t)blockIdx.x * (psize + 1);
    for (int i = threadIdx.x; i < psize; i += blockDim.x) {
        double s = 0.0;
        for (int ww = 0; ww < kFusedWarps; ++ww) s += parts[(size_t)ww * psize + i];
        out[i] = s;
    }
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int ww = 0; ww < kFusedWarps; ++ww) s += s_cost[ww];
        out[psize] = s;
    }
    rechecked = (unsigned long long)warp_sum((double)rechecked);
    if (lane == 0 && rechecked) atomicAdd(recheck_counter, rechecked);
}

// ---------------------------------------------------------------------------------------------------------------
// k <= 16, ld = 32 (BASELINE config 5): the same one-read step with the distances on the tensor cores.
// 32 rows x 16 centres x 32 columns per warp and batch is a 32 x 16 x 32 product: 16 x 32 FMAs per row on the CUDA cores
// (724 of the ~1300 instructions of a batch).  Here it is 24 warp-level mma.m16n8k16 on the two-term FP16 splits of
// assign_pair_h16.cuh (x sx = h1 + h2, -2c sw = g1 + g2, products h2.g1 + h1.g2 + h1.g1, FP32 accumulation started from
// the scaled centre norms): the same arithmetic and error bound as the tcgen05 kernel.  Why not tcgen05 here: a 32-row
// batch per warp is a fraction of its smallest tile (M = 128 per CTA, operands through TMEM, mbarrier hand-offs that
// cost ~3000 cycles per row tile on this shape -- measured with the pair kernel, 33 ms for this pass alone); the per-warp
// MMA needs no hand-off at all and the tensor work is < 1 % of this HBM-bound kernel either way.
// Fragment layout of mma.m16n8k16 (PTX ISA): g = lane / 4, t = lane % 4; A regs {(g, 2t..2t+1), (g+8, 2t..), (g, 2t+8..),
// (g+8, 2t+8..)}; B regs {(k = 2t..2t+1, n = g), (k = 2t+8.., n = g)}; C {(g, 2t), (g, 2t+1), (g+8, 2t), (g+8, 2t+1)}.
struct FusedMmaPrep {            // written by fused_mma_prepare_kernel, read by every CTA
    uint32_t g1[16 * 16];        // [centre][column pair] FP16 x2: -2 c sw, high part
    uint32_t g2[16 * 16];        //                                   residual
    float cn[16];                // ||c||^2 sx sw (scaled like the products), +inf for padding centres
    float c32[16 * 32];          // FP32 centres (direct-form cost)
    float unscale, sx, abs_x, abs_c, cmax;
};

__global__ void fused_mma_prepare_kernel(const double *__restrict__ cent, int k, const float *__restrict__ xabs,
                                         FusedMmaPrep *__restrict__ out) {
    __shared__ float red[32];
    float m = 0.f;
    for (int i = threadIdx.x; i < k * 32; i += blockDim.x) m = fmaxf(m, __double2float_ru(fabs(cent[i])));
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) m = fmaxf(m, red[w]);
    const int ex = 15 - pow2_above(*xabs), ew = 14 - pow2_above(m);
    const double sw = ldexp(1.0, ew);
    for (int i = threadIdx.x; i < 16 * 16; i += blockDim.x) {
        const int j = i >> 4, cp = i & 15;
        __half2 a = __floats2half2_rn(0.f, 0.f), b = a;
        if (j < k) {
            const double w0 = -2.0 * cent[j * 32 + 2 * cp] * sw, w1 = -2.0 * cent[j * 32 + 2 * cp + 1] * sw;
            const __half h0 = __double2half(w0), h1 = __double2half(w1);
            a = __halves2half2(h0, h1);
            b = __halves2half2(__double2half(w0 - (double)__half2float(h0)), __double2half(w1 - (double)__half2float(h1)));
        }
        out->g1[i] = *reinterpret_cast<uint32_t *>(&a);
        out->g2[i] = *reinterpret_cast<uint32_t *>(&b);
    }
    for (int i = threadIdx.x; i < 16 * 32; i += blockDim.x) out->c32[i] = i < k * 32 ? (float)cent[i] : 0.f;
    float nmax = 0.f;
    if (threadIdx.x < 16) {
        const int j = threadIdx.x;
        double s = 0.0;
        if (j < k)
            for (int c = 0; c < 32; ++c) s += cent[j * 32 + c] * cent[j * 32 + c];
        out->cn[j] = j < k ? (float)(s * ldexp(1.0, ex + ew)) : __int_as_float(0x7f800000);
        nmax = j < k ? (float)sqrt(s) * 1.000001f : 0.f;
    }
    for (int o = 8; o > 0; o >>= 1) nmax = fmaxf(nmax, __shfl_xor_sync(0xffffffffu, nmax, o));
    if (threadIdx.x == 0) {
        out->unscale = ldexpf(1.f, -(ex + ew));
        out->sx = ldexpf(1.f, ex);
        out->abs_x = ldexpf(1.f, -25 - ex);
        out->abs_c = ldexpf(1.f, -25 - ew);
        out->cmax = nmax;
    }
}

__device__ __forceinline__ void mma_f16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// smem (dynamic): per warp a 32 x 36-float X tile + 32 labels, per warp FP64 partials [k*32 + k]: 72 KB, three CTAs per SM.
// cta_part: [cta][k*32 sums | k counts | cost] FP64, reduced by fused_reduce_kernel as for the CUDA-core kernel.
__global__ void __launch_bounds__(kFusedWarps * 32, 3)
fused_mma_kernel(const float *__restrict__ X, const float *__restrict__ xnorm, const double *__restrict__ cent,
                 const FusedMmaPrep *__restrict__ prep, int64_t n, int k, float eps, int32_t *__restrict__ labels,
                 double *__restrict__ cta_part, unsigned long long *__restrict__ recheck_counter) {
    constexpr int LD = 32, STRIDE = LD + 4, WARP_FLOATS = 32 * STRIDE + 32;
    extern __shared__ __align__(16) uint8_t fsm[];
    __shared__ int s_sort[kFusedWarps][72];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    float *tile = reinterpret_cast<float *>(fsm) + (size_t)w * WARP_FLOATS;
    int *lab = reinterpret_cast<int *>(tile + 32 * STRIDE);
    const float *c32 = prep->c32;   // [16][32] FP32 centres, read through L1 (three CTAs per SM leave no shared memory for them)
    double *parts = reinterpret_cast<double *>(reinterpret_cast<float *>(fsm) + (size_t)kFusedWarps * WARP_FLOATS);
    const int psize = k * LD + k;
    double *mypart = parts + (size_t)w * psize;
    for (int i = lane; i < psize; i += 32) mypart[i] = 0.0;
    // ---- per-kernel constants in registers: B fragments and the scaled norms of this lane's columns
    uint32_t bg1[2][2][2], bg2[2][2][2];   // [n tile][k step][reg]
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int ks = 0; ks < 2; ++ks)
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const int idx = (8 * nt + g) * 16 + 8 * ks + 4 * r + t;   // centre 8nt + g, column pair (16ks + 8r + 2t) / 2
                bg1[nt][ks][r] = prep->g1[idx];
                bg2[nt][ks][r] = prep->g2[idx];
            }
    float cnv[2][2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) { cnv[nt][0] = prep->cn[8 * nt + 2 * t]; cnv[nt][1] = prep->cn[8 * nt + 2 * t + 1]; }
    const float sx = prep->sx, unscale = prep->unscale, cmax = prep->cmax;
    const float abs_term = 5.656854f /* sqrt(32) */ * prep->abs_x * 2.0f * cmax, abs_c = 5.656854f * prep->abs_c;
    double cost = 0.0;
    unsigned long long rechecked = 0;
    __syncthreads();

    const int64_t nbatches = (n + 31) / 32;
    const int64_t gw = (int64_t)blockIdx.x * kFusedWarps + w, gstride = (int64_t)gridDim.x * kFusedWarps;
    for (int64_t b = gw; b < nbatches; b += gstride) {
        const int64_t row0 = b * 32;
        const int nrows = (int)min((int64_t)32, n - row0);
        // ---- 1. coalesced load of 32 rows into the padded tile (8 float4 per row)
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), q = lane & 7;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < nrows) v = __ldg(reinterpret_cast<const float4 *>(X + (row0 + r) * LD) + q);
            *reinterpret_cast<float4 *>(tile + r * STRIDE + q * 4) = v;
        }
        float xn[4];   // squared norms of this lane's four rows (g, g + 8, g + 16, g + 24): the near-tie bound
#pragma unroll
        for (int i = 0; i < 4; ++i) xn[i] = (g + 8 * i) < nrows ? __ldg(xnorm + row0 + g + 8 * i) : 0.f;
        __syncwarp();
        // ---- 2. scores: C starts at the scaled norms; three FP16 products per (row tile, centre tile, k step)
        float acc[2][2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                acc[mt][nt][0] = acc[mt][nt][2] = cnv[nt][0];
                acc[mt][nt][1] = acc[mt][nt][3] = cnv[nt][1];
            }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                uint32_t h1[4], h2[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) {   // reg r: row 16mt + g + 8 (r & 1), columns 16ks + 8 (r >> 1) + 2t, +1
                    const float2 v = *reinterpret_cast<const float2 *>(tile + (16 * mt + g + 8 * (r & 1)) * STRIDE + 16 * ks + 8 * (r >> 1) + 2 * t);
                    const float v0 = v.x * sx, v1 = v.y * sx;
                    const __half2 a = __floats2half2_rn(v0, v1);
                    const float2 af = __half22float2(a);
                    const __half2 bb = __floats2half2_rn(v0 - af.x, v1 - af.y);
                    h1[r] = *reinterpret_cast<const uint32_t *>(&a);
                    h2[r] = *reinterpret_cast<const uint32_t *>(&bb);
                }
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {   // small terms first
                    mma_f16_16816(acc[mt][nt], h2, bg1[nt][ks][0], bg1[nt][ks][1]);
                    mma_f16_16816(acc[mt][nt], h1, bg2[nt][ks][0], bg2[nt][ks][1]);
                    mma_f16_16816(acc[mt][nt], h1, bg1[nt][ks][0], bg1[nt][ks][1]);
                }
            }
        // ---- 3. per row: best / runner-up over this lane's 4 scores, then over the 4 lanes of the quad
        unsigned flagged = 0;   // bit i: row g + 8 i is inside the near-tie bound (the same in all lanes of a quad)
        int rowlab[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // row g + 8 i = (row tile i >> 1, upper / lower half i & 1)
            const int mt = i >> 1, hf = (i & 1) * 2;
            float best = acc[mt][0][hf], second = __int_as_float(0x7f800000);
            int bi = 2 * t;
            auto put = [&](float sc, int j) {
                second = fminf(second, fmaxf(sc, best));
                if (sc < best) bi = j;          // strict: the columns of a lane come in ascending order
                best = fminf(best, sc);
            };
            put(acc[mt][0][hf + 1], 2 * t + 1);
            put(acc[mt][1][hf], 8 + 2 * t);
            put(acc[mt][1][hf + 1], 8 + 2 * t + 1);
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o), os = __shfl_xor_sync(0xffffffffu, second, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                const bool win = ob < best || (ob == best && oi < bi);   // lowest index among equal minima
                second = fminf(win ? best : ob, fminf(second, os));
                bi = win ? oi : bi;
                best = fminf(best, ob);
            }
            rowlab[i] = bi;
            const float xr = sqrtf(xn[i]);
            const float bound = 2.0f * (eps * (2.0f * xr * cmax + cmax * cmax) + abs_term + abs_c * xr);
            if ((second - best) * unscale <= bound) flagged |= 1u << i;
        }
        if (t == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) lab[g + 8 * i] = rowlab[i];
        }
        __syncwarp();
        // ---- 4. near-ties: FP64 over all centres, the warp shares a row (lanes = columns); rare
        unsigned anyflag = __ballot_sync(0xffffffffu, t == 0 && flagged != 0);
        while (anyflag) {
            const int src = __ffs(anyflag) - 1;          // a quad leader with flagged rows
            anyflag &= anyflag - 1;
            const unsigned fl = __shfl_sync(0xffffffffu, flagged, src);
            for (int i = 0; i < 4; ++i) {
                const int r = (src >> 2) + 8 * i;
                if (!(fl & (1u << i)) || r >= nrows) continue;
                const double xv = (double)tile[r * STRIDE + lane];
                double dbest = INFINITY;
                int di = 0;
                for (int j = 0; j < k; ++j) {
                    const double cv = cent[(size_t)j * LD + lane];
                    const double dot = warp_sum(xv * cv), cn = warp_sum(cv * cv);
                    const double sc = cn - 2.0 * dot;
                    if (sc < dbest) { dbest = sc; di = j; }
                }
                if (lane == 0) { lab[r] = di; ++rechecked; }
            }
        }
        __syncwarp();
        const int bidx = lane < nrows ? lab[lane] : 0;
        if (lane < nrows) labels[row0 + lane] = bidx;
        // ---- 5. accumulate (lanes own columns; see accumulate_sorted)
        accumulate_sorted<LD, STRIDE>(tile, c32, mypart, k, nrows, bidx, lane, s_sort[w], cost);
        __syncwarp();
    }
    // ---- per-CTA reduction of the 8 warp partials (fixed order), then one record per CTA in global memory
    cost = warp_sum(cost);
    __shared__ double s_cost[kFusedWarps];
    if (lane == 0) s_cost[w] = cost;
    __syncthreads();
    double *out = cta_part + (size_t)blockIdx.x * (psize + 1);
    for (int i = threadIdx.x; i < psize; i += blockDim.x) {
        double s = 0.0;
        for (int ww = 0; ww < kFusedWarps; ++ww) s += parts[(size_t)ww * psize + i];
        out[i] = s;
    }
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int ww = 0; ww < kFusedWarps; ++ww) s += s_cost[ww];
        out[psize] = s;
    }
    if (lane == 0 && rechecked) atomicAdd(recheck_counter, rechecked);
}
__host__ __device__ inline size_t fused_mma_smem_bytes(int k) {
    return (size_t)kFusedWarps * (32 * 36 + 32) * 4 + (size_t)kFusedWarps * ((size_t)k * 32 + k) * 8;
}

// red[i] = sum over CTAs (in CTA order) of cta_part[cta][i], i < k*ld + k + 1
__global__ void fused_reduce_kernel(const double *__restrict__ cta_part, int ncta, int psize1, double *__restrict__ red) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= psize1) return;
    double s = 0.0;
    for (int c = 0; c < ncta; ++c) s += cta_part[(size_t)c * psize1 + i];
    red[i] = s;
}

}  // namespace kmb
