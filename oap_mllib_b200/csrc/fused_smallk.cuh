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
        // ---- 5. accumulate: lanes own columns.  The 32 rows of the batch are first ordered by label (a counting sort
        //         inside the warp: __match_any_sync gives a row's rank among the rows of its label, a 32-entry shared-
        //         memory histogram and one warp scan give the labels' offsets), then walked in that order by a fully
        //         unrolled loop: the running FP64 sum of a column stays in a register while the label does not change
        //         and goes to the shared-memory partial once per (label, batch).  No data-dependent loop control: the
        //         only branch is the warp-uniform "label changed" test.  (Round 1 looped over the distinct labels with
        //         ballot / ffs loops: about 1000 of the 1860 instructions of a batch.)
        {
            int *sc = s_sort[w], *sp = s_sort[w] + 40;
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

// red[i] = sum over CTAs (in CTA order) of cta_part[cta][i], i < k*ld + k + 1
__global__ void fused_reduce_kernel(const double *__restrict__ cta_part, int ncta, int psize1, double *__restrict__ red) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= psize1) return;
    double s = 0.0;
    for (int c = 0; c < ncta; ++c) s += cta_part[(size_t)c * psize1 + i];
    red[i] = s;
}

}  // namespace kmb
