// CUDA-core kernels around the tensor-core assign step: ingest, centroid preparation, FP64 re-check
// of near-tie rows, label bucketing, per-cluster FP64 accumulation, finalize / convergence, the
// empty-cluster candidate search and the synthetic-blob generator.
#pragma once
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace kmb {

constexpr int kSegChunk = 512;     // rows per accumulate work unit (one CTA)
constexpr int kBucketRows = 8192;  // rows per CTA in the bucketing kernels
constexpr int kHistSmemMax = 12288;  // largest k whose histogram is kept in shared memory

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------ ingest
// staging (dense nrows x d, T) -> table rows (leading dimension ld, FP32); padding stays zero.
template <typename T>
__global__ void convert_rows_kernel(const T *__restrict__ src, float *__restrict__ dst, int64_t nrows, int d,
                                    int ld) {
    const int64_t total = nrows * d;
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / d;
        const int c = (int)(e - r * d);
        dst[r * ld + c] = (float)src[e];
    }
}

// ---- columnar and sparse ingest
// src: ncols columns of h values each, column c at src + c * col_stride (Arrow-style contiguous columns staged in
// device memory); dst: rows [0, h) of the row-major FP32 table.  32 x 32 tiles through shared memory so that both
// the column reads and the row writes are coalesced.
template <typename T>
__global__ void transpose_columns_kernel(const T *__restrict__ src, int64_t col_stride, float *__restrict__ dst,
                                         int64_t h, int d, int ld) {
    __shared__ float tile[32][33];
    const int64_t tiles_r = (h + 31) / 32;
    const int tiles_c = (d + 31) / 32;
    for (int64_t tix = blockIdx.x; tix < tiles_r * tiles_c; tix += gridDim.x) {
        const int64_t r0 = (tix / tiles_c) * 32;
        const int c0 = (int)(tix % tiles_c) * 32;
        // threadIdx.x walks ROWS here (contiguous inside a column), threadIdx.y the columns
        for (int cy = threadIdx.y; cy < 32; cy += blockDim.y) {
            const int64_t r = r0 + threadIdx.x;
            const int c = c0 + cy;
            tile[cy][threadIdx.x] = (r < h && c < d) ? (float)src[(size_t)c * col_stride + r] : 0.f;
        }
        __syncthreads();
        // threadIdx.x walks COLUMNS here (contiguous inside a row)
        for (int ry = threadIdx.y; ry < 32; ry += blockDim.y) {
            const int64_t r = r0 + ry;
            const int c = c0 + threadIdx.x;
            if (r < h && c < d) dst[r * ld + c] = tile[threadIdx.x][ry];
        }
        __syncthreads();
    }
}

// CSR rows -> dense FP32 rows (the reference densifies sparse vectors with Vector.toArray on the JVM,
// scala/.../ml/util/OneDAL.scala:135-139).  dst rows must have been zeroed.  One warp per row; indptr is relative
// to `base` (the first nonzero of the staged chunk).  An index outside [0, d) sets *bad and is skipped.
__global__ void csr_scatter_kernel(const int64_t *__restrict__ indptr, int64_t base, const int32_t *__restrict__ indices,
                                   const double *__restrict__ values, float *__restrict__ dst, int64_t h, int d, int ld,
                                   int32_t *bad) {
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp_global; r < h; r += nwarps) {
        const int64_t b = indptr[r] - base, e = indptr[r + 1] - base;
        for (int64_t q = b + lane; q < e; q += 32) {
            const int32_t c = indices[q];
            if (c < 0 || c >= d) { *bad = 1; continue; }
            dst[r * ld + c] = (float)values[q];
        }
    }
}

// ||x_i||^2 for rows [row0, row0 + nrows): G = min(32, ld/4) lanes per row, float4 loads.  Also folds max |x_ij|
// into *xabs (atomicMax on the bits of a non-negative float).
__global__ void rownorm_kernel(const float *__restrict__ X, float *__restrict__ xnorm, float *xabs, int64_t row0,
                               int64_t nrows, int ld) {
    const int vec = ld >> 2;
    const int G = vec < 32 ? vec : 32;  // ld is a multiple of 32 -> vec is a multiple of 8
    const int rows_per_warp = 32 / G;
    const int lane = threadIdx.x & 31;
    const int sub = lane / G, gl = lane % G;
    const int64_t warp_global = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    float amax = 0.f;
    for (int64_t rb = warp_global * rows_per_warp; rb < nrows; rb += nwarps * rows_per_warp) {
        const int64_t r = rb + sub;
        float acc = 0.f;
        if (r < nrows) {
            const float4 *p = reinterpret_cast<const float4 *>(X + (row0 + r) * ld);
            for (int v = gl; v < vec; v += G) {
                const float4 q = __ldg(p + v);
                acc += q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
                amax = fmaxf(amax, fmaxf(fmaxf(fabsf(q.x), fabsf(q.y)), fmaxf(fabsf(q.z), fabsf(q.w))));
            }
        }
        for (int o = G >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (r < nrows && gl == 0) xnorm[row0 + r] = acc;
    }
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0 && amax > 0.f) atomicMax(reinterpret_cast<int *>(xabs), __float_as_int(amax));
}

// ------------------------------------------------------------------ centroid preparation
// One CTA per padded centroid row j.  For j < k: B_hi/B_lo = TF32 round-to-nearest split of -2*c_j,
// cnorm = ||c_j||^2 (FP64 sum, rounded to FP32), cstats[0] = max ||c_j|| (atomicMax on the bits of
// a non-negative float).  For j >= k: zero rows and cnorm = +inf so the column never wins.
// bcorr (BF16, [kpad x 2*ld]) is the B operand of the BF16 correction MMAs: for every 32-column chunk
// 64 values per centre, [bf16(hi) x 32 | bf16(lo) x 32], so that one 128-byte row pairs with the A
// chunk [bf16(x_lo) x 32 | bf16(x) x 32] and sums x_lo.w_hi + x.w_lo.
__global__ void split_centers_kernel(const double *__restrict__ cent, float *__restrict__ bhi,
                                     float *__restrict__ blo, __nv_bfloat16 *__restrict__ bcorr,
                                     float *__restrict__ cnorm, double *__restrict__ cnorm64, float *cstats, int k,
                                     int ld) {
    const int j = blockIdx.x;
    __shared__ double red[32];
    double acc = 0.0;
    for (int c = threadIdx.x; c < ld; c += blockDim.x) {
        float h = 0.f, l = 0.f;
        if (j < k) {
            const double v = cent[(size_t)j * ld + c];
            acc += v * v;
            const float f = (float)(-2.0 * v);
            uint32_t hb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(f));
            h = __uint_as_float(hb);
            const float rem = (float)(-2.0 * v - (double)h);
            uint32_t lb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(rem));
            l = __uint_as_float(lb);
        }
        bhi[(size_t)j * ld + c] = h;
        blo[(size_t)j * ld + c] = l;
        const size_t cb = (size_t)j * 2 * ld + (size_t)(c >> 5) * 64 + (c & 31);
        bcorr[cb] = __float2bfloat16_rn(h);
        bcorr[cb + 32] = __float2bfloat16_rn(l);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < (blockDim.x + 31) / 32; ++w) s += red[w];
        cnorm64[j] = j < k ? s : INFINITY;   // FP64 norms for the full re-check
        if (j < k) {
            cnorm[j] = (float)s;
            // round the norm up a little so the bound stays an upper bound
            const float nrm = (float)sqrt(s) * 1.000001f;
            atomicMax(reinterpret_cast<int *>(cstats), __float_as_int(nrm));
        } else {
            cnorm[j] = __int_as_float(0x7f800000);
        }
    }
}

// ---- FP16-split operand (assign_pair_h16.cuh)
// cstats[1] = max |c_ij| (launch before split16_kernel; cstats zeroed by the caller)
__global__ void cent_absmax_kernel(const double *__restrict__ cent, int64_t count, float *cstats) {
    float m = 0.f;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        m = fmaxf(m, __double2float_ru(fabs(cent[i])));
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<int *>(cstats + 1), __float_as_int(m));
}

// smallest E (clamped to +-60) with m < 2^E
__device__ __forceinline__ int pow2_above(float m) {
    if (!(m > 0.f)) return 0;
    int e;
    frexpf(m, &e);   // m = f * 2^e, f in [0.5, 1)
    return max(-60, min(60, e));
}

// One CTA per padded centroid row j: b16[j] = [g1 (ld64 values) | g2 (ld64 values)] with g1 + g2 ~= -2 c_j sw in FP16
// (sw = 2^(14 - Ec), |c| < 2^Ec, so |g| < 2^15), zero beyond ld / k.  Also cnorm, cnorm64, cstats[0] exactly as
// split_centers_kernel, and (block 0) the scales the assign kernel needs: cstats[2] = 1 / (sx sw), cstats[3] = sx
// (= 2^(15 - Ex), |x| < 2^Ex), cstats[4] / cstats[5] = the absolute error of one x / -2c element that falls into the
// FP16 subnormal range (2^-25 in scaled units).
__global__ void split16_kernel(const double *__restrict__ cent, __half *__restrict__ b16, float *__restrict__ cnorm,
                               double *__restrict__ cnorm64, float *cstats, const float *__restrict__ xabs, int k,
                               int ld, int ld64) {
    const int j = blockIdx.x;
    __shared__ double red[32];
    const int Ex = pow2_above(*xabs), Ec = pow2_above(cstats[1]);
    const int ex = 15 - Ex, ew = 14 - Ec;
    const double sw = ldexp(1.0, ew);
    if (j == 0 && threadIdx.x == 0) {
        cstats[2] = ldexpf(1.f, -(ex + ew));
        cstats[3] = ldexpf(1.f, ex);
        cstats[4] = ldexpf(1.f, -25 - ex);
        cstats[5] = ldexpf(1.f, -25 - ew);
    }
    double acc = 0.0;
    for (int c = threadIdx.x; c < ld64; c += blockDim.x) {
        __half g1 = __float2half_rn(0.f), g2 = g1;
        if (j < k && c < ld) {
            const double v = cent[(size_t)j * ld + c];
            acc += v * v;
            const double ws = -2.0 * v * sw;
            g1 = __double2half(ws);
            g2 = __double2half(ws - (double)__half2float(g1));
        }
        b16[(size_t)j * 2 * ld64 + c] = g1;
        b16[(size_t)j * 2 * ld64 + ld64 + c] = g2;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < (blockDim.x + 31) / 32; ++w) s += red[w];
        cnorm64[j] = j < k ? s : INFINITY;
        if (j < k) {
            cnorm[j] = (float)s;
            const float nrm = (float)sqrt(s) * 1.000001f;
            atomicMax(reinterpret_cast<int *>(cstats), __float_as_int(nrm));
        } else {
            cnorm[j] = __int_as_float(0x7f800000);
        }
    }
}

// ------------------------------------------------------------------ exact (FP64) assignment
// score_j = ||c_j||^2 - 2 x.c_j in FP64 for one row held by a warp (lanes split the columns).
__device__ __forceinline__ double exact_score(const float *__restrict__ xrow, const double *__restrict__ cj, int ld,
                                              int lane) {
    double dot = 0.0, cn = 0.0;
    for (int c0 = lane; c0 < ld; c0 += 32) {
        const double c = __ldg(cj + c0);
        dot = fma((double)__ldg(xrow + c0), c, dot);
        cn = fma(c, c, cn);
    }
    dot = warp_sum(dot);
    cn = warp_sum(cn);
    return cn - 2.0 * dot;
}

// One warp per listed entry.  The tensor-core epilogue lists every row whose best / second-best gap
// is inside the score error bound as (row, runner-up centre) -- then the winner is decided between
// labels[row] and the runner-up in FP64 -- or as (row, -1) when a third centre is inside the bound
// too, in which case all k centres are re-scored.  list == nullptr: every row, all centres (the
// CUDA-core reference path behind option "force_simt").  First minimum (lowest index) wins ties.
__global__ void recheck_kernel(const float *__restrict__ X, const double *__restrict__ cent,
                               const int2 *__restrict__ list, const int32_t *__restrict__ list_count,
                               int32_t list_cap, int64_t n, int k, int ld, int32_t *__restrict__ labels) {
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int64_t count = n;
    if (list) {
        count = *list_count;
        if (count > list_cap) count = list_cap;
    }
    const int per_lane = ld >> 5;        // ld is a multiple of 32
    const bool cached = per_lane <= 16;  // rows up to 512 floats are held in registers for the full scan
    for (int64_t w = warp_global; w < count; w += nwarps) {
        int64_t row = w;
        int runner = -1;
        if (list) {
            const int2 e = list[w];
            row = e.x;
            runner = e.y;
        }
        const float *xrow = X + row * ld;
        if (runner != -1) {   // (-1 cannot be a packed pair: winner != runner)
            // entry.y = (winner << 16) | runner-up, both chosen by the tensor-core epilogue
            const int cur = (runner >> 16) & 0xffff;
            runner &= 0xffff;
            const double *cc = cent + (size_t)cur * ld, *cr = cent + (size_t)runner * ld;
            double s_cur, s_run;
            if (per_lane <= 8) {
                // all loads of the row and of both centres are issued before any arithmetic
                float xv[8];
                double cv[8], rv[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const bool on = i < per_lane;
                    xv[i] = on ? __ldg(xrow + i * 32 + lane) : 0.f;
                    cv[i] = on ? __ldg(cc + i * 32 + lane) : 0.0;
                    rv[i] = on ? __ldg(cr + i * 32 + lane) : 0.0;
                }
                double d0 = 0.0, n0 = 0.0, d1 = 0.0, n1 = 0.0;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const double xd = (double)xv[i];
                    d0 = fma(xd, cv[i], d0);
                    n0 = fma(cv[i], cv[i], n0);
                    d1 = fma(xd, rv[i], d1);
                    n1 = fma(rv[i], rv[i], n1);
                }
                d0 = warp_sum(d0); n0 = warp_sum(n0); d1 = warp_sum(d1); n1 = warp_sum(n1);
                s_cur = n0 - 2.0 * d0;
                s_run = n1 - 2.0 * d1;
            } else {
                s_cur = exact_score(xrow, cc, ld, lane);
                s_run = exact_score(xrow, cr, ld, lane);
            }
            const bool take = s_run < s_cur || (s_run == s_cur && runner < cur);
            if (lane == 0 && take) labels[row] = runner;
            continue;
        }
        double xr[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) xr[i] = (cached && i < per_lane) ? (double)__ldg(xrow + i * 32 + lane) : 0.0;
        double best = INFINITY;
        int bidx = 0;
        for (int j = 0; j < k; ++j) {
            const double *cj = cent + (size_t)j * ld;
            double sc;
            if (cached) {
                double dot = 0.0, cn = 0.0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    if (i < per_lane) {
                        const double c = __ldg(cj + i * 32 + lane);
                        dot = fma(xr[i], c, dot);
                        cn = fma(c, c, cn);
                    }
                }
                dot = warp_sum(dot);
                cn = warp_sum(cn);
                sc = cn - 2.0 * dot;
            } else {
                sc = exact_score(xrow, cj, ld, lane);
            }
            if (sc < best) {
                best = sc;
                bidx = j;
            }
        }
        if (lane == 0) labels[row] = bidx;
    }
}

// Full re-score of the rows listed at the BACK of flag_list (entries cap-1, cap-2, ...): a CTA of 8
// warps takes 8 rows, one per warp, and all warps walk the k centres in the same order, so a centre
// row fetched from L2 by one warp is an L1 hit for the other seven.  FP64 scores ||c||^2 - 2 x.c with
// the norms precomputed (cn64, padded to a multiple of 4 with +inf); four centres per step: the four
// per-lane partial dots are reduced together by a folding butterfly (6 shuffles instead of 20).
// First minimum wins.
// When only a few rows are listed (large k*d, e.g. k = 10000, d = 512) the k centres of a row group are split
// over several CTAs (nsplit, derived from the list length on the device); the partial winners go to a small
// scratch and recheck_merge_kernel picks the final one.  With many rows nsplit = 1 and labels are written directly.
constexpr int kRecheckMaxSplit = 64;
__device__ __forceinline__ int recheck_nsplit(int count, int grid) {
    const int ngroups = (count + 7) / 8;
    if (ngroups <= 0) return 1;
    return max(1, min(kRecheckMaxSplit, grid / ngroups));
}

__global__ void __launch_bounds__(256)
recheck_full_kernel(const float *__restrict__ X, const double *__restrict__ cent, const double *__restrict__ cn64,
                    const int2 *__restrict__ list, const int32_t *__restrict__ full_count, int32_t list_cap, int k,
                    int ld, int32_t *__restrict__ labels, double *__restrict__ part_score,
                    int32_t *__restrict__ part_idx) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int count = *full_count;
    if (count > list_cap) count = list_cap;
    const int per_lane = ld >> 5;
    const bool b16 = lane & 16, b8 = lane & 8;
    const int mine = (b16 ? 2 : 0) + (b8 ? 1 : 0);   // which of the four centres this lane ends up holding
    const int ngroups = (count + 7) / 8;
    const int nsplit = recheck_nsplit(count, gridDim.x);
    const int kk = (((k + nsplit - 1) / nsplit) + 3) & ~3;   // centres per split, multiple of 4
    for (int u = blockIdx.x; u < ngroups * nsplit; u += gridDim.x) {
        const int e = (u / nsplit) * 8 + w, c = u % nsplit;
        if (e >= count) continue;
        const int row = list[list_cap - 1 - e].x;
        const float *xrow = X + (size_t)row * ld;
        const int j0 = c * kk, j1 = min(k, j0 + kk);
        double best = INFINITY;
        int bidx = 0x7fffffff;
        if (per_lane <= 16) {
            double xr[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) xr[i] = i < per_lane ? (double)__ldg(xrow + i * 32 + lane) : 0.0;
            for (int j = j0; j < j1; j += 4) {
                double v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int jj = min(j + q, k - 1);   // clamp the address; the padded norm (+inf) removes it
                    const double *cj = cent + (size_t)jj * ld;
                    double dot = 0.0;
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (i < per_lane) dot = fma(xr[i], cj[i * 32 + lane], dot);   // plain loads: keep them in L1
                    v[q] = dot;
                }
                // fold: lanes with bit 4 clear keep centres 0,1, the others 2,3; then bit 3 picks one of the two
                double k0 = (b16 ? v[2] : v[0]) + __shfl_xor_sync(0xffffffffu, b16 ? v[0] : v[2], 16);
                double k1 = (b16 ? v[3] : v[1]) + __shfl_xor_sync(0xffffffffu, b16 ? v[1] : v[3], 16);
                double t = (b8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, b8 ? k0 : k1, 8);
                t += __shfl_xor_sync(0xffffffffu, t, 4);
                t += __shfl_xor_sync(0xffffffffu, t, 2);
                t += __shfl_xor_sync(0xffffffffu, t, 1);
                double sc = cn64[j + mine] - 2.0 * t;   // cn64 is padded with +inf beyond k
                int si = j + mine;
                // minimum of the four (lower index wins ties): partners at lane distance 8 and 16
                double o = __shfl_xor_sync(0xffffffffu, sc, 8);
                int oi = si ^ 1;
                if (o < sc || (o == sc && oi < si)) { sc = o; si = oi; }
                o = __shfl_xor_sync(0xffffffffu, sc, 16);
                oi = __shfl_xor_sync(0xffffffffu, si, 16);
                if (o < sc || (o == sc && oi < si)) { sc = o; si = oi; }
                if (sc < best) { best = sc; bidx = si; }
            }
        } else {
            for (int j = j0; j < j1; ++j) {
                const double sc = exact_score(xrow, cent + (size_t)j * ld, ld, lane);
                if (sc < best) { best = sc; bidx = j; }
            }
        }
        if (lane == 0) {
            if (nsplit == 1) {
                labels[row] = bidx;
            } else {
                part_score[(size_t)e * kRecheckMaxSplit + c] = best;
                part_idx[(size_t)e * kRecheckMaxSplit + c] = bidx;
            }
        }
    }
}

// grid0 = the grid size recheck_full_kernel was launched with (same nsplit derivation)
__global__ void recheck_merge_kernel(const int2 *__restrict__ list, const int32_t *__restrict__ full_count,
                                     int32_t list_cap, int grid0, const double *__restrict__ part_score,
                                     const int32_t *__restrict__ part_idx, int32_t *__restrict__ labels) {
    int count = *full_count;
    if (count > list_cap) count = list_cap;
    const int nsplit = recheck_nsplit(count, grid0);
    if (nsplit == 1) return;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < count; e += gridDim.x * blockDim.x) {
        double b = INFINITY;
        int bi = 0x7fffffff;
        for (int c = 0; c < nsplit; ++c) {   // ascending centre ranges: strict '<' keeps the lowest index
            const double sc = part_score[(size_t)e * kRecheckMaxSplit + c];
            if (sc < b) { b = sc; bi = part_idx[(size_t)e * kRecheckMaxSplit + c]; }
        }
        labels[list[list_cap - 1 - e].x] = bi;
    }
}

// ---- two-level bookkeeping of the FP16-split assign kernel
// Dense copy of the candidate rows (one warp per row, float4 moves) and of their squared norms.
__global__ void gather_rows_kernel(const float *__restrict__ X, const float *__restrict__ xnorm,
                                   const int32_t *__restrict__ cand, const int32_t *__restrict__ cand_count,
                                   int32_t cand_cap, int ld, float *__restrict__ xc, float *__restrict__ xnc) {
    int count = *cand_count;
    if (count > cand_cap) count = cand_cap;
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const int vec = ld >> 2;
    for (int i = warp_global; i < count; i += nwarps) {
        const int row = cand[i];
        const float4 *src = reinterpret_cast<const float4 *>(X + (size_t)row * ld);
        float4 *dst = reinterpret_cast<float4 *>(xc + (size_t)i * ld);
        for (int v = lane; v < vec; v += 32) dst[v] = __ldg(src + v);
        if (lane == 0) xnc[i] = xnorm[row];
    }
}

__global__ void scatter_labels_kernel(const int32_t *__restrict__ cand, const int32_t *__restrict__ cand_count,
                                      int32_t cand_cap, const int32_t *__restrict__ labels_c,
                                      int32_t *__restrict__ labels) {
    int count = *cand_count;
    if (count > cand_cap) count = cand_cap;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) labels[cand[i]] = labels_c[i];
}

// ------------------------------------------------------------------ bucketing rows by label
// counts[j] = number of local rows with label j.
__global__ void label_hist_kernel(const int32_t *__restrict__ labels, int64_t n, int k, int32_t *counts) {
    extern __shared__ int32_t sh[];
    const bool use_sh = k <= kHistSmemMax;
    if (use_sh) {
        for (int j = threadIdx.x; j < k; j += blockDim.x) sh[j] = 0;
        __syncthreads();
    }
    const int64_t r0 = (int64_t)blockIdx.x * kBucketRows;
    const int64_t r1 = min(n, r0 + kBucketRows);
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
        const int l = labels[r];
        if (use_sh) atomicAdd(&sh[l], 1);
        else atomicAdd(&counts[l], 1);
    }
    if (use_sh) {
        __syncthreads();
        for (int j = threadIdx.x; j < k; j += blockDim.x)
            if (sh[j]) atomicAdd(&counts[j], sh[j]);
    }
}

// Single CTA: offsets = exclusive scan of counts (k+1 entries), cursor = offsets, chunk_off =
// exclusive scan of ceil(counts / kSegChunk) (k+1 entries).
__global__ void scan_kernel(const int32_t *__restrict__ counts, int k, int32_t *offsets, int32_t *cursor,
                            int32_t *chunk_off) {
    __shared__ int32_t s_cnt[1024], s_chk[1024];
    __shared__ int32_t base_cnt, base_chk;
    if (threadIdx.x == 0) { base_cnt = 0; base_chk = 0; }
    __syncthreads();
    for (int j0 = 0; j0 < k; j0 += 1024) {
        const int j = j0 + threadIdx.x;
        const int c = j < k ? counts[j] : 0;
        const int h = (c + kSegChunk - 1) / kSegChunk;
        s_cnt[threadIdx.x] = c;
        s_chk[threadIdx.x] = h;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {  // Hillis-Steele inclusive scan
            int a = 0, b = 0;
            if ((int)threadIdx.x >= o) { a = s_cnt[threadIdx.x - o]; b = s_chk[threadIdx.x - o]; }
            __syncthreads();
            s_cnt[threadIdx.x] += a;
            s_chk[threadIdx.x] += b;
            __syncthreads();
        }
        if (j < k) {
            const int ex = base_cnt + s_cnt[threadIdx.x] - c;
            offsets[j] = ex;
            cursor[j] = ex;
            chunk_off[j] = base_chk + s_chk[threadIdx.x] - h;
        }
        __syncthreads();
        if (threadIdx.x == 1023) { base_cnt += s_cnt[1023]; base_chk += s_chk[1023]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { offsets[k] = base_cnt; chunk_off[k] = base_chk; }
}

// order[] = row indices grouped by label (order inside a group follows atomic arrival): k > kHistSmemMax only.
__global__ void scatter_kernel(const int32_t *__restrict__ labels, int64_t n, int k, int32_t *cursor,
                               int32_t *__restrict__ order) {
    extern __shared__ int32_t sh[];  // [k] counts -> bases (only when k <= kHistSmemMax)
    const bool use_sh = k <= kHistSmemMax;
    const int64_t r0 = (int64_t)blockIdx.x * kBucketRows;
    const int64_t r1 = min(n, r0 + kBucketRows);
    if (!use_sh) {
        for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
            const int pos = atomicAdd(&cursor[labels[r]], 1);
            order[pos] = (int32_t)r;
        }
        return;
    }
    for (int j = threadIdx.x; j < k; j += blockDim.x) sh[j] = 0;
    __syncthreads();
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) atomicAdd(&sh[labels[r]], 1);
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const int c = sh[j];
        sh[j] = c ? atomicAdd(&cursor[j], c) : 0;  // reserve a contiguous range per (block, label)
    }
    __syncthreads();
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
        const int pos = atomicAdd(&sh[labels[r]], 1);
        order[pos] = (int32_t)r;
    }
}

// ---- stable bucketing (k <= kHistSmemMax): the rows of a cluster appear in order[] in ROW ORDER, so every FP64 sum
// downstream adds the same values in the same order on every run -- a fit is bit-reproducible.
// (1) block_hist_kernel: hb[b][j] = rows of block b (kBucketRows consecutive rows) with label j;
// (2) the column scan over blocks in three small kernels (segments of kScanSeg blocks): per-segment column sums,
//     a scan over the segments, then the exclusive prefix inside every segment written back in place; counts[j]
//     receives the column total;
// (3) scan_kernel (above) turns counts into offsets / chunk offsets;
// (4) scatter_stable_kernel: block b starts label j at offsets[j] + hb[b][j] and ranks its rows in row order.
constexpr int kScanSeg = 128;   // blocks per segment of the column scan

__global__ void block_hist_kernel(const int32_t *__restrict__ labels, int64_t n, int k, int32_t *__restrict__ hb) {
    extern __shared__ int32_t sh[];
    for (int j = threadIdx.x; j < k; j += blockDim.x) sh[j] = 0;
    __syncthreads();
    const int64_t r0 = (int64_t)blockIdx.x * kBucketRows;
    const int64_t r1 = min(n, r0 + kBucketRows);
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) atomicAdd(&sh[labels[r]], 1);
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += blockDim.x) hb[(size_t)blockIdx.x * k + j] = sh[j];
}

// grid (ceil(k / 256), nseg): segsum[s][j] = sum of hb[b][j] over the blocks of segment s
__global__ void colsum_seg_kernel(const int32_t *__restrict__ hb, int nblocks, int k, int32_t *__restrict__ segsum) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    const int b0 = blockIdx.y * kScanSeg, b1 = min(nblocks, b0 + kScanSeg);
    int s = 0;
    for (int b = b0; b < b1; ++b) s += hb[(size_t)b * k + j];
    segsum[(size_t)blockIdx.y * k + j] = s;
}
// one thread per label: exclusive scan of segsum over the segments (in place); counts[j] = total
__global__ void seg_scan_kernel(int32_t *__restrict__ segsum, int nseg, int k, int32_t *__restrict__ counts) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    int run = 0;
    for (int s = 0; s < nseg; ++s) {
        const int v = segsum[(size_t)s * k + j];
        segsum[(size_t)s * k + j] = run;
        run += v;
    }
    counts[j] = run;
}
// grid (ceil(k / 256), nseg): hb[b][j] <- rows with label j in the blocks before b (exclusive, in place)
__global__ void colscan_seg_kernel(int32_t *__restrict__ hb, int nblocks, int k, const int32_t *__restrict__ segsum) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    const int b0 = blockIdx.y * kScanSeg, b1 = min(nblocks, b0 + kScanSeg);
    int run = segsum[(size_t)blockIdx.y * k + j];
    for (int b = b0; b < b1; ++b) {
        const int v = hb[(size_t)b * k + j];
        hb[(size_t)b * k + j] = run;
        run += v;
    }
}

// 256 threads, one block of kBucketRows rows.  The rows are taken 256 at a time; inside such a slice the rank of a
// row among the rows of its label is: (lanes of the same warp with the same label and a lower lane id) + what the
// lower warps of the slice added, which the warps publish one after the other through the shared cursor.
__global__ void __launch_bounds__(256)
scatter_stable_kernel(const int32_t *__restrict__ labels, int64_t n, int k, const int32_t *__restrict__ offsets,
                      const int32_t *__restrict__ hb, int32_t *__restrict__ order) {
    extern __shared__ int32_t cur[];   // [k] next position of label j for this block
    for (int j = threadIdx.x; j < k; j += blockDim.x) cur[j] = offsets[j] + hb[(size_t)blockIdx.x * k + j];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * kBucketRows;
    const int64_t r1 = min(n, r0 + kBucketRows);
    for (int64_t s0 = r0; s0 < r1; s0 += 256) {
        const int64_t r = s0 + threadIdx.x;
        const bool valid = r < r1;
        const int l = valid ? labels[r] : -1 - lane;   // (distinct negatives: lanes without a row match nobody)
        const unsigned peers = __match_any_sync(0xffffffffu, l);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        const int leader = __ffs(peers) - 1;
        int pos = 0;
        for (int w = 0; w < 8; ++w) {
            if (warp == w) {
                int base = 0;
                if (valid && lane == leader) {   // one lane per distinct label of this warp: no conflicts
                    base = cur[l];
                    cur[l] = base + __popc(peers);
                }
                base = __shfl_sync(0xffffffffu, base, leader);
                pos = base + rank;
            }
            __syncthreads();
        }
        if (valid) order[pos] = (int32_t)r;
    }
}

// ------------------------------------------------------------------ per-cluster accumulation
// One CTA per chunk (<= kSegChunk rows of ONE cluster, taken from the bucketed order[]).  Thread
// layout: TPR = ld/4 threads per row (one float4 each), RPP = 256/TPR rows in flight per pass.
// Sums and the direct-form cost sum ||x - c_j||^2 are accumulated in FP64 registers; no atomics.
// part[chunk][ld] and cpart[chunk] receive the chunk partials.
__global__ void __launch_bounds__(256)
segment_sums_kernel(const float *__restrict__ X, const double *__restrict__ cent,
                    const int32_t *__restrict__ order, const int32_t *__restrict__ offsets,
                    const int32_t *__restrict__ chunk_off, int k, int ld, double *__restrict__ part,
                    double *__restrict__ cpart) {
    const int chunk = blockIdx.x;
    if (chunk >= chunk_off[k]) return;
    // binary search: largest j with chunk_off[j] <= chunk
    int lo = 0, hi = k;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (chunk_off[mid] <= chunk) lo = mid; else hi = mid;
    }
    const int j = lo;
    const int start = offsets[j] + (chunk - chunk_off[j]) * kSegChunk;
    const int end = min(offsets[j + 1], start + kSegChunk);

    __shared__ double sred[1024];  // [RPP][4 * TPR] of the current pass, for the cross-row reduction
    const int vec = ld >> 2;
    double cost = 0.0;
    for (int v0 = 0; v0 < vec; v0 += 256) {  // ld <= 1024 per pass of 256 float4 lanes
        const int tpr = min(vec - v0, 256);
        const int rpp = 256 / tpr;
        const int tv = threadIdx.x % tpr + v0;
        const int tr = threadIdx.x / tpr;
        double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
        if (tr < rpp) {
            const double c0 = cent[(size_t)j * ld + 4 * tv + 0], c1 = cent[(size_t)j * ld + 4 * tv + 1];
            const double c2 = cent[(size_t)j * ld + 4 * tv + 2], c3 = cent[(size_t)j * ld + 4 * tv + 3];
            int i = start + tr;
            // 4 rows in flight per thread
            for (; i + 3 * rpp < end; i += 4 * rpp) {
                const int ra = order[i], rb = order[i + rpp], rc = order[i + 2 * rpp], rd = order[i + 3 * rpp];
                const float4 a = __ldg(reinterpret_cast<const float4 *>(X + (size_t)ra * ld) + tv);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(X + (size_t)rb * ld) + tv);
                const float4 c = __ldg(reinterpret_cast<const float4 *>(X + (size_t)rc * ld) + tv);
                const float4 d = __ldg(reinterpret_cast<const float4 *>(X + (size_t)rd * ld) + tv);
#define KMB_ACC(q)                                                                  \
    {                                                                               \
        const double x0 = q.x, x1 = q.y, x2 = q.z, x3 = q.w;                        \
        s0 += x0; s1 += x1; s2 += x2; s3 += x3;                                     \
        const double e0 = x0 - c0, e1 = x1 - c1, e2 = x2 - c2, e3 = x3 - c3;        \
        cost = fma(e0, e0, cost); cost = fma(e1, e1, cost);                         \
        cost = fma(e2, e2, cost); cost = fma(e3, e3, cost);                         \
    }
                KMB_ACC(a) KMB_ACC(b) KMB_ACC(c) KMB_ACC(d)
            }
            for (; i < end; i += rpp) {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(X + (size_t)order[i] * ld) + tv);
                KMB_ACC(a)
            }
#undef KMB_ACC
        }
        // reduce the rpp row-lanes that share a column group (fixed order -> deterministic per chunk)
        // (scratch laid out per pass: [rpp][4 * tpr] <= 1024 doubles whatever ld is)
        if (tr < rpp) {
            double *dst = sred + (size_t)tr * (4 * tpr) + 4 * (tv - v0);
            dst[0] = s0; dst[1] = s1; dst[2] = s2; dst[3] = s3;
        }
        __syncthreads();
        for (int c = threadIdx.x; c < 4 * tpr; c += blockDim.x) {
            double s = 0.0;
            for (int r = 0; r < rpp; ++r) s += sred[(size_t)r * (4 * tpr) + c];
            part[(size_t)chunk * ld + 4 * v0 + c] = s;
        }
        __syncthreads();
    }
    // cost: block reduction in a fixed tree
    __shared__ double cred[8];
    cost = warp_sum(cost);
    if ((threadIdx.x & 31) == 0) cred[threadIdx.x >> 5] = cost;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += cred[w];
        cpart[chunk] = s;
    }
}

// One CTA per cluster: red[j*ld + c] = sum over the cluster's chunks (in chunk order) of part;
// red[k*ld + j] = count_j (as FP64, exact below 2^53); costj[j] = sum of the chunk costs.
__global__ void segment_reduce_kernel(const double *__restrict__ part, const double *__restrict__ cpart,
                                      const int32_t *__restrict__ offsets, const int32_t *__restrict__ chunk_off,
                                      int k, int ld, double *__restrict__ red, double *__restrict__ costj) {
    const int j = blockIdx.x;
    const int c0 = chunk_off[j], c1 = chunk_off[j + 1];
    // (eight loads in flight, added in chunk order: the same sum as a plain loop, without waiting for each load)
    for (int c = threadIdx.x; c < ld; c += blockDim.x) {
        double s = 0.0;
        int ch = c0;
        for (; ch + 8 <= c1; ch += 8) {
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = part[(size_t)(ch + u) * ld + c];
#pragma unroll
            for (int u = 0; u < 8; ++u) s += v[u];
        }
        for (; ch < c1; ++ch) s += part[(size_t)ch * ld + c];
        red[(size_t)j * ld + c] = s;
    }
    if (threadIdx.x == blockDim.x - 1) {   // (the last warp: with ld <= 96 it has no column work)
        double s = 0.0;
        int ch = c0;
        for (; ch + 8 <= c1; ch += 8) {
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = cpart[ch + u];
#pragma unroll
            for (int u = 0; u < 8; ++u) s += v[u];
        }
        for (; ch < c1; ++ch) s += cpart[ch];
        costj[j] = s;
        red[(size_t)k * ld + j] = (double)(offsets[j + 1] - offsets[j]);
    }
}

// Single CTA: red[k*ld + k] = sum_j costj[j] in a fixed order.
__global__ void cost_sum_kernel(const double *__restrict__ costj, int k, int ld, double *__restrict__ red) {
    __shared__ double sh[1024];
    double s = 0.0;
    for (int j = threadIdx.x; j < k; j += 1024) s += costj[j];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) red[(size_t)k * ld + k] = sh[0];
}

// ------------------------------------------------------------------ finalize + convergence
// One CTA per cluster, after the all-reduce: c_new = sum * (1/count) (reciprocal multiply, as oneDAL
// and Spark's scal(1/weightSum)); moved^2 = sequential sum_d (new-old)^2 (KMeansDALImpl.cpp:135-145);
// flags[0] += 1 when the centre moved more than tol, flags[1] += 1 when the cluster is empty (its old
// centre is copied; the host-driven candidate substitution replaces it).  The new centres go to a
// second buffer so the old ones stay available for that substitution.
__global__ void finalize_kernel(const double *__restrict__ red, const double *__restrict__ cent,
                                double *__restrict__ cent_new, int k, int ld, int d, double tol, int32_t *flags) {
    const int j = blockIdx.x;
    constexpr int kCols = 2048;   // columns per pass (any d: the sum below continues across passes)
    __shared__ double sdiff[kCols];
    const double cnt = red[(size_t)k * ld + j];
    if (cnt <= 0.0) {
        for (int c = threadIdx.x; c < d; c += blockDim.x) cent_new[(size_t)j * ld + c] = cent[(size_t)j * ld + c];
        if (threadIdx.x == 0) atomicAdd(&flags[1], 1);
        return;
    }
    const double inv = 1.0 / cnt;
    double s = 0.0;   // thread 0 only: sequential over the columns, like the reference's loop
    for (int c0 = 0; c0 < d; c0 += kCols) {
        const int c1 = min(d, c0 + kCols);
        for (int c = c0 + threadIdx.x; c < c1; c += blockDim.x) {
            const double nv = red[(size_t)j * ld + c] * inv;
            const double df = nv - cent[(size_t)j * ld + c];
            sdiff[c - c0] = df * df;
            cent_new[(size_t)j * ld + c] = nv;
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int c = c0; c < c1; ++c) s += sdiff[c - c0];
        __syncthreads();
    }
    if (threadIdx.x == 0 && !(s <= tol * tol)) atomicAdd(&flags[0], 1);
}

// ------------------------------------------------------------------ empty-cluster candidates
// dist[i] = ||x_i - c_label(i)||^2 in FP64, one warp per row.
__global__ void rowdist_kernel(const float *__restrict__ X, const double *__restrict__ cent,
                               const int32_t *__restrict__ labels, int64_t n, int ld, double *__restrict__ dist) {
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = warp_global; row < n; row += nwarps) {
        const double *cj = cent + (size_t)labels[row] * ld;
        double s = 0.0;
        for (int c = lane; c < ld; c += 32) {
            const double e = (double)X[row * ld + c] - cj[c];
            s = fma(e, e, s);
        }
        s = warp_sum(s);
        if (lane == 0) dist[row] = s;
    }
}

// ---- the m largest distances without a host round trip per candidate: radix select on the bits of the (non-negative)
// FP64 distances, 12 bits per round.  Round r looks at the rows whose key agrees with `prefix` above bit shift + 12 and
// histograms the digit (key >> shift) & 0xfff; digit_select_kernel then finds the digit that holds the `need`-th
// largest of them.  Typically two rounds narrow the threshold bin to a handful of rows; dist_collect_kernel lists every
// row at or above the bin and the host orders that short list (distance descending, row ascending).
constexpr int kDigitBits = 12;
__global__ void dist_digit_hist_kernel(const double *__restrict__ dist, int64_t n, unsigned long long prefix, int shift,
                                       int first_round, unsigned int *__restrict__ hist) {
    __shared__ unsigned int sh[1 << kDigitBits];
    for (int i = threadIdx.x; i < (1 << kDigitBits); i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = (unsigned long long)__double_as_longlong(dist[r]);
        if (first_round || (key >> (shift + kDigitBits)) == prefix)
            atomicAdd(&sh[(unsigned)(key >> shift) & ((1u << kDigitBits) - 1u)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < (1 << kDigitBits); i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}
// One CTA of 1024 threads.  out[0] = the digit d whose bin holds the need-th largest key, out[1] = keys in higher
// bins, out[2] = keys in bin d.
__global__ void digit_select_kernel(const unsigned int *__restrict__ hist, unsigned int need, unsigned int *__restrict__ out) {
    __shared__ unsigned int part[1024];
    constexpr int PER = (1 << kDigitBits) / 1024;   // digits per thread, highest digits in thread 0
    const int t = threadIdx.x;
    unsigned int mine = 0;
    for (int i = 0; i < PER; ++i) mine += hist[(1 << kDigitBits) - 1 - (t * PER + i)];
    part[t] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {   // inclusive scan over the threads (descending digits)
        const unsigned int a = t >= o ? part[t - o] : 0u;
        __syncthreads();
        part[t] += a;
        __syncthreads();
    }
    const unsigned int before = part[t] - mine;   // keys in the digits above this thread's
    if (before < need && need <= part[t]) {
        unsigned int above = before;
        for (int i = 0; i < PER; ++i) {
            const int d = (1 << kDigitBits) - 1 - (t * PER + i);
            const unsigned int c = hist[d];
            if (need <= above + c) {
                out[0] = (unsigned)d;
                out[1] = above;
                out[2] = c;
                break;
            }
            above += c;
        }
    }
}
// Lists (key, row) of every row whose key is at or above the threshold bin: (key >> shift) >= prefix.
__global__ void dist_collect_kernel(const double *__restrict__ dist, int64_t n, unsigned long long prefix, int shift,
                                    unsigned long long *__restrict__ keys, int32_t *__restrict__ rows, unsigned int *count,
                                    unsigned int cap) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = (unsigned long long)__double_as_longlong(dist[r]);
        if ((key >> shift) >= prefix) {
            const unsigned int pos = atomicAdd(count, 1u);
            if (pos < cap) { keys[pos] = key; rows[pos] = (int32_t)r; }
        }
    }
}
// out[i][0..d) = X[rows[i]][0..d)
__global__ void gather_list_kernel(const float *__restrict__ X, const int32_t *__restrict__ rows, int m, int d, int ld,
                                   float *__restrict__ out) {
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < (int64_t)m * d; e += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / d), c = (int)(e % d);
        out[e] = X[(size_t)rows[i] * ld + c];
    }
}

// ------------------------------------------------------------------ k-means|| initialisation passes
// cost[i] = min(cost[i], dist[i]) (first: cost[i] = dist[i]); part[block] = sum of the block's new costs in a fixed order.
__global__ void init_min_kernel(double *__restrict__ cost, const double *__restrict__ dist, int64_t n, int first,
                                double *__restrict__ part) {
    __shared__ double sh[256];
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = blockIdx.x * per, r1 = min(n, r0 + per);
    double s = 0.0;
    for (int64_t r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
        const double c = first ? dist[r] : fmin(cost[r], dist[r]);
        cost[r] = c;
        s += c;
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}
__global__ void sum_parts_kernel(const double *__restrict__ part, int nparts, double *__restrict__ out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < nparts; ++i) s += part[i];
        out[0] = s;
    }
}

// ------------------------------------------------------------------ cluster sizes (kmb200_assign)
__global__ void counts_to_i64_kernel(const int32_t *__restrict__ counts, int k, unsigned long long *out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < k) out[j] = (unsigned long long)counts[j];
}

// ------------------------------------------------------------------ synthetic Gaussian blobs
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// Row i (global index row_offset + i) belongs to blob splitmix64(seed ^ gi) mod k_true; each thread
// produces 4 consecutive columns: two Box-Muller pairs from one 64-bit hash per pair.
__global__ void blobs_kernel(float *__restrict__ X, int64_t nrows, int d, int ld, const float *__restrict__ centres,
                             int k_true, uint64_t seed, int64_t row_offset) {
    const int vec = ld >> 2;
    const int64_t total = nrows * vec;
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / vec;
        const int v = (int)(e - r * vec);
        const uint64_t gi = (uint64_t)(row_offset + r);
        const int blob = (int)(splitmix64(seed ^ (gi * 0x9E3779B97F4A7C15ull)) % (uint64_t)k_true);
        float out[4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint64_t bits = splitmix64(seed + 0x632BE59BD9B4E019ull * (gi * (uint64_t)vec * 2 + (uint64_t)(2 * v + h) + 1));
            const float u1 = ((uint32_t)(bits >> 40) + 1u) * (1.0f / 16777217.0f);  // (0,1]
            const float u2 = (uint32_t)(bits & 0xFFFFFFu) * (1.0f / 16777216.0f);   // [0,1)
            const float rad = sqrtf(-2.0f * __logf(u1));
            float sn, cs;
            __sincosf(6.283185307179586f * u2, &sn, &cs);
            out[2 * h] = rad * cs;
            out[2 * h + 1] = rad * sn;
        }
        float4 q;
        const int c = 4 * v;
        q.x = c + 0 < d ? out[0] + centres[(size_t)blob * d + c + 0] : 0.f;
        q.y = c + 1 < d ? out[1] + centres[(size_t)blob * d + c + 1] : 0.f;
        q.z = c + 2 < d ? out[2] + centres[(size_t)blob * d + c + 2] : 0.f;
        q.w = c + 3 < d ? out[3] + centres[(size_t)blob * d + c + 3] : 0.f;
        reinterpret_cast<float4 *>(X + r * ld)[v] = q;
    }
}

}  // namespace kmb
