// K11 — Pearson correlation of count profiles.
//
// Replaces `correlation` of kmerdb/distance.pyx:108-152, which kmerdb.distances calls once per pair of samples
// (kmerdb/__init__.py:733-742):   x_bar = sum(a) / N;  ssxx = Σ (a[i] - x_bar)^2;  ssyy likewise;  ssxy = Σ (a[i] - x_bar)(b[i] - y_bar);
// r = ssxy / sqrt(ssxx * ssyy).  The reference declares the means `float` and the sums `long double`, so as compiled its residuals and
// products are float32 operations whose results are summed without further rounding that matters.  MODE 0 restates exactly that
// (float32 conversions, subtractions and products with round-to-nearest intrinsics — no contraction — and compensated float64 sums,
// ~106 bits); MODE 1 is the same formula in float64 throughout (the textbook Pearson r).
// One launch handles a 4 x 4 block of sample pairs: every thread streams bins, keeps 16 running sums, and the CTA writes one partial
// per pair; the host adds the partials in a fixed order (results do not depend on scheduling).
#pragma once
#include "kdb_common.cuh"

namespace kdb {

constexpr int kCorrThreads = 256;
constexpr int kCorrTile = 4;

struct DD { double hi, lo; };                                   // unevaluated sum hi + lo
__device__ __forceinline__ void dd_add(DD& a, double p) {       // a += p  (Knuth two-sum; the error term goes to lo)
    const double s = __dadd_rn(a.hi, p);
    const double bb = __dadd_rn(s, -a.hi);
    const double err = __dadd_rn(__dadd_rn(a.hi, -__dadd_rn(s, -bb)), __dadd_rn(p, -bb));
    a.hi = s;
    a.lo = __dadd_rn(a.lo, err);
}
__device__ __forceinline__ void dd_merge(DD& a, const DD b) { dd_add(a, b.hi); a.lo = __dadd_rn(a.lo, b.lo); }

// sums[s] = Σ_t x[s][t]   (exact in uint64)
__global__ void __launch_bounds__(kCorrThreads) k11_sums(const unsigned long long* __restrict__ x, size_t n_bins, unsigned long long* __restrict__ sums) {
    __shared__ unsigned long long s_w[kCorrThreads / 32];
    const unsigned long long* row = x + (size_t)blockIdx.y * n_bins;
    unsigned long long acc = 0;
    for (size_t t = (size_t)blockIdx.x * kCorrThreads + threadIdx.x; t < n_bins; t += (size_t)gridDim.x * kCorrThreads) acc += row[t];
    for (int d = 16; d; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long s = 0;
        for (int w = 0; w < kCorrThreads / 32; w++) s += s_w[w];
        if (s) atomicAdd(sums + blockIdx.y, s);
    }
}

// part[blockIdx.x][a * 4 + b] = Σ over this CTA's bins of  res(i0 + a, t) * res(j0 + b, t),  res = x - mean in the arithmetic of MODE
template <int MODE>
__global__ void __launch_bounds__(kCorrThreads) k11_products(const unsigned long long* __restrict__ x, size_t n_bins, int n_samples,
                                                             const double* __restrict__ mean, int i0, int j0, DD* __restrict__ part) {
    __shared__ DD s_w[kCorrThreads / 32][kCorrTile * kCorrTile];
    const unsigned long long* ri[kCorrTile];
    const unsigned long long* rj[kCorrTile];
    double mi[kCorrTile], mj[kCorrTile];
#pragma unroll
    for (int a = 0; a < kCorrTile; a++) {
        const int si = min(i0 + a, n_samples - 1), sj = min(j0 + a, n_samples - 1);       // rows beyond the last sample repeat it (ignored by the host)
        ri[a] = x + (size_t)si * n_bins; rj[a] = x + (size_t)sj * n_bins;
        mi[a] = mean[si]; mj[a] = mean[sj];
    }
    DD acc[kCorrTile * kCorrTile];
#pragma unroll
    for (int q = 0; q < kCorrTile * kCorrTile; q++) acc[q] = DD{0.0, 0.0};
    for (size_t t = (size_t)blockIdx.x * kCorrThreads + threadIdx.x; t < n_bins; t += (size_t)gridDim.x * kCorrThreads) {
        double di[kCorrTile], dj[kCorrTile];
#pragma unroll
        for (int a = 0; a < kCorrTile; a++) {
            if (MODE == 0) {
                di[a] = (double)__fsub_rn(__ull2float_rn(ri[a][t]), (float)mi[a]);
                dj[a] = (double)__fsub_rn(__ull2float_rn(rj[a][t]), (float)mj[a]);
            } else {
                di[a] = __dadd_rn(__ull2double_rn(ri[a][t]), -mi[a]);
                dj[a] = __dadd_rn(__ull2double_rn(rj[a][t]), -mj[a]);
            }
        }
#pragma unroll
        for (int a = 0; a < kCorrTile; a++)
#pragma unroll
            for (int b = 0; b < kCorrTile; b++) {
                const double p = MODE == 0 ? (double)__fmul_rn((float)di[a], (float)dj[b]) : __dmul_rn(di[a], dj[b]);
                dd_add(acc[a * kCorrTile + b], p);
            }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < kCorrTile * kCorrTile; q++) {
        DD v = acc[q];
        for (int d = 16; d; d >>= 1) {
            DD o;
            o.hi = __shfl_down_sync(0xffffffffu, v.hi, d); o.lo = __shfl_down_sync(0xffffffffu, v.lo, d);
            if (lane < d) dd_merge(v, o);
        }
        if (lane == 0) s_w[warp][q] = v;
    }
    __syncthreads();
    if (threadIdx.x < kCorrTile * kCorrTile) {
        DD v = s_w[0][threadIdx.x];
        for (int w = 1; w < kCorrThreads / 32; w++) dd_merge(v, s_w[w][threadIdx.x]);
        part[(size_t)blockIdx.x * (kCorrTile * kCorrTile) + threadIdx.x] = v;
    }
}

}  // namespace kdb
