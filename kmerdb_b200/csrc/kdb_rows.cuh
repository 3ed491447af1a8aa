// K9 — .kdb body rows on the GPU.
//
// Replaces the per-row Python loop of kmerdb/__init__.py:1980-1990:
//     kdb_out.write("{0}\t{1}\t{2}\t{3}\n".format(i, kmer_id, c, f))          i == kmer_id, c = counts[i], f = c / total
// Integer work only: the text of "c\tf" (np.uint64 / np.float64 formatting, shortest round-trip repr) is made by the
// host once per DISTINCT count value (a few thousand values for 4^k rows) and handed over as a string table; the
// kernels look every row's count up in that table, size the rows, and assemble the bytes.
//   k9_row_len   : bytes of every row + per-block sums
//   k9_row_write : block-local scan, rows assembled in shared memory, coalesced copy-out
#pragma once
#include "kdb_common.cuh"

namespace kdb {

constexpr int kRowBlock = 256;                  // rows per CTA
constexpr int kRowMax = 96;                     // 2 x 20 digits + 3 separators + "count\tfreq" (<= 20 + 1 + 25)

__device__ __forceinline__ int ndigits10(unsigned long long v) {
    int n = 1;
    while (v >= 10000ull) { v /= 10000ull; n += 4; }
    return n + (v >= 1000ull ? 3 : v >= 100ull ? 2 : v >= 10ull ? 1 : 0);
}
__device__ __forceinline__ void put_decimal(char* dst, unsigned long long v, int nd) {
    for (int j = nd - 1; j >= 0; j--) { dst[j] = (char)('0' + v % 10ull); v /= 10ull; }
}
// index of `c` in the ascending table uniq[0..n)  (every count of the vector is in it)
__device__ __forceinline__ unsigned int uniq_index(const unsigned long long* __restrict__ uniq, unsigned int n, unsigned long long c) {
    unsigned int lo = 0, hi = n;
    while (hi - lo > 1) { const unsigned int mid = (lo + hi) >> 1; if (uniq[mid] <= c) lo = mid; else hi = mid; }
    return lo;
}

__global__ void __launch_bounds__(kRowBlock) k9_row_len(const unsigned long long* __restrict__ counts, unsigned long long first, unsigned long long n,
                                                        const unsigned long long* __restrict__ uniq, const unsigned int* __restrict__ str_off,
                                                        unsigned int n_uniq, unsigned int* __restrict__ row_idx, unsigned int* __restrict__ block_len,
                                                        unsigned int* bad) {
    __shared__ unsigned int s_sum;
    if (threadIdx.x == 0) s_sum = 0;
    __syncthreads();
    const unsigned long long i = (unsigned long long)blockIdx.x * kRowBlock + threadIdx.x;
    unsigned int len = 0;
    if (i < n) {
        const unsigned long long c = counts[i];
        const unsigned int u = uniq_index(uniq, n_uniq, c);
        if (uniq[u] != c) atomicOr(bad, 1u);                              // table does not cover the vector
        row_idx[i] = u;
        len = 2 * ndigits10(first + i) + 3 + (str_off[u + 1] - str_off[u]);
        if (len > kRowMax) { atomicOr(bad, 2u); len = 0; }
    }
    for (int d = 16; d; d >>= 1) len += __shfl_xor_sync(0xffffffffu, len, d);
    if ((threadIdx.x & 31) == 0 && len) atomicAdd(&s_sum, len);
    __syncthreads();
    if (threadIdx.x == 0) block_len[blockIdx.x] = s_sum;
}

__global__ void __launch_bounds__(kRowBlock) k9_row_write(unsigned long long first, unsigned long long n, const unsigned int* __restrict__ row_idx,
                                                          const unsigned int* __restrict__ str_off, const char* __restrict__ str_bytes,
                                                          const unsigned long long* __restrict__ block_off, char* __restrict__ out) {
    __shared__ __align__(16) char s_rows[kRowBlock * kRowMax];
    __shared__ unsigned int s_warp[kRowBlock / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long i = (unsigned long long)blockIdx.x * kRowBlock + threadIdx.x;
    unsigned int len = 0, u = 0, nd = 0;
    if (i < n) { u = row_idx[i]; nd = ndigits10(first + i); len = 2 * nd + 3 + (str_off[u + 1] - str_off[u]); }
    unsigned int inc = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned int base = inc - len, total = 0;
#pragma unroll
    for (int w = 0; w < kRowBlock / 32; w++) { const unsigned int c = s_warp[w]; if (w < warp) base += c; total += c; }
    if (i < n) {
        char* p = s_rows + base;
        put_decimal(p, first + i, nd); p[nd] = '\t';
        put_decimal(p + nd + 1, first + i, nd); p[2 * nd + 1] = '\t';
        const unsigned int s0 = str_off[u], sl = str_off[u + 1] - s0;
        for (unsigned int j = 0; j < sl; j++) p[2 * nd + 2 + j] = str_bytes[s0 + j];
        p[2 * nd + 2 + sl] = '\n';
    }
    __syncthreads();
    char* dst = out + block_off[blockIdx.x];
    for (unsigned int j = threadIdx.x; j < total; j += kRowBlock) dst[j] = s_rows[j];
}

}  // namespace kdb
