// Finalisation kernels: u32->u64 fold, unique / nullomer reductions (record lengths are checked inside the tokeniser, kdb_encode2.cuh).
// Replaces kmerdb/parse.py:139-146 (nullomer_array, unique_kmers, min/max read length) and the
// record-length abort of kmerdb/kmer.py:461-463.
#pragma once
#include "kdb_common.cuh"

namespace kdb {

// tab64 += tab32; tab32 = 0.  Run before any bin of tab32 can wrap (see DESIGN.md "fold rule").
__global__ void __launch_bounds__(256) k3_fold(unsigned int* __restrict__ tab32, unsigned long long* __restrict__ tab64,
                                               unsigned long long nbins) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * 4;
    for (unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < nbins; i += stride) {
        if (i + 4 <= nbins) {
            uint4 v = *reinterpret_cast<uint4*>(tab32 + i);
            if (v.x | v.y | v.z | v.w) {
                ulonglong2 a = *reinterpret_cast<ulonglong2*>(tab64 + i);
                ulonglong2 b = *reinterpret_cast<ulonglong2*>(tab64 + i + 2);
                a.x += v.x; a.y += v.y; b.x += v.z; b.y += v.w;
                *reinterpret_cast<ulonglong2*>(tab64 + i) = a;
                *reinterpret_cast<ulonglong2*>(tab64 + i + 2) = b;
                *reinterpret_cast<uint4*>(tab32 + i) = make_uint4(0, 0, 0, 0);
            }
        } else {
            for (unsigned long long j = i; j < nbins; j++) { tab64[j] += tab32[j]; tab32[j] = 0; }
        }
    }
}

// The fold that ends a job: FIRST = tab64 holds nothing yet (it is written, not read: no zeroing pass, no read);
// COUNT = also count the non-zero bins of the result (unique_kmers, parse.py:141) so the table is not read again.
template <bool FIRST, bool COUNT>
__global__ void __launch_bounds__(256) k3_fold_final(unsigned int* __restrict__ tab32, unsigned long long* __restrict__ tab64,
                                                     unsigned long long nbins, Accum* acc) {
    unsigned long long nz = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * 4;
    for (unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < nbins; i += stride) {
        if (i + 4 <= nbins) {
            const uint4 v = *reinterpret_cast<uint4*>(tab32 + i);
            ulonglong2 a = make_ulonglong2(0, 0), b = make_ulonglong2(0, 0);
            if (!FIRST) { a = *reinterpret_cast<ulonglong2*>(tab64 + i); b = *reinterpret_cast<ulonglong2*>(tab64 + i + 2); }
            a.x += v.x; a.y += v.y; b.x += v.z; b.y += v.w;
            if (FIRST || (v.x | v.y | v.z | v.w)) {
                *reinterpret_cast<ulonglong2*>(tab64 + i) = a;
                *reinterpret_cast<ulonglong2*>(tab64 + i + 2) = b;
            }
            if (v.x | v.y | v.z | v.w) *reinterpret_cast<uint4*>(tab32 + i) = make_uint4(0, 0, 0, 0);
            if (COUNT) nz += (a.x != 0) + (a.y != 0) + (b.x != 0) + (b.y != 0);
        } else {
            for (unsigned long long j = i; j < nbins; j++) {
                const unsigned long long x = (FIRST ? 0ull : tab64[j]) + tab32[j];
                tab64[j] = x; tab32[j] = 0;
                if (COUNT) nz += x != 0;
            }
        }
    }
    if (COUNT) {
        for (int d = 16; d; d >>= 1) nz += __shfl_xor_sync(0xffffffffu, nz, d);
        if ((threadIdx.x & 31) == 0 && nz) atomicAdd(&acc->unique, nz);
    }
}

// unique_kmers = count_nonzero(counts)  (parse.py:141)
__global__ void __launch_bounds__(256) k5_count_nonzero(const unsigned long long* __restrict__ tab64, unsigned long long nbins, Accum* acc) {
    unsigned long long c = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < nbins; i += stride) c += tab64[i] != 0;
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&acc->unique, c);
}

// nullomer_array = arange(4^k)[counts == 0]  (parse.py:139-140), ascending.  Two passes over
// fixed 2048-bin blocks: per-block zero counts, host-side exclusive scan of the (small) block
// counts, then an ordered write.
constexpr int kNullBlock = 2048;
__global__ void __launch_bounds__(256) k6_null_count(const unsigned long long* __restrict__ tab64, unsigned long long nbins,
                                                     unsigned int* __restrict__ block_zero) {
    __shared__ unsigned int s;
    if (threadIdx.x == 0) s = 0;
    __syncthreads();
    const unsigned long long base = (unsigned long long)blockIdx.x * kNullBlock;
    unsigned int c = 0;
    for (int j = threadIdx.x; j < kNullBlock; j += 256) { const unsigned long long i = base + j; if (i < nbins && tab64[i] == 0) c++; }
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s, c);
    __syncthreads();
    if (threadIdx.x == 0) block_zero[blockIdx.x] = s;
}
__global__ void __launch_bounds__(256) k6_null_write(const unsigned long long* __restrict__ tab64, unsigned long long nbins,
                                                     const unsigned long long* __restrict__ block_off, unsigned long long* __restrict__ out,
                                                     unsigned long long cap) {
    __shared__ unsigned int s_warp[8];
    const unsigned long long base = (unsigned long long)blockIdx.x * kNullBlock;
    unsigned long long off = block_off[blockIdx.x];
    for (int j0 = 0; j0 < kNullBlock; j0 += 256) {
        const unsigned long long i = base + j0 + threadIdx.x;
        const bool z = i < nbins && tab64[i] == 0;
        const unsigned int bal = __ballot_sync(0xffffffffu, z);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        unsigned int before = 0, total = 0;
        for (int w = 0; w < 8; w++) { const unsigned int x = s_warp[w]; if (w < warp) before += x; total += x; }
        if (z) { const unsigned long long pos = off + before + __popc(bal & ((1u << lane) - 1u)); if (pos < cap) out[pos] = i; }
        off += total;
        __syncthreads();
    }
}

}  // namespace kdb
