// Synthetic benchmark input: reads of i.i.d. uniform ACGT letters from a counter-based
// generator (Philox-4x32-10, key = seed, counter = (read index, 64-letter block)), so any read can
// be regenerated bit-identically on the GPU and on the host.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace kdb {

struct Philox4 { uint32_t v[4]; };

__host__ __device__ inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    Philox4 o; o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}

// Byte `off` of record `read` (0-based) in the synthetic text.
struct SynthCache { uint64_t read; uint32_t blk; Philox4 p; bool valid; };

__host__ __device__ inline uint8_t synth_byte(uint64_t read, uint32_t off, int read_len, int fastq, uint64_t seed, SynthCache& cache) {
    const uint32_t title = 11;   // marker + 9 digits + LF
    if (off == 0) return fastq ? '@' : '>';
    if (off < 10) {
        uint64_t v = read % 1000000000ull;
        uint32_t div = 1;
        for (uint32_t i = 0; i < 9 - off; i++) div *= 10;
        return (uint8_t)('0' + (v / div) % 10);
    }
    if (off == 10) return '\n';
    uint32_t j = off - title;
    if (j < (uint32_t)read_len) {
        const uint32_t blk = j >> 6;
        if (!cache.valid || cache.read != read || cache.blk != blk) {
            cache.p = philox4x32_10((uint32_t)read, (uint32_t)(read >> 32), blk, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
            cache.read = read; cache.blk = blk; cache.valid = true;
        }
        const uint32_t w = cache.p.v[(j & 63) >> 4];
        const uint32_t code = (w >> (2 * (j & 15))) & 3u;
        return (uint8_t)("ACGT"[code]);
    }
    j -= (uint32_t)read_len;
    if (j == 0) return '\n';
    // FASTQ tail: "+\n" quality "\n"
    if (j == 1) return '+';
    if (j == 2) return '\n';
    if (j < 3 + (uint32_t)read_len) return 'I';
    return '\n';
}

__host__ __device__ inline uint32_t synth_record_bytes(int read_len, int fastq) {
    return fastq ? (uint32_t)(11 + read_len + 1 + 2 + read_len + 1) : (uint32_t)(11 + read_len + 1);
}

__global__ void __launch_bounds__(256) k_synth(uint8_t* __restrict__ out, uint64_t first_read, uint64_t n_reads, int read_len, int fastq,
                                               uint64_t seed) {
    const uint32_t rb = synth_record_bytes(read_len, fastq);
    const uint64_t total = n_reads * rb;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * 16;
    SynthCache cache; cache.valid = false; cache.read = 0; cache.blk = 0;
    for (uint64_t g = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16; g < total; g += stride) {
        uint32_t w[4] = {0, 0, 0, 0};
        uint64_t rec = g / rb;
        uint32_t off = (uint32_t)(g - rec * rb);
        const int m = total - g >= 16 ? 16 : (int)(total - g);
        for (int j = 0; j < m; j++) {
            w[j >> 2] |= (uint32_t)synth_byte(first_read + rec, off, read_len, fastq, seed, cache) << (8 * (j & 3));
            if (++off == rb) { off = 0; rec++; }
        }
        if (m == 16 && ((uintptr_t)(out + g) & 15) == 0) *reinterpret_cast<uint4*>(out + g) = make_uint4(w[0], w[1], w[2], w[3]);
        else for (int j = 0; j < m; j++) out[g + j] = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
    }
}

// ---- microbenchmarks behind the roofline discussion ------------------------
// mode 0: random red.global.add.u32 into a table; mode 1: random shared-memory atomicAdd (64 KiB
// table per CTA); mode 2: streaming read only (sum), for the input side.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__global__ void __launch_bounds__(256) k_micro_red_global(unsigned int* tab, uint32_t mask, uint32_t per_thread) {
    uint32_t x = mix32(blockIdx.x * blockDim.x + threadIdx.x + 1u);
    for (uint32_t i = 0; i < per_thread; i++) {
        x = x * 1664525u + 1013904223u;
        atomicAdd(tab + (mix32(x) & mask), 1u);
    }
}
__global__ void __launch_bounds__(256) k_micro_atoms_shared(unsigned int* sink, uint32_t per_thread) {
    extern __shared__ unsigned int s_tab[];
    for (int i = threadIdx.x; i < 16384; i += blockDim.x) s_tab[i] = 0;
    __syncthreads();
    uint32_t x = mix32(blockIdx.x * blockDim.x + threadIdx.x + 1u);
    for (uint32_t i = 0; i < per_thread; i++) {
        x = x * 1664525u + 1013904223u;
        atomicAdd(s_tab + (mix32(x) & 16383u), 1u);
    }
    __syncthreads();
    unsigned int s = 0;
    for (int i = threadIdx.x; i < 16384; i += blockDim.x) s += s_tab[i];
    if (s == 0xffffffffu) sink[0] = s;
}

// mode 2 / 3: the same random shared-memory atomicAdd, but the table is spread over the shared memories of a thread-block
// cluster (8 / 16 CTAs x 64 KiB): red.shared::cluster on distributed shared memory.  Decides whether a cluster-wide count
// table (an 8- or 16-way partition instead of a 512-way one) could beat partition + count.
template <int CSIZE>
__global__ void __launch_bounds__(256) k_micro_atoms_cluster(unsigned int* sink, uint32_t per_thread) {
    extern __shared__ unsigned int s_tab[];
    for (int i = threadIdx.x; i < 16384; i += blockDim.x) s_tab[i] = 0;
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    uint32_t x = mix32(blockIdx.x * blockDim.x + threadIdx.x + 1u);
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(s_tab);
    for (uint32_t i = 0; i < per_thread; i++) {
        x = x * 1664525u + 1013904223u;
        const uint32_t h = mix32(x);
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(base + 4u * (h & 16383u)), "r"((h >> 14) & (uint32_t)(CSIZE - 1)));
        asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" ::"r"(remote), "r"(1u) : "memory");
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    unsigned int s = 0;
    for (int i = threadIdx.x; i < 16384; i += blockDim.x) s += s_tab[i];
    if (s == 0xffffffffu) sink[0] = s;
}

}  // namespace kdb