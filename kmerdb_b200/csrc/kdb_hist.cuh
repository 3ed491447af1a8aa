// K3 — rolling (canonical) k-mer ids from the packed 2-bit stream + dense 4^k histogram, sm_100a.
//
// Replaces kmer.kmer_to_id (kmerdb/kmer.py:297-317: fwd id, reverse-complement id, min) and the
// count loop `counts[kmer_id] += 1` (kmerdb/parse.py:133-136).  One thread owns 32 consecutive
// letters of a tile's packed region (two code words + the word before them as the k-1 letter halo)
// and emits one update per set W bit.  K is a template parameter so every shift is an immediate.
//   k <= 7 : per-CTA shared-memory table (ATOMS), flushed once per CTA with RED
//   k = 8  : the same with the 65536 bins as packed 16-bit counters that hand their overflow to the global table (add16)
//   k >= 9 : red.global.add.u32 into the u32 working table (L2-resident up to k = 12)
// N windows in expansion mode (kmer.py:550-565) take a generic per-letter path.
#pragma once
#include "kdb_common.cuh"
#include "kdb_encode.cuh"

namespace kdb {

constexpr int kHistThreads = 512;     // 512 threads x 32 letters = one full tile

struct HistParams {
    const uint32_t* packed;
    unsigned int n_tiles;
    int skip_n;
    int smear;                  // k - (k the packed stream was encoded for): see smear_w
    unsigned int* tab32;
    Accum* acc;
};

// W bits of a larger k from the W bits of the k the stream was encoded for: a window of k+d letters is complete
// and clean exactly when the d+1 windows of k letters ending at its last d+1 letters are (k >= 2: consecutive
// windows overlap, so they lie in one record).  wb holds 32 letters MSB first, prevw the 32 letters before them.
__device__ __forceinline__ uint32_t smear_w(uint32_t prevw, uint32_t wb, int d) {
    if (d == 0) return wb;
    unsigned long long x = ((unsigned long long)prevw << 32) | wb;
    const uint32_t n = (uint32_t)d + 1;
    uint32_t w = 1;
    while (2 * w <= n) { x &= x >> w; w *= 2; }
    if (n > w) x &= x >> (n - w);
    return (uint32_t)x;
}

// reverse complement of 16 packed letters (first letter in the top bits)
__device__ __forceinline__ uint32_t rc_word(uint32_t x) {
    uint32_t y = __brev(~x);
    return ((y >> 1) & 0x55555555u) | ((y & 0x55555555u) << 1);
}

// low 32 bits of the 96-bit value (w2:w1:w0) >> S, S a compile-time constant in [0, 94]
template <int S>
__device__ __forceinline__ uint32_t shr96(uint32_t w2, uint32_t w1, uint32_t w0) {
    if constexpr (S < 32) return __funnelshift_r(w0, w1, S);
    else if constexpr (S < 64) return __funnelshift_r(w1, w2, S - 32);
    else return w2 >> (S - 64);
}

// 32 bits of the 96-bit string w0:w1:w2 (most significant first) starting at bit o; o is a compile-time constant after unrolling.
__device__ __forceinline__ uint32_t ext96(uint32_t w0, uint32_t w1, uint32_t w2, int o) {
    return o < 32 ? (o == 0 ? w0 : __funnelshift_l(w1, w0, o)) : (o < 64 ? (o == 32 ? w1 : __funnelshift_l(w2, w1, o - 32)) : (w2 << (o - 64)));
}

// 65536 bins as 16-bit counters, two per shared word (128 KiB): k = 8 directly, k = 13 / 14 per bucket (k3p_count16).  The counters are exact: every update is one 32-bit atomicAdd that returns the word it found, so the
// thread whose +1 wraps the low counter (it found 0xFFFF there) knows it — it moves 65536 to the global bin and takes the
// carry back out of the high counter; an update that carries out of (or borrows into) the 32-bit word knows that too from
// the value it found and moves 65536 to / from the global bin of the high counter.  The word is an integer mod 2^32 and
// every crossing of 2^32 is seen by exactly the update that caused it, so  bin = counter + global part  holds in any
// interleaving.
// what the update that found `old` in word w owes when its +1 (at bit sh) wrapped its 16-bit counter
__device__ __forceinline__ void add16_fix(unsigned int* s_tab, uint32_t w, uint32_t sh, uint32_t old, unsigned int* gdst) {
    if (sh == 0) {
        atomicAdd(gdst + 2 * w, 65536u);
        if (old == 0xFFFFFFFFu) atomicAdd(gdst + 2 * w + 1, 65536u);       // the carry also ran out of the word
        const uint32_t old2 = atomicSub(&s_tab[w], 0x10000u);               // take the carry back out of the high counter
        if ((old2 >> 16) == 0) atomicSub(gdst + 2 * w + 1, 65536u);        // (borrowed from beyond the word)
    } else atomicAdd(gdst + 2 * w + 1, 65536u);
}
__device__ __forceinline__ void add16(unsigned int* s_tab, uint32_t bin, unsigned int* gdst) {
    const uint32_t w = bin >> 1, sh = (bin & 1u) << 4;
    const uint32_t old = atomicAdd(&s_tab[w], 1u << sh);
    if (((old >> sh) & 0xFFFFu) == 0xFFFFu) add16_fix(s_tab, w, sh, old, gdst);    // once per 65536 updates of a bin
}

template <int K, bool CANON, bool SMEM, int E>
__device__ __forceinline__ void emit_window(unsigned int* tab, unsigned int* gtab, uint32_t prev, uint32_t cur0, uint32_t cur1,
                                            uint32_t r2, uint32_t r1, uint32_t r0, uint32_t wb) {
    constexpr uint32_t mask = K == 16 ? 0xffffffffu : ((1u << (2 * K)) - 1u);
    if (wb & (0x80000000u >> E)) {
        uint32_t id = (E < 16 ? __funnelshift_r(cur0, prev, 30 - 2 * (E & 15)) : __funnelshift_r(cur1, cur0, 30 - 2 * (E & 15))) & mask;
        if (CANON) {
            const uint32_t rc = shr96<34 + 2 * E - 2 * K>(r2, r1, r0) & mask;
            id = rc < id ? rc : id;
        }
        if (SMEM && K == 8) add16(tab, id, gtab);
        else atomicAdd(tab + id, 1u);
    }
}

template <int K, bool CANON, bool SMEM, int E>
struct EmitAll {
    static __device__ __forceinline__ void run(unsigned int* tab, unsigned int* gtab, uint32_t prev, uint32_t cur0, uint32_t cur1,
                                               uint32_t r2, uint32_t r1, uint32_t r0, uint32_t wb) {
        emit_window<K, CANON, SMEM, E>(tab, gtab, prev, cur0, cur1, r2, r1, r0, wb);
        EmitAll<K, CANON, SMEM, E + 1>::run(tab, gtab, prev, cur0, cur1, r2, r1, r0, wb);
    }
};
template <int K, bool CANON, bool SMEM>
struct EmitAll<K, CANON, SMEM, 32> {
    static __device__ __forceinline__ void run(unsigned int*, unsigned int*, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {}
};

// kmer.py:550-565, :586-621: every ACGT substitution of the N positions gets +1.
__device__ __noinline__ unsigned long long expand_window(unsigned int* tab, int k, int canon, uint32_t base, uint32_t nmask) {
    const int nn = __popc(nmask);
    const unsigned long long combos = 1ull << (2 * nn);
    for (unsigned long long m = 0; m < combos; m++) {
        uint32_t f = base, rest = nmask;
        unsigned long long mm = m;
        while (rest) {
            const int b = __ffs(rest) - 1;
            rest &= rest - 1;
            f |= (uint32_t)(mm & 3) << (2 * b);
            mm >>= 2;
        }
        uint32_t id = f;
        if (canon) { const uint32_t r = revcomp32(f, k); id = r < f ? r : f; }
        atomicAdd(tab + id, 1u);
    }
    return combos;
}

template <int K, bool CANON, bool SMEM>
__global__ void __launch_bounds__((SMEM && K == 8) ? 2 * kHistThreads : kHistThreads) k3_count(const HistParams P) {          // blockDim.x = kHistThreads, or twice that (two tiles side by side)
    extern __shared__ unsigned int s_tab[];
    __shared__ unsigned long long s_red[2 * kHistThreads / 32];
    constexpr uint32_t nbins = 1u << (2 * (K < 16 ? K : 15));     // SMEM only for K <= 8
    constexpr bool PACK = SMEM && K == 8;                         // k = 8: 65536 bins as packed 16-bit counters (add16)
    constexpr uint32_t nwords = PACK ? nbins / 2 : nbins;
    constexpr uint32_t mask = K == 16 ? 0xffffffffu : ((1u << (2 * K)) - 1u);
    const int tid = threadIdx.x & (kHistThreads - 1), side = threadIdx.x / kHistThreads, sides = blockDim.x / kHistThreads;
    unsigned int* tab = SMEM ? s_tab : P.tab32;
    if (SMEM) {
        for (uint32_t i = threadIdx.x; i < nwords; i += blockDim.x) s_tab[i] = 0;
        __syncthreads();
    }
    unsigned long long kmers = 0;
    for (unsigned int tile = blockIdx.x * sides + side; tile < P.n_tiles; tile += gridDim.x * sides) {
        const uint32_t* reg = P.packed + (size_t)tile * kPackStride;
        const uint32_t n = reg[0];
        if ((uint32_t)tid * 32 >= n) continue;
        const uint32_t* codes = reg + kPackHdr;
        const uint32_t cur0 = codes[2 * tid];
        const uint32_t cur1 = (uint32_t)tid * 32 + 16 < n ? codes[2 * tid + 1] : 0u;
        const uint32_t prev = tid ? codes[2 * tid - 1] : reg[1];
        uint32_t wb = reg[kPackHdr + kPackCodes + tid];
        if (P.smear) wb = smear_w(tid ? reg[kPackHdr + kPackCodes + tid - 1] : reg[3], wb, P.smear);
        bool generic = false;
        uint32_t nn_cur = 0, nn_prev = 0;
        if (!P.skip_n) {                                   // expansion mode: windows may hold N
            const uint32_t* nnw = reg + kPackHdr + kPackCodes + kPackBits;
            nn_cur = nnw[tid];
            nn_prev = tid ? nnw[tid - 1] : reg[2];
            generic = (nn_cur | (nn_prev & ((1u << (K - 1)) - 1u))) != 0;
        }
        if (!generic) {
            const uint32_t r2 = CANON ? rc_word(cur1) : 0u, r1 = CANON ? rc_word(cur0) : 0u, r0 = CANON ? rc_word(prev) : 0u;
            if (SMEM) {
                // shared tables: ids top-aligned (no masks), and a letter without a window bumps a per-lane dummy word behind the
                // table, so that the atomics are unconditional and pipeline (see k3p_partition, kdb_part.cuh)
                unsigned int* const dummy = s_tab + nwords + (threadIdx.x & 31);
#pragma unroll
                for (int e = 0; e < 32; e++) {
                    const bool valid = (wb & (0x80000000u >> e)) != 0;
                    uint32_t m = ext96(prev, cur0, cur1, 2 * (17 + e - K));
                    if (CANON) { const uint32_t rc = ext96(r2, r1, r0, 62 - 2 * e); m = rc < m ? rc : m; }
                    uint32_t id = m >> (32 - 2 * K);
                    asm volatile("" : "+r"(id));                         // (keeps the id arithmetic out of a branch built for the select)
                    if (!PACK) atomicAdd(valid ? s_tab + id : dummy, 1u);
                    else {
                        const uint32_t w = id >> 1, sh = (id & 1u) << 4;
                        const uint32_t old = atomicAdd(valid ? s_tab + w : dummy, 1u << sh);
                        if (valid && ((old >> sh) & 0xFFFFu) == 0xFFFFu) add16_fix(s_tab, w, sh, old, P.tab32);
                    }
                }
            } else EmitAll<K, CANON, SMEM, 0>::run(tab, P.tab32, prev, cur0, cur1, r2, r1, r0, wb);
            kmers += __popc(wb);
        } else {
            const unsigned long long S0 = ((unsigned long long)prev << 32) | cur0, S1 = ((unsigned long long)cur0 << 32) | cur1;
            const unsigned long long NS = ((unsigned long long)nn_prev << 32) | nn_cur;
            for (int e = 0; e < 32; e++) {
                if (!(wb & (0x80000000u >> e))) continue;
                const uint32_t f = (uint32_t)((e < 16 ? S0 : S1) >> (30 - 2 * (e & 15))) & mask;
                const uint32_t nm = (uint32_t)(NS >> (31 - e)) & ((1u << K) - 1u);
                unsigned int* const gen = PACK ? P.tab32 : tab;       // (the generic path counts straight into the global table when the shared one is packed)
                if (nm) kmers += expand_window(gen, K, CANON, f, nm);
                else {
                    uint32_t id = f;
                    if (CANON) { const uint32_t r = revcomp32(f, K); id = r < f ? r : f; }
                    atomicAdd(gen + id, 1u);
                    kmers++;
                }
            }
        }
    }
    for (int d = 16; d; d >>= 1) kmers += __shfl_xor_sync(0xffffffffu, kmers, d);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = kmers;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long s = 0;
        for (int w = 0; w < (int)blockDim.x / 32; w++) s += s_red[w];
        if (s) atomicAdd(&P.acc->total_kmers, s);
    }
    if (SMEM) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < nwords; i += blockDim.x) {
            const unsigned int x = s_tab[i];
            if (PACK) { if (x & 0xFFFFu) atomicAdd(P.tab32 + 2 * i, x & 0xFFFFu); if (x >> 16) atomicAdd(P.tab32 + 2 * i + 1, x >> 16); }
            else if (x) atomicAdd(P.tab32 + i, x);
        }
    }
}

typedef void (*k3_fn)(const HistParams);

template <int K>
__host__ inline k3_fn k3_select(bool canon) {
    constexpr bool smem = K <= 8;
    return canon ? (k3_fn)k3_count<K, true, smem> : (k3_fn)k3_count<K, false, smem>;
}

__host__ inline k3_fn k3_kernel(int k, bool canon) {
    switch (k) {
        case 1: return k3_select<1>(canon);   case 2: return k3_select<2>(canon);
        case 3: return k3_select<3>(canon);   case 4: return k3_select<4>(canon);
        case 5: return k3_select<5>(canon);   case 6: return k3_select<6>(canon);
        case 7: return k3_select<7>(canon);   case 8: return k3_select<8>(canon);
        case 9: return k3_select<9>(canon);   case 10: return k3_select<10>(canon);
        case 11: return k3_select<11>(canon); case 12: return k3_select<12>(canon);
        case 13: return k3_select<13>(canon); case 14: return k3_select<14>(canon);
        case 15: return k3_select<15>(canon); case 16: return k3_select<16>(canon);
    }
    return nullptr;
}

}  // namespace kdb
