// K1 — line / record boundary scan (FASTA and FASTQ), sm_100a.
//
// Replaces the tokenising half of Bio.SeqIO.parse as called at kmerdb/parse.py:70-72: it finds,
// for every 16 KiB tile of text, how many lines and title lines begin before it and what kind of
// line the tile starts in, so that the count kernel can classify every byte without looking at
// anything but its own tile (plus a short backward halo).
//   k1_tile_summary : one CTA per tile, streaming read (1 B/B), block reduction
//   k1_carry_scan   : one CTA, exclusive scan of the summaries with the carry of earlier passes
#pragma once
#include "kdb_common.cuh"

namespace kdb {

struct LineAgg {
    unsigned int ls, hdr, last;
};
__device__ __forceinline__ LineAgg combine(LineAgg a, LineAgg b) {
    LineAgg r;
    r.ls = a.ls + b.ls;
    r.hdr = a.hdr + b.hdr;
    r.last = b.last ? b.last : a.last;
    return r;
}
__device__ __forceinline__ LineAgg shfl_up_agg(LineAgg v, int d) {
    LineAgg r;
    r.ls = __shfl_up_sync(0xffffffffu, v.ls, d);
    r.hdr = __shfl_up_sync(0xffffffffu, v.hdr, d);
    r.last = __shfl_up_sync(0xffffffffu, v.last, d);
    return r;
}

// Inclusive scan over the CTA (kThreads threads); returns this thread's inclusive value and the
// CTA total through *total.  `sm` needs kThreads/32 entries.
__device__ __forceinline__ LineAgg block_scan_inclusive(LineAgg v, LineAgg* sm, LineAgg* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        LineAgg o = shfl_up_agg(v, d);
        if (lane >= d) v = combine(o, v);
    }
    if (lane == 31) sm[warp] = v;
    __syncthreads();
    if (warp == 0) {
        LineAgg w = lane < (int)(blockDim.x >> 5) ? sm[lane] : LineAgg{0, 0, 0};
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            LineAgg o = shfl_up_agg(w, d);
            if (lane >= d) w = combine(o, w);
        }
        if (lane < (int)(blockDim.x >> 5)) sm[lane] = w;
    }
    __syncthreads();
    if (warp > 0) v = combine(sm[warp - 1], v);
    *total = sm[(blockDim.x >> 5) - 1];
    return v;
}

// 0x80 in every byte of w that equals the byte replicated in pat4 (exact, no false positives).
__device__ __forceinline__ uint32_t eqmask(uint32_t w, uint32_t pat4) {
    const uint32_t x = w ^ pat4;
    return ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);
}

// Line aggregate of the bytes [a, a+lim) given the byte before them.  Full chunks go four bytes at a
// time (SWAR byte compares), ragged tails byte by byte.
__device__ __forceinline__ LineAgg chunk_lines(const uint32_t (&w)[kChunk / 4], int lim, uint8_t prev) {
    LineAgg g{0, 0, 0};
    if (lim == kChunk) {
        uint32_t nl_prev = prev == '\n' ? 0x80000000u : 0u;
#pragma unroll
        for (int i = 0; i < kChunk / 4; i++) {
            const uint32_t nl = eqmask(w[i], 0x0A0A0A0Au);
            const uint32_t lsm = (nl << 8) | (nl_prev >> 24);          // 0x80 flags of the bytes that begin a line
            if (lsm) {
                const uint32_t hm = lsm & eqmask(w[i], 0x3E3E3E3Eu);   // ... with '>'
                g.ls += __popc(lsm);
                g.hdr += __popc(hm);
                g.last = (hm & (0x80000000u >> __clz(lsm))) ? 2u : 1u;
            }
            nl_prev = nl;
        }
        return g;
    }
#pragma unroll
    for (int j = 0; j < kChunk; j++) {
        const uint8_t c = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
        if (j < lim && prev == '\n') {
            g.ls++;
            const unsigned int cl = (c == '>') ? 2u : 1u;
            g.hdr += cl >> 1;
            g.last = cl;
        }
        prev = c;
    }
    return g;
}

__device__ __forceinline__ void load_chunk_global(const uint8_t* text, unsigned long long a, unsigned long long n,
                                                  uint32_t (&w)[kChunk / 4], int* lim) {
    if (a + kChunk <= n) {
        const uint4* p = reinterpret_cast<const uint4*>(text + a);
#pragma unroll
        for (int q = 0; q < kChunk / 16; q++) {
            uint4 v = __ldg(p + q);
            w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
        }
        *lim = kChunk;
    } else {
        const int m = a < n ? (int)(n - a) : 0;
#pragma unroll
        for (int i = 0; i < kChunk / 4; i++) w[i] = 0;
        for (int j = 0; j < m; j++) w[j >> 2] |= (uint32_t)text[a + j] << (8 * (j & 3));
        *lim = m;
    }
}

// One CTA per tile.  text+lo is the first byte of tile 0 of this pass.
__global__ void __launch_bounds__(kThreads) k1_tile_summary(const uint8_t* __restrict__ text, unsigned long long lo,
                                                            unsigned long long hi, TileSummary* __restrict__ out) {
    __shared__ LineAgg sm[kThreads / 32];
    const unsigned long long a = lo + (unsigned long long)blockIdx.x * kTile + (unsigned long long)threadIdx.x * kChunk;
    uint32_t w[kChunk / 4];
    int lim;
    load_chunk_global(text, a, hi, w, &lim);
    const uint8_t prev = (a == 0) ? (uint8_t)'\n' : (a <= hi ? text[a - 1] : (uint8_t)0);
    LineAgg g = chunk_lines(w, lim, prev);
    LineAgg total;
    block_scan_inclusive(g, sm, &total);
    if (threadIdx.x == 0) out[blockIdx.x] = TileSummary{total.ls, total.hdr, total.last, 0};
}

// One CTA of 1024 threads: exclusive scan of the tile summaries, seeded with (and updating) the
// carry of the passes before this one.
__global__ void __launch_bounds__(1024) k1_carry_scan(const TileSummary* __restrict__ in, TileState* __restrict__ out,
                                                      unsigned int n_tiles, ScanCarry* carry) {
    __shared__ LineAgg sm[32];
    __shared__ unsigned long long s_ls, s_hdr;
    __shared__ unsigned int s_last;
    if (threadIdx.x == 0) { s_ls = carry->ls_total; s_hdr = carry->hdr_total; s_last = carry->last; }
    __syncthreads();
    for (unsigned int base = 0; base < n_tiles; base += blockDim.x) {
        const unsigned int t = base + threadIdx.x;
        LineAgg g{0, 0, 0};
        if (t < n_tiles) { TileSummary s = in[t]; g = LineAgg{s.ls, s.hdr, s.last}; }
        LineAgg total;
        LineAgg inc = block_scan_inclusive(g, sm, &total);
        // exclusive value = inclusive of the previous thread
        LineAgg exc = shfl_up_agg(inc, 1);
        if ((threadIdx.x & 31) == 0) exc = (threadIdx.x >> 5) ? sm[(threadIdx.x >> 5) - 1] : LineAgg{0, 0, 0};
        const unsigned long long ls0 = s_ls, hdr0 = s_hdr;
        const unsigned int last0 = s_last;
        if (t < n_tiles) out[t] = TileState{ls0 + exc.ls, hdr0 + exc.hdr, exc.last ? exc.last : last0, 0};
        __syncthreads();
        if (threadIdx.x == 0) { s_ls = ls0 + total.ls; s_hdr = hdr0 + total.hdr; s_last = total.last ? total.last : last0; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { carry->ls_total = s_ls; carry->hdr_total = s_hdr; carry->last = s_last; }
}

}  // namespace kdb
