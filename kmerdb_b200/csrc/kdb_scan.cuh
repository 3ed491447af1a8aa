// Line / record boundary scan helpers (FASTA and FASTQ), sm_100a: the line aggregate (line starts, title lines, class of the
// right-most line start) and its scan operator, used by the single-pass tokeniser (kdb_encode2.cuh), which replaces the
// tokenising half of Bio.SeqIO.parse as called at kmerdb/parse.py:70-72.  (The r01 summary pass k1_tile_summary / k1_carry_scan —
// a second read of the whole text — is gone: the prefix now comes from a decoupled look-back inside the tokeniser.)
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

}  // namespace kdb
