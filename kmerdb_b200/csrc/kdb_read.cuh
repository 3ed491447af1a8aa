// K10 — the body of a .kdb file back into a count vector.
//
// Replaces the row loop of KDBReader._slurp, sort=False (kmerdb/fileutil.py:333-377): for row j of the body,
//   l = line.rstrip().split("\t"); assert len(l) == 4; x, kmer_id, count = int(l[0]), int(l[1]), int(l[2]);
//   assert j == x; kmer_ids[j] = kmer_id; counts[kmer_id] = count          (frequency = count / 4^k is host arithmetic)
// Rows are found by their line feeds: row r starts after the r-th '\n' of the body.  Pass 1 counts the line feeds of
// each 16 KiB block; the host scans the few thousand block counts; pass 2 parses one row per line feed.
// The text is read straight from global memory (a warp's rows are neighbours, the lines stay in L1).
#pragma once
#include "kdb_common.cuh"

namespace kdb {

constexpr int kReadThreads = 256;
constexpr int kReadPerThread = 64;                                   // text bytes per thread: four 16-byte loads
constexpr int kReadBlock = kReadThreads * kReadPerThread;

enum : unsigned int { R_COLUMNS = 1, R_INT = 2, R_INDEX = 3, R_RANGE = 4, R_DUP = 5 };   // first failing check of a row, in the reference's order

// 64-bit mask of the line feeds among the 64 bytes at `base` (bit i = byte base + i); bytes at or after n count as absent.
__device__ __forceinline__ unsigned long long newline_mask(const uint8_t* __restrict__ text, size_t n, size_t base) {
    unsigned long long m = 0;
    if (base + kReadPerThread <= n && (reinterpret_cast<uintptr_t>(text + base) & 15) == 0) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + base) + q);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t x = w[j] ^ 0x0A0A0A0Au;                 // zero byte where the text holds '\n'
                const uint32_t z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);   // 0x80 in exactly those bytes
                const unsigned long long bits = ((z >> 7) & 1u) | ((z >> 14) & 2u) | ((z >> 21) & 4u) | ((z >> 28) & 8u);
                m |= bits << (16 * q + 4 * j);
            }
        }
    } else {
        for (int i = 0; i < kReadPerThread; i++)
            if (base + i < n && text[base + i] == '\n') m |= 1ull << i;
    }
    return m;
}

__global__ void __launch_bounds__(kReadThreads) k10_count_rows(const uint8_t* __restrict__ text, size_t n, unsigned int* __restrict__ blk_rows) {
    __shared__ unsigned int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const size_t base = (size_t)blockIdx.x * kReadBlock + (size_t)threadIdx.x * kReadPerThread;
    unsigned int c = base < n ? __popcll(newline_mask(text, n, base)) : 0u;
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) blk_rows[blockIdx.x] = s_cnt;
}

struct ReadParams {
    const uint8_t* text;
    size_t n;
    const unsigned long long* blk_off;       // line feeds before each block (exclusive scan of k10_count_rows)
    unsigned long long first_row;            // ordinal of the row that starts at text[0]
    unsigned long long n_bins;               // 4^k: rows beyond it are never read by the reference (for j in range(N))
    unsigned long long* kmer_ids;            // [n_bins]  kmer_ids[j] = id of row j
    unsigned long long* counts;              // [n_bins]  counts[id] = count
    unsigned int* seen;                      // [n_bins / 32] one bit per id that a row has set
    unsigned long long* err;                 // min over failing rows of (row << 3 | class)
};

__device__ __forceinline__ void parse_row(const ReadParams& P, size_t start, unsigned long long row) {
    if (start >= P.n || row >= P.n_bins) return;
    unsigned long long v[3] = {0, 0, 0};
    unsigned int nd[3] = {0, 0, 0};
    bool bad[3] = {false, false, false};
    unsigned int tabs = 0, tabs_nz = 0;
    bool any = false;
    for (size_t p = start; p < P.n; p++) {
        const unsigned int c = __ldg(P.text + p);
        if (c == '\n') break;
        if (c == '\t') { tabs++; continue; }
        const bool ws = c == ' ' || c == '\r' || c == 0x0B || c == 0x0C;
        if (!ws) { any = true; tabs_nz = tabs; }                       // rstrip(): tabs after the last visible byte do not split
        if (tabs < 3) {
            const unsigned int d = c - '0';
            if (d > 9u || v[tabs] > (0xFFFFFFFFFFFFFFFFull - d) / 10ull) bad[tabs] = true;
            else { v[tabs] = v[tabs] * 10ull + d; nd[tabs]++; }
        }
    }
    unsigned int cls = 0;
    const unsigned int columns = any ? tabs_nz + 1 : 1;
    if (columns != 4) cls = R_COLUMNS;                                  // fileutil.py:351
    else if (bad[0] || !nd[0] || bad[1] || !nd[1]) cls = R_INT;         // :357-358
    else if (v[0] != row) cls = R_INDEX;                                // :359
    else if (bad[2] || !nd[2]) cls = R_INT;                             // :361
    else if (v[1] >= P.n_bins) cls = R_RANGE;                           // :365 (numpy IndexError)
    if (!cls) {
        const unsigned int old = atomicOr(P.seen + (v[1] >> 5), 1u << (v[1] & 31));
        if (old & (1u << (v[1] & 31))) cls = R_DUP;                      // the later row wins in the reference: not a layout kmerdb writes
    }
    if (cls) { atomicMin(P.err, (row << 3) | cls); return; }
    P.kmer_ids[row] = v[1];
    P.counts[v[1]] = v[2];
}

__global__ void __launch_bounds__(kReadThreads) k10_parse_rows(const ReadParams P) {
    __shared__ unsigned int s_warp[kReadThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t base = (size_t)blockIdx.x * kReadBlock + (size_t)tid * kReadPerThread;
    unsigned long long m = base < P.n ? newline_mask(P.text, P.n, base) : 0ull;
    const unsigned int c = __popcll(m);
    unsigned int inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned int before = inc - c;
    for (int w = 0; w < warp; w++) before += s_warp[w];
    unsigned long long row = P.first_row + P.blk_off[blockIdx.x] + before;   // line feeds before my bytes = rows that started before them, minus row 0
    if (blockIdx.x == 0 && tid == 0) parse_row(P, 0, P.first_row);
    while (m) {
        const int i = __ffsll((long long)m) - 1;
        m &= m - 1;
        row++;
        parse_row(P, base + i + 1, row);
    }
}

}  // namespace kdb
