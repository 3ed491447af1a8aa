// K3p — bucket-partitioned histogram for 8 <= k <= 14, sm_100a.
//
// Same job as K3 (kmerdb/kmer.py:297-317 ids, kmerdb/parse.py:133-136 counts) but built to get past
// the scattered-RED ceiling (1.9e11 updates/s): shared-memory atomics are ~11x faster
// than global ones, so ids are first partitioned by their high bits and each bucket is then counted
// in a 64 KiB shared-memory table.
//   k3p_partition : one CTA per packed tile.  ids -> tile-local counting sort by bucket
//                   (rank = shared atomicAdd on a bucket histogram, CTA scan, scatter into a sorted
//                   shared array), one global cursor bump per non-empty bucket, then the sorted
//                   array leaves as 14-bit payloads in runs of consecutive 2-byte stores.
//   k3p_count     : work item = (bucket, part): zero a 16384-bin shared table, stream the bucket's
//                   payloads with 16-byte loads, ATOMS each, flush the table into tab32 with RED.
// A bucket stream that would overflow its capacity falls back to direct REDs, so the result is exact
// whatever the id distribution.
#pragma once
#include "kdb_common.cuh"
#include "kdb_encode.cuh"
#include "kdb_hist.cuh"

#ifndef KDB_PB14
#define KDB_PB14 17      // payload bits at k = 14 (measured, 10 M reads: 17 -> partition 5.93 + count 3.24 ms; 18 -> 3.93 + 5.71 ms)
#endif

namespace kdb {


// payload bits of a stream entry: 14 for k <= 11 and 15 for k = 12 (one 16384- / 32768-bin table per bucket, 16-bit entries: 512
// buckets at k = 12 sort faster than 1024 — fewer cursors to scan and bump per tile, longer runs); 16 for k = 13 (65536 packed
// 16-bit counters per bucket, k3p_count16); 17 for k = 14 (32-bit entries, 2048 buckets, two sub-passes of 16 bits: with 16-bit
// payloads k = 14 needs 4096 buckets, which leaves 4 entries per bucket per tile — 8-byte runs and one cursor bump per 4 keys)
__host__ __device__ constexpr int payload_bits(int k) { return k <= 11 ? 14 : (k == 12 ? 15 : (k == 13 ? 16 : KDB_PB14)); }
// k3s_partition (below): the bucket of an id is its bits [kBucketShift, kBucketShift + nbits), the payload the bits above and below
// it squeezed together; this puts them back
constexpr int kBucketShift = 8;
__host__ __device__ constexpr uint32_t slotted_id(uint32_t payload, uint32_t bucket, int nbits) {
    return ((payload >> kBucketShift) << (kBucketShift + nbits)) | (bucket << kBucketShift) | (payload & ((1u << kBucketShift) - 1u));
}
constexpr int kPartThreads = 512;                      // 512 threads x 32 letters = one full tile
constexpr int kBucketThreads = 256;

struct PartParams {
    const uint32_t* packed;
    unsigned int n_tiles;
    int skip_n;
    int smear;                   // k - (k the packed stream was encoded for), kdb_hist.cuh smear_w
    unsigned int* tab32;
    uint16_t* stream;            // n_buckets * cap entries (32-bit entries when payload_bits > 16)
    unsigned int* cursor;        // n_buckets: entries claimed so far
    unsigned int* fail_base;     // n_buckets: lowest base of a run that did not fit (0xFFFFFFFF = none)
    unsigned int cap;            // entries per bucket (multiple of 8)
    Accum* acc;
};

__device__ __forceinline__ uint32_t shr96_rt(uint32_t w2, uint32_t w1, uint32_t w0, int s) {
    return s < 32 ? __funnelshift_r(w0, w1, s) : (s < 64 ? __funnelshift_r(w1, w2, s - 32) : (w2 >> (s - 64)));
}

// Direct update of the u32 table for the entries the sort cannot take (a bucket holding more entries in one pass than a
// rank has bits for: low-complexity sequence).  Equal ids are merged across the warp first — such input repeats the same
// few k-mers, and same-address REDs serialise in L2 (1.5e9/s measured on poly-A before this).
__device__ __forceinline__ void red_aggregated(unsigned int* tab32, uint32_t id) {
    const unsigned int peers = __match_any_sync(__activemask(), id);
    if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(tab32 + id, (unsigned int)__popc(peers));
}

// Two sweeps over the 32 letters of a thread: sweep A only counts (bucket histogram), the CTA scans the histogram into
// per-bucket cursors, sweep B recomputes each id and takes its place from the cursor.  Nothing but the packed words is
// kept in registers between the sweeps, so three CTAs fit one SM (the register array of a one-sweep rank sort allows two),
// and a crowded bucket needs no special case (no rank is packed beside the id).
template <int K, bool CANON>
__global__ void __launch_bounds__(kPartThreads, 3) k3p_partition(const PartParams P) {
    constexpr int PB = payload_bits(K);
    constexpr int NBITS = 2 * K - PB, NB = 1 << NBITS;              // buckets
    constexpr uint32_t mask = (1u << (2 * K)) - 1u;
    constexpr int PER = (NB + kPartThreads - 1) / kPartThreads;     // buckets per thread in the scan
    extern __shared__ uint32_t s_dyn_part[];                        // (kTile + 2 NB + 32) words, see k3p_smem_bytes
    uint32_t* const s_sorted = s_dyn_part;                          // ids of the tile, grouped by bucket
    uint32_t* const s_hist = s_dyn_part + kTile;                    // sweep A: entries per bucket; sweep B: next free sorted slot
    uint32_t* const s_delta = s_hist + NB + 32;                     // stream index of sorted entry j = s_delta[bucket] + j (32 dummy counters sit between)
    __shared__ uint32_t s_warp[kPartThreads / 32];
    __shared__ unsigned long long s_red[kPartThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long kmers = 0;

    // Work unit = two consecutive packed tiles.  Tiles that are at most half full (FASTQ: half of the text is quality
    // lines) share one sort pass — tile A on threads 0..255, tile B on threads 256..511 — so the per-pass costs (bucket
    // histogram, scan, cursor bumps) and the stream runs do not degrade with sparse tiles.
    const unsigned int n_units = (P.n_tiles + 1) / 2;
    for (unsigned int unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
      const unsigned int tA = 2 * unit, tB = tA + 1;
      const uint32_t nA = P.packed[(size_t)tA * kPackStride], nB = tB < P.n_tiles ? P.packed[(size_t)tB * kPackStride] : 0u;
      const bool both = nA <= kTile / 2 && nB <= kTile / 2;
      const int n_half = (both || tB >= P.n_tiles) ? 1 : 2;
      for (int half = 0; half < n_half; half++) {
        const bool onB = both ? tid >= kPartThreads / 2 : half == 1;
        const int ltid = both ? (tid & (kPartThreads / 2 - 1)) : tid;
        const uint32_t* reg = P.packed + (size_t)(onB ? tB : tA) * kPackStride;
        const uint32_t n = onB ? nB : nA;
        __syncthreads();
        for (int i = tid; i < NB; i += kPartThreads) s_hist[i] = 0;
        __syncthreads();

        // ---- sweep A: bucket histogram of my 32 letters -------------------------------------------
        const bool active = (uint32_t)ltid * 32 < n;
        uint32_t cur0 = 0, cur1 = 0, prev = 0, wb = 0;
        if (active) {
            const uint32_t* codes = reg + kPackHdr;
            cur0 = codes[2 * ltid];
            cur1 = (uint32_t)ltid * 32 + 16 < n ? codes[2 * ltid + 1] : 0u;
            prev = ltid ? codes[2 * ltid - 1] : reg[1];
            wb = reg[kPackHdr + kPackCodes + ltid];
            if (P.smear) wb = smear_w(ltid ? reg[kPackHdr + kPackCodes + ltid - 1] : reg[3], wb, P.smear);
            if (!P.skip_n) {
                const uint32_t* nnw = reg + kPackHdr + kPackCodes + kPackBits;
                const uint32_t nn_cur = nnw[ltid], nn_prev = ltid ? nnw[ltid - 1] : reg[2];
                if ((nn_cur | (nn_prev & ((1u << (K - 1)) - 1u))) != 0) {
                    // expansion mode with N near by: generic path, direct REDs; nothing of this thread enters the sort
                    const unsigned long long S0 = ((unsigned long long)prev << 32) | cur0, S1 = ((unsigned long long)cur0 << 32) | cur1;
                    const unsigned long long NS = ((unsigned long long)nn_prev << 32) | nn_cur;
                    for (int e = 0; e < 32; e++) {
                        if (!(wb & (0x80000000u >> e))) continue;
                        const uint32_t f = (uint32_t)((e < 16 ? S0 : S1) >> (30 - 2 * (e & 15))) & mask;
                        const uint32_t nm = (uint32_t)(NS >> (31 - e)) & ((1u << K) - 1u);
                        if (nm) kmers += expand_window(P.tab32, K, CANON, f, nm);
                        else {
                            uint32_t id = f;
                            if (CANON) { const uint32_t r = revcomp32(f, K); id = r < f ? r : f; }
                            atomicAdd(P.tab32 + id, 1u);
                            kmers++;
                        }
                    }
                    wb = 0;
                }
            }
        }
        // Ids stay top-aligned (the 2K id bits on top, later letters below them): min() of the two strands is decided by
        // the id bits alone, and bucket / payload / id come out by shifts without a mask.
        const uint32_t r2 = CANON ? rc_word(cur1) : 0u, r1 = CANON ? rc_word(cur0) : 0u, r0 = CANON ? rc_word(prev) : 0u;
        auto window_top = [&](int e) -> uint32_t {                     // window ending at my letter e
            uint32_t m = ext96(prev, cur0, cur1, 2 * (17 + e - K));
            if (CANON) { const uint32_t rc = ext96(r2, r1, r0, 62 - 2 * e); m = rc < m ? rc : m; }
            return m;
        };
        // A letter without a window bumps a per-lane dummy counter behind the histogram instead of branching around the
        // atomic: the compiler wraps every conditional shared atomic in its own branch + reconvergence barrier (three
        // extra instructions per entry, and no overlap between entries); 93 % of the letters of a read have a window.
        uint32_t* const dummy = s_hist + NB + lane;
        if (wb) {
#pragma unroll
            for (int e = 0; e < 32; e++) {
                const bool valid = (wb & (0x80000000u >> e)) != 0;
                uint32_t b = window_top(e) >> (32 - NBITS);
                asm volatile("" : "+r"(b));                              // (keeps the id arithmetic out of a branch the compiler would build for the select)
                atomicAdd(valid ? s_hist + b : dummy, 1u);
            }
            kmers += __popc(wb);
        }
        __syncthreads();

        // ---- scan: bucket counts -> cursors; claim room in the global streams ------------------------
        uint32_t cnt[PER], sum = 0;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            const int b = tid * PER + j;
            cnt[j] = b < NB ? s_hist[b] : 0u;
            sum += cnt[j];
        }
        uint32_t inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        uint32_t base = inc - sum, total = 0;
#pragma unroll
        for (int w = 0; w < kPartThreads / 32; w++) { const uint32_t c = s_warp[w]; if (w < warp) base += c; total += c; }
#pragma unroll
        for (int j = 0; j < PER; j++) {
            const int b = tid * PER + j;
            if (b < NB) {
                s_hist[b] = base;
                uint32_t dl = 0xFFFFFFFFu;
                if (cnt[j]) {
                    const uint32_t gb = atomicAdd(P.cursor + b, cnt[j]);
                    if (gb + cnt[j] > P.cap) atomicMin(P.fail_base + b, gb);                          // does not fit: RED path
                    else dl = (uint32_t)b * P.cap + gb - base;                                        // (host keeps n_buckets * cap < 2^32)
                }
                s_delta[b] = dl;
                base += cnt[j];
            }
        }
        __syncthreads();

        // ---- sweep B: every id takes the next slot of its bucket ------------------------------------------
        if (wb) {
#pragma unroll
            for (int e = 0; e < 32; e++) {
                const bool valid = (wb & (0x80000000u >> e)) != 0;
                const uint32_t m = window_top(e);
                const uint32_t pos = atomicAdd(valid ? s_hist + (m >> (32 - NBITS)) : dummy, 1u);
                if (valid) s_sorted[pos] = m;
            }
        }
        __syncthreads();

        // ---- the sorted array leaves as payload runs, one entry per thread ----------------------
#pragma unroll 4
        for (uint32_t j = tid; j < total; j += kPartThreads) {
            const uint32_t m = s_sorted[j];
            const uint32_t dl = s_delta[m >> (32 - NBITS)];
            if (dl != 0xFFFFFFFFu) {
                if (PB > 16) reinterpret_cast<uint32_t*>(P.stream)[dl + j] = (m >> (32 - 2 * K)) & ((1u << PB) - 1u);
                else P.stream[dl + j] = (uint16_t)((m >> (32 - 2 * K)) & ((1u << PB) - 1u));
            }
            else red_aggregated(P.tab32, m >> (32 - 2 * K));         // stream of this bucket is full (low-complexity input): equal ids sit side by side
        }
      }
    }
    for (int d = 16; d; d >>= 1) kmers += __shfl_xor_sync(0xffffffffu, kmers, d);
    __syncthreads();
    if (lane == 0) s_red[warp] = kmers;
    __syncthreads();
    if (tid == 0) {
        unsigned long long s = 0;
        for (int w = 0; w < kPartThreads / 32; w++) s += s_red[w];
        if (s) atomicAdd(&P.acc->total_kmers, s);
    }
}

// One work item = (bucket, part): grid = n_buckets * parts.  The payloads of one bucket are counted into a shared table of
// 2^TB bins (TB = payload bits: 14 -> 64 KiB, three CTAs of 256 threads per SM; 15 -> 128 KiB, one CTA of 1024 threads) and the
// table is flushed with coalesced REDs.
template <int TB>
__global__ void __launch_bounds__(TB == 14 ? kBucketThreads : 1024) k3p_count(const uint16_t* __restrict__ stream, const unsigned int* __restrict__ cursor,
                                                                              const unsigned int* __restrict__ fail_base, unsigned int cap, int parts,
                                                                              unsigned int* __restrict__ tab32, int low_bits) {   // low_bits != 0: the streams come from k3s_partition (bucket = id bits [kBucketShift, kBucketShift + low_bits))
    extern __shared__ unsigned int s_tab[];
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t b = blockIdx.x / parts, part = blockIdx.x % parts;
    uint32_t end = cursor[b];
    { const uint32_t fb = fail_base[b]; end = end < fb ? end : fb; end = end < cap ? end : cap; }
    if (end == 0) return;
    // split [0, end) into `parts` ranges on 8-entry boundaries
    const uint32_t units = (end + 7) >> 3, per = (units + parts - 1) / parts;
    const uint32_t u0 = part * per, u1 = min(units, u0 + per);
    if (u0 >= u1) return;
    for (uint32_t i = tid; i < (1u << TB); i += nthr) s_tab[i] = 0;
    __syncthreads();
    const uint4* src = reinterpret_cast<const uint4*>(stream + (size_t)b * cap);
    // four 16-byte loads in flight per thread (the kernel sat in long-scoreboard stalls with one), and every atomic unconditional:
    // 0xFFFF — the padding of k3s_partition's runs, and what entries past the end of the stream are turned into — lands in a per-lane
    // dummy bin behind the table instead of being branched around
    const uint32_t dummy = (1u << TB) + (tid & 31u);
    for (uint32_t ub = u0; ub < u1; ub += 4 * nthr) {
        uint4 v[4];
        uint32_t valid[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t u = ub + r * nthr + tid;
            valid[r] = u < u1 ? min(8u, end - 8 * u) : 0u;
            v[r] = valid[r] ? __ldcs(src + u) : make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const uint32_t e = (w[q >> 1] >> (16 * (q & 1))) & 0xFFFFu;
                atomicAdd(&s_tab[(uint32_t)q < valid[r] ? min(e, dummy) : dummy], 1u);
            }
        }
    }
    __syncthreads();
    if (low_bits) {                                                   // bin p of bucket b (k3s_partition): runs of 2^kBucketShift consecutive ids
        for (uint32_t i = tid; i < (1u << TB); i += nthr) { const unsigned int x = s_tab[i]; if (x) atomicAdd(tab32 + slotted_id(i, b, low_bits), x); }
        return;
    }
    unsigned int* dst = tab32 + ((size_t)b << TB);
    for (uint32_t i = tid; i < (1u << TB); i += nthr) { const unsigned int x = s_tab[i]; if (x) atomicAdd(dst + i, x); }
}

// 16-bit payloads (k = 13, 14): the 65536 bins of a bucket as packed 16-bit counters (add16, kdb_hist.cuh), counted in ONE pass.
constexpr int kCount16Threads = 1024;
// grid = n_buckets; one CTA counts one bucket stream
__global__ void __launch_bounds__(kCount16Threads) k3p_count16(const uint16_t* __restrict__ stream, const unsigned int* __restrict__ cursor,
                                                               const unsigned int* __restrict__ fail_base, unsigned int cap,
                                                               unsigned int* __restrict__ tab32) {
    extern __shared__ unsigned int s_tab[];                           // 32768 words
    const int tid = threadIdx.x;
    const uint32_t b = blockIdx.x;
    uint32_t end = cursor[b];
    { const uint32_t fb = fail_base[b]; end = end < fb ? end : fb; end = end < cap ? end : cap; }
    if (end == 0) return;
    for (int i = tid; i < 32768; i += kCount16Threads) s_tab[i] = 0;
    __syncthreads();
    unsigned int* dst = tab32 + ((size_t)b << 16);
    const uint4* src = reinterpret_cast<const uint4*>(stream + (size_t)b * cap);
    const uint32_t units = (end + 7) >> 3;
    for (uint32_t u = tid; u < units; u += kCount16Threads) {
        const uint4 v = __ldcs(src + u);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        const uint32_t valid = min(8u, end - 8 * u);
#pragma unroll
        for (int q = 0; q < 8; q++)
            if ((uint32_t)q < valid) add16(s_tab, (w[q >> 1] >> (16 * (q & 1))) & 0xFFFFu, dst);
    }
    __syncthreads();
    for (int i = tid; i < 32768; i += kCount16Threads) {
        const unsigned int x = s_tab[i];
        if (x & 0xFFFFu) atomicAdd(dst + 2 * i, x & 0xFFFFu);
        if (x >> 16) atomicAdd(dst + 2 * i + 1, x >> 16);
    }
}

// pb-bit payloads (pb = KDB_PB14 = 17) in 32-bit entries (k = 14): grid = n_buckets * 2^(pb - 16); CTA (b, sub) counts the entries of
// bucket b whose payload bits above 16 equal sub (the 2^(pb - 16) CTAs of a bucket run next to each other: the re-reads hit L2)
__global__ void __launch_bounds__(kCount16Threads) k3p_count16w(const uint32_t* __restrict__ stream, const unsigned int* __restrict__ cursor,
                                                                const unsigned int* __restrict__ fail_base, unsigned int cap, int pb,
                                                                unsigned int* __restrict__ tab32) {
    extern __shared__ unsigned int s_tab[];                           // 32768 words
    const int tid = threadIdx.x;
    const uint32_t subs = 1u << (pb - 16);
    const uint32_t b = blockIdx.x / subs, sub = blockIdx.x % subs;
    uint32_t end = cursor[b];
    { const uint32_t fb = fail_base[b]; end = end < fb ? end : fb; end = end < cap ? end : cap; }
    if (end == 0) return;
    for (int i = tid; i < 32768; i += kCount16Threads) s_tab[i] = 0;
    __syncthreads();
    unsigned int* dst = tab32 + ((size_t)b << pb) + ((size_t)sub << 16);
    const uint4* src = reinterpret_cast<const uint4*>(stream + (size_t)b * cap);
    const uint32_t units = (end + 3) >> 2;
    // four 16-byte loads in flight per thread: one CTA per SM has to cover the L2 latency on its own
    for (uint32_t u0 = 0; u0 < units; u0 += 4 * kCount16Threads) {
        uint4 v[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t u = u0 + r * kCount16Threads + tid;
            v[r] = u < units ? __ldg(src + u) : make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);    // (payloads have pb < 32 bits: no sub matches)
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t u = u0 + r * kCount16Threads + tid;
            const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
            const uint32_t valid = u < units ? min(4u, end - 4 * u) : 0u;
#pragma unroll
            for (int q = 0; q < 4; q++)
                if ((uint32_t)q < valid && (w[q] >> 16) == sub) add16(s_tab, w[q] & 0xFFFFu, dst);
        }
    }
    __syncthreads();
    for (int i = tid; i < 32768; i += kCount16Threads) {
        const unsigned int x = s_tab[i];
        if (x & 0xFFFFu) atomicAdd(dst + 2 * i, x & 0xFFFFu);
        if (x >> 16) atomicAdd(dst + 2 * i + 1, x >> 16);
    }
}

// ---- single-sweep slotted partition (r02) ----------------------------------------------------------------------
// One returning shared atomic + one predicated 2-byte store per key (the two-sweep sort above pays a histogram atomic, a cursor
// atomic and a scatter store, and builds every id twice).  Every bucket owns a fixed range of C slots in shared memory;
// pos = atomicAdd(&cnt[b], 1) is the slot, so there is no scan and no second pass over the letters.
// Which bits make the bucket: min(fwd, rc) is decided by the first letters of the two strands, so both ends of a canonical id
// are skewed (the top NBITS bits 2:0 over the range, the lowest ones with a std of 0.58 of the mean); from bit 6 to bit
// 2K - NBITS - 6 the bucket loads of uniform sequence are flat (max / mean <= 1.05, measured for k = 9..12).  The bucket is id
// bits [kBucketShift, kBucketShift + NBITS), the payload the bits above and below it squeezed together (slotted_id puts them
// back), so one slot count fits every bucket and the slot address is b * C + pos with C an immediate: C = 32768 / NB - 4, nearly twice
// the load of a tile of dense text (+4.9 sigma at 512 buckets).
// A bucket that gets more than C keys in one unit (low-complexity sequence) is not slotted at all in that unit: a second sweep
// (only run when some bucket overflowed) sends every key of such a bucket to the global table directly (warp-aggregated REDs),
// so the result is exact for any input and the common case carries no per-key overflow bookkeeping.
// Runs leave padded to 4 entries with 0xFFFF (no bin: k3p_count skips it) as aligned 8-byte stores; the slots are 0xFFFF
// whenever a sweep begins (filled at kernel start, restored by the copy-out), so the padding costs nothing.
// (C / 2 = -2 mod 32: the bucket bases spread over 16 banks.  With C = 64 every bucket starts in bank 0 and, since all buckets
// fill at the same rate, the stores of a warp pile onto the bank of the current fill level: 105 M conflicts per 324 MB instead of 48 M.)
__host__ __device__ constexpr int slot_count(int nb) { return 32768 / nb - 4; }
__host__ inline size_t k3s_smem_bytes(int k) {
    const int nb = 1 << (2 * k - payload_bits(k));
    return (size_t)(2 * nb + 32) * sizeof(uint32_t) + (size_t)nb * slot_count(nb) * sizeof(uint16_t);
}

template <int K, bool CANON>
__global__ void __launch_bounds__(kPartThreads, 3) k3s_partition(const PartParams P) {
    constexpr int PB = payload_bits(K) <= 16 ? payload_bits(K) : 16;   // (instantiated, never launched, for k = 14)
    constexpr int NBITS = 2 * K - PB, NB = 1 << NBITS;
    constexpr int C = slot_count(NB), Q = C / 4;
    constexpr int S = 32 - 2 * K;                                   // the id sits in the top 2K bits of a window word
    constexpr uint32_t mask = (1u << (2 * K)) - 1u;
    extern __shared__ uint32_t s_dyn_part[];                        // NB counters, 32 dummies, NB global bases, NB * C 16-bit slots
    uint32_t* const s_cnt = s_dyn_part;
    uint32_t* const s_gb = s_dyn_part + NB + 32;
    uint16_t* const s_slot = reinterpret_cast<uint16_t*>(s_gb + NB);
    unsigned long long* const slot64 = reinterpret_cast<unsigned long long*>(s_slot);
    __shared__ unsigned long long s_red[kPartThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long kmers = 0;
    for (int i = tid; i < NB * Q; i += kPartThreads) slot64[i] = ~0ull;
    if (tid < 32) s_cnt[NB + tid] = 0x80000000u;                    // per-lane dummy counters: never below C, so what they return never "fits"

    const unsigned int n_units = (P.n_tiles + 1) / 2;               // work unit = two packed tiles (see k3p_partition)
    for (unsigned int unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
      const unsigned int tA = 2 * unit, tB = tA + 1;
      const uint32_t nA = P.packed[(size_t)tA * kPackStride], nB = tB < P.n_tiles ? P.packed[(size_t)tB * kPackStride] : 0u;
      const bool both = nA <= kTile / 2 && nB <= kTile / 2;
      const int n_half = (both || tB >= P.n_tiles) ? 1 : 2;
      for (int half = 0; half < n_half; half++) {
        const bool onB = both ? tid >= kPartThreads / 2 : half == 1;
        const int ltid = both ? (tid & (kPartThreads / 2 - 1)) : tid;
        const uint32_t* reg = P.packed + (size_t)(onB ? tB : tA) * kPackStride;
        const uint32_t n = onB ? nB : nA;
        __syncthreads();                                            // the copy-out of the unit before is done with s_cnt / s_slot
        for (int i = tid; i < NB; i += kPartThreads) s_cnt[i] = 0;
        __syncthreads();

        const bool active = (uint32_t)ltid * 32 < n;
        uint32_t cur0 = 0, cur1 = 0, prev = 0, wb = 0;
        if (active) {
            const uint32_t* codes = reg + kPackHdr;
            cur0 = codes[2 * ltid];
            cur1 = (uint32_t)ltid * 32 + 16 < n ? codes[2 * ltid + 1] : 0u;
            prev = ltid ? codes[2 * ltid - 1] : reg[1];
            wb = reg[kPackHdr + kPackCodes + ltid];
            if (P.smear) wb = smear_w(ltid ? reg[kPackHdr + kPackCodes + ltid - 1] : reg[3], wb, P.smear);
            if (!P.skip_n) {
                const uint32_t* nnw = reg + kPackHdr + kPackCodes + kPackBits;
                const uint32_t nn_cur = nnw[ltid], nn_prev = ltid ? nnw[ltid - 1] : reg[2];
                if ((nn_cur | (nn_prev & ((1u << (K - 1)) - 1u))) != 0) {
                    // expansion mode with N near by: generic path, direct REDs; nothing of this thread enters the slots
                    const unsigned long long S0 = ((unsigned long long)prev << 32) | cur0, S1 = ((unsigned long long)cur0 << 32) | cur1;
                    const unsigned long long NS = ((unsigned long long)nn_prev << 32) | nn_cur;
                    for (int e = 0; e < 32; e++) {
                        if (!(wb & (0x80000000u >> e))) continue;
                        const uint32_t f = (uint32_t)((e < 16 ? S0 : S1) >> (30 - 2 * (e & 15))) & mask;
                        const uint32_t nm = (uint32_t)(NS >> (31 - e)) & ((1u << K) - 1u);
                        if (nm) kmers += expand_window(P.tab32, K, CANON, f, nm);
                        else {
                            uint32_t id = f;
                            if (CANON) { const uint32_t r = revcomp32(f, K); id = r < f ? r : f; }
                            atomicAdd(P.tab32 + id, 1u);
                            kmers++;
                        }
                    }
                    wb = 0;
                }
            }
        }
        const uint32_t r2 = CANON ? rc_word(cur1) : 0u, r1 = CANON ? rc_word(cur0) : 0u, r0 = CANON ? rc_word(prev) : 0u;
        auto window_top = [&](int e) -> uint32_t {                     // window ending at my letter e, id bits on top
            uint32_t m = ext96(prev, cur0, cur1, 2 * (17 + e - K));
            if (CANON) { const uint32_t rc = ext96(r2, r1, r0, 62 - 2 * e); m = rc < m ? rc : m; }
            return m;
        };
        uint32_t* const dummy = s_cnt + NB + lane;                     // a letter without a window bumps a per-lane dummy (no branch)
        if (wb) {
#pragma unroll
            for (int e = 0; e < 32; e++) {
                const bool valid = (wb & (0x80000000u >> e)) != 0;
                const uint32_t m = window_top(e);
                uint32_t b = (m >> (S + kBucketShift)) & (uint32_t)(NB - 1);
                asm volatile("" : "+r"(b));                              // (keeps the id arithmetic out of a branch the compiler would build for the select)
                const uint32_t pos = atomicAdd(valid ? s_cnt + b : dummy, 1u);
                const uint32_t pay = ((m >> (S + NBITS)) & ~((1u << kBucketShift) - 1u)) | ((m >> S) & ((1u << kBucketShift) - 1u));
                if (pos < (uint32_t)C) s_slot[b * C + pos] = (uint16_t)pay;
            }
            kmers += __popc(wb);
        }
        __syncthreads();

        // ---- one global cursor bump per non-empty bucket: runs are padded to 4 entries so that they leave as 8-byte stores ----
        bool over = false;
        for (int b = tid; b < NB; b += kPartThreads) {
            const uint32_t c = s_cnt[b], c4 = (c + 3u) & ~3u;
            uint32_t gb = 0xFFFFFFFFu;                                  // nothing to copy: empty, overflowed (second sweep), or no room in the stream
            if (c > (uint32_t)C) over = true;
            else if (c) {
                const uint32_t g = atomicAdd(P.cursor + b, c4);
                if (g + c4 > P.cap) { atomicMin(P.fail_base + b, g); over = true; s_cnt[b] = C + 1; }   // no room: treat like an overflow
                else gb = ((uint32_t)b * P.cap + g) >> 2;               // (host keeps n_buckets * cap < 2^32; cap is a multiple of 8)
            }
            s_gb[b] = gb;
        }
        const int any_over = __syncthreads_or(over);

        // ---- the slots leave as payload runs, four entries per thread, and are reset to 0xFFFF ------------------------------
#pragma unroll 4
        for (int i = 0; i < (NB * Q + kPartThreads - 1) / kPartThreads; i++) {
            const uint32_t j = (uint32_t)tid + (uint32_t)i * kPartThreads, b = min(j / (uint32_t)Q, (uint32_t)(NB - 1)), r = 4u * (j % (uint32_t)Q);
            const uint32_t c = j < (uint32_t)(NB * Q) ? s_cnt[b] : 0u;
            if (r < c) {
                const unsigned long long v = slot64[j];
                slot64[j] = ~0ull;
                const uint32_t gb = s_gb[b];
                if (gb != 0xFFFFFFFFu) reinterpret_cast<unsigned long long*>(P.stream)[gb + (r >> 2)] = v;
            }
        }
        if (any_over) {
            // ---- second sweep (rare): every key of an overflowed bucket goes straight to the table ---------------------------
            if (wb) {
                const unsigned long long S0 = ((unsigned long long)prev << 32) | cur0, S1 = ((unsigned long long)cur0 << 32) | cur1;
                for (int e = 0; e < 32; e++) {
                    if (!(wb & (0x80000000u >> e))) continue;
                    const uint32_t f = (uint32_t)((e < 16 ? S0 : S1) >> (30 - 2 * (e & 15))) & mask;
                    uint32_t id = f;
                    if (CANON) { const uint32_t r = revcomp32(f, K); id = r < f ? r : f; }
                    if (s_cnt[(id >> kBucketShift) & (uint32_t)(NB - 1)] > (uint32_t)C) red_aggregated(P.tab32, id);
                }
            }
        }
      }
    }
    for (int d = 16; d; d >>= 1) kmers += __shfl_xor_sync(0xffffffffu, kmers, d);
    __syncthreads();
    if (lane == 0) s_red[warp] = kmers;
    __syncthreads();
    if (tid == 0) {
        unsigned long long s = 0;
        for (int w = 0; w < kPartThreads / 32; w++) s += s_red[w];
        if (s) atomicAdd(&P.acc->total_kmers, s);
    }
}

__host__ inline size_t k3p_smem_bytes(int k) { return (size_t)(kTile + 2 * (1 << (2 * k - payload_bits(k))) + 32) * sizeof(uint32_t); }

typedef void (*k3p_fn)(const PartParams);
template <int K>
__host__ inline k3p_fn k3p_select(bool canon) { return canon ? (k3p_fn)k3p_partition<K, true> : (k3p_fn)k3p_partition<K, false>; }
template <int K>
__host__ inline k3p_fn k3s_select(bool canon) { return canon ? (k3p_fn)k3s_partition<K, true> : (k3p_fn)k3s_partition<K, false>; }
__host__ inline k3p_fn k3s_kernel(int k, bool canon) {              // slotted single-sweep partition: 16-bit stream entries only
    switch (k) {
        case 9: return k3s_select<9>(canon);   case 10: return k3s_select<10>(canon);
        case 11: return k3s_select<11>(canon); case 12: return k3s_select<12>(canon);
    }
    return nullptr;
}
__host__ inline k3p_fn k3p_kernel(int k, bool canon) {
    switch (k) {
        case 8: return k3p_select<8>(canon);   case 9: return k3p_select<9>(canon);
        case 10: return k3p_select<10>(canon); case 11: return k3p_select<11>(canon);
        case 12: return k3p_select<12>(canon); case 13: return k3p_select<13>(canon);
        case 14: return k3p_select<14>(canon);
    }
    return nullptr;
}

}  // namespace kdb
