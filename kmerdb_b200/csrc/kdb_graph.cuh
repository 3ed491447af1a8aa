// K4 — graph edge rows, aggregated edge weights and minimizers, sm_100a.
//
//   k8_check_letters / k8_edge_rows : graph.make_edges_from_kmerids rows (kmerdb/graph.py:108-216):
//       for every record the ordered pairs (pos i-1, id[i-1], pos i, id[i]), i = 1 .. len-k.
//       Input is the host-tokenised record list (letters back to back + record offsets), exactly what
//       graph.make_edges_from_fasta hands to kmer.shred (graph.py:261-274).
//   k7_edge_weight_*                : the aggregated (n1, n2, weight) list (legacy .kdbg columns)
//       from the dense histogram of forward (k+1)-mers built by the profile kernels.
//   k8_min_compat                   : minimizer.select_lexicographical_minimizers rows
//       (kmerdb/minimizer.py:65-81): (i, canonical id, i % W == 0).
//   k8_min_window / k8_min_*        : true windowed minimizers (extension): minimum canonical
//       id among w consecutive k-mers, leftmost on ties, by warp shuffles over a 64-wide register window.
#pragma once
#include "kdb_common.cuh"

namespace kdb {

constexpr unsigned long long kIdNone = ~0ull;          // kmer_to_id -> None (window holds N)

// error key of the record kernels: rec << 36 | class << 34 | pos << 1 | is_iupac   (min = first raise)
enum : unsigned long long { G_ALPHABET = 0, G_SHORT = 1, G_LETTER = 2 };
__device__ __forceinline__ unsigned long long gkey(unsigned long long rec, unsigned long long cls, unsigned long long pos, unsigned long long iupac) {
    return (rec << 36) | (cls << 34) | ((pos & 0x1FFFFFFFFull) << 1) | iupac;
}

// index of the last element of off[0..n] that is <= x   (off ascending, off[0] = 0)
__device__ __forceinline__ unsigned long long upper_rec(const unsigned long long* __restrict__ off, unsigned long long n, unsigned long long x) {
    unsigned long long lo = 0, hi = n;                  // invariant: off[lo] <= x < off[hi]
    while (hi - lo > 1) {
        const unsigned long long mid = (lo + hi) >> 1;
        if (off[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// forward and reverse-complement id of the k letters at s (all ACGT); canonical = min
// (`lut`: a shared-memory copy of the class table — the __constant__ table indexed by a data-dependent byte serialises the warp)
__device__ __forceinline__ void load_lut(uint8_t* s_lut) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_lut[i] = c_class_lut[i];
    __syncthreads();
}
__device__ __forceinline__ unsigned long long window_id(const uint8_t* __restrict__ s, int k, int canonical, const uint8_t* lut) {
    unsigned long long fwd = 0, rc = 0;
    for (int j = 0; j < k; j++) {
        const unsigned long long code = lut[s[j]] & 3u;
        fwd = (fwd << 2) | code;
        rc |= (3ull - code) << (2 * j);
    }
    return (canonical && rc < fwd) ? rc : fwd;
}

// graph mode (graph.py:261-284 -> kmer.shred): alphabet (NameError) < short (AttributeError) < first N
// (ValueError) or RYSWKMBDHV (KeyError) in letter order.  One thread per letter, then one per record.
__global__ void __launch_bounds__(256) k8_check_letters(const uint8_t* __restrict__ letters, unsigned long long n_letters,
                                                        const unsigned long long* __restrict__ rec_off, unsigned long long n_recs,
                                                        int k, unsigned long long* err_key) {
    __shared__ uint8_t s_lut[256];
    load_lut(s_lut);
    const unsigned long long g = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long key = ~0ull;
    if (g < n_letters) {
        const uint8_t cl = s_lut[letters[g]];
        if (cl > C_T) {
            const unsigned long long r = upper_rec(rec_off, n_recs, g);
            if (cl == C_N) key = gkey(r, G_LETTER, g - rec_off[r], 0);
            else if (cl == C_IUPAC) key = gkey(r, G_LETTER, g - rec_off[r], 1);
            else key = gkey(r, G_ALPHABET, 0, 0);
        }
    }
    if (g < n_recs && rec_off[g + 1] - rec_off[g] < (unsigned long long)k) { const unsigned long long k2 = gkey(g, G_SHORT, 0, 0); key = k2 < key ? k2 : key; }
    for (int d = 16; d; d >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, key, d); key = o < key ? o : key; }
    if ((threadIdx.x & 31) == 0 && key != ~0ull) atomicMin(err_key, key);
}

// One thread per edge.  edge_off[r] = number of edges of the records before r.
__global__ void __launch_bounds__(256) k8_edge_rows(const uint8_t* __restrict__ letters, const unsigned long long* __restrict__ rec_off,
                                                    const unsigned long long* __restrict__ edge_off, unsigned long long n_recs,
                                                    unsigned long long n_edges, int k, int canonical,
                                                    unsigned long long* __restrict__ rec, unsigned long long* __restrict__ pos1,
                                                    unsigned long long* __restrict__ id1, unsigned long long* __restrict__ pos2,
                                                    unsigned long long* __restrict__ id2) {
    __shared__ uint8_t s_lut[256];
    load_lut(s_lut);
    const unsigned long long e = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const unsigned long long r = upper_rec(edge_off, n_recs, e);
    const unsigned long long i = e - edge_off[r] + 1;                 // graph.py:150: i >= 1
    const uint8_t* s = letters + rec_off[r] + i - 1;
    rec[e] = r; pos1[e] = i - 1; pos2[e] = i;
    // both windows from one walk over the k + 1 letters: the second id is the first one rolled by a letter
    unsigned long long fwd = 0, rc = 0;
    for (int j = 0; j < k; j++) { const unsigned long long code = s_lut[s[j]] & 3u; fwd = (fwd << 2) | code; rc |= (3ull - code) << (2 * j); }
    id1[e] = (canonical && rc < fwd) ? rc : fwd;
    const unsigned long long c2 = s_lut[s[k]] & 3u, mask = k == 32 ? ~0ull : ((1ull << (2 * k)) - 1ull);
    fwd = ((fwd << 2) | c2) & mask;
    rc = (rc >> 2) | ((3ull - c2) << (2 * (k - 1)));
    id2[e] = (canonical && rc < fwd) ? rc : fwd;
}

// ---- aggregated edge weights from the forward (k+1)-mer histogram -------------------------------------
__device__ __forceinline__ unsigned long long revcomp64(unsigned long long f, int k) {
    unsigned long long x = __brevll(~f);
    x = ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);
    return x >> (64 - 2 * k);
}

// The ordered pair (n1, n2) = (id of the k-prefix, id of the k-suffix) of the (k+1)-mer x (graph.py:150-214
// pairs kmer_ids[i-1] -> kmer_ids[i]).  With canonical ids several (k+1)-mers fall on one pair: the pair is
// emitted by its smallest member, weight = sum over the members.  Returns true when x is that member.
__device__ __forceinline__ bool edge_group(const unsigned long long* __restrict__ tab, unsigned long long x, int k, int canonical,
                                           unsigned long long* n1, unsigned long long* n2, unsigned long long* w) {
    const unsigned long long mk = (1ull << (2 * k)) - 1ull, mk1 = mk >> 2;
    const unsigned long long p = x >> 2, s = x & mk;
    if (!canonical) { *n1 = p; *n2 = s; *w = tab[x]; return true; }
    const unsigned long long rp = revcomp64(p, k), rs = revcomp64(s, k);
    const unsigned long long a = rp < p ? rp : p, b = rs < s ? rs : s;
    const unsigned long long pc[2] = {a, revcomp64(a, k)}, sc[2] = {b, revcomp64(b, k)};
    unsigned long long cand[4];
    int nc = 0;
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++)
            if ((pc[i] & mk1) == (sc[j] >> 2)) {
                const unsigned long long y = (pc[i] << 2) | (sc[j] & 3ull);
                bool dup = false;
                for (int q = 0; q < nc; q++) dup |= cand[q] == y;
                if (!dup) cand[nc++] = y;
            }
    unsigned long long sum = 0;
    bool first = true;
    for (int q = 0; q < nc; q++) { const unsigned long long c = tab[cand[q]]; sum += c; first &= c == 0 || x <= cand[q]; }   // empty members never emit
    *n1 = a; *n2 = b; *w = sum;
    return first;
}

constexpr int kEdgeBlock = 2048;
__global__ void __launch_bounds__(256) k7_edge_weight_count(const unsigned long long* __restrict__ tab, unsigned long long nbins, int k, int canonical,
                                                            unsigned int* __restrict__ block_n) {
    __shared__ unsigned int s;
    if (threadIdx.x == 0) s = 0;
    __syncthreads();
    const unsigned long long base = (unsigned long long)blockIdx.x * kEdgeBlock;
    unsigned int c = 0;
    for (int j = threadIdx.x; j < kEdgeBlock; j += 256) {
        const unsigned long long x = base + j;
        if (x < nbins && tab[x] != 0) { unsigned long long a, b, w; c += edge_group(tab, x, k, canonical, &a, &b, &w) ? 1u : 0u; }
    }
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s, c);
    __syncthreads();
    if (threadIdx.x == 0) block_n[blockIdx.x] = s;
}

__global__ void __launch_bounds__(256) k7_edge_weight_write(const unsigned long long* __restrict__ tab, unsigned long long nbins, int k, int canonical,
                                                            const unsigned long long* __restrict__ block_off, unsigned long long* __restrict__ n1,
                                                            unsigned long long* __restrict__ n2, unsigned long long* __restrict__ wt, unsigned long long cap) {
    __shared__ unsigned int s_warp[8];
    const unsigned long long base = (unsigned long long)blockIdx.x * kEdgeBlock;
    unsigned long long off = block_off[blockIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int j0 = 0; j0 < kEdgeBlock; j0 += 256) {
        const unsigned long long x = base + j0 + threadIdx.x;
        unsigned long long a = 0, b = 0, w = 0;
        const bool z = x < nbins && tab[x] != 0 && edge_group(tab, x, k, canonical, &a, &b, &w);
        const unsigned int bal = __ballot_sync(0xffffffffu, z);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        unsigned int before = 0, total = 0;
        for (int q = 0; q < 8; q++) { const unsigned int v = s_warp[q]; if (q < warp) before += v; total += v; }
        if (z) {
            const unsigned long long pos = off + before + __popc(bal & ((1u << lane) - 1u));
            if (pos < cap) { n1[pos] = a; n2[pos] = b; wt[pos] = w; }
        }
        off += total;
        __syncthreads();
    }
}

// ---- minimizers ---------------------------------------------------------------------------------------
// id of the window at s: kIdNone when it holds N (kmer.py:287-289); *bad is raised for other non-ACGT letters
__device__ __forceinline__ unsigned long long window_id_checked(const uint8_t* __restrict__ s, int k, unsigned int* bad, const uint8_t* lut) {
    unsigned long long fwd = 0, rc = 0;
    bool has_n = false;
    for (int j = 0; j < k; j++) {
        const uint8_t cl = lut[s[j]];
        if (cl == C_N) has_n = true;
        else if (cl == C_IUPAC) *bad |= 2u;
        else if (cl > C_T) *bad |= 1u;
        const unsigned long long code = cl & 3u;
        fwd = (fwd << 2) | code;
        rc |= (3ull - code) << (2 * j);
    }
    return has_n ? kIdNone : (rc < fwd ? rc : fwd);
}

// minimizer.py:65-81: row i = (canonical id of seq[i:i+k], i % W == 0), i < n_rows
__global__ void __launch_bounds__(256) k8_min_compat(const uint8_t* __restrict__ seq, unsigned long long n_rows, int k, unsigned long long window,
                                                     unsigned long long* __restrict__ ids, uint8_t* __restrict__ is_min, unsigned int* bad_flag) {
    __shared__ uint8_t s_lut[256];
    load_lut(s_lut);
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int bad = 0;
    if (i < n_rows) {
        ids[i] = window_id_checked(seq + i, k, &bad, s_lut);
        is_min[i] = (uint8_t)(i % window == 0);
    }
    if (bad) atomicOr(bad_flag, bad);
}

// canonical ids of all windows of one sequence (true minimizers, step 1)
__global__ void __launch_bounds__(256) k8_window_ids(const uint8_t* __restrict__ seq, unsigned long long n_win, int k,
                                                     unsigned long long* __restrict__ ids, unsigned int* bad_flag) {
    __shared__ uint8_t s_lut[256];
    load_lut(s_lut);
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int bad = 0;
    if (i < n_win) { const unsigned long long v = window_id_checked(seq + i, k, &bad, s_lut); if (v == kIdNone) bad |= 2u; ids[i] = v; }
    if (bad) atomicOr(bad_flag, bad);
}

struct IdPos { unsigned long long id, pos; };
__device__ __forceinline__ IdPos min_left(IdPos a, IdPos b) { return (b.id < a.id || (b.id == a.id && b.pos < a.pos)) ? b : a; }
// element (lane + d) of the 64-wide window held as (lo: elements 0..31, hi: elements 32..63), d in [0, 32]
__device__ __forceinline__ IdPos shfl64(IdPos lo, IdPos hi, int d, int lane) {
    const int src = (lane + d) & 31;
    IdPos a, b;
    a.id = __shfl_sync(0xffffffffu, lo.id, src); a.pos = __shfl_sync(0xffffffffu, lo.pos, src);
    b.id = __shfl_sync(0xffffffffu, hi.id, src); b.pos = __shfl_sync(0xffffffffu, hi.pos, src);
    return lane + d < 32 ? a : b;
}

// best[j] = position of the minimum id among ids[j .. j+w-1], leftmost on ties.  w <= 32: each warp
// holds 64 consecutive ids in two registers per lane and builds the window minimum by doubling
// (sparse-table steps through shuffles); lane j answers window base+j.  w > 32: per-thread loop.
__global__ void __launch_bounds__(256) k8_min_window(const unsigned long long* __restrict__ ids, unsigned long long n_win, int w,
                                                     unsigned long long n_out, unsigned long long* __restrict__ best) {
    const int lane = threadIdx.x & 31;
    const unsigned long long warp_global = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long base = warp_global * 32;
    if (base >= n_out) return;
    if (w > 32) {
        const unsigned long long j = base + lane;
        if (j < n_out) {
            IdPos m{ids[j], j};
            for (unsigned long long q = j + 1; q < j + (unsigned long long)w; q++) m = min_left(m, IdPos{ids[q], q});
            best[j] = m.pos;
        }
        return;
    }
    const unsigned long long i0 = base + lane, i1 = base + 32 + lane;
    IdPos lo{i0 < n_win ? ids[i0] : ~0ull, i0}, hi{i1 < n_win ? ids[i1] : ~0ull, i1};
    // after the step with span p: element e holds the minimum of [e, e+p)
    int p = 1;
    while (2 * p <= w) {
        const IdPos lo_s = shfl64(lo, hi, p, lane);
        IdPos hi_s;                                            // element 32+lane+p: only needed while it stays below 64
        { const int src = (lane + p) & 31; hi_s.id = __shfl_sync(0xffffffffu, hi.id, src); hi_s.pos = __shfl_sync(0xffffffffu, hi.pos, src);
          if (lane + p >= 32) { hi_s.id = ~0ull; hi_s.pos = ~0ull; } }
        lo = min_left(lo, lo_s);
        hi = min_left(hi, hi_s);
        p *= 2;
    }
    IdPos r = lo;
    if (w > p) r = min_left(lo, shfl64(lo, hi, w - p, lane));
    const unsigned long long j = base + lane;
    if (j < n_out) best[j] = r.pos;
}

// keep[j] = best[j] != best[j-1]: consecutive windows that pick the same k-mer collapse to one row.
__global__ void __launch_bounds__(256) k8_min_count(const unsigned long long* __restrict__ best, unsigned long long n_out, unsigned int* __restrict__ block_n) {
    __shared__ unsigned int s;
    if (threadIdx.x == 0) s = 0;
    __syncthreads();
    const unsigned long long base = (unsigned long long)blockIdx.x * kEdgeBlock;
    unsigned int c = 0;
    for (int q = threadIdx.x; q < kEdgeBlock; q += 256) { const unsigned long long j = base + q; if (j < n_out && (j == 0 || best[j] != best[j - 1])) c++; }
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s, c);
    __syncthreads();
    if (threadIdx.x == 0) block_n[blockIdx.x] = s;
}
__global__ void __launch_bounds__(256) k8_min_write(const unsigned long long* __restrict__ best, const unsigned long long* __restrict__ ids,
                                                    unsigned long long n_out, const unsigned long long* __restrict__ block_off,
                                                    unsigned long long* __restrict__ out_pos, unsigned long long* __restrict__ out_id, unsigned long long cap) {
    __shared__ unsigned int s_warp[8];
    const unsigned long long base = (unsigned long long)blockIdx.x * kEdgeBlock;
    unsigned long long off = block_off[blockIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int j0 = 0; j0 < kEdgeBlock; j0 += 256) {
        const unsigned long long j = base + j0 + threadIdx.x;
        const bool z = j < n_out && (j == 0 || best[j] != best[j - 1]);
        const unsigned int bal = __ballot_sync(0xffffffffu, z);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        unsigned int before = 0, total = 0;
        for (int q = 0; q < 8; q++) { const unsigned int v = s_warp[q]; if (q < warp) before += v; total += v; }
        if (z) { const unsigned long long pos = off + before + __popc(bal & ((1u << lane) - 1u)); if (pos < cap) { out_pos[pos] = best[j]; out_id[pos] = ids[best[j]]; } }
        off += total;
        __syncthreads();
    }
}

// ---- batched minimizers: many records (reads) in one call ------------------------------------------------
// ids and window minima are computed over the concatenated letters with the kernels above (a window that crosses a
// record boundary yields garbage that is never emitted); these two kernels decide, per letter position p, whether the
// window group starting at p lies inside its record and whether it names a new k-mer, and write the rows
// (record, position inside the record, id).  The record of a position comes from a binary search in a shared-memory
// copy of the slice of rec_off that covers the block.
constexpr int kMinbSlice = kEdgeBlock + 2;
struct RecSlice {
    unsigned long long first;                      // index of the first record of the block
    unsigned int n;                                // entries of rec_off held in shared memory (0 = search global memory)
};
__device__ __forceinline__ RecSlice load_rec_slice(const unsigned long long* __restrict__ rec_off, unsigned long long n_recs,
                                                   unsigned long long p0, unsigned long long p1, unsigned long long* s_off, RecSlice* s_hdr) {
    if (threadIdx.x == 0) {
        const unsigned long long r0 = upper_rec(rec_off, n_recs, p0), r1 = upper_rec(rec_off, n_recs, p1);
        s_hdr->first = r0;
        s_hdr->n = (r1 - r0 + 2 <= (unsigned long long)kMinbSlice) ? (unsigned int)(r1 - r0 + 2) : 0u;
    }
    __syncthreads();
    const RecSlice h = *s_hdr;
    for (unsigned int i = threadIdx.x; i < h.n; i += blockDim.x) s_off[i] = rec_off[h.first + i];
    __syncthreads();
    return h;
}
// record of position p and its bounds
__device__ __forceinline__ unsigned long long rec_of(const unsigned long long* __restrict__ rec_off, unsigned long long n_recs, const RecSlice h,
                                                     const unsigned long long* s_off, unsigned long long p, unsigned long long* start, unsigned long long* end) {
    if (h.n) {
        unsigned int lo = 0, hi = h.n - 1;          // s_off[lo] <= p < s_off[hi]
        while (hi - lo > 1) { const unsigned int mid = (lo + hi) >> 1; if (s_off[mid] <= p) lo = mid; else hi = mid; }
        *start = s_off[lo]; *end = s_off[lo + 1];
        return h.first + lo;
    }
    const unsigned long long r = upper_rec(rec_off, n_recs, p);
    *start = rec_off[r]; *end = rec_off[r + 1];
    return r;
}

template <bool WRITE>
__global__ void __launch_bounds__(256) k8_minb_rows(const unsigned long long* __restrict__ best, const unsigned long long* __restrict__ ids,
                                                    unsigned long long n_out, const unsigned long long* __restrict__ rec_off, unsigned long long n_recs,
                                                    int k, int w, unsigned int* __restrict__ block_n, const unsigned long long* __restrict__ block_off,
                                                    unsigned long long* __restrict__ out_rec, unsigned long long* __restrict__ out_pos,
                                                    unsigned long long* __restrict__ out_id, unsigned long long cap) {
    __shared__ unsigned long long s_off[kMinbSlice];
    __shared__ RecSlice s_hdr;
    __shared__ unsigned int s_warp[8];
    __shared__ unsigned int s_total;
    const unsigned long long base = (unsigned long long)blockIdx.x * kEdgeBlock;
    const unsigned long long last = base + kEdgeBlock - 1 < n_out ? base + kEdgeBlock - 1 : n_out - 1;
    const RecSlice h = load_rec_slice(rec_off, n_recs, base, last, s_off, &s_hdr);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long off = WRITE ? block_off[blockIdx.x] : 0ull;
    unsigned int mine = 0;
    for (int j0 = 0; j0 < kEdgeBlock; j0 += 256) {
        const unsigned long long p = base + j0 + threadIdx.x;
        bool keep = false;
        unsigned long long r = 0, start = 0, end = 0;
        if (p < n_out) {
            r = rec_of(rec_off, n_recs, h, s_off, p, &start, &end);
            keep = p + (unsigned long long)(k + w - 1) <= end && (p == start || best[p] != best[p - 1]);
        }
        if (!WRITE) { mine += keep; continue; }
        const unsigned int bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        unsigned int before = 0, total = 0;
        for (int q = 0; q < 8; q++) { const unsigned int v = s_warp[q]; if (q < warp) before += v; total += v; }
        if (keep) {
            const unsigned long long o = off + before + __popc(bal & ((1u << lane) - 1u));
            if (o < cap) { out_rec[o] = r; out_pos[o] = best[p] - start; out_id[o] = ids[best[p]]; }
        }
        off += total;
        __syncthreads();
    }
    if (!WRITE) {
        if (threadIdx.x == 0) s_total = 0;
        __syncthreads();
        for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
        if (lane == 0 && mine) atomicAdd(&s_total, mine);
        __syncthreads();
        if (threadIdx.x == 0) block_n[blockIdx.x] = s_total;
    }
}

// ---- batched minimizers, fused (r02): letters -> rows in one kernel, nothing staged in HBM ---------------------------------
// One CTA owns kEdgeBlock window-group starts.  Its letters (with the k + w - 1 halo) are classified through a shared-memory
// copy of the class table (the __constant__ table indexed by a data-dependent byte serialises the warp), every thread builds
// the canonical 32-bit ids of a run of consecutive windows by rolling (k <= 16), the window minimum is a doubling pass over
// (id << 32 | position) keys in shared memory — the 64-bit minimum IS the leftmost minimum — and the rows leave compacted.
// Replaces k8_window_ids + k8_min_window + k8_minb_rows and their two uint64 arrays (16 B of HBM per window before any row).
constexpr int kFusedMaxW = 64;
constexpr int kFusedIds = kEdgeBlock + 1 + kFusedMaxW;             // ids a CTA may need (one group before the block, w - 1 after)
template <bool WRITE>
__global__ void __launch_bounds__(256) k8_minb_fused(const uint8_t* __restrict__ letters, unsigned long long n_letters, unsigned long long n_out,
                                                     const unsigned long long* __restrict__ rec_off, unsigned long long n_recs, int k, int w,
                                                     unsigned int* __restrict__ block_n, const unsigned long long* __restrict__ block_off,
                                                     unsigned long long* __restrict__ out_rec, unsigned long long* __restrict__ out_pos,
                                                     unsigned long long* __restrict__ out_id, unsigned long long cap, unsigned int* bad_flag) {
    __shared__ unsigned long long s_key[2][kFusedIds];
    __shared__ uint8_t s_cls[kFusedIds + 16];
    __shared__ uint8_t s_lut[256];
    __shared__ unsigned int s_off[kMinbSlice];                      // record offsets of the block's slice, relative to its first record
    __shared__ RecSlice s_hdr;
    __shared__ unsigned long long s_off0;
    __shared__ unsigned int s_warp[8];
    __shared__ unsigned int s_total;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long base = (unsigned long long)blockIdx.x * kEdgeBlock;
    const unsigned long long last = base + kEdgeBlock - 1 < n_out ? base + kEdgeBlock - 1 : n_out - 1;
    s_lut[tid] = c_class_lut[tid];
    if (tid == 0) {
        const unsigned long long r0 = upper_rec(rec_off, n_recs, base), r1 = upper_rec(rec_off, n_recs, last);
        s_hdr.first = r0;
        s_hdr.n = (r1 - r0 + 2 <= (unsigned long long)kMinbSlice && rec_off[r1 + 1] - rec_off[r0] < 0xFFFFFFFFull) ? (unsigned int)(r1 - r0 + 2) : 0u;
        s_off0 = rec_off[r0];
    }
    __syncthreads();
    const RecSlice h = s_hdr;
    const unsigned long long off0 = s_off0;
    for (unsigned int i = tid; i < h.n; i += 256) s_off[i] = (unsigned int)(rec_off[h.first + i] - off0);
    __syncthreads();
    auto rec_of32 = [&](unsigned long long pq, unsigned long long* start, unsigned long long* end) -> unsigned long long {
        if (h.n) {
            const unsigned int x = (unsigned int)(pq - off0);
            unsigned int lo = 0, hi = h.n - 1;          // s_off[lo] <= x < s_off[hi]
            while (hi - lo > 1) { const unsigned int mid = (lo + hi) >> 1; if (s_off[mid] <= x) lo = mid; else hi = mid; }
            *start = off0 + s_off[lo]; *end = off0 + s_off[lo + 1];
            return h.first + lo;
        }
        const unsigned long long r = upper_rec(rec_off, n_recs, pq);
        *start = rec_off[r]; *end = rec_off[r + 1];
        return r;
    };
    // local index q <-> global position g0 + q; q = 0 is the group before the block (its minimum decides whether base opens a new row)
    const long long g0 = (long long)base - 1;
    const int n_ids = kEdgeBlock + w;                               // ids [0, n_ids): groups 0 .. kEdgeBlock, each spanning w ids
    const int n_let = n_ids + k - 1;
    unsigned int bad = 0;
    for (int i = tid; i < n_let; i += 256) {
        const long long g = g0 + i;
        uint8_t cl = C_A;
        if (g >= 0 && (unsigned long long)g < n_letters) {
            cl = s_lut[letters[g]];
            if (cl == C_N || cl == C_IUPAC) bad |= 2u; else if (cl > C_T) bad |= 1u;
        }
        s_cls[i] = cl & 3u;
    }
    __syncthreads();
    {   // ids: a run of consecutive windows per thread, rolled
        const int per = (n_ids + 255) / 256, i0 = tid * per, i1 = min(n_ids, i0 + per);
        if (i0 < i1) {
            const uint32_t mask = k == 16 ? 0xFFFFFFFFu : ((1u << (2 * k)) - 1u);
            uint32_t fwd = 0, rc = 0;
            for (int j = 0; j < k; j++) { const uint32_t c = s_cls[i0 + j]; fwd = (fwd << 2) | c; rc |= (3u - c) << (2 * j); }
            for (int i = i0; ; ) {
                const uint32_t id = rc < fwd ? rc : fwd;
                s_key[0][i] = ((unsigned long long)id << 32) | (unsigned int)i;
                if (++i >= i1) break;
                const uint32_t c = s_cls[i + k - 1];
                fwd = ((fwd << 2) | c) & mask;
                rc = (rc >> 2) | ((3u - c) << (2 * (k - 1)));
            }
        }
    }
    __syncthreads();
    // window minimum by doubling: after the pass with span p, key[i] = min of ids [i, i + p)
    int cur = 0, p = 1;
    while (2 * p <= w) {
        for (int i = tid; i < n_ids; i += 256) {
            const unsigned long long a = s_key[cur][i], b = i + p < n_ids ? s_key[cur][i + p] : ~0ull;
            s_key[cur ^ 1][i] = b < a ? b : a;
        }
        __syncthreads();
        cur ^= 1; p *= 2;
    }
    if (w > p) {
        for (int i = tid; i <= kEdgeBlock; i += 256) {
            const unsigned long long a = s_key[cur][i], b = s_key[cur][i + w - p];
            s_key[cur ^ 1][i] = b < a ? b : a;
        }
        __syncthreads();
        cur ^= 1;
    }
    const unsigned long long* best = s_key[cur];                    // best[q]: (id, local position) of the minimum of group q
    unsigned long long off = WRITE ? block_off[blockIdx.x] : 0ull;
    unsigned int mine = 0;
    for (int j0 = 0; j0 < kEdgeBlock; j0 += 256) {
        const int q = 1 + j0 + tid;
        const unsigned long long pg = base + j0 + tid;
        bool keep = false;
        unsigned long long r = 0, start = 0, end = 0;
        if (pg < n_out) {
            r = rec_of32(pg, &start, &end);
            keep = pg + (unsigned long long)(k + w - 1) <= end && (pg == start || (unsigned int)best[q] != (unsigned int)best[q - 1]);
        }
        if (!WRITE) { mine += keep; continue; }
        const unsigned int bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        unsigned int before = 0, total = 0;
        for (int x = 0; x < 8; x++) { const unsigned int v = s_warp[x]; if (x < warp) before += v; total += v; }
        if (keep) {
            const unsigned long long o = off + before + __popc(bal & ((1u << lane) - 1u));
            if (o < cap) { out_rec[o] = r; out_pos[o] = (unsigned long long)(g0 + (long long)(unsigned int)best[q]) - start; out_id[o] = best[q] >> 32; }
        }
        off += total;
        __syncthreads();
    }
    if (!WRITE) {
        if (tid == 0) s_total = 0;
        __syncthreads();
        for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
        if (lane == 0 && mine) atomicAdd(&s_total, mine);
        __syncthreads();
        if (tid == 0) block_n[blockIdx.x] = s_total;
        if (bad) atomicOr(bad_flag, bad);
    }
}

}  // namespace kdb
