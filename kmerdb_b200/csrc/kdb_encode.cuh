// K2 — tokenise + 2-bit encode (FASTA / FASTQ text -> packed letter stream), sm_100a.
//
// Replaces the tokenising half of kmer.shred (kmerdb/kmer.py:517-534: alphabet check, N / IUPAC
// detection, window completeness) and the record bookkeeping of parse.parsefile
// (kmerdb/parse.py:128-137: seq_lengths).  For every 16 KiB text tile it writes a packed region
//   [0] n_letters  [1] halo codes  [2] halo N bits  [3] halo W bits (W of the <=15 letters before the tile,
//       most recent at bit 0: lets K3 derive the W bits of a larger k from this stream, see smear_w)
//   codes : 2 bits / letter, 16 letters per u32, first letter in the top bits (kmer.py:44-49 codes)
//   W     : 1 bit / letter — the window ENDING at this letter is complete (k letters of one record)
//           and free of RYSWKMBDHV; in skip mode (replace_with_none) also free of N
//   NN    : 1 bit / letter — the letter is N (only consumed in expansion mode)
// so that the count kernel (K3) runs with uniform control flow on clean 2-bit data.
#pragma once
#include "kdb_common.cuh"
#include "kdb_scan.cuh"

namespace kdb {

constexpr int kPackHdr = 4;
constexpr int kPackCodes = kTile / 16;                 // 1024 words
constexpr int kPackBits = kTile / 32;                  // 512 words (W) + 512 words (NN)
constexpr int kPackStride = kPackHdr + kPackCodes + 2 * kPackBits;   // 2052 words per tile

struct EncodeParams {
    const uint8_t* text;
    unsigned long long lo, hi, n;
    const TileState* tiles;
    unsigned int n_tiles;
    int k, skip_n;
    int k_halo;                       // >= k: letters of history the tile header carries (largest k counted from this stream)
    uint32_t* packed;                 // n_tiles * kPackStride words
    unsigned long long* rec_len;
    unsigned long long rec_first;
    unsigned long long fq_rec_base;
    Accum* acc;
};

// Window-tracking state of one thread (no ids here: K3 builds them from the packed stream).
struct Track {
    uint32_t nmask, imask;   // bit j: the letter j positions back is N / RYSWKMBDHV (or illegal)
    uint32_t have;           // letters of the current record so far, saturated at k
};

// Per-thread output accumulators: letters right-aligned, most recent letter in the low bits.
struct Emit {
    unsigned long long chi, clo;   // 2-bit codes
    unsigned long long wacc, nacc; // W / NN bits
    uint32_t n;
};

struct Tally {
    unsigned int mono[4];
    unsigned int nwin;       // complete windows holding N
    unsigned int err;        // bit P_ALPHABET / bit P_IUPAC for the current record
};

__device__ __forceinline__ void emit1(Emit& e, uint32_t code, uint32_t wbit, uint32_t nbit) {
    e.chi = (e.chi << 2) | (e.clo >> 62);
    e.clo = (e.clo << 2) | code;
    e.wacc = (e.wacc << 1) | wbit;
    e.nacc = (e.nacc << 1) | nbit;
    e.n++;
}

// One sequence letter of class cl (C_A..C_ILLEGAL) through the window tracker.
template <bool EMIT>
__device__ __forceinline__ void track_letter(Track& t, Emit& e, Tally& y, uint32_t cl, int k, uint32_t kmask, int skip_n) {
    const uint32_t isn = cl == C_N ? 1u : 0u;
    t.nmask = ((t.nmask << 1) | isn) & kmask;
    t.imask = ((t.imask << 1) | (cl >= C_IUPAC ? 1u : 0u)) & kmask;
    t.have = t.have < (uint32_t)k ? t.have + 1 : t.have;
    if (!EMIT) return;
    uint32_t wbit = 0;
    if (cl <= C_T) y.mono[cl]++;
    if (t.have >= (uint32_t)k) {
        if (t.nmask) y.nwin++;
        if (t.nmask && skip_n) wbit = 0;                                  // kmer.py:542-544
        else if (t.imask) y.err |= 1u << P_IUPAC;                         // kmer.py:309
        else wbit = 1;
    }
    emit1(e, cl <= C_T ? cl : 0u, wbit, isn);
}

// Classes of the up-to-(k-1) letters that precede byte `a` inside the same record, oldest first.
template <bool FASTQ>
__device__ __noinline__ int halo_classes(const uint8_t* text, unsigned long long a, int k, uint8_t* out) {
    uint8_t cls[16];
    int cnt = 0;
    unsigned long long p = a;
    while (cnt < k - 1 && p > 0) {
        const uint8_t ch = text[p - 1];
        if (ch == '\n') {
            if (FASTQ) break;
            unsigned long long s = p - 1;                       // start of the line this LF ends
            while (s > 0 && text[s - 1] != '\n') s--;
            if (text[s] == '>') break;                          // title line: the record starts after it
            p--;
            continue;
        }
        const uint8_t cl = c_class_lut[ch];
        if (cl <= C_ILLEGAL) cls[cnt++] = cl;
        p--;
    }
    for (int i = 0; i < cnt; i++) out[i] = cls[cnt - 1 - i];
    return cnt;
}

}  // namespace kdb

namespace kdb {

// OR a left-aligned bit string (x[0] holds the first bits; nwords words) into the big-endian bit
// array `dst` at bit offset `bitoff`.  Shared-memory atomics: neighbouring threads share words.
template <int NW>
__device__ __forceinline__ void or_bits(uint32_t* dst, uint32_t bitoff, const uint32_t (&x)[NW], uint32_t nbits) {
    if (nbits == 0) return;
    const uint32_t w0 = bitoff >> 5, b = bitoff & 31;
    const uint32_t last = (bitoff + nbits - 1) >> 5;
#pragma unroll
    for (int q = 0; q <= NW; q++) {
        if (w0 + q > last) break;
        const uint32_t cur = q < NW ? x[q < NW ? q : 0] : 0u;
        const uint32_t prv = q > 0 ? x[q > 0 ? q - 1 : 0] : 0u;
        const uint32_t v = __funnelshift_r(cur, prv, b);     // (prv:cur) >> b, low word
        if (v) atomicOr(dst + w0 + q, v);
    }
}

// 4-bit gather of the 0x80 byte flags of m (byte j -> bit j)
// (flags at bits 7, 15, 23, 31: the product puts flag j at bit 28 + j and every other partial product below bit 28 or beyond
// bit 31, each on a bit of its own, so nothing carries: one IMAD + one shift)
__device__ __forceinline__ uint32_t nib(uint32_t m) { return (m * 0x00204081u) >> 28; }
__device__ __forceinline__ unsigned long long ones64(uint32_t n) { return n >= 64 ? ~0ull : ((1ull << n) - 1ull); }
// x | x>>1 | ... | x>>(n-1)   (n >= 1)
__device__ __forceinline__ unsigned long long smear_down(unsigned long long x, uint32_t n) {
    uint32_t w = 1;
    while (2 * w <= n) { x |= x >> w; w *= 2; }
    if (n > w) x |= x >> (n - w);
    return x;
}

// Everything the bit-parallel path knows about one 64-byte chunk (bit i <-> byte i).
struct ChunkMasks {
    unsigned long long lf;      // byte == LF
    unsigned long long acgt;    // byte in ACGT
    uint32_t x[4];              // 2-bit codes of all 64 bytes, byte 0 in the top bits of x[0] (garbage for non-letters, 0 for N)
    uint32_t any_cr, any_hi;    // some byte is CR / >= 0x80
};

template <typename Quad>
__device__ __forceinline__ ChunkMasks chunk_masks(Quad quad) {
    ChunkMasks M;
    M.lf = 0; M.acgt = 0; M.any_cr = 0; M.any_hi = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint32_t xc = 0, lf16 = 0, ok16 = 0;
        const uint4 v4 = quad(q);                                                  // one conflict-free 16-byte shared load
        const uint32_t w4[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t wv = w4[r];
            const uint32_t t = ((wv >> 1) ^ (wv >> 2)) & 0x03030303u;              // A0 C1 G2 T3 (N -> 0)
            const uint32_t s = t | (t >> 4);
            const uint32_t sel = (s & 0xFFu) | ((s >> 8) & 0xFF00u);
            const uint32_t d = __byte_perm(0x54474341u, 0u, sel) ^ wv;            // zero byte <=> ACGT
            const uint32_t z = ~(((d & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | d | 0x7F7F7F7Fu);
            ok16 |= nib(z) << (4 * r);
            lf16 |= nib(eqmask(wv, 0x0A0A0A0Au)) << (4 * r);
            xc = (xc << 8) | ((t * 0x40100401u) >> 24);
            M.any_cr |= eqmask(wv, 0x0D0D0D0Du);
            M.any_hi |= wv;
        }
        M.x[q] = xc;
        M.lf |= (unsigned long long)lf16 << (16 * q);
        M.acgt |= (unsigned long long)ok16 << (16 * q);
    }
    M.any_hi &= 0x80808080u;
    return M;
}

// The two halves of chunk_masks for the single-pass tokeniser: the line-feed mask first (all the line scan needs), the letter
// masks and codes later — in k2s_encode they fill the time between publishing a tile's aggregate and looking back for its
// predecessors' (kdb_encode2.cuh).
template <typename Quad>
__device__ __forceinline__ unsigned long long chunk_lf(Quad quad, uint32_t* any_cr, uint32_t* any_hi) {
    unsigned long long lf = 0;
    uint32_t cr = 0, hi = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const uint4 v4 = quad(q);
        const uint32_t m16 = nib(eqmask(v4.x, 0x0A0A0A0Au)) | (nib(eqmask(v4.y, 0x0A0A0A0Au)) << 4) | (nib(eqmask(v4.z, 0x0A0A0A0Au)) << 8) |
                             (nib(eqmask(v4.w, 0x0A0A0A0Au)) << 12);
        lf |= (unsigned long long)m16 << (16 * q);
        cr |= eqmask(v4.x, 0x0D0D0D0Du) | eqmask(v4.y, 0x0D0D0D0Du) | eqmask(v4.z, 0x0D0D0D0Du) | eqmask(v4.w, 0x0D0D0D0Du);
        hi |= v4.x | v4.y | v4.z | v4.w;
    }
    *any_cr = cr; *any_hi = hi & 0x80808080u;
    return lf;
}
// letter validity + 2-bit codes: only chunks that hold sequence-line bytes need them (FASTQ: about half)
template <typename Quad>
__device__ __forceinline__ ChunkMasks chunk_masks_rest(Quad quad, unsigned long long lf, uint32_t any_cr, uint32_t any_hi) {
    ChunkMasks M;
    M.lf = lf; M.acgt = 0; M.any_cr = any_cr; M.any_hi = any_hi;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint32_t xc = 0, ok16 = 0;
        const uint4 v4 = quad(q);
        const uint32_t w4[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint32_t wv = w4[r];
            const uint32_t t = ((wv >> 1) ^ (wv >> 2)) & 0x03030303u;              // A0 C1 G2 T3 (N -> 0)
            const uint32_t s = t | (t >> 4);
            const uint32_t sel = (s & 0xFFu) | ((s >> 8) & 0xFF00u);
            const uint32_t d = __byte_perm(0x54474341u, 0u, sel) ^ wv;            // zero byte <=> ACGT
            const uint32_t z = ~(((d & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | d | 0x7F7F7F7Fu);
            ok16 |= nib(z) << (4 * r);
            xc = (xc << 8) | ((t * 0x40100401u) >> 24);
        }
        M.x[q] = xc;
        M.acgt |= (unsigned long long)ok16 << (16 * q);
    }
    return M;
}

// 64-bit byte-equality mask, computed only when needed (N letters, CR).
template <typename ByteWord>
__device__ __noinline__ unsigned long long lazy_eq_mask(ByteWord word, uint32_t pat4) {
    unsigned long long m = 0;
    for (int i = 0; i < kChunk / 4; i++) m |= (unsigned long long)nib(eqmask(word(i), pat4)) << (4 * i);
    return m;
}
// FASTQ quality bytes outside chr 33..126 (Bio.SeqIO raises ValueError)
template <typename ByteWord>
__device__ __noinline__ unsigned long long lazy_qbad_mask(ByteWord word) {
    unsigned long long m = 0;
    for (int i = 0; i < kChunk / 4; i++) {
        const uint32_t wv = word(i);
        const uint32_t lo = ~(((wv & 0x7F7F7F7Fu) + 0x5F5F5F5Fu) | wv) & 0x80808080u;     // byte < 0x21 (and < 0x80)
        m |= (unsigned long long)nib(lo | (wv & 0x80808080u) | eqmask(wv, 0x7F7F7F7Fu)) << (4 * i);
    }
    return m;
}

// bits [2*(s+len)-1 .. 2*s] counted from the top of the 128-bit code string x, right-aligned
__device__ __forceinline__ void extract_codes(const uint32_t (&x)[4], uint32_t s, uint32_t len, unsigned long long* hi, unsigned long long* lo) {
    unsigned long long h = ((unsigned long long)x[0] << 32) | x[1], l = ((unsigned long long)x[2] << 32) | x[3];
    const uint32_t sh = 128 - 2 * (s + len);                 // right shift, 0..126
    if (sh >= 64) { l = h >> (sh - 64); h = 0; }
    else if (sh) { l = (l >> sh) | (h << (64 - sh)); h >>= sh; }
    const uint32_t nb = 2 * len;                              // keep the low nb bits
    if (nb < 64) { h = 0; l &= (1ull << nb) - 1ull; }
    else if (nb < 128) h &= (1ull << (nb - 64)) - 1ull;
    *hi = h; *lo = l;
}

// What a chunk hands to the chunks after it: the last (up to 15) letters of the record that is open
// at its end.  Associative, so the start state of every thread is one CTA scan away.
struct TailAgg {
    uint32_t codes;    // 2-bit codes, most recent letter in the low bits
    uint32_t ni;       // N flags (low 16) and RYSWKMBDHV/illegal flags (high 16), most recent letter at bit 0
    uint32_t meta;     // letters held (0..15) | reset << 8 (a record started inside the span)
    uint32_t letters;  // all letters emitted by the span (for the output offsets)
};
__device__ __forceinline__ TailAgg tail_combine(const TailAgg a, const TailAgg b) {
    TailAgg r;
    r.letters = a.letters + b.letters;
    if (b.meta >> 8) { r.codes = b.codes; r.ni = b.ni; r.meta = b.meta; return r; }
    const uint32_t bc = b.meta & 0xFF, ac = a.meta & 0xFF;
    const uint32_t tot = ac + bc < 15 ? ac + bc : 15;
    r.codes = ((a.codes << (2 * bc)) | b.codes) & 0x3FFFFFFFu;
    const uint32_t an = a.ni & 0xFFFF, ai = a.ni >> 16, bn = b.ni & 0xFFFF, bi = b.ni >> 16;
    r.ni = (((an << bc) | bn) & 0x7FFFu) | ((((ai << bc) | bi) & 0x7FFFu) << 16);
    r.meta = tot | (a.meta & 0x100);
    return r;
}
__device__ __forceinline__ TailAgg tail_shfl_up(const TailAgg v, int d) {
    TailAgg r;
    r.codes = __shfl_up_sync(0xffffffffu, v.codes, d); r.ni = __shfl_up_sync(0xffffffffu, v.ni, d);
    r.meta = __shfl_up_sync(0xffffffffu, v.meta, d); r.letters = __shfl_up_sync(0xffffffffu, v.letters, d);
    return r;
}
// Exclusive scan over the CTA; *total receives the CTA aggregate.  sm: kThreads/32 entries.
__device__ __forceinline__ TailAgg tail_scan_exclusive(TailAgg v, TailAgg* sm, TailAgg* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const TailAgg ident{0, 0, 0, 0};
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const TailAgg o = tail_shfl_up(v, d); if (lane >= d) v = tail_combine(o, v); }
    if (lane == 31) sm[warp] = v;
    __syncthreads();
    if (warp == 0) {                                         // inclusive scan of the warp aggregates by the first warp
        TailAgg w = lane < kThreads / 32 ? sm[lane] : ident;
#pragma unroll
        for (int d = 1; d < kThreads / 32; d <<= 1) { const TailAgg o = tail_shfl_up(w, d); if (lane >= d) w = tail_combine(o, w); }
        if (lane < kThreads / 32) sm[lane] = w;
    }
    __syncthreads();
    const TailAgg pre = warp ? sm[warp - 1] : ident;
    *total = sm[kThreads / 32 - 1];
    TailAgg e = tail_shfl_up(v, 1);
    if (lane == 0) e = ident;
    return tail_combine(pre, e);
}

// W bits of the E.n letters gathered by the bit-parallel path (all ACGT or N), given the state the
// record was in when the chunk began.  Returns true when a window must raise KeyError (kmer.py:309).
__device__ __forceinline__ bool finalize_w(Emit& E, Tally& Y, unsigned long long rs_acc, uint32_t have0, uint32_t nmask0, uint32_t imask0, int k, int skip_n) {
    const uint32_t n = E.n;
    nmask0 &= (1u << (k - 1)) - 1u; imask0 &= (1u << (k - 1)) - 1u;     // only the k-1 letters before the chunk matter
    unsigned long long incp = k > 1 ? smear_down(rs_acc, (uint32_t)k - 1) : 0ull;
    if (have0 + 1 < (uint32_t)k) { const uint32_t c = min(n, (uint32_t)k - 1 - have0); incp |= ones64(c) << (n - c); }
    unsigned long long nsm = smear_down(E.nacc, (uint32_t)k);
    if (nmask0) { const uint32_t c2 = min(n, (uint32_t)k - 1 - (uint32_t)(__ffs(nmask0) - 1)); nsm |= ones64(c2) << (n - c2); }
    const unsigned long long complete = ones64(n) & ~incp;
    Y.nwin += __popcll(complete & nsm);
    unsigned long long w = skip_n ? complete & ~nsm : complete;
    bool raise = false;
    if (imask0) {                         // a RYSWKMBDHV letter sits in the k-1 letters before the chunk
        const uint32_t c3 = min(n, (uint32_t)k - 1 - (uint32_t)(__ffs(imask0) - 1));
        const unsigned long long aff = (ones64(c3) << (n - c3)) & w;
        raise = aff != 0;
        w &= ~aff;
    }
    E.wacc = w;
    const unsigned long long o5 = 0x5555555555555555ull;
    const uint32_t lo = __popcll(E.clo & o5) + __popcll(E.chi & o5), hi = __popcll(E.clo & (o5 << 1)) + __popcll(E.chi & (o5 << 1));
    const uint32_t both = __popcll(E.clo & (E.clo >> 1) & o5) + __popcll(E.chi & (E.chi >> 1) & o5);
    Y.mono[3] += both; Y.mono[2] += hi - both; Y.mono[1] += lo - both; Y.mono[0] += n - lo - hi + both - __popcll(E.nacc);
    return raise;
}

template <bool FASTQ>
__global__ void __launch_bounds__(kThreads, 4) k2_encode(const EncodeParams P) {
    // two text buffers (cp.async double buffering).  Chunk t (64 bytes) lives at words [16t, 16t+16) with its four
    // 16-byte groups XOR-swizzled by (t >> 1) & 3, so that the per-thread LDS.128 of a quarter warp hits 8 distinct
    // bank groups.
    __shared__ __align__(16) uint32_t s_text2[2][kTile / 4];
    __shared__ uint32_t s_out[kPackCodes + 2 * kPackBits];
    __shared__ uint8_t s_lut[256];
    __shared__ LineAgg s_scan[kThreads / 32];
    __shared__ TailAgg s_tail[kThreads / 32];
    __shared__ TailAgg s_seed;
    __shared__ unsigned long long s_red[kThreads / 32][8];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = P.k, skip_n = P.skip_n;
    const uint32_t kmask = (1u << k) - 1u;
    s_lut[tid] = c_class_lut[tid];

    Tally Y;
    Y.mono[0] = Y.mono[1] = Y.mono[2] = Y.mono[3] = 0; Y.nwin = 0; Y.err = 0;
    unsigned long long letters_total = 0, err_key = ~0ull;
    unsigned int layout = 0;
    const uint32_t* s_text = s_text2[0];
    const uint32_t swz = ((uint32_t)tid >> 1) & 3u;
    auto word = [&](int i) -> uint32_t { return s_text[16 * tid + 4 * (((uint32_t)i >> 2) ^ swz) + (i & 3)]; };
    auto quad = [&](int q) -> uint4 { return *reinterpret_cast<const uint4*>(s_text + 16 * tid + 4 * ((uint32_t)q ^ swz)); };
    auto byte_at = [&](uint32_t p) -> uint8_t { return (uint8_t)(s_text[16 * tid + 4 * ((p >> 4) ^ swz) + ((p >> 2) & 3)] >> (8 * (p & 3))); };
    // asynchronous copy of my own 64-byte chunk of one tile into buffer `b` (four 16-byte requests, zero fill past
    // P.hi).  A thread only ever reads the chunk it copied, so cp.async.wait_group is all the synchronisation the text
    // needs; .ca keeps the 32-byte sectors in L1 between the two requests that share them.
    auto stage = [&](unsigned int tile, int b) {
        const unsigned long long c0 = P.lo + (unsigned long long)tile * kTile + (unsigned long long)tid * kChunk;
#pragma unroll
        for (int r = 0; r < kChunk / 16; r++) {
            const unsigned long long g = c0 + 16ull * r;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&s_text2[b][16 * tid + 4 * ((uint32_t)r ^ swz)]);
            const uint32_t nb = g + 16 <= P.hi ? 16u : (g < P.hi ? (uint32_t)(P.hi - g) : 0u);
            const uint8_t* src = P.text + (nb ? g : P.lo);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(nb) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    int buf = 0;
    if (blockIdx.x < P.n_tiles) stage(blockIdx.x, 0);
    for (int i = tid; i < kPackCodes + 2 * kPackBits; i += kThreads) s_out[i] = 0;
    __syncthreads();
    for (unsigned int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, buf ^= 1) {
        const unsigned long long tile0 = P.lo + (unsigned long long)tile * kTile;
        // ---- stage: the next tile streams into the other buffer while this one is processed ----------------
        if (tile + gridDim.x < P.n_tiles) { stage(tile + gridDim.x, buf ^ 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        s_text = s_text2[buf];

        const unsigned long long a = tile0 + (unsigned long long)tid * kChunk;
        const int lim = a >= P.hi ? 0 : (P.hi - a >= (unsigned long long)kChunk ? kChunk : (int)(P.hi - a));
        uint8_t prev = '\n';
        {   // last byte of the chunk before mine: from the lane below, or from the text for the first lane of a warp
            const uint32_t up = __shfl_up_sync(0xffffffffu, word(15) >> 24, 1);
            if (lim > 0 && a > 0) prev = lane > 0 ? (uint8_t)up : P.text[a - 1];
        }

        // ---- phase A: bit masks of my chunk; phase B: line starts -> CTA scan ---------------------
        const ChunkMasks M = chunk_masks(quad);      // zero fill beyond lim is neither LF nor a letter
        LineAgg g{0, 0, 0};
        if (lim > 0) {
            unsigned long long lsm = ((M.lf << 1) | (prev == '\n' ? 1ull : 0ull)) & ones64((uint32_t)lim);
            g.ls = __popcll(lsm);
            if (!FASTQ && lsm) {
                const uint32_t top = 63 - __clzll(lsm);
                g.last = byte_at(top) == '>' ? 2u : 1u;
                while (lsm) { const uint32_t p = __ffsll(lsm) - 1; lsm &= lsm - 1; g.hdr += byte_at(p) == '>' ? 1u : 0u; }
            } else if (lsm) g.last = 1u;
        }
        // CTA scan with one barrier: warp scans by shuffles, then every warp folds the (<= 8) warp totals below it itself
        LineAgg inc = g, warp_pre{0, 0, 0};
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const LineAgg o = shfl_up_agg(inc, d); if (lane >= d) inc = combine(o, inc); }
        if (lane == 31) s_scan[warp] = inc;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < kThreads / 32; w++) { const LineAgg t = s_scan[w]; if (w < warp) warp_pre = combine(warp_pre, t); }
        inc = combine(warp_pre, inc);
        const TileState ts = P.tiles[tile];
        const unsigned long long ls_before = ts.ls_base + (inc.ls - g.ls);
        unsigned long long hdr_before = ts.hdr_base + (inc.hdr - g.hdr);
        unsigned int last_before;
        {
            LineAgg e = shfl_up_agg(inc, 1);
            if (lane == 0) e = warp_pre;
            last_before = e.last ? e.last : ts.last;
        }

        Emit E; E.chi = E.clo = E.wacc = E.nacc = 0; E.n = 0;
        unsigned long long rs_acc = 0, rec0 = 0;        // record-start marks (accumulator orientation); record open at `a`
        bool finalized = true;                          // W bits still to be derived after the tail scan?
        TailAgg tail{0, 0, 0, 0}, seed{0, 0, 0x100, 0};
        if (lim > 0) {
            unsigned long long line = ls_before;      // FASTQ: index of the next line start
            uint32_t kind;
            if (FASTQ) kind = ls_before ? (uint32_t)((ls_before - 1) & 3) : K_HEADER;
            else kind = last_before == 2 ? K_HEADER : (hdr_before ? K_SEQ : K_IGNORE);
            unsigned long long rec = FASTQ ? P.fq_rec_base + (ls_before ? (ls_before - 1) >> 2 : 0) : (hdr_before ? hdr_before - 1 : 0);
            Track T; T.nmask = T.imask = T.have = 0;
            bool need_halo;
            if (FASTQ) need_halo = prev != '\n' && kind == 1;
            else if (prev != '\n') need_halo = kind == K_SEQ;
            else need_halo = hdr_before != 0 && byte_at(0) != '>';
            uint32_t halo_codes = 0, halo_nn = 0, halo_w = 0;
            bool walked = false;
            rec0 = rec;
            // k-1 letters before my chunk, read back from the text: only the first thread of a tile, ragged
            // tail chunks and chunks that fall to the byte-wise path need it; everyone else gets the same
            // state from the CTA scan of the chunk tails below
            auto walk = [&]() {
                walked = true;
                if (!need_halo) return;
                uint8_t hc[16];
                const int cnt = halo_classes<FASTQ>(P.text, a, P.k_halo, hc);
                Emit dummyE; Tally dummyY;
                for (int i = 0; i < cnt; i++) {
                    track_letter<false>(T, dummyE, dummyY, hc[i], k, kmask, skip_n);
                    halo_w = (halo_w << 1) | ((T.have >= (uint32_t)k && !(T.nmask && skip_n) && !T.imask) ? 1u : 0u);
                    halo_codes = (halo_codes << 2) | (hc[i] <= C_T ? hc[i] : 0u);
                    halo_nn = (halo_nn << 1) | (hc[i] == C_N ? 1u : 0u);
                }
            };
            if (tid == 0 || lim != kChunk) walk();
            if (tid == 0) {
                uint32_t* hdr = P.packed + (size_t)tile * kPackStride;
                hdr[1] = halo_codes; hdr[2] = halo_nn; hdr[3] = halo_w;
                seed.codes = halo_codes; seed.ni = T.nmask | (T.imask << 16); seed.meta = T.have | 0x100; seed.letters = 0;
            }
            unsigned long long len = 0, qlen = 0;
            Y.err = 0;

            auto flush = [&]() {
                if (len | qlen | Y.err) {
                    const unsigned long long idx = rec - P.rec_first;
                    if (FASTQ) { if (len) atomicAdd(P.rec_len + 2 * idx, len); if (qlen) atomicAdd(P.rec_len + 2 * idx + 1, qlen); }
                    else if (len) atomicAdd(P.rec_len + idx, len);
                    letters_total += len;
                    if (Y.err) {
                        const unsigned long long key = rec * 4 + ((Y.err & 1u) ? P_ALPHABET : P_IUPAC);
                        err_key = key < err_key ? key : err_key;
                    }
                }
                len = 0; qlen = 0; Y.err = 0;
            };
            // line-start bookkeeping shared by both paths; ch is the first byte of the line
            bool rs_in_chunk = false;
            auto begin_line = [&](uint8_t ch) {
                if (FASTQ ? (line & 3) == 0 : ch == '>') rs_in_chunk = true;
                if (FASTQ) {
                    kind = (uint32_t)(line & 3);
                    if (kind == 0) { flush(); rec = P.fq_rec_base + (line >> 2); T.have = 0; T.nmask = T.imask = 0; if (ch != '@') layout |= 1; }
                    else if (kind == 2 && ch != '+') layout |= 1;
                    line++;
                } else {
                    if (ch == '>') { flush(); hdr_before++; rec = hdr_before - 1; kind = K_HEADER; T.have = 0; T.nmask = T.imask = 0; }
                    else kind = hdr_before ? K_SEQ : K_IGNORE;
                }
            };

            // ---- phase C: bit-parallel walk over the line segments of my chunk ------------------
            uint32_t pos = 0;
            bool slow = lim != kChunk;
            if (M.any_hi) layout |= 1;                                  // UTF-8 decoding is the host tokeniser's business
            uint32_t have_run = 0;                                      // letters gathered since the last record start in the chunk
            bool seen_rs = false, pending_rs = false;
            unsigned long long nm = 0, crm = 0, qbad = 0;
            bool have_nm = false, have_q = false;
            if (!slow) {
                if (M.any_cr) {
                    crm = lazy_eq_mask(word, 0x0D0D0D0Du);
                    if (crm & ~(M.lf >> 1) & ~(1ull << 63)) layout |= 1;   // CR not followed by LF: a line break in text mode
                }
                if (prev == '\r' && !(M.lf & 1ull)) layout |= 1;
                unsigned long long m = M.lf;
                bool at_ls = prev == '\n';
                while (true) {
                    const uint32_t e = m ? (uint32_t)__ffsll(m) - 1 : 64u;      // segment [pos, e), LF at e
                    if (at_ls) {
                        const uint8_t ch = byte_at(pos);
                        begin_line(ch);
                        if (FASTQ ? kind == 0 : (ch == '>')) pending_rs = true;
                        prev = 0;                                           // line start consumed (the byte-wise path must not redo it)
                    }
                    const unsigned long long seg = (ones64(e) & ~ones64(pos)) & ~crm;
                    if (kind == 1) {                                        // sequence line
                        const unsigned long long bad = seg & ~M.acgt;
                        if (bad) {
                            if (!have_nm) { nm = lazy_eq_mask(word, 0x4E4E4E4Eu); have_nm = true; }
                            if (bad & ~nm) { slow = true; break; }          // IUPAC / lowercase / space / ...: byte-wise
                        }
                        const uint32_t L = __popcll(seg);
                        if (L) {
                            unsigned long long h, l;
                            extract_codes(M.x, pos, L, &h, &l);
                            const unsigned long long nrun = have_nm ? (__brevll((nm >> pos) & ones64(L)) >> (64 - L)) : 0ull;
                            if (L == 64) { E.chi = h; E.clo = l; E.nacc = nrun; rs_acc = 0; }
                            else {
                                if (2 * L >= 64) { E.chi = (E.clo << (2 * L - 64)) | h; E.clo = l; }      // 32 <= L < 64
                                else { E.chi = (E.chi << (2 * L)) | (E.clo >> (64 - 2 * L)); E.clo = (E.clo << (2 * L)) | l; }
                                E.nacc = (E.nacc << L) | nrun;
                                rs_acc <<= L;
                            }
                            if (pending_rs) { rs_acc |= 1ull << (L - 1); pending_rs = false; seen_rs = true; have_run = 0; }
                            have_run += L;
                            E.n += L;
                            len += L;
                        }
                    } else if (FASTQ && kind == 3) {
                        if (!have_q) { qbad = lazy_qbad_mask(word); have_q = true; }
                        if (seg & qbad) layout |= 1;
                        qlen += __popcll(seg);
                    } else if (FASTQ && kind == 2) {
                        if (__popcll(seg) > (at_ls ? 1 : 0)) layout |= 1;  // "+title" repeats go to the host tokeniser
                    }
                    if (e >= 64) { prev = byte_at(63); break; }
                    m &= m - 1;
                    pos = e + 1;
                    at_ls = true;
                    prev = '\n';
                    if (pos >= 64) break;
                }
            }
            if (slow) {
                // ---- byte-wise path from `pos`: anything the masks do not describe ------------------
                if (!walked) walk();
                const uint32_t have0 = T.have, nmask0 = T.nmask, imask0 = T.imask;
                if (E.n && finalize_w(E, Y, rs_acc, have0, nmask0, imask0, k, skip_n)) { const unsigned long long key = rec0 * 4 + P_IUPAC; err_key = key < err_key ? key : err_key; }
                if (pos > 0 || E.n) {                                   // window state after the gathered letters
                    const bool carry = !seen_rs && !pending_rs && E.n < 32;
                    const uint32_t hv = seen_rs ? have_run : have0 + E.n;
                    T.have = hv < (uint32_t)k ? hv : (uint32_t)k;
                    T.nmask = (uint32_t)(E.nacc | (carry ? ((unsigned long long)nmask0 << E.n) : 0ull)) & kmask & (uint32_t)ones64(T.have);
                    T.imask = (uint32_t)(carry ? ((unsigned long long)imask0 << E.n) : 0ull) & kmask & (uint32_t)ones64(T.have);
                    if (pending_rs) { T.have = 0; T.nmask = T.imask = 0; }
                }
                for (uint32_t j = pos; j < (uint32_t)lim; j++) {
                    const uint8_t ch = byte_at(j);
                    const bool at_ls = prev == '\n';
                    if (at_ls) begin_line(ch);
                    else if (prev == '\r' && ch != '\n') layout |= 1;
                    prev = ch;
                    if (ch & 0x80) layout |= 1;
                    if (kind == 1) {
                        const uint8_t cl = s_lut[ch];
                        if (cl <= C_ILLEGAL) { if (cl == C_ILLEGAL) Y.err |= 1u << P_ALPHABET; len++; track_letter<true>(T, E, Y, cl, k, kmask, skip_n); }
                        else if (cl == C_WEIRD || (FASTQ && cl == C_SPACE)) layout |= 1;
                    } else if (FASTQ && kind == 3) {
                        const uint8_t cl = s_lut[ch];
                        if (cl <= C_ILLEGAL) qlen++; else if (cl == C_SPACE || cl == C_WEIRD) layout |= 1;
                    } else if (FASTQ && kind == 2) {
                        if (!at_ls && ch != '\n' && ch != '\r') layout |= 1;
                    }
                }
            }
            // ---- what the chunks after me need to know about the record open at my end ------------
            {
                const uint32_t since = slow ? T.have : (pending_rs ? 0u : have_run);
                const uint32_t cnt = min(15u, rs_in_chunk ? since : E.n);
                tail.codes = (uint32_t)E.clo & (uint32_t)ones64(2 * cnt);
                tail.ni = ((uint32_t)E.nacc & (uint32_t)ones64(cnt)) | (slow ? (T.imask & (uint32_t)ones64(cnt)) << 16 : 0u);
                tail.meta = cnt | (rs_in_chunk ? 0x100u : 0u);
                tail.letters = E.n;
            }
            finalized = slow || E.n == 0;
            flush();
            if (a + lim == P.n && prev == '\r') layout |= 1;              // file ends in a lone CR
        }

        // ---- CTA scan of the chunk tails: start state of every chunk + output offsets -------------
        if (tid == 0) s_seed = seed;
        TailAgg tile_total;
        TailAgg start = tail_scan_exclusive(tail, s_tail, &tile_total);      // (syncs: s_seed is visible after it)
        const uint32_t off = start.letters, tile_letters = tile_total.letters;
        if (!finalized) {
            if (tid) start = tail_combine(s_seed, start); else start = s_seed;
            const uint32_t have0 = start.meta & 0xFF, nmask0 = start.ni & 0xFFFF, imask0 = start.ni >> 16;
            if (finalize_w(E, Y, rs_acc, have0, nmask0, imask0, k, skip_n)) { const unsigned long long key = rec0 * 4 + P_IUPAC; err_key = key < err_key ? key : err_key; }
        }
        if (E.n) {
            // left-align the right-aligned accumulators, then OR them in at my letter offset
            const uint32_t sh = 128 - 2 * E.n;                      // 0..126
            uint32_t r[4] = {(uint32_t)(E.chi >> 32), (uint32_t)E.chi, (uint32_t)(E.clo >> 32), (uint32_t)E.clo};
            if (sh & 64) { r[0] = r[2]; r[1] = r[3]; r[2] = 0; r[3] = 0; }
            if (sh & 32) { r[0] = r[1]; r[1] = r[2]; r[2] = r[3]; r[3] = 0; }
            const uint32_t c5 = sh & 31;
            uint32_t x[4];
            x[0] = __funnelshift_l(r[1], r[0], c5); x[1] = __funnelshift_l(r[2], r[1], c5);
            x[2] = __funnelshift_l(r[3], r[2], c5); x[3] = r[3] << c5;
            or_bits<4>(s_out, 2 * off, x, 2 * E.n);
            const uint32_t sh1 = 64 - E.n;                          // 0..63
            uint32_t y[2], z[2];
            {
                const unsigned long long wl = sh1 < 64 ? E.wacc << sh1 : 0ull, nl = sh1 < 64 ? E.nacc << sh1 : 0ull;
                y[0] = (uint32_t)(wl >> 32); y[1] = (uint32_t)wl; z[0] = (uint32_t)(nl >> 32); z[1] = (uint32_t)nl;
            }
            or_bits<2>(s_out + kPackCodes, off, y, E.n);
            or_bits<2>(s_out + kPackCodes + kPackBits, off, z, E.n);
        }
        __syncthreads();
        // ---- write the packed region: only the words the letters of this tile occupy ------------
        {
            uint32_t* dst = P.packed + (size_t)tile * kPackStride;
            if (tid == 0) dst[0] = tile_letters;
            const uint32_t cw = (tile_letters + 15) >> 4, bw = (tile_letters + 31) >> 5;
            // (the words are cleared as they leave: the next tile ORs into them after its own scans' barriers)
            for (uint32_t i = tid; i < cw; i += kThreads) { dst[kPackHdr + i] = s_out[i]; s_out[i] = 0; }
            for (uint32_t i = tid; i < bw; i += kThreads) {
                dst[kPackHdr + kPackCodes + i] = s_out[kPackCodes + i]; s_out[kPackCodes + i] = 0;
                dst[kPackHdr + kPackCodes + kPackBits + i] = s_out[kPackCodes + kPackBits + i]; s_out[kPackCodes + kPackBits + i] = 0;
            }
        }
    }

    // ---- CTA reduction of the fused scalars --------------------------------------------------------
    unsigned long long v[8] = {0, letters_total, Y.mono[0], Y.mono[1], Y.mono[2], Y.mono[3], Y.nwin, 0};
#pragma unroll
    for (int i = 1; i < 7; i++)
        for (int d = 16; d; d >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], d);
    for (int d = 16; d; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, err_key, d);
        err_key = o < err_key ? o : err_key;
        layout |= __shfl_xor_sync(0xffffffffu, layout, d);
    }
    __syncthreads();
    if (lane == 0) { for (int i = 1; i < 7; i++) s_red[warp][i] = v[i]; s_red[warp][7] = err_key; }
    __syncthreads();
    if (tid >= 1 && tid < 7) {
        unsigned long long s = 0;
        for (int wp = 0; wp < kThreads / 32; wp++) s += s_red[wp][tid];
        unsigned long long* dst = tid == 1 ? &P.acc->sum_len : tid <= 5 ? &P.acc->mono[tid - 2] : &P.acc->n_windows_with_n;
        if (s) atomicAdd(dst, s);
    } else if (tid == 7) {
        unsigned long long m = ~0ull;
        for (int wp = 0; wp < kThreads / 32; wp++) m = s_red[wp][7] < m ? s_red[wp][7] : m;
        if (m != ~0ull) atomicMin(&P.acc->err_key, m);
    }
    if (layout && lane == 0) atomicOr(&P.acc->layout_flag, 1u);
}

}  // namespace kdb
