// K2 helpers — tokenise + 2-bit encode (FASTA / FASTQ text -> packed letter stream), sm_100a.  The kernel is k2s_encode (kdb_encode2.cuh).
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

// The chunk masks in two halves: the line-feed mask first (all the line scan needs), the letter
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

}  // namespace kdb
