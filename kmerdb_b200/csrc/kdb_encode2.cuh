// K2s — single-pass tokenise + 2-bit encode (FASTA / FASTQ text -> packed letter stream), sm_100a.
//
// Same output as k2_encode (kdb_encode.cuh: packed codes / W bits / NN bits per 16 KiB tile; kmerdb/kmer.py:517-534,
// kmerdb/parse.py:128-137) but the text is read ONCE (kmerdb/parse.py:70-72 is one tokenising pass too):
//   * tiles are taken in ticket order (atomicAdd on a global counter), so every predecessor of a tile is running or done;
//   * the line / title-line prefix a tile needs (which kind of line it starts in, which record) comes from a decoupled
//     look-back over per-tile status words (one 64-bit word each, written and read whole: no fences): a tile publishes its own
//     aggregate right after its cheap line scan and waits only for published aggregates / prefixes of its predecessors — no
//     summary kernel, no second read of the text, no host synchronisation in the middle of a pass;
//   * record lengths (min / max read length, the short-record raise of kmer.py:461-463, FASTQ sequence/quality agreement)
//     ride on the CTA scan of the chunk tails: a record that closes inside the tile is checked on the spot, the two records
//     that cross the tile borders are left in a TileRec for k4_stitch_records — no per-record array, no global atomics;
//   * the text tile is staged by TMA: one elected thread issues cp.async.bulk.tensor.2d into the double buffer, completion
//     on an mbarrier; the tensor map views the text as [n / 64][64 B] with CU_TENSOR_MAP_SWIZZLE_64B, which is exactly the
//     conflict-free LDS.128 layout (16-byte group q of chunk t at q ^ ((t >> 1) & 3)).
// Tried and rejected (r02g / r02h, measured on the same workload, encode ms per 10 M reads; this kernel: 3.38): a warp-
// specialised CTA — a ninth warp owning tickets, TMA, the tile's line aggregate and the look-back one tile ahead, eight
// consumer warps on a named barrier with two barriers per tile.  Correct (all parity tests), but the producer's chain
// (ticket atomic -> TMA -> 16 KiB aggregate by ONE warp -> look-back) is longer than a tile: 4.51 ms with two text buffers
// (consumers 18 % of their samples in the ready-wait), 7.46 ms with three (tickets two tiles ahead: the early-ticket wait is back).
#pragma once
#include <cuda.h>

#include "kdb_common.cuh"
#include "kdb_encode.cuh"
#include "kdb_scan.cuh"

namespace kdb {

struct EncodeParams2 {
    const uint8_t* text;              // first byte of the pass (feed base + lo), 16-byte aligned
    unsigned long long lo, hi, n;     // byte range of the pass inside the feed, feed length
    const uint8_t* feed;              // feed base (halo walks read back before the pass)
    unsigned int n_tiles;
    int k, skip_n, k_halo, short_k;
    uint32_t* packed;                 // n_tiles * kPackStride words
    TileLook* look;                   // n_tiles status words, zeroed before the pass
    TileRec* trec;                    // n_tiles
    unsigned int* ticket;             // zeroed before the pass
    const PassCarry* carry_in;
    PassCarry* carry_out;
    Accum* acc;
};

// ---- PTX helpers: mbarrier + TMA ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
// one 2-D box of the tensor map -> shared memory, completion counted in bytes on `bar`
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar) : "memory");
}

// ---- chunk tails with the open-record lengths -------------------------------------------------------------------
struct TailAgg2 {
    uint32_t codes;    // 2-bit codes of the last (up to 15) letters, most recent letter in the low bits
    uint32_t ni;       // N flags (low 16) and RYSWKMBDHV/illegal flags (high 16), most recent letter at bit 0
    uint32_t meta;     // letters held (0..15) | window state restarts << 8 (record start in the span, or the tile seed) | record start in the span << 9
    uint32_t letters;  // all letters emitted by the span (output offsets)
    uint32_t olen;     // letters of the record open at the end of the span (since its start, or since the span began)
    uint32_t oqlen;    // FASTQ: its quality characters
};
template <bool FASTQ>
__device__ __forceinline__ TailAgg2 tail2_combine(const TailAgg2 a, const TailAgg2 b) {
    TailAgg2 r;
    r.letters = a.letters + b.letters;
    const bool brs = (b.meta & 0x200u) != 0;
    r.olen = brs ? b.olen : a.olen + b.olen;
    r.oqlen = FASTQ ? (brs ? b.oqlen : a.oqlen + b.oqlen) : 0u;
    if (b.meta & 0x100u) { r.codes = b.codes; r.ni = b.ni; r.meta = b.meta | (a.meta & 0x200u); return r; }
    const uint32_t bc = b.meta & 0xFF, ac = a.meta & 0xFF;
    const uint32_t tot = ac + bc < 15 ? ac + bc : 15;
    r.codes = ((a.codes << (2 * bc)) | b.codes) & 0x3FFFFFFFu;
    const uint32_t an = a.ni & 0xFFFF, ai = a.ni >> 16, bn = b.ni & 0xFFFF, bi = b.ni >> 16;
    r.ni = (((an << bc) | bn) & 0x7FFFu) | ((((ai << bc) | bi) & 0x7FFFu) << 16);
    r.meta = tot | (a.meta & 0x300u) | (b.meta & 0x200u);
    return r;
}
template <bool FASTQ>
__device__ __forceinline__ TailAgg2 tail2_shfl_up(const TailAgg2 v, int d) {
    TailAgg2 r;
    r.codes = __shfl_up_sync(0xffffffffu, v.codes, d); r.ni = __shfl_up_sync(0xffffffffu, v.ni, d);
    r.meta = __shfl_up_sync(0xffffffffu, v.meta, d); r.letters = __shfl_up_sync(0xffffffffu, v.letters, d);
    r.olen = __shfl_up_sync(0xffffffffu, v.olen, d);
    r.oqlen = FASTQ ? __shfl_up_sync(0xffffffffu, v.oqlen, d) : 0u;
    return r;
}
// Exclusive scan over the CTA; *total receives the CTA aggregate.  sm: kThreads/32 entries.  One barrier: every warp folds the
// (<= 8) warp aggregates below it itself.
template <bool FASTQ>
__device__ __forceinline__ TailAgg2 tail2_scan_exclusive(TailAgg2 v, TailAgg2* sm, TailAgg2* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const TailAgg2 ident{0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const TailAgg2 o = tail2_shfl_up<FASTQ>(v, d); if (lane >= d) v = tail2_combine<FASTQ>(o, v); }
    if (lane == 31) sm[warp] = v;
    __syncthreads();
    TailAgg2 pre = ident, tot = ident;
#pragma unroll
    for (int w = 0; w < kThreads / 32; w++) {
        const TailAgg2 t = sm[w];
        if (w < warp) pre = w ? tail2_combine<FASTQ>(pre, t) : t;
        tot = w ? tail2_combine<FASTQ>(tot, t) : t;
    }
    *total = tot;
    TailAgg2 e = tail2_shfl_up<FASTQ>(v, 1);
    if (lane == 0) return pre;
    return warp ? tail2_combine<FASTQ>(pre, e) : e;
}

template <bool FASTQ>
// (launch bounds: 4 CTAs per SM at 64 registers with 52 bytes of spills; 3 CTAs at 80 registers measured 3.27 ms against 3.19)
__global__ void __launch_bounds__(kThreads, 4) k2s_encode(const EncodeParams2 P, const __grid_constant__ CUtensorMap tmap) {
    __shared__ __align__(1024) uint32_t s_text2[2][kTile / 4];     // TMA destinations (64-byte swizzle: 512-byte pattern)
    __shared__ uint32_t s_out[kPackCodes + 2 * kPackBits];
    __shared__ uint8_t s_lut[256];
    __shared__ LineAgg s_scan[kThreads / 32];
    __shared__ TailAgg2 s_tail[kThreads / 32];
    __shared__ TailAgg2 s_seed;
    __shared__ unsigned long long s_red[kThreads / 32][10];
    __shared__ __align__(8) unsigned long long s_bar[2];
    __shared__ unsigned int s_next[2];
    __shared__ unsigned long long s_head_rec, s_carry_ls, s_carry_hdr, s_grp_sum[kThreads / 32];
    __shared__ unsigned int s_head[3], s_carry_last, s_grp_flag[kThreads / 32];
    __shared__ unsigned int s_prev[6];                               // the tile whose packed words still wait in s_out: index, letters, olen, oqlen, meta, valid

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = P.k, skip_n = P.skip_n;
    const uint32_t kmask = (1u << k) - 1u;
    s_lut[tid] = c_class_lut[tid];

    Tally Y;
    Y.mono[0] = Y.mono[1] = Y.mono[2] = Y.mono[3] = 0; Y.nwin = 0; Y.err = 0;
    unsigned long long letters_total = 0, err_key = ~0ull;
    unsigned int layout = 0, mn_len = 0xFFFFFFFFu, mx_len = 0, any_rec = 0;
    const uint32_t* s_text = s_text2[0];
    const uint32_t swz = ((uint32_t)tid >> 1) & 3u;
    auto word = [&](int i) -> uint32_t { return s_text[16 * tid + 4 * (((uint32_t)i >> 2) ^ swz) + (i & 3)]; };
    auto quad = [&](int q) -> uint4 { return *reinterpret_cast<const uint4*>(s_text + 16 * tid + 4 * ((uint32_t)q ^ swz)); };
    auto byte_at = [&](uint32_t p) -> uint8_t { return (uint8_t)(s_text[16 * tid + 4 * ((p >> 4) ^ swz) + ((p >> 2) & 3)] >> (8 * (p & 3))); };
    const unsigned long long fq_base = P.carry_in->fq_rec_base;
    const uint32_t bar0 = smem_u32(&s_bar[0]);
    // the whole tile in one TMA request; rows past the end of the pass arrive as zeros
    auto stage = [&](unsigned int tile, int b) {
        const uint32_t bar = bar0 + 8u * (uint32_t)b;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(bar, kTile);
        tma_load_2d(smem_u32(&s_text2[b][0]), &tmap, 0, (int)(tile * (unsigned int)kThreads), bar);
    };

    // The next tile's index travels with its text: thread 0 writes s_next[slot] BEFORE it arrives on the slot's mbarrier
    // (release), every thread reads it AFTER the wait on that mbarrier (acquire) — no CTA barrier is spent on the hand-over.
    // Past the last tile nothing is loaded: a plain arrive completes the phase so that the waiters see the end.
    auto take_ticket = [&](unsigned int slot, int b) {
        const unsigned int nxt = atomicAdd(P.ticket, 1u);
        s_next[slot] = nxt;
        if (nxt < P.n_tiles) stage(nxt, b);
        else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8u * (uint32_t)b) : "memory");
    };
    if (tid == 0) {
        mbar_init(bar0, 1); mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        s_carry_ls = P.carry_in->ls_total; s_carry_hdr = P.carry_in->hdr_total; s_carry_last = P.carry_in->last;
        s_prev[5] = 0;
    }
    for (int i = tid; i < kPackCodes + 2 * kPackBits; i += kThreads) s_out[i] = 0;
    __syncthreads();
    if (tid == 0) take_ticket(0, 0);
    // The packed words of a tile leave shared memory one iteration later, while warp 0 looks back for the next tile
    // (see below); after the loop the last tile is flushed by everybody.
    auto copy_out = [&](int first_thread, int n_threads) {
        const uint32_t letters = s_prev[1];
        uint32_t* dst = P.packed + (size_t)s_prev[0] * kPackStride;
        const int t = tid - first_thread;
        if (t == 0) {
            dst[0] = letters;
            TileRec tr;
            const bool has_b = (s_prev[4] & 0x200u) != 0;
            tr.head_len = has_b ? s_head[0] : s_prev[2]; tr.head_qlen = has_b ? s_head[1] : s_prev[3];
            tr.tail_len = s_prev[2]; tr.tail_qlen = s_prev[3];
            tr.flags = has_b ? s_head[2] : 0u; tr.pad = 0; tr.head_rec = has_b ? s_head_rec : 0ull;
            P.trec[s_prev[0]] = tr;
        }
        const uint32_t cw = (letters + 15) >> 4, bw = (letters + 31) >> 5;
        // (the words are cleared as they leave: the tile in work ORs into them after two more barriers)
        for (uint32_t i = t; i < cw; i += n_threads) { dst[kPackHdr + i] = s_out[i]; s_out[i] = 0; }
        for (uint32_t i = t; i < bw; i += n_threads) {
            dst[kPackHdr + kPackCodes + i] = s_out[kPackCodes + i]; s_out[kPackCodes + i] = 0;
            dst[kPackHdr + kPackCodes + kPackBits + i] = s_out[kPackCodes + kPackBits + i]; s_out[kPackCodes + kPackBits + i] = 0;
        }
    };
    int buf = 0;
    for (unsigned int it = 0;; it++, buf ^= 1) {
        mbar_wait(bar0 + 8u * (uint32_t)buf, (it >> 1) & 1u);
        const unsigned int tile = s_next[it & 1];
        if (tile >= P.n_tiles) break;
        const unsigned long long tile0 = P.lo + (unsigned long long)tile * kTile;
        s_text = s_text2[buf];

        const unsigned long long a = tile0 + (unsigned long long)tid * kChunk;
        const int lim = a >= P.hi ? 0 : (P.hi - a >= (unsigned long long)kChunk ? kChunk : (int)(P.hi - a));
        if (lim > 0 && lim < kChunk) {
            // the ragged last row of the feed is not part of the tensor (zeros arrived): its owner copies it by hand
            uint32_t* mine = s_text2[buf] + 16 * tid;
            for (int j = 0; j < lim; j++) {
                uint32_t* w = mine + 4 * (((uint32_t)j >> 4) ^ swz) + ((j >> 2) & 3);
                *w |= (uint32_t)P.feed[a + j] << (8 * (j & 3));
            }
        }
        uint8_t prev = '\n';
        if (lim > 0 && a > 0) {
            // last byte of the chunk before mine: my neighbour's chunk in shared memory, or the text before the tile
            if (tid > 0) { const uint32_t sz = ((uint32_t)(tid - 1) >> 1) & 3u; prev = (uint8_t)(s_text[16 * (tid - 1) + 4 * (3u ^ sz) + 3] >> 24); }
            else prev = P.feed[a - 1];
        }

        // ---- phase A: bit masks of my chunk; phase B: line starts -> CTA scan ---------------------
        uint32_t any_cr, any_hi;
        const unsigned long long lfm = chunk_lf(quad, &any_cr, &any_hi);      // (zero fill beyond lim is neither LF nor a letter)
        LineAgg g{0, 0, 0};
        if (lim > 0) {
            unsigned long long lsm = ((lfm << 1) | (prev == '\n' ? 1ull : 0ull)) & ones64((uint32_t)lim);
            g.ls = __popcll(lsm);
            if (!FASTQ && lsm) {
                const uint32_t top = 63 - __clzll(lsm);
                g.last = byte_at(top) == '>' ? 2u : 1u;
                while (lsm) { const uint32_t p = __ffsll(lsm) - 1; lsm &= lsm - 1; g.hdr += byte_at(p) == '>' ? 1u : 0u; }
            } else if (lsm) g.last = 1u;
        }
        LineAgg inc = g, warp_pre{0, 0, 0}, tot{0, 0, 0};
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const LineAgg o = shfl_up_agg(inc, d); if (lane >= d) inc = combine(o, inc); }
        if (lane == 31) s_scan[warp] = inc;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < kThreads / 32; w++) { const LineAgg t = s_scan[w]; if (w < warp) warp_pre = combine(warp_pre, t); tot = combine(tot, t); }
        inc = combine(warp_pre, inc);
        // The tile's aggregate goes out NOW, and every thread asks for the status word of ONE predecessor (thread t: tile - 1 - t:
        // a window of 256 tiles per round, the load issued here and looked at later).  The letter masks and codes of my chunk
        // (the larger half of phase A) and the copy-out of the tile before fill the time until the answers are in; the
        // predecessors — which took their tickets moments before this tile and are at the same point of their own iteration —
        // have published by then.
        if (tid == 0)
            *reinterpret_cast<volatile unsigned long long*>(&P.look[tile].w0) =
                (1ull << 62) | ((unsigned long long)tot.last << 60) | ((unsigned long long)tot.ls << 30) | (unsigned long long)tot.hdr;
        long long pidx = (long long)tile - 1 - tid;
        unsigned long long pw = 2ull << 62;                                  // before tile 0 of the pass: an empty prefix
        if (pidx >= 0) pw = *reinterpret_cast<const volatile unsigned long long*>(&P.look[pidx].w0);
        // (tried for FASTQ, where half of the chunks lie in quality / title lines: building the letter masks on demand inside the
        // segment walk — 9.23 ms per 10 M reads against 7.91 with the masks built here for every chunk: divergent, and more spills)
        const ChunkMasks M = chunk_masks_rest(quad, lfm, any_cr, any_hi);
        if (s_prev[5]) copy_out(0, kThreads);                                // the packed words of the tile before leave

        // ---- decoupled look-back, all eight warps: what came before this tile ------------------------------------------
        // Warp w holds predecessors [32 w, 32 w + 32) of the window.  The sum runs back to the nearest tile that already holds an
        // inclusive prefix; everything nearer contributes its aggregate.  All of them hold earlier tickets, so they are running
        // or done: the spin cannot starve.  One round (256 tiles) is nearly always enough.
        unsigned long long ex = 0;                                           // [59:30] line starts, [29:0] title lines before this tile, in this pass
        unsigned int ex_last = 0;
        while (true) {
            if (pidx >= 0) {
                const volatile unsigned long long* src = reinterpret_cast<const volatile unsigned long long*>(&P.look[pidx].w0);
                while ((pw >> 62) == 0) pw = *src;                           // that predecessor has not published yet
            }
            __syncwarp();
            const unsigned int m2 = __ballot_sync(0xffffffffu, (pw >> 62) == 2);
            const int first = m2 ? __ffs(m2) - 1 : 32;                       // nearest tile of my group that already holds an inclusive prefix
            const bool use = lane <= first;
            const unsigned int last = use ? (unsigned int)(pw >> 60) & 3u : 0u;
            const unsigned int ml = __ballot_sync(0xffffffffu, last != 0);
            const unsigned int glast = ml ? __shfl_sync(0xffffffffu, last, __ffs(ml) - 1) : 0u;
            unsigned long long mine = use ? pw & ((1ull << 60) - 1ull) : 0ull;
#pragma unroll
            for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
            if (lane == 0) { s_grp_sum[warp] = mine; s_grp_flag[warp] = glast | (m2 ? 4u : 0u); }
            __syncthreads();
            bool found = false;
#pragma unroll
            for (int w = 0; w < kThreads / 32; w++) {
                if (found) break;
                const unsigned int f = s_grp_flag[w];
                ex += s_grp_sum[w];
                if (ex_last == 0) ex_last = f & 3u;
                found = (f & 4u) != 0;
            }
            if (found) break;
            __syncthreads();                                                 // (rare: more than 256 predecessors without a prefix)
            pidx -= kThreads;
            pw = 2ull << 62;
            if (pidx >= 0) pw = *reinterpret_cast<const volatile unsigned long long*>(&P.look[pidx].w0);
        }
        const unsigned long long c_ls = s_carry_ls, c_hdr = s_carry_hdr;
        const unsigned int c_last = s_carry_last;
        if (tid == 0) {
            const unsigned long long in = ex + (((unsigned long long)tot.ls << 30) | (unsigned long long)tot.hdr);
            const unsigned int in_last = tot.last ? tot.last : ex_last;
            *reinterpret_cast<volatile unsigned long long*>(&P.look[tile].w0) = (2ull << 62) | ((unsigned long long)in_last << 60) | in;
            if (tile == P.n_tiles - 1) {
                P.carry_out->ls_total = c_ls + (in >> 30); P.carry_out->hdr_total = c_hdr + (in & 0x3FFFFFFFull);
                P.carry_out->last = in_last ? in_last : c_last; P.carry_out->fq_rec_base = fq_base;
            }
        }
        const unsigned long long s_pre_ls = c_ls + (ex >> 30), s_pre_hdr = c_hdr + (ex & 0x3FFFFFFFull);
        const unsigned int s_pre_last = ex_last ? ex_last : c_last;
        const unsigned long long ls_before = s_pre_ls + (inc.ls - g.ls);
        unsigned long long hdr_before = s_pre_hdr + (inc.hdr - g.hdr);
        unsigned int last_before;
        {
            LineAgg e = shfl_up_agg(inc, 1);
            if (lane == 0) e = warp_pre;
            last_before = e.last ? e.last : s_pre_last;
        }

        Emit E; E.chi = E.clo = E.wacc = E.nacc = 0; E.n = 0;
        unsigned long long rs_acc = 0, rec0 = 0;        // record-start marks (accumulator orientation); record open at `a`
        bool finalized = true;                          // W bits still to be derived after the tail scan?
        TailAgg2 tail{0, 0, 0, 0, 0, 0}, seed{0, 0, 0x100, 0, 0, 0};
        // records: the one open when my chunk began is closed by my first record start (first_*); later ones are mine alone
        uint32_t first_len = 0, first_qlen = 0;
        bool seen_b = false, first_open = false;
        unsigned long long first_rec = 0;
        auto check_record = [&](unsigned long long rec_id, uint32_t l, uint32_t q) {       // kmerdb/kmer.py:461-463, parse.py:142-146
            mn_len = l < mn_len ? l : mn_len; mx_len = l > mx_len ? l : mx_len; any_rec = 1;
            if (l < (uint32_t)P.short_k) { const unsigned long long key = rec_id * 4 + P_SHORT; err_key = key < err_key ? key : err_key; }
            if (FASTQ && q != l) layout |= 1;                                               // Bio.SeqIO raises ValueError: host tokeniser
        };
        if (lim > 0) {
            unsigned long long line = ls_before;      // FASTQ: index of the next line start
            uint32_t kind;
            if (FASTQ) kind = ls_before ? (uint32_t)((ls_before - 1) & 3) : K_HEADER;
            else kind = last_before == 2 ? K_HEADER : (hdr_before ? K_SEQ : K_IGNORE);
            unsigned long long rec = FASTQ ? fq_base + (ls_before ? (ls_before - 1) >> 2 : 0) : (hdr_before ? hdr_before - 1 : 0);
            Track T; T.nmask = T.imask = T.have = 0;
            bool need_halo;
            if (FASTQ) need_halo = prev != '\n' && kind == 1;
            else if (prev != '\n') need_halo = kind == K_SEQ;
            else need_halo = hdr_before != 0 && byte_at(0) != '>';
            uint32_t halo_codes = 0, halo_nn = 0, halo_w = 0;
            bool walked = false;
            rec0 = rec;
            auto walk = [&]() {
                walked = true;
                if (!need_halo) return;
                uint8_t hc[16];
                const int cnt = halo_classes<FASTQ>(P.feed, a, P.k_halo, hc);
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
                seed.codes = halo_codes; seed.ni = T.nmask | (T.imask << 16); seed.meta = T.have | 0x100;
            }
            uint32_t len = 0, qlen = 0;
            Y.err = 0;

            // at a record start (boundary) and at the end of my chunk
            auto flush = [&](bool boundary, bool open) {
                letters_total += len;
                if (Y.err) {
                    const unsigned long long key = rec * 4 + ((Y.err & 1u) ? P_ALPHABET : P_IUPAC);
                    err_key = key < err_key ? key : err_key;
                    Y.err = 0;
                }
                if (boundary) {
                    if (!seen_b) { first_len = len; first_qlen = qlen; first_open = open; first_rec = rec; seen_b = true; }
                    else if (open) check_record(rec, len, qlen);
                    len = 0; qlen = 0;
                }
            };
            auto begin_line = [&](uint8_t ch) {
                if (FASTQ) {
                    kind = (uint32_t)(line & 3);
                    if (kind == 0) { flush(true, line != 0); rec = fq_base + (line >> 2); T.have = 0; T.nmask = T.imask = 0; if (ch != '@') layout |= 1; }
                    else if (kind == 2 && ch != '+') layout |= 1;
                    line++;
                } else {
                    if (ch == '>') { flush(true, hdr_before != 0); hdr_before++; rec = hdr_before - 1; kind = K_HEADER; T.have = 0; T.nmask = T.imask = 0; }
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
                        const unsigned long long was = FASTQ ? 0 : hdr_before;
                        begin_line(ch);
                        if (FASTQ ? kind == 0 : (hdr_before != was)) pending_rs = true;
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
                const uint32_t cnt = min(15u, seen_b ? since : E.n);
                tail.codes = (uint32_t)E.clo & (uint32_t)ones64(2 * cnt);
                tail.ni = ((uint32_t)E.nacc & (uint32_t)ones64(cnt)) | (slow ? (T.imask & (uint32_t)ones64(cnt)) << 16 : 0u);
                tail.meta = cnt | (seen_b ? 0x300u : 0u);
                tail.letters = E.n;
                tail.olen = len; tail.oqlen = qlen;
            }
            finalized = slow || E.n == 0;
            flush(false, false);
            if (a + lim == P.n && prev == '\r') layout |= 1;              // file ends in a lone CR
        }

        // ---- CTA scan of the chunk tails: start state of every chunk + output offsets -------------
        if (tid == 0) s_seed = seed;
        TailAgg2 tile_total;
        TailAgg2 start = tail2_scan_exclusive<FASTQ>(tail, s_tail, &tile_total);      // (syncs: s_seed is visible after it)
        // ---- stage: thread 0 takes the next ticket and starts its TMA load into the other buffer (the barrier of the tail
        // scan is behind us: nobody reads text any more in this iteration).  The ticket is taken LATE — one copy-out before the tile is worked on, not a whole tile earlier: a
        // ticket holder publishes its aggregate a few microseconds after taking the ticket, so a tile rarely finds a predecessor
        // that has not published yet (with the ticket taken a tile ahead, 42 % of the kernel's samples sat in that wait). -----------
        if (tid == 0) {
            take_ticket((it + 1) & 1, buf ^ 1);
            s_prev[0] = tile; s_prev[1] = tile_total.letters; s_prev[2] = tile_total.olen; s_prev[3] = tile_total.oqlen; s_prev[4] = tile_total.meta; s_prev[5] = 1;
        }
        const uint32_t off = start.letters;
        if (seen_b) {
            // my first record start closes the record that was open when my chunk began
            const uint32_t cl = start.olen + first_len, cq = start.oqlen + first_qlen;
            if (start.meta & 0x200u) { if (first_open) check_record(first_rec, cl, cq); }
            else { s_head[0] = cl; s_head[1] = cq; s_head[2] = first_open ? 3u : 1u; s_head_rec = first_rec; }   // ... which began before this tile
        }
        if (!finalized) {
            if (tid) start = tail2_combine<FASTQ>(s_seed, start); else start = s_seed;
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
    }
    __syncthreads();
    if (s_prev[5]) copy_out(0, kThreads);

    // ---- CTA reduction of the fused scalars --------------------------------------------------------
    unsigned long long v[8] = {0, letters_total, Y.mono[0], Y.mono[1], Y.mono[2], Y.mono[3], Y.nwin, 0};
#pragma unroll
    for (int i = 1; i < 7; i++)
        for (int d = 16; d; d >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], d);
    for (int d = 16; d; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, err_key, d);
        err_key = o < err_key ? o : err_key;
        layout |= __shfl_xor_sync(0xffffffffu, layout, d);
        const unsigned int o1 = __shfl_xor_sync(0xffffffffu, mn_len, d); mn_len = o1 < mn_len ? o1 : mn_len;
        const unsigned int o2 = __shfl_xor_sync(0xffffffffu, mx_len, d); mx_len = o2 > mx_len ? o2 : mx_len;
        any_rec |= __shfl_xor_sync(0xffffffffu, any_rec, d);
    }
    __syncthreads();
    if (lane == 0) {
        for (int i = 1; i < 7; i++) s_red[warp][i] = v[i];
        s_red[warp][7] = err_key; s_red[warp][8] = any_rec ? mn_len : ~0ull; s_red[warp][9] = mx_len;
    }
    __syncthreads();
    if (tid >= 1 && tid < 7) {
        unsigned long long s = 0;
        for (int wp = 0; wp < kThreads / 32; wp++) s += s_red[wp][tid];
        unsigned long long* dst = tid == 1 ? &P.acc->sum_len : tid <= 5 ? &P.acc->mono[tid - 2] : &P.acc->n_windows_with_n;
        if (s) atomicAdd(dst, s);
    } else if (tid == 7 || tid == 8) {
        unsigned long long m = ~0ull;
        for (int wp = 0; wp < kThreads / 32; wp++) m = s_red[wp][tid] < m ? s_red[wp][tid] : m;
        if (m != ~0ull) atomicMin(tid == 7 ? &P.acc->err_key : &P.acc->min_len, m);
    } else if (tid == 9) {
        unsigned long long m = 0;
        for (int wp = 0; wp < kThreads / 32; wp++) m = s_red[wp][9] > m ? s_red[wp][9] : m;
        if (m) atomicMax(&P.acc->max_len, m);
    }
    if (layout && lane == 0) atomicOr(&P.acc->layout_flag, 1u);
}

// ---- the records that cross tile borders ---------------------------------------------------------------------------
// One thread per tile.  The length of the record a tile's first record start closes is  (letters since the last record
// start before the tile) + head_len; the scan operator is (record start seen, tail_len): a tile with a record start
// replaces what came before it, one without adds to it.  Before the first record start of a pass stands the carry of the
// passes before.  The last tile leaves the new carry.
struct RecAgg { unsigned int rs; unsigned long long len, qlen; };
__device__ __forceinline__ RecAgg rec_combine(const RecAgg a, const RecAgg b) {
    RecAgg r;
    r.rs = a.rs | b.rs;
    r.len = b.rs ? b.len : a.len + b.len;
    r.qlen = b.rs ? b.qlen : a.qlen + b.qlen;
    return r;
}
__device__ __forceinline__ RecAgg rec_shfl_up(const RecAgg v, int d) {
    RecAgg r;
    r.rs = __shfl_up_sync(0xffffffffu, v.rs, d); r.len = __shfl_up_sync(0xffffffffu, v.len, d); r.qlen = __shfl_up_sync(0xffffffffu, v.qlen, d);
    return r;
}
constexpr int kStitchThreads = 1024;
template <bool FASTQ>
__global__ void __launch_bounds__(kStitchThreads) k4_stitch_records(const TileRec* __restrict__ trec, unsigned int n_tiles, const PassCarry* carry_in,
                                                                    PassCarry* carry_out, int short_k, Accum* acc) {
    __shared__ RecAgg s_w[kStitchThreads / 32];
    __shared__ RecAgg s_pre;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int seg0 = blockIdx.x * kStitchThreads, t = seg0 + tid;
    TileRec tr{};
    if (t < n_tiles) tr = trec[t];
    RecAgg mine{t < n_tiles ? (tr.flags & 1u) : 0u, tr.tail_len, tr.tail_qlen}, inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const RecAgg o = rec_shfl_up(inc, d); if (lane >= d) inc = rec_combine(o, inc); }
    if (lane == 31) s_w[warp] = inc;
    // what stands before this CTA's segment: walk back to the nearest tile with a record start (almost always the tile before)
    if (tid == 0) {
        RecAgg pre{0, 0, 0};
        long long j = (long long)seg0 - 1;
        unsigned long long l = 0, q = 0;
        while (j >= 0) {
            const TileRec p = trec[j];
            l += p.tail_len; q += p.tail_qlen;
            if (p.flags & 1u) { pre.rs = 1; break; }
            j--;
        }
        if (!pre.rs) { l += carry_in->olen; q += carry_in->oqlen; }
        pre.len = l; pre.qlen = q;
        s_pre = pre;
    }
    __syncthreads();
    RecAgg ex = s_pre;
    for (int w = 0; w < warp; w++) ex = rec_combine(ex, s_w[w]);
    {
        RecAgg e = rec_shfl_up(inc, 1);
        if (lane) ex = rec_combine(ex, e);
    }
    unsigned long long mn = ~0ull, mx = 0, err = ~0ull;
    unsigned int layout = 0;
    if (t < n_tiles && (tr.flags & 2u)) {                         // my first record start closes a record that was open
        const unsigned long long l = ex.len + tr.head_len, q = ex.qlen + tr.head_qlen;
        mn = l; mx = l;
        if (l < (unsigned long long)short_k) err = tr.head_rec * 4 + P_SHORT;
        if (FASTQ && q != l) layout = 1;
    }
    for (int d = 16; d; d >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xffffffffu, mn, d); mn = o < mn ? o : mn;
        o = __shfl_xor_sync(0xffffffffu, mx, d); mx = o > mx ? o : mx;
        o = __shfl_xor_sync(0xffffffffu, err, d); err = o < err ? o : err;
        layout |= __shfl_xor_sync(0xffffffffu, layout, d);
    }
    if (lane == 0) {
        if (mn != ~0ull) { atomicMin(&acc->min_len, mn); atomicMax(&acc->max_len, mx); }
        if (err != ~0ull) atomicMin(&acc->err_key, err);
        if (layout) atomicOr(&acc->layout_flag, 1u);
    }
    if (t == n_tiles - 1) {
        const RecAgg fin = rec_combine(ex, mine);
        carry_out->olen = fin.len; carry_out->oqlen = fin.qlen;
        carry_out->open = (carry_in->open || fin.rs) ? 1u : 0u;
    }
}

// The record still open is complete (end of a FASTQ feed; end of the job): check it.  FASTQ feeds also fold their line count
// into the record base; line indices restart with the next feed.   mode bit 0: close the open record; bit 1: end of a feed.
template <bool FASTQ>
__global__ void k4_close_carry(PassCarry* c, int mode, int short_k, Accum* acc) {
    if (threadIdx.x || blockIdx.x) return;
    if (mode & 2) {
        if (FASTQ) {
            if (c->ls_total & 3ull) atomicOr(&acc->layout_flag, 1u);      // truncated / blank-line tail: the host tokeniser decides
            c->fq_rec_base += c->ls_total >> 2;
        }
        c->ls_total = 0; c->last = 0;
    }
    if ((mode & 1) && c->open) {
        const unsigned long long l = c->olen, rec = FASTQ ? c->fq_rec_base - 1 : c->hdr_total - 1;
        atomicMin(&acc->min_len, l); atomicMax(&acc->max_len, l);
        if (l < (unsigned long long)short_k) atomicMin(&acc->err_key, rec * 4 + P_SHORT);
        if (FASTQ && c->oqlen != l) atomicOr(&acc->layout_flag, 1u);
        c->open = 0; c->olen = 0; c->oqlen = 0;
    }
}

}  // namespace kdb
