// K2+K3 — 2-bit encode, rolling (canonical) k-mer id, dense 4^k histogram, fused reductions.
//
// Replaces kmer.shred (kmerdb/kmer.py:489-577) + kmer.kmer_to_id (:234-317) + the count loop
// of parse.parsefile (kmerdb/parse.py:128-137) for one pass of text.  Persistent CTAs walk the
// tiles of the pass; every thread owns kChunk consecutive text bytes and emits the windows whose
// LAST letter lies in its chunk (the k-1 letters before the chunk are re-read by a short backward
// walk, so there is no inter-thread dependency).
#pragma once
#include "kdb_common.cuh"
#include "kdb_scan.cuh"

namespace kdb {

// Rolling window state of one thread.
struct Win {
    uint32_t fwd, rc;      // 2-bit packed ids, first letter most significant (kmer.py:307-312)
    uint32_t nmask, imask; // bit j set: the letter j positions back is N / RYSWKMBDHV
    uint32_t have;         // letters of the current record seen so far, saturated at k
};

struct Ctl {
    int k, canonical, skip_n;
    uint32_t idmask, kmask, rcshift;
};

template <bool SMEM_HIST>
__device__ __forceinline__ void bump(unsigned int* tab, uint32_t id) {
    atomicAdd(tab + id, 1u);   // result unused -> RED.E.ADD (global) / ATOMS (shared)
}

// kmer.py:550-565, :586-621: every ACGT substitution of the N positions gets +1.
template <bool SMEM_HIST>
__device__ __noinline__ unsigned long long expand_n(unsigned int* tab, const Ctl& c, uint32_t base, uint32_t nmask) {
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
        if (c.canonical) { const uint32_t r = revcomp32(f, c.k); id = r < f ? r : f; }
        bump<SMEM_HIST>(tab, id);
    }
    return combos;
}

// Feed one sequence-line byte class into the window; emit a count when a full window is present.
struct Local {
    unsigned long long kmers;   // counted ids (parse.py:133-136)
    unsigned int mono[4];
    unsigned int nwin;          // windows holding N
    unsigned int err;           // bit0 alphabet, bit2 iupac (priority bits), for the current record
};

template <bool SMEM_HIST, bool EMIT>
__device__ __forceinline__ void push_letter(Win& w, const Ctl& c, uint32_t cl, unsigned int* tab, Local& L) {
    const uint32_t code = cl <= 3 ? cl : 0;
    w.fwd = ((w.fwd << 2) | code) & c.idmask;
    w.rc = (w.rc >> 2) | ((3u - code) << c.rcshift);
    w.nmask = ((w.nmask << 1) | (cl == C_N ? 1u : 0u)) & c.kmask;
    w.imask = ((w.imask << 1) | (cl >= C_IUPAC ? 1u : 0u)) & c.kmask;   // C_ILLEGAL raises on its own
    w.have = w.have < (uint32_t)c.k ? w.have + 1 : w.have;
    if (!EMIT) return;
    if (cl <= 3) L.mono[cl]++;
    if (w.have < (uint32_t)c.k) return;
    if (w.nmask == 0) {
        if (w.imask) { L.err |= 1u << P_IUPAC; return; }
        const uint32_t id = (c.canonical && w.rc < w.fwd) ? w.rc : w.fwd;
        bump<SMEM_HIST>(tab, id);
        L.kmers++;
    } else {
        L.nwin++;
        if (!c.skip_n) {
            if (w.imask) { L.err |= 1u << P_IUPAC; return; }
            L.kmers += expand_n<SMEM_HIST>(tab, c, w.fwd, w.nmask);
        }
    }
}

// Position of the first byte of the line that holds byte p (p itself is not a line feed).
__device__ __forceinline__ unsigned long long line_start(const uint8_t* text, unsigned long long p) {
    while (p > 0 && text[p - 1] != '\n') p--;
    return p;
}

// Re-read up to k-1 letters that precede byte `a` inside the same record and replay them into w.
// FASTA (kmerdb/parse.py:70): the record's lines are joined, so the walk crosses line feeds until
// it meets the title line.  FASTQ: the sequence is one line.
template <bool FASTQ>
__device__ __noinline__ void halo(const uint8_t* text, unsigned long long a, const Ctl& c, Win& w) {
    uint8_t cls[16];
    int cnt = 0;
    unsigned long long p = a;
    while (cnt < c.k - 1 && p > 0) {
        const uint8_t ch = text[p - 1];
        if (ch == '\n') {
            if (FASTQ) break;
            // previous line is [s, p-1): a title line ends the record
            const unsigned long long s = line_start(text, p - 1);
            if (text[s] == '>') break;
            p--;
            continue;
        }
        const uint8_t cl = c_class_lut[ch];
        if (cl <= C_ILLEGAL) cls[cnt++] = cl;
        p--;
    }
    for (int i = cnt - 1; i >= 0; i--) {
        Local dummy;
        push_letter<false, false>(w, c, cls[i], nullptr, dummy);
    }
}

template <bool FASTQ, bool SMEM_HIST>
__global__ void __launch_bounds__(kThreads) k2_count(const CountParams P) {
    extern __shared__ unsigned int s_dyn[];
    __shared__ uint32_t s_text[(kChunk / 4) * (kThreads + 1)];   // transposed: word i of thread t at [i*(T+1)+t]
    __shared__ uint8_t s_lut[256];
    __shared__ LineAgg s_scan[kThreads / 32];
    __shared__ unsigned long long s_red[kThreads / 32][8];

    const int tid = threadIdx.x;
    Ctl c;
    c.k = P.k; c.canonical = P.canonical; c.skip_n = P.skip_n;
    c.idmask = P.k == 16 ? 0xffffffffu : ((1u << (2 * P.k)) - 1u);
    c.kmask = (1u << P.k) - 1u;
    c.rcshift = 2 * (P.k - 1);
    const uint32_t nbins = SMEM_HIST ? 1u << (2 * P.k) : 0u;      // SMEM_HIST only for k <= 7
    unsigned int* tab = SMEM_HIST ? s_dyn : P.tab32;

    s_lut[tid] = c_class_lut[tid];
    if (SMEM_HIST) for (uint32_t i = tid; i < nbins; i += kThreads) s_dyn[i] = 0;
    __syncthreads();

    Local L;
    L.kmers = 0; L.mono[0] = L.mono[1] = L.mono[2] = L.mono[3] = 0; L.nwin = 0; L.err = 0;
    unsigned long long letters_total = 0;
    unsigned int layout = 0;
    unsigned long long err_key = ~0ull;

    for (unsigned int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        const unsigned long long tile0 = P.lo + (unsigned long long)tile * kTile;
        // ---- stage the tile: coalesced 16-byte loads, transposed into shared memory ----------
        __syncthreads();
#pragma unroll
        for (int r = 0; r < kChunk / 16; r++) {
            const int q = tid + kThreads * r;                 // 16-byte unit inside the tile
            const unsigned long long g = tile0 + 16ull * q;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (g + 16 <= P.hi) v = __ldcs(reinterpret_cast<const uint4*>(P.text + g));
            else if (g < P.hi) {
                uint32_t t4[4] = {0, 0, 0, 0};
                for (int j = 0; j < (int)(P.hi - g); j++) t4[j >> 2] |= (uint32_t)P.text[g + j] << (8 * (j & 3));
                v = make_uint4(t4[0], t4[1], t4[2], t4[3]);
            }
            const int owner = q >> 2, w0 = (q & 3) * 4;
            s_text[(w0 + 0) * (kThreads + 1) + owner] = v.x;
            s_text[(w0 + 1) * (kThreads + 1) + owner] = v.y;
            s_text[(w0 + 2) * (kThreads + 1) + owner] = v.z;
            s_text[(w0 + 3) * (kThreads + 1) + owner] = v.w;
        }
        __syncthreads();

        const unsigned long long a = tile0 + (unsigned long long)tid * kChunk;
        const int lim = a >= P.hi ? 0 : (P.hi - a >= (unsigned long long)kChunk ? kChunk : (int)(P.hi - a));
        uint8_t prev = '\n';
        if (lim > 0 && a > 0) prev = tid > 0 ? (uint8_t)(s_text[(kChunk / 4 - 1) * (kThreads + 1) + tid - 1] >> 24) : P.text[a - 1];

        // ---- line starts in my chunk, CTA scan -> line / record index and line kind at `a` ---
        LineAgg g{0, 0, 0};
        {
            uint8_t pv = prev;
            for (int i = 0; i < kChunk / 4; i++) {
                const uint32_t wv = s_text[i * (kThreads + 1) + tid];
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const uint8_t ch = (uint8_t)(wv >> (8 * b));
                    if (4 * i + b < lim && pv == '\n') { g.ls++; const unsigned int cl = ch == '>' ? 2u : 1u; g.hdr += cl >> 1; g.last = cl; }
                    pv = ch;
                }
            }
        }
        LineAgg total;
        const LineAgg inc = block_scan_inclusive(g, s_scan, &total);
        const TileState ts = P.tiles[tile];
        // exclusive prefix: everything before my chunk
        const unsigned long long ls_before = ts.ls_base + (inc.ls - g.ls);
        unsigned long long hdr_before = ts.hdr_base + (inc.hdr - g.hdr);
        unsigned int last_before;
        {
            LineAgg e = shfl_up_agg(inc, 1);
            if ((tid & 31) == 0) e = (tid >> 5) ? s_scan[(tid >> 5) - 1] : LineAgg{0, 0, 0};
            last_before = e.last ? e.last : ts.last;
        }

        if (lim > 0) {
            // ---- state at the first byte of my chunk ----------------------------------------
            unsigned long long line = ls_before;          // FASTQ: index of the NEXT line start; current line = line-1
            uint32_t kind;
            if (FASTQ) kind = ls_before ? (uint32_t)((ls_before - 1) & 3) : K_HEADER;   // 0 title,1 seq,2 plus,3 qual
            else kind = last_before == 2 ? K_HEADER : (hdr_before ? K_SEQ : K_IGNORE);
            // record id of the current line
            unsigned long long rec = FASTQ ? P.hdr_file_base + (ls_before ? (ls_before - 1) >> 2 : 0) : (hdr_before ? hdr_before - 1 : 0);
            Win w; w.fwd = w.rc = w.nmask = w.imask = w.have = 0;
            // Does my first byte continue a sequence whose earlier letters lie before my chunk?
            bool need_halo;
            if (FASTQ) need_halo = prev != '\n' && kind == 1;
            else if (prev != '\n') need_halo = kind == K_SEQ;
            else need_halo = hdr_before != 0 && (uint8_t)s_text[tid] != '>';   // FASTA continuation line
            if (need_halo) halo<FASTQ>(P.text, a, c, w);
            unsigned long long len = 0, qlen = 0;     // letters (quality chars) of `rec` seen by this thread
            L.err = 0;

            auto flush = [&]() {
                if (len | qlen | L.err) {
                    const unsigned long long idx = rec - P.rec_first;
                    if (FASTQ) { if (len) atomicAdd(P.rec_len + 2 * idx, len); if (qlen) atomicAdd(P.rec_len + 2 * idx + 1, qlen); }
                    else if (len) atomicAdd(P.rec_len + idx, len);
                    letters_total += len;
                    if (L.err) {
                        const unsigned int pr = (L.err & 1u) ? P_ALPHABET : P_IUPAC;
                        const unsigned long long key = rec * 4 + pr;
                        err_key = key < err_key ? key : err_key;
                    }
                }
                len = 0; qlen = 0; L.err = 0;
            };

            for (int i = 0; i < kChunk / 4; i++) {
                if (4 * i >= lim) break;
                const uint32_t wv = s_text[i * (kThreads + 1) + tid];
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    if (4 * i + b >= lim) break;
                    const uint8_t ch = (uint8_t)(wv >> (8 * b));
                    const bool at_ls = prev == '\n';
                    if (at_ls) {                              // `ch` begins a line
                        if (FASTQ) {
                            kind = (uint32_t)(line & 3);
                            if (kind == 0) { flush(); rec = P.hdr_file_base + (line >> 2); w.have = 0; w.nmask = w.imask = 0; if (ch != '@') layout |= 1; }
                            else if (kind == 2 && ch != '+') layout |= 1;
                            line++;
                        } else {
                            if (ch == '>') { flush(); hdr_before++; rec = hdr_before - 1; kind = K_HEADER; w.have = 0; w.nmask = w.imask = 0; }
                            else kind = hdr_before ? K_SEQ : K_IGNORE;
                        }
                    } else if (prev == '\r' && ch != '\n') layout |= 1;   // lone CR is a line break in text mode
                    prev = ch;
                    if (ch & 0x80) layout |= 1;               // UTF-8 decoding is the host tokeniser's business
                    if (FASTQ) {
                        if (kind == 1) {
                            const uint8_t cl = s_lut[ch];
                            if (cl <= C_ILLEGAL) { if (cl == C_ILLEGAL) L.err |= 1u << P_ALPHABET; len++; push_letter<SMEM_HIST, true>(w, c, cl, tab, L); }
                            else if (cl == C_SPACE || cl == C_WEIRD) layout |= 1;
                        } else if (kind == 3) {
                            const uint8_t cl = s_lut[ch];
                            if (cl <= C_ILLEGAL) qlen++; else if (cl == C_SPACE || cl == C_WEIRD) layout |= 1;
                        } else if (kind == 2) {
                            if (!at_ls && ch != '\n' && ch != '\r') layout |= 1;   // "+title" repeats go to the host tokeniser
                        }
                    } else if (kind == K_SEQ) {
                        const uint8_t cl = s_lut[ch];
                        if (cl <= C_ILLEGAL) { if (cl == C_ILLEGAL) L.err |= 1u << P_ALPHABET; len++; push_letter<SMEM_HIST, true>(w, c, cl, tab, L); }
                        else if (cl == C_WEIRD) layout |= 1;
                    }
                }
            }
            flush();
            if (a + lim == P.n && prev == '\r') layout |= 1;   // file ends in a lone CR
        }
    }

    // ---- CTA reduction of the fused scalars, one atomic per CTA per scalar ----------------------
    unsigned long long v[8] = {L.kmers, letters_total, L.mono[0], L.mono[1], L.mono[2], L.mono[3], L.nwin, 0};
#pragma unroll
    for (int i = 0; i < 7; i++)
        for (int d = 16; d; d >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], d);
    for (int d = 16; d; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, err_key, d);
        err_key = o < err_key ? o : err_key;
        layout |= __shfl_xor_sync(0xffffffffu, layout, d);
    }
    __syncthreads();
    if ((tid & 31) == 0) { for (int i = 0; i < 7; i++) s_red[tid >> 5][i] = v[i]; s_red[tid >> 5][7] = err_key; }
    __syncthreads();
    if (tid < 7) {
        unsigned long long s = 0;
        for (int wp = 0; wp < kThreads / 32; wp++) s += s_red[wp][tid];
        unsigned long long* dst = tid == 0 ? &P.acc->total_kmers : tid == 1 ? &P.acc->sum_len : tid <= 5 ? &P.acc->mono[tid - 2] : &P.acc->n_windows_with_n;
        if (s) atomicAdd(dst, s);
    } else if (tid == 7) {
        unsigned long long m = ~0ull;
        for (int wp = 0; wp < kThreads / 32; wp++) m = s_red[wp][7] < m ? s_red[wp][7] : m;
        if (m != ~0ull) atomicMin(&P.acc->err_key, m);
    }
    if (layout && (tid & 31) == 0) atomicOr(&P.acc->layout_flag, 1u);
    if (SMEM_HIST) {
        __syncthreads();
        for (uint32_t i = tid; i < nbins; i += kThreads) { const unsigned int x = s_dyn[i]; if (x) atomicAdd(P.tab32 + i, x); }
    }
}

}  // namespace kdb
