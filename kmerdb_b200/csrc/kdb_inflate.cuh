// GPU inflate of BGZF members (reference: gzip.open at kmerdb/parse.py:63-68, sniffed by
// kmerdb/util.py:80-88).  BGZF (bgzip, Bio.bgzf) is a series of independent gzip members of at most 64 KiB of text each,
// so the COMPRESSED bytes cross PCIe and every member is inflated by one warp straight into the text buffer the
// tokeniser reads.  RFC 1951 restated for a warp:
//   * all 32 lanes run the bit-stream decoder redundantly (no divergence); the compressed bytes arrive a 128-byte line at a time
//     in registers (one word per lane) and a refill is a shuffle;
//   * 32 symbols are decoded into per-lane tokens, a warp scan gives their output positions, then the literals (one byte per
//     lane) and the matches whose source lies before the batch are copied side by side, the rest in token order
//     (out[pos + i] = out[pos - d + i mod d] covers overlapping matches);
//   * Huffman tables per warp in shared memory: a 10-bit (literal/length) and an 8-bit (distance) first-level table filled by
//     the lanes together, longer codes by the canonical count/offset walk (RFC 1951 3.2.2);
//   * stored and fixed-Huffman blocks included (bgzip's EOF marker is an empty fixed block).
// Tried and rejected (r02, 10 M reads re-blocked at zlib level 1 / 6): loading the bytes of a short match at once and storing them
// only when the next match needs them (to overlap the L2 round trip with the following decode), together with length / distance
// bases by arithmetic instead of table loads: 23.6 / 35.9 GB/s of text against 28.1 / 40.0 without.
// Checks per member: the stream ends exactly at its last block, the output length equals ISIZE, and CRC-32 of the output
// equals the trailer (32 lane-chunk CRCs by table, joined with the GF(2) shift operator of zlib's crc32_combine).
#pragma once
#include "kdb_common.cuh"

namespace kdb {

struct BgzfMember {
    unsigned long long in_off;    // first byte of the deflate payload inside the compressed buffer
    unsigned long long out_off;   // where its text goes
    unsigned int in_len;          // payload bytes (member size - header - 8 trailer bytes)
    unsigned int out_len;         // ISIZE
    unsigned int crc;             // CRC-32 of the text (gzip trailer)
    unsigned int pad;
};

constexpr int kInfWarps = 8;                 // warps (members in flight) per CTA
constexpr int kLitBits = 10, kDistBits = 8;

struct InfTables {
    unsigned short lit[1 << kLitBits];       // first level: symbol | length << 9  (0: longer code / invalid)
    unsigned short dist[1 << kDistBits];     // symbol | length << 5
    unsigned short lit_sym[288], dist_sym[32];      // symbols in canonical order
    unsigned short lit_cnt[16], dist_cnt[16];       // codes per length
    unsigned char lens[352];                 // code lengths while a block header is read
};

__constant__ unsigned short c_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
__constant__ unsigned char c_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ unsigned short c_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ unsigned char c_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ unsigned char c_clen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
__constant__ unsigned int c_crc_table[256];  // filled by the host (kdb_api.cu: build_crc_table)

// LSB-first bit reader, identical in every lane.  The payload is fetched a 128-byte line at a time: lane l holds word l of the
// current line in a register (one coalesced load per 128 bytes of compressed data, the next line requested before it is needed),
// and a refill is a shuffle from the lane that holds the next word — no memory latency on the decode chain and no dependence on
// L1, which the per-warp tables leave little of.  The payload may start at any byte: the stream starts at the aligned line below
// it and the first word is shifted into place.  A line may reach past the payload (the trailer and the next member follow; the
// buffer carries slack behind the last one): `used()` tells how many payload bits were really consumed.
struct BitReader {
    const uint32_t* line;             // next 128-byte line to request
    uint32_t cur, nxt;                // my word of the line being consumed / of the line after it
    unsigned long long buf;
    long long bits_in;                // payload bits loaded so far
    int cnt;
    int widx;                         // next word of the current line (0..32)
    __device__ __forceinline__ void init(const uint8_t* p) {
        const int lane = threadIdx.x & 31;
        const uintptr_t a = (uintptr_t)p;
        line = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)127);
        cur = __ldg(line + lane); line += 32;
        nxt = __ldg(line + lane); line += 32;
        const unsigned int skip = (unsigned int)(a & 3u);
        widx = (int)((a & 127u) >> 2);
        const uint32_t w0 = __shfl_sync(0xffffffffu, cur, widx);
        widx++;
        buf = (unsigned long long)(w0 >> (8 * skip));
        cnt = 32 - 8 * (int)skip;
        bits_in = cnt;
    }
    __device__ __forceinline__ void refill() {                      // at least 33 bits afterwards; called by all lanes together
        if (cnt <= 32) {
            if (widx == 32) { cur = nxt; nxt = __ldg(line + (threadIdx.x & 31)); line += 32; widx = 0; }
            const uint32_t w = __shfl_sync(0xffffffffu, cur, widx);
            widx++;
            buf |= (unsigned long long)w << cnt; cnt += 32; bits_in += 32;
        }
    }
    __device__ __forceinline__ unsigned int peek(int n) const { return (unsigned int)buf & ((1u << n) - 1u); }
    __device__ __forceinline__ void drop(int n) { buf >>= n; cnt -= n; }
    __device__ __forceinline__ unsigned int take(int n) { const unsigned int v = peek(n); drop(n); return v; }
    __device__ __forceinline__ long long used() const { return bits_in - cnt; }
};

// Canonical decode, one bit at a time (codes longer than the first-level table; the code-length alphabet).
__device__ __forceinline__ int slow_decode(BitReader& br, const unsigned short* cnt, const unsigned short* sym) {
    int code = 0, first = 0, index = 0;
    for (int len = 1; len <= 15; len++) {
        code |= (int)br.take(1);
        const int c = cnt[len];
        if (code - c < first) return sym[index + (code - first)];
        index += c; first += c; first <<= 1; code <<= 1;
    }
    return -1;
}

// Build the canonical order + first-level table of `n` code lengths (lens[0..n)).  All lanes call it; returns false on an
// over-subscribed set.  tbl_bits = width of the first level, sym_shift = where the length sits in an entry.
__device__ bool build_table(const unsigned char* lens, int n, unsigned short* cnt, unsigned short* sym, unsigned short* tbl, int tbl_bits, int sym_shift, int lane) {
    if (lane < 16) cnt[lane] = 0;
    __syncwarp();
    if (lane == 0) {
        for (int i = 0; i < n; i++) cnt[lens[i]]++;
        cnt[0] = 0;
    }
    __syncwarp();
    int left = 1;
    unsigned short offs[16];
    offs[1] = 0;
    for (int len = 1; len <= 15; len++) {
        left <<= 1;
        left -= cnt[len];
        if (left < 0) return false;
        if (len < 15) offs[len + 1] = offs[len] + cnt[len];
    }
    for (int i = lane; i < (1 << tbl_bits); i += 32) tbl[i] = 0;
    __syncwarp();
    if (lane == 0) {
        unsigned short o[16];
        for (int len = 1; len <= 15; len++) o[len] = offs[len];
        for (int i = 0; i < n; i++) if (lens[i]) sym[o[lens[i]]++] = (unsigned short)i;
    }
    __syncwarp();
    // first level: symbol j of canonical order with length L and code c fills every index whose low L bits are c reversed
    int code = 0, idx = 0;
    for (int len = 1; len <= tbl_bits; len++) {
        const int c = cnt[len];
        for (int j = lane; j < c; j += 32) {
            const unsigned int rev = __brev((unsigned int)(code + j)) >> (32 - len);
            const unsigned short e = (unsigned short)(sym[idx + j] | (len << sym_shift));
            for (unsigned int i = rev; i < (1u << tbl_bits); i += 1u << len) tbl[i] = e;
        }
        code = (code + c) << 1;
        idx += c;
    }
    __syncwarp();
    return true;
}

__device__ __forceinline__ int decode_sym(BitReader& br, const unsigned short* tbl, int tbl_bits, int sym_shift, const unsigned short* cnt, const unsigned short* sym) {
    const unsigned short e = tbl[br.peek(tbl_bits)];
    if (e) { br.drop(e >> sym_shift); return e & ((1 << sym_shift) - 1); }
    return slow_decode(br, cnt, sym);
}

// CRC-32 (reflected, 0xEDB88320) helpers: byte table walk, and the GF(2) operator that appends n zero bytes
__device__ __forceinline__ unsigned int gf2_times(const unsigned int* mat, unsigned int vec) {
    unsigned int sum = 0;
    for (int i = 0; vec; vec >>= 1, i++) if (vec & 1u) sum ^= mat[i];
    return sum;
}

// status per member: 0 ok, 1 malformed stream, 2 length mismatch, 3 CRC mismatch
// BATCH: the symbols of a block are decoded 32 at a time into per-lane tokens (lane t keeps token t), their output positions come
// from a warp scan, and the copies of a batch run together: literals one byte per lane, matches whose source lies before the
// batch one per lane side by side (ONE L2 round trip per batch instead of one per match); a match that reads what the same batch
// writes waits for its turn.  !BATCH: one symbol at a time, the warp copies each match together (r02 first version).
template <bool BATCH>
__global__ void __launch_bounds__(kInfWarps * 32) k_inflate_bgzf(const uint8_t* __restrict__ comp, const BgzfMember* __restrict__ members, unsigned int n_members,
                                                                 uint8_t* __restrict__ out, unsigned int* __restrict__ status, int check_crc) {
    __shared__ InfTables s_tab[kInfWarps];
    __shared__ unsigned int s_mat[kInfWarps][2][32];
    __shared__ unsigned int s_crc[256];                              // (data-dependent index: a __constant__ table would serialise the lanes)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_crc[i] = c_crc_table[i];
    __syncthreads();
    InfTables& T = s_tab[warp];
    for (unsigned int mi = blockIdx.x * kInfWarps + warp; mi < n_members; mi += gridDim.x * kInfWarps) {
        const BgzfMember M = members[mi];
        uint8_t* const dst = out + M.out_off;
        const uint8_t* const payload = comp + M.in_off;
        BitReader br;
        br.init(payload);
        unsigned int pos = 0, err = 0;
        bool last = false;
        while (!last && !err) {
            if (br.used() > 8ll * (long long)M.in_len) { err = 1; break; }      // (a run of empty blocks in garbage must not read on for ever)
            br.refill();
            last = br.take(1) != 0;
            const unsigned int type = br.take(2);
            if (type == 0) {                                    // stored
                br.drop(br.cnt & 7);                            // (whole bytes are loaded, so the bit count says where the byte border is)
                br.refill();
                const unsigned int len = br.take(16);
                br.refill();
                const unsigned int nlen = br.take(16);
                if ((len ^ 0xFFFFu) != nlen) { err = 1; break; }
                const unsigned long long src = (unsigned long long)(br.used() >> 3);
                if (src + len > M.in_len || pos + len > M.out_len) { err = 1; break; }
                for (unsigned int i = lane; i < len; i += 32) dst[pos + i] = payload[src + i];
                pos += len;
                br.init(payload + src + len);
                br.bits_in += 8ll * (long long)(src + len);
                __syncwarp();
                continue;
            }
            if (type == 3) { err = 1; break; }
            if (type == 1) {                                    // fixed codes (RFC 1951 3.2.6)
                for (int i = lane; i < 288; i += 32) T.lens[i] = i < 144 ? 8 : (i < 256 ? 9 : (i < 280 ? 7 : 8));
                __syncwarp();
                build_table(T.lens, 288, T.lit_cnt, T.lit_sym, T.lit, kLitBits, 9, lane);
                for (int i = lane; i < 30; i += 32) T.lens[i] = 5;
                __syncwarp();
                build_table(T.lens, 30, T.dist_cnt, T.dist_sym, T.dist, kDistBits, 5, lane);
            } else {                                            // dynamic codes (3.2.7)
                const int nlen = (int)br.take(5) + 257, ndist = (int)br.take(5) + 1, ncode = (int)br.take(4) + 4;
                if (nlen > 286 || ndist > 30) { err = 1; break; }
                if (lane < 19) T.lens[lane] = 0;
                __syncwarp();
                for (int i = 0; i < ncode; i++) {
                    br.refill();
                    const unsigned int v = br.take(3);
                    if (lane == 0) T.lens[c_clen_order[i]] = (unsigned char)v;
                }
                __syncwarp();
                // the code-length alphabet (19 symbols, at most 7 bits): canonical walk only; its tables sit where the distance
                // tables go later
                if (!build_table(T.lens, 19, T.dist_cnt, T.dist_sym, T.dist, 7, 5, lane)) { err = 1; break; }
                __syncwarp();
                int idx = 0;
                while (idx < nlen + ndist) {
                    br.refill();
                    const int sym = decode_sym(br, T.dist, 7, 5, T.dist_cnt, T.dist_sym);
                    if (sym < 0) { err = 1; break; }
                    if (sym < 16) { if (lane == 0) T.lens[32 + idx] = (unsigned char)sym; idx++; }
                    else {
                        int rep, val = 0;
                        if (sym == 16) {
                            if (idx == 0) { err = 1; break; }
                            __syncwarp();
                            val = T.lens[32 + idx - 1];
                            rep = 3 + (int)br.take(2);
                        } else if (sym == 17) rep = 3 + (int)br.take(3);
                        else rep = 11 + (int)br.take(7);
                        if (idx + rep > nlen + ndist) { err = 1; break; }
                        __syncwarp();
                        for (int i = lane; i < rep; i += 32) T.lens[32 + idx + i] = (unsigned char)val;
                        idx += rep;
                        __syncwarp();
                    }
                }
                if (err) break;
                __syncwarp();
                if (T.lens[32 + 256] == 0) { err = 1; break; }          // no end-of-block code
                if (!build_table(T.lens + 32, nlen, T.lit_cnt, T.lit_sym, T.lit, kLitBits, 9, lane)) { err = 1; break; }
                if (!build_table(T.lens + 32 + nlen, ndist, T.dist_cnt, T.dist_sym, T.dist, kDistBits, 5, lane)) { err = 1; break; }
            }
            // ---- the symbols of the block ---------------------------------------------------------------------
            if constexpr (BATCH) {
                bool eob = false;
                while (!eob && !err) {
                    // phase 1: up to 32 tokens; every lane decodes, lane t keeps token t:  (length, distance or literal byte)
                    unsigned int my_len = 0, my_d = 0, n_tok = 0;
                    while (n_tok < 32) {
                        br.refill();
                        int sym = decode_sym(br, T.lit, kLitBits, 9, T.lit_cnt, T.lit_sym);
                        if (sym < 0) { err = 1; break; }
                        unsigned int tl = 1, td = 0x80000000u | (unsigned int)sym;         // literal: top bit set, the byte below
                        if (sym >= 256) {
                            if (sym == 256) { eob = true; break; }
                            sym -= 257;
                            if (sym >= 29) { err = 1; break; }
                            tl = c_len_base[sym] + br.take(c_len_extra[sym]);
                            br.refill();
                            const int ds = decode_sym(br, T.dist, kDistBits, 5, T.dist_cnt, T.dist_sym);
                            if (ds < 0 || ds >= 30) { err = 1; break; }
                            td = c_dist_base[ds] + br.take(c_dist_extra[ds]);
                        }
                        if ((unsigned int)lane == n_tok) { my_len = tl; my_d = td; }
                        n_tok++;
                    }
                    if (err) break;
                    // phase 2: where every token's output goes
                    unsigned int incl = my_len;
#pragma unroll
                    for (int dd = 1; dd < 32; dd <<= 1) { const unsigned int o = __shfl_up_sync(0xffffffffu, incl, dd); if (lane >= dd) incl += o; }
                    const unsigned int total = __shfl_sync(0xffffffffu, incl, 31);
                    const unsigned int my_pos = pos + incl - my_len;
                    const bool is_match = my_len != 0 && !(my_d & 0x80000000u);
                    if (pos + total > M.out_len) { err = 2; break; }
                    if (__any_sync(0xffffffffu, is_match && my_d > my_pos)) { err = 1; break; }
                    // phase 3: literals, and the matches that read only text written before this batch
                    const unsigned int span = my_len < my_d ? my_len : my_d;                 // bytes of source a match really reads
                    const bool indep = is_match && my_pos - my_d + span <= pos;
                    if (my_len != 0 && !is_match) dst[my_pos] = (uint8_t)my_d;
                    if (indep) {
                        const uint8_t* from = dst + my_pos - my_d;
                        for (unsigned int i = 0; i < my_len; i += 8) {
                            uint8_t v[8];
#pragma unroll
                            for (int q = 0; q < 8; q++) { const unsigned int j = i + q; v[q] = j < my_len ? from[my_d >= my_len ? j : j % my_d] : (uint8_t)0; }
#pragma unroll
                            for (int q = 0; q < 8; q++) if (i + q < my_len) dst[my_pos + i + q] = v[q];
                        }
                    }
                    __syncwarp();
                    // the rest in token order: each may read what the one before wrote
                    unsigned int dep = __ballot_sync(0xffffffffu, is_match && !indep);
                    while (dep) {
                        const int t = __ffs(dep) - 1;
                        dep &= dep - 1;
                        const unsigned int tl = __shfl_sync(0xffffffffu, my_len, t), td = __shfl_sync(0xffffffffu, my_d, t), tp = __shfl_sync(0xffffffffu, my_pos, t);
                        const uint8_t* from = dst + tp - td;
                        if (td >= tl) { for (unsigned int i = lane; i < tl; i += 32) dst[tp + i] = from[i]; }
                        else { for (unsigned int i = lane; i < tl; i += 32) dst[tp + i] = from[i % td]; }
                        __syncwarp();
                    }
                    pos += total;
                }
                continue;                                           // next block of the member
            }
            while (true) {
                br.refill();
                int sym = decode_sym(br, T.lit, kLitBits, 9, T.lit_cnt, T.lit_sym);
                if (sym < 0) { err = 1; break; }
                if (sym < 256) {
                    if (pos >= M.out_len) { err = 2; break; }
                    if (lane == 0) dst[pos] = (uint8_t)sym;
                    pos++;
                    continue;
                }
                if (sym == 256) break;
                sym -= 257;
                if (sym >= 29) { err = 1; break; }
                const unsigned int len = c_len_base[sym] + br.take(c_len_extra[sym]);
                br.refill();
                const int ds = decode_sym(br, T.dist, kDistBits, 5, T.dist_cnt, T.dist_sym);
                if (ds < 0 || ds >= 30) { err = 1; break; }
                const unsigned int d = c_dist_base[ds] + br.take(c_dist_extra[ds]);
                if (d > pos) { err = 1; break; }
                if (pos + len > M.out_len) { err = 2; break; }
                __syncwarp();                                       // the literals before the match are in place
                const uint8_t* from = dst + pos - d;
                if (d >= len) { for (unsigned int i = lane; i < len; i += 32) dst[pos + i] = from[i]; }
                else { for (unsigned int i = lane; i < len; i += 32) dst[pos + i] = from[i % d]; }
                pos += len;
                __syncwarp();
            }
        }
        if (!err && br.used() > 8ll * (long long)M.in_len) err = 1;   // ran past the end of the payload
        if (!err && pos != M.out_len) err = 2;
        __syncwarp();
        if (!err && check_crc) {
            // every lane takes one chunk of `per` bytes, then the partial values are joined front to back:
            // crc(A || B) = shift(crc(A), |B|) ^ crc(B) on the raw register values (zlib crc32_combine)
            const unsigned int per = (M.out_len + 31) / 32;
            const unsigned int b0 = min(M.out_len, lane * per), b1 = min(M.out_len, b0 + per);
            unsigned int c = 0xFFFFFFFFu;
            {   // bytes up to a 4-byte border, then whole words (one load per four table steps), then the tail
                unsigned int i = b0;
                for (; i < b1 && ((uintptr_t)(dst + i) & 3u); i++) c = s_crc[(c ^ dst[i]) & 0xFFu] ^ (c >> 8);
                for (; i + 4 <= b1; i += 4) {
                    c ^= *reinterpret_cast<const uint32_t*>(dst + i);
                    c = s_crc[c & 0xFFu] ^ (c >> 8); c = s_crc[c & 0xFFu] ^ (c >> 8);
                    c = s_crc[c & 0xFFu] ^ (c >> 8); c = s_crc[c & 0xFFu] ^ (c >> 8);
                }
                for (; i < b1; i++) c = s_crc[(c ^ dst[i]) & 0xFFu] ^ (c >> 8);
            }
            c ^= 0xFFFFFFFFu;
            // operator for `per` zero bytes: start from the one-bit operator, square to one byte, then square-and-multiply
            unsigned int* even = s_mat[warp][0];
            unsigned int* odd = s_mat[warp][1];
            odd[lane] = lane == 0 ? 0xEDB88320u : (1u << (lane - 1));                       // 1 zero bit
            __syncwarp();
            even[lane] = gf2_times(odd, odd[lane]); __syncwarp();                           // 2 bits
            odd[lane] = gf2_times(even, even[lane]); __syncwarp();                          // 4 bits
            unsigned int total = 0;
            // join: total = crc of lanes [0, L]; every lane's chunk but the last has `per` bytes
            for (int L = 0; L < 32; L++) {
                const unsigned int cl = __shfl_sync(0xffffffffu, c, L);
                const unsigned int lb0 = min(M.out_len, L * per), lb1 = min(M.out_len, lb0 + per);
                unsigned int len2 = lb1 - lb0;
                if (len2 == 0) continue;
                if (L == 0) { total = cl; continue; }
                // shift `total` by len2 bytes: square-and-multiply, every lane owns one row of the operator.  B starts as the
                // 4-bit operator; each round squares it (first round: one byte) and applies it when the bit of len2 is set.
                unsigned int v = total, nb = len2;
                const unsigned int four_bits = odd[lane];
                unsigned int* A = even;
                unsigned int* B = odd;
                __syncwarp();
                do {
                    A[lane] = gf2_times(B, B[lane]);
                    __syncwarp();
                    if (nb & 1u) v = gf2_times(A, v);
                    nb >>= 1;
                    unsigned int* t = A; A = B; B = t;
                    __syncwarp();
                } while (nb);
                total = v ^ cl;
                odd[lane] = four_bits;                               // back to the 4-bit operator for the next chunk
                __syncwarp();
            }
            if (total != M.crc) err = 3;
        }
        if (lane == 0) status[mi] = err;
        __syncwarp();
    }
}

}  // namespace kdb
