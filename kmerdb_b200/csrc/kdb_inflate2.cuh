// GPU inflate of BGZF members in two kernels (reference: gzip.open at kmerdb/parse.py:63-68).
//
// The warp-per-member kernel of kdb_inflate.cuh runs ONE serial Huffman decode in all 32 lanes of a warp and is bound by
// instruction issue (32-40 warps share four issue slots).  Here the two halves of DEFLATE are split by what they parallelise over:
//   A  k_inflate_decode   one THREAD per member: the bit stream is decoded into match tokens; literals go straight to their final
//                         place in the text (the decoder knows the output position).  32 different members per warp, no redundant
//                         lanes: the kernel is bound by the latency of one decode chain, not by issue.  First-level tables
//                         (9-bit literal/length, 6-bit distance: 1152 bytes per thread, six warps per SM) live in shared memory, interleaved so that thread t only ever
//                         touches bank t (no bank conflicts whatever the indices); canonical order for longer codes, code lengths
//                         and counts are thread-local.
//   B  k_inflate_resolve  one WARP per member: 32 tokens at a time, output positions by a warp scan, matches that read text
//                         before the batch are copied side by side (one per lane), the others in token order; then ISIZE / CRC-32
//                         as before.
// Token (u32): bit 31 clear: [30:23] literals before the match (0..255), [22:15] length - 3, [14:0] distance - 1;
//              bit 31 set:   [23:0] literals only (a run longer than 255, stored blocks, the tail of the member).
// A member of n text bytes yields at most n/3 + n/256 + 1 tokens.
#pragma once
#include "kdb_inflate.cuh"

namespace kdb {

constexpr int kDecThreads = 64;                       // members per CTA of kernel A (two warps: three CTAs = six warps per SM)
constexpr int kDecLitBits = 9, kDecDistBits = 6;     // (8 / 6 bits: 10 warps per SM instead of 6, waves of 47 360 members — measured 13.5 ms for 10 M reads and 59.1 ms for 40 M against 13.2 / 59.6: the steps get slower as fast as the waves get wider)
constexpr int kDecHalfWords = (1 << kDecLitBits) + (1 << kDecDistBits);      // u16 entries per thread
constexpr size_t kDecSmemBytes = (size_t)kDecThreads * kDecHalfWords * 2 + (size_t)kDecThreads * 32;     // tables + the bit readers' rings

__host__ __device__ inline unsigned int inflate_token_cap(unsigned int out_len) { return (out_len / 3 + out_len / 256 + 8 + 3) & ~3u; }

// Entry i of thread t's table is half-word i * kDecThreads + t: one shift gives the address, and whatever the indices, the lanes
// of a warp touch 16 different banks two by two (a row of the table = 128 bytes = all 32 banks).
struct LaneTable {
    unsigned short* base;                              // this thread's entry 0
    __device__ __forceinline__ unsigned short& at(unsigned int i) const { return base[i * kDecThreads]; }
};
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }

// LSB-first bit reader of ONE thread.  The payload is staged through a ring of four 8-byte chunks per thread in shared memory,
// filled by cp.async: while chunk i is read, chunks i+1 .. i+3 hold the bytes that follow (the last two possibly still in
// flight), so a request has two chunks' worth of symbols to arrive and NO register waits for it.  (A register read-ahead does
// not work here: the warp stalls on the pending register at the first instruction that names it, even a predicated-off move —
// 29 % of the samples of that version sat on one `@!P MOV`.)  A refill is a predicated shared-memory load; finished chunks are
// re-requested by roll(), ONCE per step of the symbol loop (a step reads at most 48 bits = one chunk), so the refills themselves
// carry no branch.
struct BitReader1 {
    const uint8_t* gp;                // next 8 bytes to request
    uint32_t ring;                    // shared-memory address of this thread's chunk 0; chunk c lies c * kChunkStride further
    uint32_t ro;                      // ring offset of the next word (0, 4, .. 28)
    uint32_t rc;                      // the chunk being read as of the last roll()
    unsigned long long buf;
    int cnt;
    unsigned int bits_in;
    static constexpr uint32_t kChunkStride = kDecThreads * 8;
    __device__ __forceinline__ void fetch(uint32_t chunk) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\ncp.async.commit_group;" ::"r"(ring + chunk * kChunkStride), "l"(gp) : "memory");
        gp += 8;
    }
    __device__ __forceinline__ void refill_on(bool on) {            // at least 33 bits afterwards (when `on`)
        if (on && cnt <= 32) {
            uint32_t w;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(ring + (ro >> 3) * kChunkStride + (ro & 7u)) : "memory");
            buf |= (unsigned long long)w << cnt; cnt += 32; bits_in += 32;
            ro = (ro + 4u) & 31u;
        }
    }
    __device__ __forceinline__ void roll() {                        // after at most two refills
        if ((ro >> 3) != rc) {
            fetch(rc);
            rc = (rc + 1u) & 3u;
            asm volatile("cp.async.wait_group 2;" ::: "memory");   // all but the two latest requests: the chunk being read and the next
        }
    }
    __device__ __forceinline__ void refill() { refill_on(true); roll(); }
    __device__ __forceinline__ void init(const uint8_t* a) {
        const uintptr_t u = (uintptr_t)a;
        gp = reinterpret_cast<const uint8_t*>(u & ~(uintptr_t)7);
        asm volatile("cp.async.wait_group 0;" ::: "memory");       // (a re-start after a stored block: nothing of the old position in flight)
        fetch(0); fetch(1); fetch(2); fetch(3);
        asm volatile("cp.async.wait_group 2;" ::: "memory");
        rc = 0;
        ro = (uint32_t)(u & 4u);
        buf = 0; cnt = 0; bits_in = 0;
        refill();
        const int skip = 8 * (int)(u & 3u);
        buf >>= skip; cnt -= skip; bits_in -= (unsigned int)skip;
    }
    __device__ __forceinline__ unsigned int peek(int n) const { return (unsigned int)buf & ((1u << n) - 1u); }
    __device__ __forceinline__ void drop(int n) { buf >>= n; cnt -= n; }
    __device__ __forceinline__ unsigned int take(int n) { const unsigned int v = peek(n); drop(n); return v; }
    __device__ __forceinline__ unsigned int used() const { return bits_in - (unsigned int)cnt; }
};

// Canonical order + first-level table of n code lengths, by one thread (RFC 1951 3.2.2).  false: over-subscribed set.
// lim[L] (L = 1..15) = (first code of length L + codes of length L) | (place of that first code in `sym` - first code) << 16:
// what long_decode needs for the codes the first level does not hold.
__device__ __noinline__ bool build_table1(const unsigned char* lens, int n, unsigned int* lim, unsigned short* sym, LaneTable tbl, unsigned int tbl_off,
                                          int tbl_bits, int sym_shift) {
    unsigned short cnt[16], offs[16];
    for (int i = 0; i < 16; i++) cnt[i] = 0;
    for (int i = 0; i < n; i++) cnt[lens[i]]++;
    cnt[0] = 0;
    int left = 1;
    unsigned int first = 0;
    offs[1] = 0;
    for (int len = 1; len <= 15; len++) {
        left <<= 1;
        left -= cnt[len];
        if (left < 0) return false;
        if (len < 15) offs[len + 1] = offs[len] + cnt[len];
        lim[len] = (first + cnt[len]) | (((unsigned int)offs[len] - first) << 16);
        first = (first + cnt[len]) << 1;
    }
    for (int i = 0; i < n; i++) if (lens[i]) sym[offs[lens[i]]++] = (unsigned short)i;
    for (unsigned int i = 0; i < (1u << tbl_bits); i++) tbl.at(tbl_off + i) = 0;
    int code = 0, idx = 0;
    for (int len = 1; len <= tbl_bits; len++) {
        const int c = cnt[len];
        for (int j = 0; j < c; j++) {
            const unsigned int rev = __brev((unsigned int)(code + j)) >> (32 - len);
            const unsigned short e = (unsigned short)(sym[idx + j] | (len << sym_shift));
            for (unsigned int i = rev; i < (1u << tbl_bits); i += 1u << len) tbl.at(tbl_off + i) = e;
        }
        code = (code + c) << 1;
        idx += c;
    }
    return true;
}

// A code longer than the first level (lengths L0 .. L0 + NL - 1): the next 15 bits, first bit on top, against the canonical limit
// of every length in turn (no shorter code matches: the first level would have held it).  n_sym bounds the look-up for garbage.
template <int L0, int NL>
__device__ __forceinline__ int long_decode(BitReader1& br, const unsigned int (&lim)[NL], const unsigned short* sym, unsigned int n_sym) {
    const unsigned int x = __brev((unsigned int)br.buf) >> 17;
#pragma unroll
    for (int i = 0; i < NL; i++) {
        const int L = L0 + i;
        const unsigned int c = x >> (15 - L);
        if (c < (lim[i] & 0xFFFFu)) { br.drop(L); return sym[min((c + (lim[i] >> 16)) & 0xFFFFu, n_sym - 1u)]; }
    }
    return -1;
}

// ---- kernel A: one thread per member ----------------------------------------------------------------------------------------------
// members[i].pad = first token of member i in `tokens`, in units of four tokens.  status: 0 ok, 1 malformed stream, 2 length mismatch.
//
// The lanes of a warp run in lock step (left to themselves they drift apart at the literal / match branch and the warp ends up
// running one lane at a time: 301 ms instead of 39 for the same input).  A round = every unfinished lane reads one block header
// and builds its tables, then all lanes decode side by side until the last of them has reached the end of its block.  With 32
// streams side by side every path is taken by some lane in every step, so a step costs the UNION of all paths: the step is
// therefore one whole symbol (a literal, or a length with its distance) written as straight-line predicated code, not a state
// machine with one look-up per iteration (measured: the same union of ~175 instructions, twice as many iterations: 12.8 ms
// against 9.4 for 1.62 GB of text).  A code longer than the first level parks its lane; parked lanes are served together once
// eight of them wait (the canonical walk and its thread-local symbol array would otherwise be on the path of nearly every step)
// and re-enter the step with the symbol in hand.
__global__ void __launch_bounds__(kDecThreads) k_inflate_decode(const uint8_t* __restrict__ comp, const BgzfMember* __restrict__ members, unsigned int n_members,
                                                                uint8_t* __restrict__ out, uint32_t* __restrict__ tokens, unsigned int* __restrict__ n_tokens,
                                                                unsigned int* __restrict__ status) {
    extern __shared__ __align__(16) unsigned short s_dec[];
    __shared__ unsigned int s_base[64];                              // base | extra bits << 16 (RFC 1951 3.2.5): lengths at 0, distances at 32
    if (threadIdx.x < 29) s_base[threadIdx.x] = c_len_base[threadIdx.x] | ((unsigned int)c_len_extra[threadIdx.x] << 16);
    if (threadIdx.x < 30) s_base[32 + threadIdx.x] = c_dist_base[threadIdx.x] | ((unsigned int)c_dist_extra[threadIdx.x] << 16);
    __syncthreads();
    const unsigned int mi = blockIdx.x * kDecThreads + threadIdx.x;
    const bool have = mi < n_members;
    LaneTable T;
    T.base = s_dec + threadIdx.x;
    const uint32_t tab_sa = (uint32_t)__cvta_generic_to_shared(T.base);          // entry i: tab_sa + i * 2 * kDecThreads
    const uint32_t base_sa = (uint32_t)__cvta_generic_to_shared(s_base);
    constexpr unsigned int kLitOff = 0, kDistOff = 1u << kDecLitBits;
    constexpr int kShift = 9;                                        // every table entry: symbol | code length << 9
    static_assert(kDecLitBits >= 7, "the code-length alphabet borrows the literal table");
    unsigned char lens[320];
    unsigned short lit_sym[288], dist_sym[32];
    unsigned int lim16[16];                                          // build_table1's limits; the ones in use are copied to registers
    unsigned int lit_lim[15 - kDecLitBits], dist_lim[15 - kDecDistBits];

    BgzfMember M;
    M.in_off = 0; M.out_off = 0; M.in_len = 0; M.out_len = 0; M.crc = 0; M.pad = 0;
    if (have) M = members[mi];
    uint8_t* const dst = out + M.out_off;
    const uint8_t* const payload = comp + M.in_off;
    uint32_t* const tok0 = tokens + 4ull * M.pad;
    unsigned int n_tok = 0;
    BitReader1 br;
    br.ring = (uint32_t)__cvta_generic_to_shared(s_dec + (size_t)kDecThreads * kDecHalfWords) + threadIdx.x * 8u;
    br.gp = comp; br.ro = 0; br.rc = 0; br.buf = 0; br.cnt = 0; br.bits_in = 0;
    if (have) br.init(payload);
    unsigned int pos = 0, err = 0, run = 0, mlen = 0;
    bool done = !have;
    while (__any_sync(0xffffffffu, !done)) {
        bool in_block = false;
        if (!done) {
            bool last = false;
            do {                                                    // one block header
                if (br.used() > 8u * M.in_len) { err = 1; break; }   // (a run of empty blocks in garbage must not read on for ever)
                br.refill();
                last = br.take(1) != 0;
                const unsigned int type = br.take(2);
                if (type == 0) {                                    // stored: bytes copied as literals
                    br.drop(br.cnt & 7);
                    br.refill();
                    const unsigned int len = br.take(16);
                    br.refill();
                    const unsigned int nlen = br.take(16);
                    if ((len ^ 0xFFFFu) != nlen) { err = 1; break; }
                    const unsigned int src = br.used() >> 3;
                    if (src + len > M.in_len || pos + len > M.out_len) { err = 1; break; }
                    for (unsigned int i = 0; i < len; i++) dst[pos + i] = payload[src + i];
                    pos += len; run += len;
                    br.init(payload + src + len);
                    br.bits_in += 8u * (src + len);
                    break;
                }
                if (type == 3) { err = 1; break; }
                if (type == 1) {                                    // fixed codes (3.2.6)
                    for (int i = 0; i < 288; i++) lens[i] = i < 144 ? 8 : (i < 256 ? 9 : (i < 280 ? 7 : 8));
                    build_table1(lens, 288, lim16, lit_sym, T, kLitOff, kDecLitBits, kShift);
#pragma unroll
                    for (int i = 0; i < 15 - kDecLitBits; i++) lit_lim[i] = lim16[kDecLitBits + 1 + i];
                    for (int i = 0; i < 30; i++) lens[i] = 5;
                    build_table1(lens, 30, lim16, dist_sym, T, kDistOff, kDecDistBits, kShift);
#pragma unroll
                    for (int i = 0; i < 15 - kDecDistBits; i++) dist_lim[i] = lim16[kDecDistBits + 1 + i];
                    in_block = true;
                    break;
                }
                // dynamic codes (3.2.7)
                const int nlen = (int)br.take(5) + 257, ndist = (int)br.take(5) + 1, ncode = (int)br.take(4) + 4;
                if (nlen > 286 || ndist > 30) { err = 1; break; }
                unsigned char cl[19];
                for (int i = 0; i < 19; i++) cl[i] = 0;
                for (int i = 0; i < ncode; i++) { br.refill(); cl[c_clen_order[i]] = (unsigned char)br.take(3); }
                // the code-length alphabet (at most 7 bits: one level covers it) uses the literal table's place for now
                if (!build_table1(cl, 19, lim16, dist_sym, T, kLitOff, 7, kShift)) { err = 1; break; }
                int idx = 0;
                while (idx < nlen + ndist) {
                    br.refill();
                    const unsigned int ce = T.at(kLitOff + br.peek(7));
                    if (ce == 0) { err = 1; break; }
                    br.drop((int)(ce >> kShift));
                    const int sym = (int)(ce & 31u);
                    if (sym < 16) { lens[idx++] = (unsigned char)sym; continue; }
                    int rep, val = 0;
                    if (sym == 16) {
                        if (idx == 0) { err = 1; break; }
                        val = lens[idx - 1];
                        rep = 3 + (int)br.take(2);
                    } else if (sym == 17) rep = 3 + (int)br.take(3);
                    else rep = 11 + (int)br.take(7);
                    if (idx + rep > nlen + ndist) { err = 1; break; }
                    for (int i = 0; i < rep; i++) lens[idx + i] = (unsigned char)val;
                    idx += rep;
                }
                if (err) break;
                if (lens[256] == 0) { err = 1; break; }             // no end-of-block code
                if (!build_table1(lens, nlen, lim16, lit_sym, T, kLitOff, kDecLitBits, kShift)) { err = 1; break; }
#pragma unroll
                for (int i = 0; i < 15 - kDecLitBits; i++) lit_lim[i] = lim16[kDecLitBits + 1 + i];
                if (!build_table1(lens + nlen, ndist, lim16, dist_sym, T, kDistOff, kDecDistBits, kShift)) { err = 1; break; }
#pragma unroll
                for (int i = 0; i < 15 - kDecDistBits; i++) dist_lim[i] = lim16[kDecDistBits + 1 + i];
                in_block = true;
            } while (false);
            if (err) in_block = false;
            if (err || last) done = true;                           // (`done` only stops the next round: the symbols of a last block follow)
        }
        __syncwarp();
        asm volatile("" ::: "memory");                           // (the tables are read by volatile asm loads below)
        // ---- the symbols of the block ----
        // pre1 / pre2: a literal-length / distance symbol decoded by the long-code pass (0x8000 | symbol: an entry of code length 0)
        unsigned int pre1 = 0, pre2 = 0;
        bool dist_next = false, parked = false;                     // a length is in hand (mlen), its distance comes next; waiting for the long-code pass
        unsigned int ib = __ballot_sync(0xffffffffu, in_block);
        while (ib) {
            const bool act = in_block && !parked;
            // literal / length symbol
            const bool act1 = act && !dist_next;
            br.refill_on(act1);
            unsigned int e1 = lds_u16(tab_sa + (kLitOff + br.peek(kDecLitBits)) * (2u * kDecThreads));
            if (pre1) e1 = pre1;
            pre1 = 0;
            const bool ok1 = act1 && e1 != 0;
            if (act1 && e1 == 0) parked = true;
            br.drop(ok1 ? (int)((e1 >> kShift) & 31u) : 0);
            const unsigned int sym = e1 & 511u;
            const bool lit = ok1 && sym < 256u, len_sym = ok1 && sym > 256u;
            if (ok1 && sym == 256u) in_block = false;
            if (lit) {
                if (pos < M.out_len) dst[pos] = (uint8_t)sym; else { err = 2; in_block = false; }
                pos++; run++;
            }
            const unsigned int li = sym - 257u;
            if (len_sym && li >= 29u) { err = 1; in_block = false; }
            const unsigned int be = lds_u32(base_sa + 4u * min(li, 28u));
            if (len_sym) { mlen = (be & 0xFFFFu) + br.take((int)(be >> 16)); dist_next = true; }
            // its distance
            const bool act2 = in_block && !parked && dist_next;
            br.refill_on(act2);
            unsigned int e2 = lds_u16(tab_sa + (kDistOff + br.peek(kDecDistBits)) * (2u * kDecThreads));
            if (pre2) e2 = pre2;
            pre2 = 0;
            const bool ok2 = act2 && e2 != 0;
            if (act2 && e2 == 0) parked = true;
            br.drop(ok2 ? (int)((e2 >> kShift) & 31u) : 0);
            const unsigned int ds = e2 & 511u;
            const unsigned int de = lds_u32(base_sa + 128u + 4u * min(ds, 29u));
            if (ok2) {
                const unsigned int d = (de & 0xFFFFu) + br.take((int)(de >> 16));
                dist_next = false;
                if (ds >= 30u || d > pos) { err = 1; in_block = false; }
                else if (pos + mlen > M.out_len) { err = 2; in_block = false; }
                else {
                    if (run > 255u) { tok0[n_tok++] = 0x80000000u | run; run = 0; }
                    tok0[n_tok++] = (run << 23) | ((mlen - 3u) << 15) | (d - 1u);
                    run = 0;
                    pos += mlen;
                }
            }
            br.roll();
            ib = __ballot_sync(0xffffffffu, in_block);
            const unsigned int pb = __ballot_sync(0xffffffffu, in_block && parked);
            if (pb && (__popc(pb) >= 8 || pb == ib)) {              // long codes: eight lanes wait, or nobody else can go on
                if (in_block && parked) {
                    parked = false;
                    const int ls = dist_next ? long_decode<kDecDistBits + 1, 15 - kDecDistBits>(br, dist_lim, dist_sym, 32u)
                                             : long_decode<kDecLitBits + 1, 15 - kDecLitBits>(br, lit_lim, lit_sym, 288u);
                    if (ls < 0) { err = 1; in_block = false; }
                    else if (dist_next) pre2 = 0x8000u | (unsigned int)ls;
                    else pre1 = 0x8000u | (unsigned int)ls;
                }
                ib = __ballot_sync(0xffffffffu, in_block);
            }
        }
        if (err) done = true;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (!have) return;
    if (!err && br.used() > 8u * M.in_len) err = 1;               // ran past the end of the payload
    if (!err && pos != M.out_len) err = 2;
    if (run) tok0[n_tok++] = 0x80000000u | run;
    n_tokens[mi] = n_tok;
    status[mi] = err;
}

// CRC-32 of dst[0, n) by one warp: lane 0 walks the first n - 31 * per bytes, every other lane `per` = n / 32 bytes (word loads, four
// table steps per word); the partial values are joined front to back with crc(A || B) = shift(crc(A), |B|) ^ crc(B) (zlib
// crc32_combine).  All 31 joins shift by the same `per` bytes, so the GF(2) operator of that shift is built ONCE (square-and-
// multiply from the one-byte operator, every lane owning one column) and applied 31 times.  (The r02 kernel rebuilt the chain of
// squares for every join: two thirds of the instructions of the whole resolve kernel.)   P, Mx: two 32-word operators of this warp.
__device__ __forceinline__ unsigned int warp_crc32(const uint8_t* dst, unsigned int n, int lane, unsigned int* P, unsigned int* Mx, const unsigned int* s_crc) {
    const unsigned int per = n / 32, head = n - 31 * per;
    const unsigned int b0 = lane == 0 ? 0u : head + (lane - 1) * per, b1 = lane == 0 ? head : b0 + per;
    unsigned int c = 0xFFFFFFFFu;
    {
        unsigned int i = b0;
        for (; i < b1 && ((uintptr_t)(dst + i) & 15u); i++) c = s_crc[(c ^ dst[i]) & 0xFFu] ^ (c >> 8);
        if (i + 16 <= b1) {                                      // 16 bytes per load, the next load in flight during the table steps
            uint4 v = *reinterpret_cast<const uint4*>(dst + i);
            for (; i + 16 <= b1; i += 16) {
                uint4 nv = v;
                if (i + 32 <= b1) nv = *reinterpret_cast<const uint4*>(dst + i + 16);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    c ^= w[q];
                    c = s_crc[c & 0xFFu] ^ (c >> 8); c = s_crc[c & 0xFFu] ^ (c >> 8);
                    c = s_crc[c & 0xFFu] ^ (c >> 8); c = s_crc[c & 0xFFu] ^ (c >> 8);
                }
                v = nv;
            }
        }
        for (; i < b1; i++) c = s_crc[(c ^ dst[i]) & 0xFFu] ^ (c >> 8);
    }
    c ^= 0xFFFFFFFFu;
    // P = operator for one zero bit, squared three times = one byte; Mx = identity
    P[lane] = lane == 0 ? 0xEDB88320u : (1u << (lane - 1));
    Mx[lane] = 1u << lane;
    __syncwarp();
    for (int r = 0; r < 3; r++) { const unsigned int t = gf2_times(P, P[lane]); __syncwarp(); P[lane] = t; __syncwarp(); }
    for (unsigned int nb = per; nb; ) {                          // Mx = product of the squares of P picked by the bits of `per`
        if (nb & 1u) { const unsigned int t = gf2_times(P, Mx[lane]); __syncwarp(); Mx[lane] = t; __syncwarp(); }
        nb >>= 1;
        if (nb) { const unsigned int t = gf2_times(P, P[lane]); __syncwarp(); P[lane] = t; __syncwarp(); }
    }
    unsigned int total = __shfl_sync(0xffffffffu, c, 0);
    for (int L = 1; L < 32; L++) total = gf2_times(Mx, total) ^ __shfl_sync(0xffffffffu, c, L);
    return total;
}

// ---- kernel B: one warp per member ------------------------------------------------------------------------------------------------
// status (from kernel A) stays unless the CRC fails (3).
constexpr int kResWarps = 8;
constexpr unsigned int kSpanBytes = 1024;             // the stretch of text one batch of 32 tokens is assembled in (per warp, shared memory)
__global__ void __launch_bounds__(kResWarps * 32) k_inflate_resolve(const BgzfMember* __restrict__ members, unsigned int n_members, const uint32_t* __restrict__ tokens,
                                                                   const unsigned int* __restrict__ n_tokens, uint8_t* __restrict__ out,
                                                                   unsigned int* __restrict__ status, int check_crc) {
    __shared__ unsigned int s_mat[kResWarps][2][32];
    __shared__ unsigned int s_crc[256];
    __shared__ __align__(16) uint32_t s_span[kResWarps][kSpanBytes / 4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_crc[i] = c_crc_table[i];
    __syncthreads();
    const unsigned int mi = blockIdx.x * kResWarps + warp;
    if (mi >= n_members) return;
    if (status[mi]) return;
    const BgzfMember M = members[mi];
    uint8_t* const dst = out + M.out_off;
    const uint32_t* const tok = tokens + 4ull * M.pad;
    const unsigned int n = n_tokens[mi];
    unsigned int pos = 0;
    uint32_t t_next = lane < (int)n ? __ldg(tok + lane) : 0x80000000u;
    for (unsigned int base = 0; base < n; base += 32) {
        const uint32_t t = t_next;
        t_next = base + 32 + lane < n ? __ldg(tok + base + 32 + lane) : 0x80000000u;
        const bool filler = (t & 0x80000000u) != 0;
        const unsigned int lit = filler ? (t & 0xFFFFFFu) : (t >> 23);
        const unsigned int my_len = filler ? 0u : ((t >> 15) & 255u) + 3u;
        const unsigned int my_d = (t & 0x7FFFu) + 1u;
        unsigned int incl = lit + my_len;
#pragma unroll
        for (int dd = 1; dd < 32; dd <<= 1) { const unsigned int o = __shfl_up_sync(0xffffffffu, incl, dd); if (lane >= dd) incl += o; }
        const unsigned int total = __shfl_sync(0xffffffffu, incl, 31);
        const unsigned int my_pos = pos + incl - my_len;                          // where my match starts
        const unsigned int any = __ballot_sync(0xffffffffu, my_len != 0);
        // matches that read only text in place before this batch (kernel A's literals are all in place; what a batch can lack is the
        // output of its own matches, so anything that ends at or before the batch start is safe)
        const unsigned int span = my_len < my_d ? my_len : my_d;
        const bool indep = my_len != 0 && my_pos - my_d + span <= pos;
        const uintptr_t A = (uintptr_t)(dst + pos);
        const unsigned int lead = (unsigned int)(A & 3u), nbytes = lead + total;
        if (any && nbytes <= kSpanBytes) {
            // The batch's stretch of text is assembled in shared memory and leaves as whole words: (a) the stretch as it is (the
            // literals are in place), (b) the matches — sources before the stretch by aligned word loads and a funnel shift, sources
            // inside it from shared memory in token order, (c) one coalesced store.  (Byte-wise copies in global memory, one match per
            // lane, cost 32 sectors per instruction: the kernel was at 86 % of the LSU wavefront peak.)
            uint32_t* const sw = s_span[warp];
            uint8_t* const sp = reinterpret_cast<uint8_t*>(sw);
            uint32_t* const g = reinterpret_cast<uint32_t*>(A - lead);
            const unsigned int nwords = (nbytes + 3u) >> 2;
            // (the first four source words of every match are requested before the stretch itself: one round trip for both)
            const uint8_t* const from = dst + my_pos - my_d;
            const uintptr_t S = (uintptr_t)from;
            const uint32_t* q = reinterpret_cast<const uint32_t*>(S & ~(uintptr_t)3);
            const bool plain = indep && my_d >= my_len;
            uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0;
            if (plain) {
                r0 = q[0]; r1 = q[1];
                if (my_len > 4) r2 = q[2];
                if (my_len > 8) r3 = q[3];
            }
            for (unsigned int w = lane; w < nwords; w += 32) sw[w] = g[w];
            __syncwarp();
            if (indep) {
                uint8_t* to = sp + lead + (my_pos - pos);
                if (plain) {
                    const unsigned int sh = 8u * (unsigned int)(S & 3u);
                    uint32_t lo = r0;
                    for (unsigned int i = 0; i < my_len; i += 4) {
                        const uint32_t hi = i == 0 ? r1 : (i == 4 ? r2 : (i == 8 ? r3 : q[(i >> 2) + 1]));
                        const uint32_t v = __funnelshift_r(lo, hi, sh);
                        lo = hi;
                        to[i] = (uint8_t)v;
                        if (i + 1 < my_len) to[i + 1] = (uint8_t)(v >> 8);
                        if (i + 2 < my_len) to[i + 2] = (uint8_t)(v >> 16);
                        if (i + 3 < my_len) to[i + 3] = (uint8_t)(v >> 24);
                    }
                } else {
                    for (unsigned int i = 0; i < my_len; i++) to[i] = i < my_d ? from[i] : to[i - my_d];      // period my_d
                }
            }
            __syncwarp();
            unsigned int dep = any & ~__ballot_sync(0xffffffffu, indep);
            while (dep) {                                                        // in token order: each may read what the one before wrote
                const int w = __ffs(dep) - 1;
                dep &= dep - 1;
                const unsigned int tl = __shfl_sync(0xffffffffu, my_len, w), td = __shfl_sync(0xffffffffu, my_d, w), tp = __shfl_sync(0xffffffffu, my_pos, w);
                for (unsigned int i = lane; i < tl; i += 32) {
                    const unsigned int p = tp - td + (td >= tl ? i : i % td);    // source position in the member (before tp)
                    sp[lead + (tp - pos) + i] = p < pos ? dst[p] : sp[lead + (p - pos)];
                }
                __syncwarp();
            }
            for (unsigned int w = lane; w < nwords; w += 32) {
                const uint32_t v = sw[w];
                const unsigned int b0 = 4u * w;
                if (b0 >= lead && b0 + 4u <= nbytes) g[w] = v;
                else {                                                           // the first / last word of the stretch: only its own bytes
                    uint8_t* gb = reinterpret_cast<uint8_t*>(g + w);
#pragma unroll
                    for (unsigned int q = 0; q < 4; q++) if (b0 + q >= lead && b0 + q < nbytes) gb[q] = (uint8_t)(v >> (8u * q));
                }
            }
            __syncwarp();
        } else if (any) {                                                        // a stretch too long to stage: copies in global memory
            if (indep) {
                const uint8_t* from = dst + my_pos - my_d;
                uint8_t* to = dst + my_pos;
                if (my_d >= my_len) {
                    for (unsigned int i = 0; i < my_len; i += 8) {
                        uint8_t v[8];
#pragma unroll
                        for (int q = 0; q < 8; q++) v[q] = i + q < my_len ? from[i + q] : (uint8_t)0;
#pragma unroll
                        for (int q = 0; q < 8; q++) if (i + q < my_len) to[i + q] = v[q];
                    }
                } else {
                    for (unsigned int i = 0; i < my_len; i++) to[i] = from[i];    // period my_d: reads what this thread has just written
                }
            }
            __syncwarp();
            unsigned int dep = any & ~__ballot_sync(0xffffffffu, indep);
            while (dep) {
                const int w = __ffs(dep) - 1;
                dep &= dep - 1;
                const unsigned int tl = __shfl_sync(0xffffffffu, my_len, w), td = __shfl_sync(0xffffffffu, my_d, w), tp = __shfl_sync(0xffffffffu, my_pos, w);
                const uint8_t* from = dst + tp - td;
                if (td >= tl) { for (unsigned int i = lane; i < tl; i += 32) dst[tp + i] = from[i]; }
                else { for (unsigned int i = lane; i < tl; i += 32) dst[tp + i] = from[i % td]; }
                __syncwarp();
            }
        }
        pos += total;
    }
    __syncwarp();
    if (check_crc && warp_crc32(dst, M.out_len, lane, s_mat[warp][0], s_mat[warp][1], s_crc) != M.crc && lane == 0) status[mi] = 3;
}

}  // namespace kdb
