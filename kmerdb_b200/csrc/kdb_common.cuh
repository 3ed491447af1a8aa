// kmerdb_b200 device-side common definitions (sm_100a only).
//
// Data layout in HBM:
//   text      : the uncompressed FASTA/FASTQ bytes of one feed, contiguous, 16-byte aligned
//   look/trec : decoupled look-back status word and border-record lengths per 16 KiB tile (k2s_encode)
//   tab32     : uint32[4^k] working histogram (red.global.add.u32 target; 64 MiB at k=12, L2-resident)
//   tab64     : uint64[4^k] result histogram (tab32 is folded into it before it can overflow)
//   acc       : Accum, the fused scalar reductions
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace kdb {

constexpr int kThreads = 256;          // threads per CTA in the scan and count kernels
constexpr int kChunk = 64;             // text bytes owned by one thread (4 x 16-byte loads)
constexpr int kTile = kThreads * kChunk;  // 16 KiB of text per tile

// Byte classes inside a sequence line.  kmerdb/kmer.py:44-49 (ACGT codes), :66-80 (permitted set).
enum : uint8_t {
    C_A = 0, C_C = 1, C_G = 2, C_T = 3,
    C_N = 4,        // ambiguous: kmer.py:287-289 / :550-565
    C_IUPAC = 5,    // RYSWKMBDHV: KeyError at kmer.py:309
    C_ILLEGAL = 6,  // any other printable ASCII: NameError at kmer.py:168-170
    C_SPACE = 7,    // dropped by the FASTA tokeniser (Bio.SeqIO: .replace(" ", ""))
    C_LF = 8,
    C_CR = 9,       // only legal directly before LF
    C_WEIRD = 10    // control / DEL / >=0x80: host re-tokenises (KDB_E_LAYOUT)
};

// Line kinds.
enum : uint32_t { K_HEADER = 0, K_SEQ = 1, K_IGNORE = 2, K_QUAL = 3 };

// Priority of a raise inside one record, in the order the reference meets them:
// is_sequence_na over the whole record (kmer.py:519) -> length check (:461-463) -> windows (:309).
enum : uint32_t { P_ALPHABET = 0, P_SHORT = 1, P_IUPAC = 2 };

// ---- single-pass tokeniser (k2s_encode): decoupled look-back over the tiles of a pass --------------------------
// One status word per tile, written and read whole (no fence is needed around it): [63:62] state (0 nothing, 1 tile aggregate,
// 2 inclusive prefix of the pass), [61:60] `last`, [59:30] line starts, [29:0] title lines.  A pass is at most 2^29 bytes
// (kPassBytes), so both counts fit their 30 bits and the two fields add without carrying into each other.
struct TileLook {
    unsigned long long w0;
};
constexpr unsigned long long kPassBytes = 1ull << 29;
// What a tile knows about the records that cross its borders (k2s_encode -> k4_stitch_records).
struct TileRec {
    unsigned int head_len, head_qlen;   // letters / quality chars from the tile start to its first record start
    unsigned int tail_len, tail_qlen;   // ... of the record open at the tile end (whole tile when no record starts in it)
    unsigned int flags;                 // bit 0: a record starts in the tile; bit 1: a record was open before that first start
    unsigned int pad;
    unsigned long long head_rec;        // id of the record the first start closes
};
// Device-resident state carried from pass to pass and feed to feed (two copies: pass p reads [p & 1], writes [(p + 1) & 1]).
struct PassCarry {
    unsigned long long ls_total;        // line starts so far in this feed
    unsigned long long hdr_total;       // FASTA title lines so far in the file
    unsigned long long fq_rec_base;     // FASTQ records in earlier feeds
    unsigned long long olen, oqlen;     // letters / quality chars of the record that is still open
    unsigned int last;                  // class of the last line start (0 none, 1 other, 2 '>')
    unsigned int open;                  // a record is open
};

// Fused scalar reductions (one instance per context, device resident).
struct Accum {
    unsigned long long total_kmers;
    unsigned long long sum_len;
    unsigned long long mono[4];
    unsigned long long n_windows_with_n;
    unsigned long long min_len;
    unsigned long long max_len;
    unsigned long long err_key;      // min over raises of (record index * 4 + priority); ~0 = none
    unsigned long long unique;
    unsigned int layout_flag;        // != 0: the text needs the host tokeniser
    unsigned int pad;
};

__device__ __forceinline__ uint32_t revcomp32(uint32_t f, int k) {
    uint32_t x = __brev(~f);
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    return x >> (32 - 2 * k);
}

// Filled by the host at context creation (kdb_api.cu: build_class_lut).  One translation unit
// includes every kernel header, so a plain __constant__ definition is enough.
__constant__ uint8_t c_class_lut[256];

}  // namespace kdb
