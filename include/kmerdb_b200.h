/* kmerdb_b200 — C-ABI of the B200-native `profile` hot path of kmerdb v0.9.6.
 *
 * The reference has no FFI: its seam for this path is the Python function
 * kmerdb.parse.parsefile (kmerdb/parse.py:90-163) and the helpers in kmerdb/kmer.py.  The entry
 * points below are what a ctypes binding inside kmerdb/parse.py would call instead of the
 * per-window Python loop (INTEGRATION.md shows that stub).  Plain C types only, caller-allocated
 * outputs, int status returns, no exceptions across the boundary.
 *
 * Threading: one context per GPU; a context is not thread-safe; different contexts are
 * independent.  All calls return after the work they describe is complete (synchronous), like the
 * reference (SURVEY §8b "Threading").
 */
#ifndef KMERDB_B200_H
#define KMERDB_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Status codes.  1..5 map 1:1 onto the exception classes the reference raises for this path. */
enum {
    KDB_OK = 0,
    KDB_E_ALPHABET = 1, /* NameError      kmerdb/kmer.py:168-170 via :519 (letter outside ACGTRYSWKMBDHVN)  */
    KDB_E_IUPAC = 2,    /* KeyError       kmerdb/kmer.py:309 (RYSWKMBDHV reaches the 2-bit encoder)          */
    KDB_E_SHORT = 3,    /* AttributeError kmerdb/kmer.py:461-463 (record shorter than k)                    */
    KDB_E_FORMAT = 4,   /* ValueError     Bio.SeqIO grammar violation (kmerdb/parse.py:70-72)               */
    KDB_E_EMPTY = 5,    /* ValueError     kmerdb/parse.py:144 max([]) on a file with no records             */
    KDB_E_ARG = 6,      /* TypeError/ValueError: bad argument                                               */
    KDB_E_LAYOUT = 7,   /* input uses a text layout the device tokeniser does not take natively (lone CR,
                           multi-line FASTQ, control/non-ASCII bytes ...): the host re-tokenises the file into
                           the plain layout and feeds it again — counting always happens on the GPU.        */
    KDB_E_CUDA = 8,     /* RuntimeError: CUDA failure, see kdb_last_error                                   */
    KDB_E_STATE = 9     /* call order violated (e.g. kdb_counts before kdb_finish)                          */
};

/* Flags for kdb_create. */
enum {
    KDB_CANONICAL = 1, /* id = min(fwd, revcomp)  kmerdb/kmer.py:310-317 (reference default)               */
    KDB_SKIP_N = 2,    /* replace_with_none=True  kmerdb/kmer.py:542-544; absent: 4^n expansion :550-565    */
    KDB_FASTQ = 4,     /* input is FASTQ (kmerdb/util.py:128-134), else FASTA (:120-126)                    */
    KDB_EDGE_KPLUS1 = 8 /* the context counts forward (k+1)-mers for kdb_edge_weights: k passed to kdb_create
                           is k+1, records of exactly k letters are legal (kmerdb/kmer.py:461-463 with the
                           graph's k), so the short-record raise uses k-1                                   */
};

typedef struct kdb_ctx kdb_ctx;

/* Per-file reductions of kmerdb/parse.py:139-160 plus the legacy mononucleotide counts. */
typedef struct {
    uint64_t total_reads;  /* parse.py:153 */
    uint64_t total_kmers;  /* parse.py:133-136 (N expansions included) */
    uint64_t unique_kmers; /* parse.py:141 */
    uint64_t sum_len;      /* sum of record lengths; avg_read_length = int(mean) parse.py:146 */
    uint64_t min_len;      /* parse.py:145 */
    uint64_t max_len;      /* parse.py:144 */
    uint64_t mono[4];      /* A,C,G,T letter counts */
    uint64_t err_record;   /* 0-based index of the first record that raises (valid when status != 0) */
    uint64_t n_windows_with_n; /* windows holding at least one N (a follower of a kdb_attach sweep reports its LEADER's k here) */
    int32_t status;        /* KDB_OK or the class of the first failing record, in the reference's order */
    int32_t pad_;
} kdb_stats_t;

/* Device timing of the passes run so far (CUDA events on the context's stream). */
typedef struct {
    float scan_ms;    /* pass set-up (zeroing the look-back status words); the old K1 scan when KDB_TWO_PASS is set */
    float encode_ms;  /* K2: tokenise + 2-bit encode into the packed letter stream           */
    float count_ms;   /* K3: rolling id + histogram (the dominant kernel)                    */
    float partition_ms; /* part of count_ms: first kernel of K3 (the whole of it when k is not partitioned) */
    float fold_ms;    /* u32 -> u64 fold of the count table                                  */
    float final_ms;   /* record-length check, unique/nullomer reductions                     */
    float h2d_ms;     /* host -> device staging                                              */
    float job_ms;     /* kdb_reset/kdb_create (table zeroing) .. end of kdb_finish, one event pair    */
    float reduce_ms;  /* kdb_reduce_nccl / kdb_reduce_peers: first collective .. root holds the sum   */
    float inflate_ms; /* kdb_feed_bgzf_host: DEFLATE of the BGZF members on the device                */
    uint64_t count_launches; /* number of count-kernel launches                              */
    uint64_t total_launches; /* all kernel launches issued by this context                   */
    uint64_t bytes_fed;
} kdb_timing_t;

/* Replaces the allocation at kmerdb/parse.py:117-121.  k in [1,16].  NULL on failure
 * (kdb_last_error(NULL) then holds the reason). */
kdb_ctx* kdb_create(int device, int k, int flags);
void kdb_destroy(kdb_ctx* ctx);
const char* kdb_last_error(const kdb_ctx* ctx);
/* Device time (CUDA events) of the kernels of this thread's last kdb_minimizers_batch call: the device-resident rate of the
 * stateless K4 entry points, whose wall time is dominated by the host -> device copy of their inputs. */
float kdb_last_kernel_ms(void);

/* Replaces the record loop of kmerdb/parse.py:128-137 (+ kmer.shred kmerdb/kmer.py:489-577).
 * `bytes` is the uncompressed file text, or a fragment of it that starts at a record boundary
 * (start of a '>' / '@' title line) and ends at one; fragments are fed in file order.
 * kdb_feed_host copies from host memory (pinned or pageable); kdb_feed_device takes a pointer
 * already in this device's HBM (read-only, no copy is made). */
int kdb_feed_host(kdb_ctx* ctx, const uint8_t* bytes, size_t n);
int kdb_feed_device(kdb_ctx* ctx, const uint8_t* dev_bytes, size_t n);

/* Compressed input (the reference opens every input through gzip.open, kmerdb/parse.py:63-68).  BGZF
 * (bgzip / Bio.bgzf: independent gzip members of <= 64 KiB of text, size in the 'BC' extra subfield) is inflated ON THE
 * DEVICE (one thread per member decodes, one warp per member resolves the matches), so the compressed bytes are what crosses PCIe.
 * kdb_bgzf_index: host walk of the member headers (RFC 1952 subfields); KDB_E_FORMAT when `raw` is not BGZF.
 * kdb_inflate_bgzf: the text of a BGZF buffer (host -> device -> inflate -> host), for tests and tools; with host_out == NULL
 *   it only reports the text size.  Every member is checked against its ISIZE and CRC-32.
 * kdb_feed_bgzf_host: kdb_feed_host for a whole BGZF file: compressed H2D in pieces, inflate as they arrive, then the text
 *   is tokenised and counted in place.  KDB_E_FORMAT: not BGZF / corrupt member (use the host's gzip, which raises what the
 *   reference raises); the job of the context is reset then (nothing of this or earlier feeds is kept).  `raw` may be
 *   page-locked (the copy then starts before the member headers are walked) or pageable. */
typedef struct { uint64_t in_off, out_off; uint32_t in_len, out_len, crc, pad_; } kdb_bgzf_member_t;
int kdb_bgzf_index(const uint8_t* raw, size_t n, kdb_bgzf_member_t* members, size_t cap, size_t* n_members, uint64_t* text_bytes);
int kdb_inflate_bgzf(int device, const uint8_t* raw, size_t n, uint8_t* host_out, size_t out_cap, size_t* out_len);
int kdb_feed_bgzf_host(kdb_ctx* ctx, const uint8_t* raw, size_t n);

/* Replaces kmerdb/parse.py:139-146: folds partial tables, computes the reductions.  After this
 * call kdb_stats/kdb_counts/kdb_nullomers are valid.  Returns stats.status. */
int kdb_finish(kdb_ctx* ctx);
int kdb_stats(kdb_ctx* ctx, kdb_stats_t* out);
/* out: host uint64[4^k]  (the `counts` array returned at kmerdb/parse.py:163) */
int kdb_counts(kdb_ctx* ctx, uint64_t* out);
/* ascending ids with zero count (parse.py:139-140).  *n receives the total; at most cap are written. */
int kdb_nullomers(kdb_ctx* ctx, uint64_t* out_ids, size_t cap, size_t* n);
/* No counterpart in the reference.  kdb_nullomers for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): allocates nothing, copies nothing, does not
 * synchronise.  dev_counts: DEVICE uint64[n_bins] (e.g. kdb_device_table of a finished context); dev_out_ids: DEVICE
 * uint64[cap], receives the first min(cap, *dev_n) ascending ids with zero count; dev_n: DEVICE uint64[1], the number of zero
 * bins; stream: a cudaStream_t (NULL = default stream) on which three kernels are queued.  The block counts / offsets live in a
 * static device array of the library sized for n_bins <= 4^14 = 268435456 (1.5 MiB per GPU, held until exit): larger n_bins
 * -> KDB_E_ARG with kdb_last_error(NULL) saying so.  Every call shares that array: calls that overlap on the GPU race.  The
 * first call on a device resolves the array's address; make it outside a capture. */
int kdb_nullomers_static(int device, const uint64_t* dev_counts, size_t n_bins, uint64_t* dev_out_ids, size_t cap, uint64_t* dev_n, void* stream);

/* Multi-GPU: the device pointer of the uint64[4^k] partial table and of the
 * additive scalar pack, for the caller's collective (torch.distributed / NCCL sum-reduce), and
 * a call that recomputes the non-additive reductions (unique, nullomers) after the reduce. */
int kdb_device_table(kdb_ctx* ctx, void** dev_counts_u64, size_t* n_bins);
/* The reduce itself behind the C-ABI.  Reference: the cross-file sum of kmerdb/__init__.py:1888-1891 and the
 * header statistics re-derived from the summed vector (:1901-1903).
 * kdb_reduce_nccl: one process per GPU; rank r holds shard r of the input (file order) in a finished context.
 *   kdb_nccl_unique_id fills `out` (cap >= 128 bytes) on one rank; the caller hands the bytes to every rank (any channel).
 *   The call is a collective: ncclReduceScatter of the uint64[4^k] tables, non-zero count of the own slice, one
 *   ncclAllGather of the 16-word scalar packs folded on the device (sums, min / max, first raise in FILE order), slices
 *   sent to rank 0.  It queues everything on the context's stream and returns; afterwards rank 0 holds the summed table
 *   (kdb_counts / kdb_nullomers there), every rank the combined kdb_stats.  4^k must be a multiple of nranks.
 * kdb_reduce_peers: one process, one finished context per GPU with peer access, ctxs[0] the root: a single kernel per GPU
 *   reads its slice of every table over NVLink, adds, counts and stores into the root's table. */
int kdb_nccl_unique_id(void* out, size_t cap);
int kdb_reduce_nccl(kdb_ctx* ctx, const void* unique_id, int rank, int nranks);
int kdb_reduce_peers(kdb_ctx** ctxs, int n);
int kdb_export_partials(kdb_ctx* ctx, uint64_t* out16); /* [reads,kmers,sum_len,mono0..3,nwin_n,min,max,err_key,...] */
int kdb_import_partials(kdb_ctx* ctx, const uint64_t* in16);
int kdb_refinalize(kdb_ctx* ctx);

/* Multi-k sweep (`--minK/--maxK`, kmerdb/__init__.py:1846-1855: the reference runs _profile once per k and
 * re-parses the input every time).  After kdb_attach(leader, follower) — same device and flags, 2 <= k(leader) <
 * k(follower), before any feed — every feed of the leader also counts the follower's k from the leader's packed
 * letter stream: the text is scanned and encoded once for the whole sweep.  Finish the leader first, then each
 * follower; kdb_counts/kdb_stats/kdb_nullomers work on followers as usual; kdb_reset(leader) resets the sweep. */
int kdb_attach(kdb_ctx* leader, kdb_ctx* follower);

/* Start a new file on the same context (tables zeroed, allocations kept). */
int kdb_reset(kdb_ctx* ctx);
int kdb_timing(kdb_ctx* ctx, kdb_timing_t* out);

/* Batched scalar helpers (kmerdb/kmer.py:234-317 kmer_to_id, :320-363 id_to_kmer) on the GPU:
 * n k-mers of k letters each, packed back to back.  ids[i] = id, or UINT64_MAX for a k-mer
 * holding N (None), UINT64_MAX-1 for other non-ACGT letters (KeyError). */
int kdb_kmer_to_id_batch(int device, const char* kmers, size_t n, int k, int canonical, uint64_t* ids);

/* ---- graph (kmerdb/graph.py:108-354) ------------------------------------------------------------------
 * kdb_edge_rows: the rows graph.make_edges_from_kmerids builds (graph.py:150-214) for records already
 * tokenised by the host (graph.make_edges_from_fasta -> parse.parse_sequence_file, graph.py:261): `letters`
 * holds the records back to back, record r = letters[rec_off[r] .. rec_off[r+1]).  Row e = (record index,
 * pos i-1, id[i-1], pos i, id[i]); the two k-mer strings of the .kdbg row (kmerdb/__init__.py:1756-1762) are
 * id_to_kmer of the ids and are rendered by the host writer.  *n_edges receives the edge count (pass cap = 0
 * to size the outputs).  Returns KDB_OK or the class of the first raise: KDB_E_ALPHABET (NameError),
 * KDB_E_SHORT (AttributeError), KDB_E_FORMAT (ValueError: N in graph mode), KDB_E_IUPAC (KeyError),
 * KDB_E_EMPTY; *err_record = index of that record. */
int kdb_edge_rows(int device, const uint8_t* letters, size_t n_letters, const uint64_t* rec_off, size_t n_recs, int k, int canonical,
                  uint64_t* rec, uint64_t* pos1, uint64_t* id1, uint64_t* pos2, uint64_t* id2, size_t cap, size_t* n_edges,
                  uint64_t* err_record);
/* Aggregated edge list (legacy .kdbg columns n1, n2, weight): weight = multiplicity of the ordered
 * id pair over the whole input = Counter of the kdb_edge_rows rows.  `ctx` is a finished context created with
 * k+1, KDB_EDGE_KPLUS1 and without KDB_CANONICAL (it holds the forward (k+1)-mer histogram); `canonical`
 * selects canonical ids for n1/n2.  Rows come in ascending order of their smallest (k+1)-mer. */
int kdb_edge_weights(kdb_ctx* ctx, int canonical, uint64_t* n1, uint64_t* n2, uint64_t* weight, size_t cap, size_t* n);

/* ---- minimizers (kmerdb/minimizer.py:31-85) -------------------------------------------------------------
 * kdb_minimizers_compat: the reference's rows for one sequence: row i = (canonical id of seq[i:i+k] or
 * UINT64_MAX for None, i % window == 0) for i < 4^k - k + 1 (minimizer.py:65); KDB_E_FORMAT = the
 * ValueError("WHOOPS") it raises whenever len(seq) < 4^k (minimizer.py:76-78).
 * kdb_minimizers: true windowed minimizers (extension): for every w consecutive k-mers the position and
 * canonical id of the smallest id, leftmost on ties, consecutive repeats collapsed. */
int kdb_minimizers_compat(int device, const uint8_t* seq, size_t len, int k, uint64_t window, uint64_t* ids, uint8_t* is_min, size_t cap, size_t* n_rows);
int kdb_minimizers(int device, const uint8_t* seq, size_t len, int k, int w, uint64_t* pos, uint64_t* ids, size_t cap, size_t* n);
/* kdb_minimizers for many records (reads) in one call: `letters` / `rec_off` as for kdb_edge_rows; rows (record index,
 * position inside the record, canonical id), record-major; records shorter than k + w - 1 give no row; cap = 0 counts only. */
int kdb_minimizers_batch(int device, const uint8_t* letters, size_t n_letters, const uint64_t* rec_off, size_t n_recs, int k, int w,
                         uint64_t* rec, uint64_t* pos, uint64_t* ids, size_t cap, size_t* n_rows);

/* ---- .kdb body rows (kmerdb/__init__.py:1980-1990) ------------------------------------------
 * Rows "i\ti\tcount\tfrequency\n" for i = first_row .. first_row + n_rows - 1 (counts[j] belongs to row first_row + j),
 * assembled on the GPU.  The text "count\tfrequency" of every DISTINCT count value comes from the caller
 * (uniq_vals ascending; string u = str_bytes[str_off[u] .. str_off[u+1])), formatted exactly as the reference
 * formats np.uint64 and np.float64, so the rows are byte-identical to the reference's.  *n_bytes receives the size
 * of the text; call with out = NULL to size the buffer.  All pointers are host pointers. */
int kdb_format_rows(int device, const uint64_t* counts, uint64_t first_row, uint64_t n_rows, const uint64_t* uniq_vals,
                    const uint32_t* str_off, const char* str_bytes, size_t n_uniq, char* out, size_t cap, size_t* n_bytes);

/* ---- .kdb body back into a count vector (KDBReader._slurp, kmerdb/fileutil.py:308-466, sort=False) ----
 * For row j of the body: l = line.rstrip().split("\t"); 4 columns; x, id, count = int(l[0..2]); x == j;
 * kmer_ids[j] = id; counts[id] = count (fileutil.py:350-365; frequencies = count / 4^k is left to the caller).
 * The reader keeps counts[4^k], kmer_ids[4^k] in HBM; feed the inflated body in pieces cut at row ends (host pointers),
 * in file order; rows beyond 4^k are ignored like the reference's `for j in range(N)`.
 * kdb_reader_finish returns KDB_OK and copies the vectors out (NULL = skip), or the class of the first failing row
 * (*err_row = its index): */
typedef struct kdb_reader kdb_reader;
enum {
    KDB_R_BASE = 32,
    KDB_R_COLUMNS = 33, /* AssertionError fileutil.py:351 (not 4 columns)                                 */
    KDB_R_INT = 34,     /* ValueError     fileutil.py:357-361 (int() of column 1, 2 or 3)                 */
    KDB_R_INDEX = 35,   /* AssertionError fileutil.py:359 (column 1 is not the row number; also a body
                           with fewer than 4^k rows, which the reference meets as a stale line)          */
    KDB_R_RANGE = 36,   /* IndexError     fileutil.py:365 (k-mer id >= 4^k)                               */
    KDB_R_DUP = 37      /* a k-mer id on two rows: the later row wins in the reference; kmerdb never
                           writes such a file and this reader refuses it                                 */
};
kdb_reader* kdb_reader_create(int device, int k);
int kdb_reader_feed(kdb_reader* r, const uint8_t* text, size_t n_bytes);
int kdb_reader_finish(kdb_reader* r, uint64_t* counts, uint64_t* kmer_ids, uint64_t* n_rows, uint64_t* err_row);
void kdb_reader_destroy(kdb_reader* r);

/* ---- Pearson correlation of count profiles (kmerdb/distance.pyx:108-152 `correlation`, one call per pair of samples in
 * kmerdb.distances, kmerdb/__init__.py:733-742) ------------------------------------------------------
 * profiles: host, sample-major [n_samples][n_bins] (the transposed matrix of __init__.py:720).  out: [n_samples][n_samples],
 * out[i][j] = ssxy / sqrt(ssxx * ssyy); NaN where a profile is constant, as the reference's division gives.
 * mode 0: the arithmetic of the reference as compiled (float32 mean, residuals and products — `cdef float x_bar`, distance.pyx:118 —
 *         summed in extended precision); mode 1: float64 throughout (the textbook coefficient). */
int kdb_correlation_matrix(int device, const uint64_t* profiles, size_t n_samples, size_t n_bins, int mode, double* out);

/* Synthetic benchmark input: n_reads records of read_len i.i.d. uniform ACGT letters,
 * FASTA framing ">%09d\n" + letters + "\n" (or FASTQ framing), written straight into HBM.
 * read index = first_read + i; counter-based generator, reproducible on the host
 * (kdb_synth_host writes the same bytes into host memory). */
size_t kdb_synth_record_bytes(int read_len, int fastq);
int kdb_synth_device(int device, uint8_t* dev_out, uint64_t first_read, uint64_t n_reads, int read_len, int fastq, uint64_t seed);
int kdb_synth_host(uint8_t* out, uint64_t first_read, uint64_t n_reads, int read_len, int fastq, uint64_t seed);

/* Microbenchmarks backing the roofline discussion: random red.global.add.u32 into a
 * table of `table_bytes`, and random shared-memory atomics.  Return updates per second. */
int kdb_microbench_atomics(int device, size_t table_bytes, uint64_t n_updates, int mode, double* updates_per_s);

/* Raw device memory helpers so callers without a CUDA binding (ctypes) can stage inputs. */
int kdb_dev_alloc(int device, size_t n, void** out);
int kdb_dev_free(int device, void* p);
int kdb_host_alloc_pinned(size_t n, void** out);
int kdb_host_free_pinned(void* p);
int kdb_memcpy_h2d(int device, void* dst, const void* src, size_t n);
int kdb_memcpy_d2h(int device, void* dst, const void* src, size_t n);

#ifdef __cplusplus
}
#endif
#endif
