/* TEST INFRASTRUCTURE — not product code.
 *
 * CPU restatement (plain C) of kmerdb v0.9.6's `profile` hot path, used ONLY as the checker:
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * it.  Nothing under kmerdb_b200/ links, imports or calls it.
 *
 * Parity pinning: this restatement is checked by tests/test_oracle.py against
 *   (1) tests/golden/golden_v1.json — outputs of the UNMODIFIED reference modules
 *       (/root/reference/kmerdb/{kmer,parse,graph,minimizer}.py) run through tests/_bio_shim by
 *       tests/golden/make_golden.py, and
 *   (2) the reference's own fixtures when /root/reference is present (SURVEY Appendix B1/B3/B4:
 *       test/data/test_Cac_ATCC824.8.kdb, example_v0.6.3_kdb_file.8.kdb, appmap.py:271-306).
 * Record tokenising follows Biopython 1.83 (Pipfile.lock:28-60) as restated in the shim; no
 * reference test pins that boundary ("parity unpinned" for FASTA/FASTQ tokenising only).
 */
#ifndef KMERDB_ORACLE_H
#define KMERDB_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum {
    KDBO_OK = 0,
    KDBO_E_ALPHABET = 1,  /* NameError      kmer.py:168-170 via :519            */
    KDBO_E_IUPAC = 2,     /* KeyError       kmer.py:309                         */
    KDBO_E_SHORT = 3,     /* AttributeError kmer.py:461-463 (Seq has no .id)    */
    KDBO_E_FORMAT = 4,    /* ValueError     Bio.SeqIO FASTQ grammar violations  */
    KDBO_E_EMPTY = 5,     /* ValueError     parse.py:144  max([])               */
    KDBO_E_ARG = 6
};

typedef struct {
    uint64_t total_reads;   /* parse.py:153  len(seq_lengths)                     */
    uint64_t total_kmers;   /* parse.py:133-136                                   */
    uint64_t unique_kmers;  /* parse.py:141                                       */
    uint64_t sum_len;       /* for avg_read_length = int(mean) parse.py:146       */
    uint64_t min_len;       /* parse.py:145                                       */
    uint64_t max_len;       /* parse.py:144                                       */
    uint64_t mono[4];       /* A,C,G,T letter counts                 */
    uint64_t err_record;    /* 0-based index of the record that raised            */
    int32_t status;
    int32_t pad_;
} kdbo_stats;

/* kmer.py:234-317.  Returns -1 for a window holding 'N' (None), -2 for KeyError letters. */
int64_t kdbo_kmer_to_id(const char* s, int k, int canonical);

/* parse.py:90-163 on an in-memory, already gunzipped file image.
 * counts: caller-allocated uint64[4^k], zeroed here.  Single thread, reference order. */
int kdbo_profile(const uint8_t* text, size_t n, int is_fastq, int k, int skip_n, int canonical,
                 uint64_t* counts, kdbo_stats* st);

/* Same result, records distributed over `threads` pthreads (CPU baseline timing leg). */
int kdbo_profile_mt(const uint8_t* text, size_t n, int is_fastq, int k, int skip_n, int canonical,
                    uint64_t* counts, kdbo_stats* st, int threads);

/* graph.py:108-354 for ACGT-only records: rows (record index, pos1, id1, pos2, id2).
 * Returns the number of edges (writes at most cap rows), or -(error code). */
int64_t kdbo_edges(const uint8_t* text, size_t n, int is_fastq, int k, int canonical,
                   uint64_t* rec, uint64_t* pos1, uint64_t* id1, uint64_t* pos2, uint64_t* id2, size_t cap);

/* minimizer.py:31-85 reference-compatible rows for one sequence: returns number of rows
 * (4^k-k+1) or -(KDBO_E_FORMAT) for the "WHOOPS" ValueError. */
int64_t kdbo_minimizers_compat(const char* seq, size_t len, int k, int window, uint64_t* ids, uint8_t* is_min, size_t cap);

/* Standard windowed minimizers (extension): for each window start j of w consecutive
 * k-mers, position of the minimum canonical id (leftmost on ties); consecutive duplicates
 * collapsed.  Returns count. */
int64_t kdbo_minimizers_true(const char* seq, size_t len, int k, int w, uint64_t* pos, uint64_t* ids, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
