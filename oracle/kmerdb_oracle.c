/* TEST INFRASTRUCTURE — CPU restatement of kmerdb v0.9.6's profile hot path (see kmerdb_oracle.h).
 * Every function cites the reference lines it follows (paths relative to /root/reference). */
#include "kmerdb_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- kmer.py:44-49 letterToBinaryNA; :66-80 permitted_NA_characters_with_N ---------------- */
static inline int base_code(uint8_t c) {
    switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}
static inline int is_iupac_other(uint8_t c) { return c && strchr("RYSWKMBDHV", c) != NULL; }
static inline int is_permitted(uint8_t c) { return base_code(c) >= 0 || c == 'N' || is_iupac_other(c); }

/* kmer.py:297-317: fwd id MSB-first; rc id of the reverse complement; min when canonical. */
static inline uint64_t canon_of(uint64_t fwd, int k, int canonical) {
    if (!canonical) return fwd;
    uint64_t rc = 0, f = fwd;
    for (int j = 0; j < k; j++) { rc = (rc << 2) | (3 - (f & 3)); f >>= 2; }
    return rc < fwd ? rc : fwd;
}

int64_t kdbo_kmer_to_id(const char* s, int k, int canonical) {
    uint64_t fwd = 0;
    for (int j = 0; j < k; j++) if (s[j] == 'N') return -1;           /* kmer.py:287-289 */
    for (int j = 0; j < k; j++) {
        int c = base_code((uint8_t)s[j]);
        if (c < 0) return -2;                                          /* kmer.py:309 KeyError */
        fwd = (fwd << 2) | (uint64_t)c;
    }
    return (int64_t)canon_of(fwd, k, canonical);
}

/* ---- record tokenising (Biopython 1.83 behaviour as restated in tests/_bio_shim/Bio/SeqIO.py) -- */
typedef struct { const uint8_t* p; size_t n; size_t pos; } lines_t;

/* Python text mode uses universal newlines: "\n", "\r\n" and lone "\r" all end a line.
 * Returns 0 at EOF; *len excludes the terminator. */
static int next_line(lines_t* L, const uint8_t** s, size_t* len) {
    if (L->pos >= L->n) return 0;
    size_t a = L->pos, i = a;
    while (i < L->n && L->p[i] != '\n' && L->p[i] != '\r') i++;
    *s = L->p + a; *len = i - a;
    if (i < L->n) { if (L->p[i] == '\r' && i + 1 < L->n && L->p[i + 1] == '\n') i += 2; else i += 1; }
    L->pos = i;
    return 1;
}
static inline int is_py_space(uint8_t c) { return c == ' ' || (c >= 9 && c <= 13) || (c >= 0x1c && c <= 0x1f); }
static size_t rstrip_len(const uint8_t* s, size_t len) { while (len && is_py_space(s[len - 1])) len--; return len; }

typedef struct { uint8_t* d; size_t n, cap; } buf_t;
static void buf_put(buf_t* b, const uint8_t* s, size_t len, int drop_sp_cr) {
    if (b->n + len + 1 > b->cap) { b->cap = (b->n + len + 1) * 2; b->d = (uint8_t*)realloc(b->d, b->cap); }
    for (size_t i = 0; i < len; i++) { if (drop_sp_cr && (s[i] == ' ' || s[i] == '\r')) continue; b->d[b->n++] = s[i]; }
}

/* A record sink gets each record's sequence letters in file order; non-zero return aborts. */
typedef int (*rec_fn)(void* ctx, uint64_t rec_idx, const uint8_t* seq, size_t len);

static int parse_fasta(const uint8_t* text, size_t n, rec_fn fn, void* ctx, uint64_t* nrec) {
    lines_t L = {text, n, 0}; const uint8_t* s; size_t len; buf_t b = {0, 0, 0};
    int have = 0, rc = 0; uint64_t idx = 0;
    while (next_line(&L, &s, &len)) {
        if (len > 0 && s[0] == '>') {
            if (have) { rc = fn(ctx, idx++, b.d, b.n); if (rc) break; }
            have = 1; b.n = 0;
        } else if (have) {
            buf_put(&b, s, rstrip_len(s, len), 1);
        }
    }
    if (!rc && have) rc = fn(ctx, idx++, b.d, b.n);
    free(b.d); *nrec = idx; return rc;
}

static int parse_fastq(const uint8_t* text, size_t n, rec_fn fn, void* ctx, uint64_t* nrec) {
    lines_t L = {text, n, 0}; const uint8_t *s, *t; size_t len, tlen; buf_t b = {0, 0, 0};
    int rc = 0; uint64_t idx = 0;
    int more = next_line(&L, &s, &len);
    while (more) {
        if (len == 0 || s[0] != '@') { rc = KDBO_E_FORMAT; break; }
        t = s + 1; tlen = rstrip_len(s + 1, len - 1);
        b.n = 0;
        if (next_line(&L, &s, &len)) buf_put(&b, s, rstrip_len(s, len), 0);
        for (;;) {
            if (!next_line(&L, &s, &len)) { rc = KDBO_E_FORMAT; break; }
            if (len > 0 && s[0] == '+') {
                size_t l2 = rstrip_len(s + 1, len - 1);
                if (l2 && (l2 != tlen || memcmp(s + 1, t, l2) != 0)) rc = KDBO_E_FORMAT;
                break;
            }
            buf_put(&b, s, rstrip_len(s, len), 0);
        }
        if (rc) break;
        for (size_t i = 0; i < b.n; i++) if (b.d[i] == ' ' || b.d[i] == '\t') { rc = KDBO_E_FORMAT; break; }
        if (rc) break;
        size_t qlen = 0; int qbad = 0;
        if (next_line(&L, &s, &len)) { size_t l = rstrip_len(s, len); for (size_t i = 0; i < l; i++) if (s[i] < 33 || s[i] > 126) qbad = 1; qlen += l; }
        for (;;) {
            more = next_line(&L, &s, &len);
            if (!more) break;
            if (len > 0 && s[0] == '@' && qlen >= b.n) break;
            size_t l = rstrip_len(s, len); for (size_t i = 0; i < l; i++) if (s[i] < 33 || s[i] > 126) qbad = 1; qlen += l;
        }
        if (qlen != b.n || qbad) { rc = KDBO_E_FORMAT; break; }
        rc = fn(ctx, idx++, b.d, b.n); if (rc) break;
    }
    free(b.d); *nrec = idx; return rc;
}

/* ---- kmer.shred + parse.parsefile count loop ------------------------------------------------ */
typedef struct {
    int k, skip_n, canonical, atomic;
    uint64_t* counts;
    uint64_t total_kmers, n_reads, sum_len, min_len, max_len, mono[4], err_record;
} prof_t;

static inline void bump(prof_t* P, uint64_t id) {
    if (P->atomic) __atomic_fetch_add(&P->counts[id], 1, __ATOMIC_RELAXED); else P->counts[id]++;
}

/* kmer.py:550-565, :586-621: every ACGT substitution of the N positions gets +1. */
static void expand_n(prof_t* P, const uint8_t* w, int k) {
    int npos[32], nn = 0; uint64_t base = 0;
    for (int j = 0; j < k; j++) { int c = base_code(w[j]); base <<= 2; if (c < 0) npos[nn++] = k - 1 - j; else base |= (uint64_t)c; }
    uint64_t combos = 1ull << (2 * nn);
    for (uint64_t m = 0; m < combos; m++) {
        uint64_t f = base;
        for (int q = 0; q < nn; q++) f |= ((m >> (2 * q)) & 3) << (2 * npos[q]);
        bump(P, canon_of(f, k, P->canonical));
    }
    P->total_kmers += combos;
}

static int profile_record(void* ctx, uint64_t idx, const uint8_t* seq, size_t len) {
    prof_t* P = (prof_t*)ctx; const int k = P->k;
    P->err_record = idx;
    for (size_t i = 0; i < len; i++) if (!is_permitted(seq[i])) return KDBO_E_ALPHABET;   /* kmer.py:519 -> :168-170 */
    if (len < (size_t)k) return KDBO_E_SHORT;                                            /* kmer.py:461-463 */
    const uint64_t mask = (k == 32) ? ~0ull : ((1ull << (2 * k)) - 1);
    uint64_t fwd = 0; size_t since_n = (size_t)k, since_i = (size_t)k;   /* letters since the last N / other-IUPAC letter */
    for (size_t i = 0; i < len; i++) {
        int c = base_code(seq[i]);
        since_n++; since_i++;
        if (c < 0) { if (seq[i] == 'N') since_n = 0; else since_i = 0; c = 0; }
        fwd = ((fwd << 2) | (uint64_t)c) & mask;
        if (i + 1 < (size_t)k) continue;
        int has_n = since_n < (size_t)k, has_i = since_i < (size_t)k;                    /* kmer.py:526-565 */
        if (!has_n) {
            if (has_i) return KDBO_E_IUPAC;                                              /* kmer.py:309 */
            bump(P, canon_of(fwd, k, P->canonical)); P->total_kmers++;                    /* kmer.py:538-540 */
        } else if (!P->skip_n) {
            if (has_i) return KDBO_E_IUPAC;   /* mis-wired doublet branch ends in :309 */
            expand_n(P, seq + i + 1 - k, k);
        }                                                                                /* else kmer.py:542-544 */
    }
    for (size_t i = 0; i < len; i++) { int c = base_code(seq[i]); if (c >= 0) P->mono[c]++; }
    P->n_reads++; P->sum_len += len;                                                      /* parse.py:137 */
    if (len < P->min_len) P->min_len = len;
    if (len > P->max_len) P->max_len = len;
    return 0;
}

static void finish_stats(prof_t* P, int rc, uint64_t nbins, kdbo_stats* st) {
    memset(st, 0, sizeof *st);
    st->status = rc; st->err_record = P->err_record;
    if (rc) return;
    uint64_t u = 0; for (uint64_t i = 0; i < nbins; i++) u += P->counts[i] != 0;          /* parse.py:141 */
    st->total_reads = P->n_reads; st->total_kmers = P->total_kmers; st->unique_kmers = u;
    st->sum_len = P->sum_len; st->min_len = P->min_len; st->max_len = P->max_len;
    memcpy(st->mono, P->mono, sizeof P->mono);
}

int kdbo_profile(const uint8_t* text, size_t n, int is_fastq, int k, int skip_n, int canonical,
                 uint64_t* counts, kdbo_stats* st) {
    if (k < 1 || k > 31 || !counts || !st) return KDBO_E_ARG;
    uint64_t nbins = 1ull << (2 * k), nrec = 0;
    memset(counts, 0, nbins * sizeof(uint64_t));
    prof_t P; memset(&P, 0, sizeof P);
    P.k = k; P.skip_n = skip_n; P.canonical = canonical; P.counts = counts; P.min_len = ~0ull;
    int rc = is_fastq ? parse_fastq(text, n, profile_record, &P, &nrec) : parse_fasta(text, n, profile_record, &P, &nrec);
    if (!rc && nrec == 0) rc = KDBO_E_EMPTY;                                              /* parse.py:144 */
    finish_stats(&P, rc, nbins, st);
    return rc;
}

/* ---- multi-threaded variant (CPU baseline leg): same arithmetic, records split over threads ---- */
typedef struct { buf_t seq; size_t* off; size_t* len; size_t n, cap; } recs_t;
static int collect_record(void* ctx, uint64_t idx, const uint8_t* seq, size_t len) {
    recs_t* R = (recs_t*)ctx; (void)idx;
    if (R->n == R->cap) { R->cap = R->cap ? R->cap * 2 : 1024; R->off = (size_t*)realloc(R->off, R->cap * sizeof(size_t)); R->len = (size_t*)realloc(R->len, R->cap * sizeof(size_t)); }
    R->off[R->n] = R->seq.n; R->len[R->n] = len; R->n++;
    buf_put(&R->seq, seq, len, 0);
    return 0;
}
typedef struct { recs_t* R; prof_t P; size_t lo, hi; int rc; } job_t;
static void* job_main(void* arg) {
    job_t* J = (job_t*)arg;
    for (size_t r = J->lo; r < J->hi; r++) {
        int rc = profile_record(&J->P, r, J->R->seq.d + J->R->off[r], J->R->len[r]);
        if (rc) { J->rc = rc; break; }
    }
    return NULL;
}

int kdbo_profile_mt(const uint8_t* text, size_t n, int is_fastq, int k, int skip_n, int canonical,
                    uint64_t* counts, kdbo_stats* st, int threads) {
    if (k < 1 || k > 31 || !counts || !st) return KDBO_E_ARG;
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    uint64_t nbins = 1ull << (2 * k), nrec = 0;
    memset(counts, 0, nbins * sizeof(uint64_t));
    recs_t R; memset(&R, 0, sizeof R);
    int prc = is_fastq ? parse_fastq(text, n, collect_record, &R, &nrec) : parse_fasta(text, n, collect_record, &R, &nrec);
    /* a tokenising error only matters if no earlier record raises: records before it were collected */
    job_t* J = (job_t*)calloc((size_t)threads, sizeof(job_t));
    pthread_t* T = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        J[t].R = &R; J[t].lo = R.n * (size_t)t / (size_t)threads; J[t].hi = R.n * (size_t)(t + 1) / (size_t)threads;
        J[t].P.k = k; J[t].P.skip_n = skip_n; J[t].P.canonical = canonical; J[t].P.counts = counts;
        J[t].P.atomic = threads > 1; J[t].P.min_len = ~0ull;
        pthread_create(&T[t], NULL, job_main, &J[t]);
    }
    prof_t P; memset(&P, 0, sizeof P); P.counts = counts; P.min_len = ~0ull;
    int rc = 0;
    for (int t = 0; t < threads; t++) {
        pthread_join(T[t], NULL);
        if (J[t].rc && !rc) { rc = J[t].rc; P.err_record = J[t].P.err_record; }   /* lowest block = earliest record */
        P.total_kmers += J[t].P.total_kmers; P.n_reads += J[t].P.n_reads; P.sum_len += J[t].P.sum_len;
        if (J[t].P.min_len < P.min_len) P.min_len = J[t].P.min_len;
        if (J[t].P.max_len > P.max_len) P.max_len = J[t].P.max_len;
        for (int c = 0; c < 4; c++) P.mono[c] += J[t].P.mono[c];
    }
    if (!rc) rc = prc;
    if (!rc && nrec == 0) rc = KDBO_E_EMPTY;
    finish_stats(&P, rc, nbins, st);
    free(J); free(T); free(R.seq.d); free(R.off); free(R.len);
    return rc;
}

/* ---- graph.py:108-354 (ACGT-only records; any N raises) ---------------------------- */
typedef struct { int k, canonical; uint64_t *rec, *pos1, *id1, *pos2, *id2; size_t cap; int64_t n; } edge_t;
static int edges_record(void* ctx, uint64_t idx, const uint8_t* seq, size_t len) {
    edge_t* E = (edge_t*)ctx; const int k = E->k;
    for (size_t i = 0; i < len; i++) if (!is_permitted(seq[i])) return KDBO_E_ALPHABET;
    if (len < (size_t)k) return KDBO_E_SHORT;
    for (size_t i = 0; i < len; i++) { if (seq[i] == 'N') return KDBO_E_FORMAT; if (base_code(seq[i]) < 0) return KDBO_E_IUPAC; }
    const uint64_t mask = (1ull << (2 * k)) - 1; uint64_t fwd = 0, prev = 0;
    for (size_t i = 0; i < len; i++) {
        fwd = ((fwd << 2) | (uint64_t)base_code(seq[i])) & mask;
        if (i + 1 < (size_t)k) continue;
        uint64_t id = canon_of(fwd, k, E->canonical); size_t p = i + 1 - k;
        if (p > 0) {                                   /* graph.py:150-214: edge kmer_ids[i-1] -> kmer_id */
            if ((size_t)E->n < E->cap) { E->rec[E->n] = idx; E->pos1[E->n] = p - 1; E->id1[E->n] = prev; E->pos2[E->n] = p; E->id2[E->n] = id; }
            E->n++;
        }
        prev = id;
    }
    return 0;
}
int64_t kdbo_edges(const uint8_t* text, size_t n, int is_fastq, int k, int canonical,
                   uint64_t* rec, uint64_t* pos1, uint64_t* id1, uint64_t* pos2, uint64_t* id2, size_t cap) {
    edge_t E = {k, canonical, rec, pos1, id1, pos2, id2, cap, 0}; uint64_t nrec = 0;
    int rc = is_fastq ? parse_fastq(text, n, edges_record, &E, &nrec) : parse_fasta(text, n, edges_record, &E, &nrec);
    if (!rc && nrec == 0) rc = KDBO_E_EMPTY;
    return rc ? -(int64_t)rc : E.n;
}

/* ---- minimizer.py:31-85: rows (i, canonical id, i % W == 0) for i < 4^k-k+1; WHOOPS if len < 4^k -- */
int64_t kdbo_minimizers_compat(const char* seq, size_t len, int k, int window, uint64_t* ids, uint8_t* is_min, size_t cap) {
    if (len < (size_t)k) return -(int64_t)KDBO_E_FORMAT;                       /* minimizer.py:59-60 */
    uint64_t L = (1ull << (2 * k)) - (uint64_t)k + 1;                         /* minimizer.py:65 */
    if (len < L + (uint64_t)k - 1) return -(int64_t)KDBO_E_FORMAT;             /* minimizer.py:76-78 */
    for (uint64_t i = 0; i < L; i++) {
        int64_t id = kdbo_kmer_to_id(seq + i, k, 1);                          /* minimizer.py:79 */
        if (id == -2) return -(int64_t)KDBO_E_IUPAC;
        if (i < cap) { ids[i] = (uint64_t)id; is_min[i] = (uint8_t)(i % (uint64_t)window == 0); }
    }
    return (int64_t)L;
}

int64_t kdbo_minimizers_true(const char* seq, size_t len, int k, int w, uint64_t* pos, uint64_t* ids, size_t cap) {
    if (len < (size_t)(k + w - 1)) return 0;
    size_t nk = len - (size_t)k + 1; int64_t n = 0; uint64_t last = ~0ull;
    uint64_t* id = (uint64_t*)malloc(nk * sizeof(uint64_t));
    for (size_t i = 0; i < nk; i++) { int64_t v = kdbo_kmer_to_id(seq + i, k, 1); if (v < 0) { free(id); return -(int64_t)KDBO_E_IUPAC; } id[i] = (uint64_t)v; }
    for (size_t j = 0; j + (size_t)w <= nk; j++) {
        size_t best = j;
        for (size_t q = j + 1; q < j + (size_t)w; q++) if (id[q] < id[best]) best = q;
        if ((uint64_t)best != last) { if ((size_t)n < cap) { pos[n] = best; ids[n] = id[best]; } n++; last = best; }
    }
    free(id);
    return n;
}
