// kmerdb_b200 — context and C-ABI (include/kmerdb_b200.h) over the sm_100a kernels.
// One translation unit: every kernel header is included here.
#include "../../include/kmerdb_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "kdb_common.cuh"
#include "kdb_scan.cuh"
#include "kdb_encode.cuh"
#include "kdb_encode2.cuh"
#include "kdb_hist.cuh"
#include "kdb_part.cuh"
#include "kdb_final.cuh"
#include "kdb_synth.cuh"
#include "kdb_graph.cuh"
#include "kdb_rows.cuh"
#include "kdb_read.cuh"
#include "kdb_corr.cuh"
#include "kdb_multi.cuh"
#include "kdb_inflate.cuh"
#include "kdb_inflate2.cuh"

using namespace kdb;

static thread_local std::string g_last_error;
static thread_local float g_last_kernel_ms = 0.f;       // device time of the kernels of the last stateless K4 call (kdb_last_kernel_ms)

struct kdb_ctx {
    int device = 0, k = 0, flags = 0;
    int short_k = 0;                            // records shorter than this raise (k, or k-1 for a KDB_EDGE_KPLUS1 context)
    bool fastq = false, canonical = false, skip_n = false, smem_hist = false;
    unsigned long long nbins = 0;
    cudaStream_t s_compute = nullptr, s_copy = nullptr;
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_job0 = nullptr, ev_job1 = nullptr;
    std::vector<cudaEvent_t> ev_copy;
    unsigned int* tab32 = nullptr;
    unsigned long long* tab64 = nullptr;
    Accum* acc = nullptr;
    uint8_t* feed_buf = nullptr;
    size_t feed_cap = 0;
    uint32_t* packed = nullptr;                 // K2 output / K3 input, one region per tile of the pass
    size_t packed_cap = 0;
    k3_fn hist_kernel = nullptr;
    size_t hist_dyn = 0;
    int encode_grid = 0, hist_grid = 0;
    // bucket-partitioned count (8 <= k <= 12)
    bool partitioned = false;
    k3p_fn part_kernel = nullptr;
    int part_grid = 0, n_buckets = 0, parts = 1, hist_threads = 0;
    int subs = 1;
    bool wide = false;
    bool slotted = false;                       // k3s_partition (buckets = low id bits) instead of the two-sweep k3p_partition
    uint16_t* stream = nullptr;
    size_t stream_cap = 0;
    unsigned int* cursor = nullptr;             // n_buckets cursors + n_buckets fail bases
    // single-pass tokeniser (k2s_encode): look-back status words, border records, ticket, device-resident carry
    TileLook* look = nullptr;                   // n_tiles status words, then n_tiles TileRec (one allocation)
    TileRec* trec = nullptr;
    size_t look_cap = 0;
    unsigned int* ticket = nullptr;
    PassCarry* pcarry = nullptr;                // two copies: a pass reads [pc_idx], writes [pc_idx ^ 1]
    int pc_idx = 0;
    uint8_t* zero_row = nullptr;                // 64 zero bytes: tensor map base for a pass shorter than one row
    // multi-GPU (kdb_reduce_nccl / kdb_reduce_peers)
    ncclComm_t comm = nullptr;
    int comm_rank = -1, comm_n = 0;
    unsigned long long* d_pack = nullptr;       // 16 words of this rank + 16 per rank gathered
    cudaEvent_t ev_red0 = nullptr, ev_red1 = nullptr;
    bool stats_dirty = false;                   // the device scalars changed after kdb_finish (a reduce): kdb_stats re-reads them
    bool reduce_timed = false;
    bool not_root = false;                      // after a reduce only the root holds the whole table
    uint8_t* comp_buf = nullptr;                // kdb_feed_bgzf_host: the compressed bytes, member table, status words
    size_t comp_cap = 0;
    uint8_t* tok_buf = nullptr;                 // kdb_feed_bgzf_host: match tokens + token counts of the pieces in flight (kdb_inflate2.cuh)
    size_t tok_cap = 0;
    std::vector<cudaStream_t> s_inf;            // kdb_feed_bgzf_host: the streams the pieces are inflated on
    std::vector<cudaEvent_t> ev_inf;            // ... and one "piece is text" event per piece
    std::vector<cudaEvent_t> ev_chunk;          // ... the chunks of an early copy (pinned input); the last one marks its start
    void* null_scratch = nullptr;               // kdb_nullomers: per-block zero counts and offsets
    size_t null_cap = 0;
    // host mirrors
    unsigned long long total_reads_host = 0;    // records of the job (read back from the device carry at kdb_finish / after a reduce)
    unsigned long long bytes_since_fold = 0;
    bool job_started = false, finished = false, host_layout = false;
    bool tab32_clean = false;                   // tab32 is all zero (every fold leaves it so): reset skips the 4^k x 4 B memset
    bool tab64_valid = false;                   // tab64 holds folded counts (it is not zeroed at reset: the first fold writes it)
    int count_grid = 0;
    kdb_stats_t stats{};
    kdb_timing_t timing{};
    std::string err;
    // multi-k sweep (kdb_attach): followers count larger k from this context's packed stream
    std::vector<kdb_ctx*> followers;
    kdb_ctx* leader = nullptr;
    int k_halo = 0;
};

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            char b_[512];                                                                          \
            snprintf(b_, sizeof b_, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
            if (ctx) ctx->err = b_;                                                                \
            g_last_error = b_;                                                                     \
            return KDB_E_CUDA;                                                                     \
        }                                                                                          \
    } while (0)

struct DevBuf {                      // scoped device allocation: freed on every exit path, CK() returns included
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 16); }
    template <typename T> T* as() { return static_cast<T*>(p); }
};

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*tensor_map_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                         const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tensor_map_encode_fn tensor_map_encoder() {
    static tensor_map_encode_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (tensor_map_encode_fn)p;
    }
    return fn;
}
// The text of one pass as rows of 64 bytes; box = one 16 KiB tile (256 rows); 64-byte swizzle = the layout k2s_encode reads with
// conflict-free LDS.128.  Only whole rows are part of the tensor: rows past the end arrive as zeros.
static bool make_text_map(CUtensorMap* m, const uint8_t* base, unsigned long long rows) {
    const cuuint64_t dims[2] = {(cuuint64_t)kChunk, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)kChunk};
    const cuuint32_t box[2] = {(cuuint32_t)kChunk, (cuuint32_t)kThreads};
    const cuuint32_t estr[2] = {1, 1};
    return tensor_map_encoder()(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static void build_class_lut(uint8_t* lut) {
    for (int c = 0; c < 256; c++) {
        uint8_t v;
        if (c >= 0x80 || c == 0x7f) v = C_WEIRD;
        else if (c == '\n') v = C_LF;
        else if (c == '\r') v = C_CR;
        else if (c == ' ') v = C_SPACE;
        else if (c < 0x20) v = C_WEIRD;
        else v = C_ILLEGAL;
        lut[c] = v;
    }
    lut['A'] = C_A; lut['C'] = C_C; lut['G'] = C_G; lut['T'] = C_T; lut['N'] = C_N;
    for (const char* p = "RYSWKMBDHV"; *p; p++) lut[(unsigned char)*p] = C_IUPAC;   // kmerdb/kmer.py:66-80
}

static int set_device(kdb_ctx* ctx, int device) {
    CK(cudaSetDevice(device));
    return KDB_OK;
}

template <typename K>
static int occupancy_grid(K kernel, int threads, size_t dyn_smem, int* grid) {
    kdb_ctx* ctx = nullptr;
    int per_sm = 0, sms = 0, dev = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, dyn_smem));
    *grid = std::max(1, per_sm) * sms;   // persistent CTAs: a whole number of CTAs per SM
    return KDB_OK;
}

extern "C" float kdb_last_kernel_ms(void) { return g_last_kernel_ms; }
extern "C" const char* kdb_last_error(const kdb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

extern "C" void kdb_destroy(kdb_ctx* ctx) {
    if (!ctx) return;
    for (kdb_ctx* f : ctx->followers) f->leader = nullptr;
    if (ctx->leader) { auto& v = ctx->leader->followers; v.erase(std::remove(v.begin(), v.end(), ctx), v.end()); }
    cudaSetDevice(ctx->device);
    cudaFree(ctx->tab32); cudaFree(ctx->tab64); cudaFree(ctx->acc);
    cudaFree(ctx->feed_buf); cudaFree(ctx->packed); cudaFree(ctx->stream); cudaFree(ctx->cursor);
    cudaFree(ctx->look); /* trec lives inside the look allocation */ cudaFree(ctx->ticket); cudaFree(ctx->pcarry); cudaFree(ctx->zero_row); cudaFree(ctx->null_scratch); cudaFree(ctx->d_pack); cudaFree(ctx->comp_buf); cudaFree(ctx->tok_buf);
    if (ctx->comm && nccl_api().ok) nccl_api().CommDestroy(ctx->comm);
    if (ctx->ev_red0) cudaEventDestroy(ctx->ev_red0);
    if (ctx->ev_red1) cudaEventDestroy(ctx->ev_red1);
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    for (auto& e : ctx->ev_copy) cudaEventDestroy(e);
    for (auto& e : ctx->ev_inf) cudaEventDestroy(e);
    for (auto& e : ctx->ev_chunk) cudaEventDestroy(e);
    for (auto& st : ctx->s_inf) cudaStreamDestroy(st);
    if (ctx->ev_job0) cudaEventDestroy(ctx->ev_job0);
    if (ctx->ev_job1) cudaEventDestroy(ctx->ev_job1);
    if (ctx->s_compute) cudaStreamDestroy(ctx->s_compute);
    if (ctx->s_copy) cudaStreamDestroy(ctx->s_copy);
    delete ctx;
}

static int reset_state(kdb_ctx* ctx) {
    CK(cudaEventRecord(ctx->ev_job0, ctx->s_compute));   // a job starts with zeroing its tables
    if (!ctx->tab32_clean) CK(cudaMemsetAsync(ctx->tab32, 0, ctx->nbins * sizeof(unsigned int), ctx->s_compute));
    ctx->tab32_clean = true;
    Accum a{};
    a.min_len = ~0ull; a.err_key = ~0ull;
    CK(cudaMemcpyAsync(ctx->acc, &a, sizeof a, cudaMemcpyHostToDevice, ctx->s_compute));
    CK(cudaMemsetAsync(ctx->pcarry, 0, 2 * sizeof(PassCarry), ctx->s_compute));
    ctx->pc_idx = 0;
    CK(cudaStreamSynchronize(ctx->s_compute));
    ctx->total_reads_host = 0; ctx->bytes_since_fold = 0;
    ctx->job_started = true; ctx->finished = false; ctx->host_layout = false; ctx->tab64_valid = false;
    ctx->stats = kdb_stats_t{}; ctx->timing = kdb_timing_t{};
    ctx->stats_dirty = false; ctx->reduce_timed = false; ctx->not_root = false;
    return KDB_OK;
}

extern "C" int kdb_reset(kdb_ctx* ctx) {
    if (!ctx) return KDB_E_ARG;
    if (ctx->leader) { ctx->err = "kdb_reset on a follower: reset its leader"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    for (kdb_ctx* f : ctx->followers) if (int rc = reset_state(f)) return rc;
    return reset_state(ctx);
}

// Multi-k sweep (kmerdb/__init__.py:1853-1855 runs _profile once per k and re-parses the file each time): a
// follower counts its larger k from the leader's packed letter stream, so the text is scanned and encoded once.
extern "C" int kdb_attach(kdb_ctx* leader, kdb_ctx* follower) {
    kdb_ctx* ctx = leader;
    if (!leader || !follower || leader == follower) return KDB_E_ARG;
    if (leader->leader || follower->leader || !follower->followers.empty()) { ctx->err = "kdb_attach: contexts are already part of a sweep"; return KDB_E_STATE; }
    if (leader->device != follower->device || (leader->flags & 7) != (follower->flags & 7)) { ctx->err = "kdb_attach: device and CANONICAL/SKIP_N/FASTQ flags must match"; return KDB_E_ARG; }
    if (leader->k < 2 || follower->k <= leader->k) { ctx->err = "kdb_attach: needs 2 <= k(leader) < k(follower)"; return KDB_E_ARG; }
    if (leader->timing.bytes_fed || follower->timing.bytes_fed || leader->finished || follower->finished) { ctx->err = "kdb_attach after data was fed"; return KDB_E_STATE; }
    leader->followers.push_back(follower);
    follower->leader = leader;
    leader->k_halo = std::max(leader->k_halo, follower->k);
    return KDB_OK;
}

extern "C" kdb_ctx* kdb_create(int device, int k, int flags) {
    if (k < 1 || k > 16) { g_last_error = "k must be in [1,16] for the dense 4^k device table"; return nullptr; }
    kdb_ctx* ctx = new kdb_ctx();
    ctx->device = device; ctx->k = k; ctx->flags = flags;
    ctx->fastq = flags & KDB_FASTQ; ctx->canonical = flags & KDB_CANONICAL; ctx->skip_n = flags & KDB_SKIP_N;
    ctx->smem_hist = k <= 8;
    ctx->short_k = (flags & KDB_EDGE_KPLUS1) ? k - 1 : k;
    ctx->nbins = 1ull << (2 * k);
    auto fail = [&](const char* what, cudaError_t e) -> kdb_ctx* {
        g_last_error = std::string(what) + ": " + cudaGetErrorString(e);
        kdb_destroy(ctx);
        return nullptr;
    };
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
    uint8_t lut[256];
    build_class_lut(lut);
    if ((e = cudaMemcpyToSymbol(c_class_lut, lut, 256)) != cudaSuccess) return fail("cudaMemcpyToSymbol", e);
    if ((e = cudaStreamCreateWithFlags(&ctx->s_compute, cudaStreamNonBlocking)) != cudaSuccess) return fail("stream", e);
    if ((e = cudaStreamCreateWithFlags(&ctx->s_copy, cudaStreamNonBlocking)) != cudaSuccess) return fail("stream", e);
    for (auto& ev : ctx->ev) if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail("event", e);
    if ((e = cudaEventCreate(&ctx->ev_job0)) != cudaSuccess) return fail("event", e);
    if ((e = cudaEventCreate(&ctx->ev_job1)) != cudaSuccess) return fail("event", e);
    if ((e = cudaMalloc(&ctx->tab32, ctx->nbins * sizeof(unsigned int))) != cudaSuccess) return fail("cudaMalloc tab32", e);
    if ((e = cudaMalloc(&ctx->tab64, ctx->nbins * sizeof(unsigned long long))) != cudaSuccess) return fail("cudaMalloc tab64", e);
    if ((e = cudaMalloc(&ctx->acc, sizeof(Accum))) != cudaSuccess) return fail("cudaMalloc acc", e);
    if ((e = cudaMalloc(&ctx->pcarry, 2 * sizeof(PassCarry))) != cudaSuccess) return fail("cudaMalloc pcarry", e);
    if ((e = cudaMalloc(&ctx->ticket, sizeof(unsigned int))) != cudaSuccess) return fail("cudaMalloc ticket", e);
    if ((e = cudaMalloc(&ctx->zero_row, 256)) != cudaSuccess) return fail("cudaMalloc zero_row", e);
    if ((e = cudaMemset(ctx->zero_row, 0, 256)) != cudaSuccess) return fail("cudaMemset zero_row", e);
    if (!tensor_map_encoder()) { g_last_error = "cuTensorMapEncodeTiled is not available from this driver"; kdb_destroy(ctx); return nullptr; }
    // launch geometry: persistent CTAs, a whole number per SM
    int rc;
    ctx->hist_kernel = k3_kernel(k, ctx->canonical);
    ctx->hist_dyn = ctx->smem_hist ? ((k == 8 ? ctx->nbins / 2 : ctx->nbins) + 32) * sizeof(unsigned int) : 0;     // + 32 per-lane dummy words (kdb_hist.cuh)
    if (ctx->hist_dyn > 48 * 1024 &&
        (e = cudaFuncSetAttribute((const void*)ctx->hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->hist_dyn)) != cudaSuccess)
        return fail("smem attr", e);
    ctx->hist_threads = k == 8 ? 2 * kHistThreads : kHistThreads;      // k = 8: one 128 KiB table per SM, two tiles side by side
    rc = occupancy_grid(ctx->hist_kernel, ctx->hist_threads, ctx->hist_dyn, &ctx->hist_grid);
    ctx->partitioned = k >= 9 && k <= 14 && !getenv("KDB_NO_PARTITION");
    if (ctx->partitioned) {
        ctx->slotted = k <= 12 && !getenv("KDB_TWO_SWEEP");
        ctx->part_kernel = ctx->slotted ? k3s_kernel(k, ctx->canonical) : k3p_kernel(k, ctx->canonical);
        ctx->n_buckets = 1 << (2 * k - payload_bits(k));
        ctx->subs = payload_bits(k) >= 16 ? 0 : 1;                                          // 0: packed 16-bit counters (k3p_count16 / k3p_count16w)
        ctx->wide = payload_bits(k) > 16;                                                   // 32-bit stream entries (k = 14)
        ctx->parts = std::max(1, ((payload_bits(k) == 15 ? 888 : 1776) + ctx->n_buckets - 1) / ctx->n_buckets);
        const size_t dyn = ctx->slotted ? k3s_smem_bytes(k) : k3p_smem_bytes(k);
        if ((e = cudaFuncSetAttribute((const void*)ctx->part_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn)) != cudaSuccess) return fail("smem attr", e);
        if ((e = cudaFuncSetAttribute((const void*)k3p_count<14>, cudaFuncAttributeMaxDynamicSharedMemorySize, (4 << 14) + 128)) != cudaSuccess) return fail("smem attr", e);
        if ((e = cudaFuncSetAttribute((const void*)k3p_count<15>, cudaFuncAttributeMaxDynamicSharedMemorySize, (4 << 15) + 128)) != cudaSuccess) return fail("smem attr", e);
        if ((e = cudaFuncSetAttribute((const void*)k3p_count16, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024)) != cudaSuccess) return fail("smem attr", e);
        if ((e = cudaFuncSetAttribute((const void*)k3p_count16w, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024)) != cudaSuccess) return fail("smem attr", e);
        if (!rc) rc = occupancy_grid(ctx->part_kernel, kPartThreads, dyn, &ctx->part_grid);
        if ((e = cudaMalloc(&ctx->cursor, 2 * ctx->n_buckets * sizeof(unsigned int))) != cudaSuccess) return fail("cudaMalloc cursor", e);
    }
    if (!rc) rc = ctx->fastq ? occupancy_grid(k2s_encode<true>, kThreads, 0, &ctx->encode_grid) : occupancy_grid(k2s_encode<false>, kThreads, 0, &ctx->encode_grid);
    // experiment knobs (scripts/concurrency_probe.py): CTAs per SM of the tokeniser / the partition, so that two contexts can share the SMs
    if (const char* e1 = getenv("KDB_ENC_CTAS")) ctx->encode_grid = std::max(1, atoi(e1)) * 148;
    if (const char* e2 = getenv("KDB_PART_CTAS")) ctx->part_grid = std::max(1, atoi(e2)) * 148;
    if (rc) { kdb_destroy(ctx); return nullptr; }
    if (reset_state(ctx)) { g_last_error = ctx->err; kdb_destroy(ctx); return nullptr; }
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("kdb_create", e);
    return ctx;
}

// ---- passes -----------------------------------------------------------------------------------------------
static int ensure(kdb_ctx* ctx, void** p, size_t* cap, size_t need, size_t elem) {
    if (need <= *cap) return KDB_OK;
    if (*p) CK(cudaFree(*p));
    *p = nullptr; *cap = 0;
    const size_t n = need + need / 4 + 64;
    CK(cudaMalloc(p, n * elem));
    *cap = n;
    return KDB_OK;
}

// tab64 += tab32 (or tab64 = tab32 for the first fold of a job), tab32 = 0.  count_unique: the fold that ends the
// job also counts the non-zero bins (acc->unique must be zero before).
static int fold_now(kdb_ctx* ctx, bool count_unique = false, cudaStream_t st = nullptr) {
    if (!st) st = ctx->s_compute;
    if (ctx->bytes_since_fold == 0 && !(count_unique || !ctx->tab64_valid)) return KDB_OK;
    if (ctx->bytes_since_fold == 0 && !ctx->tab64_valid) {      // nothing was counted: the result is all zeros
        CK(cudaMemsetAsync(ctx->tab64, 0, ctx->nbins * sizeof(unsigned long long), st));
        ctx->tab64_valid = true;
        if (!count_unique) return KDB_OK;
    }
    CK(cudaEventRecord(ctx->ev[4], st));
    const unsigned long long quads = (ctx->nbins + 3) / 4;
    const int grid = (int)std::min<unsigned long long>((quads + 255) / 256, 148ull * 16);
    const bool first = !ctx->tab64_valid;
    if (first && count_unique) k3_fold_final<true, true><<<grid, 256, 0, st>>>(ctx->tab32, ctx->tab64, ctx->nbins, ctx->acc);
    else if (first) k3_fold_final<true, false><<<grid, 256, 0, st>>>(ctx->tab32, ctx->tab64, ctx->nbins, ctx->acc);
    else if (count_unique) k3_fold_final<false, true><<<grid, 256, 0, st>>>(ctx->tab32, ctx->tab64, ctx->nbins, ctx->acc);
    else k3_fold<<<grid, 256, 0, st>>>(ctx->tab32, ctx->tab64, ctx->nbins);
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->ev[5], st));
    CK(cudaEventSynchronize(ctx->ev[5]));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev[4], ctx->ev[5]));
    ctx->timing.fold_ms += ms;
    ctx->timing.total_launches++;
    ctx->bytes_since_fold = 0;
    ctx->tab64_valid = true;
    ctx->tab32_clean = true;
    return KDB_OK;
}

// K3 / K3p of one pass: ids from `src`'s packed stream into `ctx`'s tables (ctx == src, or a follower of src that
// counts a larger k from the same stream).  Launched on src's stream; ctx->ev[7] marks the end of the first kernel.
static int count_stage(kdb_ctx* ctx, kdb_ctx* src, unsigned int n_tiles, unsigned long long lo, unsigned long long hi) {
    const int smear = ctx->k - src->k;
    cudaStream_t st = src->s_compute;
    ctx->tab32_clean = false;
    if (ctx != src && ctx->bytes_since_fold + (hi - lo) > 0x7fffffffull) if (int rc = fold_now(ctx, false, st)) return rc;
    if (ctx->partitioned) {
        // every window ends on a text byte, so (hi - lo) bounds the entries of the pass; canonical ids of
        // uniform sequence load the lowest bucket with 2x the average, hence the 3x head room
        const unsigned long long avg = (hi - lo) / ctx->n_buckets + 1;
        // (low-bit buckets of the slotted partition are uniform on uniform sequence: 2x head room)
        const unsigned int cap = (unsigned int)(((ctx->slotted ? 2 : 3) * avg + kTile + 7) & ~7ull);   // >= kTile: keeps s_delta (kdb_part.cuh) from wrapping
        if ((unsigned long long)cap * ctx->n_buckets >= 0xFFFFFFFFull) { ctx->err = "bucket stream index exceeds 32 bits"; return KDB_E_ARG; }
        // the bucket streams live in the leader: the k of a sweep are counted one after the other on one stream
        if (int rc = ensure(src, (void**)&src->stream, &src->stream_cap, (size_t)cap * ctx->n_buckets * (ctx->wide ? 2 : 1), sizeof(uint16_t))) { ctx->err = src->err; return rc; }
        CK(cudaMemsetAsync(ctx->cursor, 0, ctx->n_buckets * sizeof(unsigned int), st));
        CK(cudaMemsetAsync(ctx->cursor + ctx->n_buckets, 0xFF, ctx->n_buckets * sizeof(unsigned int), st));
        PartParams Q{};
        Q.packed = src->packed; Q.n_tiles = n_tiles; Q.skip_n = ctx->skip_n; Q.smear = smear; Q.tab32 = ctx->tab32; Q.stream = src->stream;
        Q.cursor = ctx->cursor; Q.fail_base = ctx->cursor + ctx->n_buckets; Q.cap = cap; Q.acc = ctx->acc;
        const int grid = (int)std::min<unsigned int>((n_tiles + 1) / 2, (unsigned int)ctx->part_grid);   // work unit = two tiles
        ctx->part_kernel<<<grid, kPartThreads, ctx->slotted ? k3s_smem_bytes(ctx->k) : k3p_smem_bytes(ctx->k), st>>>(Q);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ctx->ev[7], st));
        if (ctx->wide) k3p_count16w<<<ctx->n_buckets << (payload_bits(ctx->k) - 16), kCount16Threads, 128 * 1024, st>>>(reinterpret_cast<const uint32_t*>(src->stream), Q.cursor, Q.fail_base, cap, payload_bits(ctx->k), ctx->tab32);
        else if (ctx->subs == 0) k3p_count16<<<ctx->n_buckets, kCount16Threads, 128 * 1024, st>>>(src->stream, Q.cursor, Q.fail_base, cap, ctx->tab32);
        else if (payload_bits(ctx->k) == 15) k3p_count<15><<<ctx->n_buckets * ctx->parts, 1024, (4 << 15) + 128, st>>>(src->stream, Q.cursor, Q.fail_base, cap, ctx->parts, ctx->tab32, ctx->slotted ? 2 * ctx->k - 15 : 0);
        else k3p_count<14><<<ctx->n_buckets * ctx->parts, kBucketThreads, (4 << 14) + 128, st>>>(src->stream, Q.cursor, Q.fail_base, cap, ctx->parts, ctx->tab32, ctx->slotted ? 2 * ctx->k - 14 : 0);
        CK(cudaGetLastError());
        ctx->timing.total_launches++;
    } else {
        HistParams H{};
        H.packed = src->packed; H.n_tiles = n_tiles; H.skip_n = ctx->skip_n; H.smear = smear; H.tab32 = ctx->tab32; H.acc = ctx->acc;
        const unsigned int sides = ctx->hist_threads / kHistThreads;
        const int grid = (int)std::min<unsigned int>((n_tiles + sides - 1) / sides, (unsigned int)ctx->hist_grid);
        ctx->hist_kernel<<<grid, ctx->hist_threads, ctx->hist_dyn, st>>>(H);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ctx->ev[7], st));
    }
    if (ctx != src) {
        ctx->timing.count_launches++;
        ctx->timing.total_launches++;
        ctx->bytes_since_fold += hi - lo;
    }
    return KDB_OK;
}

// One pass with the single-pass tokeniser: K2s (ticket order, look-back, TMA staging) -> count stage -> border records.
// Nothing is read back in the middle: the carry between passes lives on the device.
static int run_pass2(kdb_ctx* ctx, const uint8_t* text, unsigned long long lo, unsigned long long hi, unsigned long long n) {
    const unsigned int n_tiles = (unsigned int)((hi - lo + kTile - 1) / kTile);
    if (int rc = ensure(ctx, (void**)&ctx->look, &ctx->look_cap, n_tiles, sizeof(TileLook) + sizeof(TileRec))) return rc;
    ctx->trec = reinterpret_cast<TileRec*>(ctx->look + ctx->look_cap);
    if (ctx->bytes_since_fold + (hi - lo) > 0x7fffffffull) if (int rc = fold_now(ctx)) return rc;
    if (int rc = ensure(ctx, (void**)&ctx->packed, &ctx->packed_cap, (size_t)n_tiles * kPackStride, sizeof(uint32_t))) return rc;
    CK(cudaGetLastError());
    CUtensorMap tmap;
    const unsigned long long rows = (hi - lo) / kChunk;
    if (!make_text_map(&tmap, rows ? text + lo : ctx->zero_row, rows ? rows : 1)) { ctx->err = "cuTensorMapEncodeTiled failed"; return KDB_E_CUDA; }
    CK(cudaEventRecord(ctx->ev[0], ctx->s_compute));
    CK(cudaMemsetAsync(ctx->look, 0, (size_t)n_tiles * sizeof(TileLook), ctx->s_compute));
    CK(cudaMemsetAsync(ctx->ticket, 0, sizeof(unsigned int), ctx->s_compute));
    CK(cudaEventRecord(ctx->ev[1], ctx->s_compute));
    EncodeParams2 P{};
    P.text = text + lo; P.feed = text; P.lo = lo; P.hi = hi; P.n = n; P.n_tiles = n_tiles; P.k = ctx->k; P.skip_n = ctx->skip_n;
    P.k_halo = std::max(ctx->k, ctx->k_halo); P.short_k = ctx->short_k; P.packed = ctx->packed; P.look = ctx->look; P.trec = ctx->trec;
    P.ticket = ctx->ticket; P.carry_in = ctx->pcarry + ctx->pc_idx; P.carry_out = ctx->pcarry + (ctx->pc_idx ^ 1); P.acc = ctx->acc;
    CK(cudaEventRecord(ctx->ev[2], ctx->s_compute));
    {
        const int grid = (int)std::min<unsigned int>(n_tiles, (unsigned int)ctx->encode_grid);
        if (ctx->fastq) k2s_encode<true><<<grid, kThreads, 0, ctx->s_compute>>>(P, tmap);
        else k2s_encode<false><<<grid, kThreads, 0, ctx->s_compute>>>(P, tmap);
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(ctx->ev[6], ctx->s_compute));
    if (int rc = count_stage(ctx, ctx, n_tiles, lo, hi)) return rc;
    CK(cudaEventRecord(ctx->ev[3], ctx->s_compute));
    for (kdb_ctx* f : ctx->followers) {                                  // larger k of a sweep, from the same packed stream
        CK(cudaEventRecord(f->ev[6], ctx->s_compute));
        if (int rc = count_stage(f, ctx, n_tiles, lo, hi)) { ctx->err = f->err; return rc; }
        CK(cudaEventRecord(f->ev[3], ctx->s_compute));
    }
    CK(cudaEventRecord(ctx->ev[5], ctx->s_compute));
    {
        const int g = (int)((n_tiles + kStitchThreads - 1) / kStitchThreads);
        if (ctx->fastq) k4_stitch_records<true><<<g, kStitchThreads, 0, ctx->s_compute>>>(ctx->trec, n_tiles, P.carry_in, P.carry_out, ctx->short_k, ctx->acc);
        else k4_stitch_records<false><<<g, kStitchThreads, 0, ctx->s_compute>>>(ctx->trec, n_tiles, P.carry_in, P.carry_out, ctx->short_k, ctx->acc);
        CK(cudaGetLastError());
    }
    ctx->pc_idx ^= 1;
    CK(cudaEventRecord(ctx->ev[4], ctx->s_compute));
    CK(cudaEventSynchronize(ctx->ev[4]));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1])); ctx->timing.scan_ms += ms;
    CK(cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[6])); ctx->timing.encode_ms += ms;
    CK(cudaEventElapsedTime(&ms, ctx->ev[6], ctx->ev[3])); ctx->timing.count_ms += ms;
    CK(cudaEventElapsedTime(&ms, ctx->ev[6], ctx->ev[7])); ctx->timing.partition_ms += ms;
    CK(cudaEventElapsedTime(&ms, ctx->ev[5], ctx->ev[4])); ctx->timing.final_ms += ms;
    for (kdb_ctx* f : ctx->followers) {
        CK(cudaEventElapsedTime(&ms, f->ev[6], f->ev[3])); f->timing.count_ms += ms;
        CK(cudaEventElapsedTime(&ms, f->ev[6], f->ev[7])); f->timing.partition_ms += ms;
    }
    ctx->timing.count_launches++;
    ctx->timing.total_launches += 3;
    ctx->bytes_since_fold += hi - lo;
    return KDB_OK;
}

// mode bit 0: the open record is complete, check it; bit 1: a feed ended (k4_close_carry)
static int close_carry2(kdb_ctx* ctx, int mode) {
    if (ctx->fastq) k4_close_carry<true><<<1, 32, 0, ctx->s_compute>>>(ctx->pcarry + ctx->pc_idx, mode, ctx->short_k, ctx->acc);
    else k4_close_carry<false><<<1, 32, 0, ctx->s_compute>>>(ctx->pcarry + ctx->pc_idx, mode, ctx->short_k, ctx->acc);
    CK(cudaGetLastError());
    ctx->timing.total_launches++;
    return KDB_OK;
}

static int begin_feed(kdb_ctx* ctx) {
    if (ctx->leader) { ctx->err = "kdb_feed on a follower: feed its leader"; return KDB_E_STATE; }
    if (ctx->finished) { ctx->err = "kdb_feed after kdb_finish (call kdb_reset first)"; return KDB_E_STATE; }
    if (!ctx->job_started) { CK(cudaEventRecord(ctx->ev_job0, ctx->s_compute)); ctx->job_started = true; }
    return KDB_OK;                                                       // (line indices restart with every fragment: the end of the feed before reset them on the device)
}

static int end_feed(kdb_ctx* ctx, size_t n) {
    // line indices restart with every fragment; a FASTQ fragment ends with a complete record
    if (int rc = close_carry2(ctx, ctx->fastq ? 3 : 2)) return rc;
    ctx->timing.bytes_fed += n;
    return KDB_OK;
}

extern "C" int kdb_feed_device(kdb_ctx* ctx, const uint8_t* dev_bytes, size_t n) {
    if (!ctx) return KDB_E_ARG;
    if (int rc = set_device(ctx, ctx->device)) return rc;
    if (n == 0) return KDB_OK;
    if (!dev_bytes || ((uintptr_t)dev_bytes & 15)) { ctx->err = "device text must be 16-byte aligned"; return KDB_E_ARG; }
    if (int rc = begin_feed(ctx)) return rc;
    const unsigned long long pass = kPassBytes;                          // (TileLook: the counts of a pass fit 30 bits)
    for (unsigned long long lo = 0; lo < n; lo += pass)
        if (int rc = run_pass2(ctx, dev_bytes, lo, std::min<unsigned long long>(lo + pass, n), n)) return rc;
    return end_feed(ctx, n);
}

extern "C" int kdb_feed_host(kdb_ctx* ctx, const uint8_t* bytes, size_t n) {
    if (!ctx) return KDB_E_ARG;
    if (int rc = set_device(ctx, ctx->device)) return rc;
    if (n == 0) return KDB_OK;
    if (!bytes) return KDB_E_ARG;
    if (int rc = ensure(ctx, (void**)&ctx->feed_buf, &ctx->feed_cap, n + 16, 1)) return rc;
    if (int rc = begin_feed(ctx)) return rc;
    // stage pass by pass on the copy stream; the compute stream waits for the pass it needs, so the
    // copy of pass i+1 overlaps the kernels of pass i
    const unsigned long long pass = 96ull << 20;       // short passes: the kernels of the last pass are the only ones PCIe does not hide
    const size_t n_pass = (size_t)((n + pass - 1) / pass);
    while (ctx->ev_copy.size() < n_pass + 1) { cudaEvent_t e; CK(cudaEventCreate(&e)); ctx->ev_copy.push_back(e); }
    CK(cudaEventRecord(ctx->ev_copy[n_pass], ctx->s_copy));
    for (size_t i = 0; i < n_pass; i++) {
        const unsigned long long lo = i * pass, hi = std::min<unsigned long long>(lo + pass, n);
        CK(cudaMemcpyAsync(ctx->feed_buf + lo, bytes + lo, hi - lo, cudaMemcpyHostToDevice, ctx->s_copy));
        CK(cudaEventRecord(ctx->ev_copy[i], ctx->s_copy));
    }
    for (size_t i = 0; i < n_pass; i++) {
        const unsigned long long lo = i * pass, hi = std::min<unsigned long long>(lo + pass, n);
        CK(cudaStreamWaitEvent(ctx->s_compute, ctx->ev_copy[i], 0));
        if (int rc = run_pass2(ctx, ctx->feed_buf, lo, hi, n)) return rc;
    }
    float ms = 0;
    CK(cudaEventSynchronize(ctx->ev_copy[n_pass - 1]));
    CK(cudaEventElapsedTime(&ms, ctx->ev_copy[n_pass], ctx->ev_copy[n_pass - 1]));
    ctx->timing.h2d_ms += ms;
    return end_feed(ctx, n);
}

static int compute_unique(kdb_ctx* ctx) {
    CK(cudaMemsetAsync(&ctx->acc->unique, 0, sizeof(unsigned long long), ctx->s_compute));
    const int grid = (int)std::min<unsigned long long>((ctx->nbins + 255) / 256, 148ull * 16);
    k5_count_nonzero<<<grid, 256, 0, ctx->s_compute>>>(ctx->tab64, ctx->nbins, ctx->acc);
    CK(cudaGetLastError());
    ctx->timing.total_launches++;
    return KDB_OK;
}

static int fill_stats(kdb_ctx* ctx) {
    Accum a;
    CK(cudaMemcpyAsync(&a, ctx->acc, sizeof a, cudaMemcpyDeviceToHost, ctx->s_compute));
    CK(cudaStreamSynchronize(ctx->s_compute));
    {
        PassCarry pc;
        CK(cudaMemcpyAsync(&pc, ctx->pcarry + ctx->pc_idx, sizeof pc, cudaMemcpyDeviceToHost, ctx->s_compute));
        CK(cudaStreamSynchronize(ctx->s_compute));
        ctx->total_reads_host = ctx->fastq ? pc.fq_rec_base : pc.hdr_total;
    }
    kdb_stats_t& s = ctx->stats;
    s = kdb_stats_t{};
    s.total_reads = ctx->total_reads_host;
    s.total_kmers = a.total_kmers; s.unique_kmers = a.unique; s.sum_len = a.sum_len;
    s.min_len = a.min_len == ~0ull ? 0 : a.min_len; s.max_len = a.max_len;
    for (int i = 0; i < 4; i++) s.mono[i] = a.mono[i];
    s.n_windows_with_n = a.n_windows_with_n;
    if (a.layout_flag || ctx->host_layout) s.status = KDB_E_LAYOUT;
    else if (a.err_key != ~0ull) {
        s.err_record = a.err_key >> 2;
        const unsigned int pr = (unsigned int)(a.err_key & 3);
        s.status = pr == P_ALPHABET ? KDB_E_ALPHABET : pr == P_SHORT ? KDB_E_SHORT : KDB_E_IUPAC;
    } else if (s.total_reads == 0) s.status = KDB_E_EMPTY;
    return KDB_OK;
}

// A follower finishes after its leader: its own table is folded and counted for non-zeros; the per-record
// reductions are the leader's.  Raises: whatever the leader raises (the sweep dies at its first k); else a record
// shorter than this k (kmerdb/kmer.py:461-463) — alphabet raises do not depend on k, and a window of this k that
// holds RYSWKMBDHV contains a window of the leader's k that does, so no new KeyError can appear.
static int finish_follower(kdb_ctx* ctx) {
    kdb_ctx* L = ctx->leader;
    if (!L->finished) { ctx->err = "kdb_finish on a follower before its leader"; return KDB_E_STATE; }
    CK(cudaStreamSynchronize(L->s_compute));
    CK(cudaEventRecord(ctx->ev[0], ctx->s_compute));
    CK(cudaMemsetAsync(&ctx->acc->unique, 0, sizeof(unsigned long long), ctx->s_compute));
    if (int rc = fold_now(ctx, true)) return rc;
    CK(cudaEventRecord(ctx->ev[1], ctx->s_compute));
    if (int rc = fill_stats(ctx)) return rc;
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1])); ctx->timing.final_ms += ms;
    ctx->timing.job_ms = ctx->timing.count_ms + ctx->timing.final_ms;
    kdb_stats_t& s = ctx->stats;
    const kdb_stats_t& l = L->stats;
    s.total_reads = l.total_reads; s.sum_len = l.sum_len; s.min_len = l.min_len; s.max_len = l.max_len;
    for (int i = 0; i < 4; i++) s.mono[i] = l.mono[i];
    s.n_windows_with_n = l.n_windows_with_n;
    s.err_record = l.err_record;
    s.status = l.status ? l.status : (l.min_len < (uint64_t)ctx->k ? KDB_E_SHORT : KDB_OK);
    ctx->finished = true;
    return s.status;
}

extern "C" int kdb_finish(kdb_ctx* ctx) {
    if (!ctx) return KDB_E_ARG;
    if (int rc = set_device(ctx, ctx->device)) return rc;
    if (ctx->finished) return ctx->stats.status;
    if (ctx->leader) return finish_follower(ctx);
    if (!ctx->job_started) { CK(cudaEventRecord(ctx->ev_job0, ctx->s_compute)); ctx->job_started = true; }
    CK(cudaEventRecord(ctx->ev[0], ctx->s_compute));
    if (int rc = close_carry2(ctx, 1)) return rc;
    CK(cudaMemsetAsync(&ctx->acc->unique, 0, sizeof(unsigned long long), ctx->s_compute));
    if (int rc = fold_now(ctx, true)) return rc;
    CK(cudaEventRecord(ctx->ev[1], ctx->s_compute));
    CK(cudaEventRecord(ctx->ev_job1, ctx->s_compute));
    if (int rc = fill_stats(ctx)) return rc;
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1])); ctx->timing.final_ms += ms;
    CK(cudaEventElapsedTime(&ms, ctx->ev_job0, ctx->ev_job1)); ctx->timing.job_ms = ms;
    ctx->finished = true;
    return ctx->stats.status;
}

extern "C" int kdb_stats(kdb_ctx* ctx, kdb_stats_t* out) {
    if (!ctx || !out) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_stats before kdb_finish"; return KDB_E_STATE; }
    if (ctx->stats_dirty) {                      // a multi-GPU reduce rewrote the device scalars: read them once, here
        if (int rc = set_device(ctx, ctx->device)) return rc;
        if (int rc = fill_stats(ctx)) return rc;
        ctx->stats_dirty = false;
    }
    *out = ctx->stats;
    return KDB_OK;
}

extern "C" int kdb_counts(kdb_ctx* ctx, uint64_t* out) {
    if (!ctx || !out) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_counts before kdb_finish"; return KDB_E_STATE; }
    if (ctx->not_root) { ctx->err = "kdb_counts on a non-root rank after kdb_reduce_nccl: only the root holds the whole table"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    CK(cudaMemcpyAsync(out, ctx->tab64, ctx->nbins * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->s_compute));
    CK(cudaStreamSynchronize(ctx->s_compute));
    return KDB_OK;
}

extern "C" int kdb_nullomers(kdb_ctx* ctx, uint64_t* out_ids, size_t cap, size_t* n) {
    if (!ctx || !n) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_nullomers before kdb_finish"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    const unsigned int blocks = (unsigned int)((ctx->nbins + kNullBlock - 1) / kNullBlock);
    // block counts + offsets: per-context scratch kept between calls (12 B per 2048 bins); ids: scoped, freed on every exit path
    if (int rc = ensure(ctx, (void**)&ctx->null_scratch, &ctx->null_cap, blocks, sizeof(unsigned int) + sizeof(unsigned long long))) return rc;
    unsigned long long* d_off = reinterpret_cast<unsigned long long*>(ctx->null_scratch);
    unsigned int* d_cnt = reinterpret_cast<unsigned int*>(d_off + ctx->null_cap);
    k6_null_count<<<blocks, 256, 0, ctx->s_compute>>>(ctx->tab64, ctx->nbins, d_cnt);
    CK(cudaGetLastError());
    std::vector<unsigned int> h_cnt(blocks);
    CK(cudaMemcpyAsync(h_cnt.data(), d_cnt, blocks * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->s_compute));
    CK(cudaStreamSynchronize(ctx->s_compute));
    std::vector<unsigned long long> h_off(blocks);
    unsigned long long total = 0;
    for (unsigned int b = 0; b < blocks; b++) { h_off[b] = total; total += h_cnt[b]; }
    *n = (size_t)total;
    ctx->timing.total_launches++;
    const size_t m = std::min<size_t>(cap, (size_t)total);
    if (out_ids && m) {
        DevBuf d_out;
        CK(d_out.alloc(m * sizeof(unsigned long long)));
        CK(cudaMemcpyAsync(d_off, h_off.data(), blocks * sizeof(unsigned long long), cudaMemcpyHostToDevice, ctx->s_compute));
        k6_null_write<<<blocks, 256, 0, ctx->s_compute>>>(ctx->tab64, ctx->nbins, d_off, d_out.as<unsigned long long>(), m);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(out_ids, d_out.p, m * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->s_compute));
        CK(cudaStreamSynchronize(ctx->s_compute));
        ctx->timing.total_launches++;
    }
    return KDB_OK;
}

extern "C" int kdb_timing(kdb_ctx* ctx, kdb_timing_t* out) {
    if (!ctx || !out) return KDB_E_ARG;
    if (ctx->reduce_timed) {
        float ms = 0;
        CK(cudaEventSynchronize(ctx->ev_red1));
        CK(cudaEventElapsedTime(&ms, ctx->ev_red0, ctx->ev_red1));
        ctx->timing.reduce_ms += ms;
        ctx->reduce_timed = false;
    }
    *out = ctx->timing;
    return KDB_OK;
}

// ---- multi-GPU support: expose the partial table and the additive scalars -------------------------
extern "C" int kdb_device_table(kdb_ctx* ctx, void** dev_counts_u64, size_t* n_bins) {
    if (!ctx || !dev_counts_u64 || !n_bins) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_device_table before kdb_finish"; return KDB_E_STATE; }
    *dev_counts_u64 = ctx->tab64; *n_bins = (size_t)ctx->nbins;
    return KDB_OK;
}

// out16: 0 reads, 1 kmers, 2 sum_len, 3..6 mono, 7 windows with N, 8 min_len (~0 = none), 9 max_len,
//        10 err_key (~0 = none; min-reduce), 11 layout flag (max-reduce), 12..15 reserved
extern "C" int kdb_export_partials(kdb_ctx* ctx, uint64_t* out16) {
    if (!ctx || !out16) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_export_partials before kdb_finish"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    Accum a;
    CK(cudaMemcpy(&a, ctx->acc, sizeof a, cudaMemcpyDeviceToHost));
    memset(out16, 0, 16 * sizeof(uint64_t));
    if (ctx->leader) {                       // a follower of a sweep: the per-record scalars are its leader's
        const kdb_stats_t& st = ctx->stats;
        a.sum_len = st.sum_len; a.n_windows_with_n = st.n_windows_with_n;
        for (int i = 0; i < 4; i++) a.mono[i] = st.mono[i];
        a.min_len = st.total_reads ? st.min_len : ~0ull; a.max_len = st.max_len;
        const unsigned int pr = st.status == KDB_E_ALPHABET ? P_ALPHABET : st.status == KDB_E_SHORT ? P_SHORT : P_IUPAC;
        a.err_key = (st.status >= KDB_E_ALPHABET && st.status <= KDB_E_SHORT) ? st.err_record * 4 + pr : ~0ull;
        a.layout_flag = st.status == KDB_E_LAYOUT;
    }
    out16[0] = ctx->stats.total_reads; out16[1] = a.total_kmers; out16[2] = a.sum_len;
    for (int i = 0; i < 4; i++) out16[3 + i] = a.mono[i];
    out16[7] = a.n_windows_with_n; out16[8] = a.min_len; out16[9] = a.max_len; out16[10] = a.err_key;
    out16[11] = (a.layout_flag || ctx->host_layout) ? 1 : 0;
    return KDB_OK;
}

extern "C" int kdb_import_partials(kdb_ctx* ctx, const uint64_t* in16) {
    if (!ctx || !in16) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_import_partials before kdb_finish"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    Accum a{};
    a.total_kmers = in16[1]; a.sum_len = in16[2];
    for (int i = 0; i < 4; i++) a.mono[i] = in16[3 + i];
    a.n_windows_with_n = in16[7]; a.min_len = in16[8]; a.max_len = in16[9]; a.err_key = in16[10];
    a.layout_flag = in16[11] ? 1 : 0;
    CK(cudaMemcpy(ctx->acc, &a, sizeof a, cudaMemcpyHostToDevice));
    ctx->total_reads_host = in16[0];
    {
        PassCarry pc{};
        if (ctx->fastq) pc.fq_rec_base = in16[0]; else pc.hdr_total = in16[0];
        CK(cudaMemcpy(ctx->pcarry + ctx->pc_idx, &pc, sizeof pc, cudaMemcpyHostToDevice));
    }
    ctx->host_layout = false;
    return KDB_OK;
}

// unique_kmers / nullomers are not additive across GPUs: recompute them from the reduced table.
extern "C" int kdb_refinalize(kdb_ctx* ctx) {
    if (!ctx) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_refinalize before kdb_finish"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    CK(cudaDeviceSynchronize());   // the caller's collective ran on another stream
    if (int rc = compute_unique(ctx)) return rc;
    if (int rc = fill_stats(ctx)) return rc;
    return ctx->stats.status;
}

// ---- BGZF input: compressed bytes over PCIe, inflate on the device (kdb_inflate.cuh) ------------------------------------
// a corrupt member can make the decoder read on past its payload before a check fires (at most 15 bits per output byte of a
// 64 KiB member): the compressed buffer carries that much slack behind the last member
constexpr size_t kInflateSlack = (160u << 10) + 512;       // (+ two 128-byte lines of read-ahead)
static bool g_crc_table_loaded[64] = {false};
static int load_crc_table(kdb_ctx* ctx, int device) {
    if (device >= 0 && device < 64 && g_crc_table_loaded[device]) return KDB_OK;
    unsigned int t[256];
    for (unsigned int i = 0; i < 256; i++) {
        unsigned int c = i;
        for (int b = 0; b < 8; b++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        t[i] = c;
    }
    CK(cudaMemcpyToSymbol(c_crc_table, t, sizeof t));
    if (device >= 0 && device < 64) g_crc_table_loaded[device] = true;
    return KDB_OK;
}

// Walk the gzip members of `raw` (RFC 1952): every member must carry the BGZF 'BC' extra subfield (its total size - 1),
// found by walking the subfields of the extra field, not at a fixed offset.  Fills members[0..min(cap, n)); *n_members gets
// the count, *text_bytes the sum of ISIZE.  KDB_E_FORMAT when `raw` is not a whole number of BGZF members.
extern "C" int kdb_bgzf_index(const uint8_t* raw, size_t n, kdb_bgzf_member_t* members, size_t cap, size_t* n_members, uint64_t* text_bytes) {
    if (!raw || !n_members || !text_bytes) return KDB_E_ARG;
    size_t off = 0, cnt = 0;
    uint64_t total = 0;
    while (off < n) {
        if (n - off < 18 + 8 || raw[off] != 0x1f || raw[off + 1] != 0x8b || raw[off + 2] != 8 || !(raw[off + 3] & 4)) return KDB_E_FORMAT;
        const unsigned flg = raw[off + 3];
        const size_t xlen = raw[off + 10] | (raw[off + 11] << 8);
        size_t p = off + 12, xend = p + xlen;
        if (xend > n) return KDB_E_FORMAT;
        long bsize = -1;
        while (p + 4 <= xend) {
            const size_t slen = raw[p + 2] | (raw[p + 3] << 8);
            if (raw[p] == 'B' && raw[p + 1] == 'C' && slen == 2 && p + 6 <= xend) bsize = (raw[p + 4] | (raw[p + 5] << 8)) + 1;
            p += 4 + slen;
        }
        if (bsize < 0 || off + (size_t)bsize > n) return KDB_E_FORMAT;
        size_t q = xend;
        if (flg & 8) { while (q < off + bsize && raw[q]) q++; q++; }          // FNAME
        if (flg & 16) { while (q < off + bsize && raw[q]) q++; q++; }         // FCOMMENT
        if (flg & 2) q += 2;                                                  // FHCRC
        if (q + 8 > off + (size_t)bsize) return KDB_E_FORMAT;
        const uint8_t* tr = raw + off + bsize - 8;
        const uint32_t crc = tr[0] | (tr[1] << 8) | (tr[2] << 16) | ((uint32_t)tr[3] << 24);
        const uint32_t isize = tr[4] | (tr[5] << 8) | (tr[6] << 16) | ((uint32_t)tr[7] << 24);
        if (isize > 65536) return KDB_E_FORMAT;
        if (members && cnt < cap) members[cnt] = kdb_bgzf_member_t{(uint64_t)q, total, (uint32_t)(off + bsize - 8 - q), isize, crc, 0};
        cnt++;
        total += isize;
        off += (size_t)bsize;
    }
    *n_members = cnt; *text_bytes = total;
    return KDB_OK;
}

// comp / out: device buffers (comp readable 8 bytes past comp_len); members: host array.  Returns KDB_OK, or KDB_E_FORMAT
// when a member does not inflate to its ISIZE / CRC-32 (first_bad receives its index).
// Scratch of the two-kernel inflate for members[0, n): tokens (inflate_token_cap per member) followed by one token count per member.
static size_t inflate_scratch_bytes(const kdb_bgzf_member_t* members, size_t n_members) {
    size_t tok = 0;
    for (size_t i = 0; i < n_members; i++) tok += inflate_token_cap(members[i].out_len);
    return tok * 4 + n_members * 4 + 64;
}

// d_scratch: inflate_scratch_bytes(members, n_members) bytes (16-byte aligned).  KDB_INFLATE_WARP=1: the one-kernel path of r02
// (one warp per member, every lane decoding) for A/B.
static int inflate_members(kdb_ctx* ctx, int device, const uint8_t* d_comp, const kdb_bgzf_member_t* members, size_t n_members, uint8_t* d_out,
                           BgzfMember* d_members, unsigned int* d_status, uint8_t* d_scratch, cudaStream_t st, int check_crc) {
    if (int rc = load_crc_table(ctx, device)) return rc;
    static_assert(sizeof(kdb_bgzf_member_t) == sizeof(BgzfMember), "member layout");
    static const bool warp_kernel = getenv("KDB_INFLATE_WARP") != nullptr;
    if (warp_kernel) {
        CK(cudaMemcpyAsync(d_members, members, n_members * sizeof(BgzfMember), cudaMemcpyHostToDevice, st));
        const int grid = (int)std::min<size_t>((n_members + kInfWarps - 1) / kInfWarps, (size_t)148 * 8);
        k_inflate_bgzf<true><<<grid, kInfWarps * 32, 0, st>>>(d_comp, d_members, (unsigned int)n_members, d_out, d_status, check_crc);
        CK(cudaGetLastError());
        if (ctx) ctx->timing.total_launches++;
        return KDB_OK;
    }
    std::vector<BgzfMember> m(n_members);
    memcpy(m.data(), members, n_members * sizeof(BgzfMember));
    size_t quads = 0;
    for (size_t i = 0; i < n_members; i++) {
        if (quads > 0xFFFFFFFFull) { if (ctx) ctx->err = "BGZF input too large for one inflate call"; return KDB_E_ARG; }
        m[i].pad = (unsigned int)quads;
        quads += inflate_token_cap(m[i].out_len) / 4;
    }
    uint32_t* d_tok = reinterpret_cast<uint32_t*>(d_scratch);
    unsigned int* d_ntok = reinterpret_cast<unsigned int*>(d_tok + 4 * quads);
    CK(cudaMemcpyAsync(d_members, m.data(), n_members * sizeof(BgzfMember), cudaMemcpyHostToDevice, st));   // (pageable source: staged before the call returns)
    static bool attr_set[64] = {false};
    if (device >= 0 && device < 64 && !attr_set[device]) {
        CK(cudaFuncSetAttribute(k_inflate_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDecSmemBytes));
        attr_set[device] = true;
    }
    k_inflate_decode<<<(unsigned int)((n_members + kDecThreads - 1) / kDecThreads), kDecThreads, kDecSmemBytes, st>>>(d_comp, d_members, (unsigned int)n_members, d_out,
                                                                                                                       d_tok, d_ntok, d_status);
    CK(cudaGetLastError());
    k_inflate_resolve<<<(unsigned int)((n_members + kResWarps - 1) / kResWarps), kResWarps * 32, 0, st>>>(d_members, (unsigned int)n_members, d_tok, d_ntok, d_out, d_status,
                                                                                                            check_crc);
    CK(cudaGetLastError());
    if (ctx) ctx->timing.total_launches += 2;
    return KDB_OK;
}

extern "C" int kdb_inflate_bgzf(int device, const uint8_t* raw, size_t n, uint8_t* host_out, size_t out_cap, size_t* out_len) {
    kdb_ctx* ctx = nullptr;
    if (!raw || !out_len) return KDB_E_ARG;
    size_t nm = 0; uint64_t text = 0;
    if (int rc = kdb_bgzf_index(raw, n, nullptr, 0, &nm, &text)) return rc;
    *out_len = (size_t)text;
    if (!host_out) return KDB_OK;
    if (out_cap < text) return KDB_E_ARG;
    std::vector<kdb_bgzf_member_t> mem(nm);
    kdb_bgzf_index(raw, n, mem.data(), nm, &nm, &text);
    CK(cudaSetDevice(device));
    DevBuf d_comp, d_out, d_mem, d_st, d_scr;
    CK(d_comp.alloc(n + kInflateSlack)); CK(d_out.alloc(text + 16)); CK(d_mem.alloc(nm * sizeof(BgzfMember))); CK(d_st.alloc(nm * sizeof(unsigned int)));
    CK(d_scr.alloc(inflate_scratch_bytes(mem.data(), nm)));
    CK(cudaMemcpy(d_comp.p, raw, n, cudaMemcpyHostToDevice));
    if (int rc = inflate_members(ctx, device, d_comp.as<uint8_t>(), mem.data(), nm, d_out.as<uint8_t>(), d_mem.as<BgzfMember>(), d_st.as<unsigned int>(), d_scr.as<uint8_t>(), nullptr, 1)) return rc;
    std::vector<unsigned int> status(nm);
    CK(cudaMemcpy(status.data(), d_st.p, nm * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < nm; i++) if (status[i]) { char b[128]; snprintf(b, sizeof b, "BGZF member %zu does not inflate (code %u)", i, status[i]); g_last_error = b; return KDB_E_FORMAT; }
    CK(cudaMemcpy(host_out, d_out.p, text, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// kdb_feed_bgzf_host, pipelined: the file is cut into pieces of `piece_members` members whose text starts on a tile border.  Piece
// p: compressed bytes H2D on the copy stream -> decode + resolve on inflate stream p mod W (W pieces in flight, each with its own
// token scratch; the next piece on a stream starts after the one before has been resolved) -> ONE pass of the tokeniser / counter
// over the piece's text on the compute stream.  The decode kernel is bound by the latency of one member and leaves three quarters
// of the issue slots idle: the copies, the resolve kernels and the passes of the pieces before run underneath it.
// Member status is read once at the end: a corrupt member resets the job and returns KDB_E_FORMAT.
// Returns -1 when the members cannot be cut on tile borders (the caller takes the serial path).
constexpr int kInflateStreams = 16;
constexpr size_t kEarlyChunk = 16u << 20;                     // H2D chunks of a copy that starts before the members are indexed
static int feed_bgzf_pipelined(kdb_ctx* ctx, const uint8_t* raw, size_t n, std::vector<kdb_bgzf_member_t>& mem, uint64_t text, size_t piece_members,
                               bool early_copy) {
    const size_t nm = mem.size();
    std::vector<size_t> cut;                                      // member index where each piece starts
    cut.push_back(0);
    while (true) {
        size_t j = cut.back() + piece_members;
        if (j + piece_members / 4 >= nm) break;                   // (a short tail joins the last piece)
        const size_t lim = std::min(nm, j + piece_members);
        while (j < lim && (mem[j].out_off % kTile)) j++;
        if (j >= lim) return -1;
        cut.push_back(j);
    }
    cut.push_back(nm);
    const size_t n_piece = cut.size() - 1;
    for (size_t pi = 0; pi < n_piece; pi++) {
        const unsigned long long lo = mem[cut[pi]].out_off, hi = cut[pi + 1] < nm ? mem[cut[pi + 1]].out_off : text;
        if (hi - lo > kPassBytes) return -1;
    }
    const size_t n_slot = std::min<size_t>(n_piece, kInflateStreams);
    size_t slot_bytes = 0;
    for (size_t pi = 0; pi < n_piece; pi++) slot_bytes = std::max(slot_bytes, inflate_scratch_bytes(mem.data() + cut[pi], cut[pi + 1] - cut[pi]));
    slot_bytes = (slot_bytes + 255) & ~(size_t)255;
    const size_t meta = (nm * (sizeof(BgzfMember) + sizeof(unsigned int)) + 255) & ~(size_t)255;     // member table + status words, in front of the token slots
    if (!early_copy) if (int rc = ensure(ctx, (void**)&ctx->comp_buf, &ctx->comp_cap, n + kInflateSlack + 64, 1)) return rc;
    if (int rc = ensure(ctx, (void**)&ctx->feed_buf, &ctx->feed_cap, text + 64, 1)) return rc;
    if (int rc = ensure(ctx, (void**)&ctx->tok_buf, &ctx->tok_cap, meta + slot_bytes * n_slot, 1)) return rc;
    uint8_t* d_comp = ctx->comp_buf;
    BgzfMember* d_mem = reinterpret_cast<BgzfMember*>(ctx->tok_buf);
    unsigned int* d_st = reinterpret_cast<unsigned int*>(d_mem + nm);
    uint8_t* d_slots = ctx->tok_buf + meta;
    if (int rc = load_crc_table(ctx, ctx->device)) return rc;
    while (ctx->s_inf.size() < n_slot) { cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)); ctx->s_inf.push_back(st); }
    while (ctx->ev_inf.size() < n_piece + 1) { cudaEvent_t e; CK(cudaEventCreate(&e)); ctx->ev_inf.push_back(e); }
    while (ctx->ev_copy.size() < n_piece + 2) { cudaEvent_t e; CK(cudaEventCreate(&e)); ctx->ev_copy.push_back(e); }
    // the member table of the whole file, token offsets relative to the piece's slot
    for (size_t pi = 0; pi < n_piece; pi++) {
        size_t quads = 0;
        for (size_t i = cut[pi]; i < cut[pi + 1]; i++) { mem[i].pad_ = (uint32_t)quads; quads += inflate_token_cap(mem[i].out_len) / 4; }
    }
    if (!early_copy) {
        // everything before on the compute stream (earlier feeds read feed_buf, the reset zeroes tables) comes first
        CK(cudaEventRecord(ctx->ev_copy[n_piece + 1], ctx->s_compute));
        CK(cudaStreamWaitEvent(ctx->s_copy, ctx->ev_copy[n_piece + 1], 0));
        CK(cudaEventRecord(ctx->ev_copy[n_piece], ctx->s_copy));
    }
    CK(cudaMemcpyAsync(d_mem, mem.data(), nm * sizeof(BgzfMember), cudaMemcpyHostToDevice, ctx->s_inf[0]));   // (first in the stream of piece 0; the others wait for ev_inf[n_piece])
    static bool attr_set[64] = {false};
    if (ctx->device >= 0 && ctx->device < 64 && !attr_set[ctx->device]) {
        CK(cudaFuncSetAttribute(k_inflate_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDecSmemBytes));
        attr_set[ctx->device] = true;
    }
    const bool trace = getenv("KDB_INFLATE_TRACE") != nullptr;
    size_t byte0 = 0;
    for (size_t pi = 0; pi < n_piece; pi++) {
        const size_t byte1 = cut[pi + 1] < nm ? (size_t)mem[cut[pi + 1]].in_off : n;
        cudaStream_t st = ctx->s_inf[pi % n_slot];
        if (early_copy) {                                         // the chunk that holds the piece's last byte (and the decoder's read-ahead) has arrived
            const size_t c = std::min((n - 1) / kEarlyChunk, (byte1 + 512) / kEarlyChunk);
            CK(cudaStreamWaitEvent(st, ctx->ev_chunk[c], 0));
        } else {
            CK(cudaMemcpyAsync(d_comp + byte0, raw + byte0, byte1 - byte0, cudaMemcpyHostToDevice, ctx->s_copy));
            CK(cudaEventRecord(ctx->ev_copy[pi], ctx->s_copy));
            CK(cudaStreamWaitEvent(st, ctx->ev_copy[pi], 0));
        }
        byte0 = byte1;
        if (pi == 0) CK(cudaEventRecord(ctx->ev_inf[n_piece], st));
        else if (pi < n_slot) CK(cudaStreamWaitEvent(st, ctx->ev_inf[n_piece], 0));      // the member table is on the device
        const unsigned int cnt = (unsigned int)(cut[pi + 1] - cut[pi]);
        size_t quads = 0;
        for (size_t i = cut[pi]; i < cut[pi + 1]; i++) quads += inflate_token_cap(mem[i].out_len) / 4;
        uint32_t* d_tok = reinterpret_cast<uint32_t*>(d_slots + (pi % n_slot) * slot_bytes);
        unsigned int* d_ntok = reinterpret_cast<unsigned int*>(d_tok + 4 * quads);
        k_inflate_decode<<<(cnt + kDecThreads - 1) / kDecThreads, kDecThreads, kDecSmemBytes, st>>>(d_comp, d_mem + cut[pi], cnt, ctx->feed_buf, d_tok, d_ntok, d_st + cut[pi]);
        CK(cudaGetLastError());
        k_inflate_resolve<<<(cnt + kResWarps - 1) / kResWarps, kResWarps * 32, 0, st>>>(d_mem + cut[pi], cnt, d_tok, d_ntok, ctx->feed_buf, d_st + cut[pi], 1);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ctx->ev_inf[pi], st));
        ctx->timing.total_launches += 2;
    }
    for (size_t pi = 0; pi < n_piece; pi++) {
        const unsigned long long lo = mem[cut[pi]].out_off, hi = cut[pi + 1] < nm ? mem[cut[pi + 1]].out_off : text;
        CK(cudaStreamWaitEvent(ctx->s_compute, ctx->ev_inf[pi], 0));
        if (int rc = run_pass2(ctx, ctx->feed_buf, lo, hi, text)) return rc;
        if (trace) {                                              // (run_pass2 has waited for its last event, ev[4])
            float a = 0, b = 0, c = 0;
            cudaEvent_t t0 = early_copy ? ctx->ev_chunk.back() : ctx->ev_copy[n_piece];
            if (!early_copy) cudaEventElapsedTime(&a, t0, ctx->ev_copy[pi]);
            cudaEventElapsedTime(&b, t0, ctx->ev_inf[pi]);
            cudaEventElapsedTime(&c, t0, ctx->ev[4]);
            fprintf(stderr, "piece %zu: copied %.2f inflated %.2f counted %.2f ms\n", pi, a, b, c);
        }
    }
    std::vector<unsigned int> status(nm);
    CK(cudaMemcpyAsync(status.data(), d_st, nm * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->s_compute));
    CK(cudaStreamSynchronize(ctx->s_compute));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev_inf[n_piece], ctx->ev_inf[n_piece - 1])); ctx->timing.inflate_ms += ms;      // first decode -> last resolve (the passes run underneath)
    if (early_copy) CK(cudaEventElapsedTime(&ms, ctx->ev_chunk.back(), ctx->ev_chunk[(n - 1) / kEarlyChunk]));
    else CK(cudaEventElapsedTime(&ms, ctx->ev_copy[n_piece], ctx->ev_copy[n_piece - 1]));
    ctx->timing.h2d_ms += ms;
    for (size_t i = 0; i < nm; i++)
        if (status[i]) {
            for (kdb_ctx* f : ctx->followers) if (int rc = reset_state(f)) return rc;
            if (int rc = reset_state(ctx)) return rc;
            char b[160]; snprintf(b, sizeof b, "BGZF member %zu does not inflate (code %u); the job was reset", i, status[i]); ctx->err = b;
            return KDB_E_FORMAT;
        }
    return end_feed(ctx, (size_t)text);
}

// The profile path on compressed input: `raw` = a whole BGZF file in host memory.  The compressed bytes are copied to the device
// in pieces; each piece is inflated as soon as it has arrived (the copy of piece i+1 overlaps the inflate of piece i), then the
// text is tokenised and counted like a device feed.  KDB_E_FORMAT: not BGZF, or a member is corrupt (the caller falls back
// to the host's gzip, which raises what the reference raises).
extern "C" int kdb_feed_bgzf_host(kdb_ctx* ctx, const uint8_t* raw, size_t n) {
    if (!ctx || !raw) return KDB_E_ARG;
    if (int rc = set_device(ctx, ctx->device)) return rc;
    if (n == 0) return KDB_OK;
    if (n < 28 || raw[0] != 0x1f || raw[1] != 0x8b) return KDB_E_FORMAT;
    // Pinned input: the copy of the compressed bytes does not need the member index, so it starts first and the host walks the
    // member headers (one cache miss per member: 3 ms for 25 000 members) while the bytes travel; the kernels then run one after
    // the other.  Pageable input: a copy blocks the host and runs at half the speed, so there the file goes piece by piece — copy,
    // inflate and count of different pieces overlap (feed_bgzf_pipelined).  Measured on 10 M reads (504 MB compressed): the
    // pipeline gains 14 % from pageable memory, but loses from pinned memory — the decode kernel is bound by the latency of ONE
    // member and every kernel that shares the SMs with it stretches that latency (9 -> 18 ms with the passes underneath).
    bool pinned = false;
    {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, raw) == cudaSuccess) pinned = at.type == cudaMemoryTypeHost;
        else cudaGetLastError();
    }
    const bool warp_kernel = getenv("KDB_INFLATE_WARP") != nullptr;
    bool pipelined = !warp_kernel && !pinned;
    if (getenv("KDB_INFLATE_SERIAL")) pipelined = false;
    if (getenv("KDB_INFLATE_PIPELINE") && !warp_kernel) pipelined = true;
    const bool early = pinned && !getenv("KDB_INFLATE_LATE_COPY");
    if (early) {
        if (int rc = begin_feed(ctx)) return rc;
        if (int rc = ensure(ctx, (void**)&ctx->comp_buf, &ctx->comp_cap, n + kInflateSlack + 64, 1)) return rc;
        const size_t n_chunk = (n + kEarlyChunk - 1) / kEarlyChunk;
        while (ctx->ev_chunk.size() < n_chunk + 2) { cudaEvent_t e; CK(cudaEventCreate(&e)); ctx->ev_chunk.push_back(e); }
        CK(cudaEventRecord(ctx->ev_chunk[n_chunk], ctx->s_compute));        // everything before on the compute stream comes first
        CK(cudaStreamWaitEvent(ctx->s_copy, ctx->ev_chunk[n_chunk], 0));
        CK(cudaEventRecord(ctx->ev_chunk.back(), ctx->s_copy));              // (the last event of the vector marks the start of the copy)
        for (size_t c = 0; c < n_chunk; c++) {
            const size_t b0 = c * kEarlyChunk, b1 = std::min(n, b0 + kEarlyChunk);
            CK(cudaMemcpyAsync(ctx->comp_buf + b0, raw + b0, b1 - b0, cudaMemcpyHostToDevice, ctx->s_copy));
            CK(cudaEventRecord(ctx->ev_chunk[c], ctx->s_copy));
        }
    }
    size_t nm = 0; uint64_t text = 0;
    int irc = kdb_bgzf_index(raw, n, nullptr, 0, &nm, &text);
    if (irc || text == 0) {
        if (early) CK(cudaStreamSynchronize(ctx->s_copy));         // (`raw` may go away as soon as the call returns)
        return irc;
    }
    std::vector<kdb_bgzf_member_t> mem(nm);
    kdb_bgzf_index(raw, n, mem.data(), nm, &nm, &text);
    if (pipelined) {
        size_t pm = 2048;                                          // members per piece (64 * 65280 B = 255 tiles: bgzip's members cut on tile borders every 64)
        if (const char* e = getenv("KDB_INFLATE_PIECE")) pm = (size_t)atoi(e);
        pm = std::min<size_t>(std::max<size_t>(pm, 64), 4096);
        if (!early) if (int rc = begin_feed(ctx)) return rc;
        const int rc = feed_bgzf_pipelined(ctx, raw, n, mem, text, pm, early);
        if (rc != -1) { if (early && rc != KDB_OK) cudaStreamSynchronize(ctx->s_copy); return rc; }
        for (auto& m : mem) m.pad_ = 0;
    }
    const size_t meta = (nm * (sizeof(BgzfMember) + sizeof(unsigned int)) + 255) & ~(size_t)255;     // member table + status words, in front of the tokens
    if (!early) if (int rc = ensure(ctx, (void**)&ctx->comp_buf, &ctx->comp_cap, n + kInflateSlack + 64, 1)) return rc;
    if (int rc = ensure(ctx, (void**)&ctx->feed_buf, &ctx->feed_cap, text + 64, 1)) return rc;
    uint8_t* d_comp = ctx->comp_buf;
    if (int rc = begin_feed(ctx)) return rc;
    // pieces of two waves of members (one warp inflates one member: a piece must hold enough members to fill the GPU — with 32 MiB pieces
    // a launch had 10 warps per SM and the seventeen launches of a 558 MB file took 143 ms, 8.4 ms each, the latency of ONE member)
    static int wave_members = 0;                                  // members the GPU inflates at once
    static size_t piece_members = 0;
    if (!wave_members) {
        int per_sm = 0, sms = 0;
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        if (getenv("KDB_INFLATE_WARP")) {                         // one warp per member: two waves of resident warps
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_inflate_bgzf<true>, kInfWarps * 32, 0));
            wave_members = std::max(1, per_sm) * sms * kInfWarps;
            piece_members = 2 * (size_t)wave_members;
        } else {                                                  // one thread per member (k_inflate_decode): one wave of resident threads
            CK(cudaFuncSetAttribute(k_inflate_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDecSmemBytes));
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_inflate_decode, kDecThreads, kDecSmemBytes));
            wave_members = std::max(1, per_sm) * sms * kDecThreads;
            piece_members = (size_t)wave_members;
        }
    }
    if (const char* e = getenv("KDB_INFLATE_PIECE")) piece_members = (size_t)std::max(1, atoi(e));       // (tests: several pieces from a small file)
    else piece_members = (size_t)wave_members * (warp_kernel ? 2 : 1);
    std::vector<size_t> cut;                                      // member index where each piece starts
    for (size_t i = 0; i < nm; i += piece_members) cut.push_back(i);
    if (cut.size() > 1 && nm - cut.back() < piece_members / 4) cut.pop_back();     // (a short tail joins the piece before it)
    cut.push_back(nm);
    const size_t n_piece = cut.size() - 1;
    size_t scratch = 0;                                           // pieces run one after the other on the compute stream: one scratch serves all
    for (size_t pi = 0; pi < n_piece; pi++) scratch = std::max(scratch, inflate_scratch_bytes(mem.data() + cut[pi], cut[pi + 1] - cut[pi]));
    // (Tried: the resolve kernel of piece p on a second stream underneath the decode of piece p + 1, two token slots — 40 M reads,
    // 4 pieces: 60.5 ms against 59.6 one after the other.  Whatever shares the SMs with the decode kernel stretches its latency chain.)
    if (int rc = ensure(ctx, (void**)&ctx->tok_buf, &ctx->tok_cap, meta + scratch, 1)) return rc;
    BgzfMember* d_mem = reinterpret_cast<BgzfMember*>(ctx->tok_buf);
    unsigned int* d_st = reinterpret_cast<unsigned int*>(d_mem + nm);
    while (ctx->ev_copy.size() < n_piece + 1) { cudaEvent_t e; CK(cudaEventCreate(&e)); ctx->ev_copy.push_back(e); }
    if (!early) {
        CK(cudaEventRecord(ctx->ev_copy[n_piece], ctx->s_copy));
        size_t byte0 = 0;
        for (size_t pi = 0; pi < n_piece; pi++) {
            const size_t byte1 = cut[pi + 1] < nm ? (size_t)mem[cut[pi + 1]].in_off : n;
            CK(cudaMemcpyAsync(d_comp + byte0, raw + byte0, byte1 - byte0, cudaMemcpyHostToDevice, ctx->s_copy));
            CK(cudaEventRecord(ctx->ev_copy[pi], ctx->s_copy));
            byte0 = byte1;
        }
    }
    for (size_t pi = 0; pi < n_piece; pi++) {
        if (early) {                                              // the chunk that holds the piece's last byte (and the decoder's read-ahead) has arrived
            const size_t byte1 = cut[pi + 1] < nm ? (size_t)mem[cut[pi + 1]].in_off : n;
            CK(cudaStreamWaitEvent(ctx->s_compute, ctx->ev_chunk[std::min((n - 1) / kEarlyChunk, (byte1 + 512) / kEarlyChunk)], 0));
        } else CK(cudaStreamWaitEvent(ctx->s_compute, ctx->ev_copy[pi], 0));
        if (pi == 0) CK(cudaEventRecord(ctx->ev[0], ctx->s_compute));       // (inflate_ms: from the arrival of the first piece)
        if (int rc = inflate_members(ctx, ctx->device, d_comp, mem.data() + cut[pi], cut[pi + 1] - cut[pi], ctx->feed_buf, d_mem + cut[pi], d_st + cut[pi],
                                     ctx->tok_buf + meta, ctx->s_compute, 1)) return rc;
    }
    CK(cudaEventRecord(ctx->ev[1], ctx->s_compute));
    std::vector<unsigned int> status(nm);
    CK(cudaMemcpyAsync(status.data(), d_st, nm * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->s_compute));
    CK(cudaStreamSynchronize(ctx->s_compute));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1])); ctx->timing.inflate_ms += ms;
    if (early) CK(cudaEventElapsedTime(&ms, ctx->ev_chunk.back(), ctx->ev_chunk[(n - 1) / kEarlyChunk]));
    else CK(cudaEventElapsedTime(&ms, ctx->ev_copy[n_piece], ctx->ev_copy[n_piece - 1]));
    ctx->timing.h2d_ms += ms;
    for (size_t i = 0; i < nm; i++)
        if (status[i]) {                                          // (the same contract as the pipelined path: the job starts over)
            for (kdb_ctx* f : ctx->followers) if (int rc = reset_state(f)) return rc;
            if (int rc = reset_state(ctx)) return rc;
            char b[160]; snprintf(b, sizeof b, "BGZF member %zu does not inflate (code %u); the job was reset", i, status[i]); ctx->err = b;
            return KDB_E_FORMAT;
        }
    const unsigned long long pass = kPassBytes;
    for (unsigned long long lo = 0; lo < text; lo += pass)
        if (int rc = run_pass2(ctx, ctx->feed_buf, lo, std::min<unsigned long long>(lo + pass, text), text)) return rc;
    return end_feed(ctx, (size_t)text);
}

// ---- multi-GPU behind the C-ABI (kdb_reduce_nccl) -----------------------------------------------------------
#define NCK(call)                                                                                  \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess) {                                                                   \
            char b_[512];                                                                          \
            snprintf(b_, sizeof b_, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, nccl_api().GetErrorString(r_)); \
            if (ctx) ctx->err = b_;                                                                \
            g_last_error = b_;                                                                     \
            return KDB_E_CUDA;                                                                     \
        }                                                                                          \
    } while (0)

extern "C" int kdb_nccl_unique_id(void* out, size_t cap) {
    kdb_ctx* ctx = nullptr;
    if (!out || cap < sizeof(ncclUniqueId)) return KDB_E_ARG;
    if (!nccl_api().ok) { g_last_error = "NCCL is not available: " + nccl_api().why; return KDB_E_STATE; }
    ncclUniqueId id;
    NCK(nccl_api().GetUniqueId(&id));
    memcpy(out, &id, sizeof id);
    return KDB_OK;
}

// Collective over the `nranks` contexts that hold the shards of one input, in shard order (rank r = shard r).  After it
// rank 0 holds the summed table and every rank the combined statistics.  Everything is queued on the context's stream:
// the call returns without waiting for the device.
extern "C" int kdb_reduce_nccl(kdb_ctx* ctx, const void* unique_id, int rank, int nranks) {
    if (!ctx || !unique_id || nranks < 1 || rank < 0 || rank >= nranks) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_reduce_nccl before kdb_finish"; return KDB_E_STATE; }
    if (ctx->leader && !ctx->leader->finished) { ctx->err = "kdb_reduce_nccl on a follower before its leader finished"; return KDB_E_STATE; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    if (nranks == 1) return KDB_OK;
    NcclApi& N = nccl_api();
    if (!N.ok) { ctx->err = "NCCL is not available: " + N.why; return KDB_E_STATE; }
    if (ctx->nbins % (unsigned long long)nranks) { ctx->err = "kdb_reduce_nccl: 4^k must be a multiple of the number of ranks"; return KDB_E_ARG; }
    if (ctx->comm && (ctx->comm_rank != rank || ctx->comm_n != nranks)) { N.CommDestroy(ctx->comm); ctx->comm = nullptr; }
    if (!ctx->comm) {
        ncclUniqueId id;
        memcpy(&id, unique_id, sizeof id);
        NCK(N.CommInitRank(&ctx->comm, nranks, id, rank));
        ctx->comm_rank = rank; ctx->comm_n = nranks;
    }
    if (!ctx->d_pack) CK(cudaMalloc(&ctx->d_pack, 16 * (1 + kMaxPeers * 8) * sizeof(unsigned long long)));
    if (nranks > kMaxPeers * 8) { ctx->err = "kdb_reduce_nccl: too many ranks"; return KDB_E_ARG; }
    if (!ctx->ev_red0) { CK(cudaEventCreate(&ctx->ev_red0)); CK(cudaEventCreate(&ctx->ev_red1)); }
    cudaStream_t st = ctx->s_compute;
    // (a follower's per-record scalars are its leader's: take them from there)
    kdb_ctx* rec = ctx->leader ? ctx->leader : ctx;
    const unsigned long long slice = ctx->nbins / (unsigned long long)nranks;
    unsigned long long* mine = ctx->tab64 + (unsigned long long)rank * slice;
    CK(cudaEventRecord(ctx->ev_red0, st));
    NCK(N.ReduceScatter(ctx->tab64, mine, (size_t)slice, ncclUint64, ncclSum, ctx->comm, st));
    CK(cudaMemsetAsync(&ctx->acc->unique, 0, sizeof(unsigned long long), st));
    {
        const int grid = (int)std::min<unsigned long long>((slice + 255) / 256, 148ull * 8);
        k_count_nonzero_range<<<grid, 256, 0, st>>>(ctx->tab64, (unsigned long long)rank * slice, (unsigned long long)(rank + 1) * slice, ctx->acc);
        CK(cudaGetLastError());
    }
    if (rec != ctx) {
        // follower: per-record words from the leader's device scalars, k-mer count and unique from its own
        k_make_pack<<<1, 32, 0, st>>>(rec->acc, rec->pcarry + rec->pc_idx, ctx->fastq, rec->host_layout, ctx->d_pack);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(ctx->d_pack + 1, &ctx->acc->total_kmers, sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(ctx->d_pack + 12, &ctx->acc->unique, sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st));
    } else {
        k_make_pack<<<1, 32, 0, st>>>(ctx->acc, ctx->pcarry + ctx->pc_idx, ctx->fastq, ctx->host_layout, ctx->d_pack);
        CK(cudaGetLastError());
    }
    NCK(N.AllGather(ctx->d_pack, ctx->d_pack + 16, 16, ncclUint64, ctx->comm, st));
    k_fold_packs<<<1, 32, 0, st>>>(ctx->d_pack + 16, nranks, ctx->acc, ctx->pcarry + ctx->pc_idx, ctx->fastq);
    CK(cudaGetLastError());
    // the summed slices travel to the root
    NCK(N.GroupStart());
    if (rank == 0) { for (int r = 1; r < nranks; r++) NCK(N.Recv(ctx->tab64 + (unsigned long long)r * slice, (size_t)slice, ncclUint64, r, ctx->comm, st)); }
    else NCK(N.Send(mine, (size_t)slice, ncclUint64, 0, ctx->comm, st));
    NCK(N.GroupEnd());
    CK(cudaEventRecord(ctx->ev_red1, st));
    ctx->timing.total_launches += 3;
    ctx->host_layout = false;                    // (folded into the device flag by k_make_pack / k_fold_packs)
    ctx->stats_dirty = true; ctx->reduce_timed = true; ctx->not_root = rank != 0;
    return KDB_OK;
}

// One process, one context per GPU (shard order), peer access: see kdb_multi.cuh.  ctxs[0] is the root.
extern "C" int kdb_reduce_peers(kdb_ctx** ctxs, int n) {
    kdb_ctx* ctx = (ctxs && n > 0) ? ctxs[0] : nullptr;
    if (!ctx || n > kMaxPeers) return KDB_E_ARG;
    for (int r = 0; r < n; r++) {
        if (!ctxs[r] || !ctxs[r]->finished) { ctx->err = "kdb_reduce_peers before kdb_finish on every context"; return KDB_E_STATE; }
        if (ctxs[r]->k != ctx->k || (ctxs[r]->flags & 7) != (ctx->flags & 7) || ctxs[r]->leader) { ctx->err = "kdb_reduce_peers: contexts must agree in k and flags"; return KDB_E_ARG; }
        for (int q = 0; q < r; q++) if (ctxs[q]->device == ctxs[r]->device) { ctx->err = "kdb_reduce_peers: one context per device"; return KDB_E_ARG; }
    }
    if (n == 1) return KDB_OK;
    if (ctx->nbins % (unsigned long long)(2 * n)) { ctx->err = "kdb_reduce_peers: 4^k must be a multiple of twice the number of GPUs"; return KDB_E_ARG; }
    for (int r = 0; r < n; r++) {
        CK(cudaSetDevice(ctxs[r]->device));
        for (int q = 0; q < n; q++) {
            if (q == r) continue;
            int can = 0;
            CK(cudaDeviceCanAccessPeer(&can, ctxs[r]->device, ctxs[q]->device));
            if (!can) { ctx->err = "kdb_reduce_peers: no peer access between the devices"; return KDB_E_STATE; }
            cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[q]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
            cudaGetLastError();
        }
    }
    // every GPU's table is complete once its stream reaches this point
    for (int r = 0; r < n; r++) {
        CK(cudaSetDevice(ctxs[r]->device));
        if (!ctxs[r]->ev_red0) { CK(cudaEventCreate(&ctxs[r]->ev_red0)); CK(cudaEventCreate(&ctxs[r]->ev_red1)); }
        CK(cudaEventRecord(ctxs[r]->ev_red0, ctxs[r]->s_compute));
    }
    PeerTables T{};
    for (int r = 0; r < n; r++) T.tab[r] = ctxs[r]->tab64;
    const unsigned long long slice = ctx->nbins / (unsigned long long)n;
    for (int r = 0; r < n; r++) {
        kdb_ctx* c = ctxs[r];
        CK(cudaSetDevice(c->device));
        for (int q = 0; q < n; q++) if (q != r) CK(cudaStreamWaitEvent(c->s_compute, ctxs[q]->ev_red0, 0));
        CK(cudaMemsetAsync(&c->acc->unique, 0, sizeof(unsigned long long), c->s_compute));
        const int grid = (int)std::min<unsigned long long>((slice / 2 + 255) / 256, 148ull * 8);
        k_peer_reduce_slice<<<grid, 256, 0, c->s_compute>>>(T, n, ctx->tab64, (unsigned long long)r * slice, (unsigned long long)(r + 1) * slice, &c->acc->unique);
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->ev_red1, c->s_compute));
        c->timing.total_launches++;
    }
    // scalars: a few words per GPU, folded on the host exactly like k_fold_packs (this is the end of the job)
    Accum a{};
    a.min_len = ~0ull; a.err_key = ~0ull;
    unsigned long long reads = 0;
    for (int r = 0; r < n; r++) {
        kdb_ctx* c = ctxs[r];
        CK(cudaSetDevice(c->device));
        CK(cudaStreamSynchronize(c->s_compute));
        Accum p; PassCarry pc;
        CK(cudaMemcpy(&p, c->acc, sizeof p, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(&pc, c->pcarry + c->pc_idx, sizeof pc, cudaMemcpyDeviceToHost));
        const unsigned long long rd = c->fastq ? pc.fq_rec_base : pc.hdr_total;
        unsigned long long ek = p.err_key;
        if (ek != ~0ull) ek += reads * 4;
        reads += rd;
        a.total_kmers += p.total_kmers; a.sum_len += p.sum_len;
        for (int i = 0; i < 4; i++) a.mono[i] += p.mono[i];
        a.n_windows_with_n += p.n_windows_with_n;
        a.min_len = std::min(a.min_len, p.min_len); a.max_len = std::max(a.max_len, p.max_len); a.err_key = std::min(a.err_key, ek);
        a.layout_flag |= (p.layout_flag || c->host_layout) ? 1u : 0u;
        a.unique += p.unique;
        if (r) c->not_root = true;
    }
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(ctx->acc, &a, sizeof a, cudaMemcpyHostToDevice));
    PassCarry pc{};
    if (ctx->fastq) pc.fq_rec_base = reads; else pc.hdr_total = reads;
    CK(cudaMemcpy(ctx->pcarry + ctx->pc_idx, &pc, sizeof pc, cudaMemcpyHostToDevice));
    ctx->host_layout = false;
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ctx->ev_red0, ctx->ev_red1));
    ctx->timing.reduce_ms += ms;
    if (int rc = fill_stats(ctx)) return rc;
    return ctx->stats.status;
}

// ---- batched kmer_to_id (kmerdb/kmer.py:234-317) ----------------------------------------------------
__global__ void __launch_bounds__(256) k_kmer_to_id(const char* __restrict__ kmers, size_t n, int k, int canonical, unsigned long long* __restrict__ ids) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const char* s = kmers + i * (size_t)k;
    unsigned long long fwd = 0, rc = 0;
    bool has_n = false, bad = false;
    for (int j = 0; j < k; j++) {
        const uint8_t cl = c_class_lut[(uint8_t)s[j]];
        if (cl == C_N) has_n = true;                          // kmer.py:287-289 -> None
        else if (cl > C_T) bad = true;                        // kmer.py:309 KeyError
        const unsigned long long code = cl <= C_T ? cl : 0;
        fwd = (fwd << 2) | code;
        rc |= (3ull - code) << (2 * j);
    }
    unsigned long long id = (canonical && rc < fwd) ? rc : fwd;
    if (has_n) id = ~0ull; else if (bad) id = ~0ull - 1;
    ids[i] = id;
}

extern "C" int kdb_kmer_to_id_batch(int device, const char* kmers, size_t n, int k, int canonical, uint64_t* ids) {
    kdb_ctx* ctx = nullptr;
    if (!kmers || !ids || k < 1 || k > 31) return KDB_E_ARG;
    if (n == 0) return KDB_OK;
    CK(cudaSetDevice(device));
    uint8_t lut[256];
    build_class_lut(lut);
    CK(cudaMemcpyToSymbol(c_class_lut, lut, 256));
    DevBuf d_k, d_ids;                           // scoped: freed on every exit path
    CK(d_k.alloc(n * (size_t)k));
    CK(d_ids.alloc(n * sizeof(unsigned long long)));
    CK(cudaMemcpy(d_k.p, kmers, n * (size_t)k, cudaMemcpyHostToDevice));
    k_kmer_to_id<<<(unsigned int)((n + 255) / 256), 256>>>(d_k.as<char>(), n, k, canonical, d_ids.as<unsigned long long>());
    CK(cudaGetLastError());
    CK(cudaMemcpy(ids, d_ids.p, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// ---- synthetic input ---------------------------------------------------------------------------------
extern "C" size_t kdb_synth_record_bytes(int read_len, int fastq) { return synth_record_bytes(read_len, fastq); }

extern "C" int kdb_synth_device(int device, uint8_t* dev_out, uint64_t first_read, uint64_t n_reads, int read_len, int fastq, uint64_t seed) {
    kdb_ctx* ctx = nullptr;
    if (!dev_out || read_len < 1) return KDB_E_ARG;
    CK(cudaSetDevice(device));
    k_synth<<<148 * 8, 256>>>(dev_out, first_read, n_reads, read_len, fastq, seed);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    return KDB_OK;
}

extern "C" int kdb_synth_host(uint8_t* out, uint64_t first_read, uint64_t n_reads, int read_len, int fastq, uint64_t seed) {
    if (!out || read_len < 1) return KDB_E_ARG;
    const uint32_t rb = synth_record_bytes(read_len, fastq);
    SynthCache cache; cache.valid = false; cache.read = 0; cache.blk = 0;
    for (uint64_t r = 0; r < n_reads; r++)
        for (uint32_t off = 0; off < rb; off++) out[r * rb + off] = synth_byte(first_read + r, off, read_len, fastq, seed, cache);
    return KDB_OK;
}

// ---- microbenchmarks ---------------------------------------------------------------------------------
extern "C" int kdb_microbench_atomics(int device, size_t table_bytes, uint64_t n_updates, int mode, double* updates_per_s) {
    kdb_ctx* ctx = nullptr;
    if (!updates_per_s) return KDB_E_ARG;
    CK(cudaSetDevice(device));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int grid = 148 * 8, threads = 256;
    const uint32_t per_thread = (uint32_t)std::max<uint64_t>(1, n_updates / ((uint64_t)grid * threads));
    float ms = 0;
    if (mode == 0) {
        size_t words = 1; while (words * 2 * 4 <= table_bytes) words *= 2;
        unsigned int* tab = nullptr;
        CK(cudaMalloc(&tab, words * 4));
        CK(cudaMemset(tab, 0, words * 4));
        k_micro_red_global<<<grid, threads>>>(tab, (uint32_t)(words - 1), per_thread / 8 + 1);   // warm-up
        CK(cudaEventRecord(e0));
        k_micro_red_global<<<grid, threads>>>(tab, (uint32_t)(words - 1), per_thread);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        cudaFree(tab);
    } else if (mode == 2 || mode == 3) {
        // distributed shared memory: clusters of 8 (portable) / 16 (opt-in) CTAs, random red.shared::cluster over the cluster's tables
        DevBuf sink;
        CK(sink.alloc(4));
        const int cs = mode == 2 ? 8 : 16;
        const void* fn = mode == 2 ? (const void*)k_micro_atoms_cluster<8> : (const void*)k_micro_atoms_cluster<16>;
        CK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
        if (cs == 16) CK(cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        cudaLaunchConfig_t cfg{};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3((grid / cs) * cs); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = 65536; cfg.attrs = at; cfg.numAttrs = 1;
        unsigned int* sp = sink.as<unsigned int>();
        uint32_t warm = per_thread / 8 + 1;
        void* a0[2] = {&sp, &warm};
        CK(cudaLaunchKernelExC(&cfg, fn, a0));
        CK(cudaEventRecord(e0));
        void* a1[2] = {&sp, (void*)&per_thread};
        CK(cudaLaunchKernelExC(&cfg, fn, a1));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        CK(cudaGetLastError());
        *updates_per_s = (double)per_thread * cfg.gridDim.x * threads / (ms * 1e-3);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        return KDB_OK;
    } else {
        unsigned int* sink = nullptr;
        CK(cudaMalloc(&sink, 4));
        CK(cudaFuncSetAttribute(k_micro_atoms_shared, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
        k_micro_atoms_shared<<<grid, threads, 65536>>>(sink, per_thread / 8 + 1);
        CK(cudaEventRecord(e0));
        k_micro_atoms_shared<<<grid, threads, 65536>>>(sink, per_thread);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        cudaFree(sink);
    }
    CK(cudaGetLastError());
    *updates_per_s = (double)per_thread * grid * threads / (ms * 1e-3);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return KDB_OK;
}

// ---- raw memory helpers --------------------------------------------------------------------------------
extern "C" int kdb_dev_alloc(int device, size_t n, void** out) {
    kdb_ctx* ctx = nullptr;
    if (!out) return KDB_E_ARG;
    CK(cudaSetDevice(device));
    CK(cudaMalloc(out, n ? n : 16));
    return KDB_OK;
}
extern "C" int kdb_dev_free(int device, void* p) {
    kdb_ctx* ctx = nullptr;
    CK(cudaSetDevice(device));
    CK(cudaFree(p));
    return KDB_OK;
}
extern "C" int kdb_host_alloc_pinned(size_t n, void** out) {
    kdb_ctx* ctx = nullptr;
    if (!out) return KDB_E_ARG;
    CK(cudaHostAlloc(out, n ? n : 16, cudaHostAllocDefault));
    return KDB_OK;
}
extern "C" int kdb_host_free_pinned(void* p) {
    kdb_ctx* ctx = nullptr;
    CK(cudaFreeHost(p));
    return KDB_OK;
}
extern "C" int kdb_memcpy_h2d(int device, void* dst, const void* src, size_t n) {
    kdb_ctx* ctx = nullptr;
    CK(cudaSetDevice(device));
    CK(cudaMemcpy(dst, src, n, cudaMemcpyHostToDevice));
    return KDB_OK;
}
extern "C" int kdb_memcpy_d2h(int device, void* dst, const void* src, size_t n) {
    kdb_ctx* ctx = nullptr;
    CK(cudaSetDevice(device));
    CK(cudaMemcpy(dst, src, n, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// ---- K4: graph edge rows, aggregated edge weights, minimizers (kdb_graph.cuh) -----------------------------
namespace {
int upload_lut(kdb_ctx* ctx) {
    uint8_t lut[256];
    build_class_lut(lut);
    CK(cudaMemcpyToSymbol(c_class_lut, lut, 256));
    return KDB_OK;
}
// exclusive scan of per-block counts on the host (a few thousand entries), uploaded for the write kernel
int scan_blocks(kdb_ctx* ctx, const unsigned int* d_cnt, unsigned int blocks, DevBuf& d_off, unsigned long long* total) {
    std::vector<unsigned int> h_cnt(blocks);
    CK(cudaMemcpy(h_cnt.data(), d_cnt, blocks * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    std::vector<unsigned long long> h_off(blocks);
    unsigned long long t = 0;
    for (unsigned int b = 0; b < blocks; b++) { h_off[b] = t; t += h_cnt[b]; }
    *total = t;
    CK(d_off.alloc(blocks * sizeof(unsigned long long)));
    CK(cudaMemcpy(d_off.p, h_off.data(), blocks * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    return KDB_OK;
}
}  // namespace

extern "C" int kdb_edge_rows(int device, const uint8_t* letters, size_t n_letters, const uint64_t* rec_off, size_t n_recs, int k, int canonical,
                             uint64_t* rec, uint64_t* pos1, uint64_t* id1, uint64_t* pos2, uint64_t* id2, size_t cap, size_t* n_edges,
                             uint64_t* err_record) {
    kdb_ctx* ctx = nullptr;
    if (!rec_off || !n_edges || k < 1 || k > 31 || (n_letters && !letters)) return KDB_E_ARG;
    *n_edges = 0;
    if (n_recs == 0) return KDB_E_EMPTY;
    CK(cudaSetDevice(device));
    if (int rc = upload_lut(ctx)) return rc;
    std::vector<unsigned long long> edge_off(n_recs + 1);
    edge_off[0] = 0;
    for (size_t r = 0; r < n_recs; r++) {
        const unsigned long long len = rec_off[r + 1] - rec_off[r];
        edge_off[r + 1] = edge_off[r] + (len > (unsigned long long)k ? len - k : 0);    // len-k+1 windows -> len-k edges
    }
    const unsigned long long total = edge_off[n_recs];
    DevBuf d_let, d_off, d_eoff, d_key, d_out;
    CK(d_let.alloc(n_letters)); CK(d_off.alloc((n_recs + 1) * 8)); CK(d_eoff.alloc((n_recs + 1) * 8)); CK(d_key.alloc(8));
    CK(cudaMemcpy(d_let.p, letters, n_letters, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_off.p, rec_off, (n_recs + 1) * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_eoff.p, edge_off.data(), (n_recs + 1) * 8, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_key.p, 0xFF, 8));
    const unsigned long long work = std::max<unsigned long long>(n_letters, n_recs);
    k8_check_letters<<<(unsigned int)((work + 255) / 256), 256>>>(d_let.as<uint8_t>(), n_letters, d_off.as<unsigned long long>(), n_recs, k, d_key.as<unsigned long long>());
    CK(cudaGetLastError());
    unsigned long long key = 0;
    CK(cudaMemcpy(&key, d_key.p, 8, cudaMemcpyDeviceToHost));
    if (key != ~0ull) {
        if (err_record) *err_record = key >> 36;
        const unsigned int cls = (unsigned int)((key >> 34) & 3);
        if (cls == G_ALPHABET) return KDB_E_ALPHABET;
        if (cls == G_SHORT) return KDB_E_SHORT;
        return (key & 1) ? KDB_E_IUPAC : KDB_E_FORMAT;              // N in graph mode: ValueError
    }
    *n_edges = (size_t)total;
    const size_t m = std::min<size_t>(cap, (size_t)total);
    if (total == 0 || m == 0) return KDB_OK;
    if (m < total) { g_last_error = "kdb_edge_rows: output capacity smaller than the edge count"; return KDB_E_ARG; }
    CK(d_out.alloc(5 * total * 8));
    unsigned long long* o = d_out.as<unsigned long long>();
    k8_edge_rows<<<(unsigned int)((total + 255) / 256), 256>>>(d_let.as<uint8_t>(), d_off.as<unsigned long long>(), d_eoff.as<unsigned long long>(), n_recs, total,
                                                              k, canonical, o, o + total, o + 2 * total, o + 3 * total, o + 4 * total);
    CK(cudaGetLastError());
    uint64_t* dst[5] = {rec, pos1, id1, pos2, id2};
    for (int i = 0; i < 5; i++) if (dst[i]) CK(cudaMemcpy(dst[i], o + (size_t)i * total, total * 8, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

extern "C" int kdb_edge_weights(kdb_ctx* ctx, int canonical, uint64_t* n1, uint64_t* n2, uint64_t* weight, size_t cap, size_t* n) {
    if (!ctx || !n) return KDB_E_ARG;
    if (!ctx->finished) { ctx->err = "kdb_edge_weights before kdb_finish"; return KDB_E_STATE; }
    if (ctx->canonical || ctx->k < 2) { ctx->err = "kdb_edge_weights needs a context of forward (k+1)-mers: create it with k+1, KDB_EDGE_KPLUS1 and without KDB_CANONICAL"; return KDB_E_ARG; }
    if (int rc = set_device(ctx, ctx->device)) return rc;
    CK(cudaStreamSynchronize(ctx->s_compute));
    const int k = ctx->k - 1;
    const unsigned int blocks = (unsigned int)((ctx->nbins + kEdgeBlock - 1) / kEdgeBlock);
    DevBuf d_cnt, d_off, d_out;
    CK(d_cnt.alloc(blocks * sizeof(unsigned int)));
    k7_edge_weight_count<<<blocks, 256>>>(ctx->tab64, ctx->nbins, k, canonical, d_cnt.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned long long total = 0;
    if (int rc = scan_blocks(ctx, d_cnt.as<unsigned int>(), blocks, d_off, &total)) return rc;
    *n = (size_t)total;
    ctx->timing.total_launches++;
    const size_t m = std::min<size_t>(cap, (size_t)total);
    if (!m || !n1 || !n2 || !weight) return KDB_OK;
    CK(d_out.alloc(3 * m * 8));
    unsigned long long* o = d_out.as<unsigned long long>();
    k7_edge_weight_write<<<blocks, 256>>>(ctx->tab64, ctx->nbins, k, canonical, d_off.as<unsigned long long>(), o, o + m, o + 2 * m, m);
    CK(cudaGetLastError());
    CK(cudaMemcpy(n1, o, m * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(n2, o + m, m * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(weight, o + 2 * m, m * 8, cudaMemcpyDeviceToHost));
    ctx->timing.total_launches++;
    return KDB_OK;
}

// minimizer.select_lexicographical_minimizers (kmerdb/minimizer.py:31-85) for one sequence.
extern "C" int kdb_minimizers_compat(int device, const uint8_t* seq, size_t len, int k, uint64_t window, uint64_t* ids, uint8_t* is_min,
                                     size_t cap, size_t* n_rows) {
    kdb_ctx* ctx = nullptr;
    if (!seq || !n_rows || k < 1 || k > 31 || window == 0) return KDB_E_ARG;
    *n_rows = 0;
    if (len < (size_t)k) return KDB_E_FORMAT;                                   // minimizer.py:59-60
    const unsigned long long L = (1ull << (2 * k)) - (unsigned long long)k + 1; // minimizer.py:65 (4^k - k + 1, not len - k + 1)
    if (len < L + (unsigned long long)k - 1) return KDB_E_FORMAT;               // minimizer.py:76-78 ValueError("WHOOPS")
    *n_rows = (size_t)L;
    if (!ids || !is_min || cap < L) return cap < L && (ids || is_min) ? KDB_E_ARG : KDB_OK;
    CK(cudaSetDevice(device));
    if (int rc = upload_lut(ctx)) return rc;
    DevBuf d_seq, d_ids, d_min, d_bad;
    CK(d_seq.alloc(L + k)); CK(d_ids.alloc(L * 8)); CK(d_min.alloc(L)); CK(d_bad.alloc(4));
    CK(cudaMemcpy(d_seq.p, seq, L + k - 1, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_bad.p, 0, 4));
    k8_min_compat<<<(unsigned int)((L + 255) / 256), 256>>>(d_seq.as<uint8_t>(), L, k, window, d_ids.as<unsigned long long>(), d_min.as<uint8_t>(), d_bad.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned int bad = 0;
    CK(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
    if (bad & 1u) return KDB_E_ALPHABET;
    if (bad & 2u) return KDB_E_IUPAC;
    CK(cudaMemcpy(ids, d_ids.p, L * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(is_min, d_min.p, L, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// True windowed minimizers (extension): for every run of w consecutive k-mers the position of
// the smallest canonical id (leftmost on ties); consecutive windows choosing the same k-mer give one row.
extern "C" int kdb_minimizers(int device, const uint8_t* seq, size_t len, int k, int w, uint64_t* pos, uint64_t* ids, size_t cap, size_t* n) {
    kdb_ctx* ctx = nullptr;
    if (!seq || !n || k < 1 || k > 31 || w < 1) return KDB_E_ARG;
    *n = 0;
    if (len < (size_t)k + (size_t)w - 1) return KDB_OK;
    CK(cudaSetDevice(device));
    if (int rc = upload_lut(ctx)) return rc;
    const unsigned long long n_win = len - k + 1, n_out = n_win - w + 1;
    DevBuf d_seq, d_ids, d_best, d_bad, d_cnt, d_off, d_out;
    CK(d_seq.alloc(len)); CK(d_ids.alloc(n_win * 8)); CK(d_best.alloc(n_out * 8)); CK(d_bad.alloc(4));
    CK(cudaMemcpy(d_seq.p, seq, len, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_bad.p, 0, 4));
    k8_window_ids<<<(unsigned int)((n_win + 255) / 256), 256>>>(d_seq.as<uint8_t>(), n_win, k, d_ids.as<unsigned long long>(), d_bad.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned int bad = 0;
    CK(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
    if (bad & 1u) return KDB_E_ALPHABET;
    if (bad & 2u) return KDB_E_IUPAC;
    k8_min_window<<<(unsigned int)((n_out + 255) / 256), 256>>>(d_ids.as<unsigned long long>(), n_win, w, n_out, d_best.as<unsigned long long>());
    CK(cudaGetLastError());
    const unsigned int blocks = (unsigned int)((n_out + kEdgeBlock - 1) / kEdgeBlock);
    CK(d_cnt.alloc(blocks * sizeof(unsigned int)));
    k8_min_count<<<blocks, 256>>>(d_best.as<unsigned long long>(), n_out, d_cnt.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned long long total = 0;
    if (int rc = scan_blocks(ctx, d_cnt.as<unsigned int>(), blocks, d_off, &total)) return rc;
    *n = (size_t)total;
    const size_t m = std::min<size_t>(cap, (size_t)total);
    if (!m || !pos || !ids) return KDB_OK;
    CK(d_out.alloc(2 * m * 8));
    unsigned long long* o = d_out.as<unsigned long long>();
    k8_min_write<<<blocks, 256>>>(d_best.as<unsigned long long>(), d_ids.as<unsigned long long>(), n_out, d_off.as<unsigned long long>(), o, o + m, m);
    CK(cudaGetLastError());
    CK(cudaMemcpy(pos, o, m * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ids, o + m, m * 8, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// ---- K9: .kdb body rows (kdb_rows.cuh) ---------------------------------------------------------------------
extern "C" int kdb_format_rows(int device, const uint64_t* counts, uint64_t first_row, uint64_t n_rows, const uint64_t* uniq_vals,
                               const uint32_t* str_off, const char* str_bytes, size_t n_uniq, char* out, size_t cap, size_t* n_bytes) {
    kdb_ctx* ctx = nullptr;
    if (!counts || !uniq_vals || !str_off || !str_bytes || !n_bytes || n_uniq == 0 || n_uniq > 0x7fffffffull) return KDB_E_ARG;
    *n_bytes = 0;
    if (n_rows == 0) return KDB_OK;
    CK(cudaSetDevice(device));
    const unsigned int blocks = (unsigned int)((n_rows + kRowBlock - 1) / kRowBlock);
    const size_t str_total = str_off[n_uniq];
    DevBuf d_counts, d_uniq, d_soff, d_str, d_idx, d_blen, d_boff, d_bad, d_out;
    CK(d_counts.alloc(n_rows * 8)); CK(d_uniq.alloc(n_uniq * 8)); CK(d_soff.alloc((n_uniq + 1) * 4)); CK(d_str.alloc(str_total));
    CK(d_idx.alloc(n_rows * 4)); CK(d_blen.alloc(blocks * 4)); CK(d_bad.alloc(4));
    CK(cudaMemcpy(d_counts.p, counts, n_rows * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_uniq.p, uniq_vals, n_uniq * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_soff.p, str_off, (n_uniq + 1) * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_str.p, str_bytes, str_total, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_bad.p, 0, 4));
    k9_row_len<<<blocks, kRowBlock>>>(d_counts.as<unsigned long long>(), first_row, n_rows, d_uniq.as<unsigned long long>(), d_soff.as<unsigned int>(),
                                      (unsigned int)n_uniq, d_idx.as<unsigned int>(), d_blen.as<unsigned int>(), d_bad.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned int bad = 0;
    CK(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
    if (bad) { g_last_error = (bad & 1) ? "kdb_format_rows: a count is missing from the string table" : "kdb_format_rows: a row exceeds 96 bytes"; return KDB_E_ARG; }
    unsigned long long total = 0;
    if (int rc = scan_blocks(ctx, d_blen.as<unsigned int>(), blocks, d_boff, &total)) return rc;
    *n_bytes = (size_t)total;
    if (!out) return KDB_OK;
    if (cap < total) { g_last_error = "kdb_format_rows: output buffer too small"; return KDB_E_ARG; }
    CK(d_out.alloc(total));
    k9_row_write<<<blocks, kRowBlock>>>(first_row, n_rows, d_idx.as<unsigned int>(), d_soff.as<unsigned int>(), d_str.as<char>(),
                                        d_boff.as<unsigned long long>(), d_out.as<char>());
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, d_out.p, total, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// True windowed minimizers of many records in one call (reads): rows (record, position inside the record, canonical id),
// record-major.  Records shorter than k + w - 1 give no row.  cap = 0 (or NULL outputs) only counts.
extern "C" int kdb_minimizers_batch(int device, const uint8_t* letters, size_t n_letters, const uint64_t* rec_off, size_t n_recs, int k, int w,
                                    uint64_t* rec, uint64_t* pos, uint64_t* ids, size_t cap, size_t* n_rows) {
    kdb_ctx* ctx = nullptr;
    if (!rec_off || !n_rows || k < 1 || k > 31 || w < 1 || (n_letters && !letters)) return KDB_E_ARG;
    *n_rows = 0;
    if (n_recs == 0 || n_letters < (size_t)k + (size_t)w - 1) return KDB_OK;
    CK(cudaSetDevice(device));
    if (int rc = upload_lut(ctx)) return rc;
    const unsigned long long n_win = n_letters - k + 1, n_out = n_win - w + 1;
    DevBuf d_let, d_off, d_ids, d_best, d_bad, d_cnt, d_boff, d_out;
    CK(d_let.alloc(n_letters)); CK(d_off.alloc((n_recs + 1) * 8)); CK(d_bad.alloc(4));
    CK(cudaMemcpy(d_let.p, letters, n_letters, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_off.p, rec_off, (n_recs + 1) * 8, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_bad.p, 0, 4));
    if (k <= 16 && w <= kFusedMaxW && !getenv("KDB_MIN_UNFUSED")) {
        // fused path: letters -> rows, nothing staged in HBM (k8_minb_fused); first launch counts (and checks the letters), second writes
        const unsigned int blocks = (unsigned int)((n_out + kEdgeBlock - 1) / kEdgeBlock);
        CK(d_cnt.alloc(blocks * sizeof(unsigned int)));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0));
        k8_minb_fused<false><<<blocks, 256>>>(d_let.as<uint8_t>(), n_letters, n_out, d_off.as<unsigned long long>(), n_recs, k, w, d_cnt.as<unsigned int>(),
                                              nullptr, nullptr, nullptr, nullptr, 0, d_bad.as<unsigned int>());
        CK(cudaGetLastError());
        CK(cudaEventRecord(e1));
        unsigned int bad = 0;
        CK(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
        CK(cudaEventElapsedTime(&g_last_kernel_ms, e0, e1));
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (bad & 1u) return KDB_E_ALPHABET;
        if (bad & 2u) return KDB_E_IUPAC;
        unsigned long long total = 0;
        if (int rc = scan_blocks(ctx, d_cnt.as<unsigned int>(), blocks, d_boff, &total)) return rc;
        *n_rows = (size_t)total;
        const size_t m = std::min<size_t>(cap, (size_t)total);
        if (!m || !rec || !pos || !ids) return KDB_OK;
        CK(d_out.alloc(3 * m * 8));
        unsigned long long* o = d_out.as<unsigned long long>();
        k8_minb_fused<true><<<blocks, 256>>>(d_let.as<uint8_t>(), n_letters, n_out, d_off.as<unsigned long long>(), n_recs, k, w, nullptr,
                                             d_boff.as<unsigned long long>(), o, o + m, o + 2 * m, m, d_bad.as<unsigned int>());
        CK(cudaGetLastError());
        CK(cudaMemcpy(rec, o, m * 8, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(pos, o + m, m * 8, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(ids, o + 2 * m, m * 8, cudaMemcpyDeviceToHost));
        return KDB_OK;
    }
    CK(d_ids.alloc(n_win * 8)); CK(d_best.alloc(n_out * 8));
    k8_window_ids<<<(unsigned int)((n_win + 255) / 256), 256>>>(d_let.as<uint8_t>(), n_win, k, d_ids.as<unsigned long long>(), d_bad.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned int bad = 0;
    CK(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
    if (bad & 1u) return KDB_E_ALPHABET;
    if (bad & 2u) return KDB_E_IUPAC;
    k8_min_window<<<(unsigned int)((n_out + 255) / 256), 256>>>(d_ids.as<unsigned long long>(), n_win, w, n_out, d_best.as<unsigned long long>());
    CK(cudaGetLastError());
    const unsigned int blocks = (unsigned int)((n_out + kEdgeBlock - 1) / kEdgeBlock);
    CK(d_cnt.alloc(blocks * sizeof(unsigned int)));
    k8_minb_rows<false><<<blocks, 256>>>(d_best.as<unsigned long long>(), d_ids.as<unsigned long long>(), n_out, d_off.as<unsigned long long>(), n_recs, k, w,
                                         d_cnt.as<unsigned int>(), nullptr, nullptr, nullptr, nullptr, 0);
    CK(cudaGetLastError());
    unsigned long long total = 0;
    if (int rc = scan_blocks(ctx, d_cnt.as<unsigned int>(), blocks, d_boff, &total)) return rc;
    *n_rows = (size_t)total;
    const size_t m = std::min<size_t>(cap, (size_t)total);
    if (!m || !rec || !pos || !ids) return KDB_OK;
    CK(d_out.alloc(3 * m * 8));
    unsigned long long* o = d_out.as<unsigned long long>();
    k8_minb_rows<true><<<blocks, 256>>>(d_best.as<unsigned long long>(), d_ids.as<unsigned long long>(), n_out, d_off.as<unsigned long long>(), n_recs, k, w,
                                        nullptr, d_boff.as<unsigned long long>(), o, o + m, o + 2 * m, m);
    CK(cudaGetLastError());
    CK(cudaMemcpy(rec, o, m * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(pos, o + m, m * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ids, o + 2 * m, m * 8, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// ---- .kdb body -> count vector (KDBReader._slurp, kmerdb/fileutil.py:308-466) ------------------------------
struct kdb_reader {
    int device = 0, k = 0;
    unsigned long long n_bins = 0, rows = 0;
    DevBuf d_ids, d_counts, d_seen, d_err, d_text;
    size_t text_cap = 0;
};

extern "C" kdb_reader* kdb_reader_create(int device, int k) {
    kdb_ctx* ctx = nullptr;
    if (k < 1 || k > 16) { g_last_error = "kdb_reader_create: k outside 1..16"; return nullptr; }
    kdb_reader* r = new kdb_reader();
    r->device = device; r->k = k; r->n_bins = 1ull << (2 * k);
    auto fail = [&](cudaError_t e) { g_last_error = std::string("kdb_reader_create: ") + cudaGetErrorString(e); delete r; return (kdb_reader*)nullptr; };
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(e);
    const size_t seen_bytes = (size_t)((r->n_bins + 31) / 32) * 4;
    if ((e = r->d_ids.alloc(r->n_bins * 8)) != cudaSuccess || (e = r->d_counts.alloc(r->n_bins * 8)) != cudaSuccess ||
        (e = r->d_seen.alloc(seen_bytes)) != cudaSuccess || (e = r->d_err.alloc(8)) != cudaSuccess) return fail(e);
    if ((e = cudaMemset(r->d_ids.p, 0, r->n_bins * 8)) != cudaSuccess || (e = cudaMemset(r->d_counts.p, 0, r->n_bins * 8)) != cudaSuccess ||
        (e = cudaMemset(r->d_seen.p, 0, seen_bytes)) != cudaSuccess || (e = cudaMemset(r->d_err.p, 0xFF, 8)) != cudaSuccess) return fail(e);
    (void)ctx;
    return r;
}

extern "C" void kdb_reader_destroy(kdb_reader* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    delete r;
}

// `text` holds whole rows (it starts at a row start; a last row without '\n' is taken as complete).
extern "C" int kdb_reader_feed(kdb_reader* r, const uint8_t* text, size_t n) {
    kdb_ctx* ctx = nullptr;
    if (!r || (n && !text)) return KDB_E_ARG;
    if (n == 0) return KDB_OK;
    CK(cudaSetDevice(r->device));
    if (n > r->text_cap) {
        if (r->d_text.p) { cudaFree(r->d_text.p); r->d_text.p = nullptr; }
        CK(r->d_text.alloc(n));
        r->text_cap = n;
    }
    CK(cudaMemcpy(r->d_text.p, text, n, cudaMemcpyHostToDevice));
    const unsigned int blocks = (unsigned int)((n + kReadBlock - 1) / kReadBlock);
    DevBuf d_cnt, d_off;
    CK(d_cnt.alloc(blocks * sizeof(unsigned int)));
    k10_count_rows<<<blocks, kReadThreads>>>(r->d_text.as<uint8_t>(), n, d_cnt.as<unsigned int>());
    CK(cudaGetLastError());
    unsigned long long feeds = 0;
    if (int rc = scan_blocks(ctx, d_cnt.as<unsigned int>(), blocks, d_off, &feeds)) return rc;
    ReadParams P;
    P.text = r->d_text.as<uint8_t>(); P.n = n; P.blk_off = d_off.as<unsigned long long>(); P.first_row = r->rows; P.n_bins = r->n_bins;
    P.kmer_ids = r->d_ids.as<unsigned long long>(); P.counts = r->d_counts.as<unsigned long long>(); P.seen = r->d_seen.as<unsigned int>();
    P.err = r->d_err.as<unsigned long long>();
    k10_parse_rows<<<blocks, kReadThreads>>>(P);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    r->rows += feeds + (text[n - 1] != '\n' ? 1 : 0);
    return KDB_OK;
}

// Copies the vectors out (either may be NULL).  Return value: 0, or the KDB_R_* class of the first failing row (*err_row).
// Fewer than 4^k rows fail as KDB_R_INDEX at row n_rows: the reference re-reads its last line there (fileutil.py:339-344, :359).
extern "C" int kdb_reader_finish(kdb_reader* r, uint64_t* counts, uint64_t* kmer_ids, uint64_t* n_rows, uint64_t* err_row) {
    kdb_ctx* ctx = nullptr;
    if (!r) return KDB_E_ARG;
    CK(cudaSetDevice(r->device));
    unsigned long long err = 0;
    CK(cudaMemcpy(&err, r->d_err.p, 8, cudaMemcpyDeviceToHost));
    if (n_rows) *n_rows = r->rows;
    if (err == ~0ull && r->rows < r->n_bins) err = ((unsigned long long)r->rows << 3) | (r->rows ? R_INDEX : R_COLUMNS);
    if (err != ~0ull && (err >> 3) > r->rows && r->rows < r->n_bins) err = ((unsigned long long)r->rows << 3) | R_INDEX;
    if (err != ~0ull) { if (err_row) *err_row = err >> 3; return KDB_R_BASE + (int)(err & 7); }
    if (counts) CK(cudaMemcpy(counts, r->d_counts.p, r->n_bins * 8, cudaMemcpyDeviceToHost));
    if (kmer_ids) CK(cudaMemcpy(kmer_ids, r->d_ids.p, r->n_bins * 8, cudaMemcpyDeviceToHost));
    return KDB_OK;
}

// ---- Pearson correlation of count profiles (kmerdb/distance.pyx:108-152, called per pair at __init__.py:733-742) ----
extern "C" int kdb_correlation_matrix(int device, const uint64_t* profiles, size_t n_samples, size_t n_bins, int mode, double* out) {
    kdb_ctx* ctx = nullptr;
    if (!profiles || !out || n_samples == 0 || n_bins == 0 || n_samples > 4096 || (mode != 0 && mode != 1)) return KDB_E_ARG;
    CK(cudaSetDevice(device));
    const int n = (int)n_samples;
    const unsigned int blocks = (unsigned int)std::min<size_t>(148 * 8, (n_bins + kCorrThreads - 1) / kCorrThreads);
    constexpr int Q = kCorrTile * kCorrTile;
    DevBuf d_x, d_sums, d_mean, d_part;
    CK(d_x.alloc(n_samples * n_bins * 8)); CK(d_sums.alloc(n_samples * 8)); CK(d_mean.alloc(n_samples * 8)); CK(d_part.alloc((size_t)blocks * Q * sizeof(DD)));
    CK(cudaMemcpy(d_x.p, profiles, n_samples * n_bins * 8, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_sums.p, 0, n_samples * 8));
    k11_sums<<<dim3(blocks, (unsigned int)n), kCorrThreads>>>(d_x.as<unsigned long long>(), n_bins, d_sums.as<unsigned long long>());
    CK(cudaGetLastError());
    std::vector<unsigned long long> sums(n);
    CK(cudaMemcpy(sums.data(), d_sums.p, n_samples * 8, cudaMemcpyDeviceToHost));
    std::vector<double> mean(n);
    for (int s = 0; s < n; s++) {
        const double m = (double)sums[s] / (double)n_bins;                  // np.sum(a) / total_kmers in float64 (distance.pyx:118)
        mean[s] = mode == 0 ? (double)(float)m : m;                         // `cdef float x_bar`
    }
    CK(cudaMemcpy(d_mean.p, mean.data(), n_samples * 8, cudaMemcpyHostToDevice));
    std::vector<long double> S((size_t)n * n, 0.0L);
    std::vector<DD> part((size_t)blocks * Q);
    for (int i0 = 0; i0 < n; i0 += kCorrTile)
        for (int j0 = i0; j0 < n; j0 += kCorrTile) {
            if (mode == 0) k11_products<0><<<blocks, kCorrThreads>>>(d_x.as<unsigned long long>(), n_bins, n, d_mean.as<double>(), i0, j0, d_part.as<DD>());
            else k11_products<1><<<blocks, kCorrThreads>>>(d_x.as<unsigned long long>(), n_bins, n, d_mean.as<double>(), i0, j0, d_part.as<DD>());
            CK(cudaGetLastError());
            CK(cudaMemcpy(part.data(), d_part.p, part.size() * sizeof(DD), cudaMemcpyDeviceToHost));
            for (int a = 0; a < kCorrTile && i0 + a < n; a++)
                for (int b = 0; b < kCorrTile && j0 + b < n; b++) {
                    long double acc = 0.0L;                                  // the reference's `long double` sums (distance.pyx:111-116)
                    for (unsigned int blk = 0; blk < blocks; blk++) { const DD v = part[(size_t)blk * Q + a * kCorrTile + b]; acc += (long double)v.hi; acc += (long double)v.lo; }
                    S[(size_t)(i0 + a) * n + (j0 + b)] = acc;
                    S[(size_t)(j0 + b) * n + (i0 + a)] = acc;
                }
        }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            const double prod = (double)(S[(size_t)i * n + i] * S[(size_t)j * n + j]);       // np.sqrt receives a Python float (distance.pyx:151)
            out[(size_t)i * n + j] = (double)(S[(size_t)i * n + j] / (long double)std::sqrt(prod));
        }
    return KDB_OK;
}

// ---- an entry point that allocates nothing (no counterpart in the reference) ---------
// kdb_nullomers_static: the ascending ids with zero count (what kdb_nullomers returns) for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  The per-block zero counts and offsets that kdb_nullomers
// keeps in a cudaMalloc'ed per-context buffer live in ONE file-scope device array here, sized for 4^14 bins (1.5 MiB, held by
// every GPU that loads the library until exit); the exclusive scan of the block counts runs on the device (kdb_nullomers does it
// on the host), so the call queues three kernels on the caller's stream and returns: no allocation, no copy, no synchronisation.
// All calls share the array: every word a call reads was written by that call before (nothing has to start at zero, stale
// contents of earlier calls are never read), but two calls that overlap on the GPU (different streams, or graphs replayed
// concurrently) race on it.
constexpr unsigned long long kStaticMaxBins = 1ull << 28;                       // 4^14
constexpr unsigned int kStaticMaxBlocks = (unsigned int)(kStaticMaxBins / kNullBlock);
struct NullScratch {
    unsigned long long off[kStaticMaxBlocks];       // exclusive scan of cnt
    unsigned int cnt[kStaticMaxBlocks];             // zero bins per block of kNullBlock bins
};
static __device__ __align__(256) unsigned char g_null_scratch[sizeof(NullScratch)];

// one CTA: off[b] = sum of cnt[0, b), *total = sum of all
__global__ void __launch_bounds__(1024) k6s_null_scan(const unsigned int* __restrict__ cnt, unsigned int blocks, unsigned long long* __restrict__ off,
                                                      unsigned long long* __restrict__ total) {
    __shared__ unsigned long long s_warp[32];
    const unsigned int per = (blocks + 1023u) / 1024u;
    const unsigned int b0 = min(blocks, threadIdx.x * per), b1 = min(blocks, b0 + per);
    unsigned long long sum = 0;
    for (unsigned int b = b0; b < b1; b++) sum += cnt[b];
    unsigned long long incl = sum;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = s_warp[lane];
        for (int d = 1; d < 32; d <<= 1) { const unsigned long long o = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += o; }
        s_warp[lane] = w;
    }
    __syncthreads();
    unsigned long long run = incl - sum + (warp ? s_warp[warp - 1] : 0ull);
    for (unsigned int b = b0; b < b1; b++) { off[b] = run; run += cnt[b]; }
    if (threadIdx.x == 1023) *total = s_warp[31];
}

// dev_counts: device uint64[n_bins] (e.g. kdb_device_table of a finished context); dev_out_ids: device uint64[cap];
// dev_n: device uint64[1], receives the number of zero bins (at most `cap` ids are written); stream: a cudaStream_t (NULL: the
// default stream).  The first call on a device looks the array's address up (cudaGetSymbolAddress): make it outside a capture.
extern "C" int kdb_nullomers_static(int device, const uint64_t* dev_counts, size_t n_bins, uint64_t* dev_out_ids, size_t cap, uint64_t* dev_n,
                                    void* stream) {
    kdb_ctx* ctx = nullptr;
    if (!dev_counts || !dev_n || (cap && !dev_out_ids)) { g_last_error = "kdb_nullomers_static: null pointer"; return KDB_E_ARG; }
    if (n_bins == 0 || n_bins > kStaticMaxBins) {
        char b[192];
        snprintf(b, sizeof b, "kdb_nullomers_static: n_bins = %zu is outside 1 .. %llu (4^14), the size the static scratch is built for; use kdb_nullomers",
                 n_bins, kStaticMaxBins);
        g_last_error = b;
        return KDB_E_ARG;
    }
    if (device < 0 || device >= 64) { g_last_error = "kdb_nullomers_static: device outside 0 .. 63"; return KDB_E_ARG; }
    CK(cudaSetDevice(device));
    static NullScratch* scratch_of[64] = {nullptr};                     // the array's DEVICE address, per device (its name is a host-side stand-in)
    if (!scratch_of[device]) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_null_scratch));
        scratch_of[device] = static_cast<NullScratch*>(p);
    }
    NullScratch* sc = scratch_of[device];
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned int blocks = (unsigned int)((n_bins + kNullBlock - 1) / kNullBlock);
    const unsigned long long* tab = reinterpret_cast<const unsigned long long*>(dev_counts);
    k6_null_count<<<blocks, 256, 0, st>>>(tab, (unsigned long long)n_bins, sc->cnt);
    CK(cudaGetLastError());
    k6s_null_scan<<<1, 1024, 0, st>>>(sc->cnt, blocks, sc->off, reinterpret_cast<unsigned long long*>(dev_n));
    CK(cudaGetLastError());
    if (cap) {
        k6_null_write<<<blocks, 256, 0, st>>>(tab, (unsigned long long)n_bins, sc->off, reinterpret_cast<unsigned long long*>(dev_out_ids), (unsigned long long)cap);
        CK(cudaGetLastError());
    }
    return KDB_OK;
}
