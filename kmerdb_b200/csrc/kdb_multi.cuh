// Multi-GPU sum of the partial profiles (reference: the cross-file sum of kmerdb/__init__.py:1888-1891 and the
// header statistics re-derived from the summed vector at :1901-1903), behind the C-ABI.
//   kdb_reduce_nccl : one process per GPU.  ncclReduceScatter of the uint64[4^k] tables (in place: rank r keeps the sum of
//                     slice r), the non-zero count of the own slice, ONE ncclAllGather of the 16-word scalar packs (folded on
//                     the device by k_fold_packs, every rank the same arithmetic), then the slices travel to the root.  Nothing is
//                     read back to the host inside the step and nothing waits for the device.
//   kdb_reduce_peers: one process, several contexts with peer access: ONE kernel per GPU reads slice r of every GPU's table over
//                     NVLink, adds, counts the non-zero bins and stores the sum straight into the root's table — the fold, the
//                     count and the exchange are the same kernel.
// NCCL is bound at run time (dlopen of the libnccl.so.2 the process already holds — torch's — or the system one): the library
// has no link-time dependency on it, and a machine without NCCL can use everything else.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include "kdb_common.cuh"

namespace kdb {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
    std::string why;
};

inline NcclApi& nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);          // the copy this process already uses (torch's bundled NCCL)
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { api.why = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : ""); return api; }
    auto sym = [&](const char* name) -> void* { void* p = dlsym(h, name); if (!p) api.why = std::string("missing NCCL symbol ") + name; return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.ReduceScatter = (decltype(api.ReduceScatter))sym("ncclReduceScatter");
    api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.ReduceScatter && api.AllGather && api.Send && api.Recv &&
             api.GroupStart && api.GroupEnd && api.GetErrorString;
    return api;
}

// Scalar pack of one rank (same layout as kdb_export_partials): 0 reads, 1 kmers, 2 sum_len, 3..6 mono, 7 windows with N,
// 8 min_len (~0 none), 9 max_len, 10 err_key (~0 none), 11 layout flag, 12 non-zero bins of the rank's slice of the SUM.
__global__ void k_make_pack(const Accum* acc, const PassCarry* carry, int fastq, int host_layout, unsigned long long* pack) {
    if (threadIdx.x || blockIdx.x) return;
    pack[0] = fastq ? carry->fq_rec_base : carry->hdr_total;
    pack[1] = acc->total_kmers; pack[2] = acc->sum_len;
    for (int i = 0; i < 4; i++) pack[3 + i] = acc->mono[i];
    pack[7] = acc->n_windows_with_n; pack[8] = acc->min_len; pack[9] = acc->max_len; pack[10] = acc->err_key;
    pack[11] = (acc->layout_flag || host_layout) ? 1ull : 0ull;
    pack[12] = acc->unique;
    pack[13] = pack[14] = pack[15] = 0;
}

// Every rank folds the gathered packs the same way (kmerdb_b200/distributed.py:fold_partials is the host twin): sums wrap like
// uint64, min / max are unsigned, and a rank's first-raise key — (record index inside ITS shard) * 4 + priority — is rebased by the
// records of the shards before it, so that the minimum is the first failing record in FILE order (kmerdb/parse.py:128).
__global__ void k_fold_packs(const unsigned long long* all, int nranks, Accum* acc, PassCarry* carry, int fastq) {
    if (threadIdx.x || blockIdx.x) return;
    Accum a{};
    a.min_len = ~0ull; a.err_key = ~0ull;
    unsigned long long reads = 0;
    for (int r = 0; r < nranks; r++) {
        const unsigned long long* p = all + 16 * r;
        unsigned long long ek = p[10];
        if (ek != ~0ull) ek += reads * 4;
        reads += p[0];
        a.total_kmers += p[1]; a.sum_len += p[2];
        for (int i = 0; i < 4; i++) a.mono[i] += p[3 + i];
        a.n_windows_with_n += p[7];
        a.min_len = p[8] < a.min_len ? p[8] : a.min_len;
        a.max_len = p[9] > a.max_len ? p[9] : a.max_len;
        a.err_key = ek < a.err_key ? ek : a.err_key;
        a.layout_flag |= p[11] ? 1u : 0u;
        a.unique += p[12];
    }
    *acc = a;
    if (fastq) carry->fq_rec_base = reads; else carry->hdr_total = reads;
}

// non-zero bins of tab64[lo, hi)  ->  acc->unique (set, not added: the caller zeroes it first)
__global__ void __launch_bounds__(256) k_count_nonzero_range(const unsigned long long* __restrict__ tab64, unsigned long long lo, unsigned long long hi, Accum* acc) {
    unsigned long long c = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = lo + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) c += tab64[i] != 0;
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&acc->unique, c);
}

// ---- one process, peer access: the exchange, the sum and the non-zero count in one kernel ------------------------------------
constexpr int kMaxPeers = 16;
struct PeerTables { const unsigned long long* tab[kMaxPeers]; };
// GPU `me` owns bins [lo, hi): it reads that slice of every GPU's table (its own from HBM, the others over NVLink, 16-byte
// loads), adds, counts the non-zero sums and stores them into the root's table (a peer store unless me == root).
__global__ void __launch_bounds__(256) k_peer_reduce_slice(PeerTables T, int n, unsigned long long* __restrict__ root_tab, unsigned long long lo,
                                                           unsigned long long hi, unsigned long long* __restrict__ nz_out) {
    unsigned long long nz = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * 2;
    for (unsigned long long i = lo + ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 2; i < hi; i += stride) {
        ulonglong2 s = make_ulonglong2(0, 0);
#pragma unroll 4
        for (int r = 0; r < n; r++) {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(T.tab[r] + i);
            s.x += v.x; s.y += v.y;
        }
        *reinterpret_cast<ulonglong2*>(root_tab + i) = s;
        nz += (s.x != 0) + (s.y != 0);
    }
    for (int d = 16; d; d >>= 1) nz += __shfl_xor_sync(0xffffffffu, nz, d);
    if ((threadIdx.x & 31) == 0 && nz) atomicAdd(nz_out, nz);
}

}  // namespace kdb
