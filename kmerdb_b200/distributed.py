"""Multi-GPU plumbing of the profile path: one process per GPU, reads sharded across ranks,
one sum-reduction of the 4^k table plus a 16-word scalar pack; unique/nullomers are recomputed after the sum.

torch.distributed is plumbing only (NCCL on the GPU box, gloo in the CPU tests): the kernels never see it.
The reference's counterpart is the cross-file sum of kmerdb/__init__.py:1888-1891 and the header statistics
re-derived from the summed vector at :1901-1903."""
import numpy as np

# layout of the scalar pack (kdb_export_partials): additive words, then min / max words
SUM_WORDS = slice(0, 8)      # reads, kmers, sum_len, mono[4], windows with N
MIN_WORDS = (8, 10)          # min_len, err_key   (~0 = none)
MAX_WORDS = (9, 11)          # max_len, layout flag
_SIGN = np.uint64(1 << 63)


def _ordered_i64(u):
    """uint64 -> int64 with the unsigned order kept (flip the top bit), so MIN/MAX of int64 tensors are exact."""
    return (np.asarray(u, dtype=np.uint64) ^ _SIGN).view(np.int64)


def _unordered_u64(i):
    return np.asarray(i, dtype=np.int64).view(np.uint64) ^ _SIGN


def combine_partials(part16, dist, device, dst=0):
    """Combine the 16-word scalar packs of all ranks.  Returns the combined pack on dst, None elsewhere.
    One collective (all_gather of 128 bytes per rank) and one copy each way; the words are then folded on the host:
    sums wrap like the device's uint64 arithmetic, min / max are taken on the unsigned values.
    A rank's first-raise key is (record index inside ITS shard) * 4 + priority; ranks hold consecutive shards of the
    file, so the key is rebased by the number of records of the ranks before it: the minimum is then the first
    failing record in FILE order, which is the one the reference raises on (kmerdb/parse.py:128 walks the file once)."""
    import torch
    p = np.ascontiguousarray(part16, dtype=np.uint64)
    mine = torch.from_numpy(p.view(np.int64).copy()).to(device)
    world = dist.get_world_size()
    allp = torch.empty(world * 16, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allp, mine)
    if dist.get_rank() != dst:
        return None
    g = allp.cpu().numpy().view(np.uint64).reshape(world, 16)
    return fold_partials(g)


def fold_partials(g):
    """g: uint64[world, 16], row r = the pack of rank r (shard r of the file).  Host arithmetic of combine_partials."""
    g = np.array(g, dtype=np.uint64, copy=True)
    none = np.uint64(0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        base = np.concatenate([[np.uint64(0)], np.cumsum(g[:-1, 0], dtype=np.uint64)])      # records before each shard
        hit = g[:, 10] != none
        g[hit, 10] = g[hit, 10] + base[hit] * np.uint64(4)
        out = np.zeros(16, dtype=np.uint64)
        out[SUM_WORDS] = g[:, SUM_WORDS].sum(axis=0, dtype=np.uint64)
    out[list(MIN_WORDS)] = g[:, list(MIN_WORDS)].min(axis=0)
    out[list(MAX_WORDS)] = g[:, list(MAX_WORDS)].max(axis=0)
    return out


def reduce_table(table_i64, dist, dst=0):
    """Sum-reduce the uint64[4^k] partial tables (viewed as int64: the sums are the same bits) to `dst`."""
    dist.reduce(table_i64, dst=dst, op=dist.ReduceOp.SUM)


def shard_bounds(data, fmt, world):
    """Cut uncompressed FASTA/FASTQ text into `world` contiguous ranges at record boundaries.
    Returns world+1 offsets.  FASTA: a record starts at '>' after a line feed.  FASTQ (4-line records, the only
    layout the device tokeniser takes natively): a record starts at a line whose index is a multiple of 4,
    so the cut is found by counting line feeds from the start of the text."""
    a = data if isinstance(data, np.ndarray) else np.frombuffer(data, dtype=np.uint8)
    n = a.size
    cuts = [0]
    if fmt == "fastq":
        # line feeds are counted block by block (no index array of every line feed: 8 B per line of a 31 GB input);
        # only the block that holds a cut is indexed
        blk = 1 << 26
        starts = list(range(0, n, blk))
        per = [int(np.count_nonzero(a[s:s + blk] == 10)) for s in starts]
        n_lf = sum(per)
        n_rec = (n_lf + (1 if n and a[-1] != 10 else 0)) // 4
        cum = np.concatenate([[0], np.cumsum(per)])
        for r in range(1, world):
            rec = (n_rec * r) // world
            if rec == 0:
                cuts.append(0)
                continue
            want = 4 * rec - 1                                   # index of the line feed that ends record rec - 1
            b = int(np.searchsorted(cum, want, side="right")) - 1
            lf = np.flatnonzero(a[starts[b]:starts[b] + blk] == 10)
            cuts.append(starts[b] + int(lf[want - int(cum[b])]) + 1)
    else:
        for r in range(1, world):
            t = (n * r) // world
            t = max(t, cuts[-1])
            win = a[t:]
            hit = np.flatnonzero((win[1:] == ord(">")) & (win[:-1] == 10))
            cuts.append(t + int(hit[0]) + 1 if hit.size else n)
    cuts.append(n)
    return cuts


class CudaArray:
    """Zero-copy view of a raw device allocation for torch (CUDA array interface)."""

    def __init__(self, ptr, n, typestr="<i8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def share_unique_id(dist, device):
    """128 bytes of NCCL unique id made on rank 0 (kdb_nccl_unique_id) and broadcast over the caller's process group —
    torch.distributed is only the messenger; the collective itself runs inside the library (kdb_reduce_nccl)."""
    import torch
    from . import _native
    t = torch.zeros(128, dtype=torch.uint8, device=device)
    if dist.get_rank() == 0:
        t.copy_(torch.frombuffer(bytearray(_native.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(t, src=0)
    return bytes(t.cpu().numpy().tobytes())


def reduce_profiler(prof, dist, device, dst=0):
    """After prof.finish() on every rank (rank r = shard r of the input): kdb_reduce_nccl — reduce-scatter of the partial
    tables, per-slice non-zero count, one all-gather of the scalar packs folded on the device, slices sent to rank 0.
    Nothing comes back to the host here; prof.stats() / prof.counts() (rank 0) read the result when asked."""
    assert dst == 0, "kdb_reduce_nccl gathers on rank 0"
    if getattr(prof, "_nccl_uid", None) is None:        # one communicator per context, kept by the library after the first call
        prof._nccl_uid = share_unique_id(dist, device)
    return prof.reduce_nccl(prof._nccl_uid, dist.get_rank(), dist.get_world_size())


def reduce_profiler_torch(prof, dist, device, dst=0):
    """The r01 path, kept for comparison: torch.distributed.reduce on the device table + host-folded scalar packs +
    kdb_refinalize (a second pass over the whole table)."""
    import torch
    ptr, nb = prof.device_table()
    reduce_table(torch.as_tensor(CudaArray(ptr, nb), device=device), dist, dst)
    out = combine_partials(prof.export_partials(), dist, device, dst)
    if out is not None:
        prof.import_partials(out)
        return prof.refinalize()
    return 0


def bind_to_gpu_numa(local_device):
    """Pin this process to the CPUs next to its GPU (sysfs local_cpulist of the PCI device) so that pinned host
    buffers are allocated on the GPU's NUMA node: with 8 ranks feeding 8 GPUs the host->device copies otherwise
    cross the socket interconnect.  Silently does nothing when sysfs does not say."""
    import os
    try:
        import torch
        pr = torch.cuda.get_device_properties(local_device)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % bus) as fh:
            spec = fh.read().strip()
        cpus = set()
        for part in spec.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return sorted(cpus)
    except Exception:
        pass
    return None
