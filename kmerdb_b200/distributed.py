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
    sums wrap like the device's uint64 arithmetic, min / max are taken on the unsigned values."""
    import torch
    p = np.ascontiguousarray(part16, dtype=np.uint64)
    mine = torch.from_numpy(p.view(np.int64).copy()).to(device)
    world = dist.get_world_size()
    allp = torch.empty(world * 16, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allp, mine)
    if dist.get_rank() != dst:
        return None
    g = allp.cpu().numpy().view(np.uint64).reshape(world, 16)
    out = np.zeros(16, dtype=np.uint64)
    with np.errstate(over="ignore"):
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
        lf = np.flatnonzero(a == 10)
        n_rec = (lf.size + (1 if n and a[-1] != 10 else 0)) // 4
        for r in range(1, world):
            rec = (n_rec * r) // world
            cuts.append(int(lf[4 * rec - 1]) + 1 if rec > 0 else 0)
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


def reduce_profiler(prof, dist, device, dst=0, _cache={}):
    """After prof.finish() on every rank: NCCL sum-reduce of the partial tables and scalars to `dst`, then
    the non-additive reductions (unique, nullomers) are recomputed there (kdb_refinalize)."""
    import torch
    key = id(prof)
    if key not in _cache:
        ptr, nb = prof.device_table()
        _cache.clear()
        _cache[key] = torch.as_tensor(CudaArray(ptr, nb), device=device)
    reduce_table(_cache[key], dist, dst)
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
