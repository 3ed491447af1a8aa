"""Worker of tests/test_gpu_multi.py (launched by torch.distributed.run, one rank per GPU): parity gate T3 —
the N-GPU result (shards -> partial tables -> kdb_reduce_nccl) equals the 1-GPU result:
count vector, statistics, nullomer ids, and the class + record of the first raise in file order."""
import hashlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native, distributed as kd   # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, k = 400000, 12
    text = _native.synth_host(n)                       # every rank regenerates the same stream (counter-based RNG)
    cuts = kd.shard_bounds(text, "fasta", world)
    ok = True
    for canonical in (True, False):
        with _native.Profiler(k, canonical=canonical, device=local) as p:
            p.feed(text[cuts[rank]:cuts[rank + 1]])
            assert p.finish() == 0
            kd.reduce_profiler(p, dist, torch.device("cuda", local), dst=0)      # kdb_reduce_nccl: queued on the context's stream
            if rank == 0:
                multi_counts, multi_stats, multi_null = p.counts(), p.stats(), p.nullomers()
        if rank == 0:
            with _native.Profiler(k, canonical=canonical, device=local) as q:
                q.feed(text)
                assert q.finish() == 0
                one_counts, one_stats, one_null = q.counts(), q.stats(), q.nullomers()
            same = np.array_equal(multi_counts, one_counts) and multi_stats == one_stats and np.array_equal(multi_null, one_null)
            print("T3 canonical=%s world=%d: %s sha256=%s unique=%d" % (canonical, world, "identical" if same else "DIFFERENT",
                  hashlib.sha256(multi_counts.tobytes()).hexdigest()[:16], multi_stats["unique_kmers"]), flush=True)
            ok &= same
    # raises: the N-GPU job must name the first failing record in FILE order with its class, whichever shard
    # holds it — a KeyError (IUPAC letter) early in the file beats a NameError (lowercase) at a small index of a later shard
    recs = [">r%d\n%s\n" % (i, "ACGT" * 30) for i in range(4000)]
    for where, seq in ((40, "ACGTRCGT" * 15), (4000 * (world - 1) // world + 2, "acgt" * 30)):
        recs[where] = ">r%d\n%s\n" % (where, seq)
    etext = np.frombuffer("".join(recs).encode(), dtype=np.uint8)
    ecuts = kd.shard_bounds(etext, "fasta", world)
    with _native.Profiler(k, canonical=True, device=local) as p:
        p.feed(etext[ecuts[rank]:ecuts[rank + 1]])
        p.finish()
        kd.reduce_profiler(p, dist, torch.device("cuda", local), dst=0)
        multi_stats = p.stats()
    if rank == 0:
        with _native.Profiler(k, canonical=True, device=local) as q:
            q.feed(etext)
            q.finish()
            one_stats = q.stats()
        same = multi_stats == one_stats and one_stats["status"] == _native.KDB_E_IUPAC and one_stats["err_record"] == 40
        print("T3 raise order world=%d: %s (status %d record %d)" % (world, "identical" if same else "DIFFERENT", multi_stats["status"], multi_stats["err_record"]), flush=True)
        ok &= same
    # FASTQ shards (4-line records) through the same reduce
    qtext = _native.synth_host(60000, fastq=True)
    qcuts = kd.shard_bounds(qtext, "fastq", world)
    with _native.Profiler(9, canonical=True, fastq=True, device=local) as p:
        p.feed(qtext[qcuts[rank]:qcuts[rank + 1]])
        assert p.finish() == 0
        kd.reduce_profiler(p, dist, torch.device("cuda", local), dst=0)
        if rank == 0:
            multi_counts, multi_stats = p.counts(), p.stats()
    if rank == 0:
        with _native.Profiler(9, canonical=True, fastq=True, device=local) as q:
            q.feed(qtext)
            assert q.finish() == 0
            same = np.array_equal(multi_counts, q.counts()) and multi_stats == q.stats()
        print("T3 fastq world=%d: %s" % (world, "identical" if same else "DIFFERENT"), flush=True)
        ok &= same
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
