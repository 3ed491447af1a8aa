"""Worker of tests/test_gpu_multi.py (launched by torch.distributed.run, one rank per GPU): parity gate T3 —
the N-GPU result (shards -> partial tables -> NCCL sum-reduce -> refinalize) equals the 1-GPU result."""
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
            status = kd.reduce_profiler(p, dist, torch.device("cuda", local), dst=0)
            torch.cuda.synchronize()
            if rank == 0:
                assert status == 0
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
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
