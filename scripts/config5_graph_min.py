"""BASELINE.json configs[4]: graph k=12 edge-weight list + windowed minimizers on synthetic 150 bp reads sharded over N GPUs.
Launch: python -m torch.distributed.run --nproc-per-node N scripts/config5_graph_min.py [total_reads]   (N=1: plain python)."""
import json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native, distributed as kd

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
total = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
n = total // world
k, w = 12, 10
rb = _native.synth_record_bytes()
buf = _native.DeviceBuffer(n * rb + 16, device=local)
_native.synth_device(buf, n, first_read=rank * n)
def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
# ---- graph: forward 13-mer histogram per shard -> NCCL sum -> (n1, n2, weight) list on rank 0 ------------------
res = {}
with _native.Profiler(k + 1, canonical=False, device=local, edge_kplus1=True) as p:
    for it in range(2):
        p.reset(); sync(); t0 = time.perf_counter()
        p.feed_device(buf.ptr, n * rb)
        assert p.finish() == 0
        if world > 1:
            kd.reduce_profiler(p, dist, torch.device("cuda", local), dst=0)
        sync(); t1 = time.perf_counter()
        if rank == 0:
            n1, n2, wt = p.edge_weights(canonical=True)
            t2 = time.perf_counter()
            st = p.stats()
            assert int(wt.sum()) == total * (150 - k) == st["total_kmers"], (int(wt.sum()), st["total_kmers"])
            res["graph"] = {"count_and_reduce_ms": 1e3 * (t1 - t0), "edge_list_ms": 1e3 * (t2 - t1), "edges_total": int(wt.sum()), "distinct_pairs": int(len(wt)),
                            "bases_per_s": total * 150 / (t1 - t0)}
# ---- minimizers: ALL reads of the rank, in batched calls of 4 M reads (fused kernel: letters -> rows, nothing staged in HBM);
# rows are counted (they would be ~30 per read); the first 20 000 reads of rank 0 are written out and checked against the oracle
import oracle
chunk = 4_000_000
rows = 0
t_total = 0.0
for c0 in range(0, n, chunk):
    m = min(chunk, n - c0)
    text = buf.download(m * rb, offset=c0 * rb).reshape(m, rb)
    letters = np.ascontiguousarray(text[:, 11:161]).reshape(-1)
    off = np.arange(m + 1, dtype=np.uint64) * np.uint64(150)
    sync(); t0 = time.perf_counter()
    stt, r = _native.minimizers_batch(letters, off, k, w, device=local, count_only=True)
    sync(); t_total += time.perf_counter() - t0
    assert stt == 0
    rows += r
    kern_ms = locals().get("kern_ms", 0.0) + float(_native.lib().kdb_last_kernel_ms())
    if c0 == 0 and rank == 0:
        stt2, got = _native.minimizers_batch(letters[: 20000 * 150], off[:20001], k, w, device=local)
        want = []
        for i in range(20000):
            pos, ids = oracle.minimizers_true(letters[i * 150:(i + 1) * 150].tobytes().decode(), k, w)[1:3]
            want.extend((i, int(p), int(x)) for p, x in zip(pos, ids))
        parity = stt2 == 0 and got.tolist() == [list(t) for t in want]
        print("minimizers parity on the first 20 000 reads vs oracle:", "identical" if parity else "DIFFERENT", len(want), "rows", flush=True)
        assert parity
m = n
t0, t1 = 0.0, t_total
assert stt == 0
cnt = torch.tensor([rows], dtype=torch.int64, device="cuda")
if world > 1:
    dist.reduce(cnt, dst=0)
if rank == 0:
    res["minimizers"] = {"reads_per_rank": m, "k": k, "w": w, "rows_total": int(cnt.item()), "ms": 1e3 * (t1 - t0), "bases_per_s": world * m * 150 / (t1 - t0), "kernel_ms_rank0": kern_ms, "device_resident_bases_per_s_per_gpu": m * 150 / (kern_ms * 1e-3),
                         "note": "all reads of every rank; host letters -> H2D -> fused ids + window minima + masked count inside the timed calls; first 20 000 reads checked row by row against the oracle"}
    res.update({"n_gpus": world, "total_reads": total})
    print(json.dumps(res), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/config5_n%d.json" % world, "w"), indent=1)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
