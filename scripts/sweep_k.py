"""--minK/--maxK sweep timing (BASELINE.json configs[3]): device-resident synthetic reads, one job per k.
usage: python scripts/sweep_k.py [n_reads] [kmin] [kmax]  -> gpurun_out/sweep_k.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
kmin = int(sys.argv[2]) if len(sys.argv) > 2 else 4
kmax = int(sys.argv[3]) if len(sys.argv) > 3 else 14
rb = _native.synth_record_bytes()
buf = _native.DeviceBuffer(n * rb + 16)
_native.synth_device(buf, n)
rows = []
for k in range(kmin, kmax + 1):
    with _native.Profiler(k) as p:
        best = None
        for it in range(4):
            p.reset()
            p.feed_device(buf.ptr, n * rb)
            assert p.finish() == 0
            t = p.timing()
            if it and (best is None or t["job_ms"] < best["job_ms"]):
                best = t
        st = p.stats()
        assert st["total_kmers"] == n * (150 - k + 1)
        rows.append({"k": k, "job_ms": best["job_ms"], "scan_ms": best["scan_ms"], "encode_ms": best["encode_ms"], "count_ms": best["count_ms"],
                     "partition_ms": best["partition_ms"], "fold_ms": best["fold_ms"], "final_ms": best["final_ms"],
                     "bases_per_s": n * 150 / (best["job_ms"] * 1e-3), "unique": st["unique_kmers"]})
        print(json.dumps(rows[-1]), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"n_reads": n, "rows": rows}, open("gpurun_out/sweep_k.json", "w"), indent=1)

# ---- the same sweep with one scan + encode for all k (kdb_attach) ------------------------------------------
import time
profs = [_native.Profiler(k) for k in range(kmin, kmax + 1)]
for f in profs[1:]:
    profs[0].attach(f)
best = None
for it in range(4):
    profs[0].reset()
    t0 = time.perf_counter()
    profs[0].feed_device(buf.ptr, n * rb)
    for p in profs:
        assert p.finish() == 0
    dt = time.perf_counter() - t0
    if it and (best is None or dt < best):
        best = dt
for k, p in zip(range(kmin, kmax + 1), profs):
    assert p.stats()["total_kmers"] == n * (150 - k + 1) and p.stats()["unique_kmers"] == rows[k - kmin]["unique"]
shared = {"wall_ms": 1e3 * best, "sum_of_separate_jobs_ms": sum(r["job_ms"] for r in rows),
          "leader": profs[0].timing(), "followers_count_ms": {k: p.timing()["count_ms"] for k, p in zip(range(kmin + 1, kmax + 1), profs[1:])}}
print(json.dumps(shared), flush=True)
json.dump({"n_reads": n, "rows": rows, "shared_sweep": shared}, open("gpurun_out/sweep_k.json", "w"), indent=1)
