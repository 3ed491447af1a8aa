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
