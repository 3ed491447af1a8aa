"""Throughput on non-uniform reads (what real data looks like to the bucket sort): AT-rich letters, and the worst case of
identical letters (every k-mer in one bin)."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native
n = 2_000_000
rng = np.random.default_rng(0)
def text(p):
    L = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.choice(4, size=(n, 150), p=p)]
    rec = np.empty((n, 162), dtype=np.uint8)
    rec[:, 0] = ord(">"); rec[:, 1:10] = ord("0"); rec[:, 10] = 10; rec[:, 11:161] = L; rec[:, 161] = 10
    return rec.reshape(-1)
for name, p in (("uniform", [.25, .25, .25, .25]), ("AT-rich 40/10/10/40", [.4, .1, .1, .4]), ("AT-rich 49/1/1/49", [.49, .01, .01, .49]), ("poly-A", [1, 0, 0, 0])):
    t = text(p)
    buf = _native.DeviceBuffer(t.size + 16); buf.upload(t)
    for k in (12, 8):
        with _native.Profiler(k) as pr:
            best = None
            for it in range(3):
                pr.reset(); pr.feed_device(buf.ptr, t.size); assert pr.finish() == 0
                tm = pr.timing()
                if it and (best is None or tm["job_ms"] < best["job_ms"]): best = tm
            st = pr.stats()
            assert st["total_kmers"] == n * (150 - k + 1)
        print(json.dumps({"reads": name, "k": k, "job_ms": round(best["job_ms"], 2), "encode_ms": round(best["encode_ms"], 2), "partition_ms": round(best["partition_ms"], 2),
                          "count_ms": round(best["count_ms"], 2), "unique": st["unique_kmers"], "bases_per_s": "%.2e" % (n * 150 / (best["job_ms"] * 1e-3))}), flush=True)
    buf.free()
