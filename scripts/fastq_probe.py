"""FASTQ framing of the same synthetic reads (315 B/record, 2.10 B/base): per-kernel times at k=12."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
for fastq in (False, True):
    rb = _native.synth_record_bytes(150, fastq)
    buf = _native.DeviceBuffer(n * rb + 16)
    _native.synth_device(buf, n, fastq=fastq)
    with _native.Profiler(12, fastq=fastq) as p:
        best = None
        for it in range(4):
            p.reset(); p.feed_device(buf.ptr, n * rb); assert p.finish() == 0
            t = p.timing()
            if it and (best is None or t["job_ms"] < best["job_ms"]): best = t
        st = p.stats()
        assert st["total_kmers"] == n * 139
    print(json.dumps({"format": "fastq" if fastq else "fasta", "bytes": n * rb, "job_ms": best["job_ms"], "scan_ms": best["scan_ms"], "encode_ms": best["encode_ms"],
                      "count_ms": best["count_ms"], "bases_per_s": n * 150 / (best["job_ms"] * 1e-3), "text_GBps": n * rb / (best["job_ms"] * 1e-3) / 1e9}))
    buf.free()
