"""GPU inflate of BGZF (kdb_feed_bgzf_host) on synthetic reads: inflate time and rate.  usage: inflate_probe.py [n_reads] [level]"""
import json, os, sys, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from concurrent.futures import ThreadPoolExecutor
from kmerdb_b200 import _native, kdbfile
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
level = int(sys.argv[2]) if len(sys.argv) > 2 else 6
text = bytes(_native.synth_host(n))
with ThreadPoolExecutor(max_workers=16) as ex:
    blocks = list(ex.map(lambda o: kdbfile.bgzf_block(text[o:o + 65280], level), range(0, len(text), 65280)))
blob = np.frombuffer(b"".join(blocks) + kdbfile._BGZF_EOF, dtype=np.uint8)
with _native.Profiler(12) as p:
    for it in range(3):
        p.reset()
        t0 = time.perf_counter()
        assert p.feed_bgzf(blob) == 0
        assert p.finish() == 0
        dt = time.perf_counter() - t0
        t = p.timing()
    st = p.stats()
    assert st["total_kmers"] == n * 139
print(json.dumps({"reads": n, "level": level, "text_bytes": len(text), "compressed_bytes": int(blob.size), "members": len(blocks), "inflate_ms": t["inflate_ms"],
                  "inflate_GBps_of_text": len(text) / (t["inflate_ms"] * 1e-3) / 1e9, "wall_ms": 1e3 * dt}))
