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
pin = _native.PinnedBuffer(blob.size); pin.array[:] = blob
res = {}
with _native.Profiler(12) as p:
    for name, src in (("pinned", pin.array), ("pageable", blob)):      # pinned: copy, then the kernels one after the other; pageable: pipelined pieces
        for it in range(3):
            p.reset()
            t0 = time.perf_counter()
            assert p.feed_bgzf(src) == 0
            assert p.finish() == 0
            dt = time.perf_counter() - t0
            t = p.timing()
        st = p.stats()
        assert st["total_kmers"] == n * 139
        res[name] = {"inflate_ms": t["inflate_ms"], "h2d_ms": t["h2d_ms"], "wall_ms": 1e3 * dt}
t = res["pinned"]
print(json.dumps({"reads": n, "level": level, "text_bytes": len(text), "compressed_bytes": int(blob.size), "members": len(blocks), "inflate_ms": t["inflate_ms"],
                  "inflate_GBps_of_text": len(text) / (t["inflate_ms"] * 1e-3) / 1e9, "wall_ms": t["wall_ms"], "pageable_input": res["pageable"],
                  "note": "inflate_ms: decode + resolve kernels on pinned input (serial flow); pageable input goes through the piece pipeline, its inflate_ms is the span first decode -> last resolve with copies and passes underneath"}))
