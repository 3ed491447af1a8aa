"""Timeline of kdb_feed_bgzf_host on pinned input (KDB_INFLATE_TRACE=1 prints per piece: copied / inflated / counted, ms from the
first copy).  usage: bgzf_trace.py [n_reads] [level]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from concurrent.futures import ThreadPoolExecutor
from kmerdb_b200 import _native, kdbfile
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
level = int(sys.argv[2]) if len(sys.argv) > 2 else 6
text = bytes(_native.synth_host(n))
with ThreadPoolExecutor(max_workers=16) as ex:
    blocks = list(ex.map(lambda o: kdbfile.bgzf_block(text[o:o + 65280], level), range(0, len(text), 65280)))
blob = np.frombuffer(b"".join(blocks) + kdbfile._BGZF_EOF, dtype=np.uint8)
pin = _native.PinnedBuffer(blob.size); pin.array[:] = blob
counts = _native.PinnedBuffer(8 * 4 ** 12)
cv = counts.array.view(np.uint64)
t0 = time.perf_counter(); _native.bgzf_text_size(pin.array); t_index = time.perf_counter() - t0
with _native.Profiler(12) as p:
    for it in range(4):
        p.reset()
        if it == 3: os.environ["KDB_INFLATE_TRACE"] = "1"
        t0 = time.perf_counter()
        assert p.feed_bgzf(pin.array) == 0
        t1 = time.perf_counter()
        assert p.finish() == 0
        t2 = time.perf_counter()
        p.counts(out=cv)
        t3 = time.perf_counter()
        t = p.timing()
        print(json.dumps({"it": it, "feed_ms": 1e3 * (t1 - t0), "finish_ms": 1e3 * (t2 - t1), "counts_ms": 1e3 * (t3 - t2), "total_ms": 1e3 * (t3 - t0),
                          "inflate_span_ms": t["inflate_ms"], "h2d_ms": t["h2d_ms"], "index_ms_once": 1e3 * t_index}))
