#!/usr/bin/env python
"""Time the .kdb writer and the GPU reader at k=12 (4^12 rows, ~700 MB of row text).  Writes gpurun_out/reader_probe.json."""
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from kmerdb_b200 import kdbfile, _native  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 12
rng = np.random.default_rng(12)
counts = rng.poisson(83.0, size=4 ** k).astype(np.uint64)
total, uniq = int(counts.sum()), int(np.count_nonzero(counts))
f = {"filename": "x.fa", "md5": "a" * 32, "sha256": "b" * 64, "total_reads": 1, "total_kmers": total, "unique_kmers": uniq,
     "nullomers": 4 ** k - uniq, "min_read_length": 150, "max_read_length": 150, "avg_read_length": 150}
md = {"version": "0.9.6", "metadata_blocks": 1, "k": k, "total_kmers": total, "unique_kmers": uniq, "unique_nullomers": 4 ** k - uniq,
      "sorted": False, "tags": [], "files": [f]}
out = {"k": k, "rows": 4 ** k}
with tempfile.TemporaryDirectory() as d:
    p = os.path.join(d, "p.%d.kdb" % k)
    t = time.perf_counter(); kdbfile.write_kdb_gpu(p, md, counts, threads=16); out["write_s"] = time.perf_counter() - t
    out["file_bytes"] = os.path.getsize(p)
    for rep in range(2):
        t = time.perf_counter(); r = kdbfile.KDBReader(p, threads=16); out["read_s"] = time.perf_counter() - t
    assert np.array_equal(r.counts, counts)
    raw = open(p, "rb").read()
    t = time.perf_counter(); pieces = list(kdbfile._inflate_groups(raw, 16)); out["inflate_s"] = time.perf_counter() - t
    body = b"".join(pieces)
    body = body[body.index(kdbfile.HEADER_DELIMITER.encode()) + len(kdbfile.HEADER_DELIMITER):]
    out["body_bytes"] = len(body)
    with _native.RowReader(k) as rd:
        t = time.perf_counter(); rd.feed(body); out["feed_s"] = time.perf_counter() - t
        t = time.perf_counter(); rc = rd.finish()[0]; out["finish_s"] = time.perf_counter() - t
    # the reference's loop, on a slice (it is ~1 us/row territory times a few): rows/s of the host checker
    t = time.perf_counter()
    m = 200000
    cc = np.zeros(4 ** k, dtype=np.uint64)
    for j, ln in enumerate(body[:m * 60].decode().splitlines()[:m]):
        l = ln.rstrip().split("\t")
        assert len(l) == 4 and int(l[0]) == j
        cc[int(l[1])] = int(l[2])
    out["host_loop_rows_per_s"] = m / (time.perf_counter() - t)
out["gpu_reader_rows_per_s"] = 4 ** k / out["read_s"]
print(json.dumps(out))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/reader_probe.json", "w"))
