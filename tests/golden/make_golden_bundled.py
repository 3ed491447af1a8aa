#!/usr/bin/env python
"""Freeze configs[0] / parity gate T0: the reference's own bundled genomes profiled by the
UNMODIFIED reference (/root/reference/kmerdb/parse.py through tests/_bio_shim), plus the count columns of the
reference-held fixtures test/data/test_Cac_ATCC824.8.kdb (B1) and example_v0.6.3_kdb_file.8.kdb (B3) and the
k = 12 totals of kmerdb/appmap.py:271-306 (B4).  Run in the build container only:

    python tests/golden/make_golden_bundled.py

It copies the three input files (2.6 MB) into tests/golden/data/ — inputs, not sources — and writes
tests/golden/golden_bundled.json.  The -m gpu tests read only those committed files.
"""
import gzip
import hashlib
import json
import os
import shutil
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "_bio_shim"))
sys.path.insert(0, "/root/reference")

import logging  # noqa: E402

logging.disable(logging.CRITICAL)
import numpy as np  # noqa: E402
from kmerdb import parse  # noqa: E402

REF_DATA = "/root/reference/test/data"
DATA = os.path.join(HERE, "data")
FILES = ["Cacetobutylicum_ATCC824.fasta.gz", "Cdifficile_R3.fasta.gz", "sample.fa"]


def kdb_rows(path):
    with gzip.open(path, "rt") as fh:
        text = fh.read()
    body = text.split("\n" + "=" * 24 + "\n")[-1]
    return [ln.split("\t") for ln in body.splitlines() if ln.strip()]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def main():
    os.makedirs(DATA, exist_ok=True)
    for f in FILES:
        shutil.copyfile(os.path.join(REF_DATA, f), os.path.join(DATA, f))
    cases = []
    jobs = [(f, 8, canon) for f in FILES for canon in (True, False)]
    jobs += [("Cdifficile_R3.fasta.gz", 12, False), ("sample.fa", 12, True)]
    for f, k, canon in jobs:
        t0 = time.time()
        path = os.path.join(DATA, f)
        counts, meta, nullomers = parse.parsefile(path, k, replace_with_none=False, canonicalize=canon)
        meta = dict(meta)
        assert meta.pop("filename") == path
        cases.append({
            "file": f, "k": k, "canonicalize": canon, "replace_with_none": False,
            "meta": {a: (int(b) if not isinstance(b, str) else b) for a, b in meta.items()},
            "counts_sha256": sha(counts), "counts_sum": int(counts.sum()), "counts_head": [int(x) for x in counts[:8]],
            "nullomers_sha256": sha(nullomers), "n_nullomers": int(len(nullomers)),
        })
        print(f, k, canon, cases[-1]["meta"]["total_kmers"], cases[-1]["meta"]["unique_kmers"], "%.0f s" % (time.time() - t0), flush=True)

    # fixtures held by the reference's own tree
    b1 = np.zeros(4 ** 8, dtype=np.uint64)
    for r in kdb_rows(os.path.join(REF_DATA, "test_Cac_ATCC824.8.kdb")):
        b1[int(r[1])] = int(r[2])
    b3 = np.zeros(4 ** 8, dtype=np.uint64)
    for r in kdb_rows(os.path.join(REF_DATA, "example_v0.6.3_kdb_file.8.kdb")):
        b3[int(r[1])] = int(r[3])
    fixtures = {
        "B1": {"file": "Cacetobutylicum_ATCC824.fasta.gz", "k": 8, "canonicalize": False, "source": "test/data/test_Cac_ATCC824.8.kdb; test/test_fileutil.py:15-43",
               "counts_sha256": sha(b1), "total_kmers": 4132866, "unique_kmers": 64103, "nullomers": 1433},
        "B3": {"file": "Cdifficile_R3.fasta.gz", "k": 8, "canonicalize": False, "source": "test/data/example_v0.6.3_kdb_file.8.kdb",
               "counts_sha256": sha(b3), "total_kmers": 4093140, "unique_kmers": 63199, "nullomers": 2337,
               "mononucleotides": {"A": 1444649, "C": 601738, "G": 590294, "T": 1456466}},
        "B4": {"file": "Cdifficile_R3.fasta.gz", "k": 12, "canonicalize": False, "source": "kmerdb/appmap.py:271-306",
               "total_kmers": 4093136, "unique_kmers": 2267911, "nullomers": 14509305},
    }
    # the reference run and the reference-held fixtures must agree with each other before anything is frozen
    by = {(c["file"], c["k"], c["canonicalize"]): c for c in cases}
    assert by[("Cacetobutylicum_ATCC824.fasta.gz", 8, False)]["counts_sha256"] == fixtures["B1"]["counts_sha256"]
    assert by[("Cdifficile_R3.fasta.gz", 8, False)]["counts_sha256"] == fixtures["B3"]["counts_sha256"]
    c12 = by[("Cdifficile_R3.fasta.gz", 12, False)]["meta"]
    assert (c12["total_kmers"], c12["unique_kmers"], c12["nullomers"]) == (4093136, 2267911, 14509305)
    out = {"generator": "tests/golden/make_golden_bundled.py", "reference": "kmerdb v0.9.6 (unmodified) + tests/_bio_shim",
           "files": {f: hashlib.md5(open(os.path.join(DATA, f), "rb").read()).hexdigest() for f in FILES},
           "cases": cases, "fixtures": fixtures}
    with open(os.path.join(HERE, "golden_bundled.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print("wrote golden_bundled.json:", len(cases), "cases")


if __name__ == "__main__":
    main()
