#!/usr/bin/env python
"""Freeze golden vectors for the .kdb READER by running the UNMODIFIED reference
`kmerdb.fileutil.open(path, 'r', slurp=True)` (KDBReader._slurp, kmerdb/fileutil.py:308-466) on small files.
Run in the build container only:

    python tests/golden/make_golden_kdb.py

The files are assembled here from a YAML header and a body text (one BGZF block for the header, the body in blocks of
`block` bytes), so the test can rebuild the identical file from the JSON.  Biopython is absent from the image: the
reference runs against the test-only Bio shim (tests/_bio_shim, bgzf included).  Output: tests/golden/golden_kdb.json.
"""
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "_bio_shim"))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.join(HERE, "..", ".."))

import logging  # noqa: E402

logging.disable(logging.CRITICAL)
import numpy as np  # noqa: E402
from kmerdb import fileutil  # noqa: E402
from kmerdb_b200 import kdbfile  # noqa: E402


def metadata(k, counts):
    total, uniq = int(np.sum(counts)), int(np.count_nonzero(counts))
    f = {"filename": "x.fa", "md5": "a" * 32, "sha256": "b" * 64, "total_reads": 1, "total_kmers": total, "unique_kmers": uniq,
         "nullomers": 4 ** k - uniq, "min_read_length": 10, "max_read_length": 10, "avg_read_length": 10}
    return {"version": "0.9.6", "metadata_blocks": 1, "k": k, "total_kmers": total, "unique_kmers": uniq, "unique_nullomers": 4 ** k - uniq,
            "sorted": False, "tags": [], "files": [f]}


def assemble(path, md, body, block):
    head, _ = kdbfile.header_bytes(md)
    with open(path, "wb") as fh:
        fh.write(kdbfile.bgzf_block(head))
        for o in range(0, len(body), block):
            fh.write(kdbfile.bgzf_block(body[o:o + block]))
        fh.write(kdbfile._BGZF_EOF)


def rows(k, counts, ids=None, eol="\n", tail=""):
    n = 4 ** k
    ids = range(n) if ids is None else ids
    total = max(1, int(np.sum(counts)))
    return "".join("{0}\t{1}\t{2}\t{3}{4}{5}".format(j, i, counts[j], counts[j] / total, tail, eol) for j, i in zip(range(n), ids))


def main():
    rng = np.random.default_rng(20261104)
    cases = []

    def add(name, k, body, block=4096):
        md = metadata(k, np.zeros(1))
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "t.%d.kdb" % k)
            assemble(p, md, body.encode("ascii"), block)
            case = {"name": name, "k": k, "body": body, "block": block}
            try:
                r = fileutil.open(p, mode="r", slurp=True)
                case["counts"] = [int(x) for x in r.counts]
                case["kmer_ids"] = [int(x) for x in r.kmer_ids]
                case["frequencies"] = [float(x) for x in r.frequencies]
            except BaseException as e:                                   # noqa: BLE001  (the class is the golden value)
                case["raises"] = type(e).__name__
        cases.append(case)
        print(name, case.get("raises", "ok"))

    for k in (2, 3, 4):
        c = rng.integers(0, 5000, size=4 ** k)
        add("identity_k%d" % k, k, rows(k, c))
    k = 3
    n = 4 ** k
    c = rng.integers(0, 100000, size=n)
    add("permuted_ids", k, rows(k, c, ids=[int(x) for x in rng.permutation(n)]))
    add("tiny_blocks", k, rows(k, c), block=37)
    add("crlf", k, rows(k, c, eol="\r\n"))
    add("trailing_tab", k, rows(k, c, tail="\t"))
    add("trailing_spaces", k, rows(k, c, tail="  "))
    add("no_final_newline", k, rows(k, c)[:-1])
    add("extra_rows_ignored", k, rows(k, c) + "64\t0\t7\t0.5\n65\t1\t9\t0.5\n")
    big = [int(x) for x in c]
    big[5] = 2 ** 64 - 1
    big[6] = 2 ** 63
    add("u64_max_count", k, "".join("{0}\t{0}\t{1}\t0.0\n".format(j, big[j]) for j in range(n)))
    add("zeros", k, "".join("{0}\t{0}\t0\t0.0\n".format(j) for j in range(n)))
    base = rows(k, c).splitlines(True)

    def broken(j, line):
        b = list(base)
        b[j] = line
        return "".join(b)

    add("three_columns", k, broken(9, "9\t9\t5\n"))
    add("five_columns", k, broken(9, "9\t9\t5\t0.1\tx\n"))
    add("empty_freq", k, broken(9, "9\t9\t5\t\n"))
    add("empty_line", k, broken(9, "\n"))
    add("index_mismatch", k, broken(12, "13\t12\t5\t0.1\n"))
    add("bad_id", k, broken(12, "12\t1a\t5\t0.1\n"))
    add("bad_count", k, broken(12, "12\t12\t2.0\t0.1\n"))
    add("bad_index_text", k, broken(12, "x\t12\t2\t0.1\n"))
    add("id_out_of_range", k, broken(20, "20\t64\t5\t0.1\n"))
    add("short_file", k, "".join(base[:n - 3]))
    add("columns_before_int", k, broken(3, "3\tzz\t5\n") )
    add("first_error_wins", k, broken(30, "30\tq\t5\t0.1\n").replace(base[40], "41\t40\t1\t0.1\n"))
    with open(os.path.join(HERE, "golden_kdb.json"), "w") as fh:
        json.dump({"source": "kmerdb v0.9.6 fileutil.KDBReader._slurp (unmodified) via tests/_bio_shim", "cases": cases}, fh, indent=0, sort_keys=True)
    print("wrote golden_kdb.json with", len(cases), "reader cases")


if __name__ == "__main__":
    main()
