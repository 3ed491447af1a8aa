#!/usr/bin/env python
"""Freeze golden vectors by running the UNMODIFIED reference (/root/reference/kmerdb/{kmer,parse,
graph,minimizer}.py) on small inputs.  Run in the build container only:

    python tests/golden/make_golden.py

Biopython is absent from the image, so the reference modules are imported against the test-only
shim in tests/_bio_shim (see its docstring).  Output: tests/golden/golden_v1.json (committed).
Nothing in the product, the -m gpu tests, smoke() or bench.py reads /root/reference at run time;
they read the JSON this script wrote.
"""
import gzip
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "_bio_shim"))
sys.path.insert(0, "/root/reference")

import logging  # noqa: E402

logging.disable(logging.CRITICAL)
import numpy as np  # noqa: E402
from kmerdb import kmer, parse, graph, minimizer  # noqa: E402


def rand_reads(seed, n, length, alphabet="ACGT", p=None):
    rng = np.random.default_rng(seed)
    letters = np.array(list(alphabet))
    idx = rng.choice(len(letters), size=(n, length), p=p)
    return ["".join(letters[row]) for row in idx]


def fasta_text(reads, width=None, eol="\n", ids=None):
    out = []
    for i, r in enumerate(reads):
        out.append(">" + (ids[i] if ids else "r%06d some description" % i) + eol)
        if width:
            for j in range(0, len(r), width):
                out.append(r[j:j + width] + eol)
        else:
            out.append(r + eol)
    return "".join(out)


def fastq_text(reads, eol="\n", plus_repeat=False):
    out = []
    for i, r in enumerate(reads):
        t = "q%06d/1" % i
        out.append("@" + t + eol + r + eol + "+" + (t if plus_repeat and i % 2 else "") + eol)
        # quality strings deliberately contain '@' and '+' as first characters
        q = "".join("@+I5#"[(i + j) % 5] for j in range(len(r)))
        out.append(q + eol)
    return "".join(out)


def run_parsefile(text, suffix, k, replace_with_none, canonicalize, gz=False):
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "input" + suffix + (".gz" if gz else ""))
        data = text.encode()
        with (gzip.open(path, "wb") if gz else open(path, "wb")) as fh:
            fh.write(data)
        try:
            counts, meta, nullomers = parse.parsefile(path, k, replace_with_none=replace_with_none, canonicalize=canonicalize)
        except BaseException as e:  # noqa: BLE001  (the class name IS the golden value)
            return {"raises": type(e).__name__}
        nz = np.nonzero(counts)[0]
        meta = dict(meta)
        for key in ("filename", "md5", "sha256"):
            meta.pop(key)
        import hashlib
        return {
            "meta": {a: int(b) for a, b in meta.items()},
            "nonzero_ids": [int(x) for x in nz] if len(nz) <= 4096 else None,
            "nonzero_counts": [int(x) for x in counts[nz]] if len(nz) <= 4096 else None,
            "counts_sha256": hashlib.sha256(counts.astype("<u8").tobytes()).hexdigest(),
            "nullomers_sha256": hashlib.sha256(nullomers.astype("<u8").tobytes()).hexdigest(),
            "n_nullomers": int(len(nullomers)),
        }


def main():
    cases = []

    def add(name, text, suffix, k, modes=((True, True), (True, False), (False, True), (False, False)), gz=False):
        for rn, canon in modes:
            res = run_parsefile(text, suffix, k, rn, canon, gz=gz)
            cases.append({"name": name, "text": text, "suffix": suffix, "gz": gz, "k": k,
                          "replace_with_none": rn, "canonicalize": canon, "expect": res})
            print(name, k, rn, canon, res.get("raises") or res["meta"], flush=True)

    r150 = rand_reads(1234, 300, 150)
    add("reads150_fasta", fasta_text(r150), ".fa", 8)
    add("reads150_fasta", fasta_text(r150), ".fa", 12, modes=((True, True), (True, False)))
    add("reads150_fasta_gz", fasta_text(r150[:60]), ".fasta", 6, modes=((True, True),), gz=True)
    add("reads150_fastq", fastq_text(r150[:120]), ".fq", 8, modes=((True, True), (True, False)))
    add("reads150_fastq_crlf", fastq_text(r150[:40], eol="\r\n", plus_repeat=True), ".fastq", 5, modes=((True, True),))
    genome = rand_reads(7, 3, 2500)
    add("multiline70", fasta_text(genome, width=70, ids=["chr1 first", "chr2", "plasmid x y z"]), ".fna", 9,
        modes=((True, True), (True, False)))
    add("multiline60_crlf", fasta_text(genome[:2], width=60, eol="\r\n"), ".fa", 4, modes=((True, True), (True, False)))
    add("multiline_blank_and_spaces",
        "junk before first header\n>r1 d\nACGT TGCA\n\nAC\r\nGGT  \n>r2\n\nTTTTTTTT\n>r3\nACGTACGTAC", ".fa", 4,
        modes=((True, True), (True, False)))
    withn = rand_reads(99, 40, 80, alphabet="ACGTN", p=[0.245, 0.245, 0.245, 0.245, 0.02])
    add("reads_with_N", fasta_text(withn), ".fa", 5)
    add("reads_with_N", fasta_text(withn), ".fa", 3, modes=((False, True), (False, False)))
    add("n_runs", ">a\nACGTNNNNACGTACNNGT\n>b\nNNNNNNNN\n>c\nACGTACGT\n", ".fa", 4)
    add("exact_k", ">a\nACGTAC\n>b\nTTTTTT\n", ".fa", 6, modes=((True, True), (True, False)))
    add("B10_fastq", "@q1 x\nACGTTGCA\n+\nIIIIIIII\n@q2\nGGGGCCCC\n+q2\nIIIIIIII\n", ".fq", 4, modes=((True, True),))
    # error classes
    add("err_lowercase", ">a\nACGTACGTAC\n>b\nacgtacgtac\n", ".fa", 4, modes=((True, True), (False, True)))
    add("err_iupac", ">a\nACGTACGTAC\n>b\nACGTRCGTAC\n", ".fa", 4, modes=((True, True), (False, True)))
    add("err_iupac_hidden_by_N", ">a\nACGNRNGTACGT\n", ".fa", 3, modes=((True, True), (False, True)))
    add("err_short", ">a\nACGTACGTAC\n>b\nACG\n>c\nACGTACGT\n", ".fa", 4, modes=((True, True),))
    add("err_short_then_lower", ">a\nACG\n>b\nacgtacgt\n", ".fa", 4, modes=((True, True),))
    add("err_lower_and_short_same_record", ">a\nacg\n", ".fa", 4, modes=((True, True),))
    add("err_iupac_then_lower_same_record", ">a\nACGTRCGTACGTacgt\n", ".fa", 4, modes=((True, True),))
    add("err_iupac_rec_then_lower_rec", ">a\nACGTRCGTACGT\n>b\nacgtacgt\n", ".fa", 4, modes=((True, True),))
    add("err_empty_record", ">a\n>b\nACGTACGT\n", ".fa", 4, modes=((True, True),))
    add("err_empty_file", "", ".fa", 4, modes=((True, True),))
    add("err_no_header", "ACGTACGT\nACGT\n", ".fa", 4, modes=((True, True),))
    add("err_tab_inside", ">a\nACGT\tACGT\n", ".fa", 4, modes=((True, True),))
    add("ok_tab_trailing", ">a\nACGTACGT\t \n>b\nACGTAAAA\n", ".fa", 4, modes=((True, True),))
    add("err_suffix", ">a\nACGTACGT\n", ".txt", 4, modes=((True, True),))
    add("err_fastq_no_at", "q1\nACGTACGT\n+\nIIIIIIII\n", ".fq", 4, modes=((True, True),))
    add("err_fastq_len", "@q1\nACGTACGT\n+\nIIIIIII\n", ".fq", 4, modes=((True, True),))
    add("err_fastq_trunc", "@q1\nACGTACGT\n+\n", ".fq", 4, modes=((True, True),))

    scalars = {
        "kmer_to_id_2mers_noncanonical": [kmer.kmer_to_id(a + b, canonicalize=False) for a in "ACGT" for b in "ACGT"],
        "kmer_to_id_2mers_canonical": [kmer.kmer_to_id(a + b) for a in "ACGT" for b in "ACGT"],
        "kmer_to_id_ATCNATC": kmer.kmer_to_id("ATCNATC"),
        "kmer_to_id_misc": {s: [kmer.kmer_to_id(s), kmer.kmer_to_id(s, canonicalize=False)]
                            for s in ["ACGT", "AATT", "GATTACA", "TTTTTTTTTTTT", "ACGTACGTACGTACG", "CAGCTAGCTAGCTAGGATCGATCGATCGA"]},
        "id_to_kmer": {"%d,%d" % (i, k): kmer.id_to_kmer(i, k) for i, k in [(0, 4), (27, 4), (255, 4), (6, 3), (16777215, 12), (123456, 12)]},
        "shred": {},
    }
    for s, k in [("ACGTNACG", 3), ("ACGTNNAC", 3), ("ACNGT", 3), ("ACGTTGCAACGGT", 4)]:
        for rn in (True, False):
            for canon in (True, False):
                ids, sids, pos = kmer.shred(s, k, replace_with_none=rn, canonicalize=canon)
                scalars["shred"]["%s,%d,%d,%d" % (s, k, rn, canon)] = {"ids": ids, "pos": pos}

    # graph and reference-compatible minimizer rows (A10)
    gtext = ">r1 desc\nACGTTGCAAC\nGGT\n>r2\nTTTTTTTT\n"
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "g.fa")
        open(p, "w").write(gtext)
        res = graph.make_edges_from_fasta(p, 4, quiet=True, canonicalize=True, replace_with_none=True) \
            if "replace_with_none" in graph.make_edges_from_fasta.__code__.co_varnames else graph.make_edges_from_fasta(p, 4, True)
    edges, gmeta, gcounts = res[0], res[1], res[2]
    gmeta = {a: (int(b) if isinstance(b, (int, np.integer)) else b) for a, b in dict(gmeta).items() if a not in ("filename", "md5", "sha256")}
    graph_case = {"text": gtext, "k": 4, "edges": [[str(e[0])] + [int(x) for x in e[1:]] for e in edges],
                  "meta": gmeta, "nonzero": {int(i): int(gcounts[i]) for i in np.nonzero(gcounts)[0]}}
    print("graph", len(edges), gmeta, flush=True)

    mins = {}
    for seq, k, w in [("ACGTTGCAACGGTACGTTGCA", 2, 3), ("ACGTTGCAACGGTACGTTGCAACGGTTTGACCATGACAGTAGGGATACCAGATTTTAGGACCCAAA", 3, 5)]:
        try:
            rows = minimizer.select_lexicographical_minimizers(seq, "s1", k, w)
            mins["%s,%d,%d" % (seq, k, w)] = [[r[0], int(r[1]), int(r[2]), int(r[3])] for r in rows]
        except BaseException as e:  # noqa: BLE001
            mins["%s,%d,%d" % (seq, k, w)] = {"raises": type(e).__name__}
    print("minimizers", {a: (len(b) if isinstance(b, list) else b) for a, b in mins.items()}, flush=True)

    out = {"generator": "tests/golden/make_golden.py", "reference": "kmerdb v0.9.6 (unmodified) + tests/_bio_shim",
           "parsefile": cases, "scalars": scalars, "graph": graph_case, "minimizers": mins}
    with open(os.path.join(HERE, "golden_v1.json"), "w") as fh:
        json.dump(out, fh, indent=0, sort_keys=True)
    print("wrote golden_v1.json with", len(cases), "parsefile cases")


if __name__ == "__main__":
    main()
