#!/usr/bin/env python
"""Freeze whole-file golden outputs of the UNMODIFIED reference subcommands `kmerdb.profile(args)` and
`kmerdb.make_graph(args)` (kmerdb/__init__.py:1792-2013, :1635-1788): the gunzipped text of every output file
(YAML header + rows).  Run in the build container only:

    python tests/golden/make_golden_cli.py

`kmerdb.cli()` itself cannot start in v0.9.6 (AttributeError: appmap.command_17_description while building the
parser, kmerdb/__init__.py:2365), so the subcommand functions are called with the Namespace argparse would give.
Biopython is absent from the image: the reference runs against tests/_bio_shim.  Output: golden_cli.json.
"""
import argparse
import base64
import gzip
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "_bio_shim"))
sys.path.insert(0, "/root/reference")

import numpy as np  # noqa: E402
import kmerdb  # noqa: E402


def reads(seed, n, length, alphabet="ACGT", p=None):
    rng = np.random.default_rng(seed)
    a = np.array(list(alphabet))
    return ["".join(a[rng.choice(len(a), size=length, p=p)]) for _ in range(n)]


def fasta(rs, width=None):
    out = []
    for i, r in enumerate(rs):
        out.append(">r%d some description" % i)
        out.extend([r] if not width else [r[j:j + width] for j in range(0, len(r), width)])
    return "\n".join(out) + "\n"


def fastq(rs):
    return "".join("@q%d\n%s\n+\n%s\n" % (i, r, "I" * len(r)) for i, r in enumerate(rs))


def main():
    cases = []

    def run(name, func, files, ns, outputs):
        """files: {relative name: bytes}; ns: Namespace fields; outputs: names expected afterwards."""
        cwd = os.getcwd()
        with tempfile.TemporaryDirectory() as d:
            os.chdir(d)
            try:
                for fn, data in files.items():
                    with open(fn, "wb") as fh:
                        fh.write(data)
                kmerdb.logger = kmerdb.kdbLogger.AppLogger(logfile=os.path.join(d, "ref.log"), level=0)
                case = {"name": name, "func": func, "files": {fn: base64.b64encode(data).decode() for fn, data in files.items()}, "args": ns}
                try:
                    getattr(kmerdb, func)(argparse.Namespace(**ns))
                    case["outputs"] = {o: gzip.open(o, "rt").read() for o in outputs}
                except BaseException as e:                                  # noqa: BLE001
                    case["raises"] = type(e).__name__
                cases.append(case)
                print(name, case.get("raises", "ok"), [len(v) for v in case.get("outputs", {}).values()])
            finally:
                os.chdir(cwd)

    def prof(**kw):
        ns = dict(input=["s.fa"], k=4, minK=None, maxK=None, output_name="out", quiet=True, sorted=False, no_ambiguous=False,
                  do_not_canonicalize=False, show_hist=False, verbose=0, debug=False, log_file=None, num_log_lines=4, parallel=1)
        ns.update(kw)
        return ns

    plain = fasta(reads(1, 30, 80)).encode()
    run("profile_k4", "profile", {"s.fa": plain}, prof(), ["out.4.kdb"])
    run("profile_k6_noncanonical", "profile", {"s.fa": plain}, prof(k=6, do_not_canonicalize=True), ["out.6.kdb"])
    withn = fasta(reads(2, 20, 70, "ACGTN", [0.24, 0.24, 0.24, 0.24, 0.04]), width=30).encode()
    run("profile_k5_skip_n", "profile", {"s.fa": withn}, prof(k=5, no_ambiguous=True), ["out.5.kdb"])
    run("profile_k4_expand_n", "profile", {"s.fa": withn}, prof(k=4), ["out.4.kdb"])
    run("profile_fastq", "profile", {"r.fastq": fastq(reads(3, 25, 60)).encode()}, prof(input=["r.fastq"], k=5), ["out.5.kdb"])
    run("profile_fasta_gz", "profile", {"g.fa.gz": gzip.compress(plain, mtime=0)}, prof(input=["g.fa.gz"], k=5), ["out.5.kdb"])
    run("profile_multik", "profile", {"s.fa": plain}, prof(k=None, minK=4, maxK=6), ["out.4.kdb", "out.5.kdb", "out.6.kdb"])
    other = fasta(reads(4, 10, 120)).encode()
    run("profile_samplesheet", "profile", {"a.fa": plain, "b.fa": other, "sheet.txt": b"a.fa\nb.fa\n"}, prof(input=["sheet.txt"], k=5), ["out.5.kdb"])
    run("profile_lowercase_raises", "profile", {"s.fa": b">x\nACGTacgtACGT\n"}, prof(), ["out.4.kdb"])
    run("profile_iupac_raises", "profile", {"s.fa": b">x\nACGTRYACGTAC\n"}, prof(), ["out.4.kdb"])
    run("profile_short_raises", "profile", {"s.fa": b">x\nACGTACGTAC\n>y\nAC\n"}, prof(), ["out.4.kdb"])

    def gr(**kw):
        ns = dict(input=["s.fa"], k=5, kdbg="g.kdbg", quiet=True, sorted=False, do_not_canonicalize=False, replace_with_none=False,
                  verbose=0, debug=False, log_file=None, num_log_lines=4, parallel=1)
        ns.update(kw)
        return ns

    small = fasta(reads(5, 6, 40)).encode()
    run("graph_k5", "make_graph", {"s.fa": small}, gr(), ["g.kdbg"])
    run("graph_k4_noncanonical", "make_graph", {"s.fa": small}, gr(k=4, do_not_canonicalize=True), ["g.kdbg"])
    with open(os.path.join(HERE, "golden_cli.json"), "w") as fh:
        json.dump({"source": "kmerdb v0.9.6 kmerdb.profile / kmerdb.make_graph (unmodified) via tests/_bio_shim", "cases": cases}, fh, indent=0, sort_keys=True)
    print("wrote golden_cli.json with", len(cases), "cases")


if __name__ == "__main__":
    main()
