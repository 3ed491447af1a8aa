"""`kmerdb profile | graph | minimizers` with the reference's flag names (kmerdb/__init__.py:2084-2145), on the
B200 kernels.  `python -m kmerdb_b200.cli profile -k 12 -o out reads.fa.gz` writes `out.12.kdb`.

Only the three subcommands of the hot path exist here; everything else stays in kmerdb."""
import argparse
import os
import sys
import time
from collections import OrderedDict

import numpy as np

from . import REFERENCE_VERSION, graph, kdbfile, kmer, minimizer, parse


def _inputs(args):
    """kmerdb/__init__.py:1822-1844: exactly one positional input; a .txt/.tsv input is a samplesheet."""
    if len(args.input) != 1:
        raise ValueError("Could not determine POSIX access mode for one or more input files.")
    name = args.input[0]
    if any(s in name for s in (".fastq", ".fq", ".fasta", ".fa", ".fna")):
        return [name]
    if ".txt" in name or ".tsv" in name:
        samples = []
        with open(name, "r") as fh:
            for line in fh:
                sample = line.rstrip()
                if os.access(sample, os.R_OK):
                    samples.append(sample)
                else:
                    raise ValueError("Couldn't open sample file '{0}' for reading".format(sample))
        return samples
    raise ValueError("Could not determine file type of input")


def _profile(args, k, files, precomputed=None):
    """kmerdb/__init__.py:1862-2013 for one k.  `precomputed`: [(counts, file_metadata)] per file from the shared sweep."""
    N = 4 ** k
    counts = np.zeros(N, dtype="uint64")
    file_metadata = []
    for i, f in enumerate(files):
        if precomputed is not None:
            c, meta = precomputed[i]
        else:
            # the per-file metadata arrives as a callable: md5 / sha256 of a large file finish seconds after its counts
            c, meta, _nullomers = parse.parsefile_deferred(f, k, replace_with_none=args.no_ambiguous, canonicalize=not args.do_not_canonicalize,
                                                           device=getattr(args, "devices", None) or args.device)
        counts = counts + c                                        # __init__.py:1890
        file_metadata.append(meta)
    total = int(np.sum(counts))
    unique = int(np.count_nonzero(counts))
    unique_nullomers = N - unique if args.do_not_canonicalize else int((N / 2) - unique)      # __init__.py:1903

    def metadata():
        return OrderedDict({"version": REFERENCE_VERSION, "metadata_blocks": 1, "k": k, "total_kmers": total, "unique_kmers": unique,
                            "unique_nullomers": unique_nullomers, "sorted": args.sorted, "tags": [],
                            "files": [m() if callable(m) else m for m in file_metadata]})
    if args.sorted:
        sys.stderr.write("--sorted is broken in kmerdb v0.9.6 (undefined names at kmerdb/__init__.py:1974-1975); rows are written in id order\n")
    out = "{0}.{1}.kdb".format(args.output_name, k)                 # __init__.py:1954
    # up to k = 12 the compressed body (about 100 MB) waits in memory for the header; larger tables take the header first
    kdbfile.write_kdb_gpu(out, metadata if k <= 12 else metadata(), counts, device=args.device, quiet=args.quiet, total_kmers=total)
    sys.stderr.write("\n\n\nFinal stats:\n\n\n")
    sys.stderr.write("Total k-mers processed: {0}\n".format(total))
    sys.stderr.write("Unique nullomer count:   {0}\n".format(unique_nullomers))
    sys.stderr.write("Unique {0}-mer count:     {1}\n".format(k, unique))
    sys.stderr.write("Theoretical {0}-mer number (4^{0}):     {1}\n".format(k, N))
    return out


def profile(args):
    files = _inputs(args)
    if args.k is None:
        if args.minK is None or args.maxK is None:
            raise ValueError("In multi-k mode, arguments --minK and --maxK are required.")
        elif args.minK < 4 or args.maxK < 5 or args.minK > 25 or args.maxK > 30:
            raise ValueError("Valid --min|max between 3 and 30")
        # one scan + encode per file for the whole sweep (kdb_attach); outputs appear in increasing k like the reference loop
        sweeps, outs = [], []
        for f in files:
            md5, sha256 = parse.checksum(f)
            it = parse.iter_profile_multik(parse._read_text_bytes(f), args.minK, args.maxK, fmt=parse._file_format(f),
                                           replace_with_none=args.no_ambiguous, canonicalize=not args.do_not_canonicalize, device=args.device)
            sweeps.append((f, md5, sha256, it))
        for k in range(args.minK, args.maxK + 1):
            pre = []
            for f, md5, sha256, it in sweeps:
                kk, prof = next(it)
                assert kk == k
                pre.append((prof.counts, prof.file_metadata(f, md5, sha256)))
            outs.append(_profile(args, k, files, precomputed=pre))
        return outs
    return [_profile(args, args.k, files)]


def make_graph(args):
    """kmerdb/__init__.py:1635-1788: one 8-column row per adjacent k-mer pair (rows rendered on the host)."""
    if not args.kdbg.endswith(".kdbg"):
        raise IOError("kmerdb graph expects a '.kdbg' suffixed output file")        # __init__.py:1663-1664
    N = 4 ** args.k
    counts = np.zeros(N, dtype="uint64")
    rows, metas = [], []
    for f in args.input:
        data, meta, counts_ = graph.make_edges_from_fasta(f, args.k, quiet=args.quiet, canonicalize=not args.do_not_canonicalize,
                                                           replace_with_none=args.replace_with_none, device=args.device)
        rows += data
        counts = counts + counts_                                                  # __init__.py:1676
        metas.append(meta)
    body = "".join("{0}\t{1}\t{2}\t{3}\t{4}\t{5}\t{6}\t{7}\n".format(i, sid, p1, i1, kmer.id_to_kmer(i1, args.k), p2, i2, kmer.id_to_kmer(i2, args.k))
                   for i, (sid, p1, i1, p2, i2) in enumerate(rows))             # __init__.py:1756-1762
    if not args.quiet:
        sys.stdout.write(body)
    unique = int(np.count_nonzero(counts))
    unique_nullomers = N - unique if args.do_not_canonicalize else int((N / 2) - unique)    # __init__.py:1712
    md = OrderedDict({"version": REFERENCE_VERSION, "metadata_blocks": 1, "k": args.k, "total_kmers": sum(m["total_kmers"] for m in metas),
                      "unique_kmers": unique, "unique_nullomers": unique_nullomers, "sorted": args.sorted, "tags": [], "files": metas})   # :1714-1724
    head, _ = kdbfile.header_bytes(md)
    with open(args.kdbg, "wb") as fh:
        for o in range(0, len(head), 65536):
            fh.write(kdbfile.bgzf_block(head[o:o + 65536]))
        raw = body.encode("ascii")
        for o in range(0, len(raw), 65280):
            fh.write(kdbfile.bgzf_block(raw[o:o + 65280]))
        fh.write(kdbfile._BGZF_EOF)
    return args.kdbg


def get_minimizers(args):
    """kmerdb/__init__.py:425-450.  k comes from the .kdb header; output `<kdb stripped>.kdbi.1`."""
    md = kdbfile.KDBReader(args.kdb, slurp=False).metadata
    out = str(args.kdb).strip(".kdb") + ".kdbi.1"                   # __init__.py:437 (character strip, as in the reference)
    with open(out, "w") as fh:
        for rid, seq in parse.parse_sequence_file(args.fasta):
            for sid, i, kid, is_min in minimizer.select_lexicographical_minimizers(seq, rid, md["k"], args.window_size, device=args.device):
                fh.write("{0}\t{1}\t{2}\t{3}\n".format(sid, i, kid, is_min))    # minimizer.py:299-320
    return out


def build_parser():
    ap = argparse.ArgumentParser(prog="kmerdb")
    sub = ap.add_subparsers(dest="command", required=True)

    def common(p):
        p.add_argument("-v", "--verbose", default=0, action="count")
        p.add_argument("--quiet", action="store_true", default=False)
        p.add_argument("--debug", action="store_true", default=False)
        p.add_argument("-nl", "--num-log-lines", type=int, default=50)
        p.add_argument("-l", "--log-file", type=str, default="kmerdb.log")
        p.add_argument("--device", type=int, default=0, help="CUDA device ordinal (kmerdb_b200 extension)")

    p = sub.add_parser("profile")
    common(p)
    p.add_argument("-k", type=int)
    p.add_argument("--minK", type=int)
    p.add_argument("--maxK", type=int)
    p.add_argument("-o", "--output-name", type=str, required=True)
    p.add_argument("--no-ambiguous", action="store_true")
    p.add_argument("--do-not-canonicalize", action="store_true")
    p.add_argument("--sorted", action="store_true", default=False)
    p.add_argument("--devices", type=str, default=None,
                   help="several CUDA devices, e.g. 0-7 or 0,2,4 (kmerdb_b200 extension): the reads of each file are sharded over them, "
                        "the partial tables are summed over NVLink into the first one; the .kdb is byte-identical to the one-GPU file")
    p.add_argument("input", nargs="+", type=str)
    p.set_defaults(func=profile)

    g = sub.add_parser("graph")
    common(g)
    g.add_argument("-k", default=12, type=int, required=True)
    g.add_argument("--replace-with-none", action="store_true")
    g.add_argument("--do-not-canonicalize", action="store_true")
    g.add_argument("--sorted", action="store_true", default=False)
    g.add_argument("input", nargs="+", type=str)
    g.add_argument("kdbg", type=str)
    g.set_defaults(func=make_graph)

    m = sub.add_parser("minimizers")
    common(m)
    m.add_argument("fasta", type=str)
    m.add_argument("kdb", type=str)
    m.add_argument("-w", "--window-size", type=int, required=True)
    m.set_defaults(func=get_minimizers)
    return ap


# exception class -> exit code, as kmerdb.cli maps them (kmerdb/__init__.py:2474-2530)
EXIT_CODES = [(TypeError, 1), (ValueError, 2), (KeyError, 3), (RuntimeError, 4), (OSError, 5), (IOError, 6), (ArithmeticError, 7),
              (AssertionError, 8), (AttributeError, 9)]


def cli(argv=None):
    args = build_parser().parse_args(argv)
    t0 = time.time()
    try:
        args.func(args)
    except Exception as e:
        if args.debug:
            raise
        sys.stderr.write("{0}: {1}\n".format(type(e).__name__, e))
        for cls, code in EXIT_CODES:
            if isinstance(e, cls):
                sys.exit(code)
        sys.exit(-1)
    sys.stderr.write("Program ran for {0} seconds...\n".format(round(time.time() - t0, 2)))


if __name__ == "__main__":
    cli()
