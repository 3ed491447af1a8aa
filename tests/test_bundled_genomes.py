"""configs[0] / parity gate T0: the reference's own bundled genomes (tests/golden/data, copied from
/root/reference/test/data by tests/golden/make_golden_bundled.py) — multi-Mbp records in 80-column lines, real
gzip files — against golden_bundled.json, which holds the output of the UNMODIFIED reference's parse.parsefile on
those files and the count columns of the fixtures the reference keeps for them.

not gpu: the C oracle against those goldens.   gpu: parse.parsefile and the CLI through the C-ABI, full-vector."""
import argparse
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from kmerdb_b200 import cli, parse

HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(HERE, "golden", "data")
GOLD = json.load(open(os.path.join(HERE, "golden", "golden_bundled.json")))
CASES = GOLD["cases"]


def _id(c):
    return "%s-k%d-%s" % (c["file"].split(".")[0], c["k"], "canon" if c["canonicalize"] else "fwd")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def _raw(f):
    path = os.path.join(DATA, f)
    return gzip.open(path, "rb").read() if f.endswith(".gz") else open(path, "rb").read()


def test_inputs_are_the_reference_files():
    for f, md5 in GOLD["files"].items():
        assert hashlib.md5(open(os.path.join(DATA, f), "rb").read()).hexdigest() == md5
    # md5 of the files as the reference ships them
    assert GOLD["files"]["Cacetobutylicum_ATCC824.fasta.gz"] == "0a0f73e1c8b8285703e29279bafaabef"
    assert GOLD["files"]["Cdifficile_R3.fasta.gz"] == "ed09041ad0e991a8037edd0f05643571"


@pytest.mark.parametrize("c", CASES, ids=_id)
def test_oracle_matches_reference_on_bundled_genomes(c):
    rc, counts, st = oracle.profile(_raw(c["file"]), c["k"], canonical=c["canonicalize"], skip_n=c["replace_with_none"], threads=4)
    assert rc == 0
    assert _sha(counts) == c["counts_sha256"]
    m = c["meta"]
    assert (st["total_reads"], st["total_kmers"], st["unique_kmers"]) == (m["total_reads"], m["total_kmers"], m["unique_kmers"])
    assert (st["min_len"], st["max_len"]) == (m["min_read_length"], m["max_read_length"])
    fx = [f for f in GOLD["fixtures"].values() if (f["file"], f["k"], f["canonicalize"]) == (c["file"], c["k"], c["canonicalize"])]
    for f in fx:
        if "counts_sha256" in f:
            assert _sha(counts) == f["counts_sha256"]
        assert (st["total_kmers"], st["unique_kmers"], 4 ** c["k"] - st["unique_kmers"]) == (f["total_kmers"], f["unique_kmers"], f["nullomers"])
        if "mononucleotides" in f:
            assert st["mono"] == [f["mononucleotides"][x] for x in "ACGT"]


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES, ids=_id)
def test_parsefile_on_bundled_genomes(c):
    path = os.path.join(DATA, c["file"])
    counts, meta, nullomers = parse.parsefile(path, c["k"], replace_with_none=c["replace_with_none"], canonicalize=c["canonicalize"])
    assert counts.dtype == np.uint64 and counts.shape == (4 ** c["k"],)
    assert _sha(counts) == c["counts_sha256"]                       # the full vector
    assert int(counts.sum()) == c["counts_sum"] and [int(x) for x in counts[:8]] == c["counts_head"]
    assert meta["filename"] == path
    got = {a: b for a, b in meta.items() if a != "filename"}
    mono = got.pop("mononucleotides", None)
    assert got == c["meta"]                                         # incl. md5 / sha256 of the compressed file
    assert len(nullomers) == c["n_nullomers"] and _sha(nullomers) == c["nullomers_sha256"]
    for f in GOLD["fixtures"].values():
        if (f["file"], f["k"], f["canonicalize"]) != (c["file"], c["k"], c["canonicalize"]):
            continue
        if "counts_sha256" in f:
            assert _sha(counts) == f["counts_sha256"]               # the fixture the reference's own tests hold
        assert (meta["total_kmers"], meta["unique_kmers"], meta["nullomers"]) == (f["total_kmers"], f["unique_kmers"], f["nullomers"])
        if "mononucleotides" in f and mono is not None:
            assert mono == f["mononucleotides"]


@pytest.mark.gpu
@pytest.mark.parametrize("c", [c for c in CASES if c["k"] == 8], ids=_id)
def test_cli_profile_on_bundled_genomes(c, tmp_path):
    """`kmerdb profile -k 8 -o out <genome>`: header totals of test/test_fileutil.py:15-43 and every body row."""
    out = str(tmp_path / "out")
    ns = argparse.Namespace(device=0, input=[os.path.join(DATA, c["file"])], k=c["k"], minK=None, maxK=None, output_name=out,
                            no_ambiguous=c["replace_with_none"], do_not_canonicalize=not c["canonicalize"], sorted=False, quiet=True,
                            debug=False, verbose=0, num_log_lines=0, logfile=None)
    cli.profile(ns)
    text = gzip.open(out + ".%d.kdb" % c["k"], "rt").read()
    header, body = text.split("\n" + "=" * 24 + "\n")
    import yaml
    h = yaml.safe_load(header)
    N = 4 ** c["k"]
    assert (h["k"], h["total_kmers"], h["unique_kmers"]) == (c["k"], c["meta"]["total_kmers"], c["meta"]["unique_kmers"])
    assert h["unique_nullomers"] == (int(N / 2 - h["unique_kmers"]) if c["canonicalize"] else N - h["unique_kmers"])
    assert h["files"][0]["md5"] == c["meta"]["md5"] and h["files"][0]["total_reads"] == c["meta"]["total_reads"]
    rows = [ln.split("\t") for ln in body.splitlines() if ln]
    assert len(rows) == N
    vec = np.zeros(N, dtype=np.uint64)
    for i, r in enumerate(rows):
        assert int(r[0]) == i and int(r[1]) == i
        vec[i] = int(r[2])
    assert _sha(vec) == c["counts_sha256"]
