"""The CPU oracle (oracle/kmerdb_oracle.c) against the golden vectors produced by the UNMODIFIED
reference (tests/golden/make_golden.py) and, when /root/reference is mounted, against the
reference's own fixtures."""
import gzip
import hashlib
import os

import numpy as np
import pytest

import oracle

REF = "/root/reference"


def _case_id(c):
    return "%s-k%d-rn%d-c%d" % (c["name"], c["k"], c["replace_with_none"], c["canonicalize"])


def _golden_cases():
    import json
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "golden", "golden_v1.json")) as fh:
        return json.load(fh)["parsefile"]


def check_against_expect(c, rc, counts, stats):
    e = c["expect"]
    if "raises" in e:
        assert rc != 0, "oracle accepted an input the reference rejects"
        assert oracle.ERR_NAMES[rc] == e["raises"]
        return
    assert rc == 0
    meta = oracle.metadata(stats, c["k"])
    assert meta == e["meta"]
    assert hashlib.sha256(counts.astype("<u8").tobytes()).hexdigest() == e["counts_sha256"]
    nul = np.nonzero(counts == 0)[0].astype("<u8")
    assert len(nul) == e["n_nullomers"]
    assert hashlib.sha256(nul.tobytes()).hexdigest() == e["nullomers_sha256"]
    if e["nonzero_ids"] is not None:
        nz = np.nonzero(counts)[0]
        assert nz.tolist() == e["nonzero_ids"]
        assert counts[nz].tolist() == e["nonzero_counts"]


@pytest.mark.parametrize("c", [c for c in _golden_cases() if c["suffix"] != ".txt"], ids=_case_id)
@pytest.mark.parametrize("threads", [1, 3])
def test_profile_matches_reference_golden(c, threads):
    fastq = c["suffix"] in (".fq", ".fastq")
    rc, counts, stats = oracle.profile(c["text"].encode(), c["k"], fastq=fastq, skip_n=c["replace_with_none"],
                                       canonical=c["canonicalize"], threads=threads)
    check_against_expect(c, rc, counts, stats)


def test_scalar_ids(golden):
    s = golden["scalars"]
    two = [a + b for a in "ACGT" for b in "ACGT"]
    L = oracle.lib()
    assert [L.kdbo_kmer_to_id(x.encode(), 2, 0) for x in two] == s["kmer_to_id_2mers_noncanonical"] == list(range(16))
    assert [L.kdbo_kmer_to_id(x.encode(), 2, 1) for x in two] == s["kmer_to_id_2mers_canonical"]
    assert s["kmer_to_id_ATCNATC"] is None and L.kdbo_kmer_to_id(b"ATCNATC", 7, 1) == -1
    for kmer, (canon, fwd) in s["kmer_to_id_misc"].items():
        assert L.kdbo_kmer_to_id(kmer.encode(), len(kmer), 1) == canon
        assert L.kdbo_kmer_to_id(kmer.encode(), len(kmer), 0) == fwd


def test_graph_edges(golden):
    g = golden["graph"]
    rc, rows = oracle.edges(g["text"].encode(), g["k"])
    assert rc == 0
    names = ["r1", "r2"]
    got = [[names[int(r[0])]] + [int(x) for x in r[1:]] for r in rows]
    assert got == g["edges"]


def test_minimizers_compat(golden):
    for key, exp in golden["minimizers"].items():
        seq, k, w = key.split(",")
        rc, ids, flags = oracle.minimizers_compat(seq, int(k), int(w))
        assert rc == 0
        assert [[i, int(ids[i]), int(flags[i])] for i in range(len(ids))] == [r[1:] for r in exp]


needs_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "test", "data")), reason="reference tree not mounted")


def _kdb_rows(path):
    with gzip.open(path, "rt") as fh:
        text = fh.read()
    body = text.split("\n" + "=" * 24 + "\n")[-1]
    return [ln.split("\t") for ln in body.splitlines() if ln.strip()]


@needs_ref
def test_fixture_B1_cac_k8_noncanonical():
    raw = gzip.open(os.path.join(REF, "test/data/Cacetobutylicum_ATCC824.fasta.gz"), "rb").read()
    rc, counts, st = oracle.profile(raw, 8, canonical=False, threads=4)
    assert rc == 0
    rows = _kdb_rows(os.path.join(REF, "test/data/test_Cac_ATCC824.8.kdb"))
    exp = np.zeros(4 ** 8, dtype=np.uint64)
    for r in rows:
        exp[int(r[1])] = int(r[2])
    assert np.array_equal(counts, exp)
    assert (st["total_kmers"], st["unique_kmers"], 4 ** 8 - st["unique_kmers"]) == (4132866, 64103, 1433)


@needs_ref
def test_fixture_B3_B4_cdifficile():
    raw = gzip.open(os.path.join(REF, "test/data/Cdifficile_R3.fasta.gz"), "rb").read()
    rc, counts, st = oracle.profile(raw, 8, canonical=False, threads=4)
    assert rc == 0
    rows = _kdb_rows(os.path.join(REF, "test/data/example_v0.6.3_kdb_file.8.kdb"))
    exp = np.zeros(4 ** 8, dtype=np.uint64)
    for r in rows:
        exp[int(r[1])] = int(r[3])
    assert np.array_equal(counts, exp)
    assert st["mono"] == [1444649, 601738, 590294, 1456466]
    rc, counts, st = oracle.profile(raw, 12, canonical=False, threads=4)   # kmerdb/appmap.py:271-306
    assert (st["total_kmers"], st["unique_kmers"], 4 ** 12 - st["unique_kmers"]) == (4093136, 2267911, 14509305)


def test_numpy_synth_equals_the_product_generator():
    """oracle/synth.py (NumPy Philox-4x32-10, used by bench.py's CPU legs so that the reference arm never loads the
    product library) against kdb_synth_host (the host twin of the k_synth kernel), byte for byte."""
    from kmerdb_b200 import _native
    from oracle import synth
    assert synth.fasta(0, 700) == bytes(_native.synth_host(700))
    assert synth.fasta(5_000_000_123, 5) == bytes(_native.synth_host(5, first_read=5_000_000_123))
    assert synth.fasta(3, 4, read_len=70, seed=99) == bytes(_native.synth_host(4, 70, seed=99, first_read=3))
