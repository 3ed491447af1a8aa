"""Parity of the K4 rows (graph edge rows, aggregated edge weights, minimizers), the host kmer.shred mirror
and the multi-k sweep against the golden vectors of the unmodified reference and the CPU oracle."""
import collections

import numpy as np
import pytest

import oracle
from kmerdb_b200 import _native, graph, kmer, minimizer, parse
from test_gpu_parity import _random_fasta, _random_fastq

pytestmark = pytest.mark.gpu


def test_graph_golden_rows(golden, tmp_path):
    g = golden["graph"]
    path = str(tmp_path / "g.fa")
    open(path, "w").write(g["text"])
    data, meta, counts = graph.make_edges_from_fasta(path, g["k"])
    assert [list(r) for r in data] == g["edges"]
    got = {a: b for a, b in meta.items() if a not in ("filename", "md5", "sha256")}
    assert got == g["meta"]
    nz = np.nonzero(counts)[0]
    assert {str(int(i)): int(counts[i]) for i in nz} == g["nonzero"]


@pytest.mark.parametrize("k", [2, 5, 9, 13, 20])
@pytest.mark.parametrize("canonical", [True, False])
def test_edge_rows_vs_oracle(k, canonical, tmp_path):
    rng = np.random.default_rng(40 + k)
    text = _random_fasta(rng, 150, max(k, 25), 500, 70)
    rc, rows = oracle.edges(text, k, canonical=canonical)
    assert rc == 0
    path = str(tmp_path / "e.fasta")
    open(path, "wb").write(text)
    data, meta, counts = graph.make_edges_from_fasta(path, k, canonicalize=canonical) if k <= 13 else (None, None, None)
    recs = list(parse.parse_sequence_file(path))
    lens = np.array([len(s) for _, s in recs], dtype=np.uint64)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    st, _, got = _native.edge_rows("".join(s for _, s in recs).encode(), off, k, canonical=canonical)
    assert st == 0 and np.array_equal(got, rows)
    if data is not None:
        names = [r[0] for r in recs]
        assert data == [(names[int(r[0])], int(r[1]), int(r[2]), int(r[3]), int(r[4])) for r in rows]
        assert meta["total_kmers"] == int((lens - k + 1).sum())


@pytest.mark.parametrize("bad,exc", [("ACGTNACGTACGT", ValueError), ("ACGTRACGTACGT", KeyError), ("ACGTaACGTACGT", NameError), ("ACG", AttributeError)])
def test_graph_raises_like_reference(bad, exc, tmp_path):
    path = str(tmp_path / "b.fa")
    open(path, "w").write(">a\nACGTACGTACGTAC\n>b\n%s\n>c\nACGTTTTTTTTTGA\n" % bad)
    with pytest.raises(exc):
        graph.make_edges_from_fasta(path, 5)
    rc, _ = oracle.edges(open(path, "rb").read(), 5)
    assert oracle.ERR_NAMES[rc] == exc.__name__


@pytest.mark.parametrize("k", [1, 3, 4, 7, 11])
@pytest.mark.parametrize("canonical", [True, False])
def test_edge_weights_equal_counter_of_rows(k, canonical):
    rng = np.random.default_rng(90 + k)
    # low-complexity letters make palindromes, homopolymers and colliding canonical pairs frequent
    text = _random_fasta(rng, 300, k, 300, 60, alphabet="ACGT", p=[0.4, 0.1, 0.1, 0.4])
    rc, rows = oracle.edges(text, k, canonical=canonical, cap=1 << 18)
    assert rc == 0
    want = collections.Counter((int(r[2]), int(r[4])) for r in rows)
    n1, n2, w = graph.edge_weights_bytes(text, k, canonicalize=canonical)
    got = {(int(a), int(b)): int(c) for a, b, c in zip(n1, n2, w)}
    assert len(got) == len(n1), "a pair was emitted twice"
    assert got == dict(want)


def test_edge_weights_fastq_and_n_raise():
    rng = np.random.default_rng(2)
    text = _random_fastq(rng, 400, 30, 200)
    rc, rows = oracle.edges(text, 6, fastq=True)
    want = collections.Counter((int(r[2]), int(r[4])) for r in rows)
    n1, n2, w = graph.edge_weights_bytes(text, 6, fmt="fastq")
    assert {(int(a), int(b)): int(c) for a, b, c in zip(n1, n2, w)} == dict(want)
    with pytest.raises(ValueError):
        graph.edge_weights_bytes(b">a\nACGTACGNACGTACGT\n", 4)
    with pytest.raises(ValueError):
        graph.edge_weights_bytes(b">a\nACGTACGTACGTACGT\n>b\nACGN\n", 4)      # N in a record without any edge


def test_minimizers_compat_golden(golden):
    for key, exp in golden["minimizers"].items():
        seq, k, w = key.split(",")
        rows = minimizer.select_lexicographical_minimizers(seq, "s", int(k), int(w))
        assert [[i, kid, int(f)] for _, i, kid, f in rows] == [r[1:] for r in exp]
    with pytest.raises(ValueError, match="WHOOPS"):
        minimizer.select_lexicographical_minimizers("ACGT" * 30, "s", 12, 10)      # always, for k=12 on short reads
    with pytest.raises(ValueError):
        minimizer.select_lexicographical_minimizers("ACG", "s", 5, 10)


@pytest.mark.parametrize("k,w", [(3, 1), (5, 2), (7, 5), (12, 10), (12, 20), (15, 32), (9, 33), (11, 50), (31, 7)])
def test_true_minimizers_vs_oracle(k, w):
    rng = np.random.default_rng(k * 100 + w)
    for L in (k + w - 2, k + w - 1, k + w + 5, 150, 1000, 40007):
        if L < 1:
            continue
        seq = "".join(np.array(list("ACGT"))[rng.choice(4, size=L, p=[0.45, 0.05, 0.05, 0.45])])
        rc, pos, ids = oracle.minimizers_true(seq, k, w)
        assert rc == 0
        got = minimizer.windowed_minimizers(seq, k, w)
        assert got == [(int(p), int(i)) for p, i in zip(pos, ids)], (k, w, L)


def test_shred_matches_reference_golden(golden):
    Rec = collections.namedtuple("Rec", ["id", "seq"])
    for key, exp in golden["scalars"]["shred"].items():
        s, k, rn, canon = key.split(",")
        ids, sids, pos = kmer.shred(Rec("x", s), int(k), replace_with_none=bool(int(rn)), canonicalize=bool(int(canon)))
        assert pos == exp["pos"]
        # the reference expands N through a Python set: order inside one position is undefined there
        grp = lambda I, P: sorted(zip(P, I))
        assert grp(ids, pos) == grp(exp["ids"], exp["pos"])
        assert sids == ["x"] * len(ids)


@pytest.mark.parametrize("skip_n", [True, False])
def test_multik_shared_stream_hard_layouts(skip_n):
    """The shared-encode sweep (kdb_attach) on inputs where windows cross lines, chunks, tiles and N runs."""
    rng = np.random.default_rng(77)
    cases = [(_random_fasta(rng, 120, 40, 30000, 61, alphabet="ACGTN", p=[0.2475, 0.2475, 0.2475, 0.2475, 0.01]), False),
             (_random_fasta(rng, 300, 16, 600, 0), False),
             (_random_fastq(rng, 2000, 16, 300), True)]
    for text, fq in cases:
        profs = parse.profile_multik(text, 2, 12 if not skip_n else 14, fmt="fastq" if fq else "fasta", replace_with_none=skip_n)
        for k, pr in profs.items():
            rc, ocounts, ost = oracle.profile(text, k, fastq=fq, skip_n=skip_n, threads=8)
            assert rc == 0 and np.array_equal(pr.counts, ocounts), (k, fq)
            assert (pr.total_kmers, pr.unique_kmers, pr.total_reads, pr.min_read_length, pr.max_read_length) == \
                   (ost["total_kmers"], ost["unique_kmers"], ost["total_reads"], ost["min_len"], ost["max_len"])


def test_multik_sweep_raises_at_the_first_failing_k():
    text = b">a\nACGTACGTACGTACGT\n>b\nACGTACG\n>c\nACGTTTGACCAGTACA\n"          # record b has 7 letters
    it = parse.iter_profile_multik(text, 5, 9)
    got = []
    with pytest.raises(AttributeError):                     # kmer.py:461-463 at k = 8
        for k, pr in it:
            got.append(k)
            rc, oc, _ = oracle.profile(text, k)
            assert rc == 0 and np.array_equal(pr.counts, oc)
    assert got == [5, 6, 7]
    with pytest.raises(KeyError):
        parse.profile_multik(b">a\nACGTRCGTACGTACGT\n", 4, 6)
    with pytest.raises(NameError):
        parse.profile_multik(b">a\nACGTACGTACGTACGT\n>b\nACGTaCGTACGTACGT\n", 4, 6)


def test_multik_sweep_vs_oracle():
    host = _native.synth_host(20000)
    profs = parse.profile_multik(host, 4, 14)
    for k, pr in profs.items():
        rc, ocounts, ost = oracle.profile(host, k, threads=8)
        assert rc == 0 and np.array_equal(pr.counts, ocounts), k
        assert pr.total_kmers == ost["total_kmers"] and pr.unique_kmers == ost["unique_kmers"]


@pytest.mark.parametrize("k,w", [(5, 4), (12, 10), (15, 32), (9, 40)])
def test_batched_minimizers_equal_per_record_oracle(k, w, tmp_path):
    rng = np.random.default_rng(k + w)
    recs = []
    for i in range(700):
        L = int(rng.choice([0, 1, k, k + w - 2, k + w - 1, k + w, 150, 151, 400, 5000], p=[.02, .02, .04, .05, .05, .05, .5, .17, .08, .02]))
        recs.append("".join(np.array(list("ACGT"))[rng.choice(4, size=L, p=[0.4, 0.1, 0.1, 0.4])]))
    want = []
    for r, s in enumerate(recs):
        rc, pos, ids = oracle.minimizers_true(s, k, w)
        assert rc == 0
        want += [(r, int(p), int(i)) for p, i in zip(pos, ids)]
    lens = np.array([len(s) for s in recs], dtype=np.uint64)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    st, rows = _native.minimizers_batch("".join(recs).encode(), off, k, w)
    assert st == 0 and [tuple(int(x) for x in row) for row in rows] == want
    st, n = _native.minimizers_batch("".join(recs).encode(), off, k, w, count_only=True)
    assert st == 0 and n == len(want)
    path = str(tmp_path / "m.fa")
    open(path, "w").write("".join(">s%d\n%s\n" % (i, s) for i, s in enumerate(recs) if s))
    got = minimizer.windowed_minimizers_file(path, k, w)
    names = ["s%d" % i for i, s in enumerate(recs) if s]
    nonempty = [i for i, s in enumerate(recs) if s]
    remap = {r: j for j, r in enumerate(nonempty)}
    assert got == [(names[remap[r]], p, i) for r, p, i in want]


def test_batched_minimizers_at_read_scale():
    """2 M synthetic 150 bp reads in one call (count only): every read yields at least one minimizer."""
    n = 2_000_000
    text = _native.synth_host(n).reshape(n, 162)
    letters = np.ascontiguousarray(text[:, 11:161]).reshape(-1)
    off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(150))
    st, rows = _native.minimizers_batch(letters, off, 12, 10, count_only=True)
    assert st == 0 and n <= rows <= n * 130
    s = letters[:150 * 50].tobytes().decode()
    want = sum(len(oracle.minimizers_true(s[150 * r:150 * r + 150], 12, 10)[1]) for r in range(50))
    st, rows50 = _native.minimizers_batch(letters[:150 * 50], off[:51], 12, 10)
    assert len(rows50) == want
