"""Parity tests proper: the CUDA path (through the C-ABI, kmerdb_b200._native / kmerdb_b200.parse)
against the golden vectors of the unmodified reference and against the CPU oracle on the same
seeded inputs.  Bit-exact: integer work only."""
import gzip
import hashlib
import os

import numpy as np
import pytest

import oracle
from kmerdb_b200 import _native, parse
from test_oracle import _golden_cases, _case_id

pytestmark = pytest.mark.gpu

EXC = {"NameError": NameError, "KeyError": KeyError, "AttributeError": AttributeError, "ValueError": ValueError}


@pytest.mark.parametrize("c", _golden_cases(), ids=_case_id)
def test_parsefile_matches_reference_golden(c, tmp_path):
    path = str(tmp_path / ("input" + c["suffix"] + (".gz" if c["gz"] else "")))
    with (gzip.open(path, "wb") if c["gz"] else open(path, "wb")) as fh:
        fh.write(c["text"].encode())
    e = c["expect"]
    if "raises" in e:
        with pytest.raises(EXC[e["raises"]]) as ei:
            parse.parsefile(path, c["k"], replace_with_none=c["replace_with_none"], canonicalize=c["canonicalize"])
        assert type(ei.value).__name__ == e["raises"] or e["raises"] == "ValueError"   # UnicodeDecodeError etc. are ValueErrors
        return
    counts, meta, nullomers = parse.parsefile(path, c["k"], replace_with_none=c["replace_with_none"], canonicalize=c["canonicalize"])
    assert counts.dtype == np.uint64 and counts.shape == (4 ** c["k"],)
    assert meta["filename"] == path
    assert meta["md5"] == hashlib.md5(open(path, "rb").read()).hexdigest()
    assert meta["sha256"] == hashlib.sha256(open(path, "rb").read()).hexdigest()
    got = {a: b for a, b in meta.items() if a not in ("filename", "md5", "sha256")}
    assert got == e["meta"]
    assert hashlib.sha256(counts.astype("<u8").tobytes()).hexdigest() == e["counts_sha256"]
    assert nullomers.dtype == np.uint64 and len(nullomers) == e["n_nullomers"]
    assert hashlib.sha256(nullomers.astype("<u8").tobytes()).hexdigest() == e["nullomers_sha256"]


def _cmp_oracle(data, k, fastq=False, skip_n=True, canonical=True, threads=8):
    rc, ocounts, ost = oracle.profile(data, k, fastq=fastq, skip_n=skip_n, canonical=canonical, threads=threads)
    with _native.Profiler(k, canonical=canonical, skip_n=skip_n, fastq=fastq) as p:
        p.feed(data)
        status = p.finish()
        st = p.stats()
        if rc != 0:
            assert status == {1: 1, 2: 2, 3: 3, 4: 7, 5: 5}[rc]
            return
        assert status == 0
        counts = p.counts()
        assert np.array_equal(counts, ocounts)
        for a, b in (("total_reads", "total_reads"), ("total_kmers", "total_kmers"), ("unique_kmers", "unique_kmers"),
                     ("sum_len", "sum_len"), ("min_len", "min_len"), ("max_len", "max_len"), ("mono", "mono")):
            assert st[a] == ost[b], a
        nul = p.nullomers()
        assert np.array_equal(nul, np.nonzero(ocounts == 0)[0].astype(np.uint64))
        return counts


def _random_fasta(rng, n_rec, min_len, max_len, width, alphabet="ACGT", p=None, crlf=False):
    eol = "\r\n" if crlf else "\n"
    letters = np.array(list(alphabet))
    out = []
    for i in range(n_rec):
        L = int(rng.integers(min_len, max_len + 1))
        s = "".join(letters[rng.choice(len(letters), size=L, p=p)])
        out.append(">rec%d some text > with symbols" % i + eol)
        if width:
            for j in range(0, L, width):
                out.append(s[j:j + width] + eol)
        else:
            out.append(s + eol)
    return "".join(out).encode()


@pytest.mark.parametrize("k", [1, 2, 3, 5, 7, 8, 9, 11, 12, 13])
@pytest.mark.parametrize("canonical", [True, False])
def test_random_fasta_all_k(k, canonical):
    rng = np.random.default_rng(100 + k)
    data = _random_fasta(rng, 400, 20, 900, 70)
    _cmp_oracle(data, k, canonical=canonical)


@pytest.mark.parametrize("width", [0, 1, 7, 60, 61, 64, 1000])
def test_line_widths_cross_chunks_and_tiles(width):
    rng = np.random.default_rng(7 + width)
    data = _random_fasta(rng, 60, 1000, 40000, width, crlf=(width == 61))
    _cmp_oracle(data, 9)


@pytest.mark.parametrize("skip_n", [True, False])
@pytest.mark.parametrize("k", [3, 6, 8, 10])
def test_n_windows(k, skip_n):
    rng = np.random.default_rng(5)
    data = _random_fasta(rng, 300, 30, 400, 80, alphabet="ACGTN", p=[0.2475, 0.2475, 0.2475, 0.2475, 0.01])
    _cmp_oracle(data, k, skip_n=skip_n)
    _cmp_oracle(data, k, skip_n=skip_n, canonical=False)


def test_long_n_runs_expand():
    data = b">a\nACGTNNNNNNACGTACGTNNACGTTTGACNACGT\n>b\nNNNNNNNNNNNN\n>c\nACGTACGTACGTACGT\n"
    for k in (4, 6):
        _cmp_oracle(data, k, skip_n=False)
        _cmp_oracle(data, k, skip_n=False, canonical=False)


def _random_fastq(rng, n_rec, min_len, max_len, crlf=False):
    eol = "\r\n" if crlf else "\n"
    out = []
    for i in range(n_rec):
        L = int(rng.integers(min_len, max_len + 1))
        s = "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=L)])
        q = "".join("@+I5#>"[(i + j) % 6] for j in range(L))
        out.append("@read%d/1 x" % i + eol + s + eol + "+" + eol + q + eol)
    return "".join(out).encode()


@pytest.mark.parametrize("k", [4, 8, 12])
def test_random_fastq(k):
    rng = np.random.default_rng(11)
    _cmp_oracle(_random_fastq(rng, 3000, 30, 260), k, fastq=True)
    _cmp_oracle(_random_fastq(rng, 500, 30, 260, crlf=True), k, fastq=True, canonical=False)


def test_error_record_order_matches_oracle():
    rng = np.random.default_rng(3)
    base = _random_fasta(rng, 200, 50, 300, 60).decode()
    recs = base.split(">")[1:]
    for bad_at, bad in ((150, "ACGTacgtACGTACGTAAAA"), (20, "ACGTRCGTACGTACGTAAAA"), (77, "ACG")):
        r = list(recs)
        r[bad_at] = "bad\n" + bad + "\n"
        r[180] = "bad2\nAC\n"
        _cmp_oracle((">" + ">".join(r)).encode(), 8)


def test_synthetic_generator_host_equals_device():
    n = 5000
    host = _native.synth_host(n)
    buf = _native.DeviceBuffer(host.size)
    _native.synth_device(buf, n)
    assert np.array_equal(buf.download(), host)
    hq = _native.synth_host(n, fastq=True)
    bq = _native.DeviceBuffer(hq.size)
    _native.synth_device(bq, n, fastq=True)
    assert np.array_equal(bq.download(), hq)
    assert host[0] == ord(">") and hq[0] == ord("@")


def test_synthetic_reads_k12_vs_oracle_and_device_feed():
    n = 200000
    host = _native.synth_host(n)
    _cmp_oracle(host, 12)
    buf = _native.DeviceBuffer(host.size)
    _native.synth_device(buf, n)
    rc, ocounts, ost = oracle.profile(host, 12, threads=8)
    with _native.Profiler(12) as p:
        p.feed_device(buf.ptr, host.size)
        assert p.finish() == 0
        assert np.array_equal(p.counts(), ocounts)
        assert p.stats()["total_kmers"] == n * 139 == ost["total_kmers"]


def test_fragment_feeds_equal_single_feed():
    n = 30000
    host = _native.synth_host(n)
    rb = _native.synth_record_bytes()
    with _native.Profiler(10) as p:
        p.feed(host)
        assert p.finish() == 0
        whole, wst = p.counts(), p.stats()
    with _native.Profiler(10) as p:
        cuts = [0, 1 * rb, 7777 * rb, 20000 * rb, n * rb]
        for a, b in zip(cuts, cuts[1:]):
            p.feed(host[a:b])
        assert p.finish() == 0
        assert np.array_equal(p.counts(), whole)
        assert p.stats() == wst


def test_kmer_to_id_batch(golden):
    two = "".join(a + b for a in "ACGT" for b in "ACGT").encode()
    assert _native.kmer_to_id_batch(two, 2, canonical=False).tolist() == golden["scalars"]["kmer_to_id_2mers_noncanonical"]
    assert _native.kmer_to_id_batch(two, 2, canonical=True).tolist() == golden["scalars"]["kmer_to_id_2mers_canonical"]
    ids = _native.kmer_to_id_batch(b"ATCNATCACGRACG", 7)
    assert ids[0] == np.uint64(2 ** 64 - 1) and ids[1] == np.uint64(2 ** 64 - 2)


def test_full_size_100M_reads_properties():
    """BASELINE.json configs[2] size on one GPU (100 M x 150 bp, 16.2 GB of FASTA text, 31 passes): the
    size-independent properties of the result, identity of two different feed chunkings, and gate T2 — the full
    vector's sha256 against the C oracle over the same text (about 2 minutes of host time)."""
    n, k = 100_000_000, 12
    rb = _native.synth_record_bytes()
    buf = _native.DeviceBuffer(n * rb + 16)
    _native.synth_device(buf, n)
    with _native.Profiler(k) as p:
        p.feed_device(buf.ptr, n * rb)
        assert p.finish() == 0
        st, counts = p.stats(), p.counts()
        assert st["total_reads"] == n and st["total_kmers"] == n * (150 - k + 1) == int(counts.sum(dtype=np.uint64))
        assert st["sum_len"] == n * 150 == sum(st["mono"]) and st["min_len"] == st["max_len"] == 150
        assert st["unique_kmers"] == int(np.count_nonzero(counts)) == (4 ** k + 4 ** (k // 2)) // 2     # saturated canonical space
        # canonical ids only: a bin above its own reverse complement stays empty
        ids = np.flatnonzero(counts).astype(np.uint64)
        assert ids.size == st["unique_kmers"]
        # checksum of checksums: two feeds cut at a record boundary give the same vector
        p.reset()
        cut = 37_000_008 * rb                          # a record boundary whose byte offset is a multiple of 16 (device feeds are 16-byte aligned)
        p.feed_device(buf.ptr, cut)
        p.feed_device(buf.ptr + cut, n * rb - cut)
        assert p.finish() == 0
        assert np.array_equal(p.counts(), counts) and p.stats() == st
    # Parity gate T2 at full R: sha256 of the whole little-endian uint64 vector + the scalars against the C
    # oracle run over the SAME 16.2 GB of text (copied back from the device in 8 pieces; the oracle's vectors add up)
    threads = max(4, len(os.sched_getaffinity(0)))
    osum = np.zeros(4 ** k, dtype=np.uint64)
    ost = {"total_reads": 0, "total_kmers": 0, "sum_len": 0, "mono": [0, 0, 0, 0]}
    piece = n // 8
    host = np.empty(piece * rb, dtype=np.uint8)
    for i in range(8):
        assert _native.lib().kdb_memcpy_d2h(0, host.ctypes.data, buf.ptr + i * piece * rb, piece * rb) == 0
        rc, oc, o = oracle.profile(host, k, threads=threads)
        assert rc == 0
        osum += oc
        for f in ("total_reads", "total_kmers", "sum_len"):
            ost[f] += o[f]
        ost["mono"] = [a + b for a, b in zip(ost["mono"], o["mono"])]
        assert o["min_len"] == o["max_len"] == 150
    assert hashlib.sha256(counts.tobytes()).hexdigest() == hashlib.sha256(osum.tobytes()).hexdigest()
    for f in ("total_reads", "total_kmers", "sum_len", "mono"):
        assert st[f] == ost[f], f
    assert st["unique_kmers"] == int(np.count_nonzero(osum))


def test_configs1_full_vector_vs_oracle():
    """Parity gate T2 at BASELINE.json configs[1] size: 10 M x 150 bp, k=12 canonical — the whole 4^12 vector and the
    scalars against the CPU oracle (16 host threads, ~10 s)."""
    n, k = 10_000_000, 12
    host = _native.synth_host(n)
    rc, ocounts, ost = oracle.profile(host, k, threads=max(4, len(os.sched_getaffinity(0))))
    assert rc == 0
    with _native.Profiler(k) as p:
        p.feed(host)
        assert p.finish() == 0
        st, counts = p.stats(), p.counts()
    assert hashlib.sha256(counts.tobytes()).hexdigest() == hashlib.sha256(ocounts.tobytes()).hexdigest()
    for a in ("total_reads", "total_kmers", "unique_kmers", "sum_len", "min_len", "max_len", "mono"):
        assert st[a] == ost[a], a


@pytest.mark.gpu
@pytest.mark.parametrize("k,canonical", [(8, True), (8, False), (13, True), (13, False), (14, True)])
def test_packed_16bit_counters_wrap_exactly(k, canonical):
    """k = 8 counts its 65536 bins, k = 13, 14 the 65536 bins of a bucket, as 16-bit counters, two per shared word (add16).  Two bins far above
    65535 in the two halves of ONE word, fed concurrently, exercise every carry / borrow path; full vector equality with the
    oracle.  (The input must be large: a bucket stream takes 3x the average bucket load of the pass before the rest of a
    crowded bucket is counted by direct REDs.)"""
    rng = np.random.default_rng(k)
    n = {8: 30_000_000, 13: 9_000_000, 14: 16_000_000}[k]      # k = 8: one shared table per CTA, 148 CTAs share the reads
    pat = np.frombuffer(("A" * k + "A" * (k - 1) + "C").encode(), dtype=np.uint8).reshape(2, k)      # forward ids 0 and 1
    which = rng.choice(3, size=n, p=[0.5, 0.4, 0.1])
    rec = np.empty((n, k + 3), dtype=np.uint8)
    rec[:, 0] = ord(">"); rec[:, 1] = 10; rec[:, k + 2] = 10
    rec[:, 2:k + 2] = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=(n, k), dtype=np.uint8)]
    for j in (0, 1):
        rec[which == j, 2:k + 2] = pat[j]
    counts = _cmp_oracle(rec.reshape(-1).tobytes(), k, canonical=canonical)
    assert counts[0] > 4_000_000 and counts[1] > 3_000_000


@pytest.mark.parametrize("fmt", ["fasta", "fastq"])
@pytest.mark.parametrize("gz", [False, True])
def test_streamed_parsefile_equals_whole_file(fmt, gz, tmp_path, monkeypatch):
    """f2 streaming: the file is read / inflated in small pieces and fed fragment by fragment (record-aligned cuts);
    counts, statistics and checksums equal the whole-text profile.  Records cross many piece borders (80-column genome
    lines, 150 bp reads)."""
    rng = np.random.default_rng(21)
    if fmt == "fasta":
        recs = []
        for i in range(40):
            L = int(rng.integers(30, 60000))
            seq = "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=L)])
            recs.append(">chr%d\n" % i + "\n".join(seq[j:j + 80] for j in range(0, L, 80)) + "\n")
        text = "".join(recs).encode()
    else:
        text = bytes(_native.synth_host(9000, fastq=True))
    path = str(tmp_path / ("x." + ("fa" if fmt == "fasta" else "fq") + (".gz" if gz else "")))
    with (gzip.open(path, "wb") if gz else open(path, "wb")) as fh:
        fh.write(text)
    monkeypatch.setattr(parse, "STREAM_PIECE", 1 << 16)
    counts, meta, nullomers = parse.parsefile(path, 9, replace_with_none=True, canonicalize=True)
    whole = parse.profile_bytes(text, 9, fmt, True, True)
    assert np.array_equal(counts, whole.counts) and np.array_equal(nullomers, whole.nullomer_array)
    blob = open(path, "rb").read()
    assert meta == whole.file_metadata(path, hashlib.md5(blob).hexdigest(), hashlib.sha256(blob).hexdigest())
    rc, ocounts, _ost = oracle.profile(text, 9, fastq=(fmt == "fastq"))
    assert rc == 0 and np.array_equal(counts, ocounts)
