"""f2: BGZF input inflated on the GPU.  The reference opens inputs through gzip.open (kmerdb/parse.py:63-68);
here the compressed members cross PCIe and are inflated on the device (one thread per member decodes, one warp resolves).  Oracle: zlib on the host (the same library gzip.open
uses).  Inputs: the reference's bundled genomes re-blocked as BGZF, synthetic reads, and blocks that force every DEFLATE
block type (stored, fixed, dynamic), long matches, overlapping matches and codes longer than the first-level table."""
import gzip
import hashlib
import json
import os
import struct
import zlib

import numpy as np
import pytest

from kmerdb_b200 import _native, kdbfile, parse

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(HERE, "golden", "data")
GOLD = json.load(open(os.path.join(HERE, "golden", "golden_bundled.json")))


def _bgzf(data, block=65280, level=6):
    return b"".join(kdbfile.bgzf_block(data[o:o + block], level) for o in range(0, len(data), block)) + kdbfile._BGZF_EOF


def _member(data, level, strategy=zlib.Z_DEFAULT_STRATEGY, extra=b""):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
    comp = c.compress(data) + c.flush()
    xlen = 6 + len(extra)
    bsize = len(comp) + 12 + xlen + 8
    assert bsize <= 65536
    head = b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff" + struct.pack("<H", xlen) + extra + b"BC\x02\x00" + struct.pack("<H", bsize - 1)
    return head + comp + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))


def test_every_block_type_and_match_shape():
    rng = np.random.default_rng(11)
    texts = [
        b"",                                                        # empty member
        b"A",
        bytes(rng.integers(0, 256, size=30000, dtype=np.uint8)),   # incompressible: stored blocks at level 0 / dynamic with long codes
        b"ACGT" * 15000,                                            # distance 4, length 258 matches (overlapping copies)
        b"A" * 65000,                                               # distance 1
        bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=60000)),
        ("".join(">r%d\n%s\n" % (i, "".join("ACGTN"[j % 5] for j in range(i % 90))) for i in range(900))).encode(),
        bytes(np.minimum(255, rng.geometric(0.02, size=50000)).astype(np.uint8)),   # skewed alphabet: codes of every length up to 15
    ]
    members, want = [], []
    for t in texts:
        for level, strat in ((0, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_DEFAULT_STRATEGY), (6, zlib.Z_DEFAULT_STRATEGY), (9, zlib.Z_DEFAULT_STRATEGY),
                             (6, zlib.Z_FIXED), (6, zlib.Z_HUFFMAN_ONLY), (6, zlib.Z_RLE)):
            try:
                members.append(_member(t, level, strat, extra=b"XY\x03\x00abc" if level == 9 else b""))
            except AssertionError:
                continue                                           # does not fit one BGZF member at this level
            want.append(t)
    blob = b"".join(members) + kdbfile._BGZF_EOF
    got = _native.inflate_bgzf(blob)
    assert bytes(got) == b"".join(want) == gzip.decompress(blob)


def test_members_of_several_blocks():
    """Members whose deflate stream holds several blocks (dynamic, fixed, empty stored blocks of a sync flush, stored data):
    threads of one warp then stand in different blocks of different members."""
    rng = np.random.default_rng(5)
    members, want = [], []
    for mi in range(70):
        parts = [bytes(rng.choice(np.frombuffer(b"ACGTN\n>", dtype=np.uint8), size=int(rng.integers(1, 9000)))) for _ in range(int(rng.integers(1, 6)))]
        c = zlib.compressobj(int(rng.integers(0, 10)), zlib.DEFLATED, -15)
        comp = b""
        for part in parts:
            comp += c.compress(part) + c.flush([zlib.Z_SYNC_FLUSH, zlib.Z_FULL_FLUSH, zlib.Z_NO_FLUSH][int(rng.integers(0, 3))])
        comp += c.flush()
        data = b"".join(parts)
        head = b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff" + struct.pack("<H", 6) + b"BC\x02\x00" + struct.pack("<H", len(comp) + 12 + 6 + 8 - 1)
        members.append(head + comp + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data)))
        want.append(data)
    blob = b"".join(members) + kdbfile._BGZF_EOF
    assert gzip.decompress(blob) == b"".join(want)
    assert bytes(_native.inflate_bgzf(blob)) == b"".join(want)


@pytest.mark.parametrize("name", ["Cacetobutylicum_ATCC824.fasta.gz", "Cdifficile_R3.fasta.gz", "sample.fa"])
def test_bundled_genomes_reblocked_as_bgzf(name, tmp_path):
    path = os.path.join(DATA, name)
    text = gzip.open(path, "rb").read() if name.endswith(".gz") else open(path, "rb").read()
    blob = _bgzf(text)
    got = _native.inflate_bgzf(blob)
    assert hashlib.sha256(bytes(got)).hexdigest() == hashlib.sha256(text).hexdigest()
    # and through the product path: a BGZF .fasta.gz gives the golden counts of the reference (inflate on the device)
    p = str(tmp_path / (name.split(".")[0] + ".bgzf.fasta.gz"))
    open(p, "wb").write(blob)
    for c in GOLD["cases"]:
        if c["file"] != name or c["k"] != 8:
            continue
        prof = parse.profile_bgzf(blob, 8, "fasta", c["replace_with_none"], c["canonicalize"])
        assert prof is not None and prof.timing["inflate_ms"] > 0
        counts, meta, nullomers = parse.parsefile(p, 8, replace_with_none=c["replace_with_none"], canonicalize=c["canonicalize"])
        for vec in (prof.counts, counts):
            assert hashlib.sha256(vec.astype("<u8").tobytes()).hexdigest() == c["counts_sha256"]
        assert {a: b for a, b in meta.items() if a in ("total_reads", "total_kmers", "unique_kmers", "min_read_length", "max_read_length")} == \
               {a: b for a, b in c["meta"].items() if a in ("total_reads", "total_kmers", "unique_kmers", "min_read_length", "max_read_length")}
        assert meta["md5"] == hashlib.md5(blob).hexdigest()        # checksums are of the file as it lies on disk (kmerdb/util.py:35-50)


def test_corrupt_member_is_refused_and_the_host_path_raises():
    text = bytes(_native.synth_host(3000))
    blob = bytearray(_bgzf(text))
    blocks = parse._bgzf_blocks(bytes(blob))
    blob[blocks[1][2] + 100] ^= 0x10                               # one flipped bit in the second member's payload
    with pytest.raises(ValueError):
        _native.inflate_bgzf(bytes(blob))
    assert parse.profile_bgzf(bytes(blob), 8) is None              # the device refuses it ...
    with pytest.raises(Exception):                                  # ... and the host path raises like gzip.open
        parse._gunzip(bytes(blob))
    blob2 = bytearray(_bgzf(text))
    blob2[blocks[0][0] + blocks[0][1] - 6] ^= 0xFF                  # CRC-32 field of the first trailer
    with pytest.raises(ValueError):
        _native.inflate_bgzf(bytes(blob2))


def test_fastq_reads_bgzf_equal_uncompressed():
    text = bytes(_native.synth_host(40000, fastq=True))
    one = parse.profile_bytes(text, 12, "fastq", True, True)
    two = parse.profile_bgzf(_bgzf(text, level=1), 12, "fastq", True, True)
    assert two is not None and np.array_equal(one.counts, two.counts) and one.total_reads == two.total_reads == 40000
    assert one.mononucleotides == two.mononucleotides


@pytest.mark.parametrize("fmt,block,env", [("fasta", 65280, {"KDB_INFLATE_PIECE": "64"}), ("fastq", 65280, {"KDB_INFLATE_PIECE": "64"}),
                                           ("fasta", 60000, {"KDB_INFLATE_PIECE": "64"}), ("fasta", 65280, {"KDB_INFLATE_SERIAL": "1"}),
                                           ("fasta", 65280, {"KDB_INFLATE_SERIAL": "1", "KDB_INFLATE_PIECE": "100"}),
                                           ("fastq", 65280, {"KDB_INFLATE_SERIAL": "1", "KDB_INFLATE_PIECE": "37"})])
def test_pipelined_pieces_equal_the_uncompressed_feed(fmt, block, env, monkeypatch):
    """kdb_feed_bgzf_host cuts the file into pieces on tile borders (copy, inflate and count overlap piece by piece): many small
    pieces, records and lines crossing every cut; members that cannot be cut on tile borders (60 000-byte blocks) and the serial
    path give the same profile."""
    for a, b in env.items():
        monkeypatch.setenv(a, b)
    n = 120000 if fmt == "fastq" else 200000
    text = bytes(_native.synth_host(n, fastq=(fmt == "fastq"), seed=77))
    blob = _bgzf(text, block=block, level=1)
    one = parse.profile_bytes(text, 11, fmt, True, True)
    two = parse.profile_bgzf(blob, 11, fmt, True, True)
    assert two is not None and two.total_reads == one.total_reads == n
    assert np.array_equal(one.counts, two.counts)
    assert (one.min_read_length, one.max_read_length, one.mononucleotides) == (two.min_read_length, two.max_read_length, two.mononucleotides)
    bad = bytearray(blob)
    blocks = parse._bgzf_blocks(bytes(blob))
    bad[blocks[len(blocks) // 2][2] + 50] ^= 0x04                   # a corrupt member in a middle piece: refused, nothing is kept
    assert parse.profile_bgzf(bytes(bad), 11, fmt, True, True) is None


@pytest.mark.parametrize("env", [{}, {"KDB_INFLATE_PIPELINE": "1", "KDB_INFLATE_PIECE": "64"}, {"KDB_INFLATE_LATE_COPY": "1"}, {"KDB_INFLATE_PIECE": "50"}])
def test_pinned_input_copy_starts_before_the_index(env, monkeypatch):
    """From page-locked memory the compressed bytes are on their way before the member headers are walked; the result is the
    one of the uncompressed feed, also when the call is repeated on the same context (buffers and events are reused) and for a
    file that is not BGZF at all."""
    for a, b in env.items():
        monkeypatch.setenv(a, b)
    text = bytes(_native.synth_host(150000, seed=78))
    blob = np.frombuffer(_bgzf(text, level=6), dtype=np.uint8)
    pin = _native.PinnedBuffer(blob.size)
    try:
        pin.array[:] = blob
        want = parse.profile_bytes(text, 10, "fasta", True, True)
        with _native.Profiler(10) as p:
            for _ in range(3):
                p.reset()
                assert p.feed_bgzf(pin.array) == 0
                assert p.finish() == 0
                assert np.array_equal(p.counts(), want.counts)
                assert p.stats()["total_reads"] == 150000
            p.reset()
            pin.array[:4] = np.frombuffer(b">abc", dtype=np.uint8)          # not gzip
            assert p.feed_bgzf(pin.array) == _native.KDB_E_FORMAT
            pin.array[:4] = blob[:4]
            pin.array[blob.size // 2] ^= 0x40                               # a corrupt member
            p.reset()
            assert p.feed_bgzf(pin.array) == _native.KDB_E_FORMAT
    finally:
        pin.free()
