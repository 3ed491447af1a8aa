"""Host I/O helpers of kmerdb_b200.parse (no GPU): gzip / BGZF input decode and the raw-file checksums."""
import gzip
import hashlib
import os

from kmerdb_b200 import kdbfile, parse


def test_bgzf_input_is_inflated_blockwise_to_the_same_bytes(tmp_path):
    raw = (b">r\n" + b"ACGTTGCA" * 40000 + b"\n") * 5 + os.urandom(1000)
    bgzf = b"".join(kdbfile.bgzf_block(raw[o:o + 65280]) for o in range(0, len(raw), 65280)) + kdbfile._BGZF_EOF
    assert parse._bgzf_blocks(bgzf) is not None
    assert parse._gunzip(bgzf) == raw == gzip.decompress(bgzf)
    plain = gzip.compress(raw)
    assert parse._bgzf_blocks(plain) is None and parse._gunzip(plain) == raw
    for name, blob in (("a.fa.gz", bgzf), ("b.fa.gz", plain), ("c.fa", raw)):
        p = str(tmp_path / name)
        open(p, "wb").write(blob)
        assert parse._read_text_bytes(p) == raw
    # a text file that merely starts with the gzip magic is not gzip (kmerdb/util.py:80-88 decides by trying to read)
    p = str(tmp_path / "d.fa")
    open(p, "wb").write(b"\x1f\x8b not really gzip")
    assert parse._read_text_bytes(p) == b"\x1f\x8b not really gzip"


def test_async_checksums_match_util_checksum(tmp_path):
    raw = os.urandom(3 << 20)
    p = str(tmp_path / "x.bin")
    open(p, "wb").write(raw)
    assert parse._checksums_async(raw)() == parse.checksum(p) == (hashlib.md5(raw).hexdigest(), hashlib.sha256(raw).hexdigest())


def test_read_and_hash_streams_the_file(tmp_path):
    """parse._read_and_hash: the bytes and the two digests of kmerdb/util.py:35-50, hashed while the file is read."""
    import hashlib
    import numpy as np
    from kmerdb_b200 import parse
    rng = np.random.default_rng(0)
    for size, chunk in ((0, 1 << 20), (1, 1 << 20), (1000003, 4096), (3 << 20, 1 << 20)):
        p = tmp_path / ("f%d.bin" % size)
        data = rng.integers(0, 256, size=size, dtype=np.uint8).tobytes()
        p.write_bytes(data)
        raw, sums = parse._read_and_hash(str(p), chunk=chunk)
        assert bytes(raw) == data
        assert sums() == (hashlib.md5(data).hexdigest(), hashlib.sha256(data).hexdigest())
        assert sums() == parse.checksum(str(p))
