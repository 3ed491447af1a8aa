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


def _bgzf_with_extra_subfield(data):
    """A BGZF member whose extra field holds another subfield BEFORE 'BC' (legal RFC 1952; XLEN = 12, not 6)."""
    import struct
    import zlib
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    comp = c.compress(data) + c.flush()
    bsize = len(comp) + 12 + 12 + 8
    head = b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff" + struct.pack("<H", 12) + b"XY\x02\x00\xaa\xbb" + b"BC\x02\x00" + struct.pack("<H", bsize - 1)
    return head + comp + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))


def test_bgzf_walker_reads_xlen_and_checks_the_trailer():
    """members with extra subfields must inflate from the right offset; a corrupt member must raise like
    gzip.open does (BadGzipFile), not be counted silently."""
    import ctypes
    import numpy as np
    import pytest
    from kmerdb_b200 import _native
    raw = b">r\n" + b"ACGTTGCAAC" * 20000 + b"\n"
    pieces = [raw[o:o + 60000] for o in range(0, len(raw), 60000)]
    blob = b"".join(_bgzf_with_extra_subfield(x) for x in pieces) * 6 + kdbfile._BGZF_EOF          # > 1 MiB of members
    blocks = parse._bgzf_blocks(blob)
    assert blocks is not None and all(q - off == 24 for off, _size, q in blocks[:-1])
    assert parse._gunzip(blob) == raw * 6
    # the C-ABI walker (kdb_bgzf_index, host code) agrees
    assert _native.bgzf_text_size(blob) == 6 * len(raw)
    assert _native.bgzf_text_size(gzip.compress(raw)) is None
    a = np.frombuffer(blob, dtype=np.uint8)
    mem = np.zeros((len(blocks), 4), dtype=np.uint64)                 # 32 bytes per member
    nm, text = ctypes.c_size_t(0), ctypes.c_uint64(0)
    assert _native.lib().kdb_bgzf_index(a.ctypes.data, a.size, mem.ctypes.data, len(blocks), ctypes.byref(nm), ctypes.byref(text)) == 0
    assert nm.value == len(blocks) and [int(x) for x in mem[:, 0]] == [q for _o, _s, q in blocks]
    # flip one payload byte of the second member: CRC / stream check must fire
    bad = bytearray(blob)
    bad[blocks[1][2] + 40] ^= 0x55
    with pytest.raises((gzip.BadGzipFile, OSError, EOFError, Exception)):
        parse._gunzip(bytes(bad))


def test_streamed_text_pieces_and_record_aligned_fragments(tmp_path):
    """parse._iter_text_pieces / _iter_fragments (bounded-memory input): same bytes as gzip.open for plain, gzip and
    multi-member gzip files; fragments join back to the text and start at record borders; raw bytes are hashed on the way."""
    import numpy as np
    import pytest
    rng = np.random.default_rng(7)
    recs = []
    for i in range(300):
        L = int(rng.integers(1, 400))
        seq = "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=L)])
        recs.append(">r%d with > inside\n" % i + "\n".join(seq[j:j + 70] for j in range(0, L, 70)) + "\n")
    fasta = "".join(recs).encode()
    fastq = "".join("@q%d\n%s\n+\n%s\n" % (i, "ACGT" * int(rng.integers(1, 30)), "@>+I" * 3) for i in range(500)).encode()
    for name, text, fmt in (("a.fa", fasta, "fasta"), ("b.fq", fastq, "fastq"), ("c.fq", fastq[:-1], "fastq")):
        for variant in ("plain", "gz", "gz2"):
            p = str(tmp_path / (name + ("" if variant == "plain" else ".gz")))
            blob = text if variant == "plain" else (gzip.compress(text) if variant == "gz" else gzip.compress(text[:5000]) + gzip.compress(text[5000:]))
            open(p, "wb").write(blob)
            for piece in (257, 4096, 1 << 20):
                hs = (hashlib.md5(), hashlib.sha256())
                pieces = list(parse._iter_text_pieces(p, piece, hs))
                assert b"".join(pieces) == text and max(len(x) for x in pieces) <= max(piece, 64)
                assert hs[0].hexdigest() == hashlib.md5(blob).hexdigest() and hs[1].hexdigest() == hashlib.sha256(blob).hexdigest()
                frags = list(parse._iter_fragments(iter(pieces), fmt))
                assert b"".join(frags) == text
                off = 0
                for f in frags[:-1]:
                    off += len(f)
                    if fmt == "fasta":
                        assert text[off:off + 1] == b">" and text[off - 1:off] == b"\n"
                    else:
                        assert text[:off].count(b"\n") % 4 == 0 and text[off - 1:off] == b"\n"
    # a truncated gzip stream raises like gzip.open does
    p = str(tmp_path / "t.fa.gz")
    open(p, "wb").write(gzip.compress(fasta)[:-20])
    with pytest.raises(EOFError):
        list(parse._iter_text_pieces(p, 4096, ()))
    with pytest.raises(EOFError):
        gzip.open(p, "rb").read()


def test_streaming_keeps_host_memory_bounded(tmp_path):
    """the input is never held whole.  A 48 MiB FASTA file (and its gzip) streamed in 1 MiB pieces:
    the Python-level peak while reading, inflating and cutting fragments stays a few pieces, independent of the file size."""
    import tracemalloc
    import numpy as np
    rng = np.random.default_rng(1)
    line = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=150)) + b"\n"
    rec = b">read\n" + line
    n = (48 << 20) // len(rec)
    plain = str(tmp_path / "big.fa")
    with open(plain, "wb") as fh:
        for _ in range(64):
            fh.write(rec * (n // 64))
    gz = str(tmp_path / "big.fa.gz")
    with gzip.open(gz, "wb", compresslevel=1) as out, open(plain, "rb") as src:
        for chunk in iter(lambda: src.read(1 << 22), b""):
            out.write(chunk)
    piece = 1 << 20
    for path in (plain, gz):
        tracemalloc.start()
        total = 0
        for frag in parse._iter_fragments(parse._iter_text_pieces(path, piece, (hashlib.md5(),)), "fasta"):
            total += len(frag)
            assert frag[:1] == b">"
        _cur, peak = tracemalloc.get_traced_memory()
        tracemalloc.stop()
        assert total == os.path.getsize(plain)
        assert peak < 8 * piece, (path, peak)
