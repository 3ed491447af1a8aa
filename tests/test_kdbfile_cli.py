"""The .kdb writer (host code, CPU) and the CLI surface; the CLI end-to-end run needs the GPU."""
import gzip
import os
import struct

import numpy as np
import pytest

from kmerdb_b200 import cli, kdbfile


def _bgzf_blocks(raw):
    """Walk the BGZF members: every block carries the BC extra field with its own size."""
    off, sizes = 0, []
    while off < len(raw):
        assert raw[off:off + 4] == b"\x1f\x8b\x08\x04" and raw[off + 12:off + 16] == b"BC\x02\x00"
        bsize = struct.unpack("<H", raw[off + 16:off + 18])[0] + 1
        sizes.append(bsize)
        off += bsize
    assert off == len(raw)
    return sizes


def test_write_kdb_layout_and_rows(tmp_path):
    rng = np.random.default_rng(1)
    counts = rng.integers(0, 1000, size=4 ** 6).astype(np.uint64)
    total = int(counts.sum())
    md = {"version": "0.9.6", "metadata_blocks": 1, "k": 6, "total_kmers": total, "unique_kmers": int(np.count_nonzero(counts)),
          "unique_nullomers": 0, "sorted": False, "tags": [], "files": [{"filename": "x.fa", "total_reads": 3}]}
    path = str(tmp_path / "o.6.kdb")
    kdbfile.write_kdb(path, md, counts, rows_per_task=1000)
    raw = open(path, "rb").read()
    sizes = _bgzf_blocks(raw)
    assert sizes[-1] == 28 and raw.endswith(kdbfile._BGZF_EOF)              # BGZF EOF marker
    first = gzip.decompress(raw[:sizes[0]]).decode()
    assert first.endswith(kdbfile.HEADER_DELIMITER) and first.startswith("version: 0.9.6\nmetadata_blocks: 1\nk: 6\n")   # header alone in block 0
    text = gzip.open(path, "rt").read()
    rows = text.split(kdbfile.HEADER_DELIMITER)[1].splitlines()
    assert len(rows) == 4 ** 6
    for i in (0, 1, 77, 4095):
        # kmerdb/__init__.py:1990 formats np.uint64 and np.float64 with str.format
        assert rows[i] == "{0}\t{1}\t{2}\t{3}".format(i, i, np.uint64(counts[i]), np.divide(counts, total)[i])
    md2, c2 = kdbfile.read_kdb_host(path)
    assert md2["k"] == 6 and md2["files"][0]["filename"] == "x.fa" and np.array_equal(c2, counts)


def test_cli_flag_surface():
    ap = cli.build_parser()
    a = ap.parse_args(["profile", "-k", "12", "-o", "out", "--no-ambiguous", "--do-not-canonicalize", "--quiet", "reads.fa"])
    assert a.k == 12 and a.output_name == "out" and a.no_ambiguous and a.do_not_canonicalize and a.input == ["reads.fa"]
    a = ap.parse_args(["profile", "--minK", "4", "--maxK", "14", "-o", "out", "reads.fq.gz"])
    assert a.k is None and (a.minK, a.maxK) == (4, 14)
    a = ap.parse_args(["graph", "-k", "12", "reads.fa", "out.kdbg"])
    assert a.kdbg == "out.kdbg" and a.input == ["reads.fa"]
    a = ap.parse_args(["minimizers", "ref.fa", "ref.12.kdb", "-w", "10"])
    assert a.window_size == 10
    with pytest.raises(ValueError):
        cli.profile(ap.parse_args(["profile", "-o", "o", "a.fa"]))          # no -k and no --minK/--maxK
    with pytest.raises(ValueError):
        cli.profile(ap.parse_args(["profile", "-k", "8", "-o", "o", "a.fa", "b.fa"]))


@pytest.mark.gpu
def test_cli_profile_graph_end_to_end(tmp_path):
    import oracle
    rng = np.random.default_rng(9)
    fa = str(tmp_path / "s.fa")
    text = "".join(">r%d\n%s\n" % (i, "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=90)])) for i in range(50))
    open(fa, "w").write(text)
    cli.cli(["profile", "--minK", "4", "--maxK", "6", "-o", str(tmp_path / "out"), "--quiet", fa])
    for k in (4, 5, 6):
        md, counts = kdbfile.read_kdb(str(tmp_path / ("out.%d.kdb" % k)))
        rc, oc, ost = oracle.profile(text.encode(), k)
        assert np.array_equal(counts, oc) and md["total_kmers"] == ost["total_kmers"] and md["unique_kmers"] == ost["unique_kmers"]
        assert md["unique_nullomers"] == int(4 ** k / 2 - ost["unique_kmers"]) and md["files"][0]["total_reads"] == 50
    cli.cli(["graph", "-k", "5", "--quiet", fa, str(tmp_path / "g.kdbg")])
    body = gzip.open(str(tmp_path / "g.kdbg"), "rt").read().split(kdbfile.HEADER_DELIMITER)[1].splitlines()
    assert len(body) == 50 * (90 - 5) and body[0].split("\t")[:3] == ["0", "r0", "0"] and len(body[0].split("\t")) == 8


@pytest.mark.gpu
def test_gpu_rows_are_byte_identical_to_host_rows(tmp_path):
    """kdb_format_rows (GPU) against the host formatter, which itself is checked against str.format of numpy scalars."""
    import time
    rng = np.random.default_rng(3)
    for k, hi in ((6, 5), (9, 3000), (12, 200)):
        counts = rng.integers(0, hi, size=4 ** k).astype(np.uint64)
        counts[rng.integers(0, 4 ** k, size=7)] = np.array([0, 1, 2 ** 31, 2 ** 40 + 7, 10 ** 15, 999999, 1000000], dtype=np.uint64)
        total = int(counts.sum())
        md = {"version": "0.9.6", "metadata_blocks": 1, "k": k, "total_kmers": total, "unique_kmers": int(np.count_nonzero(counts)),
              "unique_nullomers": 0, "sorted": False, "tags": [], "files": []}
        t0 = time.perf_counter()
        kdbfile.write_kdb_gpu(str(tmp_path / "g.kdb"), md, counts)
        t1 = time.perf_counter()
        g = gzip.open(str(tmp_path / "g.kdb"), "rb").read()
        if k <= 9:
            kdbfile.write_kdb(str(tmp_path / "h.kdb"), md, counts)
            assert g == gzip.open(str(tmp_path / "h.kdb"), "rb").read()
        else:
            body = g.split(kdbfile.HEADER_DELIMITER.encode())[1]
            lo, n = 4 ** k - 5000, 5000
            assert body.endswith(kdbfile.format_rows(lo, counts[lo:], total)) and body.startswith(kdbfile.format_rows(0, counts[:n], total))
            assert body.count(b"\n") == 4 ** k
            print("k=%d: %d rows, %.1f MB of text, written in %.2f s" % (k, 4 ** k, len(body) / 1e6, t1 - t0))
    z = np.zeros(256, dtype=np.uint64)                                  # all-zero profile: 0/0 prints as nan, like np.divide
    uniq, off, blob = kdbfile.count_strings(z, 0)
    assert blob == b"0\tnan"
