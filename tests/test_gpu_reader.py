"""GPU .kdb reader (kdb_reader_*, kdb_read.cuh) against golden vectors frozen from the unmodified
reference reader (tests/golden/make_golden_kdb.py -> golden_kdb.json), plus round trips through the writers."""
import builtins
import json
import os

import numpy as np
import pytest

from kmerdb_b200 import kdbfile

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden_kdb.json")))["cases"]


def _metadata(k, counts):
    total, uniq = int(np.sum(counts)), int(np.count_nonzero(counts))
    f = {"filename": "x.fa", "md5": "a" * 32, "sha256": "b" * 64, "total_reads": 1, "total_kmers": total, "unique_kmers": uniq,
         "nullomers": 4 ** k - uniq, "min_read_length": 10, "max_read_length": 10, "avg_read_length": 10}
    return {"version": "0.9.6", "metadata_blocks": 1, "k": k, "total_kmers": total, "unique_kmers": uniq, "unique_nullomers": 4 ** k - uniq,
            "sorted": False, "tags": [], "files": [f]}


def _assemble(path, k, body, block):
    head, _ = kdbfile.header_bytes(_metadata(k, np.zeros(1)))
    with open(path, "wb") as fh:
        fh.write(kdbfile.bgzf_block(head))
        for o in range(0, len(body), block):
            fh.write(kdbfile.bgzf_block(body[o:o + block]))
        fh.write(kdbfile._BGZF_EOF)


def test_golden_file_is_from_the_reference():
    assert len(GOLD) >= 24 and sum("raises" in c for c in GOLD) >= 10


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLD, ids=[c["name"] for c in GOLD])
def test_reader_matches_reference(case, tmp_path):
    p = str(tmp_path / ("t.%d.kdb" % case["k"]))
    _assemble(p, case["k"], case["body"].encode("ascii"), case["block"])
    if "raises" in case:
        with pytest.raises(getattr(builtins, case["raises"])):
            kdbfile.KDBReader(p)
        return
    r = kdbfile.KDBReader(p)
    assert r.k == case["k"] and r.counts.dtype == np.uint64 and r.kmer_ids.dtype == np.uint64 and r.frequencies.dtype == np.float64
    assert [int(x) for x in r.counts] == case["counts"]
    assert [int(x) for x in r.kmer_ids] == case["kmer_ids"]
    assert [float(x) for x in r.frequencies] == case["frequencies"]          # count / 4^k in float64: exact


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 6, 10])
def test_writer_reader_round_trip(k, tmp_path):
    rng = np.random.default_rng(k)
    counts = rng.integers(0, 2 ** 40, size=4 ** k, dtype=np.uint64)
    counts[rng.integers(0, 4 ** k, size=4 ** k // 3)] = 0
    p = str(tmp_path / ("rt.%d.kdb" % k))
    kdbfile.write_kdb_gpu(p, _metadata(k, counts), counts)
    r = kdbfile.KDBReader(p)
    assert np.array_equal(r.counts, counts) and np.array_equal(r.kmer_ids, np.arange(4 ** k, dtype=np.uint64))
    md, c2 = kdbfile.read_kdb(p)
    assert md["k"] == k and np.array_equal(c2, counts)
    if k <= 6:
        assert np.array_equal(kdbfile.read_kdb_host(p)[1], counts)


@pytest.mark.gpu
def test_reader_chunked_feed_and_row_numbers():
    """Pieces cut at arbitrary row ends (and a last row without line feed) give the same vectors as one piece."""
    from kmerdb_b200 import _native
    k = 7
    n = 4 ** k
    rng = np.random.default_rng(3)
    counts = rng.integers(0, 10 ** 9, size=n)
    ids = rng.permutation(n)
    lines = ["%d\t%d\t%d\t%r\n" % (j, ids[j], counts[j], counts[j] / 7.0) for j in range(n)]
    want = np.zeros(n, dtype=np.uint64)
    want[ids] = counts
    for cuts in ([], [1], [5000, 5001, 9000], sorted(rng.integers(1, n, size=40).tolist())):
        with _native.RowReader(k) as rd:
            lo = 0
            for c in cuts + [n]:
                rd.feed("".join(lines[lo:c]).encode() if c < n else "".join(lines[lo:c]).encode()[:-1])
                lo = c
            rc, got, got_ids, n_rows, _ = rd.finish()
        assert rc == 0 and n_rows == n and np.array_equal(got, want) and np.array_equal(got_ids, ids.astype(np.uint64))


@pytest.mark.gpu
def test_reader_duplicate_id_and_short_body():
    from kmerdb_b200 import _native
    k = 3
    body = "".join("%d\t%d\t1\t0.1\n" % (j, j) for j in range(64))
    with _native.RowReader(k) as rd:
        rd.feed(body.replace("17\t17\t", "17\t16\t").encode())
        assert rd.finish()[0] == _native.KDB_R_DUP
    with _native.RowReader(k) as rd:
        rd.feed(body[:body.index("60\t60")].encode())
        rc, _, _, n_rows, err_row = rd.finish()
        assert (rc, n_rows, err_row) == (_native.KDB_R_INDEX, 60, 60)


@pytest.mark.gpu
def test_matrix_from(tmp_path):
    k = 5
    rng = np.random.default_rng(5)
    cols, paths = [], []
    for name in ("alpha", "beta", "gamma"):
        c = rng.integers(0, 1000, size=4 ** k, dtype=np.uint64)
        p = str(tmp_path / (name + ".%d.kdb" % k))
        kdbfile.write_kdb(p, _metadata(k, c), c)
        cols.append(c)
        paths.append(p)
    m, names = kdbfile.matrix_from(paths)
    assert names == ["alpha", "beta", "gamma"] and m.shape == (4 ** k, 3) and m.dtype == np.uint64      # __init__.py:861-872
    assert np.array_equal(m, np.stack(cols, axis=1))
    other = str(tmp_path / "delta.4.kdb")
    c4 = np.ones(4 ** 4, dtype=np.uint64)
    kdbfile.write_kdb(other, _metadata(4, c4), c4)
    with pytest.raises(TypeError):
        kdbfile.matrix_from(paths + [other])                                                              # __init__.py:852-859
