"""Minimal .kdb writer/reader so the CLI produces the reference's file format (kmerdb/fileutil.py:486-565,
kmerdb/__init__.py:1980-1990).  Host code: the north star keeps .kdb writing on the host; this is a faster
restatement of it (rows formatted per chunk, BGZF blocks deflated on a thread pool), byte-compatible in layout:
YAML header + delimiter in its own BGZF block(s), then TSV rows `i<TAB>kmer_id<TAB>count<TAB>frequency`."""
import gzip
import math
import struct
import sys
import zlib
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import yaml

HEADER_DELIMITER = "\n" + ("=" * 24) + "\n"            # kmerdb/config.py:22
_BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")
_BLOCK = 65536 - 256                                     # payload per BGZF block (Bio.bgzf writes <= 65536; leave room for incompressible data)


def _host_threads():
    """Threads for the deflate / inflate pools: the CPUs this process may run on, at most 32."""
    import os
    try:
        return max(1, min(32, len(os.sched_getaffinity(0))))
    except AttributeError:
        return max(1, min(32, os.cpu_count() or 1))


def bgzf_block(data, level=6):
    """One BGZF block (gzip member with the BC extra field holding the block size)."""
    c = zlib.compressobj(level, zlib.DEFLATED, -15, zlib.DEF_MEM_LEVEL, 0)
    comp = c.compress(data) + c.flush()
    bsize = len(comp) + 25
    if bsize > 65535:                                    # incompressible: store
        c = zlib.compressobj(0, zlib.DEFLATED, -15)
        comp = c.compress(data) + c.flush()
        bsize = len(comp) + 25
    head = b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", bsize)
    return head + comp + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data) & 0xFFFFFFFF)


def header_bytes(metadata):
    """YAML header + delimiter with `metadata_blocks` settled the way KDBWriter does (fileutil.py:530-538)."""
    md = dict(metadata)
    for _ in range(2):
        raw = bytes(yaml.dump(md, sort_keys=False), "utf-8") + bytes(HEADER_DELIMITER, "utf-8")
        md["metadata_blocks"] = math.ceil(sys.getsizeof(raw) / (2 ** 16))
    return bytes(yaml.dump(md, sort_keys=False), "utf-8") + bytes(HEADER_DELIMITER, "utf-8"), md


def format_rows(start, counts, total_kmers):
    """Rows `i\\tid\\tcount\\tfrequency\\n` (kmerdb/__init__.py:1980-1990): count as np.uint64 prints, frequency as
    np.float64 prints (shortest round-trip repr)."""
    freq = np.divide(counts, total_kmers)                # float64, __init__.py:1945
    ids = range(start, start + len(counts))
    return "".join(["%d\t%d\t%d\t%r\n" % (i, i, c, f) for i, c, f in zip(ids, counts.tolist(), freq.tolist())]).encode("ascii")


def count_strings(counts, total_kmers):
    """The text "count\tfrequency" of every distinct count value (ascending), as the reference's str.format prints
    np.uint64 and np.float64 (kmerdb/__init__.py:1945, :1990).  Returns (uniq uint64, str_off uint32, bytes)."""
    mx = int(counts.max()) if len(counts) else 0
    if mx < (1 << 24):
        uniq = np.flatnonzero(np.bincount(counts.astype(np.int64), minlength=1)).astype(np.uint64)
    else:
        uniq = np.unique(counts)
    tot = float(total_kmers)
    strs = [("%d\t%r" % (c, (c / tot) if tot else float("nan"))).encode("ascii") for c in uniq.tolist()]
    off = np.zeros(len(strs) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in strs])
    return uniq, off, b"".join(strs)


def write_kdb_gpu(path, metadata, counts, device=0, threads=None, quiet=True, rows_per_call=1 << 22, total_kmers=None):
    """write_kdb with the rows assembled on the GPU (kdb_format_rows); BGZF deflate on host threads.
    `metadata` may be a callable (with `total_kmers` given): the body is then formatted and compressed first and the
    callable is asked when the header is written (the file checksums of a large input finish later than its counts)."""
    from . import _native
    deferred = callable(metadata)
    if deferred and total_kmers is None:
        raise ValueError("write_kdb_gpu: a deferred header needs total_kmers for the frequency column")
    if not deferred:
        head, md = header_bytes(metadata)
        total_kmers = md["total_kmers"]
    uniq, off, blob = count_strings(counts, total_kmers)
    threads = threads or _host_threads()

    def deflate(buf):
        return b"".join(bgzf_block(bytes(buf[o:o + _BLOCK])) for o in range(0, len(buf), _BLOCK))

    def body(ex):
        for lo in range(0, len(counts), rows_per_call):
            text = _native.format_rows(counts[lo:lo + rows_per_call], lo, uniq, off, blob, device=device)
            if not quiet:
                sys.stdout.write(text.tobytes().decode("ascii"))
            step = _BLOCK * 32
            mv = memoryview(text)
            yield from ex.map(deflate, [mv[o:o + step] for o in range(0, len(text), step)])

    with ThreadPoolExecutor(max_workers=threads) as ex:
        held = list(body(ex)) if deferred else None            # compressed body in memory while the header is still unknown
        if deferred:
            head, md = header_bytes(metadata())
            if md["total_kmers"] != total_kmers:
                raise ValueError("write_kdb_gpu: the deferred header disagrees with total_kmers")
        with open(path, "wb") as fh:
            for o in range(0, len(head), 65536):
                fh.write(bgzf_block(head[o:o + 65536]))
            for blobz in (held if deferred else body(ex)):
                fh.write(blobz)
            fh.write(_BGZF_EOF)
    return md


def write_kdb(path, metadata, counts, threads=None, quiet=True, rows_per_task=1 << 15):
    """Write `<path>`: header block(s), 4^k rows, BGZF EOF marker (rows formatted on the host)."""
    head, md = header_bytes(metadata)
    total = md["total_kmers"]
    n = len(counts)
    threads = threads or _host_threads()

    def task(lo):
        body = format_rows(lo, counts[lo:lo + rows_per_task], total)
        if not quiet:
            sys.stdout.write(body.decode("ascii"))
        return b"".join(bgzf_block(body[o:o + _BLOCK]) for o in range(0, len(body), _BLOCK))

    with open(path, "wb") as fh:
        for o in range(0, len(head), 65536):
            fh.write(bgzf_block(head[o:o + 65536]))
        with ThreadPoolExecutor(max_workers=threads) as ex:
            for blob in ex.map(task, range(0, n, rows_per_task)):
                fh.write(blob)
        fh.write(_BGZF_EOF)
    return md


_ROW_ERRORS = {33: (AssertionError, "kmerdb.fileutil.KDBReader requires .kdb files to have 4 columns."),
               34: (ValueError, "invalid literal for int() with base 10"),
               35: (AssertionError, "Line number doesn't match file row index"),
               36: (IndexError, "index is out of bounds for axis 0"),
               37: (ValueError, "a k-mer id appears on two rows")}


def _inflate_groups(raw, threads=None, group=2048):
    """The inflated text of a BGZF file, yielded in pieces of `group` blocks (inflate on a thread pool)."""
    from .parse import _bgzf_blocks, _inflate_member
    blocks = _bgzf_blocks(raw)
    if blocks is None:                                       # ordinary gzip: one piece
        yield gzip.decompress(raw)
        return
    mv = memoryview(raw)
    with ThreadPoolExecutor(max_workers=threads or _host_threads()) as ex:
        for g in range(0, len(blocks), group):
            yield b"".join(ex.map(lambda b: _inflate_member(mv, b), blocks[g:g + group]))


class KDBReader:
    """The slurped view of a .kdb file: `.metadata`, `.k`, `.kmer_ids`, `.counts`, `.frequencies`
    (kmerdb/fileutil.py:KDBReader with slurp=True, sort=False, :308-377).  The row loop runs on the GPU
    (kdb_reader_*, kdb_read.cuh); header YAML and BGZF inflate stay on the host."""

    def __init__(self, path, slurp=True, device=0, threads=None):
        from . import _native
        self._filepath = path
        with open(path, "rb") as fh:
            raw = fh.read()
        delim = HEADER_DELIMITER.encode("ascii")
        pieces = _inflate_groups(raw, threads)
        head = b""
        for piece in pieces:
            head += piece
            if delim in head:
                break
        if delim not in head:
            raise ValueError("Requires kmerdb format YAML header. Body is .tsv format table, .bgzf file.")
        cut = head.index(delim)
        self.metadata = yaml.safe_load(head[:cut].decode("utf-8"))
        self.k = self.metadata["k"]
        self.kmer_ids = self.counts = self.frequencies = None
        if not slurp:
            return
        n = 4 ** self.k
        with _native.RowReader(self.k, device) as rd:
            rest = head[cut + len(delim):]
            for piece in pieces:
                buf = rest + piece if rest else piece
                end = buf.rfind(b"\n") + 1
                rd.feed(memoryview(buf)[:end])
                rest = buf[end:]
            rd.feed(rest)
            rc, counts, ids, n_rows, err_row = rd.finish()
        if rc:
            cls, msg = _ROW_ERRORS[rc]
            raise cls("{0} (row {1} of {2})".format(msg, err_row, path))
        self.counts, self.kmer_ids = counts, ids
        self.frequencies = counts.astype(np.float64) / n              # fileutil.py:362 float(count)/N, N = 4^k

    def slurp(self):
        return self.counts


def open_kdb(path, mode="r", slurp=True, device=0):
    """fileutil.open(path, 'r', slurp=...) for reading."""
    if mode != "r":
        raise ValueError("kmerdb_b200.kdbfile.open_kdb reads; use write_kdb to write")
    return KDBReader(path, slurp=slurp, device=device)


def read_kdb(path, device=0):
    """(metadata dict, counts uint64[4^k]) of a .kdb file written by kmerdb or by write_kdb."""
    r = KDBReader(path, device=device)
    return r.metadata, r.counts


def matrix_from(paths, column_names=None, device=0):
    """`kmerdb matrix from a.kdb b.kdb ...` collation (kmerdb/__init__.py:844-896): (profiles uint64[4^k, n], columns)."""
    import os
    ks = [KDBReader(f, slurp=False).k for f in paths]
    if not all(k == ks[0] for k in ks):
        raise TypeError("One or more files did not have k set to be equal to {0}".format(ks[0]))
    profiles = np.empty((4 ** ks[0], len(paths)), dtype=np.uint64)
    for j, f in enumerate(paths):
        profiles[:, j] = KDBReader(f, device=device).counts
    columns = list(column_names) if column_names is not None else [os.path.basename(f).split(".")[0] for f in paths]
    if len(columns) != len(paths):
        raise RuntimeError("Number of column names {0} does not match number of input files {1}...".format(len(columns), len(paths)))
    return profiles, columns


def read_kdb_host(path):
    """The reference's row loop on the host (slow; kept for the tests as the checker of the GPU reader)."""
    with gzip.open(path, "rt") as fh:
        text = fh.read()
    head, _, body = text.partition(HEADER_DELIMITER)
    md = yaml.safe_load(head)
    counts = np.zeros(4 ** md["k"], dtype=np.uint64)
    for ln in body.splitlines():
        if ln:
            f = ln.split("\t")
            counts[int(f[1])] = int(f[2])
    return md, counts
