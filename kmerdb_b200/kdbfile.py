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


def write_kdb_gpu(path, metadata, counts, device=0, threads=8, quiet=True, rows_per_call=1 << 22):
    """write_kdb with the rows assembled on the GPU (kdb_format_rows); BGZF deflate on host threads."""
    from . import _native
    head, md = header_bytes(metadata)
    uniq, off, blob = count_strings(counts, md["total_kmers"])

    def deflate(buf):
        return b"".join(bgzf_block(bytes(buf[o:o + _BLOCK])) for o in range(0, len(buf), _BLOCK))

    with open(path, "wb") as fh:
        for o in range(0, len(head), 65536):
            fh.write(bgzf_block(head[o:o + 65536]))
        with ThreadPoolExecutor(max_workers=threads) as ex:
            for lo in range(0, len(counts), rows_per_call):
                text = _native.format_rows(counts[lo:lo + rows_per_call], lo, uniq, off, blob, device=device)
                if not quiet:
                    sys.stdout.write(text.tobytes().decode("ascii"))
                step = _BLOCK * 32
                mv = memoryview(text)
                for blobz in ex.map(deflate, [mv[o:o + step] for o in range(0, len(text), step)]):
                    fh.write(blobz)
        fh.write(_BGZF_EOF)
    return md


def write_kdb(path, metadata, counts, threads=8, quiet=True, rows_per_task=1 << 15):
    """Write `<path>`: header block(s), 4^k rows, BGZF EOF marker (rows formatted on the host)."""
    head, md = header_bytes(metadata)
    total = md["total_kmers"]
    n = len(counts)

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


def read_kdb(path):
    """(metadata dict, counts uint64[4^k]) of a .kdb file written by kmerdb or by write_kdb (4-column rows)."""
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
