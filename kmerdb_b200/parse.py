"""Host-side mirror of kmerdb/parse.py for the `profile` hot path.

`parsefile` keeps the reference's signature, return tuple, metadata keys and exception classes
(kmerdb/parse.py:90-163); the per-window Python loop (parse.py:128-137 -> kmer.shred,
kmerdb/kmer.py:489-577) is replaced by the CUDA kernels behind include/kmerdb_b200.h.  Host code
only opens / gunzips the file (unchanged from the reference), checksums it and turns device status
codes back into the reference's exceptions.  There is no CPU counting path."""
import gzip
import hashlib
import io
import os
from collections import namedtuple

import numpy as np

from . import _native

SeqRecord = namedtuple("SeqRecord", ["id", "seq", "description"])

FASTA_SUFFIXES = (".fna", ".fna.gz", ".fa.gz", ".fa", ".fasta", ".fasta.gz")      # kmerdb/util.py:120-126
FASTQ_SUFFIXES = (".fastq", ".fastq.gz", ".fq.gz", ".fq")                         # kmerdb/util.py:128-134


def is_fasta(fname):
    if type(fname) is not str:
        raise TypeError("kmerdb.util.is_fasta() expects a str argument")
    return fname.endswith(FASTA_SUFFIXES)


def is_fastq(fname):
    if type(fname) is not str:
        raise TypeError("kmerdb.util.is_fasta() expects a str argument")
    return fname.endswith(FASTQ_SUFFIXES)


def is_gz_file(filepath):
    """kmerdb/util.py:80-88: gzip is decided by trying to read, not by the suffix."""
    try:
        with gzip.open(filepath, "rb") as f:
            f.readline()
        return True
    except OSError:
        return False


def checksum(filepath):
    """kmerdb/util.py:35-50: md5 and sha256 of the raw (possibly compressed) file bytes."""
    if type(filepath) is not str:
        raise TypeError("kmerdb.util.checksum() expects a str as its argument")
    elif not os.path.exists(filepath):
        raise IOError("could not find '{}' on the filesystem".format(filepath))
    md5, sha = hashlib.md5(), hashlib.sha256()
    with open(filepath, "rb") as fh:
        for chunk in iter(lambda: fh.read(1 << 20), b""):
            md5.update(chunk)
            sha.update(chunk)
    return md5.hexdigest(), sha.hexdigest()


def _bgzf_blocks(raw):
    """(offset, size, payload offset) of the members of a BGZF file (gzip members whose extra field carries their size in a
    'BC' subfield, e.g. written by bgzip), or None when `raw` is ordinary gzip.  The extra field is walked subfield by
    subfield (RFC 1952 2.3.1.1): XLEN is read, not assumed to be 6."""
    out, off, n = [], 0, len(raw)
    while off < n:
        if n - off < 26 or raw[off:off + 3] != b"\x1f\x8b\x08" or not raw[off + 3] & 4:
            return None
        flg = raw[off + 3]
        xlen = int.from_bytes(raw[off + 10:off + 12], "little")
        p, xend, size = off + 12, off + 12 + xlen, None
        while p + 4 <= xend:
            slen = int.from_bytes(raw[p + 2:p + 4], "little")
            if raw[p:p + 2] == b"BC" and slen == 2:
                size = int.from_bytes(raw[p + 4:p + 6], "little") + 1
            p += 4 + slen
        if size is None or off + size > n:
            return None
        q = xend
        for bit in (8, 16):                       # FNAME, FCOMMENT: zero-terminated
            if flg & bit:
                q = raw.index(b"\x00", q) + 1 if isinstance(raw, (bytes, bytearray)) else bytes(raw[q:off + size]).index(b"\x00") + q + 1
        if flg & 2:
            q += 2
        out.append((off, size, q))
        off += size
    return out if off == n else None


def _inflate_member(mv, b):
    """One BGZF member on the host, checked against its trailer like gzip.open checks it (CRC-32 and ISIZE)."""
    import zlib
    off, size, q = b
    text = zlib.decompress(mv[q:off + size - 8], -15)
    crc, isize = int.from_bytes(mv[off + size - 8:off + size - 4], "little"), int.from_bytes(mv[off + size - 4:off + size], "little")
    if zlib.crc32(text) & 0xFFFFFFFF != crc or len(text) & 0xFFFFFFFF != isize:
        raise gzip.BadGzipFile("CRC check failed" if len(text) & 0xFFFFFFFF == isize else "Incorrect length of data produced")
    return text


def _gunzip(raw, threads=8):
    """Host gzip decode (kmerdb/parse.py:63-68 uses gzip.open).  Block-compressed input (BGZF) is inflated on a
    thread pool, one member per task, each checked against its CRC-32 / ISIZE; plain gzip goes through zlib in one piece.
    Same bytes either way; anything odd about the BGZF framing falls back to gzip.decompress, which raises what
    gzip.open raises."""
    blocks = _bgzf_blocks(raw) if len(raw) > (1 << 20) else None
    if not blocks:
        return gzip.decompress(raw)
    from concurrent.futures import ThreadPoolExecutor
    mv = memoryview(raw)
    try:
        with ThreadPoolExecutor(max_workers=threads) as ex:
            parts = list(ex.map(lambda b: _inflate_member(mv, b), blocks))
    except Exception:
        return gzip.decompress(raw)
    return b"".join(parts)


def _read_text_bytes(filepath, raw=None):
    """Uncompressed file text (host gzip decode, as in kmerdb/parse.py:63-68)."""
    if raw is None:
        with open(filepath, "rb") as fh:
            raw = fh.read()
    if raw[:2] == b"\x1f\x8b" and is_gz_file(filepath):
        return _gunzip(raw)
    return raw


def _checksums_async(raw):
    """md5 and sha256 of the raw file bytes (kmerdb/util.py:35-50) on two host threads (hashlib releases the GIL), so
    they run beside the H2D copy and the kernels instead of in front of them.  Returns a callable giving (md5, sha256)."""
    from concurrent.futures import ThreadPoolExecutor
    ex = ThreadPoolExecutor(max_workers=2)
    f1 = ex.submit(lambda: hashlib.md5(raw).hexdigest())
    f2 = ex.submit(lambda: hashlib.sha256(raw).hexdigest())

    def result():
        try:
            return f1.result(), f2.result()
        finally:
            ex.shutdown(wait=True)
    return result


def _read_and_hash(filepath, chunk=32 << 20):
    """The raw bytes of the file, and a callable giving (md5, sha256) of them (kmerdb/util.py:35-50).  The file is read in
    pieces straight into one buffer; two host threads hash the pieces as they arrive (hashlib releases the GIL) and go on
    hashing while the caller feeds the GPU."""
    import queue
    import threading
    size = os.path.getsize(filepath)
    buf = bytearray(size)
    mv = memoryview(buf)
    hashes = (hashlib.md5(), hashlib.sha256())
    queues = (queue.Queue(), queue.Queue())

    def work(h, q):
        while True:
            piece = q.get()
            if piece is None:
                return
            h.update(piece)

    threads = [threading.Thread(target=work, args=hq, daemon=True) for hq in zip(hashes, queues)]
    for t in threads:
        t.start()
    off = 0
    tail = b""
    with open(filepath, "rb", buffering=0) as fh:
        while off < size:
            n = fh.readinto(mv[off:min(size, off + chunk)])
            if not n:
                break
            for q in queues:
                q.put(mv[off:off + n])
            off += n
        if off == size:
            tail = fh.read()                                     # the file grew while it was read
            if tail:
                for q in queues:
                    q.put(tail)
    for q in queues:
        q.put(None)
    raw = mv[:off] if not tail else bytes(mv[:off]) + tail

    def result():
        for t in threads:
            t.join()
        return hashes[0].hexdigest(), hashes[1].hexdigest()
    return raw, result


def _file_format(filepath):
    if is_fasta(filepath):
        return "fasta"
    if is_fastq(filepath):
        return "fastq"
    raise ValueError("Could not determine the format of file '{0}'".format(filepath))   # parse.py:74


# ---- host tokeniser ---------------------------------------------------------------------------------
# Used (a) by parse_sequence_file, which must yield records to Python callers, and (b) when the
# device tokeniser answers KDB_E_LAYOUT (lone CR line breaks, multi-line FASTQ, "+title" repeats,
# control or non-ASCII bytes): the records are re-laid out as plain one-line FASTA and counted on
# the GPU all the same.  Semantics follow Biopython 1.83 (Pipfile.lock:28-60) as called at
# kmerdb/parse.py:70-72: text mode, UTF-8, universal newlines.

def _iter_fasta(handle):
    title, lines = None, []
    for line in handle:
        if line[0] == ">":
            if title is not None:
                yield title, "".join(lines).replace(" ", "").replace("\r", "")
            title, lines = line[1:].rstrip(), []
        elif title is not None:
            lines.append(line.rstrip())
    if title is not None:
        yield title, "".join(lines).replace(" ", "").replace("\r", "")


def _iter_fastq(handle):
    line = handle.readline()
    while line:
        if line[0] != "@":
            raise ValueError("Records in Fastq files should start with '@' character")
        title = line[1:].rstrip()
        seq = handle.readline().rstrip()
        while True:
            line = handle.readline()
            if not line:
                raise ValueError("End of file without quality information.")
            if line[0] == "+":
                second = line[1:].rstrip()
                if second and second != title:
                    raise ValueError("Sequence and quality captions differ.")
                break
            seq += line.rstrip()
        if " " in seq or "\t" in seq:
            raise ValueError("Whitespace is not allowed in the sequence.")
        qual = handle.readline().rstrip()
        while True:
            line = handle.readline()
            if not line:
                break
            if line[0] == "@" and len(qual) >= len(seq):
                break
            qual += line.rstrip()
        if len(seq) != len(qual):
            raise ValueError("Lengths of sequence and quality values differs for %s (%i and %i)." % (title, len(seq), len(qual)))
        for ch in qual:
            if not 33 <= ord(ch) <= 126:
                raise ValueError("Invalid character in quality string")
        yield title, seq


def _iter_records(data, fmt):
    handle = io.TextIOWrapper(io.BytesIO(data), encoding="UTF-8", newline=None)
    return _iter_fasta(handle) if fmt == "fasta" else _iter_fastq(handle)


def parse_sequence_file(seq_filepath, return_tuple=True):
    """kmerdb/parse.py:50-85.  Yields (seq_id, seq) tuples, or record objects with .id/.seq."""
    if type(seq_filepath) is not str:
        raise TypeError("kmerdb.graph.parse_sequence_file() expects a fasta/fastq sequence filepath as a str as its only positional_regument")
    if os.path.exists(seq_filepath) is False or os.access(seq_filepath, os.R_OK) is False:
        raise ValueError("kmerdb.graph.parse_sequence_file() expects the filepath to be be readable on the filesystem")
    data = _read_text_bytes(seq_filepath)
    fmt = _file_format(seq_filepath)
    for title, seq in _iter_records(data, fmt):
        parts = title.split(None, 1)
        rid = parts[0] if parts else ""
        yield (rid, seq) if return_tuple else SeqRecord(rid, seq, title)


_PERMITTED = frozenset("ACGTRYSWKMBDHVN")


def _plain_layout(data, fmt):
    """Re-lay the records out as '>\\n' + letters + '\\n'.  Letters outside the permitted alphabet
    become '?' (one per character, so record lengths are kept; the device raises NameError for it,
    kmerdb/kmer.py:168-170).  Returns (bytes, pending_exception)."""
    out, pending = [], None
    try:
        for _title, seq in _iter_records(data, fmt):
            if not _PERMITTED.issuperset(seq):
                seq = "".join(ch if ch in _PERMITTED else "?" for ch in seq)
            out.append(">\n" + seq + "\n")
    except ValueError as e:          # grammar violation (or UnicodeDecodeError): raised after the records before it
        pending = e
    return "".join(out).encode("ascii"), pending


def _raise_for_status(status, st):
    if status == _native.KDB_E_ALPHABET:
        raise NameError("name 'permitted_AA_characters_extended' is not defined")        # kmer.py:170
    if status == _native.KDB_E_IUPAC:
        raise KeyError("non-ACGT IUPAC letter in record %d reached the 2-bit encoder" % st["err_record"])   # kmer.py:309
    if status == _native.KDB_E_SHORT:
        raise AttributeError("'Seq' object has no attribute 'id'")                        # kmer.py:461-463
    if status == _native.KDB_E_EMPTY:
        raise ValueError("max() iterable argument is empty")                              # parse.py:144
    if status == _native.KDB_E_FORMAT:
        raise ValueError("malformed sequence file")
    if status != 0:
        raise RuntimeError("kmerdb_b200: unexpected device status %d" % status)


class Profile:
    """Result of one file's profile: the attributes the reference spreads over parse.parsefile's
    return tuple and the .kdb header (counts, total/unique k-mers, nullomers, mononucleotides)."""

    def __init__(self, k, counts, stats, nullomer_array, timing=None):
        self.k = k
        self.counts = counts
        self.nullomer_array = nullomer_array
        self.total_reads = stats["total_reads"]
        self.total_kmers = stats["total_kmers"]
        self.unique_kmers = stats["unique_kmers"]
        self.nullomers = 4 ** k - stats["unique_kmers"]                        # parse.py:143
        self.min_read_length = stats["min_len"]
        self.max_read_length = stats["max_len"]
        self.avg_read_length = int(stats["sum_len"] / stats["total_reads"])      # int(np.mean(...)) parse.py:146
        self.mononucleotides = dict(zip("ACGT", stats["mono"]))
        self.timing = timing

    def file_metadata(self, filename, md5, sha256):
        return {
            "filename": filename, "md5": md5, "sha256": sha256,
            "total_reads": self.total_reads, "total_kmers": self.total_kmers, "unique_kmers": self.unique_kmers,
            "nullomers": self.nullomers, "min_read_length": self.min_read_length,
            "max_read_length": self.max_read_length, "avg_read_length": self.avg_read_length,
        }


def _device_list(device):
    """int | "0-3" | "0,2,5" | sequence of ints -> list of device indices."""
    if isinstance(device, (list, tuple)):
        return [int(d) for d in device]
    if isinstance(device, str):
        out = []
        for part in device.split(","):
            if "-" in part:
                a, b = part.split("-")
                out.extend(range(int(a), int(b) + 1))
            elif part.strip():
                out.append(int(part))
        return out
    return [int(device)]


def _profile_bytes_multi(data, k, fmt, replace_with_none, canonicalize, devices, want_nullomers):
    """One process, several GPUs (the reference's cross-file sum, kmerdb/__init__.py:1888-1891, applied to the shards of one
    file): the text is cut at record boundaries (distributed.shard_bounds), shard r is profiled on devices[r] by its own
    host thread (the ctypes calls release the GIL: the host->device copies and kernels of the GPUs overlap), then
    kdb_reduce_peers sums the tables over NVLink into devices[0] in one kernel per GPU.  Returns None when the device
    tokeniser wants the host layout (the caller then takes the single-GPU path)."""
    from concurrent.futures import ThreadPoolExecutor
    from . import distributed as kd
    a = data if isinstance(data, np.ndarray) else np.frombuffer(data, dtype=np.uint8)
    cuts = kd.shard_bounds(a, fmt, len(devices))
    kw = dict(canonical=canonicalize, skip_n=replace_with_none, fastq=(fmt == "fastq"))
    profs = [_native.Profiler(k, device=d, **kw) for d in devices]
    try:
        def work(r):
            profs[r].feed(a[cuts[r]:cuts[r + 1]])
            return profs[r].finish()
        with ThreadPoolExecutor(max_workers=len(devices)) as ex:
            statuses = list(ex.map(work, range(len(devices))))
        if any(s == _native.KDB_E_LAYOUT for s in statuses):
            return None
        status = _native.reduce_peers(profs)
        root = profs[0]
        st = root.stats()
        _raise_for_status(status, st)
        return Profile(k, root.counts(), st, root.nullomers() if want_nullomers else None, root.timing())
    finally:
        for p in profs:
            p.close()


def profile_bytes(data, k, fmt="fasta", replace_with_none=True, canonicalize=True, device=0, want_nullomers=True):
    """Profile uncompressed FASTA/FASTQ text held in host memory.  Raises the reference's exceptions.
    `device`: one device index, or several ("0-7", [0, 1, 2, 3]): the reads are sharded over them."""
    devices = _device_list(device)
    if len(devices) > 1:
        prof = _profile_bytes_multi(data, k, fmt, replace_with_none, canonicalize, devices, want_nullomers)
        if prof is not None:
            return prof
    device = devices[0]
    with _native.Profiler(k, canonical=canonicalize, skip_n=replace_with_none, fastq=(fmt == "fastq"), device=device) as p:
        p.feed(data)
        status = p.finish()
        if status == _native.KDB_E_LAYOUT:
            plain, pending = _plain_layout(bytes(data), fmt)
            p.close()
            with _native.Profiler(k, canonical=canonicalize, skip_n=replace_with_none, fastq=False, device=device) as q:
                q.feed(plain)
                status = q.finish()
                st = q.stats()
                if status not in (0, _native.KDB_E_EMPTY) or pending is None:
                    _raise_for_status(status, st)
                if pending is not None:
                    raise pending
                return Profile(k, q.counts(), st, q.nullomers() if want_nullomers else None, q.timing())
        st = p.stats()
        _raise_for_status(status, st)
        return Profile(k, p.counts(), st, p.nullomers() if want_nullomers else None, p.timing())


def profile_bgzf(raw, k, fmt="fasta", replace_with_none=True, canonicalize=True, device=0, want_nullomers=True):
    """BGZF input: the COMPRESSED bytes go to the device and are inflated there (a thread per member decodes, a warp per member resolves)
    (kdb_feed_bgzf_host), then tokenised and counted in place.  Returns None when `raw` is not BGZF, a member fails its
    CRC-32 / ISIZE, or the text needs the host tokeniser: the caller then takes the host gzip path, which raises what the
    reference's gzip.open raises."""
    if _native.bgzf_text_size(raw) is None:
        return None
    with _native.Profiler(k, canonical=canonicalize, skip_n=replace_with_none, fastq=(fmt == "fastq"), device=device) as p:
        if p.feed_bgzf(raw) != 0:
            return None
        status = p.finish()
        if status == _native.KDB_E_LAYOUT:
            return None
        st = p.stats()
        _raise_for_status(status, st)
        return Profile(k, p.counts(), st, p.nullomers() if want_nullomers else None, p.timing())


# ---- streaming input: bounded host memory ---------------------------------------------------------------
# The reference consumes its input as a stream (kmerdb/parse.py:63-72: gzip.open + SeqIO.parse).  Here the raw file is read in
# pieces (hashed for md5 / sha256 as they pass, kmerdb/util.py:35-50), gzip is inflated piece by piece on a host thread, the
# text is cut into fragments at record borders, and every fragment is one kdb_feed_host: the next piece is being read /
# inflated while the GPU works on the current fragment.  Host memory: a few pieces, whatever the file size.
STREAM_PIECE = 64 << 20


def _iter_raw_hashed(filepath, piece, hashes):
    with open(filepath, "rb", buffering=0) as fh:
        while True:
            b = fh.read(piece)
            if not b:
                return
            for h in hashes:
                h.update(b)
            yield b


def _iter_text_pieces(filepath, piece, hashes):
    """Uncompressed text of the file in pieces of about `piece` bytes.  gzip (any number of members, as gzip.open reads
    them) is inflated with zlib on this thread; errors are what gzip.open raises (BadGzipFile / EOFError / zlib.error)."""
    import zlib
    raw = _iter_raw_hashed(filepath, piece // 4, hashes)
    first = next(raw, b"")
    if not (first[:2] == b"\x1f\x8b" and is_gz_file(filepath)):
        if first:
            yield first
        yield from raw
        return
    d = zlib.decompressobj(31)
    pending = first
    while True:
        while pending:
            out = d.decompress(pending, piece)
            if out:
                yield out
            if d.eof:                                            # next member (concatenated gzip), or trailing garbage
                pending = d.unused_data
                if pending[:2] == b"\x1f\x8b":
                    d = zlib.decompressobj(31)
                    continue
                pending = pending.lstrip(b"\x00")
                if pending and pending[:2] != b"\x1f\x8b":
                    raise gzip.BadGzipFile("Not a gzipped file (%r)" % pending[:2])
                if pending:
                    d = zlib.decompressobj(31)
                    continue
                break
            pending = d.unconsumed_tail
        nxt = next(raw, None)
        if nxt is None:
            if not d.eof:
                raise EOFError("Compressed file ended before the end-of-stream marker was reached")
            return
        if d.eof:
            d = zlib.decompressobj(31)
        pending = nxt


def _iter_fragments(pieces, fmt):
    """Cut the text at record borders: a FASTA fragment ends before a line that starts with '>', a FASTQ fragment after a line
    whose index is a multiple of 4 (the only layout the device tokeniser takes natively).  What follows the last border of a
    piece is carried into the next fragment; the end of the file flushes it."""
    carry, carry_lines = [], 0
    for piece in pieces:
        if fmt == "fasta":
            cut = piece.rfind(b"\n>") + 1                       # 0 when there is no border in this piece
        else:
            total = carry_lines + piece.count(b"\n")
            m = total - total % 4
            cut = 0
            if m > carry_lines:
                pos = len(piece)
                for _ in range(total - m + 1):
                    pos = piece.rfind(b"\n", 0, pos)
                cut = pos + 1
                carry_lines = total - m
            else:
                carry_lines = total
        if cut:
            yield b"".join(carry + [piece[:cut]]) if carry else piece[:cut]
            carry = [piece[cut:]] if cut < len(piece) else []
        else:
            carry.append(piece)
    if carry:
        yield b"".join(carry)


def profile_stream(filepath, k, fmt, replace_with_none=True, canonicalize=True, device=0, want_nullomers=True, piece=None):
    """Profile a file of any size with bounded host memory.  Returns (Profile or None, sums) — None when the text needs the
    host tokeniser (the caller re-reads the file the slow way); sums() gives (md5, sha256) of the raw bytes."""
    import queue
    import threading
    piece = piece or STREAM_PIECE
    hashes = (hashlib.md5(), hashlib.sha256())
    q = queue.Queue(maxsize=2)
    stop = threading.Event()

    def produce():
        try:
            for frag in _iter_fragments(_iter_text_pieces(filepath, piece, hashes), fmt):
                while not stop.is_set():
                    try:
                        q.put(frag, timeout=0.1)
                        break
                    except queue.Full:
                        continue
                if stop.is_set():
                    return
            q.put(None)
        except BaseException as e:                               # surfaces on the consumer side, after the fragments before it
            q.put(e)

    t = threading.Thread(target=produce, daemon=True)
    t.start()

    def sums():
        t.join()
        return hashes[0].hexdigest(), hashes[1].hexdigest()
    try:
        with _native.Profiler(k, canonical=canonicalize, skip_n=replace_with_none, fastq=(fmt == "fastq"), device=device) as p:
            pending = None
            while True:
                item = q.get()
                if item is None:
                    break
                if isinstance(item, BaseException):
                    pending = item
                    break
                p.feed(item)
            if pending is not None and not isinstance(pending, ValueError):
                raise pending                                     # BadGzipFile / EOFError / zlib.error: what gzip.open raises
            status = p.finish()
            if status == _native.KDB_E_LAYOUT:
                return None, sums
            st = p.stats()
            if status not in (0, _native.KDB_E_EMPTY) or pending is None:
                _raise_for_status(status, st)
            if pending is not None:
                raise pending
            return Profile(k, p.counts(), st, p.nullomers() if want_nullomers else None, p.timing()), sums
    except BaseException:
        stop.set()
        raise


def _is_bgzf_file(filepath):
    with open(filepath, "rb") as fh:
        head = fh.read(18)
    return len(head) == 18 and head[:4] == b"\x1f\x8b\x08\x04" and b"BC\x02\x00" in head[12:18]


def iter_profile_multik(data, min_k, max_k, fmt="fasta", replace_with_none=True, canonicalize=True, device=0, want_nullomers=False):
    """`--minK/--maxK` sweep (kmerdb/__init__.py:1846-1855) over text held in host memory, yielding (k, Profile)
    for k = min_k .. max_k.  The reference re-parses the file for every k; here the text crosses PCIe once, is
    scanned and 2-bit encoded once (for min_k), and every larger k is counted from the same packed letter stream
    (kdb_attach).  Like the reference loop, the sweep raises at the first k that fails, after the smaller k."""
    a = data if isinstance(data, np.ndarray) else np.frombuffer(data, dtype=np.uint8)
    ks = list(range(min_k, max_k + 1))
    if not ks:
        return
    kw = dict(canonical=canonicalize, skip_n=replace_with_none, fastq=(fmt == "fastq"), device=device)
    if min_k < 2:                                          # the shared stream needs overlapping windows (k >= 2)
        yield min_k, profile_bytes(data, min_k, fmt, replace_with_none, canonicalize, device, want_nullomers)
        yield from iter_profile_multik(data, min_k + 1, max_k, fmt, replace_with_none, canonicalize, device, want_nullomers)
        return
    profs = [_native.Profiler(k, **kw) for k in ks]
    try:
        for f in profs[1:]:
            profs[0].attach(f)
        profs[0].feed(a)
        status = profs[0].finish()
        if status == _native.KDB_E_LAYOUT:                 # host tokeniser decides (same for every k)
            for p in profs:
                p.close()
            for k in ks:
                yield k, profile_bytes(data, k, fmt, replace_with_none, canonicalize, device, want_nullomers)
            return
        for k, p in zip(ks, profs):
            status = p.finish()
            st = p.stats()
            _raise_for_status(status, st)
            yield k, Profile(k, p.counts(), st, p.nullomers() if want_nullomers else None, p.timing())
    finally:
        for p in profs:
            p.close()


def profile_multik(data, min_k, max_k, **kw):
    """{k: Profile} of iter_profile_multik."""
    return dict(iter_profile_multik(data, min_k, max_k, **kw))


def parsefile(filepath, k, replace_with_none=True, canonicalize=True, device=0, with_mononucleotides=False):
    """Drop-in for kmerdb.parse.parsefile (kmerdb/parse.py:90-163).

    :returns: (counts uint64[4**k], file_metadata dict, nullomer_array uint64)"""
    counts, meta_fn, nullomer_array = parsefile_deferred(filepath, k, replace_with_none, canonicalize, device, with_mononucleotides)
    return counts, meta_fn(), nullomer_array


def parsefile_deferred(filepath, k, replace_with_none=True, canonicalize=True, device=0, with_mononucleotides=False):
    """parsefile whose file_metadata comes as a callable: the count vector is ready after the device job (tens of
    milliseconds), md5 / sha256 of a large file take the host seconds longer — the CLI formats and compresses the rows in
    the meantime and asks for the metadata when it writes the header."""
    if filepath is None or type(filepath) is not str:
        raise TypeError("kmerdb.parse.parsefile expects a str as its first positional argument")
    elif not os.path.exists(filepath):
        raise OSError("kmerdb.parse.parsefile could not find the file '{0}' on the filesystem".format(filepath))
    elif k is None or type(k) is not int:
        raise TypeError("kmerdb.parse.parsefile expects an int as its second positional argument")
    elif type(replace_with_none) is not bool:
        raise TypeError("kmerdb.parse.parsefile expects the keyword argument 'replace_with_none' to be a bool")
    fmt = _file_format(filepath)
    devices = _device_list(device)
    prof, sums = None, None
    if len(devices) == 1 and not _is_bgzf_file(filepath):
        # one GPU, plain or ordinary-gzip input: streamed with bounded host memory (pieces are read / inflated while the GPU works)
        prof, sums = profile_stream(filepath, k, fmt, replace_with_none, bool(canonicalize), devices[0])
        if prof is None:
            sums()
    if prof is None:
        raw, sums = _read_and_hash(filepath)                   # hashes of the raw bytes, computed beside the read and the GPU work
        try:
            if len(devices) == 1 and len(raw) >= 28 and bytes(raw[:2]) == b"\x1f\x8b" and is_gz_file(filepath):
                prof = profile_bgzf(raw, k, fmt, replace_with_none, bool(canonicalize), devices[0])       # BGZF: inflate on the GPU
            if prof is None:
                data = _read_text_bytes(filepath, raw)
                prof = profile_bytes(data, k, fmt, replace_with_none, bool(canonicalize), device)
        except BaseException:
            sums()                                             # (join the hashing threads before the raise leaves)
            raise
    assert prof.nullomers == len(prof.nullomer_array), "Internal Error: kmerdb.parse.parsefile() found inconsistencies between two ways of counting nullomers. Internal error."

    def file_metadata():
        md5, sha256 = sums()
        meta = prof.file_metadata(filepath, md5, sha256)
        if with_mononucleotides:
            meta["mononucleotides"] = prof.mononucleotides    # legacy header field (README.md:31)
        return meta
    return prof.counts, file_metadata, prof.nullomer_array
