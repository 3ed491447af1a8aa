"""ctypes binding of libkmerdb_b200.so (include/kmerdb_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device can be used, the
calls raise.  Build with `python -m kmerdb_b200.build` (or `__graft_entry__.build()`)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libkmerdb_b200.so")

KDB_OK, KDB_E_ALPHABET, KDB_E_IUPAC, KDB_E_SHORT, KDB_E_FORMAT, KDB_E_EMPTY, KDB_E_ARG, KDB_E_LAYOUT, KDB_E_CUDA, KDB_E_STATE = range(10)
KDB_CANONICAL, KDB_SKIP_N, KDB_FASTQ, KDB_EDGE_KPLUS1 = 1, 2, 4, 8

EXPORTS = [
    "kdb_create", "kdb_destroy", "kdb_last_error", "kdb_feed_host", "kdb_feed_device", "kdb_finish", "kdb_stats",
    "kdb_counts", "kdb_nullomers", "kdb_nullomers_static", "kdb_device_table", "kdb_export_partials", "kdb_import_partials", "kdb_refinalize",
    "kdb_reset", "kdb_timing", "kdb_kmer_to_id_batch", "kdb_synth_record_bytes", "kdb_synth_device",
    "kdb_synth_host", "kdb_microbench_atomics", "kdb_dev_alloc", "kdb_dev_free", "kdb_host_alloc_pinned",
    "kdb_host_free_pinned", "kdb_memcpy_h2d", "kdb_memcpy_d2h", "kdb_nccl_unique_id", "kdb_reduce_nccl", "kdb_reduce_peers",
    "kdb_bgzf_index", "kdb_inflate_bgzf", "kdb_feed_bgzf_host", "kdb_last_kernel_ms",
    "kdb_edge_rows", "kdb_edge_weights", "kdb_minimizers_compat", "kdb_minimizers", "kdb_attach", "kdb_format_rows", "kdb_minimizers_batch",
    "kdb_reader_create", "kdb_reader_feed", "kdb_reader_finish", "kdb_reader_destroy", "kdb_correlation_matrix",
]


class Stats(ctypes.Structure):
    _fields_ = [("total_reads", ctypes.c_uint64), ("total_kmers", ctypes.c_uint64), ("unique_kmers", ctypes.c_uint64),
                ("sum_len", ctypes.c_uint64), ("min_len", ctypes.c_uint64), ("max_len", ctypes.c_uint64),
                ("mono", ctypes.c_uint64 * 4), ("err_record", ctypes.c_uint64), ("n_windows_with_n", ctypes.c_uint64),
                ("status", ctypes.c_int32), ("pad_", ctypes.c_int32)]


class Timing(ctypes.Structure):
    _fields_ = [("scan_ms", ctypes.c_float), ("encode_ms", ctypes.c_float), ("count_ms", ctypes.c_float), ("partition_ms", ctypes.c_float), ("fold_ms", ctypes.c_float),
                ("final_ms", ctypes.c_float), ("h2d_ms", ctypes.c_float), ("job_ms", ctypes.c_float), ("reduce_ms", ctypes.c_float), ("inflate_ms", ctypes.c_float),
                ("count_launches", ctypes.c_uint64),
                ("total_launches", ctypes.c_uint64), ("bytes_fed", ctypes.c_uint64)]


class NativeError(RuntimeError):
    pass


_lib = None


def lib():
    """Load the shared library (once).  Raises NativeError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise NativeError("kmerdb_b200: %s is missing — the CUDA extension was not built (python -m kmerdb_b200.build); "
                          "there is no CPU fallback" % SO_PATH)
    L = ctypes.CDLL(SO_PATH)
    vp, sz, u64, ci = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint64, ctypes.c_int
    L.kdb_create.restype = vp
    L.kdb_create.argtypes = [ci, ci, ci]
    L.kdb_destroy.restype = None
    L.kdb_destroy.argtypes = [vp]
    L.kdb_last_error.restype = ctypes.c_char_p
    L.kdb_last_error.argtypes = [vp]
    for name, args in [
        ("kdb_feed_host", [vp, vp, sz]), ("kdb_feed_device", [vp, vp, sz]), ("kdb_finish", [vp]),
        ("kdb_stats", [vp, ctypes.POINTER(Stats)]), ("kdb_counts", [vp, vp]),
        ("kdb_nullomers", [vp, vp, sz, ctypes.POINTER(sz)]),
        ("kdb_device_table", [vp, ctypes.POINTER(vp), ctypes.POINTER(sz)]),
        ("kdb_nullomers_static", [ci, vp, sz, vp, sz, vp, vp]),
        ("kdb_export_partials", [vp, vp]), ("kdb_import_partials", [vp, vp]), ("kdb_refinalize", [vp]),
        ("kdb_attach", [vp, vp]),
        ("kdb_format_rows", [ci, vp, u64, u64, vp, vp, vp, sz, vp, sz, ctypes.POINTER(sz)]), ("kdb_reset", [vp]), ("kdb_timing", [vp, ctypes.POINTER(Timing)]),
        ("kdb_kmer_to_id_batch", [ci, ctypes.c_char_p, sz, ci, ci, vp]),
        ("kdb_bgzf_index", [vp, sz, vp, sz, ctypes.POINTER(sz), ctypes.POINTER(u64)]), ("kdb_inflate_bgzf", [ci, vp, sz, vp, sz, ctypes.POINTER(sz)]),
        ("kdb_feed_bgzf_host", [vp, vp, sz]),
        ("kdb_last_kernel_ms", []),
        ("kdb_nccl_unique_id", [vp, sz]), ("kdb_reduce_nccl", [vp, vp, ci, ci]), ("kdb_reduce_peers", [vp, ci]),
        ("kdb_synth_device", [ci, vp, u64, u64, ci, ci, u64]), ("kdb_synth_host", [vp, u64, u64, ci, ci, u64]),
        ("kdb_microbench_atomics", [ci, sz, u64, ci, ctypes.POINTER(ctypes.c_double)]),
        ("kdb_dev_alloc", [ci, sz, ctypes.POINTER(vp)]), ("kdb_dev_free", [ci, vp]),
        ("kdb_host_alloc_pinned", [sz, ctypes.POINTER(vp)]), ("kdb_host_free_pinned", [vp]),
        ("kdb_memcpy_h2d", [ci, vp, vp, sz]), ("kdb_memcpy_d2h", [ci, vp, vp, sz]),
        ("kdb_edge_rows", [ci, vp, sz, vp, sz, ci, ci, vp, vp, vp, vp, vp, sz, ctypes.POINTER(sz), ctypes.POINTER(u64)]),
        ("kdb_edge_weights", [vp, ci, vp, vp, vp, sz, ctypes.POINTER(sz)]),
        ("kdb_minimizers_compat", [ci, vp, sz, ci, u64, vp, vp, sz, ctypes.POINTER(sz)]),
        ("kdb_minimizers", [ci, vp, sz, ci, ci, vp, vp, sz, ctypes.POINTER(sz)]),
        ("kdb_minimizers_batch", [ci, vp, sz, vp, sz, ci, ci, vp, vp, vp, sz, ctypes.POINTER(sz)]),
        ("kdb_correlation_matrix", [ci, vp, sz, sz, ci, vp]),
        ("kdb_reader_feed", [vp, vp, sz]), ("kdb_reader_finish", [vp, vp, vp, ctypes.POINTER(u64), ctypes.POINTER(u64)]),
    ]:
        fn = getattr(L, name)
        fn.restype = ci
        fn.argtypes = args
    L.kdb_last_kernel_ms.restype = ctypes.c_float
    L.kdb_reader_create.restype = vp
    L.kdb_reader_create.argtypes = [ci, ci]
    L.kdb_reader_destroy.restype = None
    L.kdb_reader_destroy.argtypes = [vp]
    L.kdb_synth_record_bytes.restype = sz
    L.kdb_synth_record_bytes.argtypes = [ci, ci]
    _lib = L
    return L


def _err(ctx=None):
    msg = lib().kdb_last_error(ctx)
    return msg.decode() if msg else ""


class DeviceBuffer:
    """A raw allocation in one GPU's HBM (for device-resident feeds)."""

    def __init__(self, nbytes, device=0):
        self.device, self.nbytes = device, int(nbytes)
        p = ctypes.c_void_p()
        rc = lib().kdb_dev_alloc(device, self.nbytes, ctypes.byref(p))
        if rc:
            raise NativeError("kdb_dev_alloc(%d bytes): %s" % (self.nbytes, _err()))
        self.ptr = p.value

    def upload(self, arr, offset=0):
        a = np.ascontiguousarray(arr, dtype=np.uint8)
        if lib().kdb_memcpy_h2d(self.device, self.ptr + offset, a.ctypes.data, a.size):
            raise NativeError("kdb_memcpy_h2d: " + _err())

    def download(self, nbytes=None, offset=0):
        n = self.nbytes - offset if nbytes is None else nbytes
        out = np.empty(n, dtype=np.uint8)
        if lib().kdb_memcpy_d2h(self.device, out.ctypes.data, self.ptr + offset, n):
            raise NativeError("kdb_memcpy_d2h: " + _err())
        return out

    def free(self):
        if self.ptr:
            lib().kdb_dev_free(self.device, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PinnedBuffer:
    """Page-locked host memory exposed as a numpy uint8 array (`.array`)."""

    def __init__(self, nbytes):
        p = ctypes.c_void_p()
        if lib().kdb_host_alloc_pinned(int(nbytes), ctypes.byref(p)):
            raise NativeError("kdb_host_alloc_pinned: " + _err())
        self.ptr, self.nbytes = p.value, int(nbytes)
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * self.nbytes).from_address(self.ptr))

    def free(self):
        if self.ptr:
            self.array = None
            lib().kdb_host_free_pinned(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Profiler:
    """One device context (kdb_ctx): feed text fragments, finish, read counts and reductions."""

    def __init__(self, k, canonical=True, skip_n=True, fastq=False, device=0, edge_kplus1=False):
        flags = (KDB_CANONICAL if canonical else 0) | (KDB_SKIP_N if skip_n else 0) | (KDB_FASTQ if fastq else 0) | (KDB_EDGE_KPLUS1 if edge_kplus1 else 0)
        self.k, self.device = int(k), device
        self._ctx = lib().kdb_create(device, self.k, flags)
        if not self._ctx:
            raise NativeError("kdb_create(device=%d, k=%d): %s — a CUDA device is required, there is no CPU fallback" % (device, k, _err()))

    def _check(self, rc, what):
        if rc in (KDB_E_CUDA, KDB_E_ARG, KDB_E_STATE):
            raise NativeError("%s failed (%d): %s" % (what, rc, _err(self._ctx)))
        return rc

    def feed(self, data):
        """data: bytes / bytearray / memoryview / numpy uint8 (host memory)."""
        a = data if isinstance(data, np.ndarray) else np.frombuffer(data, dtype=np.uint8)
        a = np.ascontiguousarray(a, dtype=np.uint8)
        if a.size:
            self._check(lib().kdb_feed_host(self._ctx, a.ctypes.data, a.size), "kdb_feed_host")

    def feed_device(self, ptr, nbytes):
        self._check(lib().kdb_feed_device(self._ctx, ptr, int(nbytes)), "kdb_feed_device")

    def finish(self):
        """Returns the status code (0, or the reference exception class code of the first failing record)."""
        return self._check(lib().kdb_finish(self._ctx), "kdb_finish")

    def attach(self, follower):
        """Multi-k sweep: `follower` (a Profiler with a larger k) counts from this context's packed stream."""
        self._check(lib().kdb_attach(self._ctx, follower._ctx), "kdb_attach")

    def feed_bgzf(self, raw):
        """raw: a whole BGZF file (bytes-like, host memory).  Returns KDB_OK, or KDB_E_FORMAT when it is not BGZF / a member
        is corrupt (the job is reset then: nothing is kept)."""
        a = raw if isinstance(raw, np.ndarray) else np.frombuffer(raw, dtype=np.uint8)
        a = np.ascontiguousarray(a, dtype=np.uint8)
        return self._check(lib().kdb_feed_bgzf_host(self._ctx, a.ctypes.data, a.size), "kdb_feed_bgzf_host")

    def reduce_nccl(self, unique_id, rank, nranks):
        """Collective (kdb_reduce_nccl): sum of the shard profiles; rank 0 ends up with the whole table."""
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        return self._check(lib().kdb_reduce_nccl(self._ctx, buf, rank, nranks), "kdb_reduce_nccl")

    def refinalize(self):
        return self._check(lib().kdb_refinalize(self._ctx), "kdb_refinalize")

    def reset(self):
        self._check(lib().kdb_reset(self._ctx), "kdb_reset")

    def stats(self):
        st = Stats()
        self._check(lib().kdb_stats(self._ctx, ctypes.byref(st)), "kdb_stats")
        d = {f: int(getattr(st, f)) for f, _ in Stats._fields_ if f not in ("mono", "pad_")}
        d["mono"] = [int(x) for x in st.mono]
        return d

    def timing(self):
        t = Timing()
        self._check(lib().kdb_timing(self._ctx, ctypes.byref(t)), "kdb_timing")
        return {f: (float(getattr(t, f)) if f.endswith("_ms") else int(getattr(t, f))) for f, _ in Timing._fields_}

    def counts(self, out=None):
        """uint64[4^k] histogram; `out` may be a caller-owned (e.g. pinned) uint64 array."""
        if out is None:
            out = np.empty(4 ** self.k, dtype=np.uint64)
        assert out.dtype == np.uint64 and out.size == 4 ** self.k and out.flags.c_contiguous
        self._check(lib().kdb_counts(self._ctx, out.ctypes.data), "kdb_counts")
        return out

    def nullomers(self):
        n = ctypes.c_size_t()
        self._check(lib().kdb_nullomers(self._ctx, None, 0, ctypes.byref(n)), "kdb_nullomers")
        out = np.empty(n.value, dtype=np.uint64)
        if n.value:
            self._check(lib().kdb_nullomers(self._ctx, out.ctypes.data, n.value, ctypes.byref(n)), "kdb_nullomers")
        return out

    def device_table(self):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        self._check(lib().kdb_device_table(self._ctx, ctypes.byref(p), ctypes.byref(n)), "kdb_device_table")
        return p.value, n.value

    def export_partials(self):
        out = np.zeros(16, dtype=np.uint64)
        self._check(lib().kdb_export_partials(self._ctx, out.ctypes.data), "kdb_export_partials")
        return out

    def import_partials(self, arr):
        a = np.ascontiguousarray(arr, dtype=np.uint64)
        self._check(lib().kdb_import_partials(self._ctx, a.ctypes.data), "kdb_import_partials")

    def edge_weights(self, canonical=True):
        """(n1, n2, weight) uint64 arrays from a finished (k+1)-mer context (kdb_edge_weights)."""
        n = ctypes.c_size_t()
        self._check(lib().kdb_edge_weights(self._ctx, int(canonical), None, None, None, 0, ctypes.byref(n)), "kdb_edge_weights")
        out = [np.empty(n.value, dtype=np.uint64) for _ in range(3)]
        if n.value:
            self._check(lib().kdb_edge_weights(self._ctx, int(canonical), out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data,
                                               n.value, ctypes.byref(n)), "kdb_edge_weights")
        return out

    def close(self):
        if getattr(self, "_ctx", None):
            lib().kdb_destroy(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def synth_host(n_reads, read_len=150, fastq=False, seed=20261104, first_read=0):
    rb = lib().kdb_synth_record_bytes(read_len, int(fastq))
    out = np.empty(n_reads * rb, dtype=np.uint8)
    if lib().kdb_synth_host(out.ctypes.data, first_read, n_reads, read_len, int(fastq), seed):
        raise NativeError("kdb_synth_host failed")
    return out


def synth_device(buf, n_reads, read_len=150, fastq=False, seed=20261104, first_read=0, offset=0):
    if lib().kdb_synth_device(buf.device, buf.ptr + offset, first_read, n_reads, read_len, int(fastq), seed):
        raise NativeError("kdb_synth_device: " + _err())


def synth_record_bytes(read_len=150, fastq=False):
    return int(lib().kdb_synth_record_bytes(read_len, int(fastq)))


def microbench_atomics(table_bytes, n_updates, mode=0, device=0):
    v = ctypes.c_double()
    if lib().kdb_microbench_atomics(device, table_bytes, n_updates, mode, ctypes.byref(v)):
        raise NativeError("kdb_microbench_atomics: " + _err())
    return v.value


def kmer_to_id_batch(kmers, k, canonical=True, device=0):
    """kmers: bytes of n*k letters.  Returns uint64 ids (2^64-1 = None, 2^64-2 = KeyError letters)."""
    n = len(kmers) // k
    out = np.empty(n, dtype=np.uint64)
    if lib().kdb_kmer_to_id_batch(device, kmers, n, k, int(canonical), out.ctypes.data):
        raise NativeError("kdb_kmer_to_id_batch: " + _err())
    return out


def edge_rows(letters, rec_off, k, canonical=True, device=0):
    """kdb_edge_rows: returns (status, err_record, rows) with rows = uint64[n_edges, 5] (rec, pos1, id1, pos2, id2)."""
    a = np.ascontiguousarray(np.frombuffer(letters, dtype=np.uint8) if not isinstance(letters, np.ndarray) else letters, dtype=np.uint8)
    off = np.ascontiguousarray(rec_off, dtype=np.uint64)
    n_recs = off.size - 1
    n, err = ctypes.c_size_t(), ctypes.c_uint64()
    args = (device, a.ctypes.data if a.size else None, a.size, off.ctypes.data, n_recs, int(k), int(canonical))
    rc = lib().kdb_edge_rows(*args, None, None, None, None, None, 0, ctypes.byref(n), ctypes.byref(err))
    if rc in (KDB_E_CUDA, KDB_E_ARG):
        raise NativeError("kdb_edge_rows failed (%d): %s" % (rc, _err()))
    if rc:
        return rc, int(err.value), None
    cols = [np.empty(n.value, dtype=np.uint64) for _ in range(5)]
    if n.value:
        rc = lib().kdb_edge_rows(*args, *[c.ctypes.data for c in cols], n.value, ctypes.byref(n), ctypes.byref(err))
        if rc:
            raise NativeError("kdb_edge_rows failed (%d): %s" % (rc, _err()))
    return 0, 0, np.stack(cols, axis=1) if n.value else np.zeros((0, 5), dtype=np.uint64)


def minimizers_compat(seq, k, window, device=0):
    """kdb_minimizers_compat: (status, ids uint64[L], is_min uint8[L])."""
    a = np.frombuffer(seq if isinstance(seq, (bytes, bytearray)) else seq.encode("ascii"), dtype=np.uint8)
    n = ctypes.c_size_t()
    rc = lib().kdb_minimizers_compat(device, a.ctypes.data, a.size, int(k), int(window), None, None, 0, ctypes.byref(n))
    if rc in (KDB_E_CUDA, KDB_E_ARG):
        raise NativeError("kdb_minimizers_compat failed (%d): %s" % (rc, _err()))
    if rc:
        return rc, None, None
    ids, flags = np.empty(n.value, dtype=np.uint64), np.empty(n.value, dtype=np.uint8)
    rc = lib().kdb_minimizers_compat(device, a.ctypes.data, a.size, int(k), int(window), ids.ctypes.data, flags.ctypes.data, n.value, ctypes.byref(n))
    if rc in (KDB_E_CUDA, KDB_E_ARG):
        raise NativeError("kdb_minimizers_compat failed (%d): %s" % (rc, _err()))
    return rc, ids, flags


def minimizers(seq, k, w, device=0):
    """kdb_minimizers: (status, positions uint64, ids uint64) of the true windowed minimizers."""
    a = np.frombuffer(seq if isinstance(seq, (bytes, bytearray)) else seq.encode("ascii"), dtype=np.uint8)
    cap = max(a.size, 1)
    pos, ids = np.empty(cap, dtype=np.uint64), np.empty(cap, dtype=np.uint64)
    n = ctypes.c_size_t()
    rc = lib().kdb_minimizers(device, a.ctypes.data, a.size, int(k), int(w), pos.ctypes.data, ids.ctypes.data, cap, ctypes.byref(n))
    if rc in (KDB_E_CUDA, KDB_E_ARG):
        raise NativeError("kdb_minimizers failed (%d): %s" % (rc, _err()))
    return rc, pos[:n.value], ids[:n.value]


def format_rows(counts, first_row, uniq_vals, str_off, str_bytes, device=0):
    """kdb_format_rows: the .kdb body text (uint8 array) of rows first_row .. first_row + len(counts) - 1."""
    c = np.ascontiguousarray(counts, dtype=np.uint64)
    u = np.ascontiguousarray(uniq_vals, dtype=np.uint64)
    so = np.ascontiguousarray(str_off, dtype=np.uint32)
    sb = np.frombuffer(str_bytes, dtype=np.uint8)
    n = ctypes.c_size_t()
    args = (device, c.ctypes.data, int(first_row), c.size, u.ctypes.data, so.ctypes.data, sb.ctypes.data, u.size)
    if lib().kdb_format_rows(*args, None, 0, ctypes.byref(n)):
        raise NativeError("kdb_format_rows: " + _err())
    out = np.empty(n.value, dtype=np.uint8)
    if n.value and lib().kdb_format_rows(*args, out.ctypes.data, out.size, ctypes.byref(n)):
        raise NativeError("kdb_format_rows: " + _err())
    return out


def minimizers_batch(letters, rec_off, k, w, device=0, count_only=False):
    """kdb_minimizers_batch: (status, rows uint64[n, 3] = (record, position, id)) or (status, n_rows) when count_only."""
    a = np.ascontiguousarray(np.frombuffer(letters, dtype=np.uint8) if not isinstance(letters, np.ndarray) else letters, dtype=np.uint8)
    off = np.ascontiguousarray(rec_off, dtype=np.uint64)
    n = ctypes.c_size_t()
    args = (device, a.ctypes.data if a.size else None, a.size, off.ctypes.data, off.size - 1, int(k), int(w))
    rc = lib().kdb_minimizers_batch(*args, None, None, None, 0, ctypes.byref(n))
    if rc in (KDB_E_CUDA, KDB_E_ARG):
        raise NativeError("kdb_minimizers_batch failed (%d): %s" % (rc, _err()))
    if rc or count_only:
        return rc, (n.value if count_only else None)
    cols = [np.empty(n.value, dtype=np.uint64) for _ in range(3)]
    if n.value:
        rc = lib().kdb_minimizers_batch(*args, *[c.ctypes.data for c in cols], n.value, ctypes.byref(n))
        if rc:
            raise NativeError("kdb_minimizers_batch failed (%d): %s" % (rc, _err()))
    return 0, (np.stack(cols, axis=1) if n.value else np.zeros((0, 3), dtype=np.uint64))


KDB_R_COLUMNS, KDB_R_INT, KDB_R_INDEX, KDB_R_RANGE, KDB_R_DUP = 33, 34, 35, 36, 37


class RowReader:
    """kdb_reader: the rows of a .kdb body -> (counts, kmer_ids), parsed on the GPU."""

    def __init__(self, k, device=0):
        self.k = int(k)
        self._r = lib().kdb_reader_create(int(device), self.k)
        if not self._r:
            raise NativeError("kdb_reader_create: " + _err())

    def feed(self, text):
        """`text`: bytes-like holding whole rows, in file order."""
        a = np.frombuffer(text, dtype=np.uint8)
        if a.size and lib().kdb_reader_feed(self._r, a.ctypes.data, a.size):
            raise NativeError("kdb_reader_feed: " + _err())

    def finish(self, want_ids=True):
        """(status, counts, kmer_ids, n_rows, err_row); the arrays are None unless status == 0."""
        n = 4 ** self.k
        counts = np.empty(n, dtype=np.uint64)
        ids = np.empty(n, dtype=np.uint64) if want_ids else None
        n_rows, err_row = ctypes.c_uint64(), ctypes.c_uint64()
        rc = lib().kdb_reader_finish(self._r, counts.ctypes.data, ids.ctypes.data if want_ids else None, ctypes.byref(n_rows), ctypes.byref(err_row))
        if rc in (KDB_E_CUDA, KDB_E_ARG):
            raise NativeError("kdb_reader_finish failed (%d): %s" % (rc, _err()))
        if rc:
            return rc, None, None, n_rows.value, err_row.value
        return 0, counts, ids, n_rows.value, None

    def close(self):
        if self._r:
            lib().kdb_reader_destroy(self._r)
            self._r = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def correlation_matrix(profiles, mode=0, device=0):
    """kdb_correlation_matrix: float64[n, n] Pearson coefficients of the rows of `profiles` (uint64[n, 4^k])."""
    x = np.ascontiguousarray(profiles, dtype=np.uint64)
    if x.ndim != 2:
        raise ValueError("correlation_matrix expects a 2-d array, one profile per row")
    out = np.empty((x.shape[0], x.shape[0]), dtype=np.float64)
    if lib().kdb_correlation_matrix(int(device), x.ctypes.data, x.shape[0], x.shape[1], int(mode), out.ctypes.data):
        raise NativeError("kdb_correlation_matrix: " + _err())
    return out


def nccl_unique_id():
    """128 bytes for kdb_reduce_nccl; made on one rank, handed to the others by the caller."""
    buf = ctypes.create_string_buffer(128)
    if lib().kdb_nccl_unique_id(buf, 128):
        raise NativeError("kdb_nccl_unique_id: " + _err())
    return buf.raw


def reduce_peers(profilers):
    """One process, one finished Profiler per GPU (shard order): profilers[0] ends up with the whole table (kdb_reduce_peers)."""
    arr = (ctypes.c_void_p * len(profilers))(*[p._ctx for p in profilers])
    return profilers[0]._check(lib().kdb_reduce_peers(arr, len(profilers)), "kdb_reduce_peers")


def bgzf_text_size(raw):
    """Sum of ISIZE over the members of a BGZF buffer, or None when `raw` is not BGZF (host header walk only)."""
    a = np.frombuffer(raw, dtype=np.uint8)
    nm, text = ctypes.c_size_t(0), ctypes.c_uint64(0)
    if lib().kdb_bgzf_index(a.ctypes.data, a.size, None, 0, ctypes.byref(nm), ctypes.byref(text)):
        return None
    return int(text.value)


def inflate_bgzf(raw, device=0):
    """The text of a BGZF buffer, inflated on the GPU (every member checked against ISIZE and CRC-32)."""
    a = np.frombuffer(raw, dtype=np.uint8)
    n = ctypes.c_size_t(0)
    rc = lib().kdb_inflate_bgzf(device, a.ctypes.data, a.size, None, 0, ctypes.byref(n))
    if rc:
        raise ValueError("not a BGZF file")
    out = np.empty(n.value, dtype=np.uint8)
    rc = lib().kdb_inflate_bgzf(device, a.ctypes.data, a.size, out.ctypes.data, out.size, ctypes.byref(n))
    if rc == KDB_E_FORMAT:
        raise ValueError("BGZF: " + _err())
    if rc:
        raise NativeError("kdb_inflate_bgzf failed (%d): %s" % (rc, _err()))
    return out
