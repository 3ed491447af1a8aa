"""TEST INFRASTRUCTURE: ctypes loader for the CPU oracle (oracle/kmerdb_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  The product (kmerdb_b200/) never does."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libkmerdb_oracle.so")

ERR_NAMES = {0: None, 1: "NameError", 2: "KeyError", 3: "AttributeError", 4: "ValueError", 5: "ValueError", 6: "TypeError"}


class Stats(ctypes.Structure):
    _fields_ = [("total_reads", ctypes.c_uint64), ("total_kmers", ctypes.c_uint64), ("unique_kmers", ctypes.c_uint64),
                ("sum_len", ctypes.c_uint64), ("min_len", ctypes.c_uint64), ("max_len", ctypes.c_uint64),
                ("mono", ctypes.c_uint64 * 4), ("err_record", ctypes.c_uint64), ("status", ctypes.c_int32), ("pad_", ctypes.c_int32)]


def build(force=False):
    src = os.path.join(_HERE, "kmerdb_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libkmerdb_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        u8p, u64p = ctypes.c_void_p, ctypes.c_void_p
        L.kdbo_kmer_to_id.restype = ctypes.c_int64
        L.kdbo_kmer_to_id.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int]
        L.kdbo_profile.restype = ctypes.c_int
        L.kdbo_profile.argtypes = [u8p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, u64p, ctypes.POINTER(Stats)]
        L.kdbo_profile_mt.restype = ctypes.c_int
        L.kdbo_profile_mt.argtypes = L.kdbo_profile.argtypes + [ctypes.c_int]
        L.kdbo_edges.restype = ctypes.c_int64
        L.kdbo_edges.argtypes = [u8p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int] + [u64p] * 5 + [ctypes.c_size_t]
        L.kdbo_minimizers_compat.restype = ctypes.c_int64
        L.kdbo_minimizers_compat.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, u64p, u8p, ctypes.c_size_t]
        L.kdbo_minimizers_true.restype = ctypes.c_int64
        L.kdbo_minimizers_true.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, u64p, u64p, ctypes.c_size_t]
        _lib = L
    return _lib


def _as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def profile(data, k, *, fastq=False, skip_n=True, canonical=True, threads=1):
    """Returns (status, counts uint64[4^k], dict stats).  status 0 = ok, else see ERR_NAMES."""
    a = _as_u8(data)
    counts = np.empty(4 ** k, dtype=np.uint64)
    st = Stats()
    ptr = a.ctypes.data if a.size else None
    if threads == 1:
        rc = lib().kdbo_profile(ptr, a.size, int(fastq), k, int(skip_n), int(canonical), counts.ctypes.data, ctypes.byref(st))
    else:
        rc = lib().kdbo_profile_mt(ptr, a.size, int(fastq), k, int(skip_n), int(canonical), counts.ctypes.data, ctypes.byref(st), threads)
    stats = {f: (list(getattr(st, f)) if f == "mono" else int(getattr(st, f))) for f, _ in Stats._fields_ if f != "pad_"}
    return rc, counts, stats


def metadata(stats, k):
    """parse.py:139-160 dictionary (without filename/md5/sha256) from oracle stats."""
    n = stats["total_reads"]
    return {"total_reads": n, "total_kmers": stats["total_kmers"], "unique_kmers": stats["unique_kmers"],
            "nullomers": 4 ** k - stats["unique_kmers"], "min_read_length": stats["min_len"], "max_read_length": stats["max_len"],
            "avg_read_length": int(np.mean(np.array([stats["sum_len"] / n]))) if n else 0}


def edges(data, k, *, fastq=False, canonical=True, cap=1 << 20):
    a = _as_u8(data)
    arrs = [np.zeros(cap, dtype=np.uint64) for _ in range(5)]
    n = lib().kdbo_edges(a.ctypes.data if a.size else None, a.size, int(fastq), k, int(canonical), *[x.ctypes.data for x in arrs], cap)
    if n < 0:
        return int(-n), None
    return 0, np.stack([x[:n] for x in arrs], axis=1)


def minimizers_compat(seq, k, window):
    cap = 4 ** k
    ids = np.zeros(cap, dtype=np.uint64)
    flags = np.zeros(cap, dtype=np.uint8)
    n = lib().kdbo_minimizers_compat(seq.encode(), len(seq), k, window, ids.ctypes.data, flags.ctypes.data, cap)
    if n < 0:
        return int(-n), None, None
    return 0, ids[:n], flags[:n]


def minimizers_true(seq, k, w):
    cap = max(len(seq), 1)
    pos = np.zeros(cap, dtype=np.uint64)
    ids = np.zeros(cap, dtype=np.uint64)
    n = lib().kdbo_minimizers_true(seq.encode(), len(seq), k, w, pos.ctypes.data, ids.ctypes.data, cap)
    if n < 0:
        return int(-n), None, None
    return 0, pos[:n], ids[:n]
