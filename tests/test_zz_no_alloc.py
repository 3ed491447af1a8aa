"""kdb_nullomers_static: the entry point that allocates nothing (include/kmerdb_b200.h; no counterpart in the reference).
It must return exactly what kdb_nullomers returns — the ascending ids with zero count
(kmerdb/parse.py:139-140) — on two calls in a row (the static scratch keeps the last call's contents), for tables of different
sizes one after the other, with a capacity smaller than the result, and it must refuse tables beyond its stated maximum.
Collected last on purpose (test_zz_*)."""
import ctypes

import numpy as np
import pytest

from kmerdb_b200 import _native

pytestmark = pytest.mark.gpu


def _static_nullomers(dev_counts, n_bins, cap, scratch):
    L = _native.lib()
    d_ids, d_n = scratch
    rc = L.kdb_nullomers_static(0, dev_counts, n_bins, d_ids if cap else None, cap, d_n, None)
    assert rc == _native.KDB_OK, L.kdb_last_error(None)
    n = np.zeros(1, dtype=np.uint64)
    assert L.kdb_memcpy_d2h(0, n.ctypes.data, d_n, 8) == 0          # (cudaMemcpy on the default stream: ordered after the kernels)
    m = min(cap, int(n[0]))
    ids = np.empty(m, dtype=np.uint64)
    if m:
        assert L.kdb_memcpy_d2h(0, ids.ctypes.data, d_ids, 8 * m) == 0
    return int(n[0]), ids


def test_static_nullomers_equal_kdb_nullomers_twice_in_a_row():
    L = _native.lib()
    max_bins = 4 ** 10
    d_ids, d_n = ctypes.c_void_p(), ctypes.c_void_p()
    assert L.kdb_dev_alloc(0, 8 * max_bins, ctypes.byref(d_ids)) == 0       # the CALLER's buffers (allocated before, not by the call)
    assert L.kdb_dev_alloc(0, 8, ctypes.byref(d_n)) == 0
    try:
        # a sparse k = 10 table (most bins zero), a k = 8 table (smaller: stale offsets of the call before lie behind it), then
        # a dense k = 10 table; every one twice in a row
        for k, reads, seed in ((10, 300, 7), (8, 40, 8), (10, 20000, 9), (9, 1, 10)):
            with _native.Profiler(k) as p:
                p.feed(_native.synth_host(reads, seed=seed))
                assert p.finish() == 0
                want = p.nullomers()
                dev, n_bins = p.device_table()
                assert n_bins == 4 ** k
                for _ in range(2):
                    n, ids = _static_nullomers(dev, n_bins, n_bins, (d_ids, d_n))
                    assert n == want.size
                    assert np.array_equal(ids, want)                         # integer work: bit-exact
                # capacity below the result: the count is still whole, the first `cap` ids are written
                cap = max(1, want.size // 3)
                n, ids = _static_nullomers(dev, n_bins, cap, (d_ids, d_n))
                assert n == want.size and np.array_equal(ids, want[:cap])
                n, ids = _static_nullomers(dev, n_bins, 0, (d_ids, d_n))     # count only
                assert n == want.size and ids.size == 0
    finally:
        L.kdb_dev_free(0, d_ids)
        L.kdb_dev_free(0, d_n)


def test_static_nullomers_refuses_sizes_beyond_its_scratch():
    L = _native.lib()
    d = ctypes.c_void_p()
    assert L.kdb_dev_alloc(0, 64, ctypes.byref(d)) == 0
    try:
        rc = L.kdb_nullomers_static(0, d, 4 ** 14 + 1, d, 1, d, None)
        assert rc == _native.KDB_E_ARG
        assert b"4^14" in L.kdb_last_error(None)
        assert L.kdb_nullomers_static(0, d, 0, d, 1, d, None) == _native.KDB_E_ARG
        assert L.kdb_nullomers_static(0, None, 16, d, 1, d, None) == _native.KDB_E_ARG
    finally:
        L.kdb_dev_free(0, d)
