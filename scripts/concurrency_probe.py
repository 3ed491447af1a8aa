"""Probe: do two profile jobs on one GPU overlap (K2 is issue bound, K3p LSU bound)?  Two contexts, two host threads."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native
n = 10_000_000
rb = _native.synth_record_bytes()
buf = _native.DeviceBuffer(n * rb + 16)
_native.synth_device(buf, n)
profs = [_native.Profiler(12) for _ in range(2)]
def job(p, reps):
    for _ in range(reps):
        p.reset(); p.feed_device(buf.ptr, n * rb); assert p.finish() == 0
for p in profs: job(p, 2)
t0 = time.perf_counter(); job(profs[0], 6); t1 = time.perf_counter() - t0
ths = [threading.Thread(target=job, args=(p, 6)) for p in profs]
t0 = time.perf_counter(); [t.start() for t in ths]; [t.join() for t in ths]; t2 = time.perf_counter() - t0
print("one context: %.2f ms/job; two concurrent contexts: %.2f ms per job pair -> %.2f ms/job" % (1e3 * t1 / 6, 1e3 * t2 / 6, 1e3 * t2 / 12))
