"""Atomic-throughput microbenchmarks behind the roofline discussion in DESIGN.md.
Prints updates/s for random red.global.add.u32 into tables of several sizes and for random
shared-memory atomics."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kmerdb_b200 import _native

out = {}
for mb in (1, 16, 32, 64, 256, 1024):
    out["red_global_u32_%dMiB" % mb] = _native.microbench_atomics(mb << 20, 1 << 31, mode=0)
out["atoms_shared_64KiB"] = _native.microbench_atomics(0, 1 << 31, mode=1)
# distributed shared memory: random red.shared::cluster over the tables of a cluster of 8 / 16 CTAs
for name, mode in (("red_dsmem_cluster8", 2), ("red_dsmem_cluster16", 3)):
    try:
        out[name] = _native.microbench_atomics(0, 1 << 28, mode=mode)
    except Exception as e:      # (a cluster size the part refuses is a result too)
        out[name] = "failed: %s" % e
print(json.dumps(out))
