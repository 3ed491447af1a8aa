#!/usr/bin/env python
"""Per-source-line breakdown of one kernel of an ncu report.  usage: ncu_lines.py report.ncu-rep KERNEL_REGEX [N] [samples|inst]"""
import csv, subprocess, sys
from collections import defaultdict
rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
by = sys.argv[4] if len(sys.argv) > 4 else "samples"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + rx], capture_output=True, text=True).stdout
agg = defaultdict(lambda: [0.0, 0.0, 0.0, ""])
hdr = None; fname = ""; cur = None
for r in csv.reader(out.splitlines()):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r; i_s, i_i, i_t = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    elif hdr and r[0].isdigit(): cur = (fname, int(r[0]), r[1])
    elif hdr and len(r) >= len(hdr) - 1 and r[2].startswith("0x"):
        a = agg[cur[:2]]; a[0] += float(r[i_s] or 0); a[1] += float(r[i_i] or 0); a[2] += float(r[i_t] or 0); a[3] = cur[2]
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print("samples %d, warp instructions %d" % (ts, ti))
idx = 0 if by == "samples" else 1
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][idx])[:top]:
    print("%-20s:%4d inst %5.1f%% samp %5.1f%% lanes %4.1f | %s" % (key[0], key[1], 100 * a[1] / ti, 100 * a[0] / max(ts, 1), a[2] / max(a[1], 1), a[3].strip()[:100]))
