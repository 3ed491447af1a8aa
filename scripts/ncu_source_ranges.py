"""Aggregate `ncu --page source` of one kernel by source file + line, print top lines and cumulative per file.
usage: python scripts/ncu_source_ranges.py report.ncu-rep kernel_regex [N]"""
import csv, subprocess, sys
from collections import defaultdict
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", "regex:" + kern], capture_output=True, text=True).stdout
agg = defaultdict(lambda: [0.0, 0.0, 0.0, ""]); fname = ""; hdr = None; cur = None
for r in csv.reader(out.splitlines()):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r; i_s, i_i, i_t = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    elif hdr and r[0].isdigit(): cur = (fname, int(r[0]), r[1])
    elif hdr and cur and len(r) >= len(hdr) - 1 and r[2].startswith("0x"):
        a = agg[cur[:2]]; a[0] += float(r[i_s] or 0); a[1] += float(r[i_i] or 0); a[2] += float(r[i_t] or 0); a[3] = cur[2]
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print("kernel %s: samples %d, warp instructions %d" % (kern, ts, ti))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-16s:%4d inst %5.1f%% samp %5.1f%% lanes %4.1f | %s" % (key[0][:16], key[1], 100 * a[1] / ti, 100 * a[0] / max(ts, 1), a[2] / max(a[1], 1), a[3].strip()[:100]))
if len(sys.argv) > 4:
    # ranges "file:lo-hi,..." summary
    for spec in sys.argv[4].split(","):
        f, rng = spec.split(":"); lo, hi = map(int, rng.split("-"))
        s = sum(a[1] for k, a in agg.items() if k[0].startswith(f) and lo <= k[1] <= hi); sm = sum(a[0] for k, a in agg.items() if k[0].startswith(f) and lo <= k[1] <= hi)
        print("range %-28s inst %5.1f%% samp %5.1f%%" % (spec, 100 * s / ti, 100 * sm / max(ts, 1)))
