#!/usr/bin/env python
"""Summarise an ncu --set full report into profiles/ncu_full_<tag>.json (one entry per kernel) and refresh
profiles/traffic_k3p_partition.json.   usage: python scripts/ncu_summary.py report.ncu-rep TAG [text_bytes_per_launch]"""
import csv
import json
import os
import subprocess
import sys

rep, tag = sys.argv[1], sys.argv[2]
text_bytes = int(sys.argv[3]) if len(sys.argv) > 3 else 324000000
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
out = {}
for r in rows[2:]:
    d = dict(zip(hdr, r))
    u = dict(zip(hdr, units))
    name = d["Kernel Name"].replace("void ", "").split("(")[0]
    e = {}
    for m in KEEP:
        if m in d and d[m] != "":
            v = float(d[m].replace(",", ""))
            if m.startswith("dram__bytes"):
                v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u[m], 1.0)      # MB
            if m == "gpu__time_duration.sum":
                v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(u[m], 1.0)   # us
            e[m] = v
    stalls = {}
    for k, v in d.items():
        if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and v:
            stalls[k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = float(v.replace(",", ""))
    tot = sum(stalls.values()) or 1.0
    e["stalls"] = {k: round(100 * v / tot, 1) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:5]}
    out.setdefault(name, e)
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
json.dump(out, open(os.path.join(root, "profiles", "ncu_full_%s.json" % tag), "w"), indent=1)
for name, e in out.items():
    print(name, json.dumps(e))
traffic = {}
for name, e in out.items():
    if "--no-traffic" in sys.argv or "dram__bytes_read.sum" not in e:
        continue
    rd, wr = e["dram__bytes_read.sum"] * 1e6, e["dram__bytes_write.sum"] * 1e6
    traffic[name.split("<")[0]] = {"kernel": name.replace(" ", ""), "text_bytes_per_launch": text_bytes, "dram_bytes_read": rd, "dram_bytes_write": wr,
                                   "dram_bytes_per_launch": rd + wr, "dram_bytes_per_text_byte": (rd + wr) / text_bytes}
if traffic:
    traffic["source"] = "ncu --set full --clock-control none, profiles/ncu_full_%s.json (%d bytes of text per launch)" % (tag, text_bytes)
    json.dump(traffic, open(os.path.join(root, "profiles", "traffic_%s.json" % tag), "w"), indent=1)
    print("traffic", {k: round(v["dram_bytes_per_text_byte"], 3) for k, v in traffic.items() if isinstance(v, dict)})
