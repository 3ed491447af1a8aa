#!/usr/bin/env python
"""Brief per-kernel view of an ncu --set full report: time, instructions, issue, stall reasons, memory.  usage: ncu_brief.py report.ncu-rep"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); h = rows[0]
for r in rows[2:]:
    d = dict(zip(h, r))
    f = lambda k: float(d.get(k, "0").replace(",", "") or 0)
    print(d["Kernel Name"].split("(")[0], "ms", round(f("gpu__time_duration.sum"), 3), "grid", d.get("launch__grid_size"), "regs", d.get("launch__registers_per_thread"))
    print("  warp inst %.1fM  thread/inst %.1f  issue active %.1f%%  warps active %.1f%%  lat/inst %.2f" % (
        f("smsp__inst_executed.sum") / 1e6, f("smsp__thread_inst_executed_per_inst_executed.ratio"), f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        f("sm__warps_active.avg.pct_of_peak_sustained_active"), f("smsp__average_warp_latency_per_inst_issued.ratio")))
    st = sorted(((float(v.replace(",", "") or 0), k.split("issue_stalled_")[1].split("_per_")[0]) for k, v in d.items()
                 if "average_warps_issue_stalled" in k and k.endswith("per_issue_active.ratio")), reverse=True)
    print("  stalls:", ", ".join("%s %.2f" % (k, v) for v, k in st[:7]))
    print("  lsu wavefronts %.1f%%  l1 hit %.1f%%  l2 hit %.1f%%  dram %.1f%%  occ limits: smem %s regs %s  waves %s" % (
        f("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"), f("l1tex__t_sector_hit_rate.pct"), f("lts__t_sector_hit_rate.pct"),
        f("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), d.get("launch__occupancy_limit_shared_mem"), d.get("launch__occupancy_limit_registers"), d.get("launch__waves_per_multiprocessor")))
    print("  local ld %.2fM st %.2fM  global ld sectors %.1fM st sectors %.1fM" % (f("smsp__sass_inst_executed_op_local_ld.sum") / 1e6, f("smsp__sass_inst_executed_op_local_st.sum") / 1e6,
          f("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum") / 1e6, f("l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum") / 1e6))
