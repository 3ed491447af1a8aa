#!/usr/bin/env python
"""bench.py — bases/s of the k=12 `profile` hot path on synthetic 150 bp reads.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

A "step" is one whole profile job over one batch of synthetic input: zero the tables, K1 record
scan, K2+K3 encode/count with the fused reductions, fold, unique count (and, for N>1, the NCCL
sum-reduce of the 4^k table to rank 0 plus the re-finalisation there).
  value : whole-job bases/s with the text already resident in HBM (device-timed)
  e2e   : the same job through the host-buffer C-ABI call (pinned host text -> kdb_feed_host ->
          kdb_finish -> counts + stats back on the host), host<->device copies inside the timed region
One JSON line on stdout (rank 0)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "profile_k12_bases_per_sec"
UNIT = "bases/s"
SEED = 20261104
READ_LEN = 150


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_oracle(text, k, threads):
    import oracle
    t0 = time.perf_counter()
    rc, _counts, st = oracle.profile(text, k, threads=threads)
    dt = time.perf_counter() - t0
    assert rc == 0
    return dt, st


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path (oracle port: the reference is
    pure Python with no compilable sources) on all host threads, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from kmerdb_b200 import _native
    threads = host_threads()
    n_reads = args.cpu_reads
    text = _native.synth_host(n_reads, READ_LEN, seed=SEED)
    for _ in range(args.warmup):
        time_oracle(text, args.k, threads)
    times = [time_oracle(text, args.k, threads)[0] for _ in range(args.steps)]
    dt = sum(times)
    value = n_reads * READ_LEN * args.steps / dt
    sample = "first %d of the synthetic %d bp reads (FASTA, seed %d) per step" % (n_reads, READ_LEN, SEED)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": workload_config(args, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args, n_gpus):
    return {"workload": "profile k=%d canonical, %d synthetic %d bp reads per GPU, FASTA framing 162 B/record" % (args.k, args.reads, READ_LEN),
            "k": args.k, "reads_per_gpu": args.reads, "read_len": READ_LEN, "canonical": True, "format": "fasta", "seed": SEED,
            "l2": "input (%.2f GB per GPU) is larger than L2; no flush needed" % (args.reads * 162 / 1e9),
            "parallelism": "reads sharded over %d GPU(s); partial 4^k tables sum-reduced to rank 0 with NCCL" % n_gpus if n_gpus > 1 else "1 GPU"}
