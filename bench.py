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
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
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


_REF_PARSE = None


def reference_python():
    """The UNMODIFIED reference package (oracle/_ref, put there by oracle/make_ref.sh; it travels to the GPU box with the
    snapshot) imported against real Biopython when the box has it, else against the test-only shim tests/_bio_shim.
    Returns kmerdb.parse or None when oracle/_ref is absent."""
    global _REF_PARSE
    if _REF_PARSE is None:
        ref = os.path.join(ROOT, "oracle", "_ref")
        if not os.path.isfile(os.path.join(ref, "kmerdb", "parse.py")):
            return None
        import logging
        logging.disable(logging.CRITICAL)
        try:
            import Bio  # noqa: F401
        except Exception:
            sys.path.insert(0, os.path.join(ROOT, "tests", "_bio_shim"))
        sys.path.insert(0, ref)
        from kmerdb import parse
        _REF_PARSE = parse
    return _REF_PARSE


def time_reference_python(text, k):
    """One call of the reference's parse.parsefile (kmerdb/parse.py:90) on `text` written to a FASTA file; 1 core."""
    import tempfile
    parse = reference_python()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "sample.fa")
        with open(path, "wb") as fh:
            fh.write(text)
        t0 = time.perf_counter()
        counts, meta, _nul = parse.parsefile(path, k, replace_with_none=True, canonicalize=True)
        dt = time.perf_counter() - t0
    return dt, {"total_kmers": int(meta["total_kmers"]), "unique_kmers": int(meta["unique_kmers"]), "total_reads": int(meta["total_reads"])}


def cpu_legs(k, ref_reads, port_reads):
    """The two CPU baselines on bounded prefixes of the bench's own synthetic stream (oracle/synth.py restates the
    generator in NumPy: nothing of the product is loaded): the reference's Python on 1 core (with the fixed 4^k cost of
    kmerdb/parse.py:139-140 measured on a one-read file and shown beside it) and the C port on all host threads."""
    from oracle import synth
    out = {}
    threads = host_threads()
    if reference_python() is not None:
        dt0, _ = time_reference_python(synth.fasta(0, 1, READ_LEN, SEED), k)
        dt, st = time_reference_python(synth.fasta(0, ref_reads, READ_LEN, SEED), k)
        out["cpu_baseline"] = {"value": ref_reads * READ_LEN / dt, "unit": UNIT, "cores": 1, "kind": "reference",
                               "sample": "first %d reads of the same stream through the unmodified kmerdb.parse.parsefile (oracle/_ref, Biopython %s), %.1f s"
                                         % (ref_reads, "shim tests/_bio_shim" if "tests/_bio_shim" in " ".join(sys.path) else "installed", dt),
                               "fixed_4k_setup_s": dt0, "value_without_fixed_setup": ref_reads * READ_LEN / max(dt - dt0, 1e-9), "result": st}
    text = synth.fasta(0, port_reads, READ_LEN, SEED)
    dtp, _ = time_oracle(text, k, threads)
    port = {"value": port_reads * READ_LEN / dtp, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "first %d reads of the same stream, oracle/kmerdb_oracle.c on %d threads, %.1f s" % (port_reads, threads, dtp)}
    if "cpu_baseline" in out:
        out["cpu_baseline_port"] = port
    else:
        out["cpu_baseline"] = port
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path — the unmodified Python of kmerdb v0.9.6
    (kmerdb.parse.parsefile, single process, 1 core: all it can use) — on a bounded prefix of the workload per step.
    The C port on all host threads is reported beside it (cpu_baseline_port).  Falls back to the port alone only when
    oracle/_ref is missing."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import synth
    have_ref = reference_python() is not None
    n_reads = args.ref_reads if have_ref else args.cpu_reads
    text = synth.fasta(0, n_reads, READ_LEN, SEED)
    threads = host_threads()

    def step():
        return time_reference_python(text, args.k)[0] if have_ref else time_oracle(text, args.k, threads)[0]
    for _ in range(args.warmup):
        step()
    times = [step() for _ in range(args.steps)]
    dt = sum(times)
    value = n_reads * READ_LEN * args.steps / dt
    legs = cpu_legs(args.k, args.ref_reads, args.cpu_reads)
    base = dict(legs["cpu_baseline"])
    base["value"] = value
    base["sample"] = "first %d of the synthetic %d bp reads (FASTA, seed %d) per step; %s" % (n_reads, READ_LEN, SEED, base["sample"])
    cfg = workload_config(args, args.gpus)
    cfg["timed_sample"] = "%d reads (%d bases) per step: a bounded prefix of the workload, not the %d reads of the GPU arm" % (n_reads, n_reads * READ_LEN, cfg["reads_total"])
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": cfg, "cpu_baseline": base,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if "cpu_baseline_port" in legs:
        line["cpu_baseline_port"] = legs["cpu_baseline_port"]
    _emit(line)


def workload_config(args, n_gpus):
    total = getattr(args, "reads_total", None) or args.reads or (10_000_000 if n_gpus == 1 else 100_000_000)
    per = getattr(args, "reads_per_gpu", None) or total // n_gpus
    name = "configs[1]" if n_gpus == 1 else "configs[2]"
    return {"workload": "%s: profile k=%d canonical, %d synthetic %d bp reads%s, FASTA framing 162 B/record"
                        % (name, args.k, total, READ_LEN, "" if n_gpus == 1 else " sharded over %d GPUs" % n_gpus),
            "k": args.k, "reads_total": total, "reads_per_gpu": per, "read_len": READ_LEN, "canonical": True, "format": "fasta", "seed": SEED,
            "l2": "input (%.2f GB per GPU) is larger than L2; no flush needed" % (per * 162 / 1e9),
            "parallelism": ("reads sharded over %d GPUs (rank r = shard r); partial 4^k tables summed by kdb_reduce_nccl "
                            "(reduce-scatter + per-slice finalize + slices to rank 0)" % n_gpus) if n_gpus > 1 else "1 GPU"}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from kmerdb_b200 import _native
    from kmerdb_b200 import distributed as kd

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the profile path has no CPU fallback")
    torch.cuda.set_device(local)
    numa_cpus = kd.bind_to_gpu_numa(local) if world > 1 else None      # pinned buffers on the GPU's NUMA node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    k = args.k
    # N = 1: configs[1] (10 M reads).  N > 1: configs[2] as written — 100 M reads sharded N ways (rank r holds reads
    # [r * R / N, (r + 1) * R / N) of the same stream), one reduce of the 4^12 tables per step.
    total_reads = args.reads if args.reads else (10_000_000 if world == 1 else 100_000_000)
    first_read = total_reads * rank // world
    n_reads = total_reads * (rank + 1) // world - first_read
    args.reads_total, args.reads_per_gpu = total_reads, n_reads
    rb = _native.synth_record_bytes(READ_LEN)
    nbytes = n_reads * rb
    n_bases = n_reads * READ_LEN
    dev_text = _native.DeviceBuffer(nbytes + 16, device=local)
    _native.synth_device(dev_text, n_reads, READ_LEN, seed=SEED, first_read=first_read)
    prof = _native.Profiler(k, canonical=True, skip_n=True, fastq=False, device=local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_to_rank0():
        """kdb_reduce_nccl (kmerdb_b200/csrc/kdb_multi.cuh): reduce-scatter of the 4^k uint64 partials, per-slice non-zero
        count, one all-gather of the scalar packs folded on the device, slices sent to rank 0 — all queued on the context's
        stream, nothing read back here."""
        if world == 1:
            return
        kd.reduce_profiler(prof, dist, torch.device("cuda", local), dst=0)

    def step_device():
        prof.reset()
        prof.feed_device(dev_text.ptr, nbytes)
        status = prof.finish()
        assert status == 0, "device status %d" % status
        reduce_to_rank0()
        return prof.timing()

    # ---- value: device-resident input ------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()
    job_ms = count_ms = scan_ms = fold_ms = final_ms = encode_ms = part_ms = reduce_ms = 0.0
    count_launches = launches = 0
    for _ in range(args.steps):
        t = step_device()
        reduce_ms += t["reduce_ms"]; job_ms += t["job_ms"] + t["reduce_ms"]; count_ms += t["count_ms"]; part_ms += t["partition_ms"]; encode_ms += t["encode_ms"]; scan_ms += t["scan_ms"]; fold_ms += t["fold_ms"]; final_ms += t["final_ms"]
        count_launches += t["count_launches"]; launches += t["total_launches"]
    ev1.record()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    # ONE clock for every N: CUDA events on the context's own stream — job (kdb_reset .. end of kdb_finish) plus, for
    # N > 1, the reduce (first collective .. rank 0 holds the sum), max over ranks.  The wall clock between the two
    # barriers (device time plus launch gaps) is reported beside it.
    tt = torch.tensor([wall, job_ms * 1e-3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    wall_max, job_max = tt.tolist()
    stats = prof.stats()

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region ----------------------
    pin_text = _native.PinnedBuffer(nbytes)
    _native.lib().kdb_memcpy_d2h(local, pin_text.ptr, dev_text.ptr, nbytes)
    pin_counts = _native.PinnedBuffer(8 * 4 ** k)
    counts_view = pin_counts.array.view(np.uint64)

    def step_host():
        prof.reset()
        prof.feed(pin_text.array)
        status = prof.finish()
        assert status == 0
        reduce_to_rank0()
        if rank == 0:
            prof.counts(out=counts_view)
            return prof.stats()
        return None

    for _ in range(max(1, args.warmup // 2)):
        step_host()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(e2e_steps):
        st_host = step_host()
    barrier()
    e2e_wall = time.perf_counter() - t0
    te = torch.tensor([e2e_wall], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_wall = te.item()
    e2e_sum = int(counts_view.sum()) if rank == 0 else 0               # (the result the e2e leg brought back to the host)

    # ---- e2e on COMPRESSED input (f2): the same reads as a BGZF file in host memory -> kdb_feed_bgzf_host (compressed bytes
    # over PCIe, inflate on the GPU) -> counts + stats on the host; beside it the host-inflate path on the same blob ---------
    bgzf = None
    if world == 1 and not args.no_bgzf:
        import zlib
        from concurrent.futures import ThreadPoolExecutor
        from kmerdb_b200 import kdbfile, parse as kparse
        n_bz = min(n_reads, args.bgzf_reads)
        text_mv = memoryview(pin_text.array)[: n_bz * rb]
        with ThreadPoolExecutor(max_workers=host_threads()) as ex:
            blocks = list(ex.map(lambda o: kdbfile.bgzf_block(bytes(text_mv[o:o + 65280]), 6), range(0, n_bz * rb, 65280)))
        blob = np.frombuffer(b"".join(blocks) + kdbfile._BGZF_EOF, dtype=np.uint8)
        pin_blob = _native.PinnedBuffer(blob.size)
        pin_blob.array[:] = blob

        def step_bgzf():
            prof.reset()
            assert prof.feed_bgzf(pin_blob.array) == 0
            assert prof.finish() == 0
            prof.counts(out=counts_view)
            return prof.stats(), prof.timing()
        step_bgzf()
        t0 = time.perf_counter()
        for _ in range(3):
            st_bz, tm_bz = step_bgzf()
        dt_gpu = (time.perf_counter() - t0) / 3
        assert st_bz["total_kmers"] == n_bz * (READ_LEN - k + 1)
        t0 = time.perf_counter()
        text_host = kparse._gunzip(blob.tobytes(), threads=host_threads())
        prof.reset(); prof.feed(np.frombuffer(text_host, dtype=np.uint8)); assert prof.finish() == 0
        prof.counts(out=counts_view)
        dt_host = time.perf_counter() - t0
        bgzf = {"value": n_bz * READ_LEN / dt_gpu, "unit": UNIT, "reads": n_bz, "h2d_bytes_per_step": int(blob.size), "text_bytes": n_bz * rb,
                "compression_ratio": n_bz * rb / blob.size, "zlib_level": 6, "inflate_ms_per_step": tm_bz["inflate_ms"], "inflate_GBps_of_text": n_bz * rb / (tm_bz["inflate_ms"] * 1e-3) / 1e9,
                "host_inflate_path": {"value": n_bz * READ_LEN / dt_host, "unit": UNIT, "threads": host_threads(),
                                      "note": "the same BGZF blob inflated by zlib on the host threads, then fed uncompressed"}}
        del pin_blob

    # ---- parity gate T3, driver-visible: the N-GPU result equals a 1-GPU job over the same reads -----------------
    t3 = None
    if world > 1:
        import hashlib
        prof.reset()
        prof.feed_device(dev_text.ptr, nbytes)
        assert prof.finish() == 0
        reduce_to_rank0()
        multi_stats = prof.stats()
        if rank == 0:
            multi_sha = hashlib.sha256(prof.counts(out=counts_view).tobytes()).hexdigest()
            chunk_reads = 12_500_000                                     # regenerate the whole stream on this GPU, piece by piece
            one = _native.Profiler(k, canonical=True, skip_n=True, fastq=False, device=local)
            buf = _native.DeviceBuffer(chunk_reads * rb + 16, device=local)
            for r0 in range(0, total_reads, chunk_reads):
                m = min(chunk_reads, total_reads - r0)
                _native.synth_device(buf, m, READ_LEN, seed=SEED, first_read=r0)
                one.feed_device(buf.ptr, m * rb)                         # (fragments at record boundaries)
            assert one.finish() == 0
            one_stats = one.stats()
            one_sha = hashlib.sha256(one.counts(out=counts_view).tobytes()).hexdigest()
            one.close()
            t3 = {"t3_identical": bool(one_sha == multi_sha and one_stats == multi_stats), "counts_sha256_n_gpu": multi_sha,
                  "counts_sha256_1_gpu": one_sha, "stats_equal": bool(one_stats == multi_stats)}
            stats = multi_stats
    if rank == 0:
        assert stats["total_kmers"] == total_reads * (READ_LEN - k + 1), stats
        assert st_host["total_kmers"] == stats["total_kmers"] and e2e_sum == stats["total_kmers"]
        peak, peak_src = peaks()
        # the kernels of one pass (one launch each per pass of <= 512 MiB of text), timed by CUDA events on the context's
        # stream inside the library (kdb_timing).  The DOMINANT one is reported as `roofline`: algorithmic bytes per launch
        # (the text bytes of a pass) / its average launch time, against the measured HBM copy bandwidth.
        n_launch = count_launches                                        # launches of each per-pass kernel over the K steps
        per_launch_bytes = nbytes * args.steps / n_launch
        kern_ms = {"k2s_encode": encode_ms, "k3s_partition": part_ms, "k3p_count": count_ms - part_ms}
        if scan_ms > 0.5 * args.steps:                                   # (the two-pass tokeniser, KDB_TWO_PASS=1)
            kern_ms["k1_scan"] = scan_ms
        dom = max(kern_ms, key=kern_ms.get)
        avg_launch_s = kern_ms[dom] * 1e-3 / n_launch
        achieved = per_launch_bytes / avg_launch_s / 1e9
        traffic, traffic_src = None, None
        for tname in sorted(os.listdir(os.path.join(ROOT, "profiles")), reverse=True):
            if tname.startswith("traffic_r") and tname.endswith(".json"):
                try:
                    tj = json.load(open(os.path.join(ROOT, "profiles", tname)))
                    if dom in tj:
                        traffic = tj[dom]["dram_bytes_per_text_byte"] * per_launch_bytes     # ncu --set full capture, scaled to this launch size
                        traffic_src = "profiles/" + tname
                        break
                except Exception:
                    pass
        per_kernel = {}
        for name, ms in kern_ms.items():
            gbs = nbytes * args.steps / (ms * 1e-3) / 1e9 if ms > 0 else None
            per_kernel[name] = {"ms_per_step": ms / args.steps, "GBps_of_text": gbs, "frac_of_hbm_peak": gbs / peak if gbs else None}
        line = {
            "metric": METRIC, "value": total_reads * READ_LEN * args.steps / job_max,
            "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * job_max / args.steps, "higher_is_better": True, "scaling": "weak" if world == 1 else "strong",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic", "config": workload_config(args, world),
            "wall_ms_per_step": 1e3 * wall_max / args.steps,
            "e2e": {"value": total_reads * READ_LEN * e2e_steps / e2e_wall, "unit": UNIT, "h2d_bytes_per_step": nbytes,
                    "d2h_bytes_per_step": 8 * 4 ** k + 112, "steps": e2e_steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_base": rb / READ_LEN, "bytes_per_launch": per_launch_bytes,
                         "avg_launch_ms": 1e3 * avg_launch_s,
                         "whole_job_frac": (nbytes * args.steps / job_max / 1e9) / peak,
                         "kernels": per_kernel,
                         "note": "no kernel of the step is HBM bound (ncu: DRAM 8-16 %): the tokeniser is issue / barrier bound, the partition is bound by "
                                 "shared-memory bank wavefronts (80 % of the LSU peak), the count by shared atomics: DESIGN.md section 4 and profiles/"},
            "kernel_ms_per_step": {"scan": scan_ms / args.steps, "partition": part_ms / args.steps, "count": count_ms / args.steps, "fold": fold_ms / args.steps, "final": final_ms / args.steps, "encode": encode_ms / args.steps, "reduce": reduce_ms / args.steps},
            "clocks": clocks,
            "result_check": {"total_kmers": stats["total_kmers"], "unique_kmers": stats["unique_kmers"], "total_reads": stats["total_reads"]},
        }
        if t3 is not None:
            line["result_check"].update(t3)
        if bgzf is not None:
            line["e2e_bgzf"] = bgzf
        if not args.no_cpu_baseline:
            line.update(cpu_legs(k, args.ref_reads, args.cpu_reads))
        _emit(line)
    if world > 1:
        ok = torch.tensor([0 if (t3 is not None and not t3["t3_identical"]) else 1], device="cuda")
        dist.broadcast(ok, 0)
        dist.barrier()
        dist.destroy_process_group()
        if not ok.item():
            raise SystemExit("T3 failed: the %d-GPU count vector / statistics differ from the 1-GPU job" % world)


_JSON_OUT = None


def _emit(line):
    """The one JSON line goes to the process's real stdout; everything a library prints (NCCL's version banner, warnings)
    was moved to stderr by main()."""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)                                   # fd 1 -> stderr for the rest of the run (C libraries included)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--reads", type=int, default=0, help="total reads of the job (default: 10 M at N = 1 = configs[1], 100 M sharded N ways at N > 1 = configs[2])")
    ap.add_argument("--k", type=int, default=12)
    ap.add_argument("--cpu-reads", type=int, default=2_000_000, help="bounded sample for the C-port CPU leg")
    ap.add_argument("--ref-reads", type=int, default=2_000, help="bounded sample (reads per step) for the Python reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-bgzf", action="store_true", help="skip the compressed-input leg (e2e_bgzf)")
    ap.add_argument("--bgzf-reads", type=int, default=10_000_000, help="reads of the compressed-input leg")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
