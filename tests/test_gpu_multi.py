"""Parity gate T3: 1 GPU vs N GPUs give identical results.  Needs >= 2 visible GPUs."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least 2 GPUs")
def test_multi_gpu_equals_single_gpu():
    n = min(_n_gpus(), 8)
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_multi_gpu_worker.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                        "--master-port", "29655", worker], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("identical") == 4 and "DIFFERENT" not in r.stdout


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least 2 GPUs")
def test_reduce_peers_one_process_equals_single_gpu():
    """kdb_reduce_peers: one process, one context per GPU, the sum / count / exchange in one kernel per GPU over NVLink."""
    import numpy as np
    from kmerdb_b200 import _native, distributed as kd, parse
    n = min(_n_gpus(), 8)
    text = _native.synth_host(300000)
    for k, canonical in ((12, True), (8, False)):
        one = parse.profile_bytes(text, k, "fasta", True, canonical, device=0)
        many = parse.profile_bytes(text, k, "fasta", True, canonical, device=list(range(n)))
        assert np.array_equal(one.counts, many.counts) and np.array_equal(one.nullomer_array, many.nullomer_array)
        for f in ("total_reads", "total_kmers", "unique_kmers", "min_read_length", "max_read_length", "avg_read_length", "mononucleotides"):
            assert getattr(one, f) == getattr(many, f), f
    # first raise in file order (shard 0 holds a KeyError at record 40, the last shard a NameError at its record 2)
    recs = [">r%d\n%s\n" % (i, "ACGT" * 30) for i in range(4000)]
    recs[40] = ">r40\n%s\n" % ("ACGTRCGT" * 15)
    recs[4000 * (n - 1) // n + 2] = ">x\n%s\n" % ("acgt" * 30)
    with pytest.raises(KeyError):
        parse.profile_bytes("".join(recs).encode(), 12, "fasta", True, True, device=list(range(n)))


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least 2 GPUs")
def test_cli_devices_writes_a_byte_identical_kdb(tmp_path):
    """`kmerdb profile --devices 0-(N-1)` (kmerdb_b200 extension) vs `--device 0`: the .kdb files are the same bytes after
    gunzip (header incl. per-file metadata, and every row)."""
    import argparse
    import gzip
    from kmerdb_b200 import _native, cli
    n = min(_n_gpus(), 8)
    fa = tmp_path / "reads.fa"
    fa.write_bytes(bytes(_native.synth_host(120000)))
    outs = []
    for tag, devices in (("one", None), ("many", "0-%d" % (n - 1))):
        out = str(tmp_path / tag)
        ns = argparse.Namespace(device=0, devices=devices, input=[str(fa)], k=10, minK=None, maxK=None, output_name=out, no_ambiguous=False,
                                do_not_canonicalize=False, sorted=False, quiet=True, debug=False, verbose=0, num_log_lines=0, logfile=None)
        cli.profile(ns)
        outs.append(gzip.open(out + ".10.kdb", "rb").read())
    assert outs[0] == outs[1]
