"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: shard cuts at record boundaries, the
sum-reduction of partial tables and the scalar pack (sum / min / max words).  The per-shard partials come from
the oracle (test infrastructure) — on the GPU box the same logic runs on NCCL with the CUDA partials."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from kmerdb_b200 import distributed as kd

NONE = 2 ** 64 - 1


def _pack(st):
    p = np.zeros(16, dtype=np.uint64)
    p[0], p[1], p[2] = st["total_reads"], st["total_kmers"], st["sum_len"]
    p[3:7] = st["mono"]
    p[8] = st["min_len"] if st["total_reads"] else NONE
    p[9] = st["max_len"]
    p[10] = NONE
    return p


def _worker(rank, world, port, text, k, fmt, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cuts = kd.shard_bounds(text, fmt, world)
        shard = text[cuts[rank]:cuts[rank + 1]]
        rc, counts, st = oracle.profile(shard, k, fastq=(fmt == "fastq"))
        assert rc == 0
        table = torch.from_numpy(counts.view(np.int64).copy())
        kd.reduce_table(table, dist, dst=0)
        out = kd.combine_partials(_pack(st), dist, torch.device("cpu"), dst=0)
        if rank == 0:
            q.put((table.numpy().view(np.uint64).copy(), out, cuts))
    finally:
        dist.destroy_process_group()


def _run(text, k, fmt, world=2, port=29611):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, text, k, fmt, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res


def _fasta(rng, n):
    out = []
    for i in range(n):
        L = int(rng.integers(12, 300))
        s = "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=L)])
        out.append(">r%d >x\n" % i + "\n".join(s[j:j + 60] for j in range(0, L, 60)) + "\n")
    return "".join(out).encode()


@pytest.mark.parametrize("fmt", ["fasta", "fastq"])
def test_two_rank_reduce_equals_single_profile(fmt):
    rng = np.random.default_rng(5)
    if fmt == "fasta":
        text = _fasta(rng, 500)
    else:
        text = "".join("@q%d\n%s\n+\n%s\n" % (i, "".join(np.array(list("ACGT"))[rng.integers(0, 4, size=80)]), "@" * 80) for i in range(333)).encode()
    k = 6
    table, pack, cuts = _run(text, k, fmt, port=29611 if fmt == "fasta" else 29612)
    rc, counts, st = oracle.profile(text, k, fastq=(fmt == "fastq"))
    assert rc == 0
    assert np.array_equal(table, counts)
    assert pack[:8].tolist() == [st["total_reads"], st["total_kmers"], st["sum_len"]] + st["mono"] + [0]
    assert int(pack[8]) == st["min_len"] and int(pack[9]) == st["max_len"] and int(pack[10]) == NONE
    assert int(np.count_nonzero(table)) == st["unique_kmers"]          # unique is recomputed after the sum, never summed
    assert 0 < cuts[1] < len(text) and text[cuts[1]:cuts[1] + 1] in (b">", b"@")


def test_shard_bounds_edge_cases():
    assert kd.shard_bounds(b">a\nACGT\n", "fasta", 4) == [0, 8, 8, 8, 8]
    t = b">a\nAC>GT\n>b\nAAAA\n>c\nCCCC\n"
    cuts = kd.shard_bounds(t, "fasta", 3)
    assert cuts[0] == 0 and cuts[-1] == len(t) and all(t[c:c + 1] == b">" and t[c - 1:c] == b"\n" for c in cuts[1:-1] if c < len(t))
    q = b"@a\nAC\n+\n@@\n@b\nGG\n+\n@@\n@c\nTT\n+\n@@"
    assert kd.shard_bounds(q, "fastq", 3) == [0, 11, 22, len(q)]


def test_order_preserving_map_roundtrip():
    u = np.array([0, 1, 2 ** 63 - 1, 2 ** 63, NONE], dtype=np.uint64)
    i = kd._ordered_i64(u)
    assert np.all(np.diff(i) > 0) and np.array_equal(kd._unordered_u64(i), u)


def test_first_raise_is_taken_in_file_order():
    """a rank's first-raise key counts records inside its own shard.  Shard 1 raising at its record 2 must
    not beat shard 0 raising at its record 40: the reference meets shard 0's record first (kmerdb/parse.py:128)."""
    g = np.zeros((3, 16), dtype=np.uint64)
    g[:, 0] = [100, 50, 70]                       # records per shard
    g[:, 8] = NONE
    g[:, 10] = [40 * 4 + 2, 2 * 4 + 0, NONE]      # shard 0: KeyError at its record 40; shard 1: NameError at its record 2
    out = kd.fold_partials(g)
    assert int(out[10]) == 40 * 4 + 2 and int(out[0]) == 220
    g[0, 10] = NONE
    out = kd.fold_partials(g)
    assert int(out[10]) == (100 + 2) * 4 + 0      # rebased by the 100 records of shard 0
    g[1, 10] = NONE
    assert int(kd.fold_partials(g)[10]) == NONE


def test_fastq_shard_bounds_blockwise_equals_full_index(monkeypatch):
    rng = np.random.default_rng(3)
    text = "".join("@q%d\n%s\n+\n%s\n" % (i, "A" * int(rng.integers(1, 90)), "I" * 5) for i in range(997)).encode()
    a = np.frombuffer(text, dtype=np.uint8)
    lf = np.flatnonzero(a == 10)
    for world in (2, 3, 8):
        want = [0] + [int(lf[4 * ((997 * r) // world) - 1]) + 1 for r in range(1, world)] + [len(text)]
        assert kd.shard_bounds(text, "fastq", world) == want
