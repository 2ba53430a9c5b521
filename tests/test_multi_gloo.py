"""The N > 1 host logic on CPU: two processes over gloo.  Each rank takes its chromosomes from the same
LPT partition bench.py uses, produces their statistics rows and calls (the oracle stands in for the GPU
context here — this test is about the plumbing around it), and the ranks then run the path's one
collective (conga_b200.multi.allreduce_stats) and the ordered gather to rank 0.  Rank 0 must end up
with exactly what a single process computes, in SONIC order."""
import os
import socket
import sys

import numpy as np
import pytest

from conga_b200 import STATS_DT, multi, synth

N_CHROM = 7


def _workload():
    rng = np.random.default_rng(77)
    lens = [500_000, 120_000, 410_000, 90_000, 300_000, 50_000, 220_000]
    return [synth._chrom(str(i + 1), i, L, rng, L // 60, 6, 3, True, cnv_hi=30_000) for i, L in enumerate(lens)]


def _rank_main(rank, world, port, q):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch
    import torch.distributed as dist
    import oracle_py
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        chroms = _workload()
        parts = multi.balanced_partition([c.length for c in chroms], world, iters=20_000, seeds=2)      # what bench.py shards by
        orc = oracle_py.load()
        table = torch.zeros((N_CHROM, multi.STATS_WORDS), dtype=torch.int64)
        local = {}
        for i in parts[rank]:
            _, st, _, d, u = orc.chromosome(chroms[i], 10, True)
            table[i] = torch.from_numpy(np.frombuffer(st.tobytes(), dtype=np.int64).copy())
            local[i] = (d, u)
        for _ in range(2):                                    # two "samples": rows must not accumulate
            t = table.clone()
            multi.allreduce_stats(t)
        rows = multi.stats_rows(t)
        calls = multi.gather_calls(local, N_CHROM, dst=0)
        ok_rows = all(int(rows[i]["window_per_gc"].sum()) == chroms[i].length for i in range(N_CHROM))
        if rank == 0:
            q.put(("rank0", parts, rows.tobytes(), [(d.tobytes(), u.tobytes()) for d, u in calls], ok_rows))
        else:
            q.put(("rank%d" % rank, calls is None, ok_rows))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_ranks_over_gloo_match_single_process(oracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    r0 = next(g for g in got if g[0] == "rank0")
    r1 = next(g for g in got if g[0] == "rank1")
    assert r1[1] is True and r1[2] is True                    # non-destination rank: no calls, but every row
    _, parts, rows_b, calls_b, ok_rows = r0
    assert ok_rows
    assert sorted(parts[0] + parts[1]) == list(range(N_CHROM)) and not set(parts[0]) & set(parts[1])
    chroms = _workload()
    rows = np.frombuffer(rows_b, dtype=STATS_DT)
    for i, ch in enumerate(chroms):                           # single-process truth, SONIC order
        _, st, _, d, u = oracle.chromosome(ch, 10, True)
        assert rows[i].tobytes() == st.tobytes()
        assert calls_b[i][0] == d.tobytes() and calls_b[i][1] == u.tobytes()


def test_partition_is_deterministic_and_balanced():
    lens = [c for c in synth.GRCH37]
    for w, floor in ((1, 1.0), (2, 0.99), (4, 0.96), (8, 0.95)):
        p = multi.lpt_partition(lens, w)
        assert p == multi.lpt_partition(list(lens), w)
        assert sorted(i for part in p for i in part) == list(range(22))
        assert all(part == sorted(part) for part in p)
        assert multi.balance(lens, p) >= floor


def test_gather_without_process_group_is_identity():
    from conga_b200 import CALL_DT
    d = np.zeros(2, dtype=CALL_DT)
    out = multi.gather_calls({1: (d, d[:0])}, 3)
    assert len(out) == 3 and len(out[0][0]) == 0 and len(out[1][0]) == 2
