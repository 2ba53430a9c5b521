"""Multi-GPU host logic of the path: which rank takes which chromosome, the one collective, and the
ordered gather of the per-CNV calls.  One process per GPU over ``torch.distributed`` (NCCL on the
GPUs; the same code runs over gloo in the CPU tests).

The reference is a single process that walks the chromosomes in SONIC order (bam_data.c:269-339).
Nothing computed for one chromosome is read by another — even expected_read_depth[] is rebuilt per
chromosome (read_distribution.c:49-84) — so chromosomes are the shards and the data path
needs no collective.  The only exchange is bookkeeping: every rank (and rank 0's conga.log writer)
receives every chromosome's int64[204] statistics row through ONE all-reduce (sum) of the
[n_chrom][204] table in which a rank's foreign rows are zero; and rank 0, which owns the three
output files, receives the calls and writes them in SONIC chromosome order, deletions before
duplications (likelihood.c:178-281).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .abi import CALL_DT, STATS_DT

STATS_WORDS = STATS_DT.itemsize // 8          # 204


def lpt_partition(lengths: Sequence[int], n_parts: int) -> List[List[int]]:
    """Longest-processing-time assignment of chromosomes to ranks by length: the depth
    pass streams 2 bytes per position, so length is the cost.  Deterministic on every rank; each
    part is returned in SONIC order."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    loads = [0] * n_parts
    parts: List[List[int]] = [[] for _ in range(n_parts)]
    for i in order:
        k = min(range(n_parts), key=lambda j: (loads[j], j))
        parts[k].append(i)
        loads[k] += lengths[i]
    for p in parts:
        p.sort()
    return parts


def balanced_partition(lengths: Sequence[int], n_parts: int, iters: int = 300_000, seeds: int = 4) -> List[List[int]]:
    """Chromosomes to ranks with the smallest largest share: longest-processing-time as the start, then seeded
    simulated annealing over single moves and pair swaps (deterministic: every rank computes the same partition).
    For the 22 GRCh37 autosomes LPT leaves the busiest rank with 1/0.972 (4 ranks) and 1/0.966 (8 ranks) of the
    average; this finds 0.9995 and 0.994 — three per cent of an 8-GPU step for a second of host time."""
    import math
    import random
    n = len(lengths)
    if n_parts <= 1 or n <= n_parts:
        return lpt_partition(lengths, n_parts)
    start = [0] * n
    for r, p in enumerate(lpt_partition(lengths, n_parts)):
        for i in p:
            start[i] = r
    best_max, best = None, start
    t0 = max(lengths) * 0.08
    for seed in range(seeds):
        rnd = random.Random(1234 + seed)
        assign = start[:]
        loads = [0] * n_parts
        for i, r in enumerate(assign):
            loads[r] += lengths[i]
        cur = max(loads)
        if best_max is None:
            best_max = cur
        for it in range(iters):
            temp = t0 * (1.0 - it / iters) + 1.0
            i = rnd.randrange(n)
            a = assign[i]
            if rnd.random() < 0.5:
                b = rnd.randrange(n_parts)
                if a == b:
                    continue
                loads[a] -= lengths[i]; loads[b] += lengths[i]
                new = max(loads)
                if new <= cur or rnd.random() < math.exp((cur - new) / temp):
                    assign[i] = b; cur = new
                else:
                    loads[a] += lengths[i]; loads[b] -= lengths[i]
            else:
                j = rnd.randrange(n)
                b = assign[j]
                if a == b:
                    continue
                d = lengths[j] - lengths[i]
                loads[a] += d; loads[b] -= d
                new = max(loads)
                if new <= cur or rnd.random() < math.exp((cur - new) / temp):
                    assign[i], assign[j] = b, a; cur = new
                else:
                    loads[a] -= d; loads[b] += d
            if cur < best_max:
                best_max, best = cur, assign[:]
    parts: List[List[int]] = [[] for _ in range(n_parts)]
    for i, r in enumerate(best):
        parts[r].append(i)
    # a stable order of the parts (by their first chromosome) so that rank numbers do not depend on the search path
    parts = [sorted(p) for p in parts if p] + [[] for p in parts if not p]
    parts.sort(key=lambda p: p[0] if p else n)
    return parts


def balance(lengths: Sequence[int], parts: List[List[int]]) -> float:
    """ideal / achieved makespan of a partition (1.0 = perfect)."""
    loads = [sum(lengths[i] for i in p) for p in parts]
    return (sum(lengths) / len(parts)) / max(loads) if max(loads) else 1.0


def allreduce_stats(table, group=None):
    """The path's one collective: in-place sum of the int64[n_chrom][204] statistics table (a torch
    tensor: CUDA over NCCL/NVLink, CPU over gloo).  Rows of chromosomes a rank does not own must be
    zero on that rank (cg_stats_bank / cg_stats_reset see to that on the device)."""
    import torch
    import torch.distributed as dist
    assert table.dtype == torch.int64 and table.shape[-1] == STATS_WORDS
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(table, op=dist.ReduceOp.SUM, group=group)
    return table


def stats_rows(table) -> np.ndarray:
    """The all-reduced table as STATS_DT records (host)."""
    a = table.detach().cpu().numpy() if hasattr(table, "detach") else np.asarray(table)
    return np.ascontiguousarray(a).view(STATS_DT).reshape(-1)


def gather_calls(local: Dict[int, Tuple[np.ndarray, np.ndarray]], n_chrom: int, dst: int = 0, group=None
                 ) -> Optional[List[Tuple[np.ndarray, np.ndarray]]]:
    """Collect every rank's {chr_index: (del_calls, dup_calls)} on rank `dst`, returned as a list in
    SONIC chromosome order (empty arrays for chromosomes nobody genotyped).  Other ranks get None.
    Calls are a few MB per genome; this is output plumbing, not the data path."""
    import torch.distributed as dist
    empty = (np.zeros(0, dtype=CALL_DT), np.zeros(0, dtype=CALL_DT))
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [local.get(i, empty) for i in range(n_chrom)]
    payload = {int(i): (np.ascontiguousarray(d).tobytes(), np.ascontiguousarray(u).tobytes()) for i, (d, u) in local.items()}
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [None] * world if rank == dst else None
    dist.gather_object(payload, box, dst=dst, group=group)
    if rank != dst:
        return None
    merged: Dict[int, Tuple[np.ndarray, np.ndarray]] = {}
    for part in box:
        for i, (d, u) in part.items():
            if i in merged:
                raise ValueError("chromosome %d was genotyped by two ranks" % i)
            merged[i] = (np.frombuffer(d, dtype=CALL_DT).copy(), np.frombuffer(u, dtype=CALL_DT).copy())
    return [merged.get(i, empty) for i in range(n_chrom)]
