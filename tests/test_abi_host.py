"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/conga_gpu.h declares, refuses to work without a GPU (no CPU fallback), and its one
host-side routine — flattening mappability BED rows into runs — agrees with the oracle's
painter (svs.c:350-371 semantics)."""
import os
import re

import numpy as np
import pytest

from conga_b200 import abi
from conga_b200.abi import MAPIV_DT

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "conga_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = abi.load_library()
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/conga_gpu.h but not exported"
    assert set(names) == set(abi.SYMBOLS), set(names) ^ set(abi.SYMBOLS)
    assert lib.cg_abi_version() == 3


def test_struct_sizes_match_header():
    assert abi.READ_DT.itemsize == 8 and abi.CNV_DT.itemsize == 8 and abi.MAPIV_DT.itemsize == 16
    assert abi.MAPRUN_DT.itemsize == 16 and abi.CALL_DT.itemsize == 56 and abi.STATS_DT.itemsize == 1632
    assert abi.CHROM_JOB_DT.itemsize == 168          # cg_chrom_job, ABI 3: + reads_pos, reads_mapq, the delta-coded stream


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(abi.CongaGpuError):
        abi.Context(0)


def test_no_alloc_entry_point_refuses_sizes_beyond_its_scratch():
    """cg_depth_stats_noalloc checks its arguments before the first CUDA call: a chromosome longer than the fixed
    scratch array is refused with a clear error (no GPU needed, nothing is computed)."""
    with pytest.raises(abi.CongaGpuError) as ei:
        abi.depth_stats_noalloc(0, 0, None, None, 0, -1, abi.CG_NOALLOC_MAX_LEN + 1, 1 << 20, 1 << 30)
    assert ei.value.code == abi.CG_EINVAL and "CG_NOALLOC_MAX_LEN" in str(ei.value)
    with pytest.raises(abi.CongaGpuError) as ei:            # a NULL mapq plane with a threshold that could reject a record
        abi.depth_stats_noalloc(0, 0, 1 << 20, None, 10, 20, 1000, 1 << 20, 10)
    assert ei.value.code == abi.CG_EINVAL
    import re
    hdr = open(os.path.join(ROOT, "include", "conga_gpu.h")).read()
    assert int(re.search(r"#define CG_NOALLOC_MAX_LEN (\d+)", hdr).group(1)) == abi.CG_NOALLOC_MAX_LEN


def test_async_host_mode_refuses_a_pageable_statistics_row():
    """cg_set_async: the pass writes the statistics row into the caller's page-locked memory itself; the Python
    mirror must not hand it a pageable temporary (checked before any foreign call, so no GPU is needed)."""
    c = abi.Context.__new__(abi.Context)          # no cg_create: only the wrapper's own argument check runs
    c.h, c.lib, c.mode, c.is_async = None, None, abi.CG_PTR_HOST, True
    with pytest.raises(abi.CongaGpuError):
        c.chrom_finish()


def test_product_does_not_touch_the_oracle():
    """Nothing under conga_b200/ may import, link or execute oracle/ (tier rule 3)."""
    for dirpath, _d, files in os.walk(os.path.join(ROOT, "conga_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cpp", ".cc", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in txt and "oracle_py" not in txt and "conga_oracle" not in txt, f


def runs_to_dense(runs, L):
    m = np.zeros(L, dtype=np.float32)
    for r in runs:
        m[r["start"]:r["end"]] = r["val"]
    return m


def rows_of(starts, ends, vals):
    r = np.zeros(len(starts), dtype=MAPIV_DT)
    r["start"], r["end"], r["val"] = starts, ends, vals
    r["line"] = np.arange(len(r))
    return r


@pytest.mark.parametrize("seed", range(6))
def test_flatten_matches_painter_random_overlaps(seed, oracle):
    rng = np.random.default_rng(seed)
    L = 5000
    n = 200
    st = rng.integers(-50, L + 20, size=n)
    en = st + rng.integers(-3, 400, size=n)          # includes empty (end < start) and out-of-range rows
    rows = rows_of(st, en, rng.choice(np.array([0, .25, .5, 1, .333333], dtype=np.float32), size=n))
    runs = abi.flatten_mappability(rows, L)
    dense = oracle.paint_mappability(L, rows)
    assert np.array_equal(runs_to_dense(runs, L), dense)
    assert np.all(runs["start"][1:] >= runs["end"][:-1]) and np.all(runs["end"] > runs["start"])


def test_flatten_bedgraph_shared_boundary_goes_to_later_row(oracle):
    # adjacent bedGraph rows share their boundary base; svs.c:368-371 paints inclusive ends in file
    # order, so the shared base carries the LATER row's value
    rows = rows_of([0, 10, 20], [10, 20, 29], np.array([1.0, 0.5, 0.25], dtype=np.float32))
    runs = abi.flatten_mappability(rows, 40)
    dense = runs_to_dense(runs, 40)
    assert dense[10] == np.float32(0.5) and dense[20] == np.float32(0.25) and dense[29] == np.float32(0.25) and dense[30] == 0
    assert np.array_equal(dense, oracle.paint_mappability(40, rows))


def test_flatten_empty_and_golden(oracle):
    assert len(abi.flatten_mappability(np.zeros(0, dtype=MAPIV_DT), 100)) == 0
    from golden_util import GoldenCase, golden_files
    g = GoldenCase([p for p in golden_files() if "map" in p][0])
    for k in g.processed():
        ch = g.chromosome(k)
        runs = abi.flatten_mappability(ch.maprows, ch.length)
        assert np.array_equal(runs_to_dense(runs, ch.length), oracle.paint_mappability(ch.length, ch.maprows))


def test_lpt_partition_balance():
    from conga_b200.synth import GRCH37, lpt_partition
    for n, bound in ((1, 1.0), (2, 0.99), (4, 0.95), (8, 0.95)):
        parts = lpt_partition(GRCH37, n)
        assert sorted(sum(parts, [])) == list(range(22))
        loads = [sum(GRCH37[i] for i in p) for p in parts]
        assert sum(GRCH37) / n / max(loads) >= bound


def test_balanced_partition_beats_lpt_and_is_deterministic():
    from conga_b200 import multi
    from conga_b200.synth import GRCH37
    for n, bound in ((2, 0.9999), (4, 0.999), (8, 0.99)):
        parts = multi.balanced_partition(GRCH37, n)
        assert sorted(sum(parts, [])) == list(range(22))
        assert multi.balance(GRCH37, parts) >= bound >= 0 and multi.balance(GRCH37, parts) >= multi.balance(GRCH37, multi.lpt_partition(GRCH37, n))
        assert parts == multi.balanced_partition(list(GRCH37), n)          # every rank computes the same partition
    assert multi.balanced_partition([5, 3], 4) == multi.lpt_partition([5, 3], 4)


def test_delta_coded_positions_round_trip():
    """cg_encode_positions_delta (host code of the library) against the NumPy statement of what k_decode_delta computes."""
    rng = np.random.default_rng(3)
    for n in (0, 1, 2, 1023, 1024, 1025, 4096, 50_001):
        pos = np.sort(rng.integers(0, max(1, n) * 60 + 9, size=n)).astype(np.int32)
        if n > 10:
            pos[n // 3:] += 1_000_000                      # one gap far above 255, equal positions elsewhere
            pos[5:8] = pos[5]
            pos = np.sort(pos)
        d, big, anc = abi.encode_positions_delta(pos)
        assert d.dtype == np.uint8 and len(d) == n and len(anc) == (n + 1023) // 1024
        assert int((d == 255).sum()) == len(big) and (len(big) == 0 or big.min() >= 255)
        assert np.array_equal(abi.decode_positions_delta(d, big, anc), pos)
    with pytest.raises(abi.CongaGpuError):
        abi.encode_positions_delta(np.array([10, 9], dtype=np.int32))
    with pytest.raises(abi.CongaGpuError):
        abi.encode_positions_delta(np.array([-1, 9], dtype=np.int32))
