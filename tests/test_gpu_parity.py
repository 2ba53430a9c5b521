"""Parity tests proper: the CUDA path, called through the C-ABI, against (a) vectors dumped
from the unmodified reference (tests/golden) and (b) the oracle on the same seeded inputs.

Bar (BASELINE.json north_star): depth arrays, read counts, GC statistics, observed /
expected sums, mappability and copy numbers BIT-EXACT; likelihoods and c-scores within
1e-9 relative (fp64; device libm vs glibc).
"""
import numpy as np
import pytest

from conga_b200 import CNV_DT, MAPIV_DT, MAPRUN_DT, READ_DT, CongaGpuError, hotpath, synth
from conga_b200.abi import CG_ERANGE, CG_TILE
from golden_util import GoldenCase, golden_files

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def assert_calls_equal(got, want, with_map, rtol=RTOL):
    assert np.array_equal(got["observed"], want["observed"])
    assert got["expected"].tobytes() == want["expected"].tobytes()          # ordered fp32 sum, bit-exact
    assert np.array_equal(got["copy_number"], want["copy_number"])
    if with_map:
        assert got["mappability"].tobytes() == want["mappability"].tobytes()
    for f in ("lhomo", "lhetero", "lnone", "score"):
        np.testing.assert_allclose(got[f], want[f], rtol=rtol, atol=0, equal_nan=True, err_msg=f)
    # -0.00 scores (zero reads, F5/A14) must keep their sign: it is printed
    assert np.array_equal(np.signbit(got["score"]), np.signbit(want["score"]))


def check_chromosome_against_oracle(ctx, oracle, ch, mq=-1, has_map=False, batch=1 << 22):
    bi = hotpath.BamInfo()
    hotpath.init_rd_per_chr(ctx, bi, ch, mq)
    hotpath.count_reads_bam(ctx, bi, ch.reads, batch=batch)
    hotpath.calc_mean_per_chr(ctx, bi, ch)
    dels, dups = hotpath.find_depths(ctx, ch, has_map)
    depth, st, exp, d_ref, u_ref = oracle.chromosome(ch, mq, has_map)
    assert np.array_equal(ctx.download_depth(), depth)
    assert np.array_equal(ctx.download_prefix(), np.cumsum(depth.astype(np.int64)).astype(np.int32))
    assert bi.total_read_count_unfiltered == int(st["reads_passed"])
    assert bi.rd_cnt == int(st["rd_cnt"])
    assert np.array_equal(bi.stats["rd_per_gc"], st["rd_per_gc"])
    assert np.array_equal(bi.stats["window_per_gc"], st["window_per_gc"])
    assert bi.expected_read_depth.tobytes() == exp.tobytes()
    assert_calls_equal(dels, d_ref, has_map)
    assert_calls_equal(dups, u_ref, has_map)
    return dels, dups


# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_files(), ids=lambda p: p.split("/")[-1][:-4])
def test_reference_golden_vectors(path, ctx):
    g = GoldenCase(path)
    z = g.z
    for k in g.processed():
        ch = g.chromosome(k)
        bi, log_line, dels, dups = hotpath.run_chromosome(ctx, ch, g.mq_threshold, g.with_map)
        assert np.array_equal(ctx.download_depth(), g.depth(k))
        assert bi.total_read_count_unfiltered == int(z["c%d_read_count" % k])
        assert bi.mean.tobytes() == np.float32(z["c%d_mean" % k]).tobytes()
        assert bi.expected_read_depth.tobytes() == z["c%d_exp" % k].tobytes()
        assert log_line in [str(s) for s in z["log_lines"]]                 # the conga.log line, byte for byte
        assert_calls_equal(dels, z["c%d_del" % k], g.with_map)
        assert_calls_equal(dups, z["c%d_dup" % k], g.with_map)


def test_config1_against_oracle(ctx, oracle):
    wl = synth.config1()
    dels, _ = check_chromosome_against_oracle(ctx, oracle, wl.chroms[0])
    assert (dels["observed"] == 0).sum() >= 90        # the fully thinned deletions
    assert np.any(np.signbit(dels["score"]))          # ... score -0.0 there (A14)


def test_config2_scaled_against_oracle(ctx, oracle):
    wl = synth.config2(scale=0.01)
    for ch in wl.chroms[:6]:
        check_chromosome_against_oracle(ctx, oracle, ch, wl.mq_threshold, True)


def test_config3_scaled_min_mapq(ctx, oracle):
    wl = synth.config3(scale=0.01)
    for ch in (wl.chroms[0], wl.chroms[21]):
        check_chromosome_against_oracle(ctx, oracle, ch, 20, False)


# ---------------------------------------------------------------------------------------
def tiny_chrom(L, reads, dels=(), dups=(), gc=None, maprows=None, seed=0):
    rng = np.random.default_rng(seed)
    if gc is None:
        gc = rng.integers(20, 70, size=(L + 99) // 100).astype(np.int8)
    r = np.zeros(len(reads), dtype=READ_DT)
    if len(reads):
        r["pos"] = [x[0] for x in reads]
        r["mapq"] = [x[1] for x in reads]
        r["len"] = 50
    mk = lambda a: np.array(list(a), dtype=CNV_DT) if len(a) else np.zeros(0, dtype=CNV_DT)
    return synth.Chromosome("t", 3, L, gc, r, mk(dels), mk(dups), maprows)


def test_empty_chromosome(ctx, oracle):
    ch = tiny_chrom(30_000, [], dels=[(0, 1000), (100, 29_999), (29_000, 30_000)], dups=[(5, 2005)])
    dels, dups = check_chromosome_against_oracle(ctx, oracle, ch)
    assert np.all(dels["observed"] == 0) and np.all(dels["expected"] == 0)


def test_mapq_boundary_is_strict(ctx, oracle):
    # bam_data.c:205: qual > threshold; --min-mapq 20 keeps 21.. only (A2)
    ch = tiny_chrom(20_000, [(10, 19), (10, 20), (10, 21), (11, 255), (12, 0)], dels=[(0, 1000)])
    check_chromosome_against_oracle(ctx, oracle, ch, mq=20)
    assert ctx.download_depth()[10] == 1
    check_chromosome_against_oracle(ctx, oracle, ch, mq=-1)
    assert ctx.download_depth()[10] == 3


@pytest.mark.parametrize("L", [100, 1, 99, 16384, 16385, 16300, 49_152, 50_001, 163_840 + 7])
def test_ragged_lengths_and_tile_edges(ctx, oracle, L):
    rng = np.random.default_rng(L)
    n = min(4 * L, 20000)
    pos = np.sort(rng.integers(0, L, size=n))
    # make sure the very first / last positions and both sides of every tile edge are hit
    edge = [p for p in (0, L - 1, CG_TILE - 1, CG_TILE, 2 * CG_TILE - 1, 2 * CG_TILE) if 0 <= p < L]
    pos = np.sort(np.concatenate([pos, edge, edge]))
    reads = list(zip(pos.tolist(), rng.integers(0, 61, size=len(pos)).tolist()))
    cnvs = [(0, L)] + [(a, b) for a, b in ((0, 1), (L // 3, L // 2 + 1), (max(0, L - 101), L), (L // 2, L)) if b > a]
    ch = tiny_chrom(L, reads, dels=sorted(set(cnvs)), dups=sorted(set(cnvs)), seed=L)
    check_chromosome_against_oracle(ctx, oracle, ch, mq=5)


def test_int16_wrap_pileup(ctx, oracle):
    # 70000 identical starts: the reference's `short` wraps twice past 32767 (A1)
    reads = [(777, 60)] * 70000 + [(778, 60)] * 32768 + [(40_000, 60)] * 65536
    ch = tiny_chrom(60_000, sorted(reads), dels=[(0, 1000), (700, 800 + 1000), (30_000, 45_000)], dups=[(0, 60_000)])
    check_chromosome_against_oracle(ctx, oracle, ch, batch=50_000)
    d = ctx.download_depth()
    assert d[777] == np.int16(70000 - 65536) and d[778] == -32768 and d[40_000] == 0


def test_pileup_with_mapq_filter_and_packed_neighbours(ctx, oracle):
    # pile-ups on even positions (low half of a packed bin word, carry must not leak into the odd
    # neighbour), across a chunk edge (2047|2048) and a tile edge, half of the records failing mapq
    rng = np.random.default_rng(9)
    reads = [(1000, int(q)) for q in rng.choice([10, 30], size=90_000)]          # ~45k pass at mq=20 -> wraps once
    reads += [(1001, 60)] * 5 + [(999, 60)] * 3
    reads += [(5000, 60)] * 140_000 + [(5001, 60)] * 7                           # low half overflows twice
    reads += [(2047, 60)] * 40_000 + [(2048, 60)] * 40_000
    reads += [(CG_TILE - 1, 60)] * 33_000 + [(CG_TILE, 30)] * 33_000 + [(CG_TILE, 10)] * 33_000
    reads += [(int(p), 60) for p in rng.integers(0, 40_000, size=3000) if p not in (1001, 5000, 5001)]
    cnvs = [(0, 1500), (900, 2500), (2047, 3047), (2048, 5001), (4000, 6000), (CG_TILE - 500, CG_TILE + 500), (0, 40_000)]
    ch = tiny_chrom(40_000, sorted(reads), dels=sorted(cnvs), dups=sorted(cnvs))
    for mq, batch in ((20, 1 << 22), (-1, 77_777), (20, 16_384)):
        check_chromosome_against_oracle(ctx, oracle, ch, mq=mq, batch=batch)
    d = ctx.download_depth()
    assert d[5000] == np.int16(140_000 - 131_072) and d[5001] == 7 and d[1001] == 5
    # the same records out of order: the global-atomic path must wrap the same way
    ch.reads = np.ascontiguousarray(ch.reads[rng.permutation(len(ch.reads))])
    check_chromosome_against_oracle(ctx, oracle, ch, mq=20, batch=100_000)


def test_multi_batch_equals_single_batch(ctx, oracle):
    wl = synth.config1(length=2_000_000)
    ch = wl.chroms[0]
    check_chromosome_against_oracle(ctx, oracle, ch, batch=1 << 22)
    a = ctx.download_depth()
    check_chromosome_against_oracle(ctx, oracle, ch, batch=1000)       # ~29 pushes, growing the device buffer
    assert np.array_equal(a, ctx.download_depth())


def test_unsorted_input_takes_global_atomic_path(ctx, oracle):
    wl = synth.config1(length=1_000_000)
    ch = wl.chroms[0]
    rng = np.random.default_rng(5)
    ch.reads = np.ascontiguousarray(ch.reads[rng.permutation(len(ch.reads))])
    check_chromosome_against_oracle(ctx, oracle, ch, mq=10, batch=3000)
    # and the context recovers for sorted input afterwards
    check_chromosome_against_oracle(ctx, oracle, synth.config1(length=1_000_000).chroms[0])


def test_out_of_range_is_reported_not_silently_wrong(ctx):
    ch = tiny_chrom(20_000, [(5, 60), (19_999, 60), (20_000, 60)], dels=[(0, 1000)])
    ctx.chrom_begin(ch.index, ch.length, ch.gc, -1)
    ctx.push_reads(ch.reads)
    with pytest.raises(CongaGpuError) as e:
        ctx.chrom_finish()
    assert e.value.code == CG_ERANGE
    ch = tiny_chrom(20_000, [(5, 60)])
    ctx.chrom_begin(ch.index, ch.length, ch.gc, -1)
    ctx.push_reads(ch.reads)
    ctx.chrom_finish()
    with pytest.raises(CongaGpuError) as e:
        ctx.genotype(np.array([(19_000, 20_001)], dtype=CNV_DT), False)
    assert e.value.code == CG_ERANGE


def test_call_order_errors(ctx):
    ch = tiny_chrom(20_000, [(5, 60)])
    ctx.chrom_begin(ch.index, ch.length, ch.gc, -1)
    with pytest.raises(CongaGpuError):
        ctx.genotype(np.array([(0, 1000)], dtype=CNV_DT), False)      # before cg_chrom_finish
    ctx.chrom_finish()
    with pytest.raises(CongaGpuError):
        ctx.push_reads(ch.reads)                                       # after cg_chrom_finish
    with pytest.raises(CongaGpuError):
        ctx.chrom_begin(0, 20_000, ch.gc[:10], -1)                     # GC profile too short


# ---------------------------------------------------------------------------------------
def literal_chain(expected, gc, start, end):
    """likelihood.c:115-119 literally: one fp32 addition per base, in order."""
    e = np.float32(0)
    for i in range(start, end):
        e = np.float32(e + expected[gc[i // 100]])
    return e


@pytest.mark.parametrize("seed", range(4))
def test_expected_chain_is_the_literal_fp32_sum_on_adversarial_tables(ctx, oracle, seed):
    """cg_set_expected with hostile tables: values spanning 40 binades, exact ties (powers of two
    times odd integers), tiny and huge magnitudes.  The run-collapsed device chain must equal the
    reference's one-add-per-base loop bit for bit."""
    rng = np.random.default_rng(100 + seed)
    L = 400_000
    gc = rng.integers(0, 101, size=L // 100).astype(np.int8)
    ch = tiny_chrom(L, [(int(p), 60) for p in np.sort(rng.integers(0, L, size=500))], gc=gc)
    ctx.chrom_begin(ch.index, ch.length, ch.gc, -1)
    ctx.push_reads(ch.reads)
    ctx.chrom_finish()
    table = np.zeros(101, dtype=np.float32)
    kind = seed % 4
    if kind == 0:
        table[1:] = np.exp2(rng.uniform(-30, 10, size=100)).astype(np.float32)
    elif kind == 1:
        table[1:] = (rng.integers(1, 64, size=100) * np.exp2(rng.integers(-24, 4, size=100))).astype(np.float32)  # ties galore
    elif kind == 2:
        table[1:] = rng.uniform(0.001, 0.2, size=100).astype(np.float32)       # realistic low-coverage values
    else:
        table[1:] = np.where(rng.random(100) < 0.3, 0, rng.uniform(1e-6, 3e4, size=100)).astype(np.float32)
    ctx.set_expected(table)
    st = np.sort(rng.integers(0, L - 1, size=300))
    en = np.minimum(L, st + np.exp(rng.uniform(0, np.log(L / 2), size=300)).astype(np.int64) + 1)
    cnvs = np.zeros(300, dtype=CNV_DT)
    cnvs["start"], cnvs["end"] = st, en
    got = ctx.genotype(cnvs, False)
    depth = ctx.download_depth()
    want = oracle.find_depths(depth, gc, table, None, cnvs, False)
    assert got["expected"].tobytes() == want["expected"].tobytes()
    for j in range(0, 300, 37):     # and the oracle's loop is itself the literal NumPy fp32 chain
        assert literal_chain(table, gc, int(st[j]), int(en[j])).tobytes() == got["expected"][j].tobytes()


def test_mappability_general_rows_and_fp64_chain(ctx, oracle):
    """Values that are NOT exactly summable (0.1, 0.3, 1/3...) with overlapping out-of-order rows:
    the fp64 chain must follow the reference's per-base order."""
    rng = np.random.default_rng(9)
    L = 300_000
    n = 800
    rows = np.zeros(n, dtype=MAPIV_DT)
    rows["start"] = rng.integers(0, L - 10, size=n)
    rows["end"] = np.minimum(L - 1, rows["start"] + rng.integers(0, 3000, size=n))
    rows["val"] = rng.choice(np.array([0.1, 0.3, 1 / 3, 0.7, 1e-8, 0.999999, 1.0], dtype=np.float32), size=n)
    rows["line"] = np.arange(n)
    pos = np.sort(rng.integers(0, L, size=3000))
    cn = 200
    st = np.sort(rng.integers(0, L - 2000, size=cn))
    cnvs = np.zeros(cn, dtype=CNV_DT)
    cnvs["start"], cnvs["end"] = st, np.minimum(L, st + rng.integers(1000, 100_000, size=cn))
    ch = tiny_chrom(L, [(int(p), 30) for p in pos], maprows=rows, seed=3)
    ch.dels, ch.dups = cnvs, cnvs[:50].copy()
    check_chromosome_against_oracle(ctx, oracle, ch, has_map=True)


def test_device_pointer_mode_matches_host_mode(ctx, oracle):
    """CG_PTR_DEVICE: inputs resident in HBM (torch tensors), nothing copied, no call blocks."""
    import torch
    from conga_b200 import CALL_DT, CG_PTR_DEVICE, CG_PTR_HOST, STATS_DT, flatten_mappability
    from conga_b200.abi import MAPRUN_DT
    wl = synth.config2(scale=0.004)
    ch = wl.chroms[0]
    dels_h, dups_h = check_chromosome_against_oracle(ctx, oracle, ch, has_map=True)
    dev = torch.device("cuda:0")
    to_dev = lambda a: torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).to(dev)
    gc_d, reads_d = to_dev(ch.gc), to_dev(ch.reads)
    dels_d, dups_d = to_dev(ch.dels), to_dev(ch.dups)
    runs = flatten_mappability(ch.maprows, ch.length)
    runs_d = to_dev(runs)
    out_del = torch.zeros(len(ch.dels) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
    out_dup = torch.zeros(len(ch.dups) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
    stats_d = torch.zeros(STATS_DT.itemsize, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        ctx.chrom_begin(ch.index, ch.length, gc_d, -1, n_windows=len(ch.gc))
        half = len(ch.reads) // 2       # two device batches: the second forces the gather path
        ctx.push_reads(reads_d.data_ptr(), half)
        ctx.push_reads(reads_d.data_ptr() + half * READ_DT.itemsize, len(ch.reads) - half)
        ctx.chrom_finish(stats_d)
        ctx.set_mappability_runs(runs_d, len(runs))
        ctx.genotype(dels_d, False, out_del, n=len(ch.dels))
        ctx.genotype(dups_d, True, out_dup, n=len(ch.dups))
        ctx.sync()
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)
    got_del = out_del.cpu().numpy().view(CALL_DT)
    got_dup = out_dup.cpu().numpy().view(CALL_DT)
    assert got_del.tobytes() == dels_h.tobytes() and got_dup.tobytes() == dups_h.tobytes()   # same kernels, same bits
    st = stats_d.cpu().numpy().view(STATS_DT)[0]
    assert int(st["reads_passed"]) == len(ch.reads)


def test_full_size_chromosome_properties(ctx):
    """chr1 at BASELINE configs[1] size (249 Mb, 0.1x): size-independent checks the oracle is too
    slow for — prefix end == rd_cnt == passing reads, GC sums add up, interval sums are additive,
    and the depth array is a histogram of the pushed positions (spot-checked)."""
    import torch
    L = synth.GRCH37[0] // 100 * 100
    rng = np.random.default_rng(77)
    gc = synth.make_gc(L, rng, 0.01)
    reads = synth.make_reads(L, int(L * 0.1 / 55), rng, adna=True, mapq_p=[.01, .04, .15, .3, .5])
    ctx.chrom_begin(0, L, gc, 0)
    ctx.push_reads(reads)
    st = ctx.chrom_finish()
    n_pass = int((reads["mapq"] > 0).sum())
    assert int(st["reads_passed"]) == n_pass and int(st["rd_cnt"]) == n_pass
    assert int(st["rd_per_gc"].sum()) == n_pass and int(st["window_per_gc"].sum()) == L
    assert np.array_equal(st["window_per_gc"], np.bincount(gc.astype(np.int64), minlength=101) * 100)
    cuts = np.sort(rng.integers(0, L, size=2001))
    cuts[0], cuts[-1] = 0, L
    cuts = np.unique(cuts)
    cnvs = np.zeros(len(cuts) - 1, dtype=CNV_DT)
    cnvs["start"], cnvs["end"] = cuts[:-1], cuts[1:]
    calls = ctx.genotype(cnvs, False)
    assert int(calls["observed"].sum()) == n_pass                     # a partition of the chromosome
    passing = reads["pos"][reads["mapq"] > 0]
    want = np.diff(np.searchsorted(passing, cuts, side="left"))
    assert np.array_equal(calls["observed"], want)
    whole = ctx.genotype(np.array([(0, L)], dtype=CNV_DT), True)
    assert int(whole["observed"][0]) == n_pass
    d_ptr, e_ptr = ctx.device_arrays()
    assert d_ptr and e_ptr
    depth = ctx.download_depth()
    assert int(depth.astype(np.int64).sum()) == n_pass
    idx = rng.integers(0, L, size=1000)
    cnt = np.searchsorted(passing, idx, side="right") - np.searchsorted(passing, idx, side="left")
    assert np.array_equal(depth[idx], cnt.astype(np.int16))


def test_async_host_mode_pipelines_chromosomes_without_mixing_them_up(ctx, oracle):
    """cg_set_async: page-locked buffers, no call blocks, results valid after cg_sync.  Five
    chromosomes of very different sizes go through back to back, so chromosome k's genotype kernel
    runs under chromosome k+1's depth pass and the two slots are reused twice."""
    from conga_b200 import CALL_DT, CG_PTR_HOST, STATS_DT, flatten_mappability
    rng = np.random.default_rng(11)
    lens = [1_500_000, 40_000, 700_000, 16_384, 2_100_000]
    chroms = [synth._chrom("c%d" % i, i, L, rng, max(10, L // 40), 30 + i, 12 + i, True, cnv_hi=max(2000, L // 8))
              for i, L in enumerate(lens)]
    ctx.set_pointer_mode(CG_PTR_HOST)
    ctx.set_async(True)
    keep = []
    try:
        for ch in chroms:
            runs = flatten_mappability(np.ascontiguousarray(ch.maprows), ch.length)
            pin = lambda a, dt: (lambda p: (p.__setitem__(slice(None), a), p)[1])(ctx.pinned_array(len(a), dt))
            pr, pg = pin(ch.reads, READ_DT), pin(ch.gc, np.int8)
            pd, pu, prn = pin(ch.dels, CNV_DT), pin(ch.dups, CNV_DT), pin(runs, runs.dtype)
            od, ou = ctx.pinned_array(len(ch.dels), CALL_DT), ctx.pinned_array(len(ch.dups), CALL_DT)
            st = ctx.pinned_array(1, STATS_DT)
            lib, h = ctx.lib, ctx.h
            ctx._ck(lib.cg_chrom_begin(h, ch.index, ch.length, pg.ctypes.data, len(pg), 10))
            half = len(pr) // 2
            ctx._ck(lib.cg_push_reads(h, pr.ctypes.data, half))
            ctx._ck(lib.cg_push_reads(h, pr.ctypes.data + half * READ_DT.itemsize, len(pr) - half))
            ctx._ck(lib.cg_chrom_finish(h, st.ctypes.data))
            ctx._ck(lib.cg_set_mappability_runs(h, prn.ctypes.data, len(prn), 1))
            ctx._ck(lib.cg_genotype(h, pd.ctypes.data, len(pd), 0, od.ctypes.data))
            ctx._ck(lib.cg_genotype(h, pu.ctypes.data, len(pu), 1, ou.ctypes.data))
            keep.append((ch, st, od, ou, (pr, pg, pd, pu, prn)))
        ctx.sync()
        for ch, st, od, ou, _ in keep:
            depth, st_ref, exp, d_ref, u_ref = oracle.chromosome(ch, 10, True)
            assert int(st[0]["rd_cnt"]) == int(st_ref["rd_cnt"]) and int(st[0]["reads_passed"]) == int(st_ref["reads_passed"])
            assert np.array_equal(st[0]["rd_per_gc"], st_ref["rd_per_gc"])
            assert_calls_equal(od, d_ref, True)
            assert_calls_equal(ou, u_ref, True)
    finally:
        ctx.set_async(False)


def test_statistics_banks(ctx, oracle):
    """cg_stats_bank: rows go to the selected bank and stay there (a row holds what the last chromosome with
    that chr_index left); cg_stats_reset zeroes the current bank."""
    rng = np.random.default_rng(5)
    a = synth._chrom("a", 3, 300_000, rng, 5000, 4, 2, False, cnv_hi=20_000)
    b = synth._chrom("b", 7, 200_000, rng, 3000, 4, 2, False, cnv_hi=20_000)
    st_a = oracle.chromosome(a, -1, False)[1]
    st_b = oracle.chromosome(b, -1, False)[1]
    same = lambda got, want: all(np.array_equal(got[f], want[f]) for f in ("rd_per_gc", "window_per_gc", "rd_cnt", "reads_passed"))
    for bank in (0, 1):
        ctx.stats_bank(bank)
        ctx.stats_reset()
    try:
        ctx.stats_bank(0)
        hotpath.run_chromosome(ctx, a)
        assert same(ctx.stats_row(3), st_a)
        ctx.stats_bank(1)
        hotpath.run_chromosome(ctx, b)
        assert same(ctx.stats_row(7), st_b)
        assert int(ctx.stats_row(3)["window_per_gc"].sum()) == 0          # bank 1 never saw chromosome a
        p1 = ctx.stats_table()[0]
        ctx.stats_bank(0)
        assert ctx.stats_table()[0] != p1
        assert same(ctx.stats_row(3), st_a)                                # still there
        assert int(ctx.stats_row(7)["window_per_gc"].sum()) == 0
        hotpath.run_chromosome(ctx, b)
        assert same(ctx.stats_row(7), st_b) and same(ctx.stats_row(3), st_a)
        ctx.stats_reset()
        assert int(ctx.stats_row(7)["window_per_gc"].sum()) == 0 and int(ctx.stats_row(3)["rd_cnt"]) == 0
        ctx.stats_bank(1)
        assert same(ctx.stats_row(7), st_b)                                # the other bank is untouched
    finally:
        for bank in (1, 0):
            ctx.stats_bank(bank)
            ctx.stats_reset()


def test_run_chromosomes_one_call_per_sample(ctx, oracle):
    """cg_run_chromosomes: one foreign call for the loop body of read_bam over several chromosomes, pinned
    host buffers in asynchronous HOST mode, then the same jobs again with device buffers in DEVICE mode."""
    import torch
    from conga_b200 import CALL_DT, CG_PTR_DEVICE, CG_PTR_HOST, STATS_DT, chrom_job, flatten_mappability
    rng = np.random.default_rng(12)
    lens = [900_000, 30_000, 1_300_000]
    chroms = [synth._chrom("r%d" % i, i, L, rng, max(10, L // 30), 25 + i, (10 + i) if i != 1 else 0, i != 2, cnv_hi=max(2000, L // 8))
              for i, L in enumerate(lens)]
    want = [oracle.chromosome(ch, 10, ch.maprows is not None) for ch in chroms]

    def check(stats, dels, dups):
        for ch, (depth, st_ref, exp, d_ref, u_ref), st, od, ou in zip(chroms, want, stats, dels, dups):
            assert int(st[0]["rd_cnt"]) == int(st_ref["rd_cnt"]) and np.array_equal(st[0]["rd_per_gc"], st_ref["rd_per_gc"])
            assert_calls_equal(od, d_ref, ch.maprows is not None)
            if len(ch.dups):
                assert_calls_equal(ou, u_ref, ch.maprows is not None)

    # HOST mode, nothing blocks
    ctx.set_pointer_mode(CG_PTR_HOST)
    ctx.set_async(True)
    try:
        keep, jobs, outs = [], [], ([], [], [])
        for ch in chroms:
            pin = lambda a, dt: (lambda p: (p.__setitem__(slice(None), a), p)[1])(ctx.pinned_array(len(a), dt))
            runs = flatten_mappability(np.ascontiguousarray(ch.maprows), ch.length) if ch.maprows is not None else np.zeros(0, dtype=MAPRUN_DT)
            bufs = (pin(ch.reads, READ_DT), pin(ch.gc, np.int8), pin(ch.dels, CNV_DT), pin(ch.dups, CNV_DT), pin(runs, MAPRUN_DT))
            st, od, ou = ctx.pinned_array(1, STATS_DT), ctx.pinned_array(len(ch.dels), CALL_DT), ctx.pinned_array(len(ch.dups), CALL_DT)
            jobs.append(chrom_job(ch.index, ch.length, bufs[1], len(bufs[1]), 10, bufs[0], len(bufs[0]), bufs[4], len(bufs[4]), ch.maprows is not None,
                                  bufs[2], len(bufs[2]), od, bufs[3], len(bufs[3]), ou, st))
            keep.append(bufs)
            outs[0].append(st); outs[1].append(od); outs[2].append(ou)
        ctx.run_chromosomes(np.concatenate(jobs))
        ctx.sync()
        check(*outs)
    finally:
        ctx.set_async(False)

    # DEVICE mode
    dev = torch.device("cuda:0")
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        keep, jobs, outs = [], [], ([], [], [])
        for ch in chroms:
            runs = flatten_mappability(np.ascontiguousarray(ch.maprows), ch.length) if ch.maprows is not None else np.zeros(0, dtype=MAPRUN_DT)
            bufs = (up(ch.reads), up(ch.gc), up(ch.dels) if len(ch.dels) else None, up(ch.dups) if len(ch.dups) else None, up(runs) if len(runs) else None)
            st = torch.zeros(STATS_DT.itemsize, dtype=torch.uint8, device=dev)
            od = torch.zeros(max(1, len(ch.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
            ou = torch.zeros(max(1, len(ch.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
            jobs.append(chrom_job(ch.index, ch.length, bufs[1], len(ch.gc), 10, bufs[0], len(ch.reads), bufs[4], len(runs), ch.maprows is not None,
                                  bufs[2], len(ch.dels), od, bufs[3], len(ch.dups), ou, st))
            keep.append(bufs)
            outs[0].append(st); outs[1].append(od); outs[2].append(ou)
        ctx.run_chromosomes(np.concatenate(jobs))
        ctx.sync()
        host = lambda t, dt, n: t.cpu().numpy()[: n * dt.itemsize].view(dt)
        check([host(t, STATS_DT, 1) for t in outs[0]], [host(t, CALL_DT, len(ch.dels)) for t, ch in zip(outs[1], chroms)],
              [host(t, CALL_DT, len(ch.dups)) for t, ch in zip(outs[2], chroms)])
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)


def test_set_slots_rotation(ctx, oracle):
    """cg_set_slots: 4, 2 and again 6 chromosomes in flight; seven chromosomes go through each setting, so every
    slot of the rotation is reused and the ones outside it stay idle."""
    rng = np.random.default_rng(33)
    lens = [300_000, 40_000, 450_000, 16_384, 120_000, 64_000, 700_000]
    chroms = [synth._chrom("t%d" % i, i, L, rng, max(10, L // 35), 6, 3, True, cnv_hi=max(1500, L // 6)) for i, L in enumerate(lens)]
    try:
        for n in (4, 2, 6):
            ctx.set_slots(n)
            for ch in chroms:
                check_chromosome_against_oracle(ctx, oracle, ch, 10, True)
        with pytest.raises(CongaGpuError):
            ctx.set_slots(1)
        with pytest.raises(CongaGpuError):
            ctx.set_slots(7)
    finally:
        ctx.set_slots(6)
