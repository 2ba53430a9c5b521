"""Split-read kernels (conga_b200/csrc/split_read.cu) through the C-ABI against the oracle restatement and the
reference-dumped golden fixture: the row MULTISET must be identical (order is irrelevant to count_ReadPairs),
and so must the mapping count and the per-CNV rp / border_rp."""
import os

import numpy as np
import pytest

import split_ref
from conga_b200 import CNV_DT, HALF_DT, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rp", "ref_rp.npz")


def canon(rows):
    r = rows.copy()
    r["pad"] = 0
    return np.sort(r.view(np.dtype((np.void, r.dtype.itemsize))).ravel())


def device_rows(ctx, L, gc, ref, halves, bases, mq, batches=1):
    ctx.chrom_begin(0, L, gc, mq)
    ctx.sr_reference(ref)
    cut = np.linspace(0, len(halves), batches + 1).astype(int)
    for a, b in zip(cut[:-1], cut[1:]):
        if b > a:
            h = halves[a:b].copy()
            lo = int(h["seq_off"][0])
            hi = int(h["seq_off"][-1]) + int(h["len"][-1])
            h["seq_off"] -= lo
            ctx.sr_push(h, bases[lo:hi])
    rows, n_map = ctx.sr_rows()
    ctx.chrom_finish()
    return rows, n_map


def test_golden_fixture_rows_and_counts(ctx, oracle):
    g = np.load(GOLDEN)
    mq, min_len = int(g["mq_threshold"]), int(g["min_read_length"])
    for k in range(len(g["chrom_names"])):
        L = int(g["chrom_lens"][k])
        rep = g["c%d_repeats" % k]
        halves, bases = split_ref.halves_of(g["c%d_reads" % k], g["c%d_seqqual" % k], rep, mq, min_len)
        want, want_map = oracle.sr_rows(g["c%d_ref" % k], halves, bases, L, mq)
        gc = np.full((L + 99) // 100 + 1, 40, dtype=np.int8)
        got, got_map = device_rows(ctx, L, gc, g["c%d_ref" % k], halves, bases, mq, batches=3)
        assert got_map == want_map
        assert np.array_equal(canon(got), canon(want))
        rows = split_ref.filter_and_widen(got, rep)
        dels, dups = np.zeros(len(g["c%d_del" % k]), dtype=CNV_DT), np.zeros(len(g["c%d_dup" % k]), dtype=CNV_DT)
        for a, src in ((dels, g["c%d_del" % k]), (dups, g["c%d_dup" % k])):
            a["start"], a["end"] = src["start"], src["end"]
        border_rp, rp = ctx.sr_count(rows, dels, dups)
        assert np.array_equal(border_rp, g["c%d_del" % k]["border_rp"])          # the reference's own numbers
        assert np.array_equal(rp, g["c%d_dup" % k]["rp"])


@pytest.mark.parametrize("seed,mq", [(1, -1), (2, 20), (3, 29)])
def test_random_repeats_against_oracle(ctx, oracle, seed, mq):
    """A repeat-rich random chromosome: tandem copies, a 400-copy unit (> 100 hits and one k-mer over 50 000 positions
    when long enough), an inverted repeat, N runs; halves of 10..120 bases incl. ones shorter than the seed."""
    rng = np.random.default_rng(seed)
    L = 600_000
    ref = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=L)
    ref[100_000:100_300] = ord("N")
    for c in range(1, 8):
        ref[200_000 + 150 * c:200_000 + 150 * (c + 1)] = ref[200_000:200_150]
    for c in range(1, 400):
        ref[300_000 + 40 * c:300_000 + 40 * (c + 1)] = ref[300_000:300_040]
    comp = np.zeros(256, dtype=np.uint8)
    for x, y in zip(b"ACGTN", b"TGCAN"):
        comp[x] = y
    ref[450_000:450_500] = comp[ref[440_000:440_500]][::-1]
    ref[500_000:560_000] = ord("A")                                  # poly-A: the AAAAAAAAAA k-mer exceeds MAX_SR_HIT
    n = 6000
    hv = np.zeros(n, dtype=HALF_DT)
    pieces, off = [], 0
    hot = np.concatenate([rng.integers(199_000, 202_000, n // 4), rng.integers(299_000, 317_000, n // 4),
                          rng.integers(439_000, 451_000, n // 4), rng.integers(1, L - 130, n - 3 * (n // 4))])
    for i in range(n):
        ln = int(rng.integers(8, 121))
        src = int(hot[i])
        far = src + int(rng.integers(-60_000, 60_000)) if rng.random() < 0.5 else src       # claimed position vs true origin
        b = ref[src:src + ln].copy()
        if rng.random() < 0.3:
            m = rng.random(ln) < 0.04
            b[m] = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=int(m.sum()))
        if rng.random() < 0.1:
            b = comp[b][::-1]
        hv[i] = (min(max(far, -5), L + 5), off, ln, int(rng.integers(ln, 2 * ln + 2)), int(rng.choice([0, 20, 21, 30, 60])), int(rng.integers(0, 2)), 0)
        pieces.append(b)
        off += ln
    bases = np.concatenate(pieces)
    want, want_map = oracle.sr_rows(ref, hv, bases, L, mq)
    gc = np.full((L + 99) // 100 + 1, 40, dtype=np.int8)
    got, got_map = device_rows(ctx, L, gc, ref, hv, bases, mq, batches=4)
    assert got_map == want_map
    assert len(want) > 50
    assert np.array_equal(canon(got), canon(want))
    dels = synth.make_cnvs(L, 300, rng, hi=80_000)
    dups = synth.make_cnvs(L, 200, rng, hi=80_000)
    rows = got.copy()
    rows["left_end"] -= 50
    rows["right_start"] += 50
    b0, r0 = oracle.sr_count(rows, dels, dups)
    b1, r1 = ctx.sr_count(rows, dels, dups)
    assert np.array_equal(b0, b1) and np.array_equal(r0, r1) and int(b0.sum()) + int(r0.sum()) > 0


def test_empty_and_order_errors(ctx):
    from conga_b200 import CongaGpuError
    gc = np.full(11, 40, dtype=np.int8)
    ctx.chrom_begin(0, 1000, gc, -1)
    ctx.sr_reference(np.frombuffer(b"ACGT" * 250, dtype=np.uint8))
    rows, n_map = ctx.sr_rows()
    assert len(rows) == 0 and n_map == 0
    b, r = ctx.sr_count(rows, np.zeros(0, dtype=CNV_DT), np.zeros(0, dtype=CNV_DT))
    assert len(b) == 0 and len(r) == 0
    ctx.chrom_finish()
