"""cg_depth_stats_noalloc (include/conga_gpu.h): the depth side of a chromosome out of the library's fixed
__device__ scratch array, for callers that cannot allocate device memory.  No counterpart in the reference,
so the checker is the library's existing entry point for the same seams:
cg_chrom_begin + cg_push_reads_compact + cg_chrom_finish + cg_get_expected + cg_download_depth.  Integer work and
an fp32 table computed by the same kernels: the comparison is bit for bit.

Collected last on purpose (test_zz_*).  Needs nothing from the tests before it: it
makes its own context, and the entry point itself needs none."""
import numpy as np
import pytest

from conga_b200 import Context, CongaGpuError, depth_stats_noalloc, synth
from conga_b200.abi import CG_EINVAL, CG_GC_BINS, CG_NOALLOC_MAX_LEN, STATS_DT

pytestmark = pytest.mark.gpu


def reference_entry_point(ctx, length, gc, pos, mapq, mq):
    """the existing per-chromosome seams, blocking HOST mode"""
    ctx.chrom_begin(0, length, gc, mq)
    ctx.push_reads_compact(pos, mapq if mq >= 0 else None)
    st = ctx.chrom_finish()
    return ctx.download_depth(), st, ctx.get_expected()


def no_alloc(torch, dev, stream, length, gc, pos, mapq, mq):
    """the new entry point: device inputs and outputs, made by torch outside the call"""
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    t_gc, t_pos = up(gc), up(pos)
    t_mapq = up(mapq) if mq >= 0 else None
    depth = torch.full((length,), -1, dtype=torch.int16, device=dev)
    stats = torch.full((STATS_DT.itemsize,), 0xEE, dtype=torch.uint8, device=dev)
    expected = torch.full((CG_GC_BINS,), float("nan"), dtype=torch.float32, device=dev)
    flags = torch.full((1,), 0x7FFFFFFF, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    depth_stats_noalloc(0, stream, t_pos, t_mapq, len(pos), mq, length, t_gc, len(gc), depth, stats, expected, flags)
    assert torch.cuda.memory_allocated(dev) == before          # nothing came out of torch's allocator either
    torch.cuda.synchronize()
    return (depth.cpu().numpy(), stats.cpu().numpy().view(STATS_DT)[0], expected.cpu().numpy(), int(flags.cpu().numpy()[0]))


def test_no_alloc_equals_existing_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(20221124)
    ctx = Context(0)                                            # the checker's context; the entry point under test needs none
    try:
        # two calls in a row on the shared scratch array: a long chromosome with a mapq filter and a pile-up past the
        # int16 range first, then a SHORTER, sparser one without a filter (whatever the first call left must not show)
        cases = []
        length = 3_000_017
        reads = synth.make_reads(length, 400_000, rng)
        pos = np.ascontiguousarray(reads["pos"]); mapq = np.ascontiguousarray(reads["mapq"])
        pile = np.full(40_000, 1_234_567, dtype=np.int32)       # 40,000 reads on one position: read_depth[] wraps (short, bam_data.c:215)
        pos = np.concatenate([pos, pile]); mapq = np.concatenate([mapq, np.full(len(pile), 60, dtype=np.uint8)])
        order = np.argsort(pos, kind="stable")
        cases.append((length, synth.make_gc(length, rng, n_frac=0.02), pos[order], mapq[order], 20))
        length = 700_001
        reads = synth.make_reads(length, 9_000, rng)
        cases.append((length, synth.make_gc(length, rng), np.ascontiguousarray(reads["pos"]), np.ascontiguousarray(reads["mapq"]), -1))

        stream = torch.cuda.Stream(device=dev)
        for k, (length, gc, pos, mapq, mq) in enumerate(cases):
            want_depth, want_st, want_exp = reference_entry_point(ctx, length, gc, pos, mapq, mq)
            # first call on a stream of the caller's, second on the default stream
            got_depth, got_st, got_exp, flags = no_alloc(torch, dev, stream.cuda_stream if k == 0 else 0, length, gc, pos, mapq, mq)
            assert flags == 0
            assert np.array_equal(got_depth, want_depth), "case %d: depth differs" % k
            assert got_st.tobytes() == np.asarray(want_st).tobytes(), "case %d: statistics row differs" % k
            assert got_exp.tobytes() == want_exp.tobytes(), "case %d: expected_read_depth differs" % k
            assert int(got_st["window_per_gc"].sum()) == length
            if k == 0:
                assert got_depth.min() < 0                      # the pile-up did wrap

        # the limit: rejected, not truncated; nothing is queued
        t = torch.zeros(16, dtype=torch.uint8, device=dev)
        with pytest.raises(CongaGpuError) as ei:
            depth_stats_noalloc(0, 0, None, None, 0, -1, CG_NOALLOC_MAX_LEN + 1, t, 1 << 30, None, None, None, None)
        assert ei.value.code == CG_EINVAL and "CG_NOALLOC_MAX_LEN" in str(ei.value)
        # unsorted records are reported, not repaired (no out-of-order path without scratch for it)
        length, gc, pos, mapq, mq = cases[1]
        bad = pos.copy(); bad[[10, 5000]] = bad[[5000, 10]]
        *_, flags = no_alloc(torch, dev, 0, length, gc, bad, mapq, mq)
        assert flags & 1
    finally:
        ctx.close()
