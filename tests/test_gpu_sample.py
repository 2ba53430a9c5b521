"""A whole sample as ONE submission (cg_run_chromosomes in DEVICE / asynchronous HOST mode): every kernel of
the path is launched once for all chromosomes of the job list.  Checked against the oracle per chromosome:
statistics rows bit-exact, observed / expected / mappability / copy number bit-exact, likelihoods 1e-9.

Also here: the compact record planes (cg_push_reads_compact and the reads_pos / reads_mapq job fields), jobs
whose input is out of order (re-run through the per-chromosome out-of-order path) or piles up past the int16
range (corrected inside the pass), more jobs than one job table holds, and the same chr_index in flight many
times (the statistics row must not be shared scratch).
"""
import numpy as np
import pytest

from conga_b200 import (CALL_DT, CG_PTR_DEVICE, CG_PTR_HOST, CNV_DT, MAPRUN_DT, READ_DT, STATS_DT, chrom_job,
                        flatten_mappability, synth)
from test_gpu_parity import assert_calls_equal, tiny_chrom

pytestmark = pytest.mark.gpu


def runs_of(ch):
    if ch.maprows is None:
        return np.zeros(0, dtype=MAPRUN_DT)
    return flatten_mappability(np.ascontiguousarray(ch.maprows), ch.length)


class DeviceSample:
    """Inputs of a job list resident on the GPU, outputs in device buffers."""

    def __init__(self, chroms, mq, fmt="aos"):
        import torch
        self.torch = torch
        dev = torch.device("cuda:0")
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)
        self.keep, self.jobs, self.out = [], [], []
        for ch in chroms:
            runs = runs_of(ch)
            b = dict(gc=up(ch.gc), dels=up(ch.dels) if len(ch.dels) else None, dups=up(ch.dups) if len(ch.dups) else None,
                     runs=up(runs) if len(runs) else None)
            extra = {}
            if fmt == "aos":
                b["reads"] = up(ch.reads) if len(ch.reads) else None
            elif fmt.startswith("delta"):
                from conga_b200 import encode_positions_delta
                d, big, anc = encode_positions_delta(ch.reads["pos"])
                b["delta"], b["big"], b["anc"] = (up(d) if len(d) else None), (up(big) if len(big) else None), (up(anc) if len(anc) else None)
                b["mapq"] = up(np.ascontiguousarray(ch.reads["mapq"])) if len(ch.reads) and fmt == "delta" else None
                extra = dict(reads_delta=b["delta"], reads_big=b["big"], n_big=len(big), reads_anchor=b["anc"])
            else:
                b["pos"] = up(np.ascontiguousarray(ch.reads["pos"])) if len(ch.reads) else None
                b["mapq"] = up(np.ascontiguousarray(ch.reads["mapq"])) if len(ch.reads) and fmt == "planes" else None
            st = torch.zeros(STATS_DT.itemsize, dtype=torch.uint8, device=dev)
            od = torch.zeros(max(1, len(ch.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
            ou = torch.zeros(max(1, len(ch.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
            self.jobs.append(chrom_job(ch.index, ch.length, b["gc"], len(ch.gc), mq, b.get("reads"), len(ch.reads), b["runs"], len(runs),
                                       ch.maprows is not None, b["dels"], len(ch.dels), od, b["dups"], len(ch.dups), ou, st,
                                       reads_pos=b.get("pos"), reads_mapq=b.get("mapq"), **extra))
            self.keep.append(b)
            self.out.append((st, od, ou))
        self.jobs = np.concatenate(self.jobs)
        self.chroms = chroms
        torch.cuda.synchronize()

    def results(self):
        host = lambda t, dt, n: t.cpu().numpy()[: n * dt.itemsize].view(dt).copy()
        return [(host(st, STATS_DT, 1)[0], host(od, CALL_DT, len(ch.dels)), host(ou, CALL_DT, len(ch.dups)))
                for ch, (st, od, ou) in zip(self.chroms, self.out)]


def check_against_oracle(oracle, chroms, results, mq):
    for ch, (st, od, ou) in zip(chroms, results):
        with_map = ch.maprows is not None
        depth, st_ref, exp, d_ref, u_ref = oracle.chromosome(ch, mq, with_map)
        for f in ("rd_per_gc", "window_per_gc", "rd_cnt", "reads_passed"):
            assert np.array_equal(st[f], st_ref[f]), (ch.name, f)
        assert_calls_equal(od, d_ref, with_map)
        assert_calls_equal(ou, u_ref, with_map)


def mixed_chromosomes(n, seed, big=1_200_000):
    """n chromosomes of very different sizes: tiny ones (1 and 99 positions, below one chunk), one without a
    read, some without mappability, some without duplications; every one carries a CNV over its whole length."""
    rng = np.random.default_rng(seed)
    lens = [big, 1, 99, 2048, 2049, 300_000, 65_536, 70_001] + [int(x) for x in rng.integers(3_000, 400_000, size=max(0, n - 8))]
    chroms = []
    for i, L in enumerate(lens[:n]):
        nd, nu = (0, 0) if L < 3000 else (12 + i % 5, 0 if i % 4 == 1 else 5 + i % 3)
        ch = synth._chrom("s%d" % i, i, L, rng, 0 if i == 5 else max(1, L // 30), nd, nu, i % 3 != 2 and L >= 3000, cnv_hi=max(1500, L // 6))
        whole = np.array([(0, L)], dtype=CNV_DT)
        ch.dels = np.concatenate([whole, ch.dels[ch.dels["start"] > 0]]) if len(ch.dels) else whole
        if len(ch.reads):
            ch.reads["pos"] = np.minimum(ch.reads["pos"], L - 1)
        chroms.append(ch)
    return chroms


def test_sample_of_40_chromosomes_device_mode(ctx, oracle):
    """More jobs than one job table holds (CG_JMAX = 32): two submissions behind one call."""
    chroms = mixed_chromosomes(40, seed=21)
    s = DeviceSample(chroms, 10)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        n0 = ctx.launch_count
        ctx.run_chromosomes(s.jobs)
        ctx.sync()
        assert ctx.launch_count - n0 <= 12          # six launches per submission, not eight per chromosome
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)
    check_against_oracle(oracle, chroms, s.results(), 10)


@pytest.mark.parametrize("fmt,mq", [("planes", 20), ("pos_only", -1), ("aos", -1), ("delta", 20), ("delta_pos_only", -1)])
def test_sample_record_formats_device_mode(ctx, oracle, fmt, mq):
    chroms = mixed_chromosomes(9, seed=5, big=600_000)
    s = DeviceSample(chroms, mq, fmt)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        ctx.run_chromosomes(s.jobs)
        ctx.sync()
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)
    check_against_oracle(oracle, chroms, s.results(), mq)


def test_sample_compact_planes_asynchronous_host_mode(ctx, oracle):
    """Page-locked host buffers, compact planes, nothing blocks; three samples back to back through the slots,
    every sample with its own result buffers."""
    chroms = mixed_chromosomes(10, seed=8, big=900_000)
    pin = lambda a, dt: (lambda p: (p.__setitem__(slice(None), a), p)[1])(ctx.pinned_array(len(a), dt))
    ctx.set_pointer_mode(CG_PTR_HOST)
    ctx.set_async(True)
    try:
        inputs = []
        for ch in chroms:
            runs = runs_of(ch)
            inputs.append((pin(np.ascontiguousarray(ch.reads["pos"]), np.int32), pin(np.ascontiguousarray(ch.reads["mapq"]), np.uint8),
                           pin(ch.gc, np.int8), pin(ch.dels, CNV_DT), pin(ch.dups, CNV_DT), pin(runs, MAPRUN_DT)))
        samples = []
        for _ in range(3):
            outs = [(ctx.pinned_array(1, STATS_DT), ctx.pinned_array(len(ch.dels), CALL_DT), ctx.pinned_array(len(ch.dups), CALL_DT)) for ch in chroms]
            jobs = np.concatenate([chrom_job(ch.index, ch.length, pg, len(pg), 10, None, len(pp), prn, len(prn), ch.maprows is not None,
                                             pd, len(pd), od, pu, len(pu), ou, st, reads_pos=pp, reads_mapq=pm)
                                   for ch, (pp, pm, pg, pd, pu, prn), (st, od, ou) in zip(chroms, inputs, outs)])
            ctx.run_chromosomes(jobs)
            samples.append(outs)
        ctx.sync()
        for outs in samples:
            check_against_oracle(oracle, chroms, [(st[0], od, ou) for st, od, ou in outs], 10)
    finally:
        ctx.set_async(False)


def test_sample_delta_coded_stream_asynchronous_host_mode(ctx, oracle):
    """Positions as one byte per record (cg_encode_positions_delta) from page-locked host memory; the chromosomes
    include gaps far above 255 (sparse ones), blocks that end inside the stream, a chromosome without a record, and a
    pile-up of equal positions (delta 0)."""
    from conga_b200 import encode_positions_delta
    from conga_b200.abi import ANCHOR_DT
    chroms = mixed_chromosomes(10, seed=18, big=1_000_000)
    rng = np.random.default_rng(4)
    sparse = synth._chrom("sp", 10, 3_000_000, rng, 900, 6, 3, False, cnv_hi=200_000)          # mean gap 3300: almost every delta escapes
    pile = tiny_chrom(50_000, sorted([(1234, 60)] * 3000 + [(int(p), 30) for p in rng.integers(0, 50_000, size=5000)]), dels=[(0, 50_000), (1000, 2500)])
    pile.index = 11
    chroms = chroms + [sparse, pile]
    pin = lambda a, dt: (lambda p: (p.__setitem__(slice(None), a), p)[1])(ctx.pinned_array(len(a), dt))
    ctx.set_pointer_mode(CG_PTR_HOST)
    ctx.set_async(True)
    try:
        jobs, outs, keep = [], [], []
        for ch in chroms:
            runs = runs_of(ch)
            d, big, anc = encode_positions_delta(ch.reads["pos"])
            bufs = (pin(d, np.uint8), pin(big, np.int32), pin(anc, ANCHOR_DT), pin(np.ascontiguousarray(ch.reads["mapq"]), np.uint8),
                    pin(ch.gc, np.int8), pin(ch.dels, CNV_DT), pin(ch.dups, CNV_DT), pin(runs, MAPRUN_DT))
            st, od, ou = ctx.pinned_array(1, STATS_DT), ctx.pinned_array(len(ch.dels), CALL_DT), ctx.pinned_array(len(ch.dups), CALL_DT)
            jobs.append(chrom_job(ch.index, ch.length, bufs[4], len(ch.gc), 10, None, len(ch.reads), bufs[7], len(runs), ch.maprows is not None,
                                  bufs[5], len(ch.dels), od, bufs[6], len(ch.dups), ou, st, reads_mapq=bufs[3],
                                  reads_delta=bufs[0], reads_big=bufs[1], n_big=len(big), reads_anchor=bufs[2]))
            keep.append(bufs)
            outs.append((st, od, ou))
        for _ in range(2):
            ctx.run_chromosomes(np.concatenate(jobs))
        ctx.sync()
        check_against_oracle(oracle, chroms, [(st[0], od, ou) for st, od, ou in outs], 10)
        ctx.set_async(False)                             # blocking mode runs chromosome by chromosome: no delta streams there
        with pytest.raises(Exception):
            ctx.run_chromosomes(np.concatenate(jobs))
    finally:
        ctx.set_async(False)


def test_sample_with_unsorted_and_piled_up_jobs(ctx, oracle):
    """Job 1's records are shuffled (flagged on the device, re-run through the out-of-order path inside cg_sync);
    job 2 piles 70000 and 32768 records on single positions (int16 wrap, corrected inside the pass, no re-run)."""
    rng = np.random.default_rng(3)
    chroms = mixed_chromosomes(4, seed=13, big=500_000)
    a = synth._chrom("u", 4, 350_000, rng, 9000, 9, 4, True, cnv_hi=40_000)
    a.reads = np.ascontiguousarray(a.reads[rng.permutation(len(a.reads))])
    reads = sorted([(777, 60)] * 70000 + [(778, 60)] * 32768 + [(40_000, 60)] * 65536 + [(2047, 30)] * 40_000 + [(2048, 60)] * 33_000
                   + [(int(p), 60) for p in rng.integers(0, 60_000, size=2000)])
    b = tiny_chrom(60_000, reads, dels=[(0, 1000), (700, 1800), (2047, 3047), (30_000, 45_000), (0, 60_000)], dups=[(0, 60_000), (500, 2600)])
    b.index = 5
    chroms = chroms + [a, b]
    s = DeviceSample(chroms, 10)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        ctx.run_chromosomes(s.jobs)
        ctx.sync()
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)
    res = s.results()
    check_against_oracle(oracle, chroms, res, 10)
    assert int(res[-1][2]["observed"][0]) < 0          # the wrapped int16 values dominate the whole-chromosome sum


def test_same_chromosome_twenty_times_in_flight(ctx, oracle):
    """consecutive samples reuse chr_index while several submissions are in flight.  The row in the
    statistics bank is only ever written whole by the finalize kernel of a submission; every partial sum lives in
    the slot.  Twenty submissions of one chromosome, alternating between two different read sets."""
    rng = np.random.default_rng(17)
    x = synth._chrom("x", 0, 800_000, rng, 30_000, 20, 8, True, cnv_hi=60_000)
    y = synth._chrom("x", 0, 800_000, rng, 11_000, 20, 8, True, cnv_hi=60_000)
    sx, sy = [DeviceSample([x], -1) for _ in range(10)], [DeviceSample([y], -1) for _ in range(10)]
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        for a, b in zip(sx, sy):
            ctx.run_chromosomes(a.jobs)
            ctx.run_chromosomes(b.jobs)
        ctx.sync()
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)
    for a, b in zip(sx, sy):
        check_against_oracle(oracle, [x], a.results(), -1)
        check_against_oracle(oracle, [y], b.results(), -1)


def test_push_reads_compact_per_chromosome(ctx, oracle):
    """cg_push_reads_compact through the per-chromosome entry points: several host batches (staged), with and
    without the mapq plane; then two device batches (the second forces the gather)."""
    import torch
    from conga_b200 import hotpath
    wl = synth.config1(length=1_500_000)
    ch = wl.chroms[0]
    pos, mapq = np.ascontiguousarray(ch.reads["pos"]), np.ascontiguousarray(ch.reads["mapq"])
    for mq, with_plane in ((20, True), (-1, True), (-1, False)):
        depth, st_ref, exp, d_ref, _ = oracle.chromosome(ch, mq, False)
        ctx.chrom_begin(ch.index, ch.length, ch.gc, mq)
        for i in range(0, len(pos), 7000):
            ctx.push_reads_compact(pos[i:i + 7000], mapq[i:i + 7000] if with_plane else None)
        st = ctx.chrom_finish()
        assert int(st["rd_cnt"]) == int(st_ref["rd_cnt"]) and np.array_equal(st["rd_per_gc"], st_ref["rd_per_gc"])
        assert np.array_equal(ctx.download_depth(), depth)
        assert ctx.get_expected().tobytes() == exp.tobytes()
        assert_calls_equal(ctx.genotype(ch.dels, False), d_ref, False)
    with pytest.raises(Exception):                       # no mapq plane although the threshold can reject records
        ctx.chrom_begin(ch.index, ch.length, ch.gc, 5)
        ctx.push_reads_compact(pos, None)
    ctx.chrom_finish()
    with pytest.raises(Exception):                       # one format per chromosome
        ctx.chrom_begin(ch.index, ch.length, ch.gc, -1)
        ctx.push_reads(ch.reads[:10])
        ctx.push_reads_compact(pos[10:], mapq[10:])
    ctx.chrom_finish()
    dev = torch.device("cuda:0")
    pos_d, mapq_d, gc_d = torch.from_numpy(pos).to(dev), torch.from_numpy(mapq).to(dev), torch.from_numpy(ch.gc).to(dev)
    torch.cuda.synchronize()
    depth, st_ref, exp, d_ref, _ = oracle.chromosome(ch, 20, False)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        half = len(pos) // 2
        ctx.chrom_begin(ch.index, ch.length, gc_d, 20, n_windows=len(ch.gc))
        ctx.push_reads_compact(pos_d.data_ptr(), mapq_d.data_ptr(), half)
        ctx.push_reads_compact(pos_d.data_ptr() + 4 * half, mapq_d.data_ptr() + half, len(pos) - half)
        ctx.chrom_finish(None)
        ctx.sync()
        assert np.array_equal(ctx.download_depth(), depth)
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)


def test_peer_exchange_two_contexts(ctx, oracle):
    """The fused collective (cg_peer_*): k_finalize stores a finished row into every rank's table, cg_stats_allreduce is
    a flag exchange.  Two contexts of this process are each other's peers (cg_peer_connect_local; on one GPU here, over
    NVLink between GPUs of a box); each runs its own chromosomes of two samples through alternating banks, and afterwards
    BOTH hold every chromosome's row of both samples."""
    from conga_b200 import Context, peer_connect_local
    rng = np.random.default_rng(31)
    chroms = [synth._chrom("p%d" % i, i, L, rng, max(10, L // 40), 5, 2, False, cnv_hi=max(1500, L // 8))
              for i, L in enumerate([400_000, 90_000, 250_000, 30_000, 120_000])]
    want = [oracle.chromosome(ch, -1, False)[1] for ch in chroms]
    other = Context(0)
    try:
        ranks = [ctx, other]
        peer_connect_local(ranks)
        mine = [[chroms[0], chroms[3]], [chroms[1], chroms[2], chroms[4]]]
        samples = [[DeviceSample(m, -1) for m in mine] for _ in range(2)]
        for c in ranks:
            c.set_pointer_mode(CG_PTR_DEVICE)
        for k, smp in enumerate(samples):
            for c, s in zip(ranks, smp):
                c.stats_bank(k)
                c.run_chromosomes(s.jobs)
                c.stats_allreduce(len(chroms))           # nothing blocks: the other rank's exchange is issued right behind
        for c in ranks:
            c.sync()
        for c in ranks:
            for bank in (0, 1):
                c.stats_bank(bank)
                for ch, st_ref in zip(chroms, want):
                    got = c.stats_row(ch.index)
                    for f in ("rd_per_gc", "window_per_gc", "rd_cnt", "reads_passed"):
                        assert np.array_equal(got[f], st_ref[f]), (bank, ch.name, f)
        for smp in samples:                              # and the calls are what a single rank computes
            for m, s in zip(mine, smp):
                check_against_oracle(oracle, m, s.results(), -1)
    finally:
        for c in (ctx, other):
            c.set_pointer_mode(CG_PTR_HOST)
            c.stats_bank(0)
        ctx.lib.cg_peer_disconnect(ctx.h)
        other.close()
