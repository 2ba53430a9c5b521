"""The oracle (oracle/conga_oracle.c) replayed against vectors dumped from the UNMODIFIED
reference (tests/golden/make_golden.py).  CPU only.  This is what pins the oracle."""
import numpy as np
import pytest

from golden_util import GoldenCase, golden_files

CASES = golden_files()


def test_golden_fixtures_present():
    assert len(CASES) >= 3


@pytest.mark.parametrize("path", CASES, ids=lambda p: p.split("/")[-1][:-4])
def test_oracle_matches_reference_dump(path, oracle):
    g = GoldenCase(path)
    for k in g.processed():
        ch = g.chromosome(k)
        z = g.z
        # K1: depth + read count (bam_data.c:201-217), bit-exact including the int16 wrap
        depth, cnt = oracle.count_reads(ch.reads, g.mq_threshold, ch.length)
        assert np.array_equal(depth, g.depth(k))
        assert cnt == int(z["c%d_read_count" % k])
        # K2: mean + expected table (read_distribution.c:27-84), bit-exact fp32
        rd, mean = oracle.calc_mu(depth)
        assert mean.tobytes() == np.float32(z["c%d_mean" % k]).tobytes()
        _, wcnt, exp = oracle.calc_mean_per_chr(depth, ch.gc)
        assert exp.tobytes() == z["c%d_exp" % k].tobytes()
        assert int(wcnt.sum()) == ch.length
        # K3/K4: per-CNV fields (likelihood.c:108-169)
        mapp = oracle.paint_mappability(ch.length, ch.maprows) if g.with_map else None
        for cnvs, ref, is_dup in ((ch.dels, z["c%d_del" % k], False), (ch.dups, z["c%d_dup" % k], True)):
            out = oracle.find_depths(depth, ch.gc, exp, mapp, cnvs, is_dup)
            assert np.array_equal(out["observed"], ref["observed"])
            assert out["expected"].tobytes() == ref["expected"].tobytes()
            assert np.array_equal(out["copy_number"], ref["copy_number"])
            if g.with_map:
                assert out["mappability"].tobytes() == ref["mappability"].tobytes()
            for f in ("lhomo", "lhetero", "lnone", "score"):
                # same expression, same glibc as the reference dump: every bit of every double (the sign of -0.0 included)
                assert np.ascontiguousarray(out[f]).tobytes() == np.ascontiguousarray(ref[f]).tobytes(), f


def test_reference_sorted_and_filtered_cnvs():
    """The dump holds what load_known_SVs + qsort kept: sorted by (start,end), >= min_sv_size (A9)."""
    for path in CASES:
        g = GoldenCase(path)
        for k in g.processed():
            for key in ("c%d_del" % k, "c%d_dup" % k):
                a = g.z[key]
                assert np.all(a["end"] - a["start"] >= 1000)
                order = np.lexsort((a["end"], a["start"]))
                assert np.array_equal(order, np.arange(len(a)))


def test_wrap_case_really_wraps():
    g = GoldenCase([p for p in CASES if p.endswith("ref_wrap.npz")][0])
    d = g.depth(0)
    assert d.min() < 0          # 40000 identical starts overflow the reference's `short`


def test_lpoisson_and_imax_quirks(oracle):
    # likelihood.c:100-102: lambda == 0 -> 0.01 ; common.c:262-268 fed doubles truncates toward zero (F5)
    assert oracle.lpoisson(0, 0.0) == -0.01
    assert oracle.imax(-0.9, -0.2) == 0 and oracle.imax(-3.7, -2.2) == -2 and oracle.imax(2.9, 1.1) == 2
