"""The discontinuous decisions of likelihood.c on knife-edge inputs (VERDICT r01 "margin_flags").

`likelihood_score = max((int)lhomo, (int)lhetero) / lnone` (likelihood.c:138,160 with common.c:262-268) and
`copy_number = lhomo > lhetero ? 2 : 1` (:164) flip when a likelihood sits on an integer, or the two on each other.
The device's fp64 log/lgamma and glibc's differ in the last bits, so for inputs closer to such a boundary than that
difference neither libm is an authority.  The kernel flags every CNV within 1e-9 relative of a boundary and takes the
decision from a double-double evaluation of the three likelihoods; this test searches (observed, expected) pairs that
land within 1e-9 (a few of them within 1e-12) of a boundary and checks the device's decisions against 50-digit
arithmetic (mpmath), with glibc's (the oracle's) answers counted beside them.
"""
import numpy as np
import pytest

from conga_b200 import CNV_DT, READ_DT, synth

pytestmark = pytest.mark.gpu
mp = pytest.importorskip("mpmath")
mp.mp.dps = 50


def lp_true(k, lam):
    """k*log(lam) - lam - log(k!) with 50 digits; lam is an exact double"""
    return mp.mpf(k) * mp.log(mp.mpf(lam)) - mp.mpf(lam) - mp.loggamma(mp.mpf(k + 1))


def f64_lp(k, lam):
    from scipy.special import gammaln
    return k * np.log(lam) - lam - gammaln(k + 1.0)


def search_near_integer(ks, per_k=4, seed=1, span=2_000_000):
    """(k, E) with float32 E near 0.15 k such that lhomo(dup) = lp(k, 2E) — about -0.5 k there, moving 4e-4 per
    float32 step of E — is within ~1e-9 relative of an integer"""
    rng = np.random.default_rng(seed)
    out = []
    for k in ks:
        e0 = np.float32(k * rng.uniform(0.13, 0.17))
        bits = np.arange(-span, span, dtype=np.int64) + int(e0.view(np.int32))
        E = bits.astype(np.int32).view(np.float32)
        lam = (np.float32(2.0) * E).astype(np.float64)
        l = f64_lp(float(k), lam)
        d = np.abs(l - np.rint(l))
        idx = np.argsort(d)[:per_k]
        out += [(int(k), float(E[i])) for i in idx if d[i] < 2e-9 * max(1.0, abs(l[i]))]
    return out


def search_near_equal(ks, seed=2):
    """(k, E) with lhomo = lp(k, 2E) within ~1e-9 relative of lhetero = lp(k, 1.5E): E near 2k ln(4/3)"""
    out = []
    for k in ks:
        e0 = np.float32(2.0 * k * np.log(4.0 / 3.0))
        bits = np.arange(-2000, 2000, dtype=np.int64) + int(e0.view(np.int32))
        E = bits.astype(np.int32).view(np.float32)
        lh = f64_lp(float(k), (np.float32(2.0) * E).astype(np.float64))
        lt = f64_lp(float(k), 1.5 * E.astype(np.float64))
        d = np.abs(lh - lt)
        i = int(np.argmin(d))
        if d[i] < 2e-9 * abs(lh[i]):
            out.append((int(k), float(E[i])))
    return out


def run_targets(ctx, targets):
    """One chromosome per 100 targets: target i gets `k` reads on position 100 i (window i, GC bin i + 1), the
    expected table says E for that bin, and the 1-position duplication [100 i, 100 i + 1) has observed = k,
    expected = E exactly."""
    calls = []
    for a in range(0, len(targets), 100):
        part = targets[a:a + 100]
        L = 100 * len(part)
        gc = (np.arange(len(part)) + 1).astype(np.int8)
        pos = np.repeat(np.arange(len(part), dtype=np.int32) * 100, [k for k, _ in part])
        reads = np.zeros(len(pos), dtype=READ_DT)
        reads["pos"], reads["mapq"], reads["len"] = pos, 60, 70
        ctx.chrom_begin(0, L, gc, -1)
        ctx.push_reads(reads)
        ctx.chrom_finish()
        table = np.zeros(101, dtype=np.float32)
        table[1:len(part) + 1] = [e for _, e in part]
        ctx.set_expected(table)
        cnvs = np.zeros(len(part), dtype=CNV_DT)
        cnvs["start"] = np.arange(len(part)) * 100
        cnvs["end"] = cnvs["start"] + 1
        got = ctx.genotype(cnvs, True)
        assert np.array_equal(got["observed"], [k for k, _ in part])
        assert got["expected"].tobytes() == table[1:len(part) + 1].tobytes()
        calls.append(got)
    return np.concatenate(calls)


def test_truncation_of_likelihoods_next_to_integers(ctx, oracle):
    ks = np.unique(np.random.default_rng(7).integers(2_000, 30_000, size=60))
    targets = search_near_integer(ks)
    assert len(targets) >= 100
    got = run_targets(ctx, targets)
    n_flag = n_glibc_differs = n_tight = 0
    for (k, E), g in zip(targets, got):
        lam_h, lam_t = float(np.float32(2.0) * np.float32(E)), 1.5 * float(np.float32(E))
        h, t, n = lp_true(k, lam_h), lp_true(k, lam_t), lp_true(k, float(np.float32(E)))
        assert h > t                                           # 2E is the likelier rate here: the max is (int)lhomo
        want_int = int(h) if h >= 0 else -int(-h)              # C truncation toward zero of the real value
        dist = abs(h - mp.nint(h)) / abs(h)
        if dist < mp.mpf("1e-9"):
            n_flag += 1
            assert g["margin_flags"] & 1, (k, E)
            n_tight += dist < mp.mpf("1e-12")
        got_int = int(round(float(g["score"]) * float(g["lnone"])))
        assert got_int == want_int, (k, E, float(h), got_int)
        assert abs(float(g["lhomo"]) - float(h)) <= 1e-9 * abs(float(h))
        assert abs(float(g["lnone"]) - float(n)) <= 1e-9 * abs(float(n))
        n_glibc_differs += int(np.trunc(oracle.lpoisson(k, lam_h))) != want_int
    assert n_flag >= 60
    print("near-integer likelihoods: %d targets, %d flagged, %d within 1e-12, glibc's truncation differs from the "
          "real value's in %d" % (len(targets), n_flag, n_tight, n_glibc_differs))


def test_copy_number_when_lhomo_meets_lhetero(ctx, oracle):
    ks = np.arange(300, 32_000, 5)
    targets = search_near_equal(ks)
    assert len(targets) >= 3
    got = run_targets(ctx, targets)
    n_glibc_differs = 0
    for (k, E), g in zip(targets, got):
        lam_h, lam_t = float(np.float32(2.0) * np.float32(E)), 1.5 * float(np.float32(E))
        h, t = lp_true(k, lam_h), lp_true(k, lam_t)
        want = 2 if h > t else 1
        if abs(h - t) / abs(h) < mp.mpf("1e-9"):
            assert g["margin_flags"] & 2, (k, E)
        assert int(g["copy_number"]) == want, (k, E, float(h - t))
        n_glibc_differs += (2 if oracle.lpoisson(k, lam_h) > oracle.lpoisson(k, lam_t) else 1) != want
    print("lhomo ~ lhetero: %d targets, glibc's comparison differs from the real values' in %d" % (len(targets), n_glibc_differs))


def test_flags_stay_rare_on_a_real_workload(ctx):
    """C2-shaped chromosome: a flag is a one-in-a-million event, not a code path the bulk of the CNVs takes; CNVs over
    N-runs (expected 0: the three rates are all 0.01 and the likelihoods bit-identical) are not flagged."""
    from conga_b200 import hotpath
    wl = synth.config2(scale=0.02)
    n = flagged = 0
    for ch in wl.chroms[:4]:
        _, _, dels, dups = hotpath.run_chromosome(ctx, ch, -1, True)
        for c in (dels, dups):
            n += len(c)
            flagged += int((c["margin_flags"] != 0).sum())
            same = c["lhomo"] == c["lhetero"]
            assert not np.any(c["margin_flags"][same] & 2)
    assert n > 1000 and flagged <= 2
