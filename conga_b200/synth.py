"""Deterministic synthetic workloads of the shapes BASELINE.json names.

Everything here is *input* generation (what htslib / SONIC / the BED parsers would
hand to the hot path): per chromosome a SONIC-style GC byte per 100-bp window, a
coordinate-sorted stream of 8-byte read records, del/dup interval lists sorted the way
likelihood.c:324-328 sorts them, and mappability BED rows in file order.  No stage of
the hot path is computed here.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from .abi import CNV_DT, MAPIV_DT, READ_DT

SEED0 = 20221124

# GRCh37 autosome lengths (chr1..chr22)
GRCH37 = [249250621, 243199373, 198022430, 191154276, 180915260, 171115067, 159138663, 146364022,
          141213431, 135534747, 135006516, 133851895, 115169878, 107349540, 102531392, 90354753,
          81195210, 78077248, 59128983, 63025520, 48129895, 51304566]

MAP_VALUES = np.array([1.0, 0.5, 0.333333, 0.25, 0.2, 0.166667, 0.142857, 0.125, 0.0], dtype=np.float32)


@dataclass
class Chromosome:
    name: str
    index: int                      # SONIC chromosome index
    length: int
    gc: np.ndarray                  # int8[ceil(length/100)]
    reads: np.ndarray               # READ_DT, sorted by pos
    dels: np.ndarray                # CNV_DT, sorted by (start, end)
    dups: np.ndarray                # CNV_DT
    maprows: Optional[np.ndarray] = None    # MAPIV_DT in file order, or None


@dataclass
class Workload:
    name: str
    chroms: List[Chromosome] = field(default_factory=list)
    mq_threshold: int = -1
    has_mappability: bool = False
    # a batch (config 5): sample_pos[s][i] = sorted int32 start positions of sample s on chromosome i; sample 0 is
    # what chroms[i].reads holds.  None: one sample.
    sample_pos: Optional[List[List[np.ndarray]]] = None

    @property
    def n_samples(self) -> int:
        return len(self.sample_pos) if self.sample_pos is not None else 1

    @property
    def n_reads(self) -> int:
        return int(sum(len(c.reads) for c in self.chroms))

    @property
    def n_cnvs(self) -> int:
        return int(sum(len(c.dels) + len(c.dups) for c in self.chroms))

    @property
    def n_positions(self) -> int:
        return int(sum(c.length for c in self.chroms))

    def algorithmic_bytes_depth_pass(self) -> int:
        """8·N + 2·L + ceil(L/100) summed over chromosomes: records read, int16 depth written once, GC bytes
        (SURVEY §8d's fused figure less the 4·L prefix array, which is not materialised)."""
        return int(sum(8 * len(c.reads) + 2 * c.length + (c.length + 99) // 100 for c in self.chroms))


def make_gc(length: int, rng: np.random.Generator, n_frac: float = 0.0) -> np.ndarray:
    """GC percent per 100-bp window: a per-kilobase level U(30,60) plus per-window noise, with a
    fraction `n_frac` of windows in N-runs (byte 0, which the reference maps to bin 0)."""
    nw = (length + 99) // 100
    level = rng.uniform(30.0, 60.0, size=(nw + 9) // 10)
    gc = np.repeat(level, 10)[:nw] + rng.normal(0.0, 4.0, size=nw)
    gc = np.clip(np.rint(gc), 1, 100).astype(np.int8)
    if n_frac > 0 and nw > 2000:
        n_runs = max(1, int(nw * n_frac / 1000))
        starts = rng.integers(0, nw - 1000, size=n_runs)
        for s in starts:
            gc[s:s + 1000] = 0
    return gc


def make_reads(length: int, n: int, rng: np.random.Generator, read_len=70, mapq_choices=(0, 10, 30, 37, 60),
               mapq_p=None, dup_frac: float = 0.0, adna: bool = False) -> np.ndarray:
    pos = np.sort(rng.integers(1, max(2, length - 100), size=n, dtype=np.int64)).astype(np.int32)
    r = np.zeros(n, dtype=READ_DT)
    r["pos"] = pos
    if adna:
        ln = np.clip(np.rint(rng.lognormal(np.log(55.0), 0.3, size=n)), 30, 100)
        r["len"] = ln.astype(np.uint16)
    else:
        r["len"] = read_len
    r["mapq"] = rng.choice(np.array(mapq_choices, dtype=np.uint8), size=n, p=mapq_p)
    if dup_frac > 0:
        r["fl"] = np.where(rng.random(n) < dup_frac, 4, 0).astype(np.uint8)
    return r


def make_cnvs(length: int, n: int, rng: np.random.Generator, lo=1000, hi=500_000, lognormal=None) -> np.ndarray:
    """n distinct intervals, sorted by (start, end) — likelihood.c:324-328 / common.c:199-215."""
    if n == 0:
        return np.zeros(0, dtype=CNV_DT)
    hi = min(hi, max(lo + 1, length // 4))
    if lognormal is not None:
        med, sigma, cap = lognormal
        ln = np.clip(rng.lognormal(np.log(med), sigma, size=n), lo, min(cap, hi)).astype(np.int64)
    else:
        ln = np.exp(rng.uniform(np.log(lo), np.log(hi), size=n)).astype(np.int64)
    ln = np.maximum(ln, lo)
    st = rng.integers(0, length - ln, dtype=np.int64)
    key = st * (1 << 32) + (st + ln)
    key = np.unique(key)                         # sorted by start then end, duplicates dropped
    out = np.zeros(len(key), dtype=CNV_DT)
    out["start"] = (key >> 32).astype(np.int32)
    out["end"] = (key & 0xFFFFFFFF).astype(np.int32)
    return out


def make_maprows(length: int, rng: np.random.Generator, mean_len: int = 2500) -> np.ndarray:
    """Contiguous bedGraph rows: row k = [b_k, b_{k+1}] inclusive (adjacent rows share the boundary
    base, which svs.c:368-371 gives to the later row), lengths U[1, 2*mean_len]."""
    n_est = int(length / mean_len * 1.2) + 16
    ln = rng.integers(1, 2 * mean_len + 1, size=n_est, dtype=np.int64)
    b = np.concatenate([[0], np.cumsum(ln)])
    b = b[b < length - 1]
    rows = np.zeros(len(b) - 1, dtype=MAPIV_DT)
    rows["start"] = b[:-1].astype(np.int32)
    rows["end"] = b[1:].astype(np.int32)
    rows["val"] = rng.choice(MAP_VALUES, size=len(rows), p=[.55, .1, .05, .05, .05, .04, .03, .03, .10])
    rows["line"] = np.arange(len(rows), dtype=np.uint32)
    return rows


def _apply_cnv_signal(reads: np.ndarray, dels: np.ndarray, dups: np.ndarray, rng: np.random.Generator) -> np.ndarray:
    """Thin reads inside deletions (odd index 50 %, even index 100 %) and add reads inside
    duplications (odd +50 %, even +100 %) so genotypes are not all 0/0."""
    keep = np.ones(len(reads), dtype=bool)
    pos = reads["pos"]
    for arr, sign in ((dels, -1), (dups, +1)):
        if len(arr) == 0:
            continue
        lo = np.searchsorted(pos, arr["start"], side="left")
        hi = np.searchsorted(pos, arr["end"], side="left")
        if sign < 0:
            for k in range(len(arr)):
                if hi[k] > lo[k]:
                    if k & 1:
                        keep[lo[k]:hi[k]] &= rng.random(hi[k] - lo[k]) >= 0.5
                    else:
                        keep[lo[k]:hi[k]] = False
        else:
            extra = []
            for k in range(len(arr)):
                m = (hi[k] - lo[k]) if not (k & 1) else (hi[k] - lo[k]) // 2
                if m > 0:
                    e = np.zeros(m, dtype=READ_DT)
                    e["pos"] = rng.integers(arr["start"][k], arr["end"][k], size=m)
                    e["len"] = 70
                    e["mapq"] = 60
                    extra.append(e)
            if extra:
                reads = np.concatenate([reads[keep]] + extra)
                reads = reads[np.argsort(reads["pos"], kind="stable")]
                return reads
    return reads[keep]


def _chrom(name, index, length, rng, n_reads, n_dels, n_dups, mappability, cnv_hi=500_000, n_frac=0.0,
           adna=False, dup_frac=0.0, mapq_p=None, lognormal=(None, None), signal_limit=4000) -> Chromosome:
    gc = make_gc(length, rng, n_frac)
    reads = make_reads(length, n_reads, rng, adna=adna, dup_frac=dup_frac, mapq_p=mapq_p)
    dels = make_cnvs(length, n_dels, rng, hi=cnv_hi, lognormal=lognormal[0])
    dups = make_cnvs(length, n_dups, rng, hi=cnv_hi, lognormal=lognormal[1])
    if len(dels) + len(dups) <= signal_limit:
        reads = np.ascontiguousarray(_apply_cnv_signal(reads, dels, dups, rng))
    rows = make_maprows(length, rng) if mappability else None
    return Chromosome(name, index, length, gc, reads, dels, dups, rows)


def config1(length: int = 10_000_000, seed: int = SEED0 + 1) -> Workload:
    """C1: one 10 Mb chromosome, 1x of 70-bp reads, 200 deletions on a 40 kb grid, no mappability."""
    rng = np.random.default_rng(seed)
    gc = make_gc(length, rng)
    reads = make_reads(length, length // 70, rng)
    n = min(200, length // 40_000 - 1)
    dels = np.zeros(n, dtype=CNV_DT)
    dels["start"] = (np.arange(n, dtype=np.int64) * 40_000 + 5_000).astype(np.int32)
    dels["end"] = dels["start"] + rng.integers(1000, 20000, size=n).astype(np.int32)
    reads = np.ascontiguousarray(_apply_cnv_signal(reads, dels, np.zeros(0, dtype=CNV_DT), rng))
    wl = Workload("c1: 10 Mb, 1x 70 bp, 200 dels")
    wl.chroms.append(Chromosome("1", 0, length, gc, reads, dels, np.zeros(0, dtype=CNV_DT), None))
    return wl


def genome(name: str, coverage: float, n_dels: int, n_dups: int, mappability: bool, seed: int,
           scale: float = 1.0, adna: bool = False, mq_threshold: int = -1, cnv_hi: int = 500_000,
           lognormal=(None, None), read_len: int = 70) -> Workload:
    """GRCh37-shaped autosomes (lengths rounded down to a multiple of 100), optionally
    scaled down by `scale` for quick tests.  Reads and CNVs are spread in proportion to length."""
    rng = np.random.default_rng(seed)
    lens = [max(200_000, int(l * scale) // 100 * 100) for l in GRCH37]
    tot = float(sum(lens))
    wl = Workload(name, mq_threshold=mq_threshold, has_mappability=mappability)
    for i, L in enumerate(lens):
        nr = int(round(L * coverage / read_len))
        nd = int(round(n_dels * L / tot))
        nu = int(round(n_dups * L / tot))
        wl.chroms.append(_chrom(str(i + 1), i, L, rng, nr, nd, nu, mappability, cnv_hi=cnv_hi, n_frac=0.01,
                                adna=adna, dup_frac=0.02 if adna else 0.0,
                                mapq_p=[.01, .04, .15, .3, .5] if adna else None, lognormal=lognormal))
    return wl


def config2(scale: float = 1.0, seed: int = SEED0 + 2) -> Workload:
    """C2: GRCh37-shaped genome, 0.1x aDNA-like coverage, 10k dels + 5k dups, mappability BED."""
    return genome("c2: GRCh37-shaped, 0.1x aDNA-like, 10k dels + 5k dups, mappability", 0.1, 10_000, 5_000,
                  True, seed, scale=scale, adna=True, read_len=55)


def config3(scale: float = 1.0, seed: int = SEED0 + 3) -> Workload:
    """C3: whole genome 1x, 1000G-Phase-3-shaped call sets (~33k dels median 3 kb, ~6k dups median 35 kb)."""
    return genome("c3: GRCh37-shaped, 1x, ~40k CNVs", 1.0, 33_000, 6_000, False, seed, scale=scale,
                  lognormal=((3000, 1.3, 1_000_000), (35000, 1.2, 1_000_000)))


def sorted_uniform_positions(length: int, n: int, rng: np.random.Generator) -> np.ndarray:
    """n start positions, uniform on [1, length - 100), already sorted: normalised partial sums of exponential
    gaps (the order statistics of a uniform sample) — no sort, so a 64-sample batch is generated in seconds."""
    if n == 0:
        return np.zeros(0, dtype=np.int32)
    g = rng.standard_exponential(n + 1)
    c = np.cumsum(g)
    pos = 1 + (c[:-1] / c[-1]) * max(1, length - 101)
    return np.minimum(pos.astype(np.int64), length - 1).astype(np.int32)


def config4(scale: float = 1.0, seed: int = SEED0 + 4) -> Workload:
    """C4's depth / genotype side: whole genome 5x of 70-bp reads, --min-mapq 20, the C3-shaped call sets.  (The
    --rp side of config 4 needs bases and a reference sequence: tests/cli_inputs.py builds those for the whole-
    program runs; this workload is what the four hot-path seams see.)"""
    wl = genome("c4: GRCh37-shaped, 5x, ~40k CNVs, --min-mapq 20 (depth + genotype seams)", 5.0, 33_000, 6_000, False, seed,
                scale=scale, mq_threshold=20, lognormal=((3000, 1.3, 1_000_000), (35000, 1.2, 1_000_000)))
    return wl


def config5(n_samples: int = 64, scale: float = 1.0, seed: int = SEED0 + 5) -> Workload:
    """C5: a batch of n_samples 0.5x samples genotyped against ONE shared C3-shaped CNV set.  GC bytes and CNVs
    are those of chroms[]; the samples differ in their reads (seed + sample)."""
    wl = genome("c5: batch of %d GRCh37-shaped 0.5x samples against one ~40k-CNV set" % n_samples, 0.5, 33_000, 6_000, False, seed,
                scale=scale, lognormal=((3000, 1.3, 1_000_000), (35000, 1.2, 1_000_000)))
    wl.sample_pos = [[np.ascontiguousarray(c.reads["pos"]) for c in wl.chroms]]
    for sidx in range(1, n_samples):
        rng = np.random.default_rng(seed + 1000 * sidx)
        wl.sample_pos.append([sorted_uniform_positions(c.length, int(round(c.length * 0.5 / 70)), rng) for c in wl.chroms])
    return wl


def lpt_partition(lengths: List[int], n_parts: int) -> List[List[int]]:
    """Longest-processing-time assignment of chromosomes to GPUs; see multi.lpt_partition."""
    from .multi import lpt_partition as _lpt
    return _lpt(lengths, n_parts)
