"""Host-side mirror of the reference's hot-path seams, over the C-ABI.

The reference drives the path through four calls inside read_bam (bam_data.c:304-338):
init_rd_per_chr, count_reads_bam, calc_mean_per_chr and find_SVs -> find_depths.  The
functions below keep those names and argument meanings (a `BamInfo` standing for
`bam_info`, common.h:88-104) so that tests read like calls into the reference; each is
a thin wrapper over libconga_gpu.so and raises CongaGpuError where the reference would
exit().  Everything numeric happens on the GPU.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from .abi import CG_GC_BINS, Context
from .synth import Chromosome, Workload


@dataclass
class BamInfo:
    """What `bam_info` (common.h:88-104) carries across the seams, minus the arrays that now
    live in HBM (read_depth, mappability)."""
    total_read_count_unfiltered: int = 0
    mean: np.float32 = np.float32(0)
    expected_read_depth: np.ndarray = field(default_factory=lambda: np.zeros(CG_GC_BINS, np.float32))
    rd_cnt: int = 0
    stats: Optional[np.void] = None


def init_rd_per_chr(ctx: Context, in_bam: BamInfo, chrom: Chromosome, mq_threshold: int = -1) -> None:
    """read_distribution.c:12-18 — zeroed depth array + read counter for this chromosome."""
    ctx.chrom_begin(chrom.index, chrom.length, chrom.gc, mq_threshold)
    in_bam.total_read_count_unfiltered = 0


def count_reads_bam(ctx: Context, in_bam: BamInfo, reads: np.ndarray, batch: int = 1 << 22) -> None:
    """bam_data.c:192-221 — every record with qual > mq_threshold bumps read_depth[pos]."""
    for i in range(0, len(reads), batch):
        ctx.push_reads(reads[i:i + batch])


def calc_mean_per_chr(ctx: Context, in_bam: BamInfo, chrom: Chromosome) -> str:
    """read_distribution.c:27-84 — chromosome mean, GC-bin table; returns the conga.log line."""
    st = ctx.chrom_finish()
    in_bam.stats = st
    in_bam.rd_cnt = int(st["rd_cnt"])
    in_bam.total_read_count_unfiltered = int(st["reads_passed"])
    in_bam.mean = np.float32(np.float64(in_bam.rd_cnt) / np.float64(chrom.length))   # read_distribution.c:39
    in_bam.expected_read_depth = ctx.get_expected()
    return "Read Count:%d  Window count:%d mean=%f\n" % (in_bam.rd_cnt, chrom.length, float(in_bam.mean))


def find_depths(ctx: Context, chrom: Chromosome, has_mappability: bool):
    """likelihood.c:290-308 (+ :352-356 for the mappability track) — all dels, then all dups."""
    ctx.set_mappability(chrom.maprows if has_mappability else None)
    if has_mappability and (chrom.maprows is None or len(chrom.maprows) == 0):
        ctx.set_mappability_runs(None, 0, enabled=True)
    dels = ctx.genotype(chrom.dels, False)
    dups = ctx.genotype(chrom.dups, True)
    return dels, dups


def run_chromosome(ctx: Context, chrom: Chromosome, mq_threshold: int = -1, has_mappability: bool = False):
    """The per-chromosome body of read_bam (bam_data.c:304-338) through the seams above."""
    bi = BamInfo()
    init_rd_per_chr(ctx, bi, chrom, mq_threshold)
    count_reads_bam(ctx, bi, chrom.reads)
    log_line = calc_mean_per_chr(ctx, bi, chrom)
    dels, dups = find_depths(ctx, chrom, has_mappability)
    return bi, log_line, dels, dups


def run_workload(ctx: Context, wl: Workload):
    return [run_chromosome(ctx, c, wl.mq_threshold, wl.has_mappability) for c in wl.chroms]
