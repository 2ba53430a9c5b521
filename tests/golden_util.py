"""Loader for tests/golden/*.npz (written by tests/golden/make_golden.py from runs of the
unmodified reference)."""
import glob
import os

import numpy as np

from conga_b200.abi import CNV_DT
from conga_b200.synth import Chromosome

HERE = os.path.dirname(os.path.abspath(__file__))


def golden_files():
    return sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))


class GoldenCase:
    def __init__(self, path):
        self.z = np.load(path)
        self.tag = os.path.basename(path)[:-4]
        self.names = [str(x) for x in self.z["chrom_names"]]
        self.lens = [int(x) for x in self.z["chrom_lens"]]
        self.with_map = bool(self.z["with_map"])
        args = [str(a) for a in self.z["args"]]
        self.mq_threshold = int(args[args.index("--min-mapq") + 1]) if "--min-mapq" in args else -1
        self.c_score = float(args[args.index("--c-score") + 1]) if "--c-score" in args else 0.5

    def processed(self):
        """Indices of chromosomes the reference processed (names with X / Y are skipped, bam_data.c:280)."""
        return [k for k, n in enumerate(self.names) if "X" not in n and "Y" not in n]

    def chromosome(self, k) -> Chromosome:
        z = self.z
        d, u = z["c%d_del" % k], z["c%d_dup" % k]

        def cnv(a):
            c = np.zeros(len(a), dtype=CNV_DT)
            c["start"], c["end"] = a["start"], a["end"]
            return c
        rows = np.ascontiguousarray(z["c%d_maprows" % k]) if self.with_map else None
        gc = np.ascontiguousarray(z["c%d_gc" % k])
        return Chromosome(self.names[k], k, self.lens[k], gc, np.ascontiguousarray(z["c%d_reads" % k]), cnv(d), cnv(u), rows)

    def depth(self, k):
        d = np.zeros(self.lens[k], dtype=np.int16)
        d[self.z["c%d_depth_idx" % k]] = self.z["c%d_depth_val" % k]
        return d

    def text(self, which):
        return self.z["file_" + which].tobytes().decode()
