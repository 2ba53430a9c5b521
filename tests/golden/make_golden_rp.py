#!/usr/bin/env python
"""Generate tests/golden/rp/ref_rp.npz: the split-read side path (--rp with --dups) of the UNMODIFIED reference
(oracle/_ref/conga_dump) on a small synthetic input with tandem / high-copy / inverted repeats, satellites,
N bases and low-quality reads.  Stored: what the path consumes (loaded reference sequence, records with
SEQ/QUAL, SONIC repeat intervals, sorted CNVs, flags) and what the reference produced (per-CNV rp and
border_rp from the --wrap dump).  Only runs where /root/reference exists; the fixture is committed.

usage: python tests/golden/make_golden_rp.py
"""
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cli_inputs  # noqa: E402
from make_golden import DUMP_DT  # noqa: E402
from conga_b200.abi import READ_DT  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")


def main():
    rng = np.random.default_rng(20221124 + 40)
    chroms = [("1", 180_700), ("2", 120_000)]
    mq, min_len, rp = 20, 60, 4
    tmp = tempfile.mkdtemp(prefix="golden_rp_")
    try:
        paths = cli_inputs.build_inputs_rp(tmp, chroms, rng, reads_per_mb=12000, n_dels=8, n_dups=8, junction_reads=12, repeats=True)
        os.makedirs(os.path.join(tmp, "out"))
        os.makedirs(os.path.join(tmp, "dump"))
        env = dict(os.environ, CONGA_DUMP_DIR=os.path.join(tmp, "dump"))
        subprocess.run([os.path.join(REF, "conga_dump"), "-i", paths["bam"], "-o", os.path.join(tmp, "out", "g"), "--ref", paths["ref"],
                        "--sonic", paths["sonic"], "--dels", paths["dels"], "--dups", paths["dups"], "--rp", str(rp), "--min-mapq", str(mq)],
                       check=True, cwd=tmp, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        seqs = cli_inputs.read_fasta_upper(paths["ref"])
        reps = {}
        for line in open(os.path.join(tmp, "rm.out")).read().splitlines()[3:]:
            f = line.split()
            if len(f) >= 11:
                reps.setdefault(f[4], []).append((int(f[5]), int(f[6]), 1 if ("Satel" in f[10] or "Satel" in f[9]) else 0))
        out = {"chrom_names": np.array([c[0] for c in chroms]), "chrom_lens": np.array([c[1] for c in chroms]),
               "mq_threshold": np.int32(mq), "min_read_length": np.int32(min_len)}
        for k, (name, L) in enumerate(chroms):
            loaded = 0                                       # readReferenceSeq's 1000-base stepping (common.c:434)
            mn, mx = 0, 999
            while mx < L:
                loaded += 1000
                if loaded >= L:
                    break
                mn += 1000
                mx += 1000
            out["c%d_ref" % k] = seqs[name][:min(loaded, L)]
            r = np.fromfile(os.path.join(tmp, "reads.%d.bin" % k), dtype=READ_DT)
            raw = np.fromfile(os.path.join(tmp, "seq.%d.bin" % k), dtype=np.uint8)
            out["c%d_reads" % k] = r
            out["c%d_seqqual" % k] = raw                    # per record: len bases then len phred bytes
            rp_list = sorted(reps.get(name, []))            # SONIC sorts repeats by start (sonic_interval.c:6-21)
            out["c%d_repeats" % k] = np.array(rp_list, dtype=np.int32).reshape(-1, 3)
            for key in ("del", "dup"):
                out["c%d_%s" % (k, key)] = np.fromfile(os.path.join(tmp, "dump", "chr%d.%s" % (k, key)), dtype=DUMP_DT)
        for nm in ("g_svs.bed", "g_dels.bed", "g_dups.bed"):
            out["file_" + nm[2:-4]] = np.frombuffer(open(os.path.join(tmp, "out", nm), "rb").read(), dtype=np.uint8)
        path = os.path.join(ROOT, "tests", "golden", "rp", "ref_rp.npz")
        np.savez_compressed(path, **out)
        print("ref_rp ->", path, os.path.getsize(path), "bytes;",
              "max rp", max(int(out["c%d_dup" % k]["rp"].max()) for k in range(len(chroms))),
              "max border_rp", max(int(out["c%d_del" % k]["border_rp"].max()) for k in range(len(chroms))))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
