"""Whole-program comparison on the same files: the unmodified reference `conga` (oracle/_ref/conga, one core —
the program is single-threaded) against `conga-b200` (conga_b200/host), BAM decode and all file I/O
included.  Checks that the three BED files and the conga.log chromosome lines are byte-identical, then
prints one JSON line with the wall-clock times.  Measurement tooling (may use oracle/), not product.

  python scripts/cli_bench.py [--workload c1|c2|c3] [--scale S] [--runs R] [--gpus N] [--skip-reference]
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np  # noqa: E402
import refdata  # noqa: E402
from conga_b200 import synth  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
HOST = os.path.join(ROOT, "conga_b200", "host", "conga-b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c1")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--skip-reference", action="store_true")
    ap.add_argument("--keep", action="store_true")
    ap.add_argument("--batch", type=int, default=0, help="also run K copies of the sample through CONGA_BATCH in one process")
    args = ap.parse_args()
    wl = {"c1": synth.config1, "c2": lambda: synth.config2(scale=args.scale), "c3": lambda: synth.config3(scale=args.scale)}[args.workload]()
    tmp = tempfile.mkdtemp(prefix="clibench_")
    try:
        t0 = time.perf_counter()
        base = refdata.write_workload(tmp, wl)                  # SONIC container, record files, BEDs
        tsv = os.path.join(tmp, "chroms.tsv")
        with open(tsv, "w") as f:
            for c in wl.chroms:
                f.write("%s\t%d\t%s\n" % (c.name, c.length, os.path.join(tmp, "reads.%d.bin" % c.index)))
        bam = os.path.join(tmp, "in.bam")
        subprocess.run([os.path.join(REF, "mkbam"), bam, tsv], check=True, stdout=subprocess.DEVNULL)
        open(os.path.join(tmp, "ref.fa"), "w").close()          # --ref is mandatory on the command line, read only by --rp
        flags = []
        it = iter(base)
        for a in it:                                            # ref_path's argument list -> conga's
            if a == "--reads-dir":
                next(it)
            elif a == "--min-sv-size":
                flags += [a, next(it)]
            else:
                flags += [a, next(it)]
        t_prep = time.perf_counter() - t0

        def run(exe, tag, env=None):
            wd = os.path.join(tmp, tag)
            os.makedirs(os.path.join(wd, "o"), exist_ok=True)
            cmd = [exe, "-i", bam, "--ref", os.path.join(tmp, "ref.fa"), "--out", os.path.join(wd, "o", "s")] + flags
            t = time.perf_counter()
            p = subprocess.run(cmd, cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, env=dict(os.environ, **(env or {})))
            dt = time.perf_counter() - t
            if p.returncode != 0:
                sys.exit("%s failed: %s" % (tag, p.stderr.decode(errors="replace")[-1500:]))
            if os.environ.get("CONGA_TIMING"):
                sys.stderr.write("".join(l + "\n" for l in p.stderr.decode(errors="replace").split("\n") if l.startswith("[timing]")))
            out = {}
            for nm in ("s_svs.bed", "s_dels.bed", "s_dups.bed"):
                fp = os.path.join(wd, "o", nm)
                out[nm] = open(fp, "rb").read() if os.path.exists(fp) else None
            out["log"] = b"\n".join(open(os.path.join(wd, "conga.log"), "rb").read().split(b"\n")[1:])
            return dt, out

        res = {"workload": wl.name, "reads": wl.n_reads, "cnvs": wl.n_cnvs, "positions": wl.n_positions,
               "bam_bytes": os.path.getsize(bam), "prep_s": round(t_prep, 1), "host_cores": len(os.sched_getaffinity(0))}
        env = {"CONGA_GPUS": str(args.gpus)}
        tb = []
        for _ in range(args.runs):
            dt, ob = run(HOST, "b200", env)
            tb.append(dt)
        res["b200_s"] = [round(x, 3) for x in tb]
        res["b200_median_s"] = round(float(np.median(tb)), 3)
        res["b200_gpus"] = args.gpus
        res["b200_reads_per_s"] = round(wl.n_reads / float(np.median(tb)), 1)
        if args.batch > 0:
            # configs[4] shape: many samples against one CNV set, one process (contexts, SONIC, BEDs, tracks resident)
            wd = os.path.join(tmp, "batch")
            os.makedirs(wd, exist_ok=True)
            with open(os.path.join(wd, "batch.tsv"), "w") as f:
                for i in range(args.batch):
                    f.write("%s\t%s\n" % (bam, os.path.join(wd, "s%d" % i)))
            cmd = [HOST, "-i", bam, "--ref", os.path.join(tmp, "ref.fa"), "--out", os.path.join(wd, "unused")] + flags
            t = time.perf_counter()
            p = subprocess.run(cmd, cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                               env=dict(os.environ, CONGA_BATCH=os.path.join(wd, "batch.tsv"), CONGA_GPUS=str(args.gpus), CONGA_TIMING="1"))
            dt = time.perf_counter() - t
            if p.returncode != 0:
                sys.exit("batch failed: %s" % p.stderr.decode(errors="replace")[-1500:])
            marks = [float(l.split()[1]) for l in p.stderr.decode(errors="replace").split("\n") if l.startswith("[timing]") and "sample done" in l]
            same = all(open(os.path.join(wd, "s%d_%s" % (i, nm)), "rb").read() == ob["s_" + nm]
                       for i in range(args.batch) for nm in ("svs.bed", "dels.bed", "dups.bed") if ob["s_" + nm] is not None)
            res["batch"] = {"samples": args.batch, "total_s": round(dt, 3), "first_sample_done_s": marks[0] if marks else None,
                            "s_per_sample_after_first": round((marks[-1] - marks[0]) / max(1, len(marks) - 1), 4) if len(marks) > 1 else None,
                            "identical_to_single_runs": same}
        if not args.skip_reference:
            dt, orf = run(os.path.join(REF, "conga"), "ref")
            res["reference_s"] = round(dt, 3)
            res["reference_cores"] = 1
            res["reference_reads_per_s"] = round(wl.n_reads / dt, 1)
            res["identical_outputs"] = all(orf[k] == ob[k] for k in orf)
            res["rows"] = {k: (v.count(b"\n") if v else 0) for k, v in orf.items()}
            res["speedup_whole_program"] = round(dt / float(np.median(tb)), 1)
        print(json.dumps(res), flush=True)
    finally:
        if not args.keep:
            shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
