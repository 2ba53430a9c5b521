"""Split-read support (--rp with --dups, BASELINE.json configs[3] shape) at scale, whole program on the same
files: reference `conga --rp` (one core) against `conga-b200 --rp`, outputs compared byte for byte.  One
chromosome of --mb megabases, --cov x coverage of 70-base reads with real bases (0.5 % substitutions), 0.2 %
chimeric reads across the breakpoints of the known CNVs.  Measurement tooling (may use oracle/), not product.

  python scripts/sr_bench.py [--mb 50] [--cov 5] [--skip-reference]
"""
import argparse, json, os, shutil, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import refdata
import cli_inputs
from conga_b200.abi import CNV_DT, READ_DT

REF = os.path.join(ROOT, "oracle", "_ref")
HOST = os.path.join(ROOT, "conga_b200", "host", "conga-b200")

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=float, default=50.0)
ap.add_argument("--cov", type=float, default=5.0)
ap.add_argument("--chroms", type=int, default=1)
ap.add_argument("--skip-reference", action="store_true")
ap.add_argument("--gpus", type=int, default=1)
args = ap.parse_args()
rng = np.random.default_rng(20221124 + 4)
tmp = tempfile.mkdtemp(prefix="srbench_")
try:
    t0 = time.perf_counter()
    L = int(args.mb * 1e6) // 1000 * 1000
    RL = 70
    names, seqs, gcs, tsv, del_rows, dup_rows = [], [], [], [], [], []
    n_reads = 0
    for k in range(args.chroms):
        name = str(k + 1)
        level = np.repeat(rng.uniform(0.30, 0.60, size=(L + 999) // 1000), 1000)[:L]
        is_gc = rng.random(L) < level
        pick = rng.random(L) < 0.5
        g = np.where(is_gc, np.where(pick, ord("G"), ord("C")), np.where(pick, ord("A"), ord("T"))).astype(np.uint8)
        g[int(L * 0.4):int(L * 0.4) + 20000] = ord("N")
        names.append(name); seqs.append(g)
        # SONIC GC bytes: window 0 = bases [0,99), window k = [100k-1, 100k+99)  (sonic_reference.c:82-100)
        isgc = np.concatenate([[0], ((g == ord("G")) | (g == ord("C"))).astype(np.int64)])
        cs = np.cumsum(isgc)
        nw = (L + 1) // 100
        ends = np.arange(1, nw + 1) * 100 - 1
        starts = np.concatenate([[0], ends[:-1]])
        gcs.append((cs[ends] - cs[starts]).astype(np.int8))
        n = int(L * args.cov / RL)
        cn = {}
        for key, rows, cnt in (("del", del_rows, 200), ("dup", dup_rows, 100)):
            size = np.exp(rng.uniform(np.log(1500), np.log(60000), size=cnt)).astype(np.int64)
            st = rng.integers(5000, L - size - 5000)
            c = np.unique(np.array(list(zip(st, st + size)), dtype=CNV_DT))
            cn[key] = c
            rows += ["%s\t%d\t%d\n" % (name, a, b) for a, b in zip(c["start"], c["end"])]
        pos = rng.integers(1, L - 200, size=n)
        idx = pos[:, None] + np.arange(RL)[None, :]
        h = RL // 2
        # 0.2 % chimeric: first half before a breakpoint, second half after its partner
        n_ch = int(n * 0.002)
        which = rng.integers(0, len(cn["del"]) + len(cn["dup"]), size=n_ch)
        for j, wsel in enumerate(which):
            if wsel < len(cn["del"]):
                a, b = cn["del"][wsel]["start"], cn["del"][wsel]["end"]
                pos[j] = a - h
                idx[j, :h] = np.arange(a - h, a); idx[j, h:] = np.arange(b, b + RL - h)
            else:
                a, b = cn["dup"][wsel - len(cn["del"])]["start"], cn["dup"][wsel - len(cn["del"])]["end"]
                pos[j] = b - h
                idx[j, :h] = np.arange(b - h, b); idx[j, h:] = np.arange(a, a + RL - h)
        order = np.argsort(pos, kind="stable")
        pos, idx = pos[order], idx[order]
        bases = g[idx]
        m = rng.random(bases.shape) < 0.005
        bases[m] = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=int(m.sum()))
        quals = rng.integers(20, 41, size=bases.shape).astype(np.uint8)
        r = np.zeros(n, dtype=READ_DT)
        r["pos"], r["len"] = pos, RL
        r["mapq"] = rng.choice(np.array([0, 10, 30, 37, 60], dtype=np.uint8), size=n, p=[.03, .07, .2, .3, .4])
        rp_, sp_ = os.path.join(tmp, "reads.%d.bin" % k), os.path.join(tmp, "seq.%d.bin" % k)
        r.tofile(rp_)
        np.concatenate([bases, quals], axis=1).tofile(sp_)
        tsv.append("%s\t%d\t%s\t%s\n" % (name, L, rp_, sp_))
        n_reads += n
    fa = os.path.join(tmp, "ref.fa")
    cli_inputs.write_fasta_from(fa, list(zip(names, seqs)), rng)
    refdata.write_sonic(os.path.join(tmp, "ref.sonic"), names, [L] * len(names), gcs)
    open(os.path.join(tmp, "chroms.tsv"), "w").writelines(tsv)
    bam = os.path.join(tmp, "in.bam")
    subprocess.run([os.path.join(REF, "mkbam"), bam, os.path.join(tmp, "chroms.tsv")], check=True, stdout=subprocess.DEVNULL)
    open(os.path.join(tmp, "dels.bed"), "w").writelines(del_rows)
    open(os.path.join(tmp, "dups.bed"), "w").writelines(dup_rows)
    t_prep = time.perf_counter() - t0
    flags = ["--sonic", os.path.join(tmp, "ref.sonic"), "--dels", os.path.join(tmp, "dels.bed"), "--dups", os.path.join(tmp, "dups.bed"),
             "--rp", "10", "--min-mapq", "20"]

    def run(exe, tag, env=None):
        wd = os.path.join(tmp, tag)
        os.makedirs(os.path.join(wd, "o"), exist_ok=True)
        t = time.perf_counter()
        p = subprocess.run([exe, "-i", bam, "--ref", fa, "--out", os.path.join(wd, "o", "s")] + flags, cwd=wd,
                           stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, env=dict(os.environ, **(env or {})))
        dt = time.perf_counter() - t
        if p.returncode != 0:
            sys.exit("%s failed: %s" % (tag, p.stderr.decode(errors="replace")[-1500:]))
        err = p.stderr.decode(errors="replace")
        out = {nm: open(os.path.join(wd, "o", "s_" + nm), "rb").read() for nm in ("svs.bed", "dels.bed", "dups.bed")}
        return dt, out, [l for l in err.split("\n") if l.startswith("[timing]") or "paired" in l]

    res = {"workload": "split reads: %d x %.0f Mb, %.1fx of %d-base reads, --rp 10 --min-mapq 20" % (args.chroms, L / 1e6, args.cov, RL),
           "reads": n_reads, "bam_bytes": os.path.getsize(bam), "prep_s": round(t_prep, 1)}
    dt, ob, tl = run(HOST, "b200", {"CONGA_TIMING": "1", "CONGA_GPUS": str(args.gpus)})
    dt2, ob, tl = run(HOST, "b200", {"CONGA_TIMING": "1", "CONGA_GPUS": str(args.gpus)})
    res["b200_s"] = [round(dt, 3), round(dt2, 3)]
    res["b200_timing"] = tl
    if not args.skip_reference:
        dtr, orf, tr = run(os.path.join(REF, "conga"), "ref")
        res["reference_s"] = round(dtr, 3)
        res["reference_note"] = tr
        res["identical_outputs"] = all(orf[k] == ob[k] for k in orf)
        res["max_read_pair"] = max(int(l.split(b"\t")[5]) for l in orf["dups.bed"].split(b"\n")[1:] if l)
        res["speedup_whole_program"] = round(dtr / min(dt, dt2), 1)
    print(json.dumps(res), flush=True)
finally:
    shutil.rmtree(tmp, ignore_errors=True)
