"""Input files for the command-line parity tests: FASTA + .fai -> SONIC (the reference's own builder,
oracle/_ref/sonic), read records -> coordinate-sorted BAM + index (oracle/_ref/mkbam over libhts), del /
dup / mappability BEDs.  Test infrastructure only; needs the prebuilt oracle/_ref binaries (built where
/root/reference exists, shipped to the GPU box with the snapshot)."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden import write_fasta  # noqa: E402
from conga_b200.abi import CNV_DT, READ_DT  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
MAP_VALUES = np.array([1, .5, .333333, .25, .2, .166667, .142857, .125, 0, 0.9, 0.37], dtype=np.float32)


def have_reference_binaries():
    return all(os.path.exists(os.path.join(REF, b)) for b in ("conga", "sonic", "mkbam"))


def build_inputs(tmp, chroms, rng, reads_per_mb=20000, with_map=False, bam_chroms=None, n_dels=30, n_dups=20, pileup=False):
    """chroms: [(name, length)] in SONIC order; bam_chroms: the subset that gets an @SQ line and reads
    (default all).  Returns the paths."""
    fa = os.path.join(tmp, "ref.fa")
    write_fasta(fa, chroms, rng)
    for nm in ("gaps.bed", "dups_sonic.bed"):
        open(os.path.join(tmp, nm), "w").close()
    with open(os.path.join(tmp, "rm.out"), "w") as f:
        f.write("   SW   perc perc perc  query     position in query    matching  repeat       position in repeat\n"
                "score   div. del. ins.  sequence  begin end   (left)   repeat    class/family begin end (left) ID\n\n")
    sonic = os.path.join(tmp, "ref.sonic")
    subprocess.run([os.path.join(REF, "sonic"), "--ref", fa, "--gaps", os.path.join(tmp, "gaps.bed"), "--dups",
                    os.path.join(tmp, "dups_sonic.bed"), "--reps", os.path.join(tmp, "rm.out"), "--make-sonic", sonic],
                   check=True, cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    tsv, del_rows, dup_rows, map_rows = [], [], [], []
    for k, (name, L) in enumerate(chroms):
        n = int(L / 1e6 * reads_per_mb)
        r = np.zeros(n, dtype=READ_DT)
        r["pos"] = np.sort(rng.integers(0, L - 100, size=n)).astype(np.int32)
        r["len"] = rng.integers(30, 101, size=n)
        r["mapq"] = rng.choice(np.array([0, 10, 20, 21, 30, 37, 60], dtype=np.uint8), size=n)
        r["fl"] = np.where(rng.random(n) < 0.05, 4, 0)
        if pileup and k == 0:
            tower = np.zeros(40000, dtype=READ_DT)
            tower["pos"], tower["len"], tower["mapq"] = 54321, 50, 60
            r = np.concatenate([r, tower])
            r = r[np.argsort(r["pos"], kind="stable")]
        cnvs = {}
        for key, rows, cnt in (("del", del_rows, n_dels), ("dup", dup_rows, n_dups)):
            ln = np.exp(rng.uniform(np.log(800), np.log(L // 5), size=cnt)).astype(np.int64)   # some below --min-sv-size
            st = rng.integers(0, L - ln)
            c = np.zeros(cnt, dtype=CNV_DT)
            c["start"], c["end"] = st, st + ln
            c = np.unique(c)
            cnvs[key] = c
            c = c[rng.permutation(len(c))]                    # BED rows are NOT sorted: the program sorts them
            rows += ["%s\t%d\t%d\textra_col\n" % (name, a, b) for a, b in zip(c["start"], c["end"])]
        # signal: every third deletion loses all of its reads, every third half; every second duplication doubles
        keep = np.ones(len(r), dtype=bool)
        for j, (a, b) in enumerate(zip(cnvs["del"]["start"], cnvs["del"]["end"])):
            inside = (r["pos"] >= a) & (r["pos"] < b)
            if j % 3 == 0:
                keep &= ~inside
            elif j % 3 == 1:
                keep &= ~(inside & (rng.random(len(r)) < 0.5))
        extra = [r[(r["pos"] >= a) & (r["pos"] < b)] for j, (a, b) in enumerate(zip(cnvs["dup"]["start"], cnvs["dup"]["end"])) if j % 2 == 0]
        r = np.concatenate([r[keep]] + extra)
        r = r[np.argsort(r["pos"], kind="stable")]
        if bam_chroms is None or name in bam_chroms:
            path = os.path.join(tmp, "reads.%d.bin" % k)
            np.ascontiguousarray(r).tofile(path)
            tsv.append("%s\t%d\t%s\n" % (name, L, path))
        del_rows += ["%s\t0\t1500\n" % name, "%s %d %d\n" % (name, L - 2000, L), "\n", "   \t \n"]   # ends, blanks, space-separated
        if with_map:
            b = np.concatenate([[0], np.cumsum(rng.integers(1, 3000, size=L // 1000))])
            b = b[b < L - 1]
            v = rng.choice(MAP_VALUES, size=len(b) - 1)
            map_rows += ["%s\t%d\t%d\t%.9g\n" % (name, s, e, float(x)) for s, e, x in zip(b[:-1], b[1:], v)]
            for _ in range(5):                                # out-of-order overlapping rows: later rows win
                s = int(rng.integers(0, L - 20000))
                map_rows.append("%s\t%d\t%d\t%.9g\n" % (name, s, s + int(rng.integers(1, 15000)), float(rng.choice(MAP_VALUES))))
    with open(os.path.join(tmp, "chroms.tsv"), "w") as f:
        f.writelines(tsv)
    bam = os.path.join(tmp, "in.bam")
    subprocess.run([os.path.join(REF, "mkbam"), bam, os.path.join(tmp, "chroms.tsv")], check=True, stdout=subprocess.DEVNULL)
    paths = {"ref": fa, "sonic": sonic, "bam": bam, "dels": os.path.join(tmp, "dels.bed"), "dups": os.path.join(tmp, "dups.bed")}
    # rows of different chromosomes interleaved, as a real call set sorted by something else would be
    for key, rows in (("dels", del_rows), ("dups", dup_rows)):
        order = rng.permutation(len(rows))
        with open(paths[key], "w") as f:
            f.writelines(rows[i] for i in order)
    if with_map:
        paths["map"] = os.path.join(tmp, "map.bed")
        with open(paths["map"], "w") as f:
            f.writelines(map_rows)
    return paths


def extra_bam(tmp, chroms, rng, tag, reads_per_mb=20000):
    """Another sample over the same chromosomes: reads only (uniform starts, random mapq), BAM + index."""
    tsv = []
    for k, (name, L) in enumerate(chroms):
        n = int(L / 1e6 * reads_per_mb)
        r = np.zeros(n, dtype=READ_DT)
        r["pos"] = np.sort(rng.integers(0, L - 100, size=n)).astype(np.int32)
        r["len"] = rng.integers(30, 101, size=n)
        r["mapq"] = rng.choice(np.array([0, 10, 20, 21, 30, 37, 60], dtype=np.uint8), size=n)
        path = os.path.join(tmp, "reads_%s.%d.bin" % (tag, k))
        r.tofile(path)
        tsv.append("%s\t%d\t%s\n" % (name, L, path))
    with open(os.path.join(tmp, "chroms_%s.tsv" % tag), "w") as f:
        f.writelines(tsv)
    bam = os.path.join(tmp, "in_%s.bam" % tag)
    subprocess.run([os.path.join(REF, "mkbam"), bam, os.path.join(tmp, "chroms_%s.tsv" % tag)], check=True, stdout=subprocess.DEVNULL)
    return bam
