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


def read_fasta_upper(path):
    """{name: uppercase bases as a uint8 array} of a FASTA written by write_fasta."""
    out, name, chunks = {}, None, []
    for line in open(path, "rb"):
        if line.startswith(b">"):
            if name is not None:
                out[name] = np.frombuffer(b"".join(chunks).upper(), dtype=np.uint8)
            name, chunks = line[1:].strip().decode(), []
        else:
            chunks.append(line.strip())
    if name is not None:
        out[name] = np.frombuffer(b"".join(chunks).upper(), dtype=np.uint8)
    return out


def write_fasta_from(path, named_seqs, rng, width=60):
    """FASTA + .fai from uppercase base arrays; a tenth of the bases are written in lower case (soft-masked)."""
    fai = []
    with open(path, "wb") as f:
        for name, seq in named_seqs:
            out = seq.copy()
            lower = (rng.random(len(out)) < 0.1) & (out != ord("N"))
            out[lower] += 32
            f.write((">%s\n" % name).encode())
            off = f.tell()
            for i in range(0, len(out), width):
                f.write(out[i:i + width].tobytes() + b"\n")
            fai.append("%s\t%d\t%d\t%d\t%d\n" % (name, len(out), off, width, width + 1))
    with open(path + ".fai", "w") as f:
        f.writelines(fai)


def build_inputs_rp(tmp, chroms, rng, reads_per_mb=15000, n_dels=12, n_dups=10, junction_reads=25, with_map=False,
                    satellites=True, repeats=False):
    """Inputs for the split-read path (--rp with --dups): reads carry real bases copied from the reference
    (0.5 % substitutions) and phred bytes; every CNV gets `junction_reads` chimeric reads whose two halves
    sit on either side of its breakpoints, so READ_PAIR columns are non-zero.
    RepeatMasker rows of class Satellite gate some reads out (bam_data.c:207)."""
    fa = os.path.join(tmp, "ref.fa")
    write_fasta(fa, chroms, rng)
    seqs = read_fasta_upper(fa)
    if repeats:
        seqs = {k: v.copy() for k, v in seqs.items()}
        comp = np.zeros(256, dtype=np.uint8)
        for x, y in zip(b"ACGTN", b"TGCAN"):
            comp[x] = y
        for name, L in chroms:
            g = seqs[name]
            p0 = int(L * 0.15)                               # tandem repeat: 6 copies of 200 bases -> several forward hits per half
            for c in range(1, 6):
                g[p0 + 200 * c:p0 + 200 * (c + 1)] = g[p0:p0 + 200]
            p1 = int(L * 0.55)                               # 150 copies of 60 bases -> >= MAX_MAPPING hits: all dropped
            for c in range(1, 150):
                g[p1 + 60 * c:p1 + 60 * (c + 1)] = g[p1:p1 + 60]
            p2 = int(L * 0.8)                                # inverted repeat 5 kb away -> reverse-strand hits
            g[p2 + 5000:p2 + 5400] = comp[g[p2:p2 + 400]][::-1]
        write_fasta_from(fa, [(name, seqs[name]) for name, _L in chroms], rng)
    for nm in ("gaps.bed", "dups_sonic.bed"):
        open(os.path.join(tmp, nm), "w").close()
    with open(os.path.join(tmp, "rm.out"), "w") as f:
        f.write("   SW   perc perc perc  query     position in query    matching  repeat       position in repeat\n"
                "score   div. del. ins.  sequence  begin end   (left)   repeat    class/family begin end (left) ID\n\n")
        rid = 1
        for name, L in chroms:
            if not satellites:
                break
            for s in sorted(rng.integers(1000, L - 3000, size=4).tolist()):
                cls = "Satellite/centr" if rid % 2 else "LINE/L1"
                f.write("  900  10.0  0.5  0.5  %s  %d  %d  (%d)  +  REP%d  %s  1  %d  (0)  %d\n" % (name, s, s + 1500, L - s - 1500, rid, cls, 1500, rid))
                rid += 1
    sonic = os.path.join(tmp, "ref.sonic")
    subprocess.run([os.path.join(REF, "sonic"), "--ref", fa, "--gaps", os.path.join(tmp, "gaps.bed"), "--dups",
                    os.path.join(tmp, "dups_sonic.bed"), "--reps", os.path.join(tmp, "rm.out"), "--make-sonic", sonic],
                   check=True, cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    sub = np.frombuffer(b"ACGT", dtype=np.uint8)
    tsv, del_rows, dup_rows, map_rows = [], [], [], []
    for k, (name, L) in enumerate(chroms):
        ref = seqs[name]
        assert len(ref) == L
        n = int(L / 1e6 * reads_per_mb)
        pos = rng.integers(1, L - 200, size=n)
        ln = rng.integers(50, 101, size=n)                       # some at or below --min-read-length 60: not split
        recs = [(int(p), ref[p:p + l].copy()) for p, l in zip(pos, ln)]
        if repeats:
            for frac, span in ((0.15, 1200), (0.55, 9000), (0.8, 5400)):      # dense coverage of the repeat structures
                for p in rng.integers(int(L * frac) - 100, int(L * frac) + span, size=150):
                    l = int(rng.integers(62, 101))
                    recs.append((int(p), ref[p:p + l].copy()))
            for p in rng.integers(L - 140, L - 100, size=20):                 # reads that end at the chromosome end
                recs.append((int(p), ref[p:p + 100].copy()))
            for p in (1, 2, 3):                                                # and at its start (pos 0 is skipped by the reference)
                recs.append((p, ref[p:p + 80].copy()))
            recs.append((0, ref[0:80].copy()))
        cn = {}
        for key, rows, cnt in (("del", del_rows, n_dels), ("dup", dup_rows, n_dups)):
            size = np.exp(rng.uniform(np.log(1200), np.log(min(60000, L // 6)), size=cnt)).astype(np.int64)
            st = rng.integers(2000, L - size - 2000)
            c = np.unique(np.array(list(zip(st, st + size)), dtype=CNV_DT))
            cn[key] = c
            rows += ["%s\t%d\t%d\n" % (name, a, b) for a, b in zip(c["start"], c["end"])]
        for a, b in zip(cn["del"]["start"], cn["del"]["end"]):   # junction exactly at len/2, a few bases of slack in some
            for _ in range(junction_reads):
                l = int(rng.integers(62, 101)); h = l // 2; j = int(rng.integers(-3, 4))
                recs.append((int(a - h + j), np.concatenate([ref[a - h + j:a + j], ref[b + j:b + j + (l - h)]])))
        for a, b in zip(cn["dup"]["start"], cn["dup"]["end"]):
            for _ in range(junction_reads):
                l = int(rng.integers(62, 101)); h = l // 2; j = int(rng.integers(-3, 4))
                recs.append((int(b - h + j), np.concatenate([ref[b - h + j:b + j], ref[a + j:a + j + (l - h)]])))
        recs.sort(key=lambda t: t[0])
        r = np.zeros(len(recs), dtype=READ_DT)
        r["pos"] = [t[0] for t in recs]
        r["len"] = [len(t[1]) for t in recs]
        r["mapq"] = rng.choice(np.array([0, 10, 21, 30, 37, 60], dtype=np.uint8), size=len(recs), p=[.05, .05, .1, .2, .3, .3])
        r["fl"] = np.where(rng.random(len(recs)) < 0.04, 4, 0) | np.where(rng.random(len(recs)) < 0.02, 1, 0)
        path, spath = os.path.join(tmp, "reads.%d.bin" % k), os.path.join(tmp, "seq.%d.bin" % k)
        r.tofile(path)
        with open(spath, "wb") as f:
            for _p, b in recs:
                b = b.copy()
                m = rng.random(len(b)) < 0.005
                b[m] = rng.choice(sub, size=int(m.sum()))
                if repeats and rng.random() < 0.05:
                    b[rng.random(len(b)) < 0.03] = ord("N")
                q = rng.integers(2, 41, size=len(b)).astype(np.uint8)
                if rng.random() < 0.1:
                    q[:] = rng.integers(2, 25)                    # low-quality reads: the average-phred test bites
                f.write(b.tobytes() + q.tobytes())
        tsv.append("%s\t%d\t%s\t%s\n" % (name, L, path, spath))
        if with_map:
            bnd = np.concatenate([[0], np.cumsum(rng.integers(1, 3000, size=L // 1000))])
            bnd = bnd[bnd < L - 1]
            v = rng.choice(MAP_VALUES, size=len(bnd) - 1, p=[.6, .1, .05, .05, .04, .03, .03, .02, .03, .03, .02])
            map_rows += ["%s\t%d\t%d\t%.9g\n" % (name, s, e, float(x)) for s, e, x in zip(bnd[:-1], bnd[1:], v)]
    with open(os.path.join(tmp, "chroms.tsv"), "w") as f:
        f.writelines(tsv)
    bam = os.path.join(tmp, "in.bam")
    subprocess.run([os.path.join(REF, "mkbam"), bam, os.path.join(tmp, "chroms.tsv")], check=True, stdout=subprocess.DEVNULL)
    paths = {"ref": fa, "sonic": sonic, "bam": bam, "dels": os.path.join(tmp, "dels.bed"), "dups": os.path.join(tmp, "dups.bed")}
    open(paths["dels"], "w").writelines(del_rows)
    open(paths["dups"], "w").writelines(dup_rows)
    if with_map:
        paths["map"] = os.path.join(tmp, "map.bed")
        open(paths["map"], "w").writelines(map_rows)
    return paths
