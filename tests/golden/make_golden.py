#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/conga_dump,
built by `make -C oracle ref` from /root/reference) on small synthetic inputs.

Only runs where /root/reference exists (this container); the .npz fixtures it writes are
committed and are what the GPU box and the CPU test-suite replay.  For each case it
  1. writes a FASTA + .fai, builds a SONIC file with the reference's own `sonic` builder,
  2. writes per-chromosome read records and turns them into a BAM + .bai (oracle/mkbam.c),
  3. writes del / dup / mappability BEDs,
  4. runs conga_dump (reference objects + --wrap dump shim) and collects: the int16 depth
     array, expected_read_depth[101], mean, read count, every per-CNV field, and the three
     BED outputs + conga.log,
  5. parses the SONIC container independently for the GC bytes the reference consumed.

usage: python tests/golden/make_golden.py            (the three base cases)
       python tests/golden/make_golden.py edge       (ref_edge only; its own seed, the others are not rewritten)
"""
import gzip
import os
import shutil
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from conga_b200.abi import CNV_DT, MAPIV_DT, READ_DT  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
DUMP_DT = np.dtype([("start", "<i4"), ("end", "<i4"), ("observed", "<i4"), ("expected", "<f4"),
                    ("rp", "<i4"), ("border_rp", "<i4"), ("copy_number", "<i4"), ("pad", "<i4"),
                    ("score", "<f8"), ("lhomo", "<f8"), ("lhetero", "<f8"), ("lnone", "<f8"),
                    ("mappability", "<f8")])


def write_fasta(path, chroms, rng):
    """chroms: list of (name, length).  Per-kilobase GC level U(.3,.6); one N-run per chromosome."""
    fai = []
    with open(path, "wb") as f:
        for name, L in chroms:
            level = np.repeat(rng.uniform(0.30, 0.60, size=(L + 999) // 1000), 1000)[:L]
            is_gc = rng.random(L) < level
            pick = rng.random(L) < 0.5
            seq = np.where(is_gc, np.where(pick, ord("G"), ord("C")), np.where(pick, ord("A"), ord("T"))).astype(np.uint8)
            lower = rng.random(L) < 0.1                         # soft-masked bases count too (case-insensitive)
            seq = np.where(lower, seq + 32, seq).astype(np.uint8)
            n0 = int(L * 0.4) // 100 * 100
            seq[n0:n0 + 5000] = ord("N")
            hdr = (">%s\n" % name).encode()
            f.write(hdr)
            off = f.tell()
            for i in range(0, L, 60):
                f.write(seq[i:i + 60].tobytes() + b"\n")
            fai.append("%s\t%d\t%d\t60\t61\n" % (name, L, off))
    with open(path + ".fai", "w") as f:
        f.writelines(fai)


def parse_sonic(path):
    """Independent reader of the SONIC container (layout: SURVEY Appendix B1)."""
    b = gzip.open(path, "rb").read()
    o = 0

    def i32():
        nonlocal o
        v = struct.unpack_from("<i", b, o)[0]
        o += 4
        return v
    assert i32() == 42
    o += 8
    skip = i32()                   # info string (note: `o += i32()` would drop i32's own advance)
    o += skip
    n = i32()
    names, lens = [], []
    for _ in range(n):
        ln = i32()
        names.append(b[o:o + ln].decode().rstrip("\0"))
        o += ln
        lens.append(i32())
    for _section in range(2):      # gaps, dups
        for _ in range(n):
            cnt = i32()
            o += 8 * cnt
    for _ in range(n):             # repeats
        for _ in range(i32()):
            o += 9
            tlen = i32()
            o += tlen
            clen = i32()
            o += clen
            o += 8
    gc = {}
    while True:
        idx = i32()
        if idx == 12345654:
            break
        nw = (lens[idx] + 1) // 100
        gc[idx] = np.frombuffer(b, dtype=np.int8, count=nw, offset=o).copy()
        o += nw
        assert struct.unpack_from("<b", b, o)[0] == -2
        o += 1
    return names, lens, gc


def run_case(tag, chroms, rng, n_reads_per_mb, extra_args, with_map, pileup=False):
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        fa = os.path.join(tmp, "ref.fa")
        write_fasta(fa, chroms, rng)
        for nm in ("gaps.bed", "dups.bed"):
            open(os.path.join(tmp, nm), "w").close()
        with open(os.path.join(tmp, "rm.out"), "w") as f:
            f.write("   SW   perc perc perc  query     position in query    matching  repeat       position in repeat\n"
                    "score   div. del. ins.  sequence  begin end   (left)   repeat    class/family begin end (left) ID\n\n")
        subprocess.run([os.path.join(REF, "sonic"), "--ref", fa, "--gaps", os.path.join(tmp, "gaps.bed"),
                        "--dups", os.path.join(tmp, "dups.bed"), "--reps", os.path.join(tmp, "rm.out"),
                        "--make-sonic", os.path.join(tmp, "ref.sonic")], check=True, cwd=tmp,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        names, lens, gcs = parse_sonic(os.path.join(tmp, "ref.sonic"))
        assert names == [c[0] for c in chroms] and lens == [c[1] for c in chroms]

        reads, dels, dups, maps = {}, {}, {}, {}
        tsv = []
        for k, (name, L) in enumerate(chroms):
            per_mb = n_reads_per_mb[k] if isinstance(n_reads_per_mb, (list, tuple)) else n_reads_per_mb
            n = int(L / 1e6 * per_mb)
            r = np.zeros(n, dtype=READ_DT)
            r["pos"] = np.sort(rng.integers(0, L - 100, size=n)).astype(np.int32)
            r["len"] = rng.integers(30, 101, size=n)
            r["mapq"] = rng.choice(np.array([0, 10, 20, 21, 30, 37, 60], dtype=np.uint8), size=n)
            r["fl"] = np.where(rng.random(n) < 0.05, 4, 0)
            if pileup and k == 0:
                # a tower of identical starts tall enough to wrap the reference's `short` (A1)
                tower = np.zeros(40000, dtype=READ_DT)
                tower["pos"] = 123457
                tower["len"] = 50
                tower["mapq"] = 60
                r = np.concatenate([r, tower])
                r = r[np.argsort(r["pos"], kind="stable")]
            reads[k] = np.ascontiguousarray(r)
            r.tofile(os.path.join(tmp, "reads.%d.bin" % k))
            tsv.append("%s\t%d\t%s\n" % (name, L, os.path.join(tmp, "reads.%d.bin" % k)))
            nd, nu = 40, 25
            for store, cnt in ((dels, nd), (dups, nu)):
                ln = np.exp(rng.uniform(np.log(1000), np.log(L // 5), size=cnt)).astype(np.int64)
                st = rng.integers(0, L - ln)
                c = np.zeros(cnt, dtype=CNV_DT)
                c["start"] = st
                c["end"] = st + ln
                c = np.unique(c)                     # structured sort = (start, end); drops duplicates
                store[k] = c
            # CNVs touching both chromosome ends (A23) and one below --min-sv-size (dropped, A9)
            dels[k] = np.unique(np.concatenate([dels[k], np.array([(0, 1500), (L - 2000, L), (5000, 5999)], dtype=CNV_DT)]))
            if with_map:
                b = np.concatenate([[0], np.cumsum(rng.integers(1, 3000, size=L // 1000))])
                b = b[b < L - 1]
                m = np.zeros(len(b) - 1, dtype=MAPIV_DT)
                m["start"] = b[:-1]
                m["end"] = b[1:]
                m["val"] = rng.choice(np.array([1, .5, .333333, .25, .2, .166667, .142857, .125, 0, 0.9, 0.37], dtype=np.float32), size=len(m))
                # a few out-of-order overlapping rows appended at the end of the file (later rows win, A10)
                extra = np.zeros(6, dtype=MAPIV_DT)
                extra["start"] = rng.integers(0, L - 20000, size=6)
                extra["end"] = extra["start"] + rng.integers(1, 15000, size=6)
                extra["val"] = [0.75, 0.1, 1.0, 0.0, 0.6, 0.45]
                m = np.concatenate([m, extra])
                m["line"] = np.arange(len(m))
                maps[k] = m
        with open(os.path.join(tmp, "chroms.tsv"), "w") as f:
            f.writelines(tsv)
        bam = os.path.join(tmp, "in.bam")
        subprocess.run([os.path.join(REF, "mkbam"), bam, os.path.join(tmp, "chroms.tsv")], check=True)

        def write_bed(path, store, fmt):
            with open(path, "w") as f:
                for k, (name, _L) in enumerate(chroms):
                    for row in store.get(k, []):
                        f.write(fmt(name, row))
        write_bed(os.path.join(tmp, "dels.bed"), dels, lambda n, r: "%s\t%d\t%d\n" % (n, r["start"], r["end"]))
        write_bed(os.path.join(tmp, "dups.bed"), dups, lambda n, r: "%s\t%d\t%d\n" % (n, r["start"], r["end"]))
        args = [os.path.join(REF, "conga_dump"), "-i", bam, "-o", os.path.join(tmp, "out", "g"), "--ref", fa,
                "--sonic", os.path.join(tmp, "ref.sonic"), "--dels", os.path.join(tmp, "dels.bed"),
                "--dups", os.path.join(tmp, "dups.bed")] + extra_args
        if with_map:
            # %g of a float32 prints enough digits for atof -> (float) to give the value back
            write_bed(os.path.join(tmp, "map.bed"), maps,
                      lambda n, r: "%s\t%d\t%d\t%.9g\n" % (n, r["start"], r["end"], float(r["val"])))
            args += ["--mappability", os.path.join(tmp, "map.bed")]
        os.makedirs(os.path.join(tmp, "out"))
        os.makedirs(os.path.join(tmp, "dump"))
        env = dict(os.environ, CONGA_DUMP_DIR=os.path.join(tmp, "dump"))
        subprocess.run(args, check=True, cwd=tmp, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)

        out = {"chrom_names": np.array([c[0] for c in chroms]), "chrom_lens": np.array([c[1] for c in chroms]),
               "args": np.array(extra_args + (["--mappability"] if with_map else [])), "with_map": np.array(with_map)}
        for nm in ("g_svs.bed", "g_dels.bed", "g_dups.bed"):
            out["file_" + nm[2:-4]] = np.frombuffer(open(os.path.join(tmp, "out", nm), "rb").read(), dtype=np.uint8)
        log = open(os.path.join(tmp, "conga.log")).read().splitlines(True)
        out["log_lines"] = np.array([l for l in log if l.startswith("Read Count")])
        for k, (name, L) in enumerate(chroms):
            if "X" in name or "Y" in name:
                assert not os.path.exists(os.path.join(tmp, "dump", "chr%d.meta" % k))
                continue
            meta = open(os.path.join(tmp, "dump", "chr%d.meta" % k)).read().split("\t")
            depth = np.fromfile(os.path.join(tmp, "dump", "chr%d.depth" % k), dtype=np.int16)
            assert len(depth) == L
            nz = np.flatnonzero(depth)
            out["c%d_gc" % k] = gcs[k]
            out["c%d_reads" % k] = reads[k]
            out["c%d_depth_idx" % k] = nz.astype(np.int32)
            out["c%d_depth_val" % k] = depth[nz]
            out["c%d_exp" % k] = np.fromfile(os.path.join(tmp, "dump", "chr%d.exp" % k), dtype=np.float32)
            out["c%d_mean" % k] = np.float32(meta[3])
            out["c%d_read_count" % k] = np.int64(meta[4])
            out["c%d_del" % k] = np.fromfile(os.path.join(tmp, "dump", "chr%d.del" % k), dtype=DUMP_DT)
            out["c%d_dup" % k] = np.fromfile(os.path.join(tmp, "dump", "chr%d.dup" % k), dtype=DUMP_DT)
            if with_map:
                out["c%d_maprows" % k] = maps[k]
        path = os.path.join(ROOT, "tests", "golden", tag + ".npz")
        np.savez_compressed(path, **out)
        print(tag, "->", path, os.path.getsize(path), "bytes")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    if not os.path.exists(os.path.join(REF, "conga_dump")):
        sys.exit("oracle/_ref/conga_dump missing: run `make -C oracle ref` (needs /root/reference)")
    if sys.argv[1:] == ["edge"]:
        # a chromosome without a single read (expected table all 0/0 -> 0, lambda 0 -> 0.01, every score -0.00),
        # a sparse one (most GC bins empty, most CNVs without reads; only mapq 60 passes --min-mapq 37) and a
        # name with a Y in it (skipped, A4)
        rng = np.random.default_rng(20221124 + 7)
        run_case("ref_edge", [("1", 150_000), ("MT", 120_000), ("GL_Y1", 50_000)], rng, [0, 2000, 500], ["--min-mapq", "37"], True)
        return
    rng = np.random.default_rng(20221124)
    run_case("ref_plain", [("1", 400_000), ("2", 250_000), ("X", 100_000)], rng, 14000, [], False)
    run_case("ref_map_mq20", [("1", 300_000), ("chr2", 200_000)], rng, 60000, ["--min-mapq", "20"], True)
    run_case("ref_wrap", [("1", 200_000)], rng, 3000, ["--c-score", "0.3"], False, pileup=True)


if __name__ == "__main__":
    main()
