"""Golden GC profiles from the UNMODIFIED reference SONIC builder (oracle/_ref/sonic = sonic/sonic_exe.c +
libsonic.a compiled from /root/reference by oracle/Makefile).  Run in the build container:

  python tests/golden/make_golden_gc.py

Writes tests/golden/gc/ref_gc.npz: for each case the FASTA text (uint8) and, per record, the GC bytes the
builder stored in the .sonic container (read back with an independent parser of the container layout,
sonic/sonic.c:47-130 and sonic/sonic_reference.c:139-159).  The FASTA texts are small (<= 60 kB), so they
are committed with their outputs and the tests need neither the builder nor /root/reference.
"""
import gzip
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "oracle", "_ref")


def read_sonic_gc(path):
    """-> [(name, length, int8 gc bytes)] from a container whose gap / dup / repeat sections are empty."""
    with gzip.open(path, "rb") as f:
        d = f.read()
    o = 0
    magic, = struct.unpack_from("<i", d, o); o += 4
    assert magic == 42
    o += 8                                                  # build time
    n, = struct.unpack_from("<i", d, o); o += 4 + n         # info string
    n_chr, = struct.unpack_from("<i", d, o); o += 4
    chroms = []
    for _ in range(n_chr):
        n, = struct.unpack_from("<i", d, o); o += 4
        name = d[o:o + n].decode(); o += n
        L, = struct.unpack_from("<i", d, o); o += 4
        chroms.append((name, L))
    for _section in range(3):
        for _ in range(n_chr):
            cnt, = struct.unpack_from("<i", d, o); o += 4
            assert cnt == 0
    out = {}
    while True:
        idx, = struct.unpack_from("<i", d, o); o += 4
        if idx == 12345654:
            break
        e = o
        while struct.unpack_from("<b", d, e)[0] != -2:
            e += 1
        out[idx] = np.frombuffer(d[o:e], dtype=np.int8).copy()
        o = e + 1
    return [(nm, L, out.get(i, np.zeros(0, dtype=np.int8))) for i, (nm, L) in enumerate(chroms)]


def fasta_cases(rng):
    """name -> FASTA text.  Lengths around the 99 / 100 window edges, both cases, N runs, other IUPAC
    letters, ragged last lines, line widths that are and are not divisors of 100.  Every record has at least
    one complete window: the builder writes no terminator after a record without one (sonic_reference.c:61-64),
    which leaves a container its own reader misparses."""
    alpha = np.frombuffer(b"ACGTacgtNnRYKMSWryBDHV", dtype=np.uint8)
    p = np.array([6, 6, 6, 6, 3, 3, 3, 3, 2, 1] + [0.25] * 12)
    p = p / p.sum()

    def rec(name, L, width):
        s = alpha[rng.choice(len(alpha), size=L, p=p)].tobytes()
        lines = [s[i:i + width] for i in range(0, L, width)]
        return b">" + name.encode() + b" some description\n" + b"\n".join(lines) + (b"\n" if L else b"")
    cases = {}
    cases["edges"] = b"".join(rec("e%d" % L, L, 60) for L in (99, 100, 198, 199, 200, 299, 1000, 1099))
    cases["widths"] = b"".join(rec("w%d" % w, 5000 + w, w) for w in (50, 60, 61, 70, 100, 7))
    big = rec("big", 40123, 60)
    gcrich = b">gc\n" + b"\n".join([b"GC" * 30] * 40) + b"\n" + b">at\n" + b"\n".join([b"at" * 30] * 40) + b"\n"
    cases["mixed"] = big + gcrich + rec("tail", 777, 80)
    return cases


def main():
    rng = np.random.default_rng(20221124 + 8)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for nm in ("gaps.bed", "dups.bed"):
            open(os.path.join(tmp, nm), "w").close()
        with open(os.path.join(tmp, "rm.out"), "w") as f:
            f.write("   SW   perc perc perc  query     position in query    matching  repeat       position in repeat\n"
                    "score   div. del. ins.  sequence  begin end   (left)   repeat    class/family begin end (left) ID\n\n")
        for case, text in fasta_cases(rng).items():
            fa = os.path.join(tmp, case + ".fa")
            with open(fa, "wb") as f:
                f.write(text)
            # samtools-style .fai (name, length, offset, linebases, linewidth); the builder reads columns 1-2
            with open(fa + ".fai", "w") as f:
                off = 0
                for block in text.split(b">")[1:]:
                    head, _, body = block.partition(b"\n")
                    lines = body.split(b"\n")
                    L = sum(len(l) for l in lines)
                    w = len(lines[0]) if lines and lines[0] else 1
                    off += 1 + len(head) + 1
                    f.write("%s\t%d\t%d\t%d\t%d\n" % (head.split()[0].decode(), L, off, w, w + 1))
                    off += len(body)
            sonic = os.path.join(tmp, case + ".sonic")
            subprocess.run([os.path.join(REF, "sonic"), "--ref", fa, "--gaps", os.path.join(tmp, "gaps.bed"), "--dups",
                            os.path.join(tmp, "dups.bed"), "--reps", os.path.join(tmp, "rm.out"), "--make-sonic", sonic],
                           check=True, cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            got = read_sonic_gc(sonic)
            out[case + "/fasta"] = np.frombuffer(text, dtype=np.uint8)
            out[case + "/names"] = np.array([g[0] for g in got])
            out[case + "/lengths"] = np.array([g[1] for g in got], dtype=np.int64)
            for g in got:
                out[case + "/gc/" + g[0]] = g[2]
                print(case, g[0], g[1], len(g[2]), file=sys.stderr)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "gc", "ref_gc.npz"), **out)


if __name__ == "__main__":
    main()
