"""On-disk inputs for the reference binaries under oracle/_ref (ref_path, conga).

TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/conga_oracle.c header): used by bench.py's
`--impl reference` leg and by tests.  Writes a SONIC container (layout: sonic/sonic.c:47-130,
sonic/sonic_reference.c:55-105 — magic, build time, info, chromosome table, empty gap / dup /
repeat sections, then the GC bytes) directly from GC arrays, per-chromosome cg_read record
files, and the del / dup / mappability BEDs of a synthetic workload.
"""
import gzip
import os
import struct
import time

import numpy as np


def write_sonic(path, names, lengths, gcs, info="conga-b200 synthetic"):
    with gzip.open(path, "wb", compresslevel=1) as f:
        f.write(struct.pack("<i", 42))
        f.write(struct.pack("<q", int(time.time())))
        b = info.encode()
        f.write(struct.pack("<i", len(b)) + b)
        f.write(struct.pack("<i", len(names)))
        for n, L in zip(names, lengths):
            nb = n.encode()
            f.write(struct.pack("<i", len(nb)) + nb + struct.pack("<i", L))
        for _section in range(3):                       # gaps, dups, repeats: none
            f.write(struct.pack("<i", 0) * len(names))
        for i, (L, gc) in enumerate(zip(lengths, gcs)):
            nw = (L + 1) // 100                          # sonic_reference.c stores floor((L+1)/100) bytes
            g = np.zeros(nw, dtype=np.int8)
            g[:min(nw, len(gc))] = gc[:nw]
            assert not np.any(g == -2)
            f.write(struct.pack("<i", i) + g.tobytes() + struct.pack("<b", -2))
        f.write(struct.pack("<i", 12345654))


def write_workload(dirpath, wl, chrom_subset=None):
    """Returns the ref_path argument list for this workload (without --first-chr/--last-chr/--out)."""
    os.makedirs(dirpath, exist_ok=True)
    chroms = wl.chroms
    write_sonic(os.path.join(dirpath, "ref.sonic"), [c.name for c in chroms], [c.length for c in chroms],
                [c.gc for c in chroms])
    sel = set(range(len(chroms))) if chrom_subset is None else set(chrom_subset)
    for k, c in enumerate(chroms):
        if k in sel:
            c.reads.tofile(os.path.join(dirpath, "reads.%d.bin" % c.index))
    args = ["--sonic", os.path.join(dirpath, "ref.sonic"), "--reads-dir", dirpath, "--min-sv-size", "1"]
    for key, flag in (("dels", "--dels"), ("dups", "--dups")):
        p = os.path.join(dirpath, key + ".bed")
        with open(p, "w") as f:
            for k, c in enumerate(chroms):
                if k in sel:
                    a = getattr(c, key)
                    f.write("".join("%s\t%d\t%d\n" % (c.name, s, e) for s, e in zip(a["start"].tolist(), a["end"].tolist())))
        args += [flag, p]
    if wl.has_mappability:
        p = os.path.join(dirpath, "map.bed")
        with open(p, "w") as f:
            for k, c in enumerate(chroms):
                if k in sel and c.maprows is not None:
                    r = c.maprows
                    f.write("".join("%s\t%d\t%d\t%.9g\n" % (c.name, s, e, v) for s, e, v in
                                    zip(r["start"].tolist(), r["end"].tolist(), r["val"].astype(np.float64).tolist())))
        args += ["--mappability", p]
    if wl.mq_threshold != -1:
        args += ["--min-mapq", str(wl.mq_threshold)]
    return args
