"""Host-side half of the split-read path, restated in plain Python for the tests: the gate of
bam_data.c:207, find_split_reads' average-phred tests and half extraction (split_read.c:205-336),
sonic_is_satellite (sonic/sonic_interval.c:51-184, :276-293), and read_SplitReads' satellite term
and widening (bam_data.c:101-125).  TEST INFRASTRUCTURE ONLY — small inputs, Python loops."""
import numpy as np

from conga_b200.abi import HALF_DT


def _intersects(ps, pe, s, e):                       # sonic_this_interval_intersects
    return (ps >= s and pe < e) or (ps <= s and pe > e) or (ps <= s and pe >= s) or (ps <= e and pe > e)


def _fraction(s1, e1, s2, e2):                       # intersection_fraction, float32 arithmetic
    if s1 >= s2 and e1 <= e2:
        return np.float32(1.0)
    if s1 >= s2 and e1 >= e2:
        return np.float32(e2 - s1) / np.float32(e1 - s1)
    if s1 <= s2 and e1 <= e2:
        return np.float32(e1 - s2) / np.float32(e1 - s1)
    return np.float32(0.0)


def is_satellite(repeats, ps, pe):
    """repeats: int32[n][3] = (start, end, class-or-type contains "Satel"), sorted by start."""
    n = len(repeats)
    start, end = 0, n - 1
    med = (start + end) // 2
    hit = None
    while True:                                      # sonic_intersect's search, sonic_interval.c:88-121
        if start > end:
            break
        if _intersects(ps, pe, repeats[med][0], repeats[med][1]):
            hit = med
            break
        if start == med or end == med:
            if _intersects(ps, pe, repeats[start][0], repeats[start][1]):
                hit = start
            elif _intersects(ps, pe, repeats[end][0], repeats[end][1]):
                hit = end
            break
        if ps < repeats[med][0]:
            end = med
        else:
            start = med
        med = (start + end) // 2
    if hit is None or not repeats[hit][2]:
        return np.float32(0.0)
    return _fraction(ps, pe, int(repeats[hit][0]), int(repeats[hit][1]))


def halves_of(reads, seqqual, repeats, mq_threshold, min_read_length):
    """-> (HALF_DT array, bases as uint8 array).  seqqual: per record `len` bases then `len` phred bytes."""
    halves, bases = [], []
    off = 0
    n_bases = 0
    for r in reads:
        l = int(r["len"])
        seq = seqqual[off:off + l]
        q = seqqual[off + l:off + 2 * l]
        off += 2 * l
        pos, mapq, fl = int(r["pos"]), int(r["mapq"]), int(r["fl"])
        if not mapq > mq_threshold:                                   # bam_data.c:205
            continue
        if not l > min_read_length or (fl & 0b1111) != 0:             # l_qseq test; is_proper: secondary/qcfail/dup/supplementary
            continue
        if is_satellite(repeats, pos, pos + 20) != 0:
            continue
        if pos == 0:                                                  # split_read.c:215
            continue
        h = l // 2
        avg = np.float32(0)
        for x in q[h:]:
            avg = np.float32(avg + np.float32(x))
        avg = np.float32(avg / np.float32(l - h))
        if int(np.floor(avg)) < mq_threshold:
            continue
        halves.append((pos, n_bases, l - h, l, mapq, 0, 0))
        bases.append(seq[h:])
        n_bases += l - h
        for x in q[:h]:                                               # the average is not reset (split_read.c:301)
            avg = np.float32(avg + np.float32(x))
        avg = np.float32(avg / np.float32(h))
        if int(np.floor(avg)) < mq_threshold:
            continue
        halves.append((pos + h, n_bases, h, l, mapq, 1, 0))
        bases.append(seq[:h])
        n_bases += h
    hv = np.array(halves, dtype=HALF_DT) if halves else np.zeros(0, dtype=HALF_DT)
    return hv, (np.concatenate(bases) if bases else np.zeros(0, dtype=np.uint8))


def filter_and_widen(rows, repeats):
    """read_SplitReads: both ends outside satellites (bam_data.c:101-104), then +-50 (bam_data.c:112-125)."""
    keep = [i for i, r in enumerate(rows)
            if np.float32(is_satellite(repeats, int(r["pos"]), int(r["pos"]) + 1) + is_satellite(repeats, int(r["pos_map"]), int(r["pos_map"]) + 1)) == 0]
    out = rows[keep].copy()
    out["left_start"] -= 50
    out["left_end"] -= 50
    out["right_start"] += 50
    out["right_end"] += 50
    return out
