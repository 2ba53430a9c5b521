"""Shared helpers of the GC-profile tests: the golden fixture of the reference SONIC builder and the record
splitter that turns FASTA text into (name, bases without white space)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gc", "ref_gc.npz")


def records(text: bytes):
    """FASTA text -> [(name, record body as written, bases with white space removed)]"""
    out = []
    for block in bytes(text).split(b">")[1:]:
        head, _, body = block.partition(b"\n")
        out.append((head.split()[0].decode(), body, b"".join(body.split())))
    return out


def golden_cases():
    g = np.load(GOLDEN)
    for case in sorted({k.split("/")[0] for k in g.files}):
        for name, body, bases in records(g[case + "/fasta"].tobytes()):
            yield case, name, body, bases, g[case + "/gc/" + name]


def window_sums(bases: bytes) -> np.ndarray:
    """Independent NumPy statement of the window rule: byte w = number of G/C/g/c among bases [100w-1, 100w+99)."""
    b = np.frombuffer(bases, dtype=np.uint8)
    is_gc = np.isin(b, np.frombuffer(b"GCgc", dtype=np.uint8)).astype(np.int64)
    n_w = (len(b) + 1) // 100
    cs = np.concatenate([[0], np.cumsum(is_gc)])
    w = np.arange(n_w)
    lo = np.maximum(100 * w - 1, 0)
    hi = 100 * w + 99
    return (cs[hi] - cs[lo]).astype(np.int8)
