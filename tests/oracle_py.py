"""ctypes access to oracle/liboracle.so — the CPU restatement of the reference hot path.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under conga_b200/ may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "liboracle.so")

from conga_b200.abi import CALL_DT, CNV_DT, HALF_DT, MAPIV_DT, READ_DT, SPLITROW_DT, STATS_DT  # noqa: E402  (struct layouts only)


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        P = C.c_void_p
        lib.orc_count_reads.argtypes = [P, C.c_size_t, C.c_int, P, C.c_int32, P]
        lib.orc_calc_mu.argtypes = [P, C.c_int32, P, P]
        lib.orc_calc_mean_per_chr.argtypes = [P, C.c_int32, P, P, P, P]
        lib.orc_paint_mappability.argtypes = [P, C.c_int32, P, C.c_size_t]
        lib.orc_lpoisson.argtypes = [C.c_int, C.c_double]
        lib.orc_lpoisson.restype = C.c_double
        lib.orc_imax.argtypes = [C.c_double, C.c_double]
        lib.orc_imax.restype = C.c_int
        lib.orc_find_depths.argtypes = [P, P, P, P, P, C.c_size_t, C.c_int, P]
        lib.orc_sort_cnvs.argtypes = [P, C.c_size_t]
        lib.orc_gc_bin.argtypes = [P, C.c_int]
        lib.orc_gc_bin.restype = C.c_int
        lib.orc_chromosome.argtypes = [P, C.c_size_t, C.c_int, C.c_int32, P, P, C.c_size_t,
                                       P, C.c_size_t, P, P, C.c_size_t, P, P, P, P, P]

    def sr_rows(self, ref, halves, bases, chrom_len, mq_threshold):
        """split_read.c / bam_data.c:29-154 minus the satellite term -> (SPLITROW_DT rows, n_mappings)"""
        P = C.c_void_p
        self.lib.orc_sr_rows.argtypes = [P, C.c_int, P, C.c_size_t, P, C.c_int, C.c_int, P, C.c_size_t, C.POINTER(C.c_uint64)]
        self.lib.orc_sr_rows.restype = C.c_long
        ref = np.ascontiguousarray(ref, dtype=np.uint8)
        halves = np.ascontiguousarray(halves, dtype=HALF_DT)
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        cap = max(1024, 64 * len(halves))
        out = np.zeros(cap, dtype=SPLITROW_DT)
        nm = C.c_uint64(0)
        n = self.lib.orc_sr_rows(ref.ctypes.data, len(ref), halves.ctypes.data, len(halves), bases.ctypes.data, chrom_len, mq_threshold,
                                 out.ctypes.data, cap, C.byref(nm))
        assert n <= cap
        return out[:n], int(nm.value)

    def sr_count(self, rows, dels, dups):
        P = C.c_void_p
        self.lib.orc_sr_count.argtypes = [P, C.c_size_t, P, C.c_size_t, P, C.c_size_t, P, P]
        rows = np.ascontiguousarray(rows, dtype=SPLITROW_DT)
        dels = np.ascontiguousarray(dels, dtype=CNV_DT)
        dups = np.ascontiguousarray(dups, dtype=CNV_DT)
        b, r = np.zeros(len(dels), dtype=np.int32), np.zeros(len(dups), dtype=np.int32)
        self.lib.orc_sr_count(rows.ctypes.data, len(rows), dels.ctypes.data, len(dels), dups.ctypes.data, len(dups), b.ctypes.data, r.ctypes.data)
        return b, r

    def gc_profile(self, text):
        """sonic/sonic_reference.c:57-104 for the text of one FASTA record (line feeds allowed) -> int8 bytes"""
        buf = np.frombuffer(bytes(text), dtype=np.uint8)
        out = np.zeros(len(buf) // 99 + 2, dtype=np.int8)
        self.lib.orc_gc_profile.restype = C.c_long
        self.lib.orc_gc_profile.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_long]
        n = self.lib.orc_gc_profile(buf.ctypes.data if len(buf) else None, len(buf), out.ctypes.data, len(out))
        assert n <= len(out)
        return out[:n].copy()

    def count_reads(self, reads, mq_threshold, chrom_len, depth=None):
        """bam_data.c:201-217 -> (int16 depth, total_read_count)"""
        assert reads.dtype == READ_DT
        if depth is None:
            depth = np.zeros(chrom_len, dtype=np.int16)
        cnt = C.c_int32(0)
        self.lib.orc_count_reads(reads.ctypes.data, len(reads), mq_threshold, depth.ctypes.data, chrom_len, C.byref(cnt))
        return depth, cnt.value

    def calc_mu(self, depth):
        rd = C.c_int64(0)
        mean = C.c_float(0)
        self.lib.orc_calc_mu(depth.ctypes.data, len(depth), C.byref(rd), C.byref(mean))
        return rd.value, np.float32(mean.value)

    def calc_mean_per_chr(self, depth, gc):
        rd = np.zeros(101, dtype=np.int64)
        cnt = np.zeros(101, dtype=np.int64)
        exp = np.zeros(101, dtype=np.float32)
        assert gc.dtype == np.int8 and len(gc) >= (len(depth) + 99) // 100
        self.lib.orc_calc_mean_per_chr(depth.ctypes.data, len(depth), gc.ctypes.data, rd.ctypes.data, cnt.ctypes.data, exp.ctypes.data)
        return rd, cnt, exp

    def paint_mappability(self, chrom_len, rows):
        m = np.zeros(chrom_len, dtype=np.float32)
        assert rows.dtype == MAPIV_DT
        self.lib.orc_paint_mappability(m.ctypes.data, chrom_len, rows.ctypes.data, len(rows))
        return m

    def find_depths(self, depth, gc, expected, mapp, cnvs, is_dup):
        assert cnvs.dtype == CNV_DT
        out = np.zeros(len(cnvs), dtype=CALL_DT)
        expected = np.ascontiguousarray(expected, dtype=np.float32)
        self.lib.orc_find_depths(depth.ctypes.data, gc.ctypes.data, expected.ctypes.data,
                                 mapp.ctypes.data if mapp is not None else None,
                                 cnvs.ctypes.data, len(cnvs), int(is_dup), out.ctypes.data)
        return out

    def lpoisson(self, k, lam):
        return self.lib.orc_lpoisson(int(k), float(lam))

    def imax(self, a, b):
        return self.lib.orc_imax(float(a), float(b))

    def sort_cnvs(self, cnvs):
        c = np.ascontiguousarray(cnvs.copy())
        self.lib.orc_sort_cnvs(c.ctypes.data, len(c))
        return c

    def chromosome(self, chrom, mq_threshold=-1, has_map=False):
        """The whole per-chromosome path (what bench.py times as the CPU baseline of kind 'port')."""
        L = chrom.length
        depth = np.zeros(L, dtype=np.int16)
        rows = chrom.maprows if (has_map and chrom.maprows is not None) else np.zeros(0, dtype=MAPIV_DT)
        mapp = np.zeros(L, dtype=np.float32) if has_map else None
        stats = np.zeros(1, dtype=STATS_DT)
        exp = np.zeros(101, dtype=np.float32)
        d = np.zeros(len(chrom.dels), dtype=CALL_DT)
        u = np.zeros(len(chrom.dups), dtype=CALL_DT)
        self.lib.orc_chromosome(chrom.reads.ctypes.data, len(chrom.reads), mq_threshold, L, chrom.gc.ctypes.data,
                                rows.ctypes.data if len(rows) else None, len(rows),
                                chrom.dels.ctypes.data, len(chrom.dels), d.ctypes.data,
                                chrom.dups.ctypes.data, len(chrom.dups), u.ctypes.data,
                                depth.ctypes.data, mapp.ctypes.data if mapp is not None else None,
                                stats.ctypes.data, exp.ctypes.data)
        return depth, stats[0], exp, d, u


_inst = None


def load():
    global _inst
    if _inst is None:
        if not os.path.exists(SO):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "all"], check=True, stdout=subprocess.DEVNULL)
        _inst = Oracle(C.CDLL(SO))
    return _inst
