"""ctypes binding of ``libconga_gpu.so`` (``include/conga_gpu.h``).

This is the only way Python code in this repository reaches the hot path: every
call below goes through the C-ABI a C host would bind.  There is no Python or
NumPy implementation of any stage here and none is substituted when the library
or a GPU is missing — ``load_library`` raises, and so does every failed call.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CG_LIB") or os.path.join(_HERE, "libconga_gpu.so")   # CG_LIB: tuning builds only

CG_OK, CG_EINVAL, CG_ECUDA, CG_ENOMEM, CG_ERANGE = 0, -1, -2, -3, -4
CG_PTR_HOST, CG_PTR_DEVICE = 0, 1
CG_GC_BINS = 101
CG_NSLOTS = 4
CG_TILE = 16384
CG_CHUNK = 2048
CG_JMAX = 32
CG_COMM_ID_BYTES = 128
CG_PEER_HANDLE_BYTES = 64
CG_JOB_ANNOTATION_RESIDENT = 1

# struct layouts of include/conga_gpu.h as NumPy dtypes
READ_DT = np.dtype([("pos", "<i4"), ("len", "<u2"), ("mapq", "u1"), ("fl", "u1")])
CNV_DT = np.dtype([("start", "<i4"), ("end", "<i4")])
MAPIV_DT = np.dtype([("start", "<i4"), ("end", "<i4"), ("val", "<f4"), ("line", "<u4")])
MAPRUN_DT = np.dtype([("start", "<i4"), ("end", "<i4"), ("val", "<f4"), ("pad", "<i4")])
CALL_DT = np.dtype([("observed", "<i4"), ("expected", "<f4"), ("mappability", "<f8"),
                    ("lhomo", "<f8"), ("lhetero", "<f8"), ("lnone", "<f8"), ("score", "<f8"),
                    ("copy_number", "<i4"), ("margin_flags", "<u4")])
STATS_DT = np.dtype([("rd_per_gc", "<i8", (CG_GC_BINS,)), ("window_per_gc", "<i8", (CG_GC_BINS,)),
                     ("rd_cnt", "<i8"), ("reads_passed", "<i8")])
HALF_DT = np.dtype([("pos", "<i4"), ("seq_off", "<u4"), ("len", "<u2"), ("read_len", "<u2"), ("mapq", "u1"), ("is_read2", "u1"), ("pad", "<u2")])
SPLITROW_DT = np.dtype([("pos", "<i4"), ("pos_map", "<i4"), ("left_start", "<i4"), ("left_end", "<i4"), ("right_start", "<i4"),
                        ("right_end", "<i4"), ("sv_type", "u1"), ("pad", "u1", (7,))])
# cg_chrom_job: pointers as integers (host or device addresses, whatever the pointer mode says)
CHROM_JOB_DT = np.dtype([("chr_index", "<i4"), ("chrom_len", "<i4"), ("gc", "<u8"), ("n_windows", "<i4"), ("mq_threshold", "<i4"),
                         ("reads", "<u8"), ("n_reads", "<u8"), ("runs", "<u8"), ("n_runs", "<u8"), ("use_mappability", "<i4"), ("flags", "<i4"),
                         ("dels", "<u8"), ("n_del", "<u8"), ("del_calls", "<u8"), ("dups", "<u8"), ("n_dup", "<u8"), ("dup_calls", "<u8"),
                         ("stats", "<u8"), ("reads_pos", "<u8"), ("reads_mapq", "<u8"),
                         ("reads_delta", "<u8"), ("reads_big", "<u8"), ("n_big", "<u8"), ("reads_anchor", "<u8")])
assert CHROM_JOB_DT.itemsize == 168
ANCHOR_DT = np.dtype([("pos_before", "<i4"), ("big_before", "<u4")])
CG_DELTA_BLOCK = 1024
assert HALF_DT.itemsize == 16 and SPLITROW_DT.itemsize == 32
assert READ_DT.itemsize == 8 and CNV_DT.itemsize == 8 and MAPIV_DT.itemsize == 16
assert MAPRUN_DT.itemsize == 16 and CALL_DT.itemsize == 56 and STATS_DT.itemsize == 204 * 8

# every symbol include/conga_gpu.h declares: name -> (restype, argtypes)
_P, _I, _Z = C.c_void_p, C.c_int, C.c_size_t
SYMBOLS = {
    "cg_abi_version": (_I, []),
    "cg_create": (_P, [_I]),
    "cg_destroy": (None, [_P]),
    "cg_last_error": (C.c_char_p, [_P]),
    "cg_set_pointer_mode": (_I, [_P, _I]),
    "cg_set_async": (_I, [_P, _I]),
    "cg_set_slots": (_I, [_P, _I]),
    "cg_join": (_I, [_P]),
    "cg_alloc_pinned": (_P, [_P, _Z]),
    "cg_free_pinned": (None, [_P, _P]),
    "cg_chrom_begin": (_I, [_P, _I, C.c_int32, _P, C.c_int32, _I]),
    "cg_push_reads": (_I, [_P, _P, _Z]),
    "cg_push_reads_compact": (_I, [_P, _P, _P, _Z]),
    "cg_chrom_finish": (_I, [_P, _P]),
    "cg_get_expected": (_I, [_P, _P]),
    "cg_set_expected": (_I, [_P, _P]),
    "cg_set_mappability": (_I, [_P, _P, _Z]),
    "cg_flatten_mappability": (_I, [_P, _Z, C.c_int32, C.POINTER(_P), C.POINTER(_Z)]),
    "cg_free_runs": (None, [_P]),
    "cg_set_mappability_runs": (_I, [_P, _P, _Z, _I]),
    "cg_genotype": (_I, [_P, _P, _Z, _I, _P]),
    "cg_sr_reference": (_I, [_P, _P, C.c_int32]),
    "cg_sr_push": (_I, [_P, _P, _Z, _P, _Z]),
    "cg_sr_rows": (_I, [_P, _P, _Z, C.POINTER(_Z), C.POINTER(C.c_uint64)]),
    "cg_sr_count": (_I, [_P, _P, _Z, _P, _Z, _P, _Z, _P, _P]),
    "cg_gc_profile": (_I, [_P, _P, C.c_int64, _P, C.c_int32]),
    "cg_depth_stats_noalloc": (_I, [_I, _P, _P, _P, C.c_size_t, _I, C.c_int32, _P, C.c_int32, _P, _P, _P, _P]),
    "cg_run_chromosomes": (_I, [_P, _P, _Z]),
    "cg_encode_positions_delta": (_I, [_P, _Z, _P, _P, _Z, C.POINTER(_Z), _P]),
    "cg_sync": (_I, [_P]),
    "cg_download_depth": (_I, [_P, _P]),
    "cg_download_prefix": (_I, [_P, _P]),
    "cg_device_arrays": (_I, [_P, C.POINTER(_P), C.POINTER(_P)]),
    "cg_stats_table": (_I, [_P, C.POINTER(_P), C.POINTER(C.c_int32)]),
    "cg_stats_reset": (_I, [_P]),
    "cg_stats_bank": (_I, [_P, _I]),
    "cg_stats_row": (_I, [_P, _I, _P]),
    "cg_peer_export": (_I, [_P, _P]),
    "cg_peer_connect": (_I, [_P, _I, _I, _P]),
    "cg_peer_connect_local": (_I, [C.POINTER(_P), _I]),
    "cg_peer_disconnect": (_I, [_P]),
    "cg_comm_unique_id": (_I, [_P]),
    "cg_comm_init_rank": (_I, [_P, _I, _I, _P]),
    "cg_comm_init_all": (_I, [C.POINTER(_P), _I]),
    "cg_comm_destroy": (_I, [_P]),
    "cg_stats_allreduce": (_I, [_P, C.c_int32]),
    "cg_stats_allreduce_all": (_I, [C.POINTER(_P), _I, C.c_int32]),
    "cg_set_profiling": (_I, [_P, _I]),
    "cg_kernel_times": (_I, [_P, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "cg_launch_count": (C.c_int64, [_P]),
    "cg_stream": (_P, [_P]),
}


class CongaGpuError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libconga_gpu error {code}: {message}")
        self.code = code


_lib = None


def load_library(path: Optional[str] = None):
    """dlopen libconga_gpu.so and type every entry point.  Raises if it is absent:
    the product has no other implementation to fall back to."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "or `make -C conga_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def _ptr(x) -> int:
    """Address of a NumPy array (host), a torch tensor (device or host) or a raw int."""
    if x is None:
        return 0
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return int(x.data_ptr())
    raise TypeError(f"cannot take the address of {type(x)}")


class Context:
    """One ``cg_ctx`` (one GPU).  Methods mirror the C entry points one to one; NumPy
    arrays are passed in HOST pointer mode, torch CUDA tensors / raw device addresses in
    DEVICE pointer mode."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.h = self.lib.cg_create(device)
        if not self.h:
            raise CongaGpuError(CG_ECUDA, self.lib.cg_last_error(None).decode())
        self.device = device
        self.mode = CG_PTR_HOST
        self.chrom_len = 0

    # -- plumbing -----------------------------------------------------------
    def close(self):
        if getattr(self, "h", None):
            self.lib.cg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, rc: int, allow=()):
        if rc != CG_OK and rc not in allow:
            raise CongaGpuError(rc, self.lib.cg_last_error(self.h).decode())
        return rc

    def set_pointer_mode(self, mode: int):
        self._ck(self.lib.cg_set_pointer_mode(self.h, mode))
        self.mode = mode

    def set_async(self, on: bool):
        self._ck(self.lib.cg_set_async(self.h, int(on)))
        self.is_async = bool(on)

    def join(self):
        self._ck(self.lib.cg_join(self.h))

    def sync(self):
        self._ck(self.lib.cg_sync(self.h))

    @property
    def stream(self) -> int:
        return int(self.lib.cg_stream(self.h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self.lib.cg_launch_count(self.h))

    def alloc_pinned(self, nbytes: int) -> int:
        p = self.lib.cg_alloc_pinned(self.h, nbytes)
        if not p:
            raise CongaGpuError(CG_ENOMEM, self.lib.cg_last_error(self.h).decode())
        return int(p)

    def free_pinned(self, p: int):
        self.lib.cg_free_pinned(self.h, p)

    def pinned_array(self, n: int, dtype) -> np.ndarray:
        """A NumPy view of n elements of page-locked memory owned by the library."""
        dt = np.dtype(dtype)
        p = self.alloc_pinned(max(1, n * dt.itemsize))
        buf = (C.c_char * (n * dt.itemsize)).from_address(p)
        arr = np.frombuffer(buf, dtype=dt, count=n)
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        return arr

    # -- the per-chromosome path ---------------------------------------------
    def chrom_begin(self, chr_index: int, chrom_len: int, gc_profile, mq_threshold: int = -1, n_windows=None):
        if isinstance(gc_profile, np.ndarray):
            assert gc_profile.dtype == np.int8 and gc_profile.flags.c_contiguous
            n_windows = len(gc_profile) if n_windows is None else n_windows
        elif n_windows is None:
            n_windows = int(gc_profile.numel())
        self._keep_gc = gc_profile
        self._ck(self.lib.cg_chrom_begin(self.h, chr_index, chrom_len, _ptr(gc_profile), n_windows, mq_threshold))
        self.chrom_len = chrom_len

    def push_reads(self, reads, n: Optional[int] = None):
        if isinstance(reads, np.ndarray):
            assert reads.dtype == READ_DT and reads.flags.c_contiguous
            n = len(reads) if n is None else n
        assert n is not None
        self._ck(self.lib.cg_push_reads(self.h, _ptr(reads), n))

    def push_reads_compact(self, pos, mapq=None, n: Optional[int] = None):
        """cg_push_reads_compact: an int32 position plane and a uint8 mapq plane (None: no plane, needs a
        negative mq_threshold)."""
        if isinstance(pos, np.ndarray):
            assert pos.dtype == np.int32 and pos.flags.c_contiguous
            assert mapq is None or (mapq.dtype == np.uint8 and mapq.flags.c_contiguous and len(mapq) == len(pos))
            n = len(pos) if n is None else n
        assert n is not None
        self._ck(self.lib.cg_push_reads_compact(self.h, _ptr(pos), _ptr(mapq), n))

    def chrom_finish(self, out=None, allow_range: bool = False):
        """HOST mode: returns the cg_chrom_stats record.  DEVICE mode: `out` is a device address or None."""
        if self.mode == CG_PTR_HOST:
            if getattr(self, "is_async", False):
                # cg_set_async: the pass writes the row into `out` itself, so it has to be page-locked
                # (Context.pinned_array) and is valid after sync(); a pageable NumPy array would fault
                if out is None:
                    raise CongaGpuError(CG_EINVAL, "asynchronous HOST mode: chrom_finish needs a pinned_array(1, STATS_DT) to write to")
                self._ck(self.lib.cg_chrom_finish(self.h, _ptr(out)))
                return None
            st = np.zeros(1, dtype=STATS_DT)
            self._ck(self.lib.cg_chrom_finish(self.h, st.ctypes.data), allow=(CG_ERANGE,) if allow_range else ())
            return st[0]
        self._ck(self.lib.cg_chrom_finish(self.h, _ptr(out)))
        return None

    def get_expected(self) -> np.ndarray:
        assert self.mode == CG_PTR_HOST
        e = np.zeros(CG_GC_BINS, dtype=np.float32)
        self._ck(self.lib.cg_get_expected(self.h, e.ctypes.data))
        return e

    def set_expected(self, expected):
        if isinstance(expected, np.ndarray):
            expected = np.ascontiguousarray(expected, dtype=np.float32)
            assert expected.shape == (CG_GC_BINS,)
        self._keep_exp = expected
        self._ck(self.lib.cg_set_expected(self.h, _ptr(expected)))
        if self.mode == CG_PTR_HOST:
            self.sync()

    def set_mappability(self, rows: Optional[np.ndarray]):
        if rows is None or len(rows) == 0:
            self._ck(self.lib.cg_set_mappability(self.h, None, 0))
            return
        assert rows.dtype == MAPIV_DT and rows.flags.c_contiguous
        self._ck(self.lib.cg_set_mappability(self.h, rows.ctypes.data, len(rows)))

    def set_mappability_runs(self, runs, n: Optional[int] = None, enabled: bool = True):
        if isinstance(runs, np.ndarray):
            assert runs.dtype == MAPRUN_DT and runs.flags.c_contiguous
            n = len(runs)
        self._keep_runs = runs
        self._ck(self.lib.cg_set_mappability_runs(self.h, _ptr(runs), n or 0, 1 if enabled else 0))
        if self.mode == CG_PTR_HOST:
            self.sync()

    def genotype(self, cnvs, is_dup: bool, out=None, n: Optional[int] = None, allow_range: bool = False):
        if self.mode == CG_PTR_HOST:
            assert cnvs.dtype == CNV_DT and cnvs.flags.c_contiguous
            res = np.zeros(len(cnvs), dtype=CALL_DT)
            if len(cnvs):
                self._ck(self.lib.cg_genotype(self.h, cnvs.ctypes.data, len(cnvs), int(is_dup), res.ctypes.data),
                         allow=(CG_ERANGE,) if allow_range else ())
            return res
        assert n is not None
        self._ck(self.lib.cg_genotype(self.h, _ptr(cnvs), n, int(is_dup), _ptr(out)))
        return None

    # -- diagnostics -----------------------------------------------------------
    def download_depth(self) -> np.ndarray:
        d = np.empty(self.chrom_len, dtype=np.int16)
        self._ck(self.lib.cg_download_depth(self.h, d.ctypes.data))
        return d

    def download_prefix(self) -> np.ndarray:
        d = np.empty(self.chrom_len, dtype=np.int32)
        self._ck(self.lib.cg_download_prefix(self.h, d.ctypes.data))
        return d

    def device_arrays(self):
        """Device addresses of the current chromosome's int16 depth array and float expected[101] table."""
        a, b = C.c_void_p(), C.c_void_p()
        self._ck(self.lib.cg_device_arrays(self.h, C.byref(a), C.byref(b)))
        return int(a.value or 0), int(b.value or 0)

    def stats_table(self):
        p, n = C.c_void_p(), C.c_int32()
        self._ck(self.lib.cg_stats_table(self.h, C.byref(p), C.byref(n)))
        return int(p.value or 0), int(n.value)

    def stats_reset(self):
        self._ck(self.lib.cg_stats_reset(self.h))

    # -- split reads (--rp) ---------------------------------------------------
    def sr_reference(self, ref: np.ndarray):
        ref = np.ascontiguousarray(ref, dtype=np.uint8)
        self._keep_ref = ref
        self._ck(self.lib.cg_sr_reference(self.h, ref.ctypes.data if len(ref) else C.cast(C.c_char_p(b""), C.c_void_p).value, len(ref)))

    def sr_push(self, halves: np.ndarray, bases: np.ndarray):
        halves = np.ascontiguousarray(halves, dtype=HALF_DT)
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        if len(halves):
            self._ck(self.lib.cg_sr_push(self.h, halves.ctypes.data, len(halves), bases.ctypes.data, len(bases)))

    def sr_rows(self):
        n, nm = C.c_size_t(0), C.c_uint64(0)
        self._ck(self.lib.cg_sr_rows(self.h, None, 0, C.byref(n), C.byref(nm)))
        out = np.zeros(n.value, dtype=SPLITROW_DT)
        if n.value:
            self._ck(self.lib.cg_sr_rows(self.h, out.ctypes.data, len(out), C.byref(n), C.byref(nm)))
        return out, int(nm.value)

    def sr_count(self, rows: np.ndarray, dels: np.ndarray, dups: np.ndarray):
        rows = np.ascontiguousarray(rows, dtype=SPLITROW_DT)
        dels = np.ascontiguousarray(dels, dtype=CNV_DT)
        dups = np.ascontiguousarray(dups, dtype=CNV_DT)
        b, r = np.zeros(len(dels), dtype=np.int32), np.zeros(len(dups), dtype=np.int32)
        self._ck(self.lib.cg_sr_count(self.h, rows.ctypes.data if len(rows) else None, len(rows), dels.ctypes.data if len(dels) else None, len(dels),
                                      dups.ctypes.data if len(dups) else None, len(dups), b.ctypes.data if len(b) else None, r.ctypes.data if len(r) else None))
        return b, r

    def run_chromosomes(self, jobs: np.ndarray):
        """cg_run_chromosomes: `jobs` is a CHROM_JOB_DT array (see chrom_job); one foreign call for all of them."""
        assert jobs.dtype == CHROM_JOB_DT and jobs.flags.c_contiguous
        self._ck(self.lib.cg_run_chromosomes(self.h, jobs.ctypes.data, len(jobs)))

    def gc_profile(self, bases, n_bases=None, out=None):
        """SONIC's GC bytes of one chromosome from its bases (cg_gc_profile).  HOST mode: bytes / uint8
        array in, int8 array out.  DEVICE mode: uint8 tensor in, `out` an int8 tensor of (n + 1) // 100."""
        if isinstance(bases, (bytes, bytearray)):
            bases = np.frombuffer(bases, dtype=np.uint8)
        n = len(bases) if n_bases is None else int(n_bases)
        need = (n + 1) // 100
        if out is None:
            out = np.empty(need, dtype=np.int8)
        self._ck(self.lib.cg_gc_profile(self.h, _ptr(bases) if n else None, n, _ptr(out) if need else None, need))
        return out

    def set_slots(self, n: int):
        """cg_set_slots: chromosomes in flight (2..6)."""
        self._ck(self.lib.cg_set_slots(self.h, int(n)))

    def stats_bank(self, bank: int):
        self._ck(self.lib.cg_stats_bank(self.h, int(bank)))

    def comm_init_rank(self, n_ranks: int, rank: int, uid: bytes):
        """cg_comm_init_rank: join the communicator identified by `uid` (comm_unique_id() of one rank)."""
        assert len(uid) == CG_COMM_ID_BYTES
        buf = C.create_string_buffer(uid, CG_COMM_ID_BYTES)
        self._ck(self.lib.cg_comm_init_rank(self.h, n_ranks, rank, C.cast(buf, C.c_void_p)))

    def peer_export(self) -> bytes:
        """cg_peer_export: the CUDA IPC handle of this context's reduced statistics tables (64 bytes)."""
        buf = C.create_string_buffer(CG_PEER_HANDLE_BYTES)
        self._ck(self.lib.cg_peer_export(self.h, C.cast(buf, C.c_void_p)))
        return buf.raw

    def peer_connect(self, n_ranks: int, rank: int, handles):
        """cg_peer_connect: `handles` = peer_export() of every rank, in rank order (one process per GPU)."""
        blob = b"".join(handles)
        assert len(blob) == n_ranks * CG_PEER_HANDLE_BYTES
        buf = C.create_string_buffer(blob, len(blob))
        self._ck(self.lib.cg_peer_connect(self.h, n_ranks, rank, C.cast(buf, C.c_void_p)))

    def peer_disconnect(self):
        self._ck(self.lib.cg_peer_disconnect(self.h))

    def comm_destroy(self):
        self._ck(self.lib.cg_comm_destroy(self.h))

    def stats_allreduce(self, n_rows: int):
        """cg_stats_allreduce: the path's one collective, queued on cg_stream behind everything queued so far."""
        self._ck(self.lib.cg_stats_allreduce(self.h, int(n_rows)))

    def stats_row(self, chr_index: int):
        st = np.zeros(1, dtype=STATS_DT)
        self._ck(self.lib.cg_stats_row(self.h, chr_index, st.ctypes.data))
        return st[0]

    def set_profiling(self, on: bool):
        self._ck(self.lib.cg_set_profiling(self.h, int(on)))

    def kernel_times(self):
        ms = (C.c_double * CG_NSLOTS)()
        n = (C.c_int64 * CG_NSLOTS)()
        self._ck(self.lib.cg_kernel_times(self.h, ms, n))
        return list(ms), list(n)


def encode_positions_delta(pos: np.ndarray):
    """cg_encode_positions_delta: sorted int32 positions -> (uint8 deltas, int32 side array, ANCHOR_DT anchors)."""
    lib = load_library()
    pos = np.ascontiguousarray(pos, dtype=np.int32)
    n = len(pos)
    delta = np.empty(n, dtype=np.uint8)
    anchors = np.zeros((n + CG_DELTA_BLOCK - 1) // CG_DELTA_BLOCK, dtype=ANCHOR_DT)
    big = np.empty(max(16, n // 16), dtype=np.int32)
    nb = C.c_size_t(0)
    for _ in range(2):
        rc = lib.cg_encode_positions_delta(pos.ctypes.data, n, delta.ctypes.data, big.ctypes.data, len(big), C.byref(nb), anchors.ctypes.data)
        if rc != CG_ERANGE:
            break
        big = np.empty(nb.value, dtype=np.int32)
    if rc != CG_OK:
        raise CongaGpuError(rc, lib.cg_last_error(None).decode())
    return delta, big[:nb.value].copy(), anchors


def decode_positions_delta(delta: np.ndarray, big: np.ndarray, anchors: np.ndarray) -> np.ndarray:
    """What k_decode_delta computes, in NumPy (tests): the int32 positions of a delta-coded stream."""
    gaps = delta.astype(np.int64)
    gaps[delta == 255] = big
    pos = np.cumsum(gaps).astype(np.int32)
    if len(pos):
        blk = np.arange(len(anchors)) * CG_DELTA_BLOCK
        before = np.concatenate([[0], pos[blk[1:] - 1]]) if len(blk) > 1 else np.zeros(1, dtype=np.int64)
        assert np.array_equal(anchors["pos_before"], before.astype(np.int32))
        esc = np.concatenate([[0], np.cumsum(delta == 255)])
        assert np.array_equal(anchors["big_before"], esc[blk].astype(np.uint32))
    return pos


def peer_connect_local(contexts):
    """cg_peer_connect_local: the contexts of ONE process become each other's peers (rank = position)."""
    lib = load_library()
    arr = (C.c_void_p * len(contexts))(*[c.h for c in contexts])
    rc = lib.cg_peer_connect_local(arr, len(contexts))
    if rc != CG_OK:
        raise CongaGpuError(rc, lib.cg_last_error(contexts[0].h).decode())


CG_NOALLOC_MAX_LEN = 8386560


def depth_stats_noalloc(device: int, stream, pos, mapq, n_reads: int, mq_threshold: int, chrom_len: int, gc, n_windows: int,
                        depth_out=None, stats_out=None, expected_out=None, flags_out=None) -> None:
    """cg_depth_stats_noalloc: the depth side of one chromosome out of the library's fixed scratch array.  Every
    pointer is a device address (torch tensor, int or None); `stream` is a cudaStream_t as an int (0: default)."""
    lib = load_library()
    rc = lib.cg_depth_stats_noalloc(device, int(stream or 0), _ptr(pos), _ptr(mapq), n_reads, mq_threshold, chrom_len,
                                    _ptr(gc), n_windows, _ptr(depth_out), _ptr(stats_out), _ptr(expected_out), _ptr(flags_out))
    if rc != CG_OK:
        raise CongaGpuError(rc, lib.cg_last_error(None).decode())


def comm_unique_id() -> bytes:
    """cg_comm_unique_id: 128 bytes one rank creates and every rank passes to Context.comm_init_rank."""
    lib = load_library()
    buf = C.create_string_buffer(CG_COMM_ID_BYTES)
    rc = lib.cg_comm_unique_id(C.cast(buf, C.c_void_p))
    if rc != CG_OK:
        raise CongaGpuError(rc, lib.cg_last_error(None).decode())
    return buf.raw


def chrom_job(chr_index, chrom_len, gc, n_windows, mq_threshold, reads, n_reads, runs=None, n_runs=0, use_mappability=False,
              dels=None, n_del=0, del_calls=None, dups=None, n_dup=0, dup_calls=None, stats=None,
              reads_pos=None, reads_mapq=None, flags=0, reads_delta=None, reads_big=None, n_big=0, reads_anchor=None) -> np.ndarray:
    """One cg_chrom_job record; array arguments may be NumPy arrays, torch tensors or raw addresses.
    reads=None with reads_pos (and optionally reads_mapq) selects the compact record planes."""
    j = np.zeros(1, dtype=CHROM_JOB_DT)
    j["chr_index"], j["chrom_len"], j["gc"], j["n_windows"], j["mq_threshold"] = chr_index, chrom_len, _ptr(gc), n_windows, mq_threshold
    j["reads"], j["n_reads"] = _ptr(reads), n_reads
    j["reads_pos"], j["reads_mapq"], j["flags"] = _ptr(reads_pos), _ptr(reads_mapq), flags
    j["reads_delta"], j["reads_big"], j["n_big"], j["reads_anchor"] = _ptr(reads_delta), _ptr(reads_big) if n_big else 0, n_big, _ptr(reads_anchor)
    j["runs"], j["n_runs"], j["use_mappability"] = _ptr(runs) if n_runs else 0, n_runs, int(bool(use_mappability))
    j["dels"], j["n_del"], j["del_calls"] = _ptr(dels) if n_del else 0, n_del, _ptr(del_calls) if n_del else 0
    j["dups"], j["n_dup"], j["dup_calls"] = _ptr(dups) if n_dup else 0, n_dup, _ptr(dup_calls) if n_dup else 0
    j["stats"] = _ptr(stats)
    return j


def flatten_mappability(rows: np.ndarray, chrom_len: int) -> np.ndarray:
    """Host-side cg_flatten_mappability: BED rows (file order, inclusive ends) -> disjoint runs."""
    lib = load_library()
    assert rows.dtype == MAPIV_DT and rows.flags.c_contiguous
    p, n = C.c_void_p(), C.c_size_t()
    rc = lib.cg_flatten_mappability(rows.ctypes.data if len(rows) else None, len(rows), chrom_len, C.byref(p), C.byref(n))
    if rc != CG_OK:
        raise CongaGpuError(rc, lib.cg_last_error(None).decode())
    try:
        buf = (C.c_char * (n.value * MAPRUN_DT.itemsize)).from_address(p.value) if n.value else b""
        return np.frombuffer(buf, dtype=MAPRUN_DT, count=n.value).copy()
    finally:
        lib.cg_free_runs(p)
