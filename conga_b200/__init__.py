"""conga_b200 — B200-native read-depth genotyping hot path of CONGA.

The product is ``libconga_gpu.so`` (C-ABI in ``include/conga_gpu.h``, CUDA sources in
``conga_b200/csrc``).  This package holds the ctypes binding (`abi`), a host-side mirror
of the reference's call seams (`hotpath`) and the synthetic-input generator (`synth`).
Nothing here computes any stage of the path on the CPU.
"""
from .abi import (CALL_DT, CHROM_JOB_DT, chrom_job, CNV_DT, HALF_DT, SPLITROW_DT, MAPIV_DT, MAPRUN_DT, READ_DT, STATS_DT, CG_PTR_DEVICE, CG_PTR_HOST,
                  CongaGpuError, Context, comm_unique_id, depth_stats_noalloc, encode_positions_delta, flatten_mappability, load_library, peer_connect_local)

__all__ = ["CALL_DT", "CHROM_JOB_DT", "chrom_job", "CNV_DT", "HALF_DT", "SPLITROW_DT", "MAPIV_DT", "MAPRUN_DT", "READ_DT", "STATS_DT", "CG_PTR_DEVICE",
           "CG_PTR_HOST", "CongaGpuError", "Context", "comm_unique_id", "depth_stats_noalloc", "encode_positions_delta", "peer_connect_local", "flatten_mappability", "load_library"]
