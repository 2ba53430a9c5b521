"""Depth pass alone (no genotype kernels, inputs resident in HBM) on the C2 chromosomes, for one or
more builds of the library:  python scripts/pass_bench.py [lib.so ...]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from conga_b200 import abi, synth, CG_PTR_DEVICE

libs = sys.argv[1:] or [abi.LIB_PATH]
wl = synth.config2()
dev = torch.device("cuda:0")
to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
res = [(c, to_dev(c.gc), to_dev(c.reads)) for c in wl.chroms]
alg = sum(8 * len(c.reads) + 2 * c.length + (c.length + 99) // 100 for c in wl.chroms)
for path in libs:
    abi._lib = abi.load_library(path)
    ctx = abi.Context(0)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    def step():
        for c, gc, rd in res:
            ctx.chrom_begin(c.index, c.length, gc, -1, n_windows=len(c.gc))
            ctx.push_reads(rd, len(c.reads))
            ctx.chrom_finish(None)
    for _ in range(3): step()
    ctx.sync()
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 5
    with torch.cuda.stream(ext): e0.record()
    for _ in range(K): step()
    if hasattr(ctx.lib, "cg_join"): ctx.join()
    with torch.cuda.stream(ext): e1.record()
    ctx.sync()
    wall = e0.elapsed_time(e1) / K
    ctx.set_profiling(True); ctx.kernel_times()
    for _ in range(K): step()
    ms, n = ctx.kernel_times()
    print(json.dumps({"lib": os.path.basename(path), "step_ms": round(wall, 3), "step_GBs": round(alg / wall / 1e6, 1),
                      "pass_ms": round(ms[1] / K, 3), "pass_GBs": round(alg / (ms[1] / K) / 1e6, 1), "prep_ms": round(ms[0] / K, 3)}))
    ctx.close()
