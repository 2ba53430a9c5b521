"""Host time per C-ABI entry point in asynchronous HOST mode (C2, one sample = 22 chromosomes): where the issue
loop of bench.py's end-to-end leg spends its time.  python scripts/host_call_times.py [--samples K]"""
import argparse, collections, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conga_b200 import abi, synth, CG_PTR_HOST, CG_PTR_DEVICE, CALL_DT, CNV_DT, MAPRUN_DT, READ_DT, STATS_DT, flatten_mappability
import torch

ap = argparse.ArgumentParser(); ap.add_argument("--samples", type=int, default=10); args = ap.parse_args()
wl = synth.config2()
ctx = abi.Context(0)
lib, h = ctx.lib, ctx.h
dev = torch.device("cuda:0")
out = {}
for mode in ("host", "device"):
    ctx.set_pointer_mode(CG_PTR_HOST if mode == "host" else CG_PTR_DEVICE)
    if mode == "host":
        ctx.set_async(True)
    bufs = []
    for c in wl.chroms:
        runs = flatten_mappability(np.ascontiguousarray(c.maprows), c.length)
        if mode == "host":
            pin = lambda a, dt: (lambda q: (q.__setitem__(slice(None), a), q)[1])(ctx.pinned_array(len(a), dt))
            b = [pin(c.gc, np.int8), pin(c.reads, READ_DT), pin(runs, MAPRUN_DT), pin(c.dels, CNV_DT), pin(c.dups, CNV_DT),
                 ctx.pinned_array(len(c.dels), CALL_DT), ctx.pinned_array(len(c.dups), CALL_DT), ctx.pinned_array(1, STATS_DT)]
            p = [x.ctypes.data for x in b]
        else:
            up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)
            b = [up(c.gc), up(c.reads), up(runs), up(c.dels), up(c.dups), torch.empty(len(c.dels) * 56, dtype=torch.uint8, device=dev),
                 torch.empty(len(c.dups) * 56, dtype=torch.uint8, device=dev), torch.empty(STATS_DT.itemsize, dtype=torch.uint8, device=dev)]
            p = [int(x.data_ptr()) for x in b]
        bufs.append((c, b, p, len(runs)))
    acc = collections.defaultdict(float)
    pc = time.perf_counter
    for k in range(args.samples + 2):
        if k == 2:
            acc.clear()
        for c, b, p, n_runs in bufs:
            t0 = pc(); lib.cg_chrom_begin(h, c.index, c.length, p[0], len(c.gc), wl.mq_threshold)
            t1 = pc(); lib.cg_push_reads(h, p[1], len(c.reads))
            t2 = pc(); lib.cg_chrom_finish(h, p[7])
            t3 = pc(); lib.cg_set_mappability_runs(h, p[2], n_runs, 1)
            t4 = pc(); lib.cg_genotype(h, p[3], len(c.dels), 0, p[5])
            t5 = pc(); lib.cg_genotype(h, p[4], len(c.dups), 1, p[6])
            t6 = pc()
            for nm, d in (("chrom_begin", t1 - t0), ("push_reads", t2 - t1), ("chrom_finish", t3 - t2), ("set_runs", t4 - t3), ("genotype_del", t5 - t4), ("genotype_dup", t6 - t5)):
                acc[nm] += d
        t0 = pc(); ctx.join(); acc["join"] += pc() - t0
        t0 = pc(); ctx.sync(); acc["sync_wait"] += pc() - t0
    out[mode] = {k: round(v / args.samples * 1e3, 4) for k, v in acc.items()}
    out[mode]["issue_ms_per_sample"] = round(sum(v for k, v in out[mode].items() if k != "sync_wait"), 4)
    if mode == "host":
        ctx.set_async(False)
print(json.dumps(out))
