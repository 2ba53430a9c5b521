"""Where a step's time goes on ONE GPU when it owns the chromosomes rank 0 of a WORLD-rank job would own
(no collective): host issue time, pipelined device time per step, and the same with stages removed.

  python scripts/step_breakdown.py [--world W] [--steps K] [lib.so ...]
"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from conga_b200 import abi, synth, CG_PTR_DEVICE, CALL_DT, flatten_mappability

ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--rank", type=int, default=0)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--workload", default="c2")
ap.add_argument("libs", nargs="*")
args = ap.parse_args()

wl = synth.config3() if args.workload == "c3" else synth.config2()
parts = synth.lpt_partition([c.length for c in wl.chroms], args.world)
mine = [wl.chroms[i] for i in parts[args.rank]]
dev = torch.device("cuda:0")
to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
res = []
for c in mine:
    runs = flatten_mappability(c.maprows, c.length) if (wl.has_mappability and c.maprows is not None) else None
    res.append(dict(c=c, gc=to_dev(c.gc), reads=to_dev(c.reads), dels=to_dev(c.dels), dups=to_dev(c.dups),
                    runs=to_dev(runs) if runs is not None and len(runs) else None, n_runs=0 if runs is None else len(runs),
                    od=torch.empty(max(1, len(c.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev),
                    ou=torch.empty(max(1, len(c.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)))
alg = sum(8 * len(c.reads) + 2 * c.length + (c.length + 99) // 100 for c in mine)
n_reads = sum(len(c.reads) for c in mine)

for path in (args.libs or [abi.LIB_PATH]):
    abi._lib = abi.load_library(path)
    ctx = abi.Context(0)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def step(geno=True, mapp=True, dups=True):
        for r in res:
            c = r["c"]
            ctx.chrom_begin(c.index, c.length, r["gc"], wl.mq_threshold, n_windows=len(c.gc))
            ctx.push_reads(r["reads"], len(c.reads))
            ctx.chrom_finish(None)
            if mapp and wl.has_mappability:
                ctx.set_mappability_runs(r["runs"], r["n_runs"], True)
            if geno:
                ctx.genotype(r["dels"], False, r["od"], n=len(c.dels))
                if dups:
                    ctx.genotype(r["dups"], True, r["ou"], n=len(c.dups))

    def run(K, **kw):
        for _ in range(3):
            step(**kw)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record()
        t0 = time.perf_counter()
        for _ in range(K):
            step(**kw)
        t_issue = (time.perf_counter() - t0) / K * 1e3
        ctx.join()
        with torch.cuda.stream(ext):
            e1.record()
        ctx.sync()
        return round(e0.elapsed_time(e1) / K, 4), round(t_issue, 4)

    out = {"lib": os.path.basename(path), "world": args.world, "chroms": len(mine), "pass_floor_ms": round(alg / 6536e6, 4)}
    out["full"] = run(args.steps)
    out["one_step_serial"] = run(1)
    out["no_dups"] = run(args.steps, dups=False)
    out["no_map"] = run(args.steps, mapp=False)
    out["no_geno"] = run(args.steps, geno=False, mapp=False)
    # HOST pointer mode from pinned buffers (bench.py's e2e leg), one sample in flight behind the one being issued
    from conga_b200 import CG_PTR_HOST, READ_DT, CNV_DT, MAPRUN_DT, STATS_DT
    ctx.set_pointer_mode(CG_PTR_HOST); ctx.set_async(True)
    host = []
    for c in mine:
        pin = lambda a, dt: (lambda q: (q.__setitem__(slice(None), a), q)[1])(ctx.pinned_array(len(a), dt))
        runs = flatten_mappability(np.ascontiguousarray(c.maprows), c.length) if (wl.has_mappability and c.maprows is not None) else np.zeros(0, dtype=MAPRUN_DT)
        host.append((c, pin(c.reads, READ_DT), pin(c.gc, np.int8), pin(c.dels, CNV_DT), pin(c.dups, CNV_DT), pin(runs, MAPRUN_DT),
                     [ctx.pinned_array(len(c.dels), CALL_DT) for _ in range(2)], [ctx.pinned_array(len(c.dups), CALL_DT) for _ in range(2)],
                     [ctx.pinned_array(1, STATS_DT) for _ in range(2)]))
    lib, h = ctx.lib, ctx.h
    state = {"n": 0, "pending": None}

    def step_host(depth=1, geno=True, mapp=True):
        b = state["n"] & 1; state["n"] += 1
        for (c, pr, pg, pd, pu, prn, od, ou, st) in host:
            ctx._ck(lib.cg_chrom_begin(h, c.index, c.length, pg.ctypes.data, len(pg), wl.mq_threshold))
            ctx._ck(lib.cg_push_reads(h, pr.ctypes.data, len(pr)))
            ctx._ck(lib.cg_chrom_finish(h, st[b].ctypes.data))
            if mapp and wl.has_mappability:
                ctx._ck(lib.cg_set_mappability_runs(h, prn.ctypes.data if len(prn) else None, len(prn), 1))
            if geno:
                if len(pd): ctx._ck(lib.cg_genotype(h, pd.ctypes.data, len(pd), 0, od[b].ctypes.data))
                if len(pu): ctx._ck(lib.cg_genotype(h, pu.ctypes.data, len(pu), 1, ou[b].ctypes.data))
        if depth == 0:
            ctx.sync(); return
        ev = torch.cuda.Event(); ctx.join()
        with torch.cuda.stream(ext): ev.record()
        if state["pending"] is not None: state["pending"].synchronize()
        state["pending"] = ev

    def run_host(K, **kw):
        for _ in range(3): step_host(**kw)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext): e0.record()
        t0 = time.perf_counter()
        for _ in range(K): step_host(**kw)
        t_issue = (time.perf_counter() - t0) / K * 1e3
        ctx.join()
        with torch.cuda.stream(ext): e1.record()
        ctx.sync()
        return round(e0.elapsed_time(e1) / K, 4), round(t_issue, 4)
    out["host_pipelined"] = run_host(args.steps)
    out["host_sync_each"] = run_host(args.steps, depth=0)
    out["host_no_geno"] = run_host(args.steps, geno=False, mapp=False)
    out["host_no_map"] = run_host(args.steps, mapp=False)
    out["h2d_MB"] = round(sum(x[1].nbytes + x[2].nbytes + x[3].nbytes + x[4].nbytes + x[5].nbytes for x in host) / 1e6, 2)
    ctx.set_async(False); ctx.set_pointer_mode(CG_PTR_DEVICE)
    ctx.set_profiling(True); ctx.kernel_times()
    for _ in range(args.steps):
        step()
    ms, n = ctx.kernel_times()
    ctx.set_profiling(False)
    out["kernels_alone_ms"] = {"prep": round(ms[0] / args.steps, 4), "pass": round(ms[1] / args.steps, 4), "geno": round(ms[2] / args.steps, 4)}
    out["reads_per_s_full"] = round(n_reads / (out["full"][0] / 1e3))
    print(json.dumps(out), flush=True)
    ctx.close()
