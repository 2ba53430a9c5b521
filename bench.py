#!/usr/bin/env python
"""bench.py — throughput of the read-depth genotyping hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c2|c1|c3]

A "step" is one pass of the whole hot path over one synthetic sample: for every chromosome
cg_chrom_begin -> cg_push_reads -> cg_chrom_finish (fused depth + GC-bin + prefix pass) ->
mappability -> cg_genotype(dels) -> cg_genotype(dups); at N > 1 the chromosomes are partitioned
over the ranks (longest-processing-time by length) and the step ends with the one
collective the path has, an NCCL all-reduce of the int64[n_chrom][204] statistics table.

One JSON line on stdout (rank 0).  `value` = reads/s of the whole job with inputs resident in
HBM; `e2e` = the same through the C-ABI in HOST pointer mode from pinned host buffers with the
H2D/D2H copies inside the timed region; `roofline` = the fused depth pass against the measured
HBM peak; `cpu_baseline` = the oracle port timed on one host core on a bounded sample.
`--impl reference` times the reference's own CPU implementation (oracle/_ref/ref_path: the
unmodified reference objects fed from record files; else the oracle port) on all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "reads/s through the read-depth genotyping hot path (depth + GC bins + per-CNV sums + CNV genotyping)"


def env_int(name, default):
    return int(os.environ.get(name, default))


def build_workload(name, scale):
    from conga_b200 import synth
    if name == "c1":
        return synth.config1()
    if name == "c3":
        return synth.config3(scale=scale)
    return synth.config2(scale=scale)


class NvmlSampler:
    """SM clock and throttle reasons polled through NVML every few milliseconds from a thread while
    the timed regions run (the nvidia-smi loop of B200_PROFILING.md cannot sample faster than its
    own start-up time, and a timed region here lasts tens of milliseconds)."""

    def __init__(self, gpu_index, period_s=0.004):
        import threading
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        self.period = period_s
        self.sm, self.reasons, self.busy = [], set(), []
        self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        self._stop = threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for nm, b in bits.items():
                    if r & b:
                        self.reasons.add(nm)
            except Exception:           # a failed query is a missing sample, nothing more
                pass
            self._stop.wait(self.period)

    def start(self):
        self._t.start()

    def stop(self):
        self._stop.set()
        self._t.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                "samples": len(self.sm), "reasons": sorted(self.reasons), "source": "nvml, %d ms period" % round(self.period * 1e3)}


def make_sampler(gpu_index):
    try:
        return NvmlSampler(gpu_index)
    except Exception:
        return ClockSampler(gpu_index)


class ClockSampler:
    """Fallback: nvidia-smi sampled every 200 ms while the timed region runs (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f,
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


class DevArr:
    """Expose a raw device address to torch (torch.as_tensor) without copying."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}


# ------------------------------------------------------------------------------------------
def own_arm(args):
    # A context drives ~26 streams (6 slots x (upload, slot, 2 genotype) + 2); with the default of 8 hardware
    # connections streams share queues and kernels of different chromosomes serialise behind each other
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    import torch
    import torch.distributed as dist
    from conga_b200 import multi
    from conga_b200 import (CALL_DT, CG_PTR_DEVICE, CG_PTR_HOST, CHROM_JOB_DT, CNV_DT, MAPIV_DT, MAPRUN_DT, READ_DT, STATS_DT, Context,
                            chrom_job, flatten_mappability, synth)

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    # stdout carries the JSON line and nothing else: libraries that print there (NCCL's version banner)
    # are sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    wl = build_workload(args.workload, args.scale)
    n_chrom = len(wl.chroms)
    parts = multi.lpt_partition([c.length for c in wl.chroms], world)
    mine = [wl.chroms[i] for i in parts[rank]]
    has_map, mq = wl.has_mappability, wl.mq_threshold

    ctx = Context(local)                      # raises if libconga_gpu.so or the GPU is missing
    # chromosomes in flight: six by default.  With inputs resident, a rank that owns two or three chromosomes per
    # sample (8 GPUs) gains nothing from more than four and has a third fewer streams to order in every cg_join;
    # the end-to-end leg keeps six everywhere (uploads of the chromosomes ahead need the depth: 12.5 against 9.9 G
    # reads/s at 8 GPUs).  DESIGN.md section 6.
    N_SLOTS = env_int("BENCH_SLOTS", 4 if len(mine) <= 3 else 6)
    N_SLOTS_E2E = env_int("BENCH_SLOTS_E2E", 6)
    ctx.set_slots(N_SLOTS)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    # the statistics table of each of the two banks (cg_stats_bank): consecutive samples alternate, so
    # the all-reduce of sample k is in flight while sample k+1's chromosomes run
    tables = []
    for b in (1, 0):
        ctx.stats_bank(b)
        tables.insert(0, torch.as_tensor(DevArr(ctx.stats_table()[0], (n_chrom, 204), "<i8"), device=dev))
    sample_no = [0]

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)

    # ---- leg 1: inputs resident in HBM (CG_PTR_DEVICE) ------------------------------------
    res = []
    for c in mine:
        runs = flatten_mappability(c.maprows, c.length) if (has_map and c.maprows is not None) else None
        res.append(dict(c=c, gc=to_dev(c.gc), reads=to_dev(c.reads), dels=to_dev(c.dels), dups=to_dev(c.dups),
                        runs=to_dev(runs) if runs is not None and len(runs) else None,
                        n_runs=0 if runs is None else len(runs),
                        out_del=torch.empty(max(1, len(c.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev),
                        out_dup=torch.empty(max(1, len(c.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)))
    torch.cuda.synchronize()

    # one cg_run_chromosomes per sample = cg_chrom_begin / cg_push_reads / cg_chrom_finish / cg_set_mappability_runs /
    # cg_genotype x2 for each chromosome, in one foreign call
    jobs_dev = np.concatenate([chrom_job(r["c"].index, r["c"].length, r["gc"], len(r["c"].gc), mq, r["reads"], len(r["c"].reads),
                                         r["runs"], r["n_runs"], has_map, r["dels"], len(r["c"].dels), r["out_del"],
                                         r["dups"], len(r["c"].dups), r["out_dup"], None) for r in res]) if res else np.zeros(0, dtype=CHROM_JOB_DT)

    def step_device():
        bank = sample_no[0] & 1
        sample_no[0] += 1
        if world > 1:
            ctx.stats_bank(bank)              # cleared on the device behind the all-reduce that last read it
        ctx.run_chromosomes(jobs_dev)
        if world > 1:
            ctx.join()                        # main stream now waits for every stream of the context
            with torch.cuda.stream(ext):
                multi.allreduce_stats(tables[bank])

    def timed(step_fn, k):
        barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record()
        for _ in range(k):
            step_fn()
        ctx.join()                            # cg_stream() now waits for every other stream of the context
        with torch.cuda.stream(ext):
            e1.record()
        ctx.sync()
        torch.cuda.synchronize()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    ctx.set_pointer_mode(CG_PTR_DEVICE)
    for _ in range(args.warmup):
        step_device()
    ctx.sync()
    sampler = make_sampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count
    ms_dev = timed(step_device, args.steps)
    launches = ctx.launch_count - launches0
    # same K steps again with per-launch CUDA events on the context's stream -> roofline of the depth pass
    ctx.set_profiling(True)
    ctx.kernel_times()
    timed(step_device, args.steps)
    k_ms, k_n = ctx.kernel_times()
    ctx.set_profiling(False)
    clocks = sampler.stop() if rank == 0 else None

    # ---- leg 2: end to end through the C-ABI from pinned host buffers (CG_PTR_HOST) -------
    # A pipelined host: page-locked buffers, asynchronous calls, one cg_sync per step — that is when the
    # results (statistics rows, genotype calls) are valid on the host.
    DEPTH = max(1, env_int("BENCH_E2E_DEPTH", 2))      # samples in flight behind the one being issued, per lane
    # BENCH_E2E_LANES > 1 issues from that many host threads, one context each (a context is driven by one thread
    # at a time; the foreign call releases the interpreter lock).  Measured r01h: two lanes give 2.26 against
    # 2.30 G reads/s with one — the end-to-end leg is not bound by the issuing thread — so one lane is the default.
    # Multi-GPU ranks always use one: their collective must be issued in one order on every rank.
    LANES = max(1, env_int("BENCH_E2E_LANES", 1)) if world == 1 else 1
    lanes = [dict(ctx=ctx, ext=ext)]
    for _ in range(LANES - 1):
        cx = Context(local)
        lanes.append(dict(ctx=cx, ext=torch.cuda.ExternalStream(cx.stream, device=dev)))
    for ln in lanes:
        ln["ctx"].set_slots(N_SLOTS_E2E)
        ln["ctx"].set_pointer_mode(CG_PTR_HOST)
        ln["ctx"].set_async(True)
    # inputs: one page-locked copy shared by the lanes; results: DEPTH + 1 buffers per lane, so that sample k's
    # calls stay intact while the next DEPTH samples of the lane are being issued
    inputs = []
    h2d = d2h = 0
    for c in mine:
        pr = ctx.pinned_array(len(c.reads), READ_DT); pr[:] = c.reads
        pg = ctx.pinned_array(len(c.gc), np.int8); pg[:] = c.gc
        pd = ctx.pinned_array(len(c.dels), CNV_DT); pd[:] = c.dels
        pu = ctx.pinned_array(len(c.dups), CNV_DT); pu[:] = c.dups
        rows = np.ascontiguousarray(c.maprows) if (has_map and c.maprows is not None) else np.zeros(0, dtype=MAPIV_DT)
        # the track is flattened once per BED file (input parsing, like the reference's per-chromosome
        # re-parse of the BED, svs.c:329-374); its upload is part of every step
        fr = flatten_mappability(rows, c.length) if len(rows) else np.zeros(0, dtype=MAPRUN_DT)
        prn = ctx.pinned_array(len(fr), MAPRUN_DT); prn[:] = fr
        inputs.append((c, pr, pg, pd, pu, prn))
        h2d += pr.nbytes + pg.nbytes + pd.nbytes + pu.nbytes + prn.nbytes
        d2h += len(c.dels) * CALL_DT.itemsize + len(c.dups) * CALL_DT.itemsize + STATS_DT.itemsize
    for ln in lanes:
        cx = ln["ctx"]
        ln["out"] = [[(cx.pinned_array(len(c.dels), CALL_DT), cx.pinned_array(len(c.dups), CALL_DT), cx.pinned_array(1, STATS_DT))
                      for (c, *_rest) in inputs] for _ in range(DEPTH + 1)]
        ln["jobs"] = [np.concatenate([chrom_job(c.index, c.length, pg, len(pg), mq, pr, len(pr), prn, len(prn), has_map,
                                                pd, len(pd), od, pu, len(pu), ou, st)
                                      for (c, pr, pg, pd, pu, prn), (od, ou, st) in zip(inputs, outs)]) if inputs else np.zeros(0, dtype=CHROM_JOB_DT)
                      for outs in ln["out"]]
        ln["pending"], ln["n"] = [], 0

    def lane_step(ln):
        cx = ln["ctx"]
        buf = ln["n"] % (DEPTH + 1)
        ln["n"] += 1
        if world > 1:
            bank = sample_no[0] & 1
            sample_no[0] += 1
            cx.stats_bank(bank)
        cx.run_chromosomes(ln["jobs"][buf])
        # cg_join orders cg_stream behind everything queued so far; the all-reduce and the event go on cg_stream
        cx.join()
        if world > 1:
            with torch.cuda.stream(ln["ext"]):
                multi.allreduce_stats(tables[bank])
        # DEPTH samples in flight behind the one being issued: a sample's statistics rows and calls are complete
        # in host memory when its event has passed (a pipelined host waits there, not on the whole device)
        ev = torch.cuda.Event()
        with torch.cuda.stream(ln["ext"]):
            ev.record()
        ln["pending"].append(ev)
        if len(ln["pending"]) > DEPTH:
            ln["pending"].pop(0).synchronize()

    def run_lanes(k):
        """k samples, dealt round-robin to the lanes; returns when every one of them has been issued."""
        if len(lanes) == 1:
            for _ in range(k):
                lane_step(lanes[0])
            return
        def work(ln, n):
            torch.cuda.set_device(local)
            for _ in range(n):
                lane_step(ln)
        ts = [threading.Thread(target=work, args=(ln, (k + len(lanes) - 1 - i) // len(lanes))) for i, ln in enumerate(lanes)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    def timed_e2e(k):
        barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record()
        for ln in lanes[1:]:
            ln["ext"].wait_event(e0)          # no lane starts before the clock does
        run_lanes(k)
        for ln in lanes[1:]:                  # ... and the clock stops behind the last sample of every lane
            ln["ctx"].join()
            done = torch.cuda.Event()
            with torch.cuda.stream(ln["ext"]):
                done.record()
            ext.wait_event(done)
        ctx.join()
        with torch.cuda.stream(ext):
            e1.record()
        for ln in lanes:
            ln["ctx"].sync()
        torch.cuda.synchronize()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    run_lanes(max(3, args.warmup) * len(lanes))      # chromosomes rotate through the slots: their upload buffers settle within three samples
    for ln in lanes:
        ln["ctx"].sync()
    ms_e2e = timed_e2e(args.steps)
    if world > 1:                             # every rank holds every chromosome's row after the all-reduce
        for tb in tables:
            t_host = tb.cpu().numpy()
            for c in wl.chroms:
                assert int(t_host[c.index, 101:202].sum()) == c.length, "statistics table row %d wrong after the all-reduce" % c.index

    # sanity: the e2e leg and the device leg produced the same bits for the last chromosome, in every result buffer
    if res:
        got = res[-1]["out_del"].cpu().numpy()[: len(mine[-1].dels) * CALL_DT.itemsize]
        for ln in lanes:
            for outs in ln["out"][: min(DEPTH + 1, ln["n"])]:
                assert got.tobytes() == outs[-1][0].tobytes(), "device-mode and host-mode results differ"
    for ln in lanes[1:]:
        ln["ctx"].close()

    n_reads, n_cnvs = wl.n_reads, wl.n_cnvs
    value = n_reads * args.steps / (ms_dev / 1e3)
    e2e = n_reads * args.steps / (ms_e2e / 1e3)
    if world > 1:
        t = torch.tensor([h2d, d2h, launches], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        h2d, d2h, launches = int(t[0].item()), int(t[1].item()), int(t[2].item())

    # roofline of the dominant kernel (fused depth pass), this rank's chromosomes
    alg = sum(8 * len(c.reads) + 2 * c.length + (c.length + 99) // 100 for c in mine)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    pass_ms, pass_n = k_ms[1], max(1, k_n[1])
    achieved = alg * args.steps / (pass_ms / 1e3) / 1e9 if pass_ms > 0 else 0.0
    # DRAM bytes per launch from the committed ncu capture of this command (profiles/traffic.json, written by
    # scripts/traffic_summary.py from scripts/gpu_traffic.sh); only meaningful for the workload it was taken on
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath) and args.workload == "c2" and args.scale == 1.0 and world == 1:
        tj = json.load(open(tpath))
        traffic, traffic_src = int(tj["traffic_bytes_per_launch"]), tj["source"]
    roofline = {"kernel": "k_depth_pass", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg // max(1, len(mine)), "launches": int(pass_n),
                "avg_launch_ms": round(pass_ms / pass_n, 4),
                "kernel_ms_per_step": {"prep": round(k_ms[0] / args.steps, 4), "depth_pass": round(k_ms[1] / args.steps, 4),
                                       "genotype": round(k_ms[2] / args.steps, 4), "expected_table": round(k_ms[3] / args.steps, 4)}}
    cnv_ms = k_ms[2] / args.steps
    out = {"metric": METRIC, "value": round(value, 1), "unit": "reads/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": round(ms_dev / args.steps, 4), "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "int16 depth, int32 sums, f32 ordered sum, f64 likelihood",
           "data": "synthetic", "impl": "b200",
           "config": {"workload": wl.name, "chromosomes": n_chrom, "positions": wl.n_positions, "reads": n_reads,
                      "cnvs": n_cnvs, "mappability": has_map, "partition": "chromosomes, LPT by length", "slots": N_SLOTS, "slots_e2e": N_SLOTS_E2E,
                      "l2": "inputs+outputs per step (%.1f GB) larger than L2" % (wl.algorithmic_bytes_depth_pass() / 1e9)},
           "cnvs_per_s": round(n_cnvs * args.steps / (ms_dev / 1e3), 1),
           "cnvs_per_s_genotype_kernels_rank0": round(sum(len(c.dels) + len(c.dups) for c in mine) / (cnv_ms / 1e3), 1) if cnv_ms > 0 else None,
           "e2e": {"value": round(e2e, 1), "unit": "reads/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "samples_in_flight": DEPTH * LANES, "host_threads": LANES},
           "gpu_launches": int(launches), "roofline": roofline, "clocks": clocks}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_port(wl, budget_s=args.cpu_budget)
    if rank == 0:
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------
def cpu_baseline_port(wl, budget_s=15.0):
    """The oracle (C restatement of the reference loops, oracle/conga_oracle.c) on ONE host core over a
    bounded sample: chromosomes from the smallest up until ~budget_s seconds of CPU work."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_py
    orc = oracle_py.load()
    order = sorted(range(len(wl.chroms)), key=lambda i: wl.chroms[i].length)
    t_tot, reads, cnvs, used = 0.0, 0, 0, []
    for i in order:
        c = wl.chroms[i]
        t0 = time.perf_counter()
        orc.chromosome(c, wl.mq_threshold, wl.has_mappability)
        t_tot += time.perf_counter() - t0
        reads += len(c.reads)
        cnvs += len(c.dels) + len(c.dups)
        used.append(c.name)
        if t_tot >= budget_s:
            break
    return {"value": round(reads / t_tot, 1), "unit": "reads/s", "cores": 1, "kind": "port",
            "cnvs_per_s": round(cnvs / t_tot, 1), "seconds": round(t_tot, 3),
            "sample": "chromosomes %s of the same workload (%d reads, %d CNVs), whole per-chromosome path"
                      % (",".join(used), reads, cnvs)}


def reference_arm(args):
    """The reference's own CPU implementation of the path on the host cores.  Preferred: oracle/_ref/ref_path
    (UNMODIFIED reference objects: init_rd_per_chr, count_reads_bam, calc_mean_per_chr, find_SVs, sonic_load,
    fed from record files instead of BGZF), one process per chromosome because the program is single-threaded.
    Else: the oracle port, one process per chromosome."""
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    wl = build_workload(args.workload, args.scale)
    ncpu = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ref_path = os.path.join(ROOT, "oracle", "_ref", "ref_path")
    kind = "reference" if os.path.exists(ref_path) else "port"
    order = sorted(range(len(wl.chroms)), key=lambda i: wl.chroms[i].length)
    tmp = tempfile.mkdtemp(prefix="conga_ref_")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))

    phases = {}                                   # CPU-seconds per reference function, summed over the processes of the last step

    def run_once(subset, base_args):
        t0 = time.perf_counter()
        if kind == "reference":
            procs = []
            for i in subset:
                c = wl.chroms[i]
                wd = os.path.join(tmp, "w%d" % i)
                os.makedirs(wd, exist_ok=True)
                procs.append(subprocess.Popen([ref_path] + base_args + ["--first-chr", str(c.index), "--last-chr", str(c.index),
                                               "--out", os.path.join(wd, "o")], cwd=wd, stdout=subprocess.PIPE,
                                              stderr=subprocess.DEVNULL))
            phases.clear()
            for p in procs:
                line = p.communicate()[0]             # one short JSON line per process: per-phase seconds
                if p.returncode != 0:
                    raise RuntimeError("ref_path failed")
                try:
                    for k, v in json.loads(line.decode().strip().splitlines()[-1]).items():
                        if k.endswith("_s"):
                            phases[k] = phases.get(k, 0.0) + float(v)
                except (ValueError, IndexError):
                    pass
        else:
            import multiprocessing as mp
            with mp.get_context("fork").Pool(min(ncpu, len(subset))) as pool:
                pool.map(_port_one, [(wl, i) for i in subset])
        return time.perf_counter() - t0

    import refdata
    # calibrate on the smallest chromosome, then take as many chromosomes (smallest first, one core each)
    # as keep one step under the per-step budget
    total_budget = 150.0
    per_step = total_budget / max(1, args.steps + args.warmup)
    base_args = refdata.write_workload(tmp, wl, chrom_subset=order[:1]) if kind == "reference" else []
    t_cal = run_once(order[:1], base_args)
    s_per_mb = t_cal / (wl.chroms[order[0]].length / 1e6)
    subset = [i for i in order if wl.chroms[i].length / 1e6 * s_per_mb * 1.3 <= per_step][:ncpu] or order[:1]
    if kind == "reference":
        base_args = refdata.write_workload(tmp, wl, chrom_subset=subset)
    reads = sum(len(wl.chroms[i].reads) for i in subset)
    cnvs = sum(len(wl.chroms[i].dels) + len(wl.chroms[i].dups) for i in subset)
    for _ in range(args.warmup):
        run_once(subset, base_args)
    t = 0.0
    for _ in range(args.steps):
        t += run_once(subset, base_args)
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)
    value = reads * args.steps / t
    cores = min(ncpu, len(subset))
    sample = ("%d of %d chromosomes (%s; %d reads, %d CNVs) per step, one single-threaded process per chromosome"
              % (len(subset), len(wl.chroms), ",".join(wl.chroms[i].name for i in sorted(subset)), reads, cnvs))
    out = {"metric": METRIC, "value": round(value, 1), "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": round(t / args.steps * 1e3, 3), "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "int16/int32 depth, f32 ordered sum, f64 likelihood",
           "data": "synthetic", "impl": "reference",
           "config": {"workload": wl.name, "chromosomes": len(wl.chroms), "positions": wl.n_positions,
                      "reads": wl.n_reads, "cnvs": wl.n_cnvs, "mappability": wl.has_mappability},
           "cnvs_per_s": round(cnvs * args.steps / t, 1),
           "cpu_baseline": {"value": round(value, 1), "unit": "reads/s", "cores": cores, "kind": kind, "sample": sample,
                            "host_cores_available": ncpu,
                            # where the reference spends its time: init_rd_per_chr / count_reads_bam /
                            # calc_mean_per_chr / find_SVs, CPU-seconds summed over the processes of the last step
                            "phase_cpu_seconds_last_step": {k: round(v, 4) for k, v in sorted(phases.items())} or None},
           "e2e": {"value": round(value, 1), "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def _port_one(a):
    wl, i = a
    import oracle_py
    oracle_py.load().chromosome(wl.chroms[i], wl.mq_threshold, wl.has_mappability)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the genome (debugging only; 1.0 = the named config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        own_arm(args)


if __name__ == "__main__":
    main()
