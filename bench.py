#!/usr/bin/env python
"""bench.py — throughput of the read-depth genotyping hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c3|c1|c2|c4|c5]

A "step" is one pass of the whole hot path over one synthetic sample, handed to the library as ONE
submission (cg_run_chromosomes): record sweep, fused depth + GC-bin pass, finalize, mappability index and
genotype kernel each launched once for all chromosomes of the rank.  At N > 1 the chromosomes are
partitioned over the ranks (longest-processing-time by length) and the step ends with the one
collective the path has, an NCCL all-reduce of the int64[n_chrom][204] statistics table, issued by the
library itself (cg_stats_allreduce).  Workload at N = 1: C3 (BASELINE.json configs[2]: whole genome, 1x,
~40k CNVs) — the largest configuration without --rp; c2 / c4 / c5 / c1 are selectable.

One JSON line on stdout (rank 0).  `value` = reads/s of the whole job with inputs resident in
HBM, the median of `--repeats` timed regions of K steps each (spread in `run`); `e2e` = the same through
the C-ABI in HOST pointer mode from pinned host buffers with the H2D/D2H copies inside the timed region;
`roofline` = the fused depth pass against the measured HBM peak; `parity` = checks made inside this very
run (host mode == device mode; N ranks == one GPU, by hash of every call; a bounded sample against the
oracle); `cpu_baseline` = the oracle port timed on one host core on a bounded sample.
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


def build_workload(name, scale, samples=64):
    from conga_b200 import synth
    if name == "c1":
        return synth.config1()
    if name == "c2":
        return synth.config2(scale=scale)
    if name == "c4":
        return synth.config4(scale=scale)
    if name == "c5":
        return synth.config5(n_samples=samples, scale=scale)
    return synth.config3(scale=scale)


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
def calls_digest(per_chrom):
    """sha256 over the genotype calls of a whole sample in SONIC chromosome order, deletions before duplications
    (the order output_SVs writes them, likelihood.c:178-281).  Every byte of every cg_call counts."""
    import hashlib
    h = hashlib.sha256()
    for d, u in per_chrom:
        h.update(np.ascontiguousarray(d).tobytes())
        h.update(np.ascontiguousarray(u).tobytes())
    return h.hexdigest()[:16]


def own_arm(args):
    # a context drives 6 slots x 4 streams; with the default of 8 hardware connections streams share queues
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    import torch
    import torch.distributed as dist
    from conga_b200 import multi
    from conga_b200 import (CALL_DT, CG_PTR_DEVICE, CG_PTR_HOST, CHROM_JOB_DT, CNV_DT, MAPRUN_DT, READ_DT, STATS_DT, Context,
                            chrom_job, comm_unique_id, flatten_mappability)
    from conga_b200.abi import ANCHOR_DT, CG_JOB_ANNOTATION_RESIDENT, encode_positions_delta

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    # stdout carries the JSON line and nothing else: libraries that print there (NCCL's version banner)
    # are sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world != args.gpus and world == 1 and args.gpus > 1:
        sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)       # barriers and max-over-ranks
        cpu_group = dist.new_group(backend="gloo")           # host-side objects: the communicator id, the gather of the calls

    wl = build_workload(args.workload, args.scale, args.samples)
    n_chrom, S = len(wl.chroms), wl.n_samples
    parts = multi.balanced_partition([c.length for c in wl.chroms], world)
    mine = [wl.chroms[i] for i in parts[rank]]
    if args.emulate_rank and world == 1:
        # tuning only: rank R's share of an N-rank job alone on one GPU (no collective) -- what one rank of the
        # scaling run executes per step, measurable without holding N GPUs
        r_, n_ = (int(x) for x in args.emulate_rank.split("/"))
        mine = [wl.chroms[i] for i in multi.balanced_partition([c.length for c in wl.chroms], n_)[r_]]
    has_map, mq = wl.has_mappability, wl.mq_threshold
    # A rank of an N-GPU job holds 22/N chromosomes: a sample is then a small submission (C3 at N = 8: 0.15 ms of
    # pass per sample) and the fixed cost of a set of launches shows.  A submission carries up to 32 jobs, so the
    # resident leg hands over several consecutive samples per cg_run_chromosomes call: the rows of the g-th sample of a
    # submission live at chr_index + g * n_chrom of the bank, results in buffers of their own.
    KSUB = args.samples_per_submission
    jobs_max = max(1, max(len(p) for p in (multi.balanced_partition([c.length for c in wl.chroms], int(args.emulate_rank.split("/")[1]))
                                             if (args.emulate_rank and world == 1) else parts)))      # the same on every rank: the exchange is per submission
    if KSUB <= 0:
        KSUB = min(8, 32 // jobs_max) if (world > 1 or args.emulate_rank) else 1
    KSUB = max(1, min(KSUB, 32 // jobs_max))

    collective = os.environ.get("BENCH_COLLECTIVE", "peer")       # "nccl": the library's ncclAllReduce instead (the checked baseline)
    ctx = Context(local)                      # raises if libconga_gpu.so or the GPU is missing
    ctx.set_slots(env_int("BENCH_SLOTS", 6))
    if world > 1:
        # the path's one collective runs inside the library (cg_stats_allreduce -> ncclAllReduce on cg_stream)
        if collective == "nccl":
            box = [comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0, group=cpu_group)
            ctx.comm_init_rank(world, rank, box[0])
        else:
            # fused: k_finalize stores every finished row into every rank's table over NVLink (CUDA IPC peer memory);
            # cg_stats_allreduce is then a one-warp flag exchange on cg_stream.  Should peer memory not be available on a
            # box (IPC handles refused), every rank falls back to the NCCL form together, and the line says so.
            handles = [None] * world
            try:
                mine_h = ctx.peer_export()
            except Exception:
                mine_h = None
            dist.all_gather_object(handles, mine_h, group=cpu_group)
            ok = all(h is not None for h in handles)
            if ok:
                try:
                    ctx.peer_connect(world, rank, handles)
                except Exception:
                    ok = False
            oks = [None] * world
            dist.all_gather_object(oks, ok, group=cpu_group)
            if not all(oks):
                ctx.peer_disconnect()
                collective = "nccl (peer memory unavailable)"
                box = [comm_unique_id() if rank == 0 else None]
                dist.broadcast_object_list(box, src=0, group=cpu_group)
                ctx.comm_init_rank(world, rank, box[0])
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    step_no = [0]
    no_coll = bool(os.environ.get("BENCH_DEBUG_NO_COLLECTIVE"))       # tuning only: what the collective costs a step (the row check below then fails)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor(x, dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().numpy().tolist()

    to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    planes = wl.sample_pos is not None        # a batch: the samples are position planes (default threshold: no mapq plane)

    def sample_records(c_pos, s):
        """records of sample s for the chromosome at position c_pos of `mine`, as the resident leg holds them"""
        c = mine[c_pos]
        return wl.sample_pos[s][c.index] if planes else c.reads

    # ---- leg 1: inputs resident in HBM (CG_PTR_DEVICE) ------------------------------------
    NOUT = 8
    res = []
    flatten_s = 0.0
    for k, c in enumerate(mine):
        t_fl = time.perf_counter()
        runs = flatten_mappability(c.maprows, c.length) if (has_map and c.maprows is not None) else None
        flatten_s += time.perf_counter() - t_fl
        res.append(dict(c=c, gc=to_dev(c.gc), dels=to_dev(c.dels), dups=to_dev(c.dups),
                        runs=to_dev(runs) if runs is not None and len(runs) else None,
                        n_runs=0 if runs is None else len(runs), runs_host=runs,
                        # result buffers: per position in a submission, NOUT sets in rotation (more than the six submissions in
                        # flight: the genotype kernels of two submissions may finish in either order)
                        out_del=[[torch.empty(max(1, len(c.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev) for _ in range(KSUB)] for _ in range(NOUT)],
                        out_dup=[[torch.empty(max(1, len(c.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev) for _ in range(KSUB)] for _ in range(NOUT)],
                        stats=[[torch.zeros(STATS_DT.itemsize, dtype=torch.uint8, device=dev) for _ in range(KSUB)] for _ in range(NOUT)],
                        reads=[to_dev(sample_records(k, s)) for s in range(S)], n_reads=[len(sample_records(k, s)) for s in range(S)]))
    torch.cuda.synchronize()

    def device_jobs(s, g, ob):
        """the jobs of sample s as the g-th sample of a submission, results into buffer set ob"""
        if not res:
            return np.zeros(0, dtype=CHROM_JOB_DT)
        return np.concatenate([chrom_job(r["c"].index + g * n_chrom, r["c"].length, r["gc"], len(r["c"].gc), mq,
                                         None if planes else r["reads"][s], r["n_reads"][s], r["runs"], r["n_runs"], has_map,
                                         r["dels"], len(r["c"].dels), r["out_del"][ob][g], r["dups"], len(r["c"].dups), r["out_dup"][ob][g], r["stats"][ob][g],
                                         reads_pos=r["reads"][s] if planes else None) for r in res])
    jobs_dev = {}

    def group_jobs(s0, n, ob):
        """one submission: samples s0 .. s0 + n - 1 (mod S)"""
        key = (s0 % S, n, ob)
        if key not in jobs_dev:
            jobs_dev[key] = np.concatenate([device_jobs((s0 + g) % S, g, ob) for g in range(n)])
        return jobs_dev[key]
    reads_of_sample = [sum(len(sample_records(k, s)) for k in range(len(mine))) for s in range(S)]
    last_group = [0, 0]                        # samples and buffer set of the last submission
    group_no = [0]

    def steps_device(k_steps):
        """k_steps samples, KSUB per submission"""
        done = 0
        while done < k_steps:
            n = min(KSUB, k_steps - done)
            k = step_no[0]
            step_no[0] += n
            done += n
            if world > 1:
                ctx.stats_bank(group_no[0] & 3)   # the exchange of a submission's rows is in flight while the next three fill the other banks
            ob = group_no[0] % NOUT
            group_no[0] += 1
            ctx.run_chromosomes(group_jobs(k, n, ob))     # ONE submission: every kernel of the path once for all chromosomes of this rank
            last_group[0], last_group[1] = n, ob
            if world > 1 and not no_coll:
                ctx.stats_allreduce(n_chrom * n)

    def timed(steps_fn, k, finish=None):
        """K steps bracketed by barrier + synchronize; device time by CUDA events on cg_stream (which cg_join orders
        behind every stream of the context), max over ranks; also the host time spent issuing the K steps."""
        barrier()
        ctx.sync()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record()
        t0 = time.perf_counter()
        steps_fn(k)
        ctx.join()
        issue_ms = (time.perf_counter() - t0) * 1e3
        with torch.cuda.stream(ext):
            e1.record()
        if finish:
            finish()
        ctx.sync()
        torch.cuda.synchronize()
        barrier()
        return max_over_ranks([e0.elapsed_time(e1), issue_ms])

    ctx.set_pointer_mode(CG_PTR_DEVICE)
    steps_device(max(3, args.warmup) * KSUB)
    ctx.sync()
    sampler = make_sampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count
    regions = [timed(steps_device, args.steps) for _ in range(args.repeats)]
    launches = (ctx.launch_count - launches0) // args.repeats
    last_sample_dev = (step_no[0] - 1) % S
    dev_ms = sorted(r[0] for r in regions)
    ms_dev = float(np.median(dev_ms))
    issue_ms_dev = float(np.median([r[1] for r in regions]))
    # the calls of the resident leg's last step, for the parity checks below
    g_last = last_group[0] - 1                 # ... which was the g_last-th sample of its submission
    ob_last = last_group[1]
    dev_calls = {r["c"].index: (r["out_del"][ob_last][g_last].cpu().numpy()[: len(r["c"].dels) * CALL_DT.itemsize].view(CALL_DT).copy(),
                                r["out_dup"][ob_last][g_last].cpu().numpy()[: len(r["c"].dups) * CALL_DT.itemsize].view(CALL_DT).copy()) for r in res}
    # same K steps again with per-launch CUDA events on the context's streams -> roofline of the depth pass.  The
    # slots are chained while profiling (kernels timed alone), so this leg is NOT the one `value` comes from.
    ctx.set_profiling(True)
    ctx.kernel_times()
    timed(steps_device, args.steps)
    k_ms, k_n = ctx.kernel_times()
    ctx.set_profiling(False)
    clocks = sampler.stop() if rank == 0 else None

    # ---- leg 2: end to end through the C-ABI from pinned host buffers (CG_PTR_HOST) -------
    # A pipelined host: page-locked buffers, asynchronous calls; a sample's statistics rows and calls are valid in
    # host memory when the event recorded behind it has passed.  Records travel as the compact planes of
    # cg_push_reads_compact (int32 pos; a uint8 mapq plane only when the threshold can reject a record).  Everything
    # a sample needs is copied every step — GC bytes, CNVs, mappability runs included — except in a batch (config 5),
    # where those are the same for all samples and stay resident (CG_JOB_ANNOTATION_RESIDENT).
    ctx.set_pointer_mode(CG_PTR_HOST)
    ctx.set_async(True)
    if args.e2e_format == "auto":              # a sparse sample (C2: one read per 600 bases) escapes almost every delta: 5 bytes per read
        args.e2e_format = "delta" if wl.n_reads * 200 > wl.n_positions else "planes"
    NBUF = 8                                   # result buffers in rotation: more than the six submissions in flight
    annot_resident = planes
    with_mapq = mq >= 0
    h2d = d2h = rec_h2d = 0
    inputs = []
    for k, c in enumerate(mine):
        r = res[k]
        per_sample = []
        for s in range(S):
            rec = sample_records(k, s)
            pos = rec if planes else np.ascontiguousarray(rec["pos"])
            if args.e2e_format == "delta":
                d_, big_, anc_ = encode_positions_delta(pos)
                pd_ = ctx.pinned_array(len(d_), np.uint8); pd_[:] = d_
                pb_ = ctx.pinned_array(len(big_), np.int32); pb_[:] = big_
                pa_ = ctx.pinned_array(len(anc_), ANCHOR_DT); pa_[:] = anc_
                recs = dict(n=len(pos), reads_delta=pd_, reads_big=pb_, n_big=len(big_), reads_anchor=pa_)
                nbytes = pd_.nbytes + pb_.nbytes + pa_.nbytes
            else:
                pp = ctx.pinned_array(len(pos), np.int32); pp[:] = pos
                recs = dict(n=len(pos), reads_pos=pp)
                nbytes = pp.nbytes
            if with_mapq:
                pm = ctx.pinned_array(len(pos), np.uint8); pm[:] = rec["mapq"]
                recs["reads_mapq"] = pm
                nbytes += pm.nbytes
            recs["nbytes"] = nbytes
            per_sample.append(recs)
        if annot_resident:
            ann = (r["gc"], r["dels"], r["dups"], r["runs"])
        else:
            pg = ctx.pinned_array(len(c.gc), np.int8); pg[:] = c.gc
            pd = ctx.pinned_array(len(c.dels), CNV_DT); pd[:] = c.dels
            pu = ctx.pinned_array(len(c.dups), CNV_DT); pu[:] = c.dups
            fr = r["runs_host"] if r["runs_host"] is not None else np.zeros(0, dtype=MAPRUN_DT)
            prn = ctx.pinned_array(len(fr), MAPRUN_DT); prn[:] = fr
            ann = (pg, pd, pu, prn)
            h2d += pg.nbytes + pd.nbytes + pu.nbytes + prn.nbytes
        h2d += per_sample[0]["nbytes"]
        rec_h2d += per_sample[0]["nbytes"]
        d2h += (len(c.dels) + len(c.dups)) * CALL_DT.itemsize + STATS_DT.itemsize
        inputs.append((c, per_sample, ann))
    outs = [[(ctx.pinned_array(len(c.dels), CALL_DT), ctx.pinned_array(len(c.dups), CALL_DT), ctx.pinned_array(1, STATS_DT))
             for c in mine] for _ in range(NBUF)]

    def host_jobs(s, buf):
        if not inputs:
            return np.zeros(0, dtype=CHROM_JOB_DT)
        return np.concatenate([chrom_job(c.index, c.length, ann[0], len(c.gc), mq, None, ps[s]["n"], ann[3], r["n_runs"], has_map,
                                         ann[1], len(c.dels), od, ann[2], len(c.dups), ou, st,
                                         flags=CG_JOB_ANNOTATION_RESIDENT if annot_resident else 0,
                                         **{k_: v_ for k_, v_ in ps[s].items() if k_ not in ("n", "nbytes")})
                               for (c, ps, ann), r, (od, ou, st) in zip(inputs, res, outs[buf])])
    cycle = int(np.lcm(S, NBUF))              # step k uses sample k % S and result buffers k % NBUF
    jobs_host = {(k % S, k % NBUF): host_jobs(k % S, k % NBUF) for k in range(cycle)} if cycle <= 256 else None
    e2e_no = [0]

    def step_e2e():
        k = e2e_no[0]
        e2e_no[0] += 1
        if world > 1:
            ctx.stats_bank(k & 3)
        s, b = k % S, k % NBUF
        ctx.run_chromosomes(jobs_host[(s, b)] if jobs_host is not None else host_jobs(s, b))
        if world > 1:
            ctx.stats_allreduce(n_chrom)

    def steps_e2e(k_steps):
        for _ in range(k_steps):
            step_e2e()

    steps_e2e(max(3, args.warmup))            # submissions rotate through the slots: their upload buffers settle within a few samples
    ctx.sync()
    regions_e2e = [timed(steps_e2e, args.steps) for _ in range(max(3, args.repeats // 2))]
    ms_e2e = float(np.median([r[0] for r in regions_e2e]))
    issue_ms_e2e = float(np.median([r[1] for r in regions_e2e]))
    e2e_ms_sorted = sorted(r[0] for r in regions_e2e)
    last_e2e = e2e_no[0] - 1
    ctx.set_async(False)

    # ---- parity inside the run --------------------------------------------------------------
    parity = {}
    # (1) the end-to-end leg (host buffers, compact planes) and the resident leg produced the same bits
    if res and last_e2e % S == last_sample_dev:
        ok = all(dev_calls[c.index][0].tobytes() == od.tobytes() and dev_calls[c.index][1].tobytes() == ou.tobytes()
                 for c, (od, ou, st) in zip(mine, outs[last_e2e % NBUF]))
        assert ok, "device-mode and host-mode results differ"
        parity["host_mode_equals_device_mode"] = True
    # (2) every rank holds every chromosome's statistics row after the all-reduce
    if world > 1:
        for b in (0, 1, 2, 3):
            ctx.stats_bank(b)
            for c in wl.chroms:
                assert int(ctx.stats_row(c.index)["window_per_gc"].sum()) == c.length, "statistics row %d wrong after the all-reduce" % c.index
        parity["allreduced_rows_on_every_rank"] = True
    # (3) the N-rank calls, gathered in SONIC order, against ONE GPU running the whole sample alone (rank 0)
    empty_calls = (np.zeros(0, dtype=CALL_DT), np.zeros(0, dtype=CALL_DT))
    gathered = multi.gather_calls(dev_calls, n_chrom, group=cpu_group) if world > 1 else [dev_calls.get(i, empty_calls) for i in range(n_chrom)]
    if rank == 0:
        digest = calls_digest(gathered)
        parity["calls_sha256_16"] = digest
        if world > 1:
            ctx.set_pointer_mode(CG_PTR_DEVICE)
            alone = []
            for c in wl.chroms:               # chromosome by chromosome: rank 0 never needs the whole genome resident twice
                rec = wl.sample_pos[last_sample_dev][c.index] if planes else c.reads
                t_gc, t_rec, t_d, t_u = to_dev(c.gc), to_dev(rec), to_dev(c.dels), to_dev(c.dups)
                runs = flatten_mappability(c.maprows, c.length) if (has_map and c.maprows is not None) else None
                t_runs = to_dev(runs) if runs is not None and len(runs) else None
                o_d = torch.empty(max(1, len(c.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
                o_u = torch.empty(max(1, len(c.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)
                torch.cuda.synchronize()
                ctx.run_chromosomes(chrom_job(c.index, c.length, t_gc, len(c.gc), mq, None if planes else t_rec, len(rec), t_runs,
                                              0 if runs is None else len(runs), has_map, t_d, len(c.dels), o_d, t_u, len(c.dups), o_u, None,
                                              reads_pos=t_rec if planes else None))
                ctx.sync()
                alone.append((o_d.cpu().numpy()[: len(c.dels) * CALL_DT.itemsize].view(CALL_DT).copy(),
                              o_u.cpu().numpy()[: len(c.dups) * CALL_DT.itemsize].view(CALL_DT).copy()))
            parity["single_gpu_sha256_16"] = calls_digest(alone)
            parity["n_ranks_equal_one_gpu"] = parity["single_gpu_sha256_16"] == digest
        flagged = int(sum(int((d["margin_flags"] != 0).sum()) + int((u["margin_flags"] != 0).sum()) for d, u in gathered))
        parity["calls_with_margin_flags"] = flagged
    if world > 1:
        barrier()

    n_reads = float(np.mean(reads_of_sample)) if S > 1 else float(reads_of_sample[0])
    if world > 1:
        t = torch.tensor([n_reads, h2d, d2h, launches, rec_h2d], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        n_reads, h2d, d2h, launches, rec_h2d = float(t[0].item()), int(t[1].item()), int(t[2].item()), int(t[3].item()), int(t[4].item())
    rec_h2d_all = float(rec_h2d)
    n_cnvs = wl.n_cnvs
    value = n_reads * args.steps / (ms_dev / 1e3)
    e2e = n_reads * args.steps / (ms_e2e / 1e3)

    # roofline of the dominant kernel (fused depth pass), this rank's chromosomes, record bytes as the resident leg holds them
    rec_bytes = 4 if planes else 8
    alg = sum(rec_bytes * int(np.mean(r["n_reads"])) + 2 * r["c"].length + (r["c"].length + 99) // 100 for r in res)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    pass_ms, pass_n = k_ms[1], max(1, k_n[1])
    achieved = alg * args.steps / (pass_ms / 1e3) / 1e9 if pass_ms > 0 else 0.0      # a launch carries KSUB samples' worth of alg
    # DRAM bytes per launch from the committed ncu capture of this command (profiles/traffic.json, written by
    # scripts/traffic_summary.py from scripts/gpu_traffic.sh); only meaningful for the workload it was taken on
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath) and args.scale == 1.0 and world == 1:
        tj = json.load(open(tpath))
        if tj.get("workload", "c2") == args.workload:
            traffic, traffic_src = int(tj["traffic_bytes_per_launch"]), tj["source"]
    roofline = {"kernel": "k_depth_pass", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "timed": "profiling leg: per-launch CUDA events on the launching stream, slots chained so that the pass runs alone; "
                         "`value` comes from the pipelined leg where consecutive samples overlap",
                "algorithmic_bytes_per_launch": int(alg * args.steps / pass_n), "launches": int(pass_n), "avg_launch_ms": round(pass_ms / pass_n, 4),
                "frac_whole_step": round(alg / (ms_dev / args.steps / 1e3) / 1e9 / peak, 4),
                "kernel_ms_per_step": {"init+mark+finalize": round(k_ms[0] / args.steps, 4), "depth_pass": round(k_ms[1] / args.steps, 4),
                                       "run_index+genotype": round(k_ms[2] / args.steps, 4)}}
    cnv_ms = k_ms[2] / args.steps
    spread = lambda xs: {"min": round(xs[0] / args.steps, 4), "median": round(float(np.median(xs)) / args.steps, 4), "max": round(xs[-1] / args.steps, 4)}
    out = {"metric": METRIC, "value": round(value, 1), "unit": "reads/s", "n_gpus": world, "steps": args.steps,
           "warmup": max(3, args.warmup), "ms_per_step": round(ms_dev / args.steps, 4), "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "int16 depth, int32 sums, f32 ordered sum, f64 likelihood",
           "data": "synthetic", "impl": "b200",
           "config": {"workload": wl.name, "chromosomes": n_chrom, "positions": wl.n_positions, "reads": int(round(n_reads)),
                      "cnvs": n_cnvs, "mappability": has_map, "min_mapq": mq, "samples": S,
                      "l2": "inputs+outputs per step (%.1f GB) larger than L2" % (wl.algorithmic_bytes_depth_pass() / 1e9)},
           "run": {"partition": "chromosomes over ranks, balanced by length (LPT + seeded local search: %.4f of ideal); one submission per rank and %s" % (multi.balance([c.length for c in wl.chroms], parts), "sample" if KSUB == 1 else "%d consecutive samples (resident leg)" % KSUB),
                   "samples_per_submission": KSUB, "slots": ctx_slots(ctx),
                   "collective": None if world == 1 else ("rows stored into every rank's table by k_finalize over NVLink peer memory + one-warp flag exchange (cg_peer_*)"
                                                         if not collective.startswith("nccl") else "ncclAllReduce(int64, sum) of the statistics table, issued by the library [%s]" % collective),
                   "repeats": args.repeats, "ms_per_step_spread": spread(dev_ms), "host_issue_ms_per_step": round(issue_ms_dev / args.steps, 4),
                   "resident_record_bytes": rec_bytes,
                   # host-side input preparation outside every timed region, once per BED file (the reference re-parses and re-paints
                   # the BED per chromosome inside its find_SVs, svs.c:317-377): BED rows -> disjoint runs, this rank's chromosomes
                   "mappability_flatten_ms_once": round(flatten_s * 1e3, 2) if has_map else None},
           "cnvs_per_s": round(n_cnvs * args.steps / (ms_dev / 1e3), 1),
           "cnvs_per_s_genotype_kernel_alone_rank0": round(sum(len(c.dels) + len(c.dups) for c in mine) / (cnv_ms / 1e3), 1) if cnv_ms > 0 else None,
           "e2e": {"value": round(e2e, 1), "unit": "reads/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                   "ms_per_step_spread": spread(e2e_ms_sorted), "host_issue_ms_per_step": round(issue_ms_e2e / args.steps, 4),
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "record_format": ("delta-coded positions (cg_encode_positions_delta)" if args.e2e_format == "delta" else "int32 position plane")
                                    + (" + uint8 mapq plane" if with_mapq else ""),
                   "bytes_per_read": round(rec_h2d_all / max(1.0, n_reads), 3), "annotation_resident": annot_resident,
                   "cnvs_per_s": round(n_cnvs * args.steps / (ms_e2e / 1e3), 1)},
           "gpu_launches": int(launches), "gpu_launches_per_step": round(launches / args.steps, 2),
           "roofline": roofline, "parity": parity, "clocks": clocks}
    if rank == 0:
        out["parity_ok"] = all(v for k, v in parity.items() if isinstance(v, bool))
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_port(wl, budget_s=args.cpu_budget, gpu_calls=dev_calls, sample=last_sample_dev)
        if "parity_vs_oracle" in out["cpu_baseline"]:
            out["parity"]["oracle"] = out["cpu_baseline"].pop("parity_vs_oracle")
            out["parity_ok"] = out["parity_ok"] and out["parity"]["oracle"]["ok"]
    if rank == 0:
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and not out.get("parity_ok", True):
        sys.exit("parity check failed: %s" % json.dumps(parity))


def ctx_slots(ctx):
    return env_int("BENCH_SLOTS", 6)


# ------------------------------------------------------------------------------------------
def sample_chromosome(wl, i, s):
    """chromosome i of sample s as a synth.Chromosome (a batch keeps samples > 0 as position planes)"""
    c = wl.chroms[i]
    if wl.sample_pos is None or s == 0:
        return c
    import copy
    from conga_b200 import READ_DT
    r = np.zeros(len(wl.sample_pos[s][i]), dtype=READ_DT)
    r["pos"], r["len"], r["mapq"] = wl.sample_pos[s][i], 70, 60
    c2 = copy.copy(c)
    c2.reads = r
    return c2


def cpu_baseline_port(wl, budget_s=15.0, gpu_calls=None, sample=0):
    """The oracle (C restatement of the reference loops, oracle/conga_oracle.c) on ONE host core over a
    bounded sample: chromosomes from the smallest up until ~budget_s seconds of CPU work.  The calls the GPU
    produced for those chromosomes in the timed run are compared with the oracle's on the way (observed,
    expected, mappability, copy number bit-exact; likelihoods and c-score within 1e-9 relative)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_py
    orc = oracle_py.load()
    order = sorted(range(len(wl.chroms)), key=lambda i: wl.chroms[i].length)
    t_tot, reads, cnvs, used, ok, checked = 0.0, 0, 0, [], True, 0
    for i in order:
        c = sample_chromosome(wl, i, sample)
        t0 = time.perf_counter()
        _depth, _st, _exp, d_ref, u_ref = orc.chromosome(c, wl.mq_threshold, wl.has_mappability)
        t_tot += time.perf_counter() - t0
        reads += len(c.reads)
        cnvs += len(c.dels) + len(c.dups)
        used.append(c.name)
        if gpu_calls is not None and c.index in gpu_calls:
            for got, want in zip(gpu_calls[c.index], (d_ref, u_ref)):
                same = (np.array_equal(got["observed"], want["observed"]) and got["expected"].tobytes() == want["expected"].tobytes()
                        and got["mappability"].tobytes() == want["mappability"].tobytes()
                        and np.array_equal(got["copy_number"], want["copy_number"])
                        and np.array_equal(np.signbit(got["score"]), np.signbit(want["score"])))
                for f in ("lhomo", "lhetero", "lnone", "score"):
                    same = same and bool(np.allclose(got[f], want[f], rtol=1e-9, atol=0, equal_nan=True))
                ok = ok and bool(same)
                checked += len(want)
        if t_tot >= budget_s:
            break
    out = {"value": round(reads / t_tot, 1), "unit": "reads/s", "cores": 1, "kind": "port",
           "cnvs_per_s": round(cnvs / t_tot, 1), "seconds": round(t_tot, 3),
           "sample": "chromosomes %s of the same workload (%d reads, %d CNVs), whole per-chromosome path"
                     % (",".join(used), reads, cnvs)}
    if gpu_calls is not None:
        out["parity_vs_oracle"] = {"ok": ok, "calls_checked": checked, "chromosomes": ",".join(used)}
    return out


def reference_arm(args):
    """The reference's own CPU implementation of the path on the host cores.  Preferred: oracle/_ref/ref_path
    (UNMODIFIED reference objects: init_rd_per_chr, count_reads_bam, calc_mean_per_chr, find_SVs, sonic_load,
    fed from record files instead of BGZF), one process per chromosome because the program is single-threaded.
    Else: the oracle port, one process per chromosome."""
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    wl = build_workload(args.workload, args.scale, 1)
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
    # One rule at every N and on every box: as many chromosomes as there are host cores (smallest first, one
    # single-threaded process each — the reference has no threads), further limited so that a step's longest
    # process stays inside the per-step share of a 150 s budget at a fixed 0.03 s per megabase (the reference's
    # measured pace on these hosts is 0.013-0.02 s/Mb; a constant, not a calibration, keeps the subset the same
    # from run to run).
    total_budget = 150.0
    per_step = total_budget / max(1, args.steps + args.warmup)
    subset = [i for i in order if wl.chroms[i].length / 1e6 * 0.03 <= per_step][:ncpu] or order[:1]
    base_args = refdata.write_workload(tmp, wl, chrom_subset=subset) if kind == "reference" else []
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
                      "reads": wl.n_reads, "cnvs": wl.n_cnvs, "mappability": wl.has_mappability, "min_mapq": wl.mq_threshold,
                      "samples": 64 if args.workload == "c5" else 1,
                      "l2": "inputs+outputs per step (%.1f GB) larger than L2" % (wl.algorithmic_bytes_depth_pass() / 1e9)},
           "run": {"host_cores_used": cores, "host_cores_available": ncpu, "processes": len(subset)},
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--repeats", type=int, default=10, help="timed regions of --steps steps each; the median is reported")
    ap.add_argument("--samples", type=int, default=64, help="c5 only: samples in the batch")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the genome (debugging only; 1.0 = the named config)")
    ap.add_argument("--e2e-format", default="auto", choices=["auto", "delta", "planes"],
                    help="record stream of the end-to-end leg: delta-coded positions (1 byte per read and 4 more per gap >= 255) or an "
                         "int32 position plane (4); auto = delta when the mean gap between reads is below 200 bases")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--emulate-rank", default="", help="R/N: run rank R's share of an N-rank partition alone on one GPU (tuning only)")
    ap.add_argument("--samples-per-submission", type=int, default=0,
                    help="samples one cg_run_chromosomes call carries in the resident leg (0 = auto: as many as fit the 32 jobs of a "
                         "submission, at most 8, when the chromosomes are shared among ranks; 1 on one GPU)")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        own_arm(args)


if __name__ == "__main__":
    main()
