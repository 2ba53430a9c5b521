"""Host-to-device copy bandwidth of the copy engine while the depth pass keeps HBM busy: the resident-mode C2
loop runs in a thread, 64 MB pinned copies are timed on another stream.  python scripts/ce_under_load.py"""
import json, os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from conga_b200 import abi, synth, CG_PTR_DEVICE, CALL_DT, chrom_job, flatten_mappability

dev = torch.device("cuda:0")
wl = synth.config2()
ctx = abi.Context(0)
ctx.set_pointer_mode(CG_PTR_DEVICE)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)
keep, jobs = [], []
for c in wl.chroms:
    runs = flatten_mappability(np.ascontiguousarray(c.maprows), c.length)
    b = (up(c.gc), up(c.reads), up(runs), up(c.dels), up(c.dups), torch.empty(len(c.dels) * 56, dtype=torch.uint8, device=dev),
         torch.empty(len(c.dups) * 56, dtype=torch.uint8, device=dev))
    keep.append(b)
    jobs.append(chrom_job(c.index, c.length, b[0], len(c.gc), wl.mq_threshold, b[1], len(c.reads), b[2], len(runs), True,
                          b[3], len(c.dels), b[5], b[4], len(c.dups), b[6], None))
jobs = np.concatenate(jobs)

def copy_bw(mb, k):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device=dev)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(2): d.copy_(h, non_blocking=True)
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k): d.copy_(h, non_blocking=True)
        e1.record(); st.synchronize()
    return round(n * k / e0.elapsed_time(e1) / 1e6, 1)

out = {"idle_GBps": {mb: copy_bw(mb, max(4, 256 // mb)) for mb in (2, 64)}}
stop = False
def load():
    torch.cuda.set_device(0)
    n = 0
    while not stop:
        ctx.run_chromosomes(jobs)
        n += 1
        if n % 4 == 0:
            ctx.sync()
t = threading.Thread(target=load); t.start()
time.sleep(0.5)
out["under_pass_GBps"] = {mb: copy_bw(mb, max(4, 256 // mb)) for mb in (2, 64)}
stop = True; t.join(); ctx.sync()
print(json.dumps(out))
