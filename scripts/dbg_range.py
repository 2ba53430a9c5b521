import sys, os, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from conga_b200 import *
from conga_b200 import synth, abi
wl = synth.config2(scale=float(os.environ.get("SCALE", "1.0")))
dev = torch.device("cuda:0")
to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
ctx = Context(0)
res = []
for c in wl.chroms:
    runs = flatten_mappability(c.maprows, c.length)
    res.append(dict(c=c, gc=to_dev(c.gc), reads=to_dev(c.reads), dels=to_dev(c.dels), dups=to_dev(c.dups), runs=to_dev(runs), n_runs=len(runs),
                    out_del=torch.zeros(max(1, len(c.dels)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev),
                    out_dup=torch.zeros(max(1, len(c.dups)) * CALL_DT.itemsize, dtype=torch.uint8, device=dev)))
torch.cuda.synchronize()
def check(tag):
    ctx.set_pointer_mode(CG_PTR_HOST)
    try:
        ctx.sync(); print(tag, "ok")
    except Exception as e:
        print(tag, "ERANGE")
    ctx.set_pointer_mode(CG_PTR_DEVICE)
ctx.set_pointer_mode(CG_PTR_DEVICE)
for step in range(4):
    for i, r in enumerate(res):
        c = r["c"]
        ctx.chrom_begin(c.index, c.length, r["gc"], -1, n_windows=len(c.gc))
        ctx.push_reads(r["reads"], len(c.reads))
        ctx.chrom_finish(None)
        ctx.set_mappability_runs(r["runs"], r["n_runs"], True)
        ctx.genotype(r["dels"], False, r["out_del"], n=len(c.dels))
        ctx.genotype(r["dups"], True, r["out_dup"], n=len(c.dups))
        if os.environ.get("EACH"): check("step %d chrom %d ndel %d ndup %d" % (step, i, len(c.dels), len(c.dups)))
    check("step %d" % step)
