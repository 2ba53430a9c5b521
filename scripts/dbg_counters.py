import ctypes, sys, os, numpy as np
sys.path.insert(0, "/root/repo")
from conga_b200 import abi, synth, hotpath
abi._lib = abi.load_library("/root/repo/build/variants/libconga_gpu_dbg.so")
lib = abi._lib
ctx = abi.Context(0)
wl = synth.config2(scale=0.25)
out = (ctypes.c_ulonglong * 16)()
lib.cg_debug_counters(out, 1)
n = 0
for ch in wl.chroms[:6]:
    hotpath.run_chromosome(ctx, ch, -1, True); n += len(ch.dels) + len(ch.dups)
lib.cg_debug_counters(out, 0)
names = ["exp iterations", "exp rounds", "exp rounds acc_odd", "exp rounds with bad lane", "map rounds", "map rounds with bad lane", "map iterations"]
print("CNVs", n)
for i, nm in enumerate(names): print("%-28s %9d  per CNV %.2f" % (nm, out[i], out[i] / n))
