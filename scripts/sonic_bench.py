"""SONIC builder: the unmodified `sonic --make-sonic` (oracle/_ref/sonic, one core) against
`sonic-b200` on the same FASTA, outputs compared after inflation; then k_gc_profile alone on resident bases
(CUDA events through cg_kernel_times).  Measurement tooling (may use oracle/), not product.

  python scripts/sonic_bench.py [--mb 600] [--records 3] [--skip-reference]
"""
import argparse, gzip, json, os, shutil, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

REF = os.path.join(ROOT, "oracle", "_ref", "sonic")
HOST = os.path.join(ROOT, "conga_b200", "host", "sonic-b200")

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=float, default=600.0)
ap.add_argument("--records", type=int, default=3)
ap.add_argument("--skip-reference", action="store_true")
ap.add_argument("--kernel-only", action="store_true")
ap.add_argument("--kernel-bases", type=int, default=249_250_621)
args = ap.parse_args()
rng = np.random.default_rng(20221124 + 9)
tmp = tempfile.mkdtemp(prefix="sonicbench_")
res = {"fasta_mb": args.mb, "records": args.records, "host_cores": len(os.sched_getaffinity(0))}
try:
    if args.kernel_only:
        raise StopIteration
    alpha = np.frombuffer(b"ACGTacgtNN", dtype=np.uint8)
    fa = os.path.join(tmp, "ref.fa")
    total = 0
    with open(fa, "wb") as f, open(fa + ".fai", "w") as fai:
        off = 0
        for k in range(args.records):
            L = int(args.mb * 1e6 / args.records) // 60 * 60 + 17 * k
            s = alpha[rng.integers(0, len(alpha), size=L)]
            full = (L // 60) * 60
            body = np.concatenate([s[:full].reshape(-1, 60), np.full((full // 60, 1), 10, dtype=np.uint8)], axis=1).tobytes()
            body += s[full:].tobytes() + (b"\n" if L > full else b"")
            head = b">%d\n" % (k + 1)
            f.write(head + body)
            off += len(head)
            fai.write("%d\t%d\t%d\t60\t61\n" % (k + 1, L, off))
            off += len(body)
            total += L
    res["bases"] = total
    for nm in ("gaps.bed", "dups.bed"):
        open(os.path.join(tmp, nm), "w").close()
    open(os.path.join(tmp, "rm.out"), "w").write(
        "   SW   perc perc perc  query     position in query    matching  repeat       position in repeat\n"
        "score   div. del. ins.  sequence  begin end   (left)   repeat    class/family begin end (left) ID\n\n")
    outs = {}
    for tag, exe in (("b200", HOST), ("reference", REF)):
        if tag == "reference" and args.skip_reference:
            continue
        sonic = os.path.join(tmp, tag + ".sonic")
        runs = []
        for _ in range(3 if tag == "b200" else 1):
            t = time.perf_counter()
            p = subprocess.run([exe, "--ref", fa, "--gaps", os.path.join(tmp, "gaps.bed"), "--dups", os.path.join(tmp, "dups.bed"),
                                "--reps", os.path.join(tmp, "rm.out"), "--info", "bench", "--make-sonic", sonic],
                               cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, env=dict(os.environ, CONGA_TIMING="1"))
            runs.append(round(time.perf_counter() - t, 3))
            if p.returncode != 0:
                sys.exit("%s failed: %s" % (tag, p.stderr.decode(errors="replace")[-1500:]))
        res[tag + "_s"] = runs
        if tag == "b200":
            res["b200_timing"] = [l for l in p.stderr.decode(errors="replace").split("\n") if l.startswith("[timing]")]
        d = bytearray(gzip.open(sonic, "rb").read()); d[4:12] = b"\0" * 8
        outs[tag] = bytes(d)
    if "reference" in outs:
        res["identical_containers"] = outs["reference"] == outs["b200"]
        res["speedup_whole_program"] = round(res["reference_s"][0] / float(np.median(res["b200_s"])), 1)
except StopIteration:
    pass
finally:
    shutil.rmtree(tmp, ignore_errors=True)

# the kernel alone, bases resident (chr1-sized record)
import torch
from conga_b200 import abi, CG_PTR_DEVICE
dev = torch.device("cuda:0")
n = args.kernel_bases
codes = torch.randint(0, 8, (n,), device=dev, dtype=torch.uint8)
bases = torch.take(torch.tensor(list(b"ACGTgcNn"), dtype=torch.uint8, device=dev), codes.to(torch.int64)); del codes
out = torch.empty((n + 1) // 100, dtype=torch.int8, device=dev)
ctx = abi.Context(0)
ctx.set_pointer_mode(CG_PTR_DEVICE)
for _ in range(3):
    ctx.gc_profile(bases, n_bases=n, out=out)
ctx.sync()
ctx.set_profiling(True); ctx.kernel_times()
K = 20
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(K):
    flush.fill_(1)                                   # L2 flush between launches (another stream: sync first)
    torch.cuda.synchronize()
    ctx.gc_profile(bases, n_bases=n, out=out)
    ctx.sync()
ms, cnt = ctx.kernel_times()
ctx.set_profiling(False)
alg = n + (n + 1) // 100
res["kernel"] = {"bases": n, "avg_ms": round(ms[0] / K, 4), "algorithmic_bytes": alg, "GBps": round(alg / (ms[0] / K * 1e-3) / 1e9, 1),
                 "frac_of_6536": round(alg / (ms[0] / K * 1e-3) / 1e9 / 6536.0, 3)}
ctx.close()
print(json.dumps(res), flush=True)
