"""gpurun_out/traffic.csv (scripts/gpu_traffic.sh) -> profiles/traffic.json: mean DRAM bytes per k_depth_pass launch."""
import csv, json, sys
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/traffic.csv"
tag = sys.argv[2] if len(sys.argv) > 2 else "r01"
lines = [l for l in open(src) if not l.startswith("==")]
per = {}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "ns": 1e-9, "us": 1e-6, "ms": 1e-3}
for row in csv.DictReader(lines):
    d = per.setdefault(row["ID"], {})
    d[row["Metric Name"]] = float(row["Metric Value"].replace(",", "")) * scale.get(row["Metric Unit"], 1)
n = len(per)
rd = sum(d["dram__bytes_read.sum"] for d in per.values())
wr = sum(d["dram__bytes_write.sum"] for d in per.values())
t = sum(d["gpu__time_duration.sum"] for d in per.values())
out = {"kernel": "k_depth_pass", "launches": n, "dram_bytes_read_per_launch": rd / n, "dram_bytes_write_per_launch": wr / n,
       "traffic_bytes_per_launch": (rd + wr) / n, "ncu_time_s_per_launch": t / n,
       "workload": sys.argv[3] if len(sys.argv) > 3 else "c3",
       "source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none, %d launches of k_depth_pass, one per sample (%s); "
                 "bytes still dirty in the 126 MB L2 when a launch ends are not counted" % (n, tag)}
json.dump(out, open("profiles/traffic.json", "w"), indent=1)
print(json.dumps(out))
