#!/usr/bin/env python
"""Summarise ncu artefacts brought back in gpurun_out/ into small text files for profiles/.

  python scripts/ncu_summary.py launches gpurun_out/launches.csv
  python scripts/ncu_summary.py raw gpurun_out/prof.ncu-rep
  python scripts/ncu_summary.py stalls gpurun_out/prof.ncu-rep
"""
import collections
import csv
import subprocess
import sys

RAW = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
       "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
       "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
       "launch__occupancy_limit_registers", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
       "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u in ("nsecond", "ns") else v * 1e3 if u in ("msecond", "ms") else v
        k = row["Kernel Name"].split("(")[0]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print("%-28s %6s %12s %10s %7s" % ("kernel", "n", "total_us", "avg_us", "share"))
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-28s %6d %12.1f %10.1f %7.3f" % (k, v[0], v[1], v[1] / v[0], v[1] / tot))


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


def raw(rep):
    r = ncu_csv(rep, "raw")
    hdr, units = r[0], r[1]
    for row in r[2:]:
        print(row[hdr.index("Kernel Name")][:70])
        for w in RAW:
            if w in hdr:
                i = hdr.index(w)
                print("  %-62s %s %s" % (w, row[i], units[i]))


def stalls(rep, top=25):
    rows = ncu_csv(rep, "source")
    hdr = rows[1]
    data = []
    for r in rows[2:]:
        if r and r[0] == "Kernel Name":
            break
        if len(r) == len(hdr):
            data.append(r)
    ix = {h: i for i, h in enumerate(hdr)}
    names = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = {s: sum(int(r[ix[s]] or 0) for r in data) for s in names}
    T = max(1, sum(tot.values()))
    print("first profiled launch: %d samples" % T)
    print("  " + "  ".join("%s=%.3f" % (k[6:], v / T) for k, v in sorted(tot.items(), key=lambda x: -x[1]) if v))
    data.sort(key=lambda r: -int(r[ix["# Samples"]] or 0))
    for r in data[:top]:
        st = {s: int(r[ix[s]] or 0) for s in names if int(r[ix[s]] or 0)}
        t3 = sorted(st.items(), key=lambda x: -x[1])[:2]
        print("  %6s smp %9s exec  %-52s %s" % (r[ix["# Samples"]], r[ix["Instructions Executed"]], r[1].strip()[:52],
                                               ",".join("%s:%d" % (a[6:], b) for a, b in t3)))


if __name__ == "__main__":
    {"launches": launches, "raw": raw, "stalls": stalls}[sys.argv[1]](sys.argv[2])
