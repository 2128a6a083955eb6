"""Summarise an `ncu --metrics ... --csv` pass (one row per launch and metric) as one line per launch: duration, tensor-pipe activity,
DRAM bytes, registers.  usage: ncu_metrics_summary.py metrics.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = None, collections.OrderedDict()
for r in rows:
    if len(r) > 5 and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        key = (int(d["ID"]), d["Kernel Name"].split("(")[0].replace("<unnamed>::", ""), d.get("Grid Size"), d.get("Block Size"))
        data.setdefault(key, {})[d["Metric Name"]] = (d["Metric Value"], d["Metric Unit"])


def f(v):
    return float(v[0].replace(",", ""))


print(f"{'id':>4} {'kernel':28} {'grid':>14} {'us':>9} {'tensor%':>8} {'dram rd MB':>11} {'dram wr MB':>11} {'dram%':>6} {'regs':>5} {'warps%':>7}")
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for (i, name, g, b), m in data.items():
    t, u = f(m["gpu__time_duration.sum"]), m["gpu__time_duration.sum"][1]
    us = t / 1e3 if u.startswith("n") else (t if u.startswith("u") else t * 1e3)
    tp = f(m["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"])
    rd, wr = f(m["dram__bytes_read.sum"]) / 1e6, f(m["dram__bytes_write.sum"]) / 1e6
    print(f"{i:4d} {name:28} {g:>14} {us:9.1f} {tp:8.2f} {rd:11.1f} {wr:11.1f} {f(m['gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']):6.1f} "
          f"{f(m['launch__registers_per_thread']):5.0f} {f(m['sm__warps_active.avg.pct_of_peak_sustained_active']):7.1f}")
    a = agg[name]
    a[0] += 1; a[1] += us; a[2] += tp * us; a[3] += rd + wr
print()
for name, (n, us, tpus, mb) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{name:28} x{n:<3d} {us / 1e3:8.3f} ms  time-weighted tensor-pipe activity {tpus / us:5.1f} %  DRAM {mb / 1e3:7.2f} GB")
