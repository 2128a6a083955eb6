"""DRAM bytes per frame of a kernel from `ncu -i rep --page raw --csv` exports -> a ncu_traffic.json entry.
usage: ncu_traffic.py out.json kernel raw.csv frames source-note [kernel raw.csv frames source-note ...]"""
import csv
import json
import sys

out, args = sys.argv[1], sys.argv[2:]
res = {}
for i in range(0, len(args), 4):
    kernel, f, frames, note = args[i], args[i + 1], int(args[i + 2]), args[i + 3]
    rows = list(csv.reader(open(f)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tot = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        v, u = d[k]
        tot += float(v.replace(",", "")) * scale[u]
    res[kernel] = {"dram_bytes_per_frame": tot / frames, "source": note,
                   "tensor_pipe_active_pct": float(d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", ("nan", ""))[0]),
                   "duration_ms": float(d["gpu__time_duration.sum"][0].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(d["gpu__time_duration.sum"][1], 1.0)}
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
