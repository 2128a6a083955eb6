"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the LAST step
(from the last `marker` kernel on).  usage: ncu_summary.py launches.csv [marker-prefix]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
marker = sys.argv[2] if len(sys.argv) > 2 else "embed_split"
hdr, seq = None, []
for r in rows:
    if len(r) > 5 and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        name = d["Kernel Name"].split("(")[0].replace("<unnamed>::", "")[:70]
        try:
            v = float(d["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        u = d["Metric Unit"]
        seq.append((name, v / 1e6 if u.startswith("n") else (v / 1e3 if u.startswith("u") else v)))
idx = [i for i, (n, _) in enumerate(seq) if n.startswith(marker)]
step = seq[idx[-1]:] if idx else seq
if len(idx) >= 2 and len(step) < 50:   # the capture window ended a few launches into a new step: take the last complete one
    step = seq[idx[-2]:idx[-1]]
agg = collections.defaultdict(lambda: [0, 0.0])
for n, ms in step:
    agg[n][0] += 1
    agg[n][1] += ms
tot = sum(v[1] for v in agg.values())
print(f"launches captured {len(seq)}; last step: {len(step)} launches, {tot:.3f} ms (serialised, cold cache)")
for n, (c, ms) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
    print(f"{ms:10.3f} ms {100 * ms / tot:5.1f}%  x{c:<4d} {n}")
