"""Static SASS size and executed counts per CUDA source line from an `ncu --page source --csv --print-source cuda,sass` export
(optionally .gz): where the executed instruction footprint of a kernel sits.  usage: ncu_code_hot.py export.csv[.gz] [N]"""
import collections
import csv
import gzip
import os
import sys

f = sys.argv[1]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 50
HOT_MIN = int(sys.argv[3]) if len(sys.argv) > 3 else 0   # count a SASS instruction only if it executed more often than this (warp level, whole launch)
rows = list(csv.reader(gzip.open(f, "rt", errors="replace") if f.endswith(".gz") else open(f, errors="replace")))
cur, hdr, key = None, None, None
static = collections.Counter()   # (file, line) -> SASS instructions
execd = collections.Counter()    # (file, line) -> warp-level executions
samples = collections.Counter()
src = {}
for r in rows:
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        cur, hdr = os.path.basename(r[1]), None
    elif r[0] == "Line No":
        hdr = r
    elif hdr and cur and len(r) == len(hdr) and r[0].isdigit():
        key = (cur, int(r[0]))
        src[key] = r[1].strip()[:70]
    elif hdr and cur and len(r) == len(hdr) and r[0] == "" and r[2].startswith("0x"):   # a SASS row of the line above
        d = dict(zip(hdr[4:], r[4:]))
        n_ex = int(d.get("Instructions Executed", "0") or 0)
        static[key] += 1 if n_ex > HOT_MIN else 0
        execd[key] += n_ex
        samples[key] += int(d.get("# Samples", "0") or 0)
tot_static = sum(static.values())
hot = {k: v for k, v in static.items() if execd[k] > 0}
print(f"SASS instructions executed more than {HOT_MIN} times: {tot_static} ({tot_static * 16 / 1024:.0f} KB)")
by_file = collections.Counter()
for (fl, ln), v in hot.items():
    by_file[fl] += v
print("executed-line SASS by file:", dict(by_file))
print("largest executed lines (static SASS instructions, executions, samples):")
for k, v in sorted(hot.items(), key=lambda kv: -kv[1])[:N]:
    print(f"{v:6d} {execd[k]:12d} {samples[k]:7d}  {k[0]}:{k[1]:<5d} {src.get(k, '')}")
