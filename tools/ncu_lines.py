"""Top warp-stall lines of an `ncu -i rep --page source --csv --print-source cuda,sass` export.
usage: ncu_lines.py export.csv [N]"""
import csv
import os
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr, out, kernel = None, None, [], ""
for r in rows:
    if not r:
        continue
    if r[0] in ("Kernel Name", "Function Name"):
        kernel = r[1].split("(")[0]
    elif r[0] in ("File Name", "File Path"):
        cur, hdr = os.path.basename(r[1]), None
    elif r[0] == "Line No":
        hdr = r
    elif hdr and cur and len(r) == len(hdr) and r[0].isdigit() and r[2] == "-":  # a CUDA source line (SASS rows carry an address)
        d = dict(zip(hdr[4:], r[4:]))
        n = int(d.get("# Samples", "0") or 0)
        if n:
            st = sorted(((int(v), k) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v)), reverse=True)[:2]
            out.append((n, f"{cur}:{r[0]}", r[1].strip()[:90], st))
tot = sum(o[0] for o in out)
print(f"{kernel}: warp-stall samples per CUDA source line, total {tot}")
for n, loc, src, st in sorted(out, reverse=True)[:N]:
    print(f"{100.0 * n / tot:5.1f}% {loc:28s} {src:90s} {[(k, v) for v, k in st]}")
