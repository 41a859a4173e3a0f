"""Summarise an `ncu --page source --print-source cuda,sass --csv` export per CUDA source line:
    ncu -i X.ncu-rep --page source --print-source cuda,sass --csv --launch-skip N --launch-count 1 > src.csv
    python tools/ncu_lines.py src.csv [min_pct]
Prints file:line, share of executed warp instructions, share of stall samples, top stall reasons."""
import csv
import sys

path = sys.argv[1]
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rows = list(csv.reader(open(path)))
cur_file, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        func = r[1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr and r[0].isdigit():
        lines.append((cur_file, r))
iI, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
num = lambda s: int(s) if s.isdigit() else 0
tot_i = sum(num(r[iI]) for _, r in lines) or 1
tot_s = sum(num(r[iS]) for _, r in lines) or 1
print(func)
print(f"total warp-instructions {tot_i}  samples {tot_s}")
agg = {}
for _, r in lines:
    for i, h in stall_cols:
        agg[h] = agg.get(h, 0) + num(r[i])
print("stalls overall:", ", ".join(f"{h[6:]} {100*v/tot_s:.1f}%" for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for f, r in lines:
    pi, ps = 100 * num(r[iI]) / tot_i, 100 * num(r[iS]) / tot_s
    if pi >= minpct or ps >= minpct:
        st = sorted(((num(r[i]), h[6:]) for i, h in stall_cols), reverse=True)[:3]
        sts = " ".join(f"{h}:{100*v/tot_s:.1f}" for v, h in st if v)
        print(f"{f}:{r[0]:>4} inst {pi:5.1f}% samp {ps:5.1f}% [{sts}]  {r[1].strip()[:100]}")
