#!/usr/bin/env python
"""Per-source-line hot spots from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`.
usage: python tools/ncu_lines.py src.csv [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
recs = []; cur_file = None; hdr = None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[2] == "-":   # source-line aggregate rows have Address '-'
        try:
            ie = float(r[hdr.index("Instructions Executed")]); te = float(r[hdr.index("Thread Instructions Executed")])
            smp = float(r[hdr.index("# Samples")] or 0)
        except ValueError:
            continue
        recs.append((ie, te, cur_file, r[0], r[1][:100], smp))
tot = sum(r[0] for r in recs); tott = sum(r[1] for r in recs); tots = sum(r[5] for r in recs)
print(f"total warp-inst {tot:.3e}, thread-inst {tott:.3e} (avg active {tott/tot:.2f}), samples {tots:.0f}")
recs.sort(key=lambda x: -x[0])
for ie, te, f, ln, src, smp in recs[:top]:
    print(f"{100*ie/tot:5.1f}% inst {100*smp/max(tots,1):5.1f}% smp act {te/max(ie,1):5.1f}  {str(f).split('/')[-1]}:{ln}  {src.strip()}")
