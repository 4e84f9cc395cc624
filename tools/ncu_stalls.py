#!/usr/bin/env python
"""Total warp-stall samples by reason from `ncu --page source --print-source cuda,sass --csv` output."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; tot = {}
for r in rows:
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[2] == "-":
        for h, v in zip(hdr, r):
            if h.startswith("stall_") and "Not Issued" not in h:
                try: tot[h] = tot.get(h, 0) + float(v)
                except ValueError: pass
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda x: -x[1]):
    if v: print(f"{k:28s} {100*v/s:5.1f}%")
