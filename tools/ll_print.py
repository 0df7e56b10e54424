#!/usr/bin/env python
"""Print the per-kernel durations of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import csv
import sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
tot = 0.0
for r in rows[hdr + 1:]:
    ms = float(r[vi].replace(",", "")) / (1e6 if r[ui] == "ns" else 1e3)
    tot += ms
    print(f"{r[ki][:44]:44s} {ms:.3f} ms")
print(f"{'sum':44s} {tot:.3f} ms")
