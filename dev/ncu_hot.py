#!/usr/bin/env python
"""usage: dev/ncu_hot.py <report.ncu-rep> <kernel regex> [min_samples]  -- hottest SASS lines (pc sampling) of the first match"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
mins = int(sys.argv[3]) if len(sys.argv) > 3 else 12
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kern}", "--launch-count", "1"] ,
                     capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
hi = [i for i, row in enumerate(r) if row and row[0] == "Address"][0]
h = r[hi]
si = h.index("# Samples")
rows = [row for row in r[hi + 1:] if len(row) > si and row[0].startswith("0x")]
tot = sum(int(row[si] or 0) for row in rows)
print("total samples", tot, "instructions", len(rows))
for i, row in enumerate(rows):
    s = int(row[si] or 0)
    if s >= mins:
        print(f"{i:6d} {s:6d}  {row[1].strip()}")
