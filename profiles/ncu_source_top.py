#!/usr/bin/env python
"""Top stall sites of one kernel launch from `ncu -i X.ncu-rep --page source --csv --print-source sass`:
   python profiles/ncu_source_top.py source.csv [launch index] [top N]
Prints, per SASS instruction with samples, the sample count, the dominant stall reasons and the executed count."""
import csv
import sys

path = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
blocks, cur = [], None
for row in csv.reader(open(path)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = row
    elif cur is not None and row:
        cur["rows"].append(row)
b = blocks[which]
h = b["hdr"]
col = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[col["# Samples"]] or 0) for r in b["rows"])
print(b["name"], "launch", which, "samples", tot, "instructions", len(b["rows"]))
agg = {s: sum(int(r[col[s]] or 0) for r in b["rows"]) for s in stalls}
print("by reason:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][col["# Samples"]] or 0))[:top]
for i in sorted(order):
    r = b["rows"][i]
    n = int(r[col["# Samples"]] or 0)
    why = sorted(((int(r[col[s]] or 0), s[6:]) for s in stalls), reverse=True)[:3]
    print(f"{i:5d} {n:6d} {100.0 * n / tot:5.1f}%  exec={r[col['Instructions Executed']]:>9s}  {r[col['Source']].strip()[:70]:70s} {[(w, c) for c, w in why if c]}")
