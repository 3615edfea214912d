#!/usr/bin/env python
"""Turns ncu exports into the small tracked summaries under profiles/.

  python profiles/ncu_summary.py launches <ncu --csv log of gpu__time_duration.sum> <out.csv>
  python profiles/ncu_summary.py screen <raw.csv of `ncu -i rep --page raw --csv`> <out.json>
"""
import csv
import json
import sys


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)   # kernel names of library kernels contain commas (template arguments)
        w.writerow(["id", "kernel", "gpu__time_duration.sum", "unit"])
        for n, r in enumerate(rows[1:]):
            name = r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
            w.writerow([n, name, r[vi].replace(",", ""), r[ui]])


def screen(src, dst):
    rows = list(csv.reader(open(src)))
    h = rows[0]
    out = []
    for v in rows[2:]:
        def g(k):
            return float(v[h.index(k)].replace(",", ""))
        unit_r = rows[1][h.index("dram__bytes_read.sum")]
        scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(unit_r, 1.0)
        unit_w = rows[1][h.index("dram__bytes_write.sum")]
        scale_w = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(unit_w, 1.0)
        tunit = rows[1][h.index("gpu__time_duration.sum")]
        tscale = {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(tunit, 1.0)
        out.append({
            "kernel": v[h.index("Kernel Name")][:60],
            "gpu__time_duration_ms": g("gpu__time_duration.sum") * tscale,
            "dram__bytes_read_MB": g("dram__bytes_read.sum") * scale / 1e6,
            "dram__bytes_write_MB": g("dram__bytes_write.sum") * scale_w / 1e6,
            "sm__pipe_tensor_cycles_active_pct": g("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
            "sm__issue_active_pct": g("sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
            "lts_hit_rate_pct": g("lts__t_sector_hit_rate.pct"),
            "smsp__inst_executed": g("smsp__inst_executed.sum"),
            "registers_per_thread": g("launch__registers_per_thread"),
        })
    tot = sum((o["dram__bytes_read_MB"] + o["dram__bytes_write_MB"]) * 1e6 for o in out)
    json.dump({"launches": out, "traffic_bytes_per_launch_avg": tot / max(len(out), 1)}, open(dst, "w"), indent=1)


if __name__ == "__main__":
    {"launches": launches, "screen": screen}[sys.argv[1]](sys.argv[2], sys.argv[3])
