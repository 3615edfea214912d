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


def hbm(src, dst):
    """ncu --csv log with dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum per launch -> per-kernel rows with
    achieved DRAM GB/s (long format: one row per launch and metric)."""
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    h = rows[0]
    ii, ki, mi, vi, ui = h.index("ID"), h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
    scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0,
             "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}
    by = {}
    for r in rows[1:]:
        d = by.setdefault(int(r[ii]), {"kernel": r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")})
        d[r[mi]] = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    peak = None
    try:
        import os
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    out = []
    for k in sorted(by):
        d = by[k]
        t = d.get("gpu__time_duration.sum", 0.0)
        b = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        if t < 20e-6:       # helpers below 20 us say nothing about bandwidth
            continue
        out.append({"id": k, "kernel": d["kernel"], "ms": t * 1e3, "dram_read_MB": d.get("dram__bytes_read.sum", 0.0) / 1e6,
                    "dram_write_MB": d.get("dram__bytes_write.sum", 0.0) / 1e6, "dram_gbs": b / t / 1e9 if t else None,
                    "frac_of_peak": (b / t / 1e9 / peak) if (t and peak) else None})
    json.dump({"peak_gbs": peak, "launches": out}, open(dst, "w"), indent=1)


if __name__ == "__main__":
    {"launches": launches, "screen": screen, "hbm": hbm}[sys.argv[1]](sys.argv[2], sys.argv[3])
