#!/usr/bin/env python
"""Per-kernel totals from `ncu -i X.ncu-rep --page raw --csv` output -> markdown table (HBM-bound kernels: achieved DRAM GB/s).
    python tools/ncu_group_table.py gpurun_out/r01_v13_hbm.raw.csv [hbm_peak_gbs]
"""
import csv
import json
import os
import sys


def num(v):
    try:
        return float(v.replace(",", ""))
    except Exception:
        return 0.0


def scale(v, u, kind):
    if kind == "us":
        return v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u.startswith("ms") else (v * 1e6 if u in ("s", "second") else v))
    return {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6) * v      # -> MB


def main(path, peak=None):
    if peak is None:
        pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
        peak = 6453.0
        if os.path.exists(pk):
            d = json.load(open(pk))
            for k in ("hbm_gbs", "hbm_copy_gbs", "hbm_GBps"):
                if k in d:
                    peak = float(d[k])
    rows = list(csv.reader(open(path, newline="")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    g = {}
    for r in data:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")[:40]
        us = scale(num(r[idx["gpu__time_duration.sum"]]), units[idx["gpu__time_duration.sum"]], "us")
        rd = scale(num(r[idx["dram__bytes_read.sum"]]), units[idx["dram__bytes_read.sum"]], "mb")
        wr = scale(num(r[idx["dram__bytes_write.sum"]]), units[idx["dram__bytes_write.sum"]], "mb")
        d = g.setdefault(name, {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0, "top": (0.0, None)})
        d["n"] += 1; d["us"] += us; d["rd"] += rd; d["wr"] += wr
        if us > d["top"][0]:
            d["top"] = (us, r)
    cols = [("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
            ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
            ("launch__registers_per_thread", "regs"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%")]
    cols = [(m, l) for m, l in cols if m in idx]
    print("| kernel | launches | total us | rd MB | wr MB | DRAM GB/s (all launches) | frac of HBM peak | longest launch: us | grid | "
          + " | ".join(l for _, l in cols) + " |")
    print("|---|---:|---:|---:|---:|---:|---:|---:|---|" + "---:|" * len(cols))
    for name, d in sorted(g.items(), key=lambda kv: -kv[1]["us"]):
        gbs = (d["rd"] + d["wr"]) * 1e6 / (d["us"] * 1e-6) / 1e9 if d["us"] > 0 else 0.0
        us, r = d["top"]
        print(f"| `{name}` | {d['n']} | {d['us']:.1f} | {d['rd']:.1f} | {d['wr']:.1f} | {gbs:.0f} | {gbs / peak:.2f} | {us:.1f} | {r[idx['Grid Size']]} | "
              + " | ".join(f"{num(r[idx[m]]):.1f}" for m, _ in cols) + " |")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else None)
