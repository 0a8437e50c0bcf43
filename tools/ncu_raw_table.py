#!/usr/bin/env python
"""Key metrics per launch from `ncu -i X.ncu-rep --page raw --csv` output -> markdown table.
    python tools/ncu_raw_table.py gpurun_out/prof_conv_raw.csv [min_us]
"""
import csv
import sys

COLS = [("gpu__time_duration.sum", "us", 1e-3), ("dram__bytes_read.sum", "rd MB", None), ("dram__bytes_write.sum", "wr MB", None),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 1), ("lts__t_sector_hit_rate.pct", "L2hit%", 1),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%", 1),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor%", 1),
        ("sm__inst_executed_pipe_tensor.sum", "tc inst", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 1), ("launch__registers_per_thread", "regs", 1),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 1)]


def num(v):
    try:
        return float(v.replace(",", ""))
    except Exception:
        return None


def main(path, min_us=0.0):
    rows = list(csv.reader(open(path, newline="")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    avail = [(m, lab, sc) for m, lab, sc in COLS if m in idx]
    print("| # | kernel | grid | " + " | ".join(lab for _, lab, _ in avail) + " |")
    print("|---|---|---|" + "---:|" * len(avail))
    for r in data:
        name = r[idx["Kernel Name"]].split("(")[0][:34]
        vals = []
        skip = False
        for m, lab, sc in avail:
            v = num(r[idx[m]])
            u = units[idx[m]]
            if v is None:
                vals.append("-")
                continue
            if lab == "us":
                v = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u.startswith("ms") else v)
                if v < min_us:
                    skip = True
            if lab.endswith("MB"):
                v = v / 1e6 if u == "byte" else (v / 1e3 if u == "Kbyte" else (v if u == "Mbyte" else v * 1e3))
            vals.append(f"{v:.1f}" if abs(v) < 1e5 else f"{v:.3g}")
        if not skip:
            print(f"| {r[idx['ID']]} | `{name}` | {r[idx['Grid Size']]} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 0.0)
