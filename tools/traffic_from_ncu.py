#!/usr/bin/env python
"""DRAM traffic per bench step from an `ncu --set full` raw CSV -> profiles/r02_traffic.json (read by bench.py for roofline.traffic).
    python tools/traffic_from_ncu.py gpurun_out/r02_<tag>_conv.raw.csv "<command that was profiled>"
Groups: the tcgen05 convolution kernels (k_conv_tc*, k_gemm_first: the kernels roofline.achieved is taken over) and, beside them,
the generative pair (k_gemm_first + k_cells_reduce) and the occupancy heads (k_conv_tc_mt with projection + k_gather_sum + their
neighbour-table builder), each with the ALGORITHMIC bytes bench.py books for it so the ratio can be read off."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def num(v):
    try:
        return float(v.replace(",", ""))
    except Exception:
        return 0.0


def to_bytes(v, unit):
    return v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)


def main(path, command):
    rows = list(csv.reader(open(path, newline="")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    rd, wr, name = idx["dram__bytes_read.sum"], idx["dram__bytes_write.sum"], idx["Kernel Name"]
    groups = {"tcgen05_conv": ("k_conv_tc", "k_gemm_first"), "gen_transpose": ("k_gemm_first", "k_cells_reduce"),
              "occupancy_heads": ("k_gather_sum", "k_cells_kernel_map3"), "conv_bwd_weight": ("k_conv_bwd_weight",)}
    out = {g: {"launches": 0, "dram_bytes": 0.0} for g in groups}
    for r in data:
        b = to_bytes(num(r[rd]), units[rd]) + to_bytes(num(r[wr]), units[wr])
        for g, pats in groups.items():
            if any(p in r[name] for p in pats):
                out[g]["launches"] += 1
                out[g]["dram_bytes"] += b
    res = {"tcgen05_conv_dram_bytes_per_step": int(out["tcgen05_conv"]["dram_bytes"]), "launches": out["tcgen05_conv"]["launches"],
           "groups": {g: {"launches": v["launches"], "dram_bytes": int(v["dram_bytes"])} for g, v in out.items()},
           "source": f"{os.path.basename(path)}: dram__bytes_read.sum + dram__bytes_write.sum per launch, summed (ncu --set full of `{command}`)"}
    json.dump(res, open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1")
