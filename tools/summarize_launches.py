#!/usr/bin/env python
"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) of `bench.py --steps 1`.

The timed resident step is the launch range between the first 512 MiB L2-flush fill (a
`FillFunctor<unsigned char>` launch with a grid of > 10^5 blocks) and the marker fill bench.py issues right after
the step (`FillFunctor<unsigned char>` with grid (1,1,1) and value 2; any 1-block uchar fill following the flush).
Writes a markdown table: ms, share, launches per kernel name.

    python tools/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.md
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"\(.*$", "", name)            # drop the argument list
    name = re.sub(r"<.*$", "", name)             # drop template arguments
    name = name.replace("void ", "").replace("at::native::", "native::").strip()
    return name


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        grid = int(r["Grid Size"].strip("()").split(",")[0])
        rows.append((r["Kernel Name"], grid, float(r["Metric Value"])))
    is_fill = lambda n: "FillFunctor<unsigned char>" in n
    start = next(i for i, (n, g, _) in enumerate(rows) if is_fill(n) and g > 100000)
    end = next((i for i in range(start + 1, len(rows)) if is_fill(rows[i][0]) and rows[i][1] == 1), None)
    nxt = next((i for i in range(start + 1, len(rows)) if is_fill(rows[i][0]) and rows[i][1] > 100000), len(rows))
    if end is None or end > nxt:       # list from a bench without the marker: resident step + one e2e step precede the next flush
        end = start + 1 + (nxt - start - 1) // 2
    step = rows[start + 1:end]
    agg = OrderedDict()
    for n, g, ns in step:
        a = agg.setdefault(short(n), [0, 0.0])
        a[0] += 1
        a[1] += ns
    total = sum(v[1] for v in agg.values())
    print(f"Timed step: {len(step)} launches, {total / 1e6:.2f} ms of kernel time (ncu, serialised, cold caches).\n")
    print("| ms | share | launches | kernel |")
    print("|---:|---:|---:|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {v[1] / 1e6:.3f} | {100 * v[1] / total:.1f}% | {v[0]} | `{k}` |")


if __name__ == "__main__":
    main(sys.argv[1])
