#!/usr/bin/env python
"""Micro-benchmark of the sparse-convolution kernels on the bench frame's coordinate sets (run on the GPU box).
    python tools/bench_conv.py            # env: PCG_TC_STAGES, PCG_TC_SMEM_KB tune k_conv_tc's operand ring
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from bench import make_frame  # noqa: E402

pkg = ge.build()
so = pkg.sparse_ops
dev = torch.device("cuda:0")
pts, _ = make_frame("frame10", 0)
C = torch.from_numpy(np.concatenate([np.zeros((len(pts), 1), np.int32), pts], axis=1)).to(dev)
order = so.sort_rows(C, morton=True)
cs1 = so.CoordSet(C[order].contiguous(), 1)
cs2 = cs1.strided(2)
cs4 = cs2.strided(2)
cs8 = cs4.strided(2)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(name, in_set, out_set, cin, cout, ksize, reps=5):
    K = ksize ** 3
    f = torch.randn(in_set.n, cin, device=dev)
    w = torch.randn(K, cin, cout, device=dev) * 0.05
    pw = so.PackedWeights()
    fb = so.to_bf16(f)
    nbr, _ = in_set.kernel_map(out_set, ksize, False)
    pairs = int((nbr >= 0).sum())
    ts = []
    for i in range(reps + 2):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        so.sparse_conv(fb, in_set, out_set, w, None, ksize, False, packed=pw, out="bf16" if cout % 16 == 0 else "f32")
        e1.record()
        torch.cuda.synchronize()
        if i >= 2 or reps == 0:
            ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(f"{name:<34} n_out={out_set.n:>8} pairs/row={pairs / out_set.n:5.1f}  {ms:7.3f} ms  {2e-9 * pairs * cin * cout / ms:7.1f} TF/s "
          f"{pairs * 2 * so.cin_padded(cin) / ms / 1e6:7.1f} GB/s gathered")


so.set_conv_mode("tc")
print("PCG_TC_GATHER=%s PCG_TC_STAGES=%s PCG_TC_SMEM_KB=%s" % (os.environ.get("PCG_TC_GATHER"), os.environ.get("PCG_TC_STAGES"), os.environ.get("PCG_TC_SMEM_KB")))
HEADS = "--heads" in sys.argv          # occupancy-head layers on the generative candidates (cell-major rows)
QUICK = "--quick" in sys.argv          # two launches per shape, first two shapes only (for ncu captures)
with torch.no_grad():
    if HEADS:
        so.set_conv_mode("auto")
        print("PCG_TC_CA=%s PCG_TC_CARVEOUT=%s PCG_TC_NARROW_KB=%s PCG_NARROW=%s" % tuple(os.environ.get(k) for k in
              ("PCG_TC_CA", "PCG_TC_CARVEOUT", "PCG_TC_NARROW_KB", "PCG_NARROW")))
        cand1, cand2, cand4 = cs2.generated_cells(), cs4.generated_cells(), cs8.generated_cells()
        run("27x32x16   level-3 candidates", cand1, cand1, 32, 16, 3)
        run("27x16x1    level-3 candidates", cand1, cand1, 16, 1, 3)
        run("27x128x64  level-2 candidates", cand2, cand2, 128, 64, 3)
        run("27x64x1    level-2 candidates", cand2, cand2, 64, 1, 3)
        run("27x128x64  level-1 candidates", cand4, cand4, 128, 64, 3)
        run("27x64x1    level-1 candidates", cand4, cand4, 64, 1, 3)
        sys.exit(0)
    if QUICK:
        run("27x128x64  stride-1 set (818k)", cs1, cs1, 128, 64, 3, reps=0)
        run("27x192x256 stride-2 set (214k)", cs2, cs2, 192, 256, 3, reps=0)
        sys.exit(0)
    run("27x128x64  stride-1 set (818k)", cs1, cs1, 128, 64, 3)
    run("27x128x128 stride-1 set (818k)", cs1, cs1, 128, 128, 3)
    run("27x192x256 stride-2 set (214k)", cs2, cs2, 192, 256, 3)
    run("125x128x128 s2 -> s4 (strided)", cs2, cs4, 128, 128, 5)
    run("125x128x128 stride-4 set (54k)", cs4, cs4, 128, 128, 5)
    run("125x128x128 s4 -> s8 (strided)", cs4, cs8, 128, 128, 5)
    run("125x128x128 stride-8 set (14k)", cs8, cs8, 128, 128, 5)
    run("27x192x192 stride-8 set (14k)", cs8, cs8, 192, 192, 3)
    run("27x32x16   stride-1 set (818k)", cs1, cs1, 32, 16, 3)
    run("27x16x1    stride-1 set (818k)", cs1, cs1, 16, 1, 3)
    run("27x64x1    stride-1 set (818k)", cs1, cs1, 64, 1, 3)
