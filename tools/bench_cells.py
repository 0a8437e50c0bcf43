#!/usr/bin/env python
"""Micro-benchmark of the octree-cell candidate construction on the bench frame's stride-2 set (run on the GPU box):
parent lists by 27 probes per cell (pcg_kernel_map) against the scatter form (pcg_cells_parent_map)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import __graft_entry__ as ge
from bench import make_frame
pkg = ge.build(); so = pkg.sparse_ops; dev = torch.device("cuda:0")
from importlib import import_module
lib = pkg._lib.load(); call, ptr, stream = pkg._lib.call, pkg._lib.ptr, pkg._lib.stream
pts, _ = make_frame("frame10", 0)
C = torch.from_numpy(np.concatenate([np.zeros((len(pts), 1), np.int32), pts], axis=1)).to(dev)
cs1 = so.CoordSet(C[so.sort_rows(C, morton=True)].contiguous(), 1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def t(fn, n=5):
    ts = []
    for i in range(n + 2):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts[2:])[n // 2]
for name, par in (("stride-2 parents (214k)", cs1.strided(2)), ("stride-4 parents (54k)", cs1.strided(2).strided(2))):
    cells = par.generated(3, 1, guess_per_row=14)
    ckeys, cvals, ccap = cells.table(); par.table()
    pitch = lib.pcg_nbr_pitch(cells.n)
    pn = torch.empty((27, pitch), dtype=torch.int32, device=dev); st = torch.zeros(1, dtype=torch.int32, device=dev)
    def probe():
        par._kmaps.clear(); par.kernel_map(cells, 3, False)
    def scatter():
        call("pcg_cells_parent_map", ptr(par.coords), par.n, par.ts, ptr(ckeys), ptr(cvals), ccap, cells.n, ptr(pn), pitch, ptr(st), stream())
    a = par.kernel_map(cells, 3, False)[0]; scatter(); torch.cuda.synchronize()
    print(f"{name}: cells={cells.n}  probe-per-cell {t(probe):.3f} ms   scatter {t(scatter):.3f} ms   equal={bool(torch.equal(a, pn))}")
    def whole():
        par._generated.clear(); par._kmaps.clear(); par.generated_cells()
    print(f"   generated_cells() end to end: {t(whole):.3f} ms")

# GEMM-first transposed convolution on the same parents: per-kernel times from the conv profiler (PCG_REDUCE_VPL=1|2 for A/B)
print("PCG_REDUCE_VPL=%s PCG_REDUCE_BLOCKS=%s" % (os.environ.get("PCG_REDUCE_VPL"), os.environ.get("PCG_REDUCE_BLOCKS")))
for name, par, cin, cout in (("up_3: stride-2 parents, 128 -> 32", cs1.strided(2), 128, 32),
                             ("up_2: stride-4 parents, 128 -> 128", cs1.strided(2).strided(2), 128, 128),
                             ("up_1: stride-8 parents, 128 -> 128", cs1.strided(2).strided(2).strided(2), 128, 128)):
    out_set = par.generated_cells()
    f = so.to_bf16(torch.randn(par.n, cin, device=dev))
    w = torch.randn(125, cin, cout, device=dev) * 0.05
    pw = so.PackedGemmWeights()
    so.profiler = so.ConvProfiler()
    with torch.no_grad():
        for i in range(7):
            flush.fill_(1)
            so.gen_conv_transpose_cells(f, par, out_set, w, None, packed=pw, out="both")
    recs = so.profiler.records
    torch.cuda.synchronize()
    so.profiler = None
    by = {}
    for kind, shape, flops, bytes_, e0, e1 in recs[2 * (len(recs) // 7):]:
        by.setdefault(kind, []).append((e0.elapsed_time(e1), bytes_))
    msg = "  ".join(f"{k} {sorted(t for t, _ in v)[len(v) // 2]:.3f} ms ({v[0][1] / sorted(t for t, _ in v)[len(v) // 2] / 1e6:.0f} GB/s)" for k, v in by.items())
    print(f"{name}: children={out_set.n}  {msg}")
