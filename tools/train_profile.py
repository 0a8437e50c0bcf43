#!/usr/bin/env python
"""Where does a training step spend its GPU time?  torch.profiler over one Trainer.step on a reduced batch.
    python tools/train_profile.py [n_blocks]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
import __graft_entry__ as ge  # noqa: E402

n_blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 2
pkg = ge.build()
dev = torch.device("cuda:0")
torch.manual_seed(0)
net = importlib.import_module(pkg.__name__ + ".model").UnifiedModel(bench.model_config()).to(dev)
training = importlib.import_module(pkg.__name__ + ".training")
c, f = bench.training_batch(0, n_blocks=n_blocks)
dc, df = torch.from_numpy(c).to(dev), torch.from_numpy(f).to(dev)
q = torch.rand((n_blocks, 2)).to(dev)
tr = training.Trainer(net)
for _ in range(2):
    tr.step(dc, df, q)
torch.cuda.synchronize()
import time
t0 = time.perf_counter(); tr.step(dc, df, q); torch.cuda.synchronize(); print(f"step wall {1e3 * (time.perf_counter() - t0):.1f} ms, {len(c)} voxels")
so = pkg.sparse_ops
so.profiler = so.ConvProfiler()
tr.step(dc, df, q)
summ = so.profiler.summary()
so.profiler = None
for kind, d in sorted(summ.items(), key=lambda kv: -kv[1]["ms"]):
    print(f"{kind:24s} {d['launches']:4d} launches {d['ms']:9.3f} ms  {d['flops'] / max(d['ms'], 1e-9) / 1e9:8.1f} TFLOP/s")
    if kind in ("conv_bwd_weight", "conv_tc"):
        for shape, v in sorted(d["by_shape"].items(), key=lambda kv: -kv[1][1]):
            print(f"    {str(shape):32s} x{v[0]} {v[1]:8.3f} ms {v[2] / max(v[1], 1e-9) / 1e9:8.1f} TFLOP/s")
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    tr.step(dc, df, q)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=35, max_name_column_width=60))
