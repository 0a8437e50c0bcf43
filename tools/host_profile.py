#!/usr/bin/env python
"""Host-side profile of one resident bench step: where the Python / dispatch time goes (cProfile + torch op counts).
    python tools/host_profile.py            (on the GPU box)"""
import cProfile
import importlib
import os
import pstats
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import __graft_entry__ as ge  # noqa: E402

pkg = ge.build()
dev = torch.device("cuda:0")
model_mod = importlib.import_module(pkg.__name__ + ".model")
torch.manual_seed(0)
net = model_mod.UnifiedModel(bench.model_config()).to(dev).eval()
net.update()
q = torch.tensor([[0.5, 0.5]], device=dev)
pts, rgb = bench.make_frame(sys.argv[1] if len(sys.argv) > 1 else "frame10", 0)
cloud = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1)).to(dev)


def step():
    enc = net.encode_block(cloud, q)
    ys = enc["y_symbols"]
    x_hat = net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: ys, enc["k"], q)
    return net.finalize(x_hat.C, x_hat.F)


with torch.no_grad():
    for _ in range(4):
        step()
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    step()
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr, stream=sys.stdout)
    st.sort_stats("tottime").print_stats(45)
    st.sort_stats("cumulative").print_stats(60)
