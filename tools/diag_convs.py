"""Per-convolution diagnostics of one encode+decode step (run on the GPU box): rows, pairs, mean active kernel
offsets per 128-row tile (the zero-padding factor of the output-stationary tiles) and CUDA-event time."""
import importlib, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
from bench import make_frame, model_config

pkg = ge.build()
so = pkg.sparse_ops
dev = torch.device("cuda:0")
torch.manual_seed(0)
net = importlib.import_module(pkg.__name__ + ".model").UnifiedModel(model_config()).to(dev).eval()
net.update()
q = torch.tensor([[0.5, 0.5]], device=dev)
pts, rgb = make_frame(sys.argv[1] if len(sys.argv) > 1 else "frame10", 0)
cloud = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1)).to(dev)

def step():
    enc = net.encode_block(cloud, q)
    ys = enc["y_symbols"]
    return net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: ys, enc["k"], q)

with torch.no_grad():
    for _ in range(2):
        step()
    so.profiler = so.ConvProfiler(detail=True)
    step()
    torch.cuda.synchronize()
    print(f"{'K x cin x cout':>16} {'T':>1} {'n_in':>9} {'n_out':>9} {'pairs/row':>9} {'act/tile':>8} {'ms':>7} {'TF/s':>7} {'MMA-eff':>7}")
    for L in so.profiler.launches:
        K, cin, cout = L["shape"]
        ms = L["e"][0].elapsed_time(L["e"][1])
        ppr = L["pairs"] / max(L["n_out"], 1)
        print(f"{K:>4}x{cin:>4}x{cout:>4}   {'T' if L['transposed'] else ' '} {L['n_in']:>9} {L['n_out']:>9} {ppr:>9.2f} {L['active_per_tile']:>8.1f} "
              f"{ms:>7.3f} {2e-9 * L['pairs'] * cin * cout / max(ms, 1e-6):>7.1f} {ppr / max(L['active_per_tile'], 1e-9):>7.2f}")
