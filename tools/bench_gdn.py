import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.build(); so = pkg.sparse_ops; dev = torch.device("cuda:0")
def t(fn, n=5):
    ts = []
    for i in range(n + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts[2:])[n // 2]
for n, c in ((214051, 128), (54352, 128), (13664, 128), (214051, 120)):
    x = torch.randn(n, c, device=dev); beta = torch.rand(c, device=dev) + .5; gamma = torch.rand(c, c, device=dev) * .1
    ms = t(lambda: so.gdn(x, beta, gamma, False))
    print(f"gdn n={n} c={c}: {ms:.3f} ms  {8e-6 * n * c / ms:.0f} GB/s  {2e-9 * n * c * c / ms:.1f} TF/s")
