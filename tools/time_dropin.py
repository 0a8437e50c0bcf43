#!/usr/bin/env python
"""Time the drop-in path: the reference's UNMODIFIED model/model.py (compress + decompress, component form, host range coder
included) over `install_shims()` on the 10-bit frame, beside the product's mirror doing the same call.
    python tools/time_dropin.py [frame10|block128] > gpurun_out/r02_dropin_timing.json"""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import __graft_entry__ as ge  # noqa: E402
from tests.helpers import reference_on_product  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "frame10"
pkg = ge.build()
dev = torch.device("cuda:0")
cfg = bench.model_config()
pts, rgb = bench.make_frame(workload, 0)
cloud = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1)).to(dev)
q = torch.tensor([[0.5, 0.5]], device=dev)
block = 1024


def timed(net, reps=3):
    out = []
    with torch.no_grad():
        for i in range(reps + 2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            bits, shapes, ks, coords, qv = net.compress(cloud.clone(), q, block_size=block)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            rec = net.decompress(coordinates=coords, strings=bits, shape=shapes, k=ks, q_vals=qv)
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            if i >= 2:
                out.append((1e3 * (t1 - t0), 1e3 * (t2 - t1)))
    nbytes = sum(len(s[0]) for blk in bits for s in blk)
    return {"compress_ms": sum(a for a, _ in out) / len(out), "decompress_ms": sum(b for _, b in out) / len(out), "bytes": nbytes,
            "decoded_voxels": int(rec.shape[0])}


torch.manual_seed(0)
mine = importlib.import_module(pkg.__name__ + ".model").UnifiedModel(cfg).to(dev).eval()
mine.update()
res = {"workload": workload, "voxels": int(len(pts)), "note": "wall clock around compress() / decompress() in the component form (host rANS coder and "
       "its device<->host copies INCLUDED, unlike bench.py's metric), one frame at a time, random-init weights, q=(0.5,0.5)"}
res["mirror"] = timed(mine)
with reference_on_product(pkg) as ref:
    torch.manual_seed(0)
    theirs = ref.model.UnifiedModel(cfg).to(dev).eval()
    theirs.update()
    res["reference_code_on_drop_in"] = timed(theirs)
for k in ("mirror", "reference_code_on_drop_in"):
    r = res[k]
    r["Mpoints_per_s"] = len(pts) / ((r["compress_ms"] + r["decompress_ms"]) * 1e3)
print(json.dumps(res))
