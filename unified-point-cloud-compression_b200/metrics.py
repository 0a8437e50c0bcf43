"""PSNR metric on the GPU -- mirror of /root/reference/metrics/metric.py (`PointCloudMetric`, :6-189) for voxel clouds.

The reference associates the two clouds with per-point KD-tree queries in Python (:36-43); here each direction is one
launch of `pcg_nn_search` (hash-probe shell search, csrc/pcg_metric.cu), which also evaluates the per-point colour error
terms (8-bit rounding, BT.709 YUV, equidistant-neighbour colour averaging, :127-166).  Only the final means / maxima /
log10 are taken with torch reductions in float64.  Same result keys as the reference (`AB_*`, `BA_*`, `sym_*`).
Colours are carried in fp32 (the codec's own precision; k/255 inputs round to the same 8-bit value as in fp64).
"""
import math

import torch

from . import sparse_ops as so
from ._lib import call, ptr, stream


def _voxel_cloud(cloud):
    """(N,6) [xyz|rgb] CUDA tensor -> (unique int32 coords (M,4), colours (M,3) fp32), first occurrence kept
    (metric.py:19-21 `remove_duplicated_points`)."""
    if not cloud.is_cuda:
        raise RuntimeError("point clouds must live on a CUDA device: the graft has no CPU path")
    xyz = cloud[:, :3]
    coords = torch.cat([torch.zeros((cloud.shape[0], 1), dtype=torch.int32, device=cloud.device),
                        torch.floor(xyz).to(torch.int32)], dim=1).contiguous()
    if not torch.equal(coords[:, 1:].to(xyz.dtype), xyz):
        raise RuntimeError("pointcloud_metrics: positions must lie on the integer voxel grid")
    uc, index, table = so.unique_first(coords)
    colors = cloud[:, 3:6].to(torch.float32)[index.long()].contiguous()
    return so.CoordSet(uc.contiguous(), 1, table=table), colors


def _direction(a_set, a_col, b_set, b_col, resolution, prefix, max_radius):
    na = a_set.n
    dev = a_set.coords.device
    keys, vals, cap = b_set.table()
    nn = torch.empty(max(na, 1), dtype=torch.int32, device=dev)
    d2 = torch.empty(max(na, 1), dtype=torch.int32, device=dev)
    err = torch.empty((max(na, 1), 3), dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    call("pcg_nn_search", ptr(keys), ptr(vals), cap, ptr(b_col), ptr(a_set.coords), ptr(a_col), na, int(max_radius),
         ptr(nn), ptr(d2), ptr(err), ptr(status), stream())
    if int(status.item()) & 1:
        raise RuntimeError(f"pointcloud_metrics: a point has no neighbour within {max_radius} voxels")
    dist = d2[:na].to(torch.float64) / 3.0                      # squared distance as the MEAN over x, y, z (metric.py:113)
    mse, hausdorff = float(dist.mean()), float(dist.max())
    cmse = err[:na].to(torch.float64).mean(dim=0).tolist()
    res = {prefix + "mse": mse, prefix + "hausdorff": hausdorff}
    log10 = lambda v: math.inf if v == 0 else 10 * math.log10(v)
    res[prefix + "psnr_mse"] = math.inf if mse == 0 else 10 * math.log10(resolution ** 2 / mse)
    res[prefix + "psnr_hausdorff"] = math.inf if hausdorff == 0 else 10 * math.log10(resolution ** 2 / hausdorff)
    for c, name in enumerate("yuv"):
        res[prefix + name + "_mse"] = cmse[c]
        res[prefix + name + "_psnr"] = math.inf if cmse[c] == 0 else log10(1 / cmse[c])
    m = sum(cmse) / 3
    res[prefix + "yuv_mse"] = m
    res[prefix + "yuv_psnr"] = math.inf if m == 0 else log10(1 / m)
    return res, nn[:na]


def pointcloud_metrics(source, reconstruction, resolution=1023, max_radius=64):
    """source / reconstruction: (N,6) CUDA tensors [xyz on the voxel grid | rgb in 0..1].  Returns the reference's
    result dict (metric.py:61-86): AB_* (source -> reconstruction), BA_*, sym_* = min over the two directions."""
    a_set, a_col = _voxel_cloud(source)
    b_set, b_col = _voxel_cloud(reconstruction)
    res = {}
    ab, _ = _direction(a_set, a_col, b_set, b_col, resolution, "AB_", max_radius)
    ba, _ = _direction(b_set, b_col, a_set, a_col, resolution, "BA_", max_radius)
    res.update(ab)
    res.update(ba)
    for key in ("mse", "hausdorff", "psnr_mse", "psnr_hausdorff", "y_mse", "u_mse", "v_mse", "y_psnr", "u_psnr", "v_psnr"):
        res["sym_" + key] = min(res["AB_" + key], res["BA_" + key])
    return res


def nearest_neighbours(query_set, target_set, max_radius=64):
    """(row of the nearest target voxel, integer squared distance) for every row of `query_set` (both CoordSets)."""
    dev = query_set.coords.device
    keys, vals, cap = target_set.table()
    nn = torch.empty(max(query_set.n, 1), dtype=torch.int32, device=dev)
    d2 = torch.empty(max(query_set.n, 1), dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    call("pcg_nn_search", ptr(keys), ptr(vals), cap, None, ptr(query_set.coords), None, query_set.n, int(max_radius),
         ptr(nn), ptr(d2), None, ptr(status), stream())
    return nn[:query_set.n], d2[:query_set.n]
