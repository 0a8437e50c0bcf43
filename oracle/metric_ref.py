"""CPU restatement of the reference's PSNR metric.  TEST INFRASTRUCTURE (the parity gate: D1-PSNR / Y-PSNR).

Follows /root/reference/metrics/metric.py:6-189 (`PointCloudMetric`): 2-NN association in both directions
(:36-43, Open3D KDTreeFlann there, scipy cKDTree here), point-to-point D1 error with the squared distance taken
as the MEAN over x,y,z (:113) and peak = resolution (:118), colours of equidistant neighbours averaged with the
first neighbour counted twice (:127-146), 8-bit rounding then BT.709 YUV through a truncating uint8 cast
(:148-149,171-189), PSNR = 10 log10(1 / mse) per channel (:161-165), symmetric = min of both directions (:72-84).
Open3D's `remove_duplicated_points` (:19-21) is a first-occurrence dedupe of exact duplicates.
"""
import numpy as np
from scipy.spatial import cKDTree


def _dedupe(points, colors):
    _, idx = np.unique(points, axis=0, return_index=True)
    idx = np.sort(idx)
    return points[idx], colors[idx]


def rgb_to_yuv(rgb):
    """metric.py:171-189 (BT.709, with the reference's truncating uint8 cast when the input is in [0,1])."""
    rgb = np.asarray(rgb, dtype=np.float64)
    scale = rgb.max() <= 1.0
    if scale:
        rgb = (rgb * 255).astype(np.uint8)
    yuv = np.empty(rgb.shape, dtype=np.float32)
    r, g, b = (rgb[..., i].astype(np.float64) for i in range(3))
    yuv[..., 0] = 0.2126 * r + 0.7152 * g + 0.0722 * b
    yuv[..., 1] = -0.1146 * r - 0.3854 * g + 0.5 * b
    yuv[..., 2] = 0.5 * r - 0.4542 * g - 0.0458 * b
    if scale:
        yuv = yuv / 255.0
        yuv[..., 1] += 0.5
        yuv[..., 2] += 0.5
    return yuv


def _direction(a_pts, a_col, b_pts, b_col, tree_b, resolution, prefix):
    k = min(2, len(b_pts))
    _, nn = tree_b.query(a_pts, k=k)
    nn = nn.reshape(len(a_pts), k)
    b0 = b_pts[nn[:, 0]]
    cols = b_col[nn[:, 0]].astype(np.float64).copy()
    d0 = ((a_pts - b0) ** 2).mean(axis=1)
    res = {prefix + "mse": d0.mean(), prefix + "hausdorff": d0.max()}
    with np.errstate(divide="ignore"):
        res[prefix + "psnr_mse"] = 10 * np.log10(resolution ** 2 / res[prefix + "mse"])
        res[prefix + "psnr_hausdorff"] = 10 * np.log10(resolution ** 2 / res[prefix + "hausdorff"])
    if k == 2:
        d1 = ((a_pts - b_pts[nn[:, 1]]) ** 2).mean(axis=1)
        for i in np.nonzero(np.abs(d0 - d1) < 1e-8)[0]:
            all_nns = nn[i]
            for max_search in range(5, 30, 5):
                _, all_nns = tree_b.query(a_pts[i], k=min(max_search, len(b_pts)))
                all_nns = np.atleast_1d(all_nns)
                dist = ((a_pts[i] - b_pts[all_nns]) ** 2).mean(axis=1)
                if (dist < 1e-8).sum() < max_search:
                    break
            same = [j for j in all_nns if ((a_pts[i] - b_pts[j]) ** 2).mean() == d0[i]]
            for j in same:
                cols[i] += b_col[j]
            cols[i] = cols[i] / (len(same) + 1)
    a_yuv = rgb_to_yuv(np.clip(np.round(a_col * 255.0) / 255.0, 0.0, 1.0))
    b_yuv = rgb_to_yuv(np.clip(np.round(cols * 255.0) / 255.0, 0.0, 1.0))
    mse = ((a_yuv - b_yuv) ** 2).mean(axis=0)
    with np.errstate(divide="ignore"):
        for c, name in enumerate("yuv"):
            res[prefix + name + "_mse"] = mse[c]
            res[prefix + name + "_psnr"] = 10 * np.log10(1 / mse[c])
        res[prefix + "yuv_mse"] = mse.mean()
        res[prefix + "yuv_psnr"] = 10 * np.log10(1 / mse.mean())
    return res


def pointcloud_metrics(source, reconstruction, resolution=1023, drop_duplicates=True):
    """source / reconstruction: (N,6) arrays [xyz | rgb in 0..1].  Returns the reference's result dict
    (metric.py:61-86): AB_* (source -> reconstruction), BA_*, sym_* = min over the two directions."""
    sp, sc = np.asarray(source[:, :3], np.float64), np.asarray(source[:, 3:6], np.float64)
    rp, rc = np.asarray(reconstruction[:, :3], np.float64), np.asarray(reconstruction[:, 3:6], np.float64)
    if drop_duplicates:
        sp, sc = _dedupe(sp, sc)
        rp, rc = _dedupe(rp, rc)
    ts, tr = cKDTree(sp), cKDTree(rp)
    res = {}
    res.update(_direction(sp, sc, rp, rc, tr, resolution, "AB_"))
    res.update(_direction(rp, rc, sp, sc, ts, resolution, "BA_"))
    for key in ("mse", "hausdorff", "psnr_mse", "psnr_hausdorff", "y_mse", "u_mse", "v_mse", "y_psnr", "u_psnr", "v_psnr"):
        res["sym_" + key] = min(res["AB_" + key], res["BA_" + key])
    return res
