"""The PSNR parity gate (oracle/metric_ref.py = /root/reference/metrics/metric.py restated) against brute force."""
import numpy as np

from oracle import metric_ref, synth


def test_d1_and_y_psnr_match_brute_force():
    pts, rgb = synth.small_cloud(seed=2, radius=7.0, centre=(10, 10, 10), n_samples=2500)
    src = np.concatenate([pts.astype(np.float64), rgb], axis=1)
    rng = np.random.default_rng(3)
    rec = src[rng.random(len(src)) < 0.9].copy()
    rec[:, :3] += rng.integers(-1, 2, size=(len(rec), 3)) * (rng.random((len(rec), 1)) < 0.3)
    rec[:, 3:] = np.clip(rec[:, 3:] + rng.normal(0, 0.03, (len(rec), 3)), 0, 1)
    rec = rec[np.unique(rec[:, :3], axis=0, return_index=True)[1]]
    res = metric_ref.pointcloud_metrics(src, rec, resolution=31)
    # brute force D1 (mean over xyz of the squared error to the nearest neighbour), both directions
    d = ((src[:, None, :3] - rec[None, :, :3]) ** 2).mean(axis=2)
    ab, ba = d.min(axis=1).mean(), d.min(axis=0).mean()
    assert np.isclose(res["AB_mse"], ab) and np.isclose(res["BA_mse"], ba)
    assert np.isclose(res["sym_psnr_mse"], min(10 * np.log10(31 ** 2 / ab), 10 * np.log10(31 ** 2 / ba)))
    # identical clouds: zero error, infinite PSNR
    same = metric_ref.pointcloud_metrics(src, src)
    assert same["sym_mse"] == 0 and np.isinf(same["sym_y_psnr"])


def test_yuv_conversion_constants():
    yuv = metric_ref.rgb_to_yuv(np.array([[1.0, 1.0, 1.0], [0.0, 0.0, 0.0], [1.0, 0.0, 0.0]]))
    assert np.allclose(yuv[0], [1.0, 0.5, 0.5], atol=1e-6) and np.allclose(yuv[1], [0.0, 0.5, 0.5], atol=1e-6)
    assert np.allclose(yuv[2], [0.2126, 0.5 - 0.1146, 1.0], atol=1e-6)


def test_unique_neighbour_colours_are_not_averaged():
    src = np.array([[0, 0, 0, 0.2, 0.4, 0.6], [4, 0, 0, 0.8, 0.1, 0.3]], dtype=np.float64)
    rec = np.array([[0, 0, 1, 0.2, 0.4, 0.6], [4, 1, 0, 0.8, 0.1, 0.3], [9, 9, 9, 0.0, 0.0, 0.0]], dtype=np.float64)
    res = metric_ref.pointcloud_metrics(src, rec, resolution=15)
    assert res["AB_y_mse"] == 0.0 and np.isclose(res["AB_mse"], 1.0 / 3.0)
