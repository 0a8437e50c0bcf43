"""oracle/data_ref.py pinned against the dependency the reference wraps (torchvision ColorJitter) and basic invariants."""
import numpy as np
import torch

from oracle import data_ref, synth


def test_color_jitter_restatement_equals_torchvision_forward():
    import torchvision
    tv = torchvision.transforms.ColorJitter(brightness=0.3, contrast=0.3, saturation=0.3, hue=0.3)
    _, rgb = synth.small_cloud(seed=1, radius=8.0, centre=(12, 12, 12), n_samples=3000)
    rgb = torch.from_numpy(rgb.astype(np.float32))
    for seed in range(5):
        torch.manual_seed(seed)
        fn_idx, b, c, s, h = tv.get_params(tv.brightness, tv.contrast, tv.saturation, tv.hue)
        torch.manual_seed(seed)
        ref = tv(rgb.T.unsqueeze(-1)).squeeze(-1).T
        assert torch.equal(data_ref.color_jitter(rgb, fn_idx.tolist(), [b, c, s, h]), ref)


def test_slicing_and_rotation_invariants():
    pts, rgb = synth.small_cloud(seed=2, radius=30.0, centre=(50, 40, 70), n_samples=20000)
    pts, rgb = torch.from_numpy(pts.astype(np.float32)), torch.from_numpy(rgb.astype(np.float32))
    cubes = data_ref.slice_into_cubes(pts, rgb, 32)
    assert sum(int(c["num_points"]) for c in cubes) == len(pts)
    for c in cubes:
        assert int(c["points"].min()) >= 0 and int(c["points"].max()) < 32
    rot = data_ref.rotation_matrix_3d(torch.tensor([0.3]), torch.tensor([1.1]))
    assert torch.allclose(rot @ rot.T, torch.eye(3), atol=1e-6)
    q, f, _ = data_ref.random_rotate(pts, rgb, torch.eye(3), 128)
    assert len(q) == len(pts) and torch.equal(q.float(), pts)           # identity rotation on voxel centres keeps every point
