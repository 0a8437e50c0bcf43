"""GPU data path (cube slicing, RandomRotate + de-duplication, ColorJitter) against oracle/data_ref.py, which
restates data/dataloader.py:167-208 and data/transform.py:33-124 and calls torchvision (the reference's own dependency) for
the colour ops.  Integer results bit-exact; colours within 2e-5 (fp32 mul/add contraction); rotated voxels identical except
for positions within float rounding of a voxel face (bounded at 0.1 % of the points)."""
import importlib
import itertools

import numpy as np
import pytest
import torch

from oracle import data_ref, synth

pytestmark = pytest.mark.gpu


def data_mod(pkg):
    return importlib.import_module(pkg.__name__ + ".data")


def cloud(seed, n=30000, radius=40.0, centre=(70, 64, 90)):
    pts, rgb = synth.small_cloud(seed=seed, radius=radius, centre=centre, n_samples=n)
    return torch.from_numpy(pts.astype(np.float32)), torch.from_numpy(rgb.astype(np.float32))


@pytest.mark.parametrize("cube_size", [64, 32])
def test_slice_into_cubes_matches_reference_order_and_rows(pkg, dev, cube_size):
    pts, rgb = cloud(1)
    perm = torch.randperm(len(pts), generator=torch.Generator().manual_seed(0))
    pts, rgb = pts[perm], rgb[perm]                                   # unsorted input: row order inside a cube must be the input's
    ref = data_ref.slice_into_cubes(pts, rgb, cube_size)
    got = data_mod(pkg).slice_into_cubes(pts.to(dev), rgb.to(dev), cube_size)
    assert len(got) == len(ref) > 1
    for g, r in zip(got, ref):
        assert torch.equal(g["offset"].cpu(), r["offset"]) and int(g["num_points"]) == int(r["num_points"])
        assert torch.equal(g["points"].cpu(), r["points"]) and torch.equal(g["colors"].cpu(), r["colors"])
    assert data_mod(pkg).slice_into_cubes(pts[:0].to(dev), rgb[:0].to(dev), cube_size) == []


def test_color_jitter_matches_torchvision_for_every_op_order(pkg, dev):
    _, rgb = cloud(2, n=8000)
    rgb[:50] = rgb[:50, :1]                                            # grey points: the hue op's degenerate branch
    rgb[50:60] = torch.tensor([0.0, 0.0, 0.0]); rgb[60:70] = torch.tensor([1.0, 1.0, 1.0])
    cj = data_mod(pkg).ColorJitter
    rng = np.random.default_rng(0)
    for order in itertools.permutations(range(4)):
        factors = [float(rng.uniform(0.7, 1.3)), float(rng.uniform(0.7, 1.3)), float(rng.uniform(0.7, 1.3)), float(rng.uniform(-0.3, 0.3))]
        ref = data_ref.color_jitter(rgb, list(order), factors)
        got = cj.apply(rgb.to(dev), list(order), factors).cpu()
        assert torch.allclose(got, ref, atol=2e-5, rtol=0), (order, float((got - ref).abs().max()))


def test_color_jitter_draws_the_reference_parameters_and_transforms_samples(pkg, dev):
    import torchvision
    dm = data_mod(pkg)
    tv = torchvision.transforms.ColorJitter(brightness=0.3, contrast=0.3, saturation=0.3, hue=0.3)
    torch.manual_seed(11)
    fn_idx, b, c, s, h = tv.get_params(tv.brightness, tv.contrast, tv.saturation, tv.hue)
    torch.manual_seed(11)
    order, factors = dm.ColorJitter().get_params()
    assert order == fn_idx.tolist() and factors == [b, c, s, h]
    _, rgb = cloud(3, n=4000)
    torch.manual_seed(5)
    ref = tv(rgb.T.unsqueeze(-1)).squeeze(-1).T                        # data/transform.py:50-53
    torch.manual_seed(5)
    sample = dm.ColorJitter()({"points": torch.zeros(len(rgb), 3, device=dev), "colors": rgb.to(dev)})
    assert torch.allclose(sample["colors"].cpu(), ref, atol=2e-5, rtol=0)
    torch.manual_seed(5)
    cubes = dm.ColorJitter()({"cubes": [{"colors": rgb.to(dev)}, {"colors": rgb[:100].to(dev)}]})["cubes"]
    assert torch.allclose(cubes[0]["colors"].cpu(), ref, atol=2e-5, rtol=0) and cubes[1]["colors"].shape == (100, 3)


@pytest.mark.parametrize("crop", [False, True])
def test_random_rotate_matches_reference(pkg, dev, crop):
    dm = data_mod(pkg)
    pts, rgb = cloud(4, n=40000, radius=50.0, centre=(64, 64, 64))
    block = 128
    torch.manual_seed(21)
    phi = torch.rand(1) * 2 * 3.141592653589793
    theta = torch.rand(1) * 2 * 3.141592653589793
    rot = data_ref.rotation_matrix_3d(phi, theta)
    assert torch.equal(dm.RandomRotate.rotation_matrix_3d(phi, theta), rot)
    ref_pts, ref_rgb, rotated = data_ref.random_rotate(pts, rgb, rot, block, crop=crop)
    torch.manual_seed(21)
    out = dm.RandomRotate(block, crop=crop)({"points": pts.to(dev), "colors": rgb.to(dev)})
    got_pts, got_rgb = out["points"].cpu(), out["colors"].cpu()
    assert got_pts.dtype == torch.int32 and got_pts.shape[1] == 3 and got_rgb.shape == (len(got_pts), 3)
    assert len(torch.unique(got_pts, dim=0)) == len(got_pts)          # de-duplicated
    key = lambda P: {tuple(r) for r in np.asarray(P).astype(np.int64).tolist()}
    a, b = key(got_pts), key(ref_pts)
    assert len(a ^ b) <= 1e-3 * len(b), (len(a ^ b), len(b))
    # colours of the shared voxels: the first point of the voxel in input order
    ref_map = {tuple(p): c for p, c in zip(ref_pts.tolist(), ref_rgb.tolist())}
    bad = sum(1 for p, c in zip(got_pts.tolist(), got_rgb.tolist()) if tuple(p) in ref_map and ref_map[tuple(p)] != c)
    assert bad <= 2e-3 * len(b), bad
    if crop:
        assert int(got_pts.min()) >= 0 and int(got_pts.max()) < block


def test_cpu_samples_are_rejected(pkg):
    dm = data_mod(pkg)
    pts, rgb = cloud(5, n=100)
    with pytest.raises(RuntimeError, match="no CPU path"):
        dm.ColorJitter()({"points": pts, "colors": rgb})
    with pytest.raises(RuntimeError, match="no CPU path"):
        dm.RandomRotate(64)({"points": pts, "colors": rgb})
    with pytest.raises(RuntimeError, match="no CPU path"):
        dm.slice_into_cubes(pts, rgb)
    assert [type(t).__name__ for t in dm.build_transforms({"2_RandomRotate": {"key": "RandomRotate", "block_size": 128},
                                                           "1_ColorJitter": {"key": "ColorJitter"}})] == ["ColorJitter", "RandomRotate"]
