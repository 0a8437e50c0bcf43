"""The drop-in boundary, exercised by the reference's OWN host code: /root/reference/model/model.py,
model/transforms.py, model/entropy_models.py (with the `Sorted*` subclasses :28-126 and `_prune_tensor`'s int64-key `torch.isin`,
transforms.py:257-282), model/blocks.py and loss.py, imported UNMODIFIED on top of `install_shims()` -- the product's
`MinkowskiEngine` / `compressai` packages over libpcg.so -- and checked against the same golden fixtures as the mirror.
The tree comes from /root/reference here and from the git-ignored copy `__graft_entry__.build()` leaves under baseline/_ref/ on the
GPU box.
"""
import importlib

import numpy as np
import pytest
import torch

from tests.helpers import (align_rows, build_product_model, golden_config, load_golden, model_config, reference_dir,
                           reference_on_product, spread_weights)

needs_reference = pytest.mark.skipif(reference_dir() is None, reason="no reference tree (/root/reference or baseline/_ref/reference)")


def _reference_net(ref, cfg, sd, dev):
    net = ref.model.UnifiedModel(cfg)
    missing, unexpected = net.load_state_dict(dict(sd), strict=False)
    assert not [k for k in list(missing) + list(unexpected) if not k.split(".")[-1].startswith("_")], (missing, unexpected)
    net = net.to(dev).eval()
    net.update()
    return net


@needs_reference
def test_reference_model_builds_on_the_drop_in_packages(pkg):
    """CPU: the reference's UnifiedModel constructed from its own configs/CVPR_inverse_scaling.yaml over the product's packages has
    the mirror's state-dict keys, shapes and (same seed) values, sub-classes the shim's modules (Sorted*, MinkowskiGDN(GDN),
    UnifiedModel(CompressionModel)), and `update()` yields identical CDF tables."""
    import yaml
    mm = importlib.import_module(pkg.__name__ + ".model")
    ME = importlib.import_module(pkg.__name__ + ".MinkowskiEngine")
    cai = importlib.import_module(pkg.__name__ + ".compressai")
    with reference_on_product(pkg) as ref:
        cfg = yaml.safe_load(open(ref.dir + "/configs/CVPR_inverse_scaling.yaml"))["model"]
        torch.manual_seed(0)
        net = ref.model.UnifiedModel(cfg)
        net.update()
        assert isinstance(net, cai.models.base.CompressionModel)
        assert isinstance(net.entropy_model.h_s[0], ME.MinkowskiGenerativeConvolutionTranspose) and type(net.entropy_model.h_s[0]).__module__ == "model.entropy_models"
        assert isinstance(net.g_a.down_conv_1[1], cai.layers.GDN) and type(net.g_a.down_conv_1[1]).__module__ == "model.blocks"
        assert [n for n, _ in net.named_parameters() if n.endswith(".quantiles")] == ["entropy_model.entropy_bottleneck.quantiles"]   # train.py:64-65
        sd = net.state_dict()
    torch.manual_seed(0)
    mine = mm.UnifiedModel(model_config(width=128))
    mine.update()
    sd2 = mine.state_dict()
    assert set(sd) == set(sd2)
    assert sum(v.numel() for k, v in sd.items() if not k.split(".")[-1].startswith("_")) > 18_000_000
    for k in sd:
        assert sd[k].shape == sd2[k].shape and torch.equal(sd[k], sd2[k]), k


@needs_reference
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["f32", "auto"])
@pytest.mark.parametrize("name", ["ref_inverse_scaling", "ref_inverse_nn", "ref_fixed_R1"])
def test_reference_code_on_drop_in_matches_golden(pkg, dev, name, mode):
    """compress / decompress of the reference's own model.py (model/model.py:94-250) over the product packages against the
    fixture the same code produced over the CPU oracle: same gates as the mirror (tests/test_gpu_model.py)."""
    from tests.test_gpu_model import _golden_roundtrip
    prev = pkg.sparse_ops.set_conv_mode(mode)
    try:
        with reference_on_product(pkg) as ref:
            _golden_roundtrip(pkg, dev, name, strict=(mode == "f32"), make_net=lambda c, s, d: _reference_net(ref, c, s, d))
    finally:
        pkg.sparse_ops.set_conv_mode(prev)


@needs_reference
@pytest.mark.gpu
def test_reference_block_loop_on_drop_in_matches_golden(pkg, dev):
    """The reference's own block loop (model/model.py:121-176, :225-238) over the product packages on a cloud of 8 blocks, against
    the fixture the same code produced over the CPU oracle (tests/golden/ref_multiblock.npz)."""
    from tests.test_gpu_model import _multiblock_roundtrip
    with reference_on_product(pkg) as ref:
        _multiblock_roundtrip(pkg, dev, make_net=lambda c, s, d: _reference_net(ref, c, s, d))


@needs_reference
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["f32", "auto"])
def test_reference_code_on_drop_in_equals_mirror(pkg, dev, mode):
    """Same weights, same cloud: the reference's host code on the drop-in packages and the product's mirror produce the SAME
    bitstrings, latent coordinates and level sizes (the kernels are deterministic and row-order independent, so the mirror's
    different row orders / fusions on the encoder side change no bit), and decode to the same cloud (fp32 mode: identical
    geometry; bf16 mode: the mirror's fused occupancy head keeps its hidden rows in fp32 where two separate convolutions
    round them to bf16, so only the count is compared there)."""
    from oracle import synth
    mm = importlib.import_module(pkg.__name__ + ".model")
    cfg = model_config(width=32)
    torch.manual_seed(5)
    mine = mm.UnifiedModel(cfg)
    spread_weights(mine)
    sd = {k: v.clone() for k, v in mine.state_dict().items()}
    mine = mine.to(dev).eval()
    mine.update()
    pts, rgb = synth.small_cloud(seed=6, radius=12.0, centre=(22, 22, 22), n_samples=12000)
    pc = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1)).to(dev)
    q = torch.tensor([[0.5, 0.5]], device=dev)
    prev = pkg.sparse_ops.set_conv_mode(mode)
    try:
        with torch.no_grad(), reference_on_product(pkg) as ref:
            theirs = _reference_net(ref, cfg, sd, dev)
            a = mine.compress(pc.clone(), q, block_size=1024)
            b = theirs.compress(pc.clone(), q, block_size=1024)
            assert a[1] == b[1] and [[int(v) for v in lvl] for lvl in a[2][0]] == [[int(v) for v in lvl] for lvl in b[2][0]]
            ia, ib = align_rows(a[3][0].cpu().numpy(), b[3][0].cpu().numpy())
            assert len(ia) == a[3][0].shape[0] == b[3][0].shape[0]
            assert a[0][0][1][0] == b[0][0][1][0], "z strings differ"
            assert a[0][0][0][0] == b[0][0][0][0], "y strings differ"
            ra = mine.decompress(coordinates=a[3], strings=a[0], shape=a[1], k=a[2], q_vals=a[4]).cpu().numpy()
            rb = theirs.decompress(coordinates=b[3], strings=b[0], shape=b[1], k=b[2], q_vals=b[4]).cpu().numpy()
    finally:
        pkg.sparse_ops.set_conv_mode(prev)
    assert len(ra) == len(rb)
    if mode == "f32":
        ia, ib = align_rows(ra[:, :3], rb[:, :3])
        assert len(ia) >= 0.995 * len(ra), f"decoded geometry overlap {len(ia)}/{len(ra)}"
        assert np.abs(ra[ia, 3:] - rb[ib, 3:]).max() <= 1.0 / 255 + 1e-6


@needs_reference
@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["default", "shepard"])
def test_reference_forward_and_losses_on_drop_in_match_golden(pkg, dev, tag):
    """The reference's forward (model/model.py:45-90) and its unmodified loss.py (:9-274, with utils.overlapping_mask) over the
    product packages in fp32 mode against tests/golden/ref_losses.npz (the same code over the CPU oracle)."""
    import copy
    import yaml
    z, sd = load_golden("ref_losses")
    terms = ("Multiscale_FocalLoss", "ColorLoss", "bpp-y", "bpp-z")
    prev = pkg.sparse_ops.set_conv_mode("f32")
    try:
        with torch.no_grad(), reference_on_product(pkg) as ref:
            import MinkowskiEngine as ME          # the product's package under its upstream name
            net = _reference_net(ref, golden_config(z), sd, dev)
            x = ME.SparseTensor(features=torch.from_numpy(z["colors"]).to(dev), coordinates=torch.from_numpy(z["coords"]).to(dev), device=dev)
            out = net(x, torch.from_numpy(z["q"]).to(dev), torch.from_numpy(z["lam"]).to(dev))
            cfg_name = "CVPR_inverse_scaling_shepard" if tag == "shepard" else "CVPR_inverse_scaling"
            loss_cfg = yaml.safe_load(open(f"{ref.dir}/configs/{cfg_name}.yaml"))["loss"]
            total, parts = ref.loss.Loss(copy.deepcopy(loss_cfg))(x, out)
    finally:
        pkg.sparse_ops.set_conv_mode(prev)
    for k in terms:
        want = float(z[f"loss/{tag}/{k}"])
        assert abs(float(parts[k]) - want) <= 1e-4 * abs(want), (k, float(parts[k]), want)
