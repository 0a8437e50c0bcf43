"""Loss-value parity: the product's loss mirror (model/loss.py over libpcg.so) against the values the
reference's UNMODIFIED loss.py (/root/reference/loss.py:9-274) produced on the training forward of its own model
(tests/golden/ref_losses.npz, generator oracle/make_golden.py::make_losses).  Two gates:
  * identical inputs  -- the fixture's complete forward output fed to the product's losses: every term within 1e-5;
  * whole chain       -- product forward (fp32 convolution mode) + product losses on the fixture's weights and input: 1e-4.
"""
import importlib
import os

import numpy as np
import pytest
import torch

from tests.helpers import GOLDEN, build_product_model, golden_config, load_golden

TERMS = ("Multiscale_FocalLoss", "ColorLoss", "bpp-y", "bpp-z")


def _loss_cfg(loss_mod, tag):
    return loss_mod.default_loss_config(shepard=tag == "shepard")


def _rel(a, b):
    return abs(float(a) - float(b)) / max(abs(float(b)), 1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["default", "shepard"])
def test_loss_terms_match_reference_on_identical_inputs(pkg, dev, tag):
    ME = importlib.import_module(pkg.__name__ + ".MinkowskiEngine")
    loss_mod = importlib.import_module(pkg.__name__ + ".model.loss")
    z, _ = load_golden("ref_losses")
    t = lambda a: torch.from_numpy(np.asarray(a)).to(dev)
    sparse = lambda C, F, ts=1: ME.SparseTensor(features=t(F), coordinates=t(C), tensor_stride=int(ts), device=dev)
    gt = sparse(z["coords"], z["colors"])
    ones = lambda C: np.ones((len(C), 1), np.float32)
    pred = {"prediction": sparse(z["pred_C"], z["pred_F"]),
            "occ_predictions": [sparse(z[f"occ{i}_C"], z[f"occ{i}_F"], z[f"occ{i}_ts"]) for i in range(3)],
            "points": [sparse(z[f"points{i}_C"], ones(z[f"points{i}_C"]), z[f"points{i}_ts"]) for i in range(3)],
            "q_map": t(z["lam"]), "likelihoods": {"y": t(z["lik_y"]), "z": t(z["lik_z"])}}
    assert all((p.C.cpu().numpy() == z[f"occ{i}_C"]).all() for i, p in enumerate(pred["occ_predictions"]))     # row order kept
    with torch.no_grad():
        total, parts = loss_mod.Loss(_loss_cfg(loss_mod, tag))(gt, pred)
    for k in TERMS:
        assert _rel(parts[k], z[f"loss/{tag}/{k}"]) < 1e-5, (k, float(parts[k]), float(z[f"loss/{tag}/{k}"]))
    assert _rel(total, z[f"loss/{tag}/total"]) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["default", "shepard"])
def test_forward_and_losses_match_reference_chain(pkg, dev, tag):
    """Product forward in fp32 mode (eval: noise-free quantisation, as the fixture) + product losses vs the reference chain."""
    ME = importlib.import_module(pkg.__name__ + ".MinkowskiEngine")
    loss_mod = importlib.import_module(pkg.__name__ + ".model.loss")
    z, sd = load_golden("ref_losses")
    net = build_product_model(pkg, golden_config(z), sd, dev)
    prev = pkg.sparse_ops.set_conv_mode("f32")
    try:
        with torch.no_grad():
            x = ME.SparseTensor(features=torch.from_numpy(z["colors"]).to(dev), coordinates=torch.from_numpy(z["coords"]).to(dev), device=dev)
            out = net(x, torch.from_numpy(z["q"]).to(dev), torch.from_numpy(z["lam"]).to(dev))
            total, parts = loss_mod.Loss(_loss_cfg(loss_mod, tag))(x, out)
    finally:
        pkg.sparse_ops.set_conv_mode(prev)
    for k in TERMS:
        assert _rel(parts[k], z[f"loss/{tag}/{k}"]) < 1e-4, (k, float(parts[k]), float(z[f"loss/{tag}/{k}"]))
    assert _rel(total, z[f"loss/{tag}/total"]) < 1e-4


def test_loss_fixture_is_what_the_reference_computes():
    """CPU (build container only): re-run the reference's own loss.py over the oracle shim on the stored forward output."""
    from oracle import shim
    if not shim.available():
        pytest.skip("needs /root/reference")
    assert os.path.exists(os.path.join(GOLDEN, "ref_losses.npz"))
    import copy
    import yaml
    shim.install()
    import MinkowskiEngine as ME          # the oracle, through the shim
    import loss as ref_loss               # /root/reference/loss.py, unmodified
    z, _ = load_golden("ref_losses")
    t = lambda a: torch.from_numpy(np.asarray(a))
    sparse = lambda C, F, ts=1: ME.SparseTensor(features=t(F), coordinates=t(C), tensor_stride=int(ts))
    gt = sparse(z["coords"], z["colors"])
    for tag, cfg_name in (("default", "CVPR_inverse_scaling"), ("shepard", "CVPR_inverse_scaling_shepard")):
        pred = {"prediction": sparse(z["pred_C"], z["pred_F"]),
                "occ_predictions": [sparse(z[f"occ{i}_C"], z[f"occ{i}_F"], z[f"occ{i}_ts"]) for i in range(3)],
                "points": [sparse(z[f"points{i}_C"], np.ones((len(z[f"points{i}_C"]), 1), np.float32), z[f"points{i}_ts"]) for i in range(3)],
                "q_map": t(z["lam"]), "likelihoods": {"y": t(z["lik_y"]), "z": t(z["lik_z"])}}
        cfg = yaml.safe_load(open(os.path.join(shim.REFERENCE, "configs", cfg_name + ".yaml")))["loss"]
        with torch.no_grad():
            total, parts = ref_loss.Loss(copy.deepcopy(cfg))(gt, pred)
        for k in TERMS:
            assert _rel(parts[k], z[f"loss/{tag}/{k}"]) < 1e-6, (tag, k)
