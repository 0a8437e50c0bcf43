"""Generate tests/golden/*.npz by running the reference's UNMODIFIED model/*.py (from /root/reference) on top
of the CPU oracle (oracle/shim.py).  TEST INFRASTRUCTURE; run in the build container only:

    python -m oracle.make_golden

The fixtures pin (a) oracle/pipeline_ref.py and (b) the product's host mirror + CUDA path against the
reference's own orchestration code.  Widths are reduced (8 channels) so the state dict fits a small fixture;
weights are random-init (torch.manual_seed) with the last analysis conv and the hyper-synthesis output scaled
up so that the quantised symbols are not all zero (the reference ships no trained weights, README.md:122).
"""
import os
import sys

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import shim, synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def small_config(name="CVPR_inverse_scaling", width=8):
    cfg = yaml.safe_load(open(os.path.join(shim.REFERENCE, "configs", name + ".yaml")))["model"]
    for sec in ("g_a", "g_s"):
        for k in ("N1", "N2", "N3", "N4"):
            cfg[sec][k] = width
    cfg["entropy_model"]["C_bottleneck"] = width
    cfg["entropy_model"]["C_hyper_bottleneck"] = width
    return cfg


def spread_weights(net):
    with torch.no_grad():
        net.g_a.down_conv_3[2].kernel.mul_(60.0)
        net.entropy_model.h_a[4].kernel.mul_(30.0)
        net.entropy_model.h_s[4].kernel.mul_(8.0)
        net.entropy_model.h_s[4].bias.add_(0.5)
        for head in (net.g_s.predict_1, net.g_s.predict_2, net.g_s.predict_3):
            head[0].bias.add_(0.3)          # keep the (few) hidden channels of the occupancy heads alive past the ReLU
            head[2].kernel.mul_(20.0)       # and the logits well separated
    net.update()


def make(name, cfg_name, seed, q):
    cfg = small_config(cfg_name)
    net = shim.reference_model(cfg, seed=seed)
    spread_weights(net)
    pts, rgb = synth.small_cloud(seed=seed + 1, radius=11.0, centre=(20, 20, 20), n_samples=9000)
    pc = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1))
    qt = torch.tensor([q], dtype=torch.float32)
    with torch.no_grad():
        bits, shapes, ks, coords, qv = net.compress(pc.clone(), qt, block_size=1024)
        rec = net.decompress(coordinates=coords, strings=bits, shape=shapes, k=ks, q_vals=qv)
        # eval-mode training forward (entropy_models.py:236-340) for likelihood parity
        import MinkowskiEngine as ME
        C4 = torch.cat([torch.zeros(len(pts), 1), torch.from_numpy(pts).float()], dim=1)
        x = ME.SparseTensor(coordinates=C4, features=torch.from_numpy(rgb))
        out = net(x, qt, torch.tensor([[1.0, 1.0]]))
    sd = {k: v.numpy() for k, v in net.state_dict().items()}
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        cfg_name=cfg_name, width=8, q=np.asarray(q, np.float32), points=pts, colors=rgb,
        y_coords=coords[0].numpy(), k=np.asarray(ks[0]).reshape(-1), z_shape=np.asarray(shapes[0]),
        y_string=np.frombuffer(bits[0][0][0], np.uint8), z_string=np.frombuffer(bits[0][1][0], np.uint8),
        reconstruction=rec.numpy(),
        fwd_pred_C=out["prediction"].C.numpy(), fwd_pred_F=out["prediction"].F.numpy(),
        fwd_lik_y=out["likelihoods"]["y"].numpy(), fwd_lik_z=out["likelihoods"]["z"].numpy(),
        fwd_occ3_C=out["occ_predictions"][2].C.numpy(), fwd_occ3_F=out["occ_predictions"][2].F.numpy(),
        **{"sd/" + k: v for k, v in sd.items()})
    print(name, "points", len(pts), "k", ks, "bytes y/z", len(bits[0][0][0]), len(bits[0][1][0]), "rec", tuple(rec.shape))


def make_multiblock(name="ref_multiblock", seed=4, q=(0.5, 0.5), block_size=24):
    """The reference's own block loop (model/model.py:121-176 compress, :225-238 decompress) on a cloud that falls into several
    blocks: partition order, per-block latent coordinates / level sizes / strings, and the concatenated reconstruction."""
    cfg = small_config("CVPR_inverse_scaling")
    net = shim.reference_model(cfg, seed=seed)
    spread_weights(net)
    pts, rgb = synth.small_cloud(seed=seed + 1, radius=20.0, centre=(32, 32, 32), n_samples=40000)
    pc = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1))
    qt = torch.tensor([q], dtype=torch.float32)
    with torch.no_grad():
        bits, shapes, ks, coords, qv = net.compress(pc.clone(), qt, block_size=block_size)
        rec = net.decompress(coordinates=coords, strings=bits, shape=shapes, k=ks, q_vals=qv)
    per_block = {}
    for i in range(len(bits)):
        per_block[f"y_coords_{i}"] = coords[i].numpy()
        per_block[f"k_{i}"] = np.asarray(ks[i]).reshape(-1)
        per_block[f"z_shape_{i}"] = np.asarray(shapes[i])
        per_block[f"y_string_{i}"] = np.frombuffer(bits[i][0][0], np.uint8)
        per_block[f"z_string_{i}"] = np.frombuffer(bits[i][1][0], np.uint8)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), cfg_name="CVPR_inverse_scaling", width=8, q=np.asarray(q, np.float32),
                        block_size=block_size, n_blocks=len(bits), points=pts, colors=rgb, reconstruction=rec.numpy(), **per_block,
                        **{"sd/" + k: v.numpy() for k, v in net.state_dict().items()})
    print(name, "points", len(pts), "blocks", len(bits), "k", [np.asarray(k).reshape(-1).tolist() for k in ks][:4], "rec", tuple(rec.shape))


def make_losses(name="ref_losses", seed=3, q=(0.4, 0.8), lam=(0.7, 1.3)):
    """Loss values of the reference's UNMODIFIED loss.py (/root/reference/loss.py:9-274: Loss, BPPLoss, ColorLoss,
    Multiscale_FocalLoss, ShepardsLoss with utils.overlapping_mask) on the training forward of its own model, both for the
    `loss:` section of configs/CVPR_inverse_scaling.yaml and of configs/CVPR_inverse_scaling_shepard.yaml.  The fixture keeps
    the complete forward output (prediction, the three occupancy predictions, the three ground-truth coordinate sets, both
    likelihood tensors), so the product's loss mirror can be checked on IDENTICAL inputs, and the weights + input so the whole
    chain (product forward -> product loss) can be checked too."""
    import copy
    cfg = small_config("CVPR_inverse_scaling")
    net = shim.reference_model(cfg, seed=seed)
    spread_weights(net)
    import MinkowskiEngine as ME
    import loss as ref_loss        # the reference's file (imports its utils.py over the open3d stand-in)
    clouds = [synth.small_cloud(seed=seed + 1 + b, radius=9.0 + b, centre=(18, 18, 18), n_samples=6000) for b in range(2)]
    coords = torch.cat([torch.cat([torch.full((len(p), 1), b, dtype=torch.int32), torch.from_numpy(p).int()], 1) for b, (p, _) in enumerate(clouds)])
    feats = torch.cat([torch.from_numpy(c.astype(np.float32)) for _, c in clouds])
    qt = torch.tensor([q, q[::-1]], dtype=torch.float32)
    lam_t = torch.tensor([lam, (1.1, 0.6)], dtype=torch.float32)
    store = {}
    with torch.no_grad():
        x = ME.SparseTensor(coordinates=coords, features=feats)
        out = net(x, qt, lam_t)
        for cfg_name in ("CVPR_inverse_scaling", "CVPR_inverse_scaling_shepard"):
            loss_cfg = yaml.safe_load(open(os.path.join(shim.REFERENCE, "configs", cfg_name + ".yaml")))["loss"]
            pred = dict(out, occ_predictions=list(out["occ_predictions"]), points=list(out["points"]))   # (the focal loss reverses the lists in place)
            total, parts = ref_loss.Loss(copy.deepcopy(loss_cfg))(x, pred)
            tag = "shepard" if cfg_name.endswith("shepard") else "default"
            store[f"loss/{tag}/total"] = np.float64(total)
            for k, v in parts.items():
                store[f"loss/{tag}/{k}"] = np.float64(v)
            print(name, tag, float(total), {k: float(v) for k, v in parts.items()})
    for i in range(3):
        store[f"occ{i}_C"], store[f"occ{i}_F"] = out["occ_predictions"][i].C.numpy(), out["occ_predictions"][i].F.numpy()
        store[f"points{i}_C"] = out["points"][i].C.numpy()
        store[f"points{i}_ts"] = np.asarray(out["points"][i].tensor_stride[0])
        store[f"occ{i}_ts"] = np.asarray(out["occ_predictions"][i].tensor_stride[0])
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"), cfg_name="CVPR_inverse_scaling", width=8, q=qt.numpy(), lam=lam_t.numpy(),
        coords=coords.numpy(), colors=feats.numpy(), pred_C=out["prediction"].C.numpy(), pred_F=out["prediction"].F.numpy(),
        lik_y=out["likelihoods"]["y"].numpy(), lik_z=out["likelihoods"]["z"].numpy(),
        **store, **{"sd/" + k: v.numpy() for k, v in net.state_dict().items()})


if __name__ == "__main__":
    assert shim.available(), "needs /root/reference"
    os.makedirs(OUT, exist_ok=True)
    what = sys.argv[1:] or ["codec", "losses", "multiblock"]
    if "codec" in what:
        make("ref_inverse_scaling", "CVPR_inverse_scaling", 0, (0.5, 0.5))
        make("ref_inverse_nn", "CVPR_inverse_nn", 1, (0.2, 0.9))
        make("ref_fixed_R1", "CVPR_inverse_scaling_fixed_R1", 2, (0.0, 0.0))
    if "losses" in what:
        make_losses()
    if "multiblock" in what:
        make_multiblock()
