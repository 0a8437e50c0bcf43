"""Shared test helpers (also used by __graft_entry__.smoke())."""
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

BASE_CFG = {
    "entropy_model": {"type": "MeanScaleHyperprior_map", "C_bottleneck": 128, "C_hyper_bottleneck": 192,
                      "quantization_mode": "ste", "inverse_rescaling": True, "quantization_offset": True,
                      "entropy_bottleneck_vbr": False},
    "g_a": {"C_in": 4, "N1": 128, "N2": 128, "N3": 128, "N4": 128},
    "g_s": {"C_out": 3, "N1": 128, "N2": 128, "N3": 128, "N4": 128},
}


def model_config(width=128, hyper=None, variant="inverse_scaling"):
    """configs/CVPR_*.yaml `model:` section (widths optionally reduced for tests)."""
    import copy
    cfg = copy.deepcopy(BASE_CFG)
    for sec in ("g_a", "g_s"):
        for k in ("N1", "N2", "N3", "N4"):
            cfg[sec][k] = width
    cfg["entropy_model"]["C_bottleneck"] = width
    cfg["entropy_model"]["C_hyper_bottleneck"] = hyper if hyper is not None else (192 if width == 128 else width)
    if variant == "inverse_nn":
        cfg["entropy_model"]["inverse_rescaling"] = False
    elif variant == "fixed":
        cfg["entropy_model"].update(inverse_rescaling=False, quantization_offset=False, adaptive_BN=False)
    return cfg


def spread_weights(net):
    """Same perturbation as oracle/make_golden.py so quantised symbols are not all zero."""
    with torch.no_grad():
        net.g_a.down_conv_3[2].kernel.mul_(60.0)
        net.entropy_model.h_a[4].kernel.mul_(30.0)
        net.entropy_model.h_s[4].kernel.mul_(8.0)
        net.entropy_model.h_s[4].bias.add_(0.5)
        for head in (net.g_s.predict_1, net.g_s.predict_2, net.g_s.predict_3):
            head[0].bias.add_(0.3)          # keep the (few) hidden channels of the occupancy heads alive past the ReLU
            head[2].kernel.mul_(20.0)       # and the logits well separated
    net.update()


def reference_dir():
    """The reference's own source tree: /root/reference in the build container, the git-ignored copy that
    `__graft_entry__.build()` leaves under baseline/_ref/ on the GPU box.  None when neither exists."""
    for d in ("/root/reference", os.path.join(ROOT, "baseline", "_ref", "reference")):
        if os.path.isfile(os.path.join(d, "model", "model.py")):
            return d
    return None


_REF_TOP = ("model", "utils", "loss", "MinkowskiEngine", "compressai", "open3d", "bitstream")


import contextlib


@contextlib.contextmanager
def reference_on_product(pkg):
    """The reference's UNMODIFIED model/*.py, loss.py and utils.py imported on top of the PRODUCT's drop-in packages
    (`pkg.install_shims()`): what a user of the reference gets after swapping MinkowskiEngine / compressai for the graft.
    Yields a namespace (model_module, loss_module, utils_module, directory).  sys.modules / sys.path are restored afterwards, so
    the oracle-backed import of the same files (oracle/shim.py) in the same process is not disturbed."""
    import sys
    import types
    ref = reference_dir()
    assert ref is not None, "no reference tree"
    mine = lambda: [m for m in list(sys.modules) if m.split(".")[0] in _REF_TOP]
    saved = {m: sys.modules.pop(m) for m in mine()}
    saved_path = list(sys.path)
    try:
        pkg.install_shims()
        for name in ("open3d", "bitstream"):          # imported at the top of model/model.py:5,8, used on out-of-scope paths only
            stub = types.ModuleType(name)
            stub.BitStream = object
            sys.modules[name] = stub
        sys.path.insert(0, ref)
        import model.model as ref_model               # noqa: the reference's files
        import loss as ref_loss
        import utils as ref_utils
        assert os.path.dirname(os.path.abspath(ref_model.__file__)) == os.path.join(ref, "model")
        yield types.SimpleNamespace(model=ref_model, loss=ref_loss, utils=ref_utils, dir=ref)
    finally:
        for m in mine():
            sys.modules.pop(m, None)
        sys.modules.update(saved)
        sys.path[:] = saved_path


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd/")}
    return z, sd


def golden_config(z):
    variant = {"CVPR_inverse_scaling": "inverse_scaling", "CVPR_inverse_nn": "inverse_nn",
               "CVPR_inverse_scaling_fixed_R1": "fixed"}[str(z["cfg_name"])]
    return model_config(width=int(z["width"]), variant=variant)


def coord_keys(C):
    C = np.asarray(C).astype(np.int64)
    if C.shape[1] == 3:
        C = np.concatenate([np.zeros((len(C), 1), np.int64), C], axis=1)
    return (C[:, 0] << 48) | ((C[:, 1] + 32768) << 32) | ((C[:, 2] + 32768) << 16) | (C[:, 3] + 32768)


def align_rows(C_a, C_b):
    """Indices (ia, ib) of the rows the two coordinate lists share, matched by coordinate."""
    ka, kb = coord_keys(C_a), coord_keys(C_b)
    common, ia, ib = np.intersect1d(ka, kb, return_indices=True)
    return ia, ib


def rel_l2(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def build_product_model(pkg, cfg, state_dict, device):
    import importlib
    model_mod = importlib.import_module(pkg.__name__ + ".model")
    net = model_mod.UnifiedModel(cfg)
    missing, unexpected = net.load_state_dict(dict(state_dict), strict=False)
    bad = [k for k in missing if not k.split(".")[-1].startswith("_")]
    assert not bad, f"state-dict keys missing in the product mirror: {bad}"
    assert not [k for k in unexpected if not k.split(".")[-1].startswith("_")], f"unexpected keys: {unexpected}"
    net = net.to(device).eval()
    net.update()
    return net


def tiny_codec_roundtrip(pkg, device, width=32, seed=0, mode="auto", return_clouds=False, cloud=None, state_dict=None,
                         min_overlap=None, max_symbol_mismatch=None):
    """Encode + decode a small sphere on the GPU and check it against oracle/pipeline_ref.py on the same
    weights: stride-8 coordinates and k bit-exact, symbols near-exact, decoded geometry/colours close.

    mode "f32": every convolution on the fp32 SIMT path -> strict agreement with the fp32 oracle.
    mode "auto": tcgen05 bf16 where eligible.  With RANDOM-INIT weights the occupancy logits are barely
    separated, so bf16 rounding legitimately flips top-k membership near the threshold (SURVEY.md §7 "top-k
    under reduced precision") and the flips compound over the three levels; the geometry bound there is a sanity
    check only -- the principled statement (selection errors confined to the numerical error band of the logits)
    is tests/test_gpu_conv.py::test_topk_selection_consistent_with_fp32_logits."""
    import importlib
    from oracle import pipeline_ref, synth
    model_mod = importlib.import_module(pkg.__name__ + ".model")
    cfg = model_config(width=width)
    torch.manual_seed(seed)
    net = model_mod.UnifiedModel(cfg)
    if state_dict is None:
        spread_weights(net)
    else:                                   # e.g. the trained fixture: occupancy logits separated by training, not by scaling
        net.load_state_dict({k: v for k, v in state_dict.items() if not k.split(".")[-1].startswith("_")}, strict=False)
        net.update()
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    net = net.to(device).eval()
    net.update()
    pts, rgb = cloud if cloud is not None else synth.small_cloud(seed=seed + 1, radius=11.0, centre=(20, 20, 20), n_samples=9000)
    q = torch.tensor([[0.5, 0.5]], device=device)
    x_block = torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1)).to(device)
    prev = pkg.sparse_ops.set_conv_mode(mode)
    try:
        with torch.no_grad():
            enc = net.encode_block(x_block, q)
            y_sym = enc["y_symbols"]
            x_hat = net.decode_block(enc["y_points"], enc["z_symbols"], lambda idx: y_sym, enc["k"], q)
            rec = net.finalize(x_hat.C, x_hat.F).cpu().numpy()
    finally:
        pkg.sparse_ops.set_conv_mode(prev)
    ref = pipeline_ref.RefCodec(sd, cfg)
    renc = ref.encode_block(pts, rgb, [0.5, 0.5])
    assert enc["k"] == renc["k"], (enc["k"], renc["k"])
    assert (enc["y_points"].cpu().numpy() == renc["y_points"]).all(), "stride-8 coordinates differ"
    assert (enc["z_points"].cpu().numpy() == renc["z_points"]).all(), "stride-32 coordinates differ"
    ys, rys = enc["y_symbols"].cpu().numpy(), renc["y_symbols"]
    frac = float((ys != rys).mean())
    sym_tol = max_symbol_mismatch if max_symbol_mismatch is not None else (0.002 if mode == "f32" else 0.03)
    assert frac <= sym_tol and np.abs(ys - rys).max() <= 1, f"y symbols differ on {frac:.3%} of elements"
    rrec = ref.decode_block(renc["y_points"], renc["z_symbols"], renc["y_symbols"], renc["k"], [0.5, 0.5])
    ia, ib = align_rows(rec[:, :3], rrec[:, :3])
    assert len(rec) == len(rrec)
    # random-init weights in bf16 mode: a sanity bound only (top-k flips near barely separated logits compound over three levels).
    # The parity gate of the timed bf16 mode is test_trained_codec_matches_fp32_oracle (trained weights: min_overlap 0.99, symbols
    # 99.5 %, PSNR within 0.05 dB), which passes min_overlap / max_symbol_mismatch explicitly.
    need = min_overlap if min_overlap is not None else (0.99 if mode == "f32" else 0.25)
    assert len(ia) >= need * len(rrec), f"decoded geometry overlap {len(ia)}/{len(rrec)}"
    assert np.abs(rec[ia, 3:] - rrec[ib, 3:]).mean() < 0.02
    stats = {"n": len(pts), "overlap": len(ia) / len(rrec), "symbol_mismatch": frac}
    if return_clouds:
        stats.update(source=np.concatenate([pts.astype(np.float64), rgb], axis=1), rec=rec.astype(np.float64), ref_rec=np.asarray(rrec, np.float64))
    return stats
