"""Run the reference's UNMODIFIED model/*.py (from /root/reference) on top of the CPU oracle.
TEST INFRASTRUCTURE; only usable in the build container (the GPU box has no /root/reference).

Registers oracle/me_ref.py as `MinkowskiEngine`, oracle/cai_ref.py as `compressai.*`, and inert stand-ins
for `open3d` / `bitstream` (imported at module top of model/model.py:5,8, used only on out-of-scope paths).
"""
import os
import sys
import types

REFERENCE = "/root/reference"


def available():
    return os.path.isdir(os.path.join(REFERENCE, "model"))


def install():
    from . import cai_ref, me_ref
    me = types.ModuleType("MinkowskiEngine")
    for name in dir(me_ref):
        if name.startswith("Minkowski") or name in ("SparseTensor", "utils"):
            setattr(me, name, getattr(me_ref, name))
    sys.modules["MinkowskiEngine"] = me
    sys.modules["MinkowskiEngine.utils"] = me_ref.utils

    cai = types.ModuleType("compressai")
    ops_pkg = types.ModuleType("compressai.ops")
    ops_mod = types.ModuleType("compressai.ops.ops")
    ops_mod.quantize_ste = cai_ref.quantize_ste
    ops_pkg.ops, ops_pkg.LowerBound, ops_pkg.quantize_ste = ops_mod, cai_ref.LowerBound, cai_ref.quantize_ste
    layers = types.ModuleType("compressai.layers")
    layers.GDN = cai_ref.GDN
    em = types.ModuleType("compressai.entropy_models")
    em.EntropyBottleneck, em.GaussianConditional, em.EntropyModel = cai_ref.EntropyBottleneck, cai_ref.GaussianConditional, cai_ref.EntropyModel
    models = types.ModuleType("compressai.models")
    base = types.ModuleType("compressai.models.base")
    base.CompressionModel = cai_ref.CompressionModel
    models.base = base
    cai.ops, cai.layers, cai.entropy_models, cai.models = ops_pkg, layers, em, models
    for name, mod in (("compressai", cai), ("compressai.ops", ops_pkg), ("compressai.ops.ops", ops_mod),
                      ("compressai.layers", layers), ("compressai.entropy_models", em), ("compressai.models", models),
                      ("compressai.models.base", base)):
        sys.modules[name] = mod
    for name in ("open3d", "bitstream"):
        if name not in sys.modules:
            stub = types.ModuleType(name)
            stub.BitStream = object
            sys.modules[name] = stub
    if REFERENCE not in sys.path:
        sys.path.insert(0, REFERENCE)


def reference_model(config, seed=0):
    """The reference's own UnifiedModel (random-init, eval, CDFs updated), running on the oracle."""
    import torch
    install()
    from model.model import UnifiedModel  # noqa: the reference's file
    torch.manual_seed(seed)
    net = UnifiedModel(config)
    net.eval()
    net.update()
    return net
