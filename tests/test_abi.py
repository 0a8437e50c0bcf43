"""The C-ABI library loads and exports every symbol include/pcg.h declares (no compute without a GPU)."""
import ctypes
import importlib
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pcg.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pcg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in pcg.h but not exported"
    assert sorted(pkg._lib.exported_symbols()) == names, "ctypes prototypes and pcg.h disagree"


def test_host_only_entry_points():
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    lib = pkg._lib.load()
    assert lib.pcg_version() == 100
    assert lib.pcg_hash_capacity(1000) == 2048 and lib.pcg_hash_capacity(3) == 1024
    assert lib.pcg_scratch_bytes(10_000) > 10_000
    assert lib.pcg_packed_weight_bytes(27, 128, 64) == 27 * 2 * 64 * 128
    assert lib.pcg_launch_count() == 0


def test_cpu_tensors_are_rejected_loudly():
    import pytest
    import torch
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    with pytest.raises(RuntimeError, match="no CPU path|CUDA"):
        pkg.sparse_ops.CoordSet(torch.zeros((4, 4), dtype=torch.int32), 1)
    me = importlib.import_module(pkg.__name__ + ".MinkowskiEngine")
    with pytest.raises(RuntimeError, match="no CPU path"):
        me.SparseTensor(torch.zeros(4, 1), torch.zeros((4, 4), dtype=torch.int32))


def test_reference_state_dict_loads_into_product_mirror():
    """Checkpoint compatibility: the golden state dict was saved by the reference's own UnifiedModel."""
    import torch
    from tests.helpers import golden_config, load_golden
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    model_mod = importlib.import_module(pkg.__name__ + ".model")
    z, sd = load_golden("ref_inverse_scaling")
    net = model_mod.UnifiedModel(golden_config(z))
    missing, unexpected = net.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    assert net.entropy_model.gaussian_conditional._quantized_cdf.shape == sd["entropy_model.gaussian_conditional._quantized_cdf"].shape
    net.update()        # host-side table rebuild must reproduce the reference's tables exactly
    assert torch.equal(net.entropy_model.gaussian_conditional._quantized_cdf, sd["entropy_model.gaussian_conditional._quantized_cdf"])
    assert torch.equal(net.entropy_model.entropy_bottleneck._quantized_cdf, sd["entropy_model.entropy_bottleneck._quantized_cdf"])
    assert torch.equal(net.entropy_model.entropy_bottleneck._offset, sd["entropy_model.entropy_bottleneck._offset"])


def test_static_kernel_selection_rules_cover_the_reference_heads():
    """Host logic: the occupancy heads of the reference configurations (model/transforms.py:146-160 with N = 128: 128->64->1 twice,
    32->16->1) take the fused GEMM-first path; the rule is static (shapes and conv mode only -- never timing),
    switches off in fp32 parity mode, and shapes outside the tested set stay on the two-convolution path."""
    import torch
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    so = pkg.sparse_ops
    fus = lambda cin, cmid, k1=3, k2=3: so.head_fusable(k1, torch.zeros(k1 ** 3, cin, cmid), k2, torch.zeros(k2 ** 3, cmid, 1))
    assert fus(128, 64) and fus(32, 16) and fus(64, 32)
    assert not fus(16, 8) and not fus(8, 4) and not fus(32, 32) and not fus(128, 128)
    assert not fus(128, 64, k1=5) and not so.head_fusable(3, torch.zeros(27, 128, 64), 3, torch.zeros(27, 64, 2))
    prev = so.set_conv_mode("f32")
    try:
        assert not fus(128, 64)
    finally:
        so.set_conv_mode(prev)
    assert so.tc_eligible(27, 128, 64) and so.tc_eligible(125, 128, 128) and not so.tc_eligible(1, 32, 3)
    assert so.cells_eligible(5, 2, 128, 128) and so.cells_eligible(5, 2, 128, 32) and not so.cells_eligible(2, 2, 192, 192)
