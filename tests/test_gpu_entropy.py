"""CUDA entropy-model kernels vs the oracle's compressai restatement on identical inputs.
Symbols / indexes: bit-exact.  Likelihoods: <= 1e-2 relative L2 (north-star), in practice ~1e-6."""
import importlib

import numpy as np
import pytest
import torch

from oracle import cai_ref
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu


def models(pkg, dev, C=12):
    pem = importlib.import_module(pkg.__name__ + ".compressai.entropy_models")
    torch.manual_seed(0)
    e_ref = cai_ref.EntropyBottleneck(C)
    with torch.no_grad():
        for p in e_ref.factors:
            p.uniform_(-0.5, 0.5)
        e_ref.quantiles[:, 0, 1] = torch.randn(C) * 0.4
    e = pem.EntropyBottleneck(C)
    e.load_state_dict({k: v for k, v in e_ref.state_dict().items() if not k.startswith("_")}, strict=False)
    g_ref, g = cai_ref.GaussianConditional(None), pem.GaussianConditional(None)
    for m in (e_ref, e):
        m.update(force=True)
    g_ref.update_scale_table(cai_ref.get_scale_table())
    g.update_scale_table(pem.get_scale_table())
    return e_ref.eval(), e.to(dev).eval(), g_ref.eval(), g.to(dev).eval()


@pytest.mark.parametrize("layout", ["cn", "nc_view"])
def test_entropy_bottleneck_forward(pkg, dev, layout):
    e_ref, e, _, _ = models(pkg, dev)
    N = 777
    z_nc = torch.randn(N, 12) * 3
    x_ref = z_nc.t().unsqueeze(0)
    x = (z_nc.to(dev).t().unsqueeze(0) if layout == "nc_view" else z_nc.t().contiguous().unsqueeze(0).to(dev))
    with torch.no_grad():
        out_ref, lik_ref = e_ref(x_ref)
        out, lik = e(x)
    assert torch.equal(out.cpu(), out_ref)
    assert rel_l2(lik.cpu().numpy(), lik_ref.numpy()) < 1e-4
    assert lik.min() >= 1e-9


def test_gaussian_conditional_forward_indexes_symbols(pkg, dev):
    _, _, g_ref, g = models(pkg, dev)
    N, C = 999, 16
    y = torch.randn(N, C) * 5
    s = torch.exp(torch.randn(N, C) * 2)
    m = torch.randn(N, C)
    v = lambda t: t.t().unsqueeze(0)
    with torch.no_grad():
        out_ref, lik_ref = g_ref(v(y), v(s), means=v(m))
        out, lik = g(v(y.to(dev)), v(s.to(dev)), means=v(m.to(dev)))
        idx_ref = g_ref.build_indexes(v(s))
        idx = g.build_indexes(v(s.to(dev)))
        sym_ref = g_ref.quantize(v(y), "symbols", v(m))
        sym = g.quantize(v(y.to(dev)), "symbols", v(m.to(dev)))
    assert torch.equal(out.cpu(), out_ref)
    assert rel_l2(lik.cpu().numpy(), lik_ref.numpy()) < 1e-4
    assert torch.equal(idx.cpu(), idx_ref) and idx.dtype == torch.int32
    assert torch.equal(sym.cpu(), sym_ref)
    # table edges: scales exactly on table entries must index identically (<= comparison)
    edge = g_ref.scale_table.repeat(3).reshape(1, 3, -1)
    assert torch.equal(g.build_indexes(edge.to(dev)).cpu(), g_ref.build_indexes(edge))


def test_compress_decompress_strings_match_oracle(pkg, dev):
    e_ref, e, g_ref, g = models(pkg, dev)
    z = torch.randn(1, 12, 50) * 4
    s_ref = e_ref.compress(z)
    s = e.compress(z.to(dev))
    assert s == s_ref
    assert torch.equal(e.decompress(s, [50]).cpu(), e_ref.decompress(s_ref, [50]))
    y = torch.randn(1, 16, 300) * 20
    sc = torch.exp(torch.randn(1, 16, 300))
    mu = torch.randn(1, 16, 300)
    idx_ref = g_ref.build_indexes(sc)
    idx = g.build_indexes(sc.to(dev))
    ys_ref = g_ref.compress(y, idx_ref, means=mu)
    ys = g.compress(y.to(dev), idx, means=mu.to(dev))
    assert ys == ys_ref
    assert torch.equal(g.decompress(ys, idx, means=mu.to(dev)).cpu(), g_ref.decompress(ys_ref, idx_ref, means=mu))


def test_uninitialised_tables_raise(pkg, dev):
    pem = importlib.import_module(pkg.__name__ + ".compressai.entropy_models")
    e = pem.EntropyBottleneck(4).to(dev)
    with pytest.raises(ValueError, match="Uninitialized CDFs"):
        e.compress(torch.zeros(1, 4, 3, device=dev))


def test_dequant_offsets_kernel(pkg, dev):
    eo = pkg.entropy_ops if hasattr(pkg, "entropy_ops") else importlib.import_module(pkg.__name__ + ".entropy_ops")
    torch.manual_seed(2)
    N, C = 500, 8
    mlp = torch.nn.Sequential(torch.nn.Linear(2, 10), torch.nn.ReLU(), torch.nn.Linear(10, 10), torch.nn.ReLU(), torch.nn.Linear(10, 1))
    q = torch.round(torch.randn(1, C, N) * 3)
    sh, mh = torch.exp(torch.randn(1, C, N)), torch.randn(1, C, N)
    gain = (torch.rand(1, C) + 0.5)[torch.zeros(N, dtype=torch.long)].t().unsqueeze(0)
    rescale = 1.0 / gain
    sd = torch.clamp(sh * gain, min=0.11)
    with torch.no_grad():
        off = -mlp(torch.cat((gain.expand_as(sd).unsqueeze(3), sd.unsqueeze(3)), dim=3)).squeeze(3)
    off[q.abs() < 1e-4] = 0
    ref = torch.sign(q) * (q.abs() + off) * rescale + mh
    out = eo.gc_dequant_offsets(q.to(dev), sh.to(dev), mh.to(dev), gain.to(dev), rescale.to(dev),
                                eo.pack_quant_mlp(mlp).to(dev), 0.11).cpu()
    assert rel_l2(out.numpy(), ref.numpy()) < 1e-5
