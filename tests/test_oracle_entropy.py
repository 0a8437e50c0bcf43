"""Pin the oracle's entropy-model restatement against scipy and by round trips."""
import importlib

import numpy as np
import pytest
import scipy.special
import scipy.stats
import torch

from oracle import cai_ref, rans_ref


def test_gaussian_likelihood_matches_scipy():
    torch.manual_seed(0)
    gc = cai_ref.GaussianConditional(None)
    x = torch.randn(1, 4, 500) * 3
    s = torch.rand(1, 4, 500) * 4 + 0.01
    m = torch.randn(1, 4, 500)
    out, lik = gc(x, s, means=m, training=False)
    assert torch.equal(out, torch.round(x - m) + m)
    v = np.abs((out - m).double().numpy())
    sd = np.maximum(s.double().numpy(), 0.11)
    ref = scipy.stats.norm.cdf((0.5 - v) / sd) - scipy.stats.norm.cdf((-0.5 - v) / sd)
    assert np.allclose(lik.numpy(), np.maximum(ref, 1e-9), rtol=2e-4, atol=1e-7)


def test_entropy_bottleneck_matches_expit_chain():
    torch.manual_seed(1)
    eb = cai_ref.EntropyBottleneck(6)
    with torch.no_grad():
        for p in eb.factors:
            p.uniform_(-0.5, 0.5)
        eb.quantiles[:, 0, 1] = torch.randn(6) * 0.3
    x = torch.randn(1, 6, 300) * 4
    eb.eval()
    out, lik = eb(x)
    med = eb.quantiles[:, 0, 1].detach().reshape(1, 6, 1)
    assert torch.allclose(out, torch.round(x - med) + med)

    def logits(v):          # numpy float64 restatement of the cumulative chain
        h = v.transpose(1, 0, 2).astype(np.float64)                       # (C,1,N)
        for i in range(5):
            M = np.log1p(np.exp(eb.matrices[i].detach().double().numpy()))
            h = M @ h + eb.biases[i].detach().double().numpy()
            if i < 4:
                h = h + np.tanh(eb.factors[i].detach().double().numpy()) * np.tanh(h)
        return h
    o = out.detach().numpy()
    ref = scipy.special.expit(logits(o + 0.5)) - scipy.special.expit(logits(o - 0.5))
    assert np.allclose(lik.detach().numpy().transpose(1, 0, 2), np.maximum(ref, 1e-9), rtol=1e-3, atol=1e-7)


def test_build_indexes_is_table_search():
    gc = cai_ref.GaussianConditional(None)
    gc.update_scale_table(cai_ref.get_scale_table())
    s = torch.exp(torch.linspace(-4, 6, 999)).reshape(1, 1, -1)
    idx = gc.build_indexes(s).numpy().ravel()
    table = gc.scale_table.numpy()
    ref = np.searchsorted(table[:-1], np.maximum(s.numpy().ravel(), np.float32(0.11)), side="left")
    assert (idx == ref).all() and idx.min() == 0 and idx.max() == 63


def test_cdf_tables_are_valid():
    gc = cai_ref.GaussianConditional(None)
    gc.update_scale_table(cai_ref.get_scale_table())
    cdf, ln = gc._quantized_cdf.numpy(), gc._cdf_length.numpy()
    for i in range(64):
        c = cdf[i, : ln[i]]
        assert c[0] == 0 and c[-1] == 65536 and (np.diff(c) > 0).all()
    assert (gc._offset.numpy() == -(ln - 3) // 2).all()
    eb = cai_ref.EntropyBottleneck(5)
    eb.update()
    for i in range(5):
        c = eb._quantized_cdf.numpy()[i, : eb._cdf_length.numpy()[i]]
        assert c[0] == 0 and c[-1] == 65536 and (np.diff(c) > 0).all()


def test_pmf_to_cdf_stealing_rule():
    cdf = cai_ref.pmf_to_quantized_cdf(np.array([0.5, 0.0, 0.0, 0.5, 1e-9], np.float32), 16)
    assert cdf[0] == 0 and cdf[-1] == 65536 and (np.diff(cdf) > 0).all()


@pytest.mark.parametrize("n", [0, 1, 5000])
def test_rans_round_trip_with_escapes(n):
    gc = cai_ref.GaussianConditional(None)
    gc.update_scale_table(cai_ref.get_scale_table())
    rng = np.random.default_rng(n)
    idx = rng.integers(0, 64, n).astype(np.int32)
    sym = np.round(rng.standard_normal(n) * np.exp(rng.uniform(-2, 6, n))).astype(np.int32)
    args = (gc._quantized_cdf.numpy(), gc._cdf_length.numpy(), gc._offset.numpy())
    s = rans_ref.encode_with_indexes(sym, idx, *args)
    assert (rans_ref.decode_with_indexes(s, idx, *args) == sym).all()


def test_product_host_coder_is_byte_identical_to_oracle():
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    hc = importlib.import_module(pkg.__name__ + ".host_coder")
    gc = cai_ref.GaussianConditional(None)
    gc.update_scale_table(cai_ref.get_scale_table())
    rng = np.random.default_rng(7)
    n = 20000
    idx = rng.integers(0, 64, n).astype(np.int32)
    sym = np.round(rng.standard_normal(n) * np.exp(rng.uniform(-2, 6, n))).astype(np.int32)
    args = (gc._quantized_cdf.numpy(), gc._cdf_length.numpy(), gc._offset.numpy())
    ref = rans_ref.encode_with_indexes(sym, idx, *args)
    assert hc.encode(sym, idx, *args) == ref
    assert (hc.decode(ref, idx, *args) == sym).all()
    for pmf in ([0.5, 0.0, 0.0, 0.5, 1e-9], list(np.random.default_rng(1).dirichlet(np.ones(40) * 0.1)) + [1e-9]):
        assert (hc.pmf_to_quantized_cdf(np.array(pmf, np.float32)) == cai_ref.pmf_to_quantized_cdf(np.array(pmf, np.float32))).all()


def test_pmf_to_cdf_rounds_like_std_round():
    """compressai rounds p * 2^16 with std::round; adding 0.5 in fp32 instead rounds 0.49999997 up to 1."""
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    hc = importlib.import_module(pkg.__name__ + ".host_coder")
    edge = np.float32(np.nextafter(np.float32(0.5), np.float32(0.0))) / np.float32(65536.0)     # p * 65536 == 0.49999997f exactly
    assert np.float32(edge * np.float32(65536.0)) + np.float32(0.5) == np.float32(1.0)            # the fp32 addition rounds up
    pmf = np.array([0.25, edge, 0.75], np.float32)
    # frequencies before renormalisation must be [16384, 0, 49152]: the zero then takes one count from the cheapest donor
    for fn in (cai_ref.pmf_to_quantized_cdf, hc.pmf_to_quantized_cdf):
        cdf = fn(pmf)
        assert list(np.diff(cdf)) == [16383, 1, 49152], list(np.diff(cdf))


def test_product_entropy_tables_match_oracle():
    """Same parameters -> same 16-bit CDF tables (they define cdf_length / offset, hence the bitstream)."""
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    pem = importlib.import_module(pkg.__name__ + ".compressai.entropy_models")
    assert abs(pem._norm_ppf(5e-10) - scipy.stats.norm.ppf(5e-10)) < 1e-9
    g_ref, g = cai_ref.GaussianConditional(None), pem.GaussianConditional(None)
    g_ref.update_scale_table(cai_ref.get_scale_table())
    g.update_scale_table(pem.get_scale_table())
    assert torch.equal(g_ref._quantized_cdf, g._quantized_cdf) and torch.equal(g_ref._offset, g._offset)
    torch.manual_seed(3)
    e_ref = cai_ref.EntropyBottleneck(7)
    e = pem.EntropyBottleneck(7)
    e.load_state_dict({k: v for k, v in e_ref.state_dict().items() if not k.startswith("_")}, strict=False)
    e_ref.update(); e.update()
    assert torch.equal(e_ref._quantized_cdf, e._quantized_cdf) and torch.equal(e_ref._cdf_length, e._cdf_length)
    assert set(e.state_dict()) == set(e_ref.state_dict())
