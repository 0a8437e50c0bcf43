"""Entropy-model arithmetic on top of libpcg.so (compressai 1.2.4 semantics).

Every function accepts tensors of any rank >= 2 laid out as (..., C, N)-like views: operands are
addressed through (row, channel) element strides, so the reference's `(1, C, N)` views of `(N, C)`
feature matrices (`y.F.t().unsqueeze(0)`) are consumed without a transposing copy.
"""
import torch

from ._lib import call, ptr, stream


def _span(kind, n, c, elems):
    """bench.py's roofline pass: CUDA events + algorithmic bytes (4-byte elements read once / written once)."""
    from . import sparse_ops
    return sparse_ops.hbm_span("entropy", (kind, c), 4.0 * n * c * elems)


def _as_cn(t):
    """View a (1,C,N) / (C,N) / (C,1,N)-style tensor as (n, c, stride_n, stride_c) without copying."""
    if t.dim() == 3:
        if t.shape[0] == 1:
            t2 = t[0]
        elif t.shape[1] == 1:
            t2 = t[:, 0]
        else:
            raise ValueError(f"unsupported entropy-model operand shape {tuple(t.shape)}")
    elif t.dim() == 2:
        t2 = t
    else:
        raise ValueError(f"unsupported entropy-model operand shape {tuple(t.shape)}")
    c, n = t2.shape
    return n, c, t2.stride(1), t2.stride(0)


def _need(t):
    if not t.is_cuda:
        raise RuntimeError("entropy ops need CUDA tensors: the graft has no CPU path")
    if t.dtype not in (torch.float32, torch.int32):
        raise RuntimeError(f"unsupported dtype {t.dtype}")


def _same_layout_out(x, dtype):
    """Output with exactly x's strides (so one (sn, sc) pair addresses both)."""
    return torch.empty_strided(x.shape, x.stride(), dtype=dtype, device=x.device)


def _expand_like(m, x):
    return m.expand(x.shape) if m.shape != x.shape else m


def eb_pack_params(matrices, biases, factors):
    """58 floats per channel for filters (3,3,3,3): softplus(M_l), b_l, tanh(f_l)  (see pcg.h)."""
    assert len(matrices) == 5 and tuple(matrices[0].shape[1:]) == (3, 1) and tuple(matrices[4].shape[1:]) == (1, 3), \
        "EntropyBottleneck kernels are specialised for filters=(3,3,3,3)"
    C = matrices[0].shape[0]
    parts = []
    for l in range(5):
        parts.append(torch.nn.functional.softplus(matrices[l].detach()).reshape(C, -1))
        parts.append(biases[l].detach().reshape(C, -1))
        if l < 4:
            parts.append(torch.tanh(factors[l].detach()).reshape(C, -1))
    p = torch.cat(parts, dim=1).contiguous().float()
    assert p.shape[1] == 58
    return p


def eb_likelihood(x, params, bound):
    _need(x)
    n, c, sn, sc = _as_cn(x)
    out = _same_layout_out(x, torch.float32)
    with _span("eb_likelihood", n, c, 2):
        call("pcg_eb_likelihood", ptr(x), n, c, sn, sc, ptr(params), float(bound), ptr(out), stream())
    return out


def quantize(x, means, mode):
    """mode 'dequantize' -> round(x - m) + m (fp32); 'symbols' -> int32 round(x - m)."""
    _need(x)
    n, c, sn, sc = _as_cn(x)
    if means is not None:
        means = _expand_like(means, x)
        _, _, msn, msc = _as_cn(means)
    else:
        msn = msc = 0
    if mode == "dequantize":
        out = _same_layout_out(x, torch.float32)
        with _span("quantize", n, c, 3):
            call("pcg_quantize", ptr(x), ptr(means), n, c, sn, sc, msn, msc, 0, ptr(out), None, stream())
    elif mode == "symbols":
        out = _same_layout_out(x, torch.int32)
        with _span("quantize", n, c, 3):
            call("pcg_quantize", ptr(x), ptr(means), n, c, sn, sc, msn, msc, 1, None, ptr(out), stream())
    else:
        raise ValueError(f'Invalid quantization mode: "{mode}"')
    return out


def dequantize(sym, means):
    _need(sym)
    n, c, sn, sc = _as_cn(sym)
    if means is not None:
        means = _expand_like(means, sym)
        _, _, msn, msc = _as_cn(means)
    else:
        msn = msc = 0
    out = _same_layout_out(sym, torch.float32)
    with _span("dequantize", n, c, 3):
        call("pcg_dequantize", ptr(sym), ptr(means), n, c, sn, sc, msn, msc, ptr(out), stream())
    return out


def _match(t, x):
    """Bring t to x's exact strides (copy only when they differ)."""
    t = _expand_like(t, x)
    if t.stride() == x.stride():
        return t
    o = torch.empty_strided(x.shape, x.stride(), dtype=t.dtype, device=x.device)
    o.copy_(t)
    return o


def gc_likelihood(x, scales, means, scale_bound, lik_bound):
    _need(x)
    n, c, sn, sc = _as_cn(x)
    scales = _match(scales, x)
    means = _match(means, x) if means is not None else None
    out = _same_layout_out(x, torch.float32)
    with _span("gc_likelihood", n, c, 4):
        call("pcg_gc_likelihood", ptr(x), ptr(scales), ptr(means), n, c, sn, sc, float(scale_bound), float(lik_bound),
             ptr(out), stream())
    return out


def gc_build_indexes(scales, table, scale_bound):
    _need(scales)
    n, c, sn, sc = _as_cn(scales)
    out = _same_layout_out(scales, torch.int32)
    with _span("gc_build_indexes", n, c, 2):
        call("pcg_gc_build_indexes", ptr(scales), n, c, sn, sc, ptr(table.contiguous()), table.numel(), float(scale_bound),
             ptr(out), stream())
    return out


def pack_quant_mlp(seq):
    """[W1(10x2) b1 W2(10x10) b2 W3(1x10) b3] from nn.Sequential(Linear, ReLU, Linear, ReLU, Linear)."""
    lin = [m for m in seq if isinstance(m, torch.nn.Linear)]
    assert [tuple(l.weight.shape) for l in lin] == [(10, 2), (10, 10), (1, 10)], "quant_nn must be 2->10->10->1"
    return torch.cat([t.detach().reshape(-1) for l in lin for t in (l.weight, l.bias)]).contiguous().float()


def gc_dequant_offsets(q, scales_hat, means_hat, gain, rescale, mlp, scale_bound):
    """Fused entropy_models.py:469-481: offsets MLP + sign/abs reconstruction + rescale + mean."""
    _need(q)
    n, c, sn, sc = _as_cn(q)
    scales_hat, means_hat = _match(scales_hat, q), _match(means_hat, q)
    gain = _expand_like(gain, q)
    rescale = _expand_like(rescale, q)
    if rescale.stride() != gain.stride():
        rescale = _match(rescale, gain.contiguous())
        gain = gain.contiguous()
    _, _, gsn, gsc = _as_cn(gain)
    out = _same_layout_out(q, torch.float32)
    with _span("gc_dequant_offsets", n, c, 4):
        call("pcg_gc_dequant_offsets", ptr(q), ptr(scales_hat), ptr(means_hat), ptr(gain), ptr(rescale), n, c, sn, sc,
             gsn, gsc, ptr(mlp), float(scale_bound), ptr(out), stream())
    return out
