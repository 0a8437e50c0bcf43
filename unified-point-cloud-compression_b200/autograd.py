"""Autograd plumbing for the training step (/root/reference/train.py:195-234): torch.autograd.Functions whose forward and
backward are libpcg kernels.  PyTorch only chains them.

  sparse conv      forward  = pcg_conv_fwd_* over the kernel map
                   d input  = the SAME forward kernels run on the inverse map with W[k]^T (a convolution's input gradient
                              is the transposed convolution and vice versa)
                   d weight = pcg_conv_bwd_weight_f32 (fixed-order fp32)
                   d bias   = column sum
  features_at      forward  = pcg_features_at; backward scatters the rows back (on-grid queries: one hit of weight 1 per query,
                   the only use on the reference's path, entropy_models.py:294 / loss.py:105)
"""
import torch

from . import sparse_ops as so
from ._lib import call, ptr, stream


def bwd_weight_tc_eligible(K, cin, cout):
    """Static rule for the tensor-core weight gradient (pcg_conv_bwd_weight_tc): every layer of the model at its configured widths
    -- the narrow occupancy-head layers and the 4-channel input layer with their rows padded to 64 channels (the fp32 kernel ran
    the input layer's 125 x 4 x 128 gradient in 18 ms on FFMA; padded 16x it is still a ~2 ms tensor-core job); the fp32 kernel
    remains the parity mode and serves the 1x1 colour head's K = 1 only through the same rule."""
    return so.get_conv_mode() != "f32" and K <= 128 and 1 <= cin <= 256 and 1 <= cout <= 256 and ((cin + 127) // 128) * ((cout + 15) // 16 * 16) <= 512


def _bf16_rows64(t):
    """fp32 (n, c) -> bf16 (n, roundup64(c)), zero padded: operand rows of the weight-gradient kernel."""
    c = t.shape[1]
    cp = (c + 63) // 64 * 64
    out = torch.empty((t.shape[0], cp), dtype=torch.bfloat16, device=t.device)
    call("pcg_f32_to_bf16_rows", ptr(t.contiguous()), t.shape[0], c, cp, ptr(out), stream())
    return out


def _act_grad(gout, out, act):
    if act == so.ACT_RELU:
        return gout * (out > 0).to(gout.dtype)
    if act == so.ACT_LEAKY:
        return gout * torch.where(out > 0, torch.ones_like(out), torch.full_like(out, 0.01))
    return gout


class _SparseConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, weight, bias, in_set, out_set, ksize, transposed, act):
        out = so.sparse_conv(feats.detach(), in_set, out_set, weight.detach(), bias.detach() if bias is not None else None,
                             ksize, transposed, act=act)
        ctx.save_for_backward(feats.detach(), weight.detach(), out if act != so.ACT_NONE else None)
        ctx.meta = (in_set, out_set, ksize, transposed, act, bias is not None)
        return out

    @staticmethod
    def backward(ctx, gout):
        feats, weight, out = ctx.saved_tensors
        in_set, out_set, ksize, transposed, act, has_bias = ctx.meta
        g = _act_grad(gout.contiguous(), out, act).contiguous()
        w3 = weight if weight.dim() == 3 else weight.unsqueeze(0)
        K, cin, cout = w3.shape
        g_feats = g_weight = g_bias = None
        if ctx.needs_input_grad[0]:
            # inverse map: rows of out_set feed rows of in_set through the mirrored offsets
            g_feats = so.sparse_conv(g, out_set, in_set, w3.transpose(1, 2).contiguous(), None, ksize, not transposed)
        if ctx.needs_input_grad[1]:
            nbr, tmask = in_set.kernel_map(out_set, ksize, transposed)
            dw = torch.empty_like(w3)
            prof = so.profiler
            if prof is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            if bwd_weight_tc_eligible(K, cin, cout) and tmask is not None and out_set.n > 0 and feats.shape[0] > 0:
                # tensor cores: bf16 operand rows of F (cached shadow of the forward pass when there is one) and of G
                fb = so.to_bf16(feats) if cin >= 33 else _bf16_rows64(feats)
                gb = _bf16_rows64(g)
                lib = so._lib.load()
                ws = so.scratch(lib.pcg_conv_bwd_weight_tc_scratch_bytes(K, cin, cout, out_set.n), g.device)
                call("pcg_conv_bwd_weight_tc", ptr(fb), cin, ptr(gb), cout, ptr(nbr), nbr.shape[1], ptr(tmask), K, out_set.n,
                     ptr(ws), ws.numel(), ptr(dw), stream())
            else:
                nb = so._lib.load().pcg_conv_bwd_weight_f32_scratch_bytes(K, cin, cout, out_set.n)
                ws = so.scratch(nb, g.device) if nb else None
                call("pcg_conv_bwd_weight_f32_sliced", ptr(feats.contiguous()), cin, ptr(nbr), nbr.shape[1], ptr(tmask), K, out_set.n, ptr(g), cout,
                     ptr(ws), nb, ptr(dw), stream())
            if prof is not None:
                e1.record()
                pairs = prof.pairs(nbr)
                prof.records.append(("conv_bwd_weight", (K, cin, cout, int(out_set.n)), 2.0 * pairs * cin * cout,
                                     2.0 * pairs * (cin + cout) + 4.0 * K * cin * cout, e0, e1))
            g_weight = dw.view_as(weight)
        if has_bias and ctx.needs_input_grad[2]:
            g_bias = g.sum(dim=0, keepdim=True)
        return g_feats, g_weight, g_bias, None, None, None, None, None


def sparse_conv_autograd(feats, in_set, out_set, weight, bias, ksize, transposed, act=so.ACT_NONE):
    return _SparseConv.apply(feats, weight, bias, in_set, out_set, ksize, transposed, act)


class _FeaturesAt(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, cset, query):
        out = so.features_at(feats.detach(), cset, query)
        ctx.cset, ctx.query, ctx.n = cset, query, feats.shape[0]
        return out

    @staticmethod
    def backward(ctx, gout):
        q = ctx.query
        qi = torch.round(q).to(torch.int32)
        if not torch.equal(qi.float(), q.float()):
            raise NotImplementedError("features_at_coordinates backward: on-grid queries only (all the reference uses)")
        rows = ctx.cset.lookup(qi).long()
        hit = rows >= 0
        g = torch.zeros((ctx.n, gout.shape[1]), dtype=gout.dtype, device=gout.device)
        g.index_add_(0, rows[hit], gout[hit])
        return g, None, None


def features_at_autograd(feats, cset, query):
    return _FeaturesAt.apply(feats, cset, query)
