"""MeanScaleHyperprior -- mirror of /root/reference/model/entropy_models.py:128-490.

Same sub-module / parameter names (h_a, h_s, scale_nn, rescale_nn, quant_nn, entropy_bottleneck,
gaussian_conditional).  What changes is the execution:
  * the `Sorted*` wrappers (entropy_models.py:28-126) exist upstream only to make ME's atomics-based
    GPU kernels reproducible (2 argsorts + 2 hash rebuilds per layer).  libpcg's convolutions accumulate
    every output row in a fixed order, so the wrapped classes are the plain modules here; encoder and
    decoder agree bit-for-bit on h_s(z_hat) because they run the same kernels on the same row order
    (both sides lexsort z first, utils.py:142-187);
  * LeakyReLUs are fused into the preceding convolution's epilogue (inference path);
  * `build_indexes` is one kernel instead of 63 compare passes; offsets MLP + sign/abs reconstruction +
    rescale + mean (entropy_models.py:218-233,469-481) is one kernel;
  * all `(1,C,N)` operands are strided views of `(N,C)` matrices -- no transposed copies.
"""
import torch
import torch.nn as nn

from .. import MinkowskiEngine as ME
from .. import entropy_ops as eo
from .. import sparse_ops as so
from ..compressai.entropy_models import EntropyBottleneck, GaussianConditional
from ..compressai.models.base import CompressionModel
from ..compressai.ops import LowerBound, ops


def quantize_noise(x):
    """x + U(-0.5, 0.5) (entropy_models.py:11-25)."""
    return x + torch.empty_like(x).uniform_(-0.5, 0.5)


class SortedMinkowskiConvolution(ME.MinkowskiConvolution):
    """Deterministic by construction in libpcg; kept as a named class for drop-in parity."""


class SortedMinkowskiGenerativeConvolutionTranspose(ME.MinkowskiGenerativeConvolutionTranspose):
    pass


class SortedMinkowskiLeakyReLU(ME.MinkowskiLeakyReLU):
    pass


def sort_tensor(x):
    """Rows in lexicographic (b,x,y,z) order -- defines the symbol order of the bitstream (utils.py:142-165)."""
    if x.F.is_cuda and x.F.dtype == torch.float32 and not (torch.is_grad_enabled() and x.F.requires_grad):
        oc, of = so.gather_rows(x.C, x.F, so.sort_rows_i32(x.C, morton=False))       # coordinates and features in one pass
        return ME.SparseTensor(of, _coord_set=so.CoordSet(oc, x.coord_set.ts))
    perm = so.lexsort(x.C)
    return ME.SparseTensor(x.F[perm], _coord_set=so.CoordSet(x.C[perm], x.coord_set.ts))


def sort_points(points):
    if points.is_cuda and points.dtype == torch.int32:
        return so.gather_rows(points, None, so.sort_rows_i32(points, morton=False))[0]
    return points[so.lexsort(points.to(torch.int32))]


def _mlp(widths, last=None):
    layers = []
    for i in range(len(widths) - 1):
        layers.append(nn.Linear(widths[i], widths[i + 1]))
        if i < len(widths) - 2:
            layers.append(nn.ReLU())
    if last is not None:
        layers.append(last)
    return nn.Sequential(*layers)


def _run_fused(seq, x):
    """Run conv(+LeakyReLU) pairs of an nn.Sequential with the activation folded into the conv epilogue."""
    mods = list(seq)
    i = 0
    while i < len(mods):
        m = mods[i]
        nxt = mods[i + 1] if i + 1 < len(mods) else None
        if isinstance(m, ME.MinkowskiConvolutionBase) and isinstance(nxt, ME.MinkowskiLeakyReLU):
            x = m(x, _act=so.ACT_LEAKY)          # per-call argument: no module state is touched (frames may run concurrently)
            i += 2
        else:
            x = m(x)
            i += 1
    return x


class MeanScaleHyperprior(CompressionModel):
    def __init__(self, config):
        super().__init__()
        C, Ch = config["C_bottleneck"], config["C_hyper_bottleneck"]
        self.inverse_rescaling = config["inverse_rescaling"]
        self.quantization_mode = config["quantization_mode"]
        self.entropy_bottleneck_vbr = config["entropy_bottleneck_vbr"]
        self.adaptive_BN = config.get("adaptive_BN", True)
        self.eps = 0.0001
        self.quantization_offset = config["quantization_offset"]
        if self.entropy_bottleneck_vbr:
            raise NotImplementedError("EntropyBottleneckVbr is commented out upstream (entropy_models.py:8); every config sets it False")
        self.gaussian_conditional = GaussianConditional(None)
        self.entropy_bottleneck = EntropyBottleneck(Ch)

        conv = lambda i, o, s=1, b=False: ME.MinkowskiConvolution(in_channels=i, out_channels=o, kernel_size=3, stride=s, bias=b, dimension=3)
        self.h_a = nn.Sequential(conv(C, Ch), ME.MinkowskiLeakyReLU(inplace=False), conv(Ch, Ch, 2),
                                 ME.MinkowskiLeakyReLU(inplace=False), conv(Ch, Ch, 2))
        gen = lambda i, o: SortedMinkowskiGenerativeConvolutionTranspose(in_channels=i, out_channels=o, kernel_size=2, stride=2, bias=True, dimension=3)
        self.h_s = nn.Sequential(gen(Ch, Ch), SortedMinkowskiLeakyReLU(inplace=False), gen(Ch, C * 3 // 2),
                                 SortedMinkowskiLeakyReLU(inplace=False),
                                 SortedMinkowskiConvolution(in_channels=C * 3 // 2, out_channels=C * 2, kernel_size=3, stride=1, dimension=3, bias=True))
        self.scale_nn = _mlp([2, 8, C // 4, C], nn.Softplus())
        self.rescale_nn = _mlp([2, 8, C // 4, C], nn.Softplus())
        self.quant_nn = _mlp([2, 10, 10, 1])

    def prepare(self):
        """Fill the lazily built caches of the entropy models (packed prior parameters, host CDF tables)."""
        if self.entropy_bottleneck._quantized_cdf.numel():
            self.entropy_bottleneck._tables()
            self.gaussian_conditional._tables()
        if next(self.entropy_bottleneck.parameters()).is_cuda:
            self.entropy_bottleneck._packed_params()

    # -- pieces ------------------------------------------------------------------------------
    def get_offsets(self, stddev, scale):
        """Per-element MLP 2->10->10->1 on (scale, stddev) (entropy_models.py:218-233); torch path for training."""
        return self.quant_nn(torch.stack((scale.expand_as(stddev), stddev), dim=3)).squeeze(dim=3)

    def _gains(self, q, batch_idx, n_rows, device, need_rescale):
        """(scale, rescale) as (1,C,N) views, gathered per row by batch index (a9).  batch_idx: tensor, or a callable that
        produces it (only evaluated when several quality settings share the tensor)."""
        C = self.scale_nn[-2].out_features
        if not self.adaptive_BN:
            one = torch.ones((1, C, n_rows), device=device)
            return one, one
        if not torch.is_grad_enabled():
            # inference: the gain vectors depend on (q, MLP weights) only -- computed once per q tensor / weight version instead of
            # ~20 tiny launches per frame; one quality setting for the whole tensor (q has one row) needs no per-row gather either
            # (keyed on the q tensor OBJECT, which the cache keeps alive: an address alone could be recycled for another q)
            key = (q._version, tuple(p._version for p in self.scale_nn.parameters()), tuple(p._version for p in self.rescale_nn.parameters()))
            hit = getattr(self, "_gain_cache", None)
            if hit is None or hit[3] is not q or hit[0] != key:
                scale_b = self.scale_nn(q) + self.eps
                hit = (key, scale_b, 1.0 / scale_b if self.inverse_rescaling else 1.0 / self.rescale_nn(q), q)
                self._gain_cache = hit
            scale_b, rescale_b = hit[1], hit[2]
            if q.shape[0] == 1:
                view = lambda t: t[0].view(1, C, 1).expand(1, C, n_rows)
            else:
                bi = batch_idx() if callable(batch_idx) else batch_idx
                view = lambda t: t[bi].t().unsqueeze(0)
            return view(scale_b), (view(rescale_b) if need_rescale else None)
        if callable(batch_idx):
            batch_idx = batch_idx()
        scale_b = self.scale_nn(q) + self.eps                       # (B, C)
        scale = scale_b[batch_idx].t().unsqueeze(0)                 # view (1, C, N)
        rescale = None
        if need_rescale:
            if self.inverse_rescaling:
                rescale = (1.0 / scale_b.detach())[batch_idx].t().unsqueeze(0)
            else:
                rescale = (1.0 / self.rescale_nn(q))[batch_idx].t().unsqueeze(0)
        return scale, rescale

    def _hyper_synthesis(self, z_hat, y_coords):
        """h_s(z_hat) sampled at the latent coordinates -> (scales_hat, means_hat) as (1,C,N) views."""
        params = _run_fused(self.h_s, z_hat) if not (torch.is_grad_enabled() and z_hat.F.requires_grad) else self.h_s(z_hat)
        feats = params.features_at_coordinates(y_coords.float())     # (N, 2C)
        C = feats.shape[1] // 2
        return feats[:, :C].t().unsqueeze(0), feats[:, C:].t().unsqueeze(0)

    def _h_a(self, y):
        return _run_fused(self.h_a, y) if not (torch.is_grad_enabled() and y.F.requires_grad) else self.h_a(y)

    # -- training forward (entropy_models.py:236-340) -------------------------------------------
    def forward(self, y, q):
        z = self._h_a(y)
        z_feats = z.F.t().unsqueeze(0)
        y_feats = y.F.t().unsqueeze(0)
        batch_idx = y.C[:, 0].long()
        scale, rescale = self._gains(q, batch_idx, y.F.shape[0], y.device, True)

        if self.quantization_mode == "uniform":
            z_hat, z_likelihoods = self.entropy_bottleneck(z_feats)
        else:
            _, z_likelihoods = self.entropy_bottleneck(z_feats)
            # upstream broadcasts (1,C,N) - (C,1,1) to (C,C,N) and keeps [0], i.e. channel 0's median for
            # every channel (entropy_models.py:283-287); reproduced as such
            z_offset = self.entropy_bottleneck._get_medians()[0:1]
            z_hat = ops.quantize_ste(z_feats - z_offset) + z_offset
        z_hat = ME.SparseTensor(z_hat[0].t(), _coord_set=z.coord_set)

        scales_hat, means_hat = self._hyper_synthesis(z_hat, y.C)
        if self.quantization_offset:
            y_tmp = scale * (y_feats - means_hat)
            signs = torch.sign(y_tmp).detach()
            y_q_abs = quantize_noise(torch.abs(y_tmp)) if self.quantization_mode == "uniform" else ops.quantize_ste(torch.abs(y_tmp))
            _, y_likelihoods = self.gaussian_conditional(y_feats * scale, scales_hat * scale, means=means_hat * scale)
            y_q_stdev = self.gaussian_conditional.lower_bound_scale(scales_hat * scale)
            q_offsets = -self.get_offsets(y_q_stdev, scale.clone().detach())
            q_offsets = torch.where(y_q_abs < 0.0001, torch.zeros_like(q_offsets), q_offsets)
            y_hat = signs * (y_q_abs + q_offsets) * rescale + means_hat
        else:
            y_hat, y_likelihoods = self.gaussian_conditional(y_feats * scale, scales_hat * scale, means=means_hat * scale)
            y_hat = y_hat * rescale
        y_hat = ME.SparseTensor(y_hat[0].t(), _coord_set=y.coord_set)
        return y_hat, (y_likelihoods, z_likelihoods)

    # -- compress (entropy_models.py:344-406) ----------------------------------------------------
    def encode_symbols(self, y, q):
        """GPU half of `compress`: everything up to (and excluding) the host range coder.

        Returns dict(y_points, z_points, z_symbols (1,Ch,Nz) int32, y_symbols (1,C,Ny) int32,
        y_indexes (1,C,Ny) int32) -- device tensors in bitstream (channel-major) order.
        """
        z = sort_tensor(self._h_a(y))
        y = sort_tensor(y)
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        z_feats = z.F.t().unsqueeze(0)
        med = eb._get_medians().detach().reshape(1, -1, 1)
        z_sym = eb.quantize(z_feats, "symbols", med)
        z_hat = ME.SparseTensor(eb.dequantize(z_sym, med.expand_as(z_feats))[0].t(), _coord_set=z.coord_set)
        scales_hat, means_hat = self._hyper_synthesis(z_hat, y.C)
        scale, _ = self._gains(q, lambda: y.C[:, 0].long(), y.F.shape[0], y.device, False)
        indexes = gc.build_indexes(scales_hat * scale)
        y_sym = gc.quantize(y.F.t().unsqueeze(0) * scale, "symbols", means_hat * scale)
        return {"y_points": y.C, "z_points": z.C, "z_symbols": z_sym, "y_symbols": y_sym, "y_indexes": indexes}

    def compress(self, y, q):
        enc = self.encode_symbols(y, q)
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        z_idx = eb._build_indexes(enc["z_symbols"].size()).to(y.device)
        z_strings = self._code(eb, enc["z_symbols"], z_idx)
        y_strings = self._code(gc, enc["y_symbols"], enc["y_indexes"])
        return [enc["y_points"], enc["z_points"]], [y_strings, z_strings], [enc["z_points"].shape[0]]

    @staticmethod
    def _code(model, symbols, indexes):
        from .. import host_coder
        cdf, lengths, offsets = model._tables()
        s = symbols.reshape(symbols.size(0), -1).cpu().numpy()
        i = indexes.reshape(indexes.size(0), -1).int().cpu().numpy()
        return [host_coder.encode(s[b], i[b], cdf, lengths, offsets) for b in range(s.shape[0])]

    # -- decompress (entropy_models.py:409-490) --------------------------------------------------
    def decode_symbols(self, y_points, z_points, z_symbols, y_symbol_source, q):
        """GPU half of `decompress`.  `y_symbol_source(indexes)` returns the (1,C,Ny) int32 symbols for the
        given CDF indexes (host range decoder, or a device tensor when the coder is bypassed)."""
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        y_points = sort_points(y_points)
        z_points = sort_points(z_points)
        med = eb._get_medians().detach().reshape(1, -1, 1)
        z_hat_f = eb.dequantize(z_symbols, med.expand(z_symbols.shape))
        z_hat = ME.SparseTensor(z_hat_f[0].t(), z_points.to(torch.int32), tensor_stride=32, device=z_points.device)
        scales_hat, means_hat = self._hyper_synthesis(z_hat, y_points)
        scale, rescale = self._gains(q, lambda: y_points[:, 0].long(), y_points.shape[0], y_points.device, True)
        indexes = gc.build_indexes(scales_hat * scale)
        sym = y_symbol_source(indexes)
        if self.quantization_offset:
            mlp = eo.pack_quant_mlp(self.quant_nn)
            y_hat = eo.gc_dequant_offsets(sym.float(), scales_hat, means_hat, scale.detach(), rescale.detach(), mlp,
                                          gc.scale_bound_value)
        else:
            y_hat = gc.dequantize(sym, means_hat * scale)
        return ME.SparseTensor(y_hat[0].t(), y_points.to(torch.int32), tensor_stride=8, device=y_points.device)

    def decompress(self, points, strings, shape, q):
        assert isinstance(strings, list) and len(strings) == 2
        assert isinstance(points, list) and len(points) == 2
        from .. import host_coder
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        y_strings, z_strings = strings
        dev = points[0].device
        z_size = (len(z_strings), eb._quantized_cdf.size(0), *shape)
        z_idx = eb._build_indexes(z_size)
        cdf, lengths, offsets = eb._tables()
        z_sym = torch.stack([torch.from_numpy(host_coder.decode(s, z_idx[b].reshape(-1).numpy(), cdf, lengths, offsets)).reshape(z_size[1:])
                             for b, s in enumerate(z_strings)]).to(dev)

        def y_source(indexes):
            cdf_y, len_y, off_y = gc._tables()
            idx = indexes.reshape(indexes.size(0), -1).cpu().numpy()
            vals = [torch.from_numpy(host_coder.decode(s, idx[b], cdf_y, len_y, off_y)) for b, s in enumerate(y_strings)]
            return torch.stack(vals).reshape(indexes.shape).to(dev)

        return self.decode_symbols(points[0], points[1], z_sym, y_source, q)
