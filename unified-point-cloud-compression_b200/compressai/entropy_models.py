"""compressai.entropy_models drop-in: EntropyBottleneck (CAI1) and GaussianConditional (CAI2).

Reference call sites: model/entropy_models.py:173-176,272,282-285,310-327,371-372,396-400,438,
468-484; model/model.py:30-41.  Parameters / buffers / state-dict keys follow compressai 1.2.4
(`matrices.N`, `biases.N`, `factors.N`, `quantiles`, `_offset`, `_quantized_cdf`, `_cdf_length`,
`scale_table`).  Inference arithmetic (quantise, likelihood, CDF index) runs in libpcg.so kernels on
strided views -- no `(1,C,N)` transposed copies; training (autograd) uses the differentiable torch
expression of the same formulas on the device.  `compress` / `decompress` move one int32 symbol
buffer per call across PCIe and run the host range coder (csrc/host_rans.cpp).
"""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import entropy_ops as eo
from .. import sparse_ops as so
from .. import host_coder
from .ops.bound_ops import LowerBound


def get_scale_table(min=0.11, max=256, levels=64):
    return torch.exp(torch.linspace(math.log(min), math.log(max), levels))


def _norm_ppf(q):
    """Inverse standard-normal CDF (scalar), Acklam's rational approximation + one Halley step on erfc."""
    a = [-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02, 1.383577518672690e+02,
         -3.066479806614716e+01, 2.506628277459239e+00]
    b = [-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02, 6.680131188771972e+01,
         -1.328068155288572e+01]
    c = [-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00, -2.549732539343734e+00,
         4.374664141464968e+00, 2.938163982698783e+00]
    d = [7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00, 3.754408661907416e+00]
    if q < 0.02425:
        t = math.sqrt(-2 * math.log(q))
        x = (((((c[0] * t + c[1]) * t + c[2]) * t + c[3]) * t + c[4]) * t + c[5]) / ((((d[0] * t + d[1]) * t + d[2]) * t + d[3]) * t + 1)
    elif q <= 1 - 0.02425:
        t = q - 0.5
        r = t * t
        x = (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * t / (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1)
    else:
        t = math.sqrt(-2 * math.log(1 - q))
        x = -(((((c[0] * t + c[1]) * t + c[2]) * t + c[3]) * t + c[4]) * t + c[5]) / ((((d[0] * t + d[1]) * t + d[2]) * t + d[3]) * t + 1)
    for _ in range(2):
        e = 0.5 * math.erfc(-x / math.sqrt(2)) - q
        u = e * math.sqrt(2 * math.pi) * math.exp(x * x / 2)
        x = x - u / (1 + x * u / 2)
    return x


class EntropyModel(nn.Module):
    def __init__(self, likelihood_bound=1e-9, entropy_coder=None, entropy_coder_precision=16):
        super().__init__()
        self.entropy_coder_precision = int(entropy_coder_precision)
        self.likelihood_bound = float(likelihood_bound)
        self.use_likelihood_bound = likelihood_bound > 0
        if self.use_likelihood_bound:
            self.likelihood_lower_bound = LowerBound(likelihood_bound)
        self.register_buffer("_offset", torch.IntTensor())
        self.register_buffer("_quantized_cdf", torch.IntTensor())
        self.register_buffer("_cdf_length", torch.IntTensor())
        self._host_tables = None

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        # checkpoints carry the CDF buffers at their updated sizes (compressai resizes them in
        # CompressionModel.load_state_dict / update_registered_buffers): adopt the incoming shapes
        for name in ("_offset", "_quantized_cdf", "_cdf_length", "scale_table"):
            src = state_dict.get(prefix + name)
            cur = getattr(self, name, None)
            if src is not None and cur is not None and tuple(src.shape) != tuple(cur.shape):
                setattr(self, name, torch.empty(src.shape, dtype=cur.dtype, device=cur.device))
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    # -- quantisation --------------------------------------------------------------------------
    def quantize(self, inputs, mode, means=None):
        if mode not in ("noise", "dequantize", "symbols"):
            raise ValueError(f'Invalid quantization mode: "{mode}"')
        if mode == "noise":
            return inputs + torch.empty_like(inputs).uniform_(-0.5, 0.5)
        if inputs.requires_grad and torch.is_grad_enabled():
            out = torch.round(inputs - means) if means is not None else torch.round(inputs)
            if mode == "dequantize":
                return out + means if means is not None else out
            return out.int()
        return eo.quantize(inputs.detach(), None if means is None else means.detach(), mode)

    @staticmethod
    def dequantize(inputs, means=None, dtype=torch.float):
        if inputs.dtype == torch.int32 and inputs.is_cuda:
            return eo.dequantize(inputs, means)
        if means is not None:
            return inputs.type_as(means) + means
        return inputs.type(dtype)

    # -- CDF tables ----------------------------------------------------------------------------
    def _pmf_to_cdf(self, pmf, tail_mass, pmf_length, max_length):
        cdf = torch.zeros((len(pmf_length), max_length + 2), dtype=torch.int32)
        pmf, tail_mass = pmf.detach().cpu(), tail_mass.detach().cpu()
        for i, p in enumerate(pmf):
            prob = torch.cat((p[: pmf_length[i]], tail_mass[i]), dim=0)
            _cdf = torch.from_numpy(host_coder.pmf_to_quantized_cdf(prob.numpy(), self.entropy_coder_precision))
            cdf[i, : _cdf.size(0)] = _cdf
        return cdf

    def _check_cdf_size(self):
        if self._quantized_cdf.numel() == 0:
            raise ValueError("Uninitialized CDFs. Run update() first")
        if len(self._quantized_cdf.size()) != 2:
            raise ValueError(f"Invalid CDF size {self._quantized_cdf.size()}")

    def _check_offsets_size(self):
        if self._offset.numel() == 0:
            raise ValueError("Uninitialized offsets. Run update() first")

    def _check_cdf_length(self):
        if self._cdf_length.numel() == 0:
            raise ValueError("Uninitialized CDF lengths. Run update() first")

    def _tables(self):
        self._check_cdf_size(); self._check_cdf_length(); self._check_offsets_size()
        key = (self._quantized_cdf.data_ptr(), self._quantized_cdf._version)
        hit = self._host_tables                      # one tuple, published whole (several frames may run concurrently)
        if hit is None or hit[0] != key:
            hit = (key, self._quantized_cdf.cpu().numpy().astype(np.int32),
                   self._cdf_length.reshape(-1).cpu().numpy().astype(np.int32),
                   self._offset.reshape(-1).cpu().numpy().astype(np.int32))
            self._host_tables = hit
        return hit[1:]

    # -- host range coder ----------------------------------------------------------------------
    def compress(self, inputs, indexes, means=None):
        symbols = self.quantize(inputs, "symbols", means)
        if len(inputs.size()) < 2:
            raise ValueError("Invalid `inputs` size. Expected a tensor with at least 2 dimensions.")
        if inputs.size() != indexes.size():
            raise ValueError("`inputs` and `indexes` should have the same size.")
        cdf, lengths, offsets = self._tables()
        sym_h = symbols.reshape(symbols.size(0), -1).cpu().numpy()
        idx_h = indexes.reshape(indexes.size(0), -1).int().cpu().numpy()
        return [host_coder.encode(sym_h[i], idx_h[i], cdf, lengths, offsets) for i in range(sym_h.shape[0])]

    def decompress(self, strings, indexes, dtype=torch.float, means=None):
        if not isinstance(strings, (tuple, list)):
            raise ValueError("Invalid `strings` parameter type.")
        if not len(strings) == indexes.size(0):
            raise ValueError("Invalid strings or indexes parameters")
        if means is not None and means.size()[:2] != indexes.size()[:2]:
            raise ValueError("Invalid means or indexes parameters")
        cdf, lengths, offsets = self._tables()
        idx_h = indexes.reshape(indexes.size(0), -1).int().cpu().numpy()
        vals = np.stack([host_coder.decode(s, idx_h[i], cdf, lengths, offsets) for i, s in enumerate(strings)])
        sym = torch.from_numpy(vals).to(indexes.device).reshape(indexes.size())
        return self.dequantize(sym, means, dtype)


class EntropyBottleneck(EntropyModel):
    """Factorised prior; likelihood chain specialised (in the kernel) for the default filters (3,3,3,3)."""

    def __init__(self, channels, *args, tail_mass=1e-9, init_scale=10, filters=(3, 3, 3, 3), **kwargs):
        super().__init__(*args, **kwargs)
        self.channels = int(channels)
        self.filters = tuple(int(f) for f in filters)
        self.init_scale, self.tail_mass = float(init_scale), float(tail_mass)
        filters = (1,) + self.filters + (1,)
        scale = self.init_scale ** (1 / (len(self.filters) + 1))
        self.matrices, self.biases, self.factors = nn.ParameterList(), nn.ParameterList(), nn.ParameterList()
        for i in range(len(self.filters) + 1):
            init = np.log(np.expm1(1 / scale / filters[i + 1]))
            self.matrices.append(nn.Parameter(torch.full((channels, filters[i + 1], filters[i]), float(init))))
            self.biases.append(nn.Parameter(torch.empty(channels, filters[i + 1], 1).uniform_(-0.5, 0.5)))
            if i < len(self.filters):
                self.factors.append(nn.Parameter(torch.zeros(channels, filters[i + 1], 1)))
        self.quantiles = nn.Parameter(torch.Tensor([-self.init_scale, 0, self.init_scale]).repeat(channels, 1, 1))
        target = np.log(2 / self.tail_mass - 1)
        self.register_buffer("target", torch.Tensor([-target, 0, target]))
        self._packed = None

    def _get_medians(self):
        return self.quantiles[:, :, 1:2]

    def _packed_params(self):
        key = tuple((p.data_ptr(), p._version) for p in list(self.matrices) + list(self.biases) + list(self.factors))
        hit = self._packed
        if hit is None or hit[0] != key:
            with so.pack_lock:                       # packed on the current stream: finished before any other stream sees it
                hit = (key, eo.eb_pack_params(list(self.matrices), list(self.biases), list(self.factors)))
                if hit[1].is_cuda:
                    torch.cuda.current_stream().synchronize()
                self._packed = hit
        return hit[1]

    def update(self, force=False, update_quantiles=False):
        """Build the 16-bit CDF tables.  Always evaluated on the host in fp32, whatever device the parameters live on:
        the tables are part of the bitstream contract, and ulp-level differences between a GPU and a CPU evaluation of
        the pmf would otherwise give encoder and decoder different tables."""
        if self._offset.numel() > 0 and not force:
            return False
        dev = self.quantiles.device
        quantiles = self.quantiles.detach().cpu()
        medians = quantiles[:, 0, 1]
        minima = torch.clamp(torch.ceil(medians - quantiles[:, 0, 0]).int(), min=0)
        maxima = torch.clamp(torch.ceil(quantiles[:, 0, 2] - medians).int(), min=0)
        pmf_start = medians - minima
        pmf_length = maxima + minima + 1
        max_length = int(pmf_length.max().item())
        samples = torch.arange(max_length)[None, :] + pmf_start[:, None, None]
        host = [[p.detach().cpu() for p in plist] for plist in (self.matrices, self.biases, self.factors)]
        pmf, lower, upper = self._likelihood(samples, stop_gradient=True, params=host)
        pmf = pmf[:, 0, :]
        tail_mass = torch.sigmoid(lower[:, 0, :1]) + torch.sigmoid(-upper[:, 0, -1:])
        self._offset = (-minima).to(dev)
        self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length).to(dev)
        self._cdf_length = (pmf_length + 2).to(dev)
        return True

    def loss(self):
        logits = self._logits_cumulative(self.quantiles, stop_gradient=True)
        return torch.abs(logits - self.target).sum()

    def _logits_cumulative(self, inputs, stop_gradient, params=None):
        matrices, biases, factors = params if params is not None else (self.matrices, self.biases, self.factors)
        logits = inputs
        for i in range(len(self.filters) + 1):
            matrix = matrices[i].detach() if stop_gradient else matrices[i]
            bias = biases[i].detach() if stop_gradient else biases[i]
            logits = torch.matmul(F.softplus(matrix), logits) + bias
            if i < len(self.filters):
                factor = factors[i].detach() if stop_gradient else factors[i]
                logits = logits + torch.tanh(factor) * torch.tanh(logits)
        return logits

    def _likelihood(self, inputs, stop_gradient=False, params=None):
        lower = self._logits_cumulative(inputs - 0.5, stop_gradient, params)
        upper = self._logits_cumulative(inputs + 0.5, stop_gradient, params)
        return torch.sigmoid(upper) - torch.sigmoid(lower), lower, upper

    def _needs_grad(self, x):
        return torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.matrices)) and self.training

    def forward(self, x, training=None):
        if training is None:
            training = self.training
        if x.dim() != 3:
            raise ValueError("EntropyBottleneck(pcg) expects (B, C, N) inputs")
        if self._needs_grad(x) or self.filters != (3, 3, 3, 3) or not x.is_cuda or x.shape[0] != 1:
            # differentiable device path (training): same formulas, torch autograd
            xp = x.permute(1, 0, 2).contiguous()
            shape = xp.size()
            values = xp.reshape(xp.size(0), 1, -1)
            med = self._get_medians()
            outputs = values + torch.empty_like(values).uniform_(-0.5, 0.5) if training else torch.round(values - med) + med
            likelihood, _, _ = self._likelihood(outputs)
            if self.use_likelihood_bound:
                likelihood = self.likelihood_lower_bound(likelihood)
            return (outputs.reshape(shape).permute(1, 0, 2).contiguous(),
                    likelihood.reshape(shape).permute(1, 0, 2).contiguous())
        med = self._get_medians().detach().reshape(1, -1, 1)
        outputs = self.quantize(x, "noise" if training else "dequantize", med)
        likelihood = eo.eb_likelihood(outputs, self._packed_params(), self.likelihood_bound if self.use_likelihood_bound else 0.0)
        return outputs, likelihood

    @staticmethod
    def _build_indexes(size):
        N, C = size[0], size[1]
        view = [1] * len(size)
        view[1] = -1
        return torch.arange(C).view(*view).int().repeat(N, 1, *size[2:])

    def _medians_like(self, n_batch, spatial_dims):
        med = self._get_medians().detach().reshape(-1, *([1] * spatial_dims))
        return med.expand(n_batch, *([-1] * (spatial_dims + 1)))

    def compress(self, x):
        indexes = self._build_indexes(x.size()).to(x.device)
        return super().compress(x, indexes, self._medians_like(x.size(0), x.dim() - 2))

    def decompress(self, strings, size):
        output_size = (len(strings), self._quantized_cdf.size(0), *size)
        indexes = self._build_indexes(output_size).to(self._quantized_cdf.device)
        medians = self._medians_like(len(strings), len(size))
        return super().decompress(strings, indexes, medians.dtype, medians)


class GaussianConditional(EntropyModel):
    def __init__(self, scale_table, *args, scale_bound=0.11, tail_mass=1e-9, **kwargs):
        super().__init__(*args, **kwargs)
        if not isinstance(scale_table, (type(None), list, tuple)):
            raise ValueError(f'Invalid type for scale_table "{type(scale_table)}"')
        if isinstance(scale_table, (list, tuple)) and len(scale_table) < 1:
            raise ValueError(f'Invalid scale_table length "{len(scale_table)}"')
        if scale_table and (scale_table != sorted(scale_table) or any(s <= 0 for s in scale_table)):
            raise ValueError(f'Invalid scale_table "({scale_table})"')
        self.tail_mass = float(tail_mass)
        if scale_bound is None and scale_table:
            scale_bound = scale_table[0]
        if scale_bound <= 0:
            raise ValueError("Invalid parameters")
        self.scale_bound_value = float(scale_bound)
        self.lower_bound_scale = LowerBound(scale_bound)
        self.register_buffer("scale_table", self._prepare_scale_table(scale_table) if scale_table else torch.Tensor())
        self.register_buffer("scale_bound", torch.Tensor([float(scale_bound)]))

    @staticmethod
    def _prepare_scale_table(scale_table):
        return torch.Tensor(tuple(float(s) for s in scale_table))

    def _standardized_cumulative(self, inputs):
        return 0.5 * torch.erfc(float(-(2 ** -0.5)) * inputs)

    @staticmethod
    def _standardized_quantile(quantile):
        return _norm_ppf(quantile)

    def update_scale_table(self, scale_table, force=False):
        if self._offset.numel() > 0 and not force:
            return False
        device = self.scale_table.device
        self.scale_table = self._prepare_scale_table(scale_table).to(device)
        self.update()
        return True

    def update(self):
        """16-bit CDF tables per scale-table entry; evaluated on the host (see EntropyBottleneck.update)."""
        dev = self.scale_table.device
        table = self.scale_table.detach().cpu()
        multiplier = -self._standardized_quantile(self.tail_mass / 2)
        pmf_center = torch.ceil(table * multiplier).int()
        pmf_length = 2 * pmf_center + 1
        max_length = int(torch.max(pmf_length).item())
        samples = torch.abs(torch.arange(max_length).int() - pmf_center[:, None]).float()
        samples_scale = table.unsqueeze(1).float()
        upper = self._standardized_cumulative((0.5 - samples) / samples_scale)
        lower = self._standardized_cumulative((-0.5 - samples) / samples_scale)
        pmf = upper - lower
        tail_mass = 2 * lower[:, :1]
        self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length).to(dev)
        self._offset = (-pmf_center).to(dev)
        self._cdf_length = (pmf_length + 2).to(dev)

    def _likelihood(self, inputs, scales, means=None):
        values = inputs - means if means is not None else inputs
        scales = self.lower_bound_scale(scales)
        values = torch.abs(values)
        upper = self._standardized_cumulative((0.5 - values) / scales)
        lower = self._standardized_cumulative((-0.5 - values) / scales)
        return upper - lower

    def forward(self, inputs, scales, means=None, training=None):
        if training is None:
            training = self.training
        grad = torch.is_grad_enabled() and (inputs.requires_grad or scales.requires_grad or (means is not None and means.requires_grad))
        if grad or not inputs.is_cuda:
            if training:
                outputs = inputs + torch.empty_like(inputs).uniform_(-0.5, 0.5)
            else:
                outputs = torch.round(inputs - means) + means if means is not None else torch.round(inputs)
            likelihood = self._likelihood(outputs, scales, means)
            if self.use_likelihood_bound:
                likelihood = self.likelihood_lower_bound(likelihood)
            return outputs, likelihood
        outputs = self.quantize(inputs, "noise" if training else "dequantize", means)
        likelihood = eo.gc_likelihood(outputs, scales, means, self.scale_bound_value,
                                      self.likelihood_bound if self.use_likelihood_bound else 0.0)
        return outputs, likelihood

    def build_indexes(self, scales):
        if self.scale_table.numel() == 0:
            raise ValueError("Uninitialized scale table. Run update_scale_table() first")
        return eo.gc_build_indexes(scales.detach(), self.scale_table.to(scales.device), self.scale_bound_value)
