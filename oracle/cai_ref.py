"""CPU restatement of the compressai surface the reference calls.  TEST INFRASTRUCTURE.

Un-vendored dependency restated: compressai==1.2.4 (/root/reference/requirements.txt:9).
Recalled from upstream `entropy_models/entropy_models.py`, `ops/{ops,bound_ops,
parametrizers}.py`, `layers/gdn.py`, `models/base.py`, `cpp_exts/ops/ops.cpp`,
`cpp_exts/rans/rans_interface.cpp` -- PARITY UNPINNED (oracle/__init__.py); independent
pins (scipy.stats.norm, scipy.special.expit) in tests/test_oracle_entropy.py.

Call sites in the reference:
  model/entropy_models.py:6-8,173-176,272,282-285,310-327,371-372,396-400,438,468-484
  model/blocks.py:5,22,40-41     model/model.py:7,15,30-41

The range coder itself (`ans` C++ extension) is the reference's host coder and outside
the graft; `oracle/rans_ref.py` restates it for round-trip tests.  Here `compress` returns
and `decompress` accepts byte strings produced by that restatement.
"""
import math

import numpy as np
import scipy.stats
import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------------------------------------------------------- ops
class _LowerBoundFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bound):
        ctx.save_for_backward(x, bound)
        return torch.max(x, bound)

    @staticmethod
    def backward(ctx, g):
        x, bound = ctx.saved_tensors
        return ((x >= bound) | (g < 0)) * g, None


class LowerBound(nn.Module):
    """ops/bound_ops.py: max(x, bound) with the pass-through gradient rule."""

    def __init__(self, bound):
        super().__init__()
        self.register_buffer("bound", torch.Tensor([float(bound)]))

    def forward(self, x):
        return _LowerBoundFn.apply(x, self.bound)


def quantize_ste(x):
    """ops/ops.py: (round(x) - x).detach() + x."""
    return (torch.round(x) - x).detach() + x


class NonNegativeParametrizer(nn.Module):
    """ops/parametrizers.py."""

    def __init__(self, minimum=0, reparam_offset=2 ** -18):
        super().__init__()
        self.minimum, self.reparam_offset = float(minimum), float(reparam_offset)
        pedestal = self.reparam_offset ** 2
        self.register_buffer("pedestal", torch.Tensor([pedestal]))
        self.lower_bound = LowerBound((self.minimum + self.reparam_offset ** 2) ** 0.5)

    def init(self, x):
        return torch.sqrt(torch.max(x + self.pedestal, self.pedestal))

    def forward(self, x):
        return self.lower_bound(x) ** 2 - self.pedestal


class GDN(nn.Module):
    """layers/gdn.py (parameters + reparam; the reference overrides forward, blocks.py:26-57)."""

    def __init__(self, in_channels, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        self.inverse = bool(inverse)
        self.beta_reparam = NonNegativeParametrizer(minimum=float(beta_min))
        self.beta = nn.Parameter(self.beta_reparam.init(torch.ones(in_channels)))
        self.gamma_reparam = NonNegativeParametrizer()
        self.gamma = nn.Parameter(self.gamma_reparam.init(float(gamma_init) * torch.eye(in_channels)))

    def forward(self, x):
        _, C, _, _ = x.size()
        beta = self.beta_reparam(self.beta)
        gamma = self.gamma_reparam(self.gamma).reshape(C, C, 1, 1)
        norm = F.conv2d(x ** 2, gamma, beta)
        norm = torch.sqrt(norm) if self.inverse else torch.rsqrt(norm)
        return x * norm


# ----------------------------------------------------------------------------- CDF tables
def pmf_to_quantized_cdf(pmf, precision=16):
    """cpp_exts/ops/ops.cpp::pmf_to_quantized_cdf, including the frequency-stealing rule."""
    pmf = np.asarray(pmf, dtype=np.float32)
    if (pmf < 0).any() or not np.isfinite(pmf).all():
        raise ValueError("Invalid `pmf`, non-finite or negative element found")
    cdf = np.zeros(len(pmf) + 1, dtype=np.uint64)
    # std::round on float: half away from zero
    # (the fp32 product widened to double before the +0.5: in fp32 the addition itself rounds, 0.49999997 + 0.5 -> 1.0)
    cdf[1:] = np.floor((pmf.astype(np.float32) * np.float32(1 << precision)).astype(np.float64) + 0.5).astype(np.uint64)
    total = int(cdf.sum())
    if total == 0:
        raise ValueError("Invalid `pmf`: at least one element must have a non-zero probability.")
    cdf = ((np.uint64(1 << precision) * cdf) // np.uint64(total)).astype(np.int64)
    cdf = np.cumsum(cdf)
    cdf[-1] = 1 << precision
    n = len(cdf)
    for i in range(n - 1):
        if cdf[i] == cdf[i + 1]:
            freq = cdf[1:] - cdf[:-1]
            cand = np.where(freq > 1, freq, np.iinfo(np.int64).max)
            best = int(np.argmin(cand))            # first minimal frequency > 1
            assert cand[best] != np.iinfo(np.int64).max
            if best < i:
                cdf[best + 1:i + 1] -= 1
            else:
                assert best > i
                cdf[i + 1:best + 1] += 1
    return cdf.astype(np.int32)


def get_scale_table(min=0.11, max=256, levels=64):
    return torch.exp(torch.linspace(math.log(min), math.log(max), levels))


# ----------------------------------------------------------------------------- entropy models
class EntropyModel(nn.Module):
    def __init__(self, likelihood_bound=1e-9, entropy_coder=None, entropy_coder_precision=16):
        super().__init__()
        self.entropy_coder_precision = int(entropy_coder_precision)
        self.use_likelihood_bound = likelihood_bound > 0
        if self.use_likelihood_bound:
            self.likelihood_lower_bound = LowerBound(likelihood_bound)
        self.register_buffer("_offset", torch.IntTensor())
        self.register_buffer("_quantized_cdf", torch.IntTensor())
        self.register_buffer("_cdf_length", torch.IntTensor())

    def quantize(self, inputs, mode, means=None):
        if mode not in ("noise", "dequantize", "symbols"):
            raise ValueError(f'Invalid quantization mode: "{mode}"')
        if mode == "noise":
            return inputs + torch.empty_like(inputs).uniform_(-0.5, 0.5)
        outputs = inputs.clone()
        if means is not None:
            outputs -= means
        outputs = torch.round(outputs)
        if mode == "dequantize":
            if means is not None:
                outputs += means
            return outputs
        return outputs.int()

    @staticmethod
    def dequantize(inputs, means=None, dtype=torch.float):
        if means is not None:
            outputs = inputs.type_as(means)
            outputs = outputs + means
        else:
            outputs = inputs.type(dtype)
        return outputs

    def _pmf_to_cdf(self, pmf, tail_mass, pmf_length, max_length):
        cdf = torch.zeros((len(pmf_length), max_length + 2), dtype=torch.int32)
        for i, p in enumerate(pmf):
            prob = torch.cat((p[: pmf_length[i]], tail_mass[i]), dim=0)
            _cdf = torch.as_tensor(pmf_to_quantized_cdf(prob.detach().numpy(), self.entropy_coder_precision))
            cdf[i, : _cdf.size(0)] = _cdf
        return cdf

    def _check_tables(self):
        if self._quantized_cdf.numel() == 0:
            raise ValueError("Uninitialized CDFs. Run update() first")

    def compress(self, inputs, indexes, means=None):
        from . import rans_ref
        symbols = self.quantize(inputs, "symbols", means)
        self._check_tables()
        strings = []
        for i in range(symbols.size(0)):
            strings.append(rans_ref.encode_with_indexes(
                symbols[i].reshape(-1).int().numpy(), indexes[i].reshape(-1).int().numpy(),
                self._quantized_cdf.numpy(), self._cdf_length.reshape(-1).int().numpy(),
                self._offset.reshape(-1).int().numpy()))
        return strings

    def decompress(self, strings, indexes, dtype=torch.float, means=None):
        from . import rans_ref
        self._check_tables()
        outputs = self._quantized_cdf.new_empty(indexes.size())
        for i, s in enumerate(strings):
            values = rans_ref.decode_with_indexes(
                s, indexes[i].reshape(-1).int().numpy(), self._quantized_cdf.numpy(),
                self._cdf_length.reshape(-1).int().numpy(), self._offset.reshape(-1).int().numpy())
            outputs[i] = torch.as_tensor(values, dtype=outputs.dtype).reshape(outputs[i].size())
        return self.dequantize(outputs, means, dtype)


class EntropyBottleneck(EntropyModel):
    """CAI1: factorised prior (Ballé et al. 2018)."""

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

    def _get_medians(self):
        return self.quantiles[:, :, 1:2]

    def update(self, force=False, update_quantiles=False):
        if self._offset.numel() > 0 and not force:
            return False
        medians = self.quantiles[:, 0, 1]
        minima = torch.clamp(torch.ceil(medians - self.quantiles[:, 0, 0]).int(), min=0)
        maxima = torch.clamp(torch.ceil(self.quantiles[:, 0, 2] - medians).int(), min=0)
        self._offset = -minima
        pmf_start = medians - minima
        pmf_length = maxima + minima + 1
        max_length = int(pmf_length.max().item())
        samples = torch.arange(max_length)[None, :] + pmf_start[:, None, None]
        pmf, lower, upper = self._likelihood(samples, stop_gradient=True)
        pmf = pmf[:, 0, :]
        tail_mass = torch.sigmoid(lower[:, 0, :1]) + torch.sigmoid(-upper[:, 0, -1:])
        self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length)
        self._cdf_length = pmf_length + 2
        return True

    def loss(self):
        logits = self._logits_cumulative(self.quantiles, stop_gradient=True)
        return torch.abs(logits - self.target).sum()

    def _logits_cumulative(self, inputs, stop_gradient):
        logits = inputs
        for i in range(len(self.filters) + 1):
            matrix = self.matrices[i].detach() if stop_gradient else self.matrices[i]
            logits = torch.matmul(F.softplus(matrix), logits)
            bias = self.biases[i].detach() if stop_gradient else self.biases[i]
            logits = logits + bias
            if i < len(self.filters):
                factor = self.factors[i].detach() if stop_gradient else self.factors[i]
                logits = logits + torch.tanh(factor) * torch.tanh(logits)
        return logits

    def _likelihood(self, inputs, stop_gradient=False):
        lower = self._logits_cumulative(inputs - 0.5, stop_gradient)
        upper = self._logits_cumulative(inputs + 0.5, stop_gradient)
        return torch.sigmoid(upper) - torch.sigmoid(lower), lower, upper

    def forward(self, x, training=None):
        if training is None:
            training = self.training
        perm = list(range(x.dim()))
        perm[0], perm[1] = 1, 0
        x = x.permute(*perm).contiguous()
        shape = x.size()
        values = x.reshape(x.size(0), 1, -1)
        outputs = self.quantize(values, "noise" if training else "dequantize", self._get_medians())
        likelihood, _, _ = self._likelihood(outputs)
        if self.use_likelihood_bound:
            likelihood = self.likelihood_lower_bound(likelihood)
        outputs = outputs.reshape(shape).permute(*perm).contiguous()
        likelihood = likelihood.reshape(shape).permute(*perm).contiguous()
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
        indexes = self._build_indexes(x.size())
        return super().compress(x, indexes, self._medians_like(x.size(0), x.dim() - 2))

    def decompress(self, strings, size):
        output_size = (len(strings), self._quantized_cdf.size(0), *size)
        indexes = self._build_indexes(output_size)
        medians = self._medians_like(len(strings), len(size))
        return super().decompress(strings, indexes, medians.dtype, medians)


class GaussianConditional(EntropyModel):
    """CAI2: conditional Gaussian with a 64-entry scale table."""

    def __init__(self, scale_table, *args, scale_bound=0.11, tail_mass=1e-9, **kwargs):
        super().__init__(*args, **kwargs)
        if scale_table is not None and len(scale_table) < 1:
            raise ValueError("Invalid scale_table")
        self.tail_mass = float(tail_mass)
        if scale_bound is None and scale_table:
            scale_bound = scale_table[0]
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
        return scipy.stats.norm.ppf(quantile)

    def update_scale_table(self, scale_table, force=False):
        if self._offset.numel() > 0 and not force:
            return False
        self.scale_table = self._prepare_scale_table(scale_table)
        self.update()
        return True

    def update(self):
        multiplier = -self._standardized_quantile(self.tail_mass / 2)
        pmf_center = torch.ceil(self.scale_table * multiplier).int()
        pmf_length = 2 * pmf_center + 1
        max_length = int(torch.max(pmf_length).item())
        samples = torch.abs(torch.arange(max_length).int() - pmf_center[:, None]).float()
        samples_scale = self.scale_table.unsqueeze(1).float()
        upper = self._standardized_cumulative((0.5 - samples) / samples_scale)
        lower = self._standardized_cumulative((-0.5 - samples) / samples_scale)
        pmf = upper - lower
        tail_mass = 2 * lower[:, :1]
        self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length)
        self._offset = -pmf_center
        self._cdf_length = pmf_length + 2

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
        outputs = self.quantize(inputs, "noise" if training else "dequantize", means)
        likelihood = self._likelihood(outputs, scales, means)
        if self.use_likelihood_bound:
            likelihood = self.likelihood_lower_bound(likelihood)
        return outputs, likelihood

    def build_indexes(self, scales):
        scales = self.lower_bound_scale(scales)
        indexes = scales.new_full(scales.size(), len(self.scale_table) - 1).int()
        for s in self.scale_table[:-1]:
            indexes -= (scales <= s).int()
        return indexes


class CompressionModel(nn.Module):
    """models/base.py: update() / aux_loss() over child entropy models."""

    def __init__(self, entropy_bottleneck_channels=None, init_weights=None):
        super().__init__()
        if entropy_bottleneck_channels is not None:
            self.entropy_bottleneck = EntropyBottleneck(entropy_bottleneck_channels)

    def aux_loss(self):
        return sum(m.loss() for m in self.modules() if isinstance(m, EntropyBottleneck))

    def update(self, scale_table=None, force=False, update_quantiles=False):
        if scale_table is None:
            scale_table = get_scale_table()
        updated = False
        for _, module in self.named_modules():
            if isinstance(module, EntropyBottleneck):
                updated |= module.update(force=force)
            if isinstance(module, GaussianConditional):
                updated |= module.update_scale_table(scale_table, force=force)
        return updated
