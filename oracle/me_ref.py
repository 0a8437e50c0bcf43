"""CPU restatement of the MinkowskiEngine surface the reference calls.  TEST INFRASTRUCTURE.

Un-vendored dependency restated: NVIDIA/MinkowskiEngine master (v0.5.4 lineage, unpinned
`git clone`, /root/reference/README.md:43-53).  Semantics are recalled from the upstream
sources (MinkowskiSparseTensor.py, MinkowskiConvolution.py, MinkowskiKernelGenerator.py,
MinkowskiPruning.py, MinkowskiInterpolation.py, MinkowskiChannelwiseConvolution.py,
utils/quantization.py, src/kernel_region.hpp, src/coordinate_map_cpu.hpp,
src/convolution_kernel.cpp) -- PARITY UNPINNED, see oracle/__init__.py.  The geometry and
weight-index conventions are pinned independently against dense conv3d / conv_transpose3d
in tests/test_oracle_dense.py.

Call sites in the reference this module serves:
  model/model.py:66-70,147-161,227   model/blocks.py:53   model/transforms.py:32-166
  model/entropy_models.py:28-126,178-191,287,294,335,374,381,439,446,486
  utils.py:159   loss.py:105,124-125,185-188,221-271   train.py:199-211

Everything is plain torch-CPU (autograd-capable, so it also yields reference gradients).
Conventions (single place; flip here if real ME ever disagrees):
  * coordinates are (N,4) int32 rows (batch, x, y, z);
  * kernel offsets: odd k -> {-(k//2)..k//2}, even k -> {0..k-1}, in units of the INPUT
    tensor stride (conv) / OUTPUT tensor stride (transposed); offset index
    i = ix + k*iy + k*k*iz (first spatial axis fastest);
  * strided conv output coords = floor(c / (s*ts)) * (s*ts), first-occurrence order;
  * generative transposed conv output coords = unique{p + off_i*ts_out}, first-occurrence
    order over (input row major, offset minor); out[f] += F[p] @ W[i];
  * duplicate input coordinates keep the first row (ME's default RANDOM_SUBSAMPLE picks an
    arbitrary one; "first" is the deterministic CPU behaviour).
"""
import math
import types

import numpy as np
import torch
import torch.nn as nn

_BIAS = 1 << 15


# ----------------------------------------------------------------------------- coordinates
def pack_keys(C):
    """(N,4) int coords -> int64 keys whose integer order is lexicographic (b,x,y,z)."""
    C = np.asarray(C, dtype=np.int64)
    if C.size and (C[:, 1:].min() < -_BIAS or C[:, 1:].max() >= _BIAS or C[:, 0].min() < 0 or C[:, 0].max() >= (1 << 15)):
        raise ValueError("coordinate out of the 16-bit packing range")
    return (C[:, 0] << 48) | ((C[:, 1] + _BIAS) << 32) | ((C[:, 2] + _BIAS) << 16) | (C[:, 3] + _BIAS)


def unique_first(C):
    """Row indices of first occurrences, in order of first occurrence."""
    keys = pack_keys(C)
    _, first = np.unique(keys, return_index=True)
    first.sort()
    return first


def kernel_offsets(kernel_size, dimension=3):
    """(K,3) integer offsets, first spatial axis fastest (ME kernel_region.hpp order)."""
    assert dimension == 3
    k = int(kernel_size)
    r = np.arange(-(k // 2), k // 2 + 1) if k % 2 == 1 else np.arange(0, k)
    iz, iy, ix = np.meshgrid(r, r, r, indexing="ij")
    return np.stack([ix.ravel(), iy.ravel(), iz.ravel()], axis=1).astype(np.int64)


class _Lookup:
    """Sorted-key row lookup (stands in for ME's coordinate hash map)."""

    def __init__(self, C):
        keys = pack_keys(C)
        self.order = np.argsort(keys, kind="stable")
        self.sorted = keys[self.order]

    def __call__(self, Cq):
        q = pack_keys(Cq)
        if self.sorted.size == 0:
            return np.full(q.shape, -1, np.int64)
        pos = np.searchsorted(self.sorted, q)
        pos = np.minimum(pos, self.sorted.size - 1)
        hit = self.sorted[pos] == q
        return np.where(hit, self.order[pos], -1)


def stride_coords(C, ts_in, stride):
    """Output coordinates of a stride-`stride` convolution (ME coordinate_map `stride`)."""
    C = np.asarray(C, dtype=np.int64)
    s = int(ts_in) * int(stride)
    out = C.copy()
    out[:, 1:] = np.floor_divide(C[:, 1:], s) * s
    return out[unique_first(out)].astype(np.int32)


def generate_coords(C, ts_out, kernel_size):
    """Output coordinates of a generative transposed convolution (ME `stride_region`)."""
    C = np.asarray(C, dtype=np.int64)
    offs = kernel_offsets(kernel_size) * int(ts_out)
    cand = np.repeat(C, len(offs), axis=0)
    cand[:, 1:] += np.tile(offs, (len(C), 1))
    return cand[unique_first(cand)].astype(np.int32)


def kernel_map(C_in, C_out, kernel_size, ts, transposed=False):
    """Per-offset (in_rows, out_rows) pair lists.

    conv:        in = out + off_i*ts      (ts = input tensor stride)
    transposed:  out = in + off_i*ts      (ts = output tensor stride)
    Pairs are ordered by ascending out row inside each offset.
    """
    C_in = np.asarray(C_in, dtype=np.int64)
    C_out = np.asarray(C_out, dtype=np.int64)
    look = _Lookup(C_in)
    maps = []
    for off in kernel_offsets(kernel_size) * int(ts):
        q = C_out.copy()
        q[:, 1:] += (-off if transposed else off)
        rows = look(q)
        out_rows = np.nonzero(rows >= 0)[0]
        maps.append((rows[out_rows], out_rows))
    return maps


def neighbour_table(C_in, C_out, kernel_size, ts, transposed=False):
    """(K, N_out) int32 table of input rows (-1 = absent): the layout the CUDA path uses."""
    K = int(kernel_size) ** 3
    nbr = np.full((K, len(C_out)), -1, np.int32)
    for i, (ir, orow) in enumerate(kernel_map(C_in, C_out, kernel_size, ts, transposed)):
        nbr[i, orow] = ir
    return nbr


def sparse_conv_features(F, maps, W, bias, n_out):
    """out[o] = bias + sum_i sum_{(n,o) in M_i} F[n] @ W[i]   (ME convolution_kernel.cpp)."""
    out = F.new_zeros((n_out, W.shape[-1]))
    for i, (ir, orow) in enumerate(maps):
        if len(ir) == 0:
            continue
        ir_t = torch.as_tensor(ir, dtype=torch.long)
        or_t = torch.as_tensor(orow, dtype=torch.long)
        out = out.index_add(0, or_t, F[ir_t] @ W[i])
    if bias is not None:
        out = out + bias
    return out


def lexsort_rows(C):
    """Permutation sorting rows lexicographically by (b,x,y,z) -- utils.py:142-187."""
    return np.argsort(pack_keys(C), kind="stable")


# ----------------------------------------------------------------------------- SparseTensor
def _as_stride(ts):
    if isinstance(ts, (list, tuple)):
        return [int(t) for t in ts]
    return [int(ts)] * 3


class SparseTensor:
    """ME.SparseTensor restatement (ME1)."""

    def __init__(self, features=None, coordinates=None, tensor_stride=1, coordinate_map_key=None,
                 coordinate_manager=None, quantization_mode=None, device=None, requires_grad=None,
                 _trusted=False):
        assert features is not None and coordinates is not None
        if not torch.is_tensor(coordinates):
            coordinates = torch.as_tensor(np.asarray(coordinates))
        if coordinates.dtype != torch.int32:
            coordinates = torch.floor(coordinates.float()).int() if coordinates.is_floating_point() else coordinates.int()
        coordinates = coordinates.cpu()
        features = features.cpu() if features.device.type != "cpu" else features
        if features.dim() == 1:
            features = features.unsqueeze(1)
        assert coordinates.shape[0] == features.shape[0], "coordinate / feature row mismatch"
        if not _trusted and coordinates.shape[0] > 0:
            first = unique_first(coordinates.numpy())
            if len(first) != coordinates.shape[0]:
                idx = torch.as_tensor(first, dtype=torch.long)
                coordinates, features = coordinates[idx], features[idx]
        self._C = coordinates.contiguous()
        self._F = features
        self.tensor_stride = _as_stride(tensor_stride)
        self.coordinate_manager = coordinate_manager
        self.coordinate_map_key = coordinate_map_key

    C = property(lambda self: self._C)
    coordinates = property(lambda self: self._C)
    F = property(lambda self: self._F)
    features = property(lambda self: self._F)
    device = property(lambda self: self._F.device)
    D = property(lambda self: 3)
    shape = property(lambda self: self._F.shape)
    dtype = property(lambda self: self._F.dtype)

    def __len__(self):
        return self._F.shape[0]

    def features_at_coordinates(self, query):
        """Trilinear interpolation at tensor-stride spacing (ME6, MinkowskiInterpolation.py)."""
        ts = self.tensor_stride[0]
        q = query.detach().cpu().double().numpy()
        b = q[:, :1].astype(np.int64)
        base = np.floor(q[:, 1:] / ts) * ts
        frac = (q[:, 1:] - base) / ts
        look = _Lookup(self._C.numpy())
        out = self._F.new_zeros((q.shape[0], self._F.shape[1]))
        for corner in range(8):
            d = np.array([(corner >> a) & 1 for a in range(3)], dtype=np.int64)
            w = np.prod(np.where(d == 1, frac, 1.0 - frac), axis=1)
            cq = np.concatenate([b, (base + d * ts).astype(np.int64)], axis=1)
            safe = np.clip(cq, [0, -_BIAS, -_BIAS, -_BIAS], [(1 << 15) - 1, _BIAS - 1, _BIAS - 1, _BIAS - 1])
            rows = look(safe)
            rows = np.where((safe == cq).all(axis=1), rows, -1)
            sel = np.nonzero((rows >= 0) & (w != 0))[0]
            if len(sel) == 0:
                continue
            sel_t = torch.as_tensor(sel, dtype=torch.long)
            w_t = torch.as_tensor(w[sel], dtype=self._F.dtype).unsqueeze(1)
            out = out.index_add(0, sel_t, self._F[torch.as_tensor(rows[sel], dtype=torch.long)] * w_t)
        return out


# ----------------------------------------------------------------------------- modules
class MinkowskiConvolutionBase(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, is_transpose=False, expand_coordinates=False, dimension=3):
        super().__init__()
        assert dimension == 3 and dilation == 1
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride = int(kernel_size), int(stride)
        self.is_transpose, self.expand_coordinates = is_transpose, expand_coordinates
        self.kernel_volume = self.kernel_size ** 3
        shape = (self.kernel_volume, in_channels, out_channels) if self.kernel_volume > 1 else (in_channels, out_channels)
        self.kernel = nn.Parameter(torch.empty(*shape))
        self.bias = nn.Parameter(torch.empty(1, out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        with torch.no_grad():
            n = (self.out_channels if self.is_transpose else self.in_channels) * self.kernel_volume
            stdv = 1.0 / math.sqrt(n)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def _weights(self):
        return self.kernel if self.kernel_volume > 1 else self.kernel.unsqueeze(0)

    def forward(self, input, coordinates=None):
        ts_in = input.tensor_stride[0]
        C_in = input.C.numpy()
        if not self.is_transpose:
            C_out = C_in if self.stride == 1 else stride_coords(C_in, ts_in, self.stride)
            ts_out = ts_in * self.stride
            maps = kernel_map(C_in, C_out, self.kernel_size, ts_in, transposed=False)
        else:
            ts_out = ts_in // self.stride
            assert ts_out >= 1, "transposed conv would produce a fractional tensor stride"
            if self.expand_coordinates:
                C_out = generate_coords(C_in, ts_out, self.kernel_size)
            else:
                assert coordinates is not None, "MinkowskiConvolutionTranspose needs output coordinates"
                C_out = (coordinates.C if isinstance(coordinates, SparseTensor) else coordinates).numpy()
            maps = kernel_map(C_in, C_out, self.kernel_size, ts_out, transposed=True)
        out = sparse_conv_features(input.F, maps, self._weights(), self.bias, len(C_out))
        return SparseTensor(out, torch.as_tensor(np.ascontiguousarray(C_out), dtype=torch.int32),
                            tensor_stride=ts_out, _trusted=True)


class MinkowskiConvolution(MinkowskiConvolutionBase):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, dimension=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, dilation, bias,
                         kernel_generator, False, expand_coordinates, dimension)


class MinkowskiConvolutionTranspose(MinkowskiConvolutionBase):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, dimension=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, dilation, bias,
                         kernel_generator, True, expand_coordinates, dimension)


class MinkowskiGenerativeConvolutionTranspose(MinkowskiConvolutionBase):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, dimension=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, dilation, bias,
                         kernel_generator, True, True, dimension)


class MinkowskiChannelwiseConvolution(nn.Module):
    """ME8: out[o] += F[n] * kernel[i]  (python loop over the kernel map upstream)."""

    def __init__(self, in_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, dimension=3):
        super().__init__()
        self.in_channels, self.kernel_size, self.stride = in_channels, int(kernel_size), int(stride)
        self.kernel = nn.Parameter(torch.empty(self.kernel_size ** 3, in_channels))
        self.bias = nn.Parameter(torch.empty(1, in_channels)) if bias else None
        with torch.no_grad():
            stdv = 1.0 / math.sqrt(in_channels * self.kernel_size ** 3)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def forward(self, input):
        assert self.stride == 1
        C = input.C.numpy()
        maps = kernel_map(C, C, self.kernel_size, input.tensor_stride[0])
        out = input.F.new_zeros(input.F.shape)
        for i, (ir, orow) in enumerate(maps):
            if len(ir):
                out = out.index_add(0, torch.as_tensor(orow, dtype=torch.long),
                                    input.F[torch.as_tensor(ir, dtype=torch.long)] * self.kernel[i])
        if self.bias is not None:
            out = out + self.bias
        return SparseTensor(out, input.C, tensor_stride=input.tensor_stride, _trusted=True)


class MinkowskiPruning(nn.Module):
    """ME5: keep rows where mask is true, order preserved."""

    def forward(self, input, mask):
        mask = mask.cpu().bool()
        return SparseTensor(input.F[mask], input.C[mask], tensor_stride=input.tensor_stride, _trusted=True)


class _Elementwise(nn.Module):
    def forward(self, input):
        return SparseTensor(self.fn(input.F), input.C, tensor_stride=input.tensor_stride, _trusted=True)


class MinkowskiReLU(_Elementwise):
    def __init__(self, inplace=False):
        super().__init__()
        self.fn = torch.relu


class MinkowskiLeakyReLU(_Elementwise):
    def __init__(self, negative_slope=0.01, inplace=False):
        super().__init__()
        self.negative_slope = negative_slope
        self.fn = lambda f: torch.nn.functional.leaky_relu(f, negative_slope)


class MinkowskiAvgPooling(nn.Module):
    """Constructed at loss.py:124-125, never called."""

    def __init__(self, kernel_size, stride=1, dilation=1, kernel_generator=None, dimension=None):
        super().__init__()
        self.kernel_size, self.stride = kernel_size, stride

    def forward(self, input):
        raise NotImplementedError("MinkowskiAvgPooling is never called by the reference")


# ----------------------------------------------------------------------------- ME.utils
def sparse_quantize(coordinates, features=None, labels=None, ignore_label=-100, return_index=False,
                    return_inverse=False, return_maps_only=False, quantization_size=None, device="cpu"):
    """ME9: floor(c / q) then first-occurrence unique (utils/quantization.py)."""
    is_t = torch.is_tensor(coordinates)
    c = coordinates.detach().cpu().numpy() if is_t else np.asarray(coordinates)
    if quantization_size is not None:
        c = c / quantization_size
    disc = np.floor(c).astype(np.int32)
    first = unique_first(_with_batch(disc))
    idx = torch.as_tensor(first, dtype=torch.long)
    out_c = torch.as_tensor(disc[first]) if is_t else disc[first]
    if features is None:
        return out_c
    return out_c, features[idx] if torch.is_tensor(features) else features[first]


def _with_batch(c):
    return c if c.shape[1] == 4 else np.concatenate([np.zeros((len(c), 1), c.dtype), c], axis=1)


def sparse_collate(coords, feats, labels=None, dtype=torch.int32, device=None):
    """ME9: prepend batch column + concatenate (utils/collation.py)."""
    cs, fs = [], []
    for b, (c, f) in enumerate(zip(coords, feats)):
        c = torch.as_tensor(c)
        cs.append(torch.cat([torch.full((c.shape[0], 1), b, dtype=dtype), torch.floor(c.float()).to(dtype) if c.is_floating_point() else c.to(dtype)], dim=1))
        fs.append(torch.as_tensor(f))
    return torch.cat(cs, 0), torch.cat(fs, 0)


utils = types.ModuleType("MinkowskiEngine.utils")
utils.sparse_quantize = sparse_quantize
utils.sparse_collate = sparse_collate
utils.collation = types.ModuleType("MinkowskiEngine.utils.collation")
utils.collation.sparse_collate = sparse_collate
