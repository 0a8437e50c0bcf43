"""Drop-in for the `MinkowskiEngine` names the reference imports, backed by libpcg.so (sm_100a).

Mirrors (same class names, constructor arguments, parameter names/shapes and error behaviour) the
surface used at model/transforms.py:32-166, model/entropy_models.py:28-191, model/blocks.py:53,
model/model.py:66-70,147-161,227, utils.py:159, loss.py:105,124-125,185-188,221-271,
train.py:199-211.  No CPU path: tensors are moved to / must live on CUDA.

State-dict compatibility: `kernel` is (K,Cin,Cout) (or (Cin,Cout) when K == 1), `bias` is (1,Cout).
"""
import math
import types

import torch
import torch.nn as nn

from .. import sparse_ops as so

__version__ = "0.5.4+pcg.b200"


def _cuda_device(device, *tensors):
    if device is not None:
        dev = torch.device(device) if not isinstance(device, torch.device) else device
        if dev.type == "cuda":
            return dev
        if dev.type == "cpu":
            raise RuntimeError("MinkowskiEngine(pcg): no CPU path -- pass device='cuda'")
    for t in tensors:
        if t is not None and torch.is_tensor(t) and t.is_cuda:
            return t.device
    raise RuntimeError("MinkowskiEngine(pcg): no CPU path -- tensors must be on a CUDA device")


def _as_stride(ts):
    if isinstance(ts, (list, tuple)):
        assert len(set(int(t) for t in ts)) == 1, "anisotropic tensor strides are not supported"
        return int(ts[0])
    return int(ts)


class SparseTensor:
    """ME.SparseTensor (ME1): `.C` (N,4) int32, `.F` (N,C) fp32, `.tensor_stride`, `.device`."""

    def __init__(self, features=None, coordinates=None, tensor_stride=1, coordinate_map_key=None,
                 coordinate_manager=None, quantization_mode=None, allocator_type=None, minkowski_algorithm=None,
                 requires_grad=None, device=None, _coord_set=None):
        if features is None:
            raise ValueError("features must be given")
        if features.dim() == 1:
            features = features.unsqueeze(1)
        ts = _as_stride(tensor_stride)
        if _coord_set is not None:                      # internal, trusted construction
            self._cs = _coord_set
            self._F = features
        else:
            if coordinates is None:
                raise ValueError("coordinates must be given")
            dev = _cuda_device(device, features, coordinates)
            cs = so.coord_set_of(coordinates, ts) if (torch.is_tensor(coordinates) and coordinates.is_cuda and coordinates.device == dev) else None
            features = features.to(dev)
            if cs is None:
                c = coordinates.to(dev)
                if c.dtype != torch.int32:
                    c = torch.floor(c).to(torch.int32) if c.is_floating_point() else c.to(torch.int32)
                if c.shape[0] != features.shape[0]:
                    raise ValueError("coordinates and features must have the same number of rows")
                uniq, index, table = so.unique_first(c)
                if uniq.shape[0] != c.shape[0]:         # duplicates: keep the first row of each coordinate
                    features = features[index.long()]
                    c = uniq
                else:
                    c = c.contiguous()
                cs = so.CoordSet(c, ts, table=table)
            self._cs = cs
            self._F = features
        if requires_grad:
            self._F.requires_grad_(True)
        self.coordinate_manager = coordinate_manager
        self.coordinate_map_key = coordinate_map_key

    C = property(lambda self: self._cs.coords)
    coordinates = property(lambda self: self._cs.coords)
    F = property(lambda self: self._F)
    features = property(lambda self: self._F)
    tensor_stride = property(lambda self: [self._cs.ts] * 3)
    device = property(lambda self: self._F.device)
    D = property(lambda self: 3)
    dimension = property(lambda self: 3)
    shape = property(lambda self: self._F.shape)
    dtype = property(lambda self: self._F.dtype)
    coord_set = property(lambda self: self._cs)

    def size(self, *a):
        return self._F.size(*a)

    def __len__(self):
        return self._F.shape[0]

    def __repr__(self):
        return f"SparseTensor(pcg, N={self._F.shape[0]}, C={self._F.shape[1]}, tensor_stride={self._cs.ts}, device={self.device})"

    def features_at_coordinates(self, query_coordinates):
        """ME6: trilinear interpolation at tensor-stride spacing; on-grid queries are exact row lookups."""
        if torch.is_grad_enabled() and self._F.requires_grad:
            from ..autograd import features_at_autograd
            return features_at_autograd(self._F, self._cs, query_coordinates.to(self._F.device))
        return so.features_at(self._F, self._cs, query_coordinates)

    def _like(self, features, coord_set=None):
        return SparseTensor(features, _coord_set=coord_set or self._cs)


# ----------------------------------------------------------------------------- convolutions
class MinkowskiConvolutionBase(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, is_transpose=False, expand_coordinates=False, dimension=-1):
        super().__init__()
        if dimension != 3:
            raise ValueError("only dimension=3 is supported")
        if dilation != 1 or kernel_generator is not None:
            raise NotImplementedError("dilation / custom kernel generators are not on the reference's path")
        self.in_channels, self.out_channels = int(in_channels), int(out_channels)
        self.kernel_size, self.stride = int(kernel_size), int(stride)
        self.is_transpose, self.expand_coordinates = bool(is_transpose), bool(expand_coordinates)
        self.dimension = 3
        self.kernel_volume = self.kernel_size ** 3
        shape = (self.kernel_volume, self.in_channels, self.out_channels) if self.kernel_volume > 1 else (self.in_channels, self.out_channels)
        self.kernel = nn.Parameter(torch.empty(*shape))
        self.bias = nn.Parameter(torch.empty(1, self.out_channels)) if bias else None
        self._packed = so.PackedWeights()
        self._packed_gemm = so.PackedGemmWeights()
        self.fused_activation = so.ACT_NONE      # host mirror may fuse the following (Leaky)ReLU into the epilogue
        self.emit = "f32"                        # "both": also write the bf16 shadow the next tensor-core layer reads
        self.reset_parameters()

    def reset_parameters(self, is_transpose=None):
        with torch.no_grad():
            n = (self.out_channels if self.is_transpose else self.in_channels) * self.kernel_volume
            stdv = 1.0 / math.sqrt(n)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def prepare(self):
        """Pack the operand images the static kernel-selection rules will ask for (sparse_ops.pack_kind) before the first
        frame, so that concurrent frames (codec.FramePipeline) only ever read these caches."""
        w3 = self.kernel.detach() if self.kernel.dim() == 3 else self.kernel.detach().unsqueeze(0)
        if not w3.is_cuda:
            return
        K, cin, cout = w3.shape
        if self.is_transpose and self.expand_coordinates and so.cells_eligible(self.kernel_size, self.stride, cin, cout):
            self._packed_gemm.get(w3)
            return
        kind = so.pack_kind(K, cin, cout, fp32_input=cin <= 8)
        if kind is not None:
            self._packed.get(w3, kind=kind)

    def forward(self, input, coordinates=None, _act=None):
        # _act: activation fused into the epilogue for THIS call (passed as an argument, not kept on the module: several
        # frames may run through the same module concurrently, codec.FramePipeline)
        act = self.fused_activation if _act is None else _act
        if not isinstance(input, SparseTensor):
            raise TypeError("input must be a SparseTensor")
        if input.F.shape[1] != self.in_channels:
            raise RuntimeError(f"expected {self.in_channels} input channels, got {input.F.shape[1]}")
        in_set = input.coord_set
        grad = torch.is_grad_enabled() and (input.F.requires_grad or self.kernel.requires_grad and self.training)
        if not self.is_transpose:
            out_set = in_set.strided(self.stride)
        elif self.expand_coordinates:
            if not grad and so.cells_eligible(self.kernel_size, self.stride, self.in_channels, self.out_channels):
                # kernel 5 / stride 2: octree-cell candidate set + GEMM-first transposed convolution
                out_set = in_set.generated_cells()
                out = so.gen_conv_transpose_cells(input.F, in_set, out_set, self.kernel, self.bias, act=act,
                                                  packed=self._packed_gemm, out=self.emit)
                return SparseTensor(out, _coord_set=out_set)
            out_set = in_set.generated(self.kernel_size, self.stride)
        else:
            if coordinates is None:
                raise ValueError("MinkowskiConvolutionTranspose needs the output coordinates (SparseTensor or (N,4) tensor)")
            ts_out = in_set.ts // self.stride
            if isinstance(coordinates, SparseTensor):
                out_set = coordinates.coord_set
            else:
                out_set = so.coord_set_of(coordinates, ts_out) or SparseTensor(
                    torch.empty((coordinates.shape[0], 0), device=input.device), coordinates, tensor_stride=ts_out,
                    device=input.device).coord_set
            if out_set.ts != ts_out:
                raise RuntimeError(f"output coordinates have tensor stride {out_set.ts}, expected {ts_out}")
        if grad:
            from ..autograd import sparse_conv_autograd
            out = sparse_conv_autograd(input.F, in_set, out_set, self.kernel, self.bias, self.kernel_size,
                                       self.is_transpose, act)
        else:
            out = so.sparse_conv(input.F, in_set, out_set, self.kernel, self.bias, self.kernel_size, self.is_transpose,
                                 act=act, packed=self._packed, out=self.emit)
        return SparseTensor(out, _coord_set=out_set)


class MinkowskiConvolution(MinkowskiConvolutionBase):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, dimension=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, dilation, bias, kernel_generator,
                         False, expand_coordinates, dimension)


class MinkowskiConvolutionTranspose(MinkowskiConvolutionBase):
    """ME4: transposed convolution onto caller-given output coordinates (no call site in the reference)."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, dimension=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, dilation, bias, kernel_generator,
                         True, expand_coordinates, dimension)


class MinkowskiGenerativeConvolutionTranspose(MinkowskiConvolutionBase):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, dimension=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, dilation, bias, kernel_generator,
                         True, True, dimension)


class MinkowskiChannelwiseConvolution(nn.Module):
    """ME8 (loss.py:185-188): out[o] += F[n] * kernel[i]; the reference swaps in a (K,1) kernel."""

    def __init__(self, in_channels, kernel_size=-1, stride=1, dilation=1, bias=False, kernel_generator=None, dimension=-1):
        super().__init__()
        if stride != 1 or dilation != 1:
            raise NotImplementedError("channel-wise convolution: stride/dilation 1 only")
        self.in_channels, self.kernel_size = int(in_channels), int(kernel_size)
        K = self.kernel_size ** 3
        self.kernel = nn.Parameter(torch.empty(K, self.in_channels))
        self.bias = nn.Parameter(torch.empty(1, self.in_channels)) if bias else None
        with torch.no_grad():
            stdv = 1.0 / math.sqrt(self.in_channels * K)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def forward(self, input):
        k = self.kernel
        if k.shape[1] not in (1, input.F.shape[1]):
            raise RuntimeError("channel-wise kernel must have 1 or C columns")
        out = so.channelwise_conv(input.F, input.coord_set, k.to(input.device), self.kernel_size)
        if self.bias is not None:
            out = out + self.bias
        return input._like(out)


# ----------------------------------------------------------------------------- pruning / activations
class MinkowskiPruning(nn.Module):
    """ME5: rows where mask is true, order preserved, same tensor stride."""

    def forward(self, input, mask):
        if mask.shape[0] != input.F.shape[0]:
            raise ValueError("mask and input must have the same number of rows")
        mask = mask.to(input.device)
        if torch.is_grad_enabled() and input.F.requires_grad:
            oc, _ = so.prune_rows(input.C, None, mask)
            return SparseTensor(input.F[mask.bool()], _coord_set=so.CoordSet(oc, input.coord_set.ts))   # differentiable row gather
        oc, of = so.prune_rows(input.C, input.F, mask)
        return SparseTensor(of, _coord_set=so.CoordSet(oc, input.coord_set.ts))


class MinkowskiReLU(nn.Module):
    def __init__(self, inplace=False):
        super().__init__()

    def forward(self, input):
        if torch.is_grad_enabled() and input.F.requires_grad:
            return input._like(torch.relu(input.F))
        return input._like(so.activation(input.F, so.ACT_RELU))


class MinkowskiLeakyReLU(nn.Module):
    def __init__(self, negative_slope=0.01, inplace=False):
        super().__init__()
        if abs(negative_slope - 0.01) > 1e-12:
            raise NotImplementedError("only the default negative_slope=0.01 is on the reference's path")
        self.negative_slope = negative_slope

    def forward(self, input):
        if torch.is_grad_enabled() and input.F.requires_grad:
            return input._like(torch.nn.functional.leaky_relu(input.F, 0.01))
        return input._like(so.activation(input.F, so.ACT_LEAKY))


class MinkowskiAvgPooling(nn.Module):
    """Constructed at loss.py:124-125 and never called."""

    def __init__(self, kernel_size, stride=1, dilation=1, kernel_generator=None, dimension=None):
        super().__init__()
        self.kernel_size, self.stride = kernel_size, stride

    def forward(self, input):
        raise NotImplementedError("MinkowskiAvgPooling has no call site in the reference")


# ----------------------------------------------------------------------------- ME.utils
def sparse_quantize(coordinates, features=None, labels=None, ignore_label=-100, return_index=False,
                    return_inverse=False, return_maps_only=False, quantization_size=None, device="cpu"):
    """ME9: floor(c / quantization_size) then first-occurrence unique, on the GPU hash."""
    if not torch.is_tensor(coordinates):
        coordinates = torch.as_tensor(coordinates)
    if not coordinates.is_cuda:
        raise RuntimeError("sparse_quantize(pcg): coordinates must be on a CUDA device (no CPU path)")
    c = coordinates
    if quantization_size is not None:
        c = c / quantization_size
    disc = torch.floor(c).to(torch.int32) if c.is_floating_point() else c.to(torch.int32)
    c4 = disc if disc.shape[1] == 4 else torch.cat([torch.zeros_like(disc[:, :1]), disc], dim=1)
    uniq, index, _ = so.unique_first(c4)
    out_c = uniq if disc.shape[1] == 4 else uniq[:, 1:]
    if return_index and features is None:
        return out_c, index.long()
    if features is None:
        return out_c
    return out_c, features[index.long()]


def sparse_collate(coords, feats, labels=None, dtype=torch.int32, device=None):
    """ME9: prepend the batch column and concatenate."""
    cs, fs = [], []
    for b, (c, f) in enumerate(zip(coords, feats)):
        c = torch.as_tensor(c)
        c = torch.floor(c).to(dtype) if c.is_floating_point() else c.to(dtype)
        cs.append(torch.cat([torch.full((c.shape[0], 1), b, dtype=dtype, device=c.device), c], dim=1))
        fs.append(torch.as_tensor(f))
    return torch.cat(cs, 0), torch.cat(fs, 0)


utils = types.ModuleType(__name__ + ".utils")
utils.sparse_quantize = sparse_quantize
utils.sparse_collate = sparse_collate
utils.collation = types.ModuleType(__name__ + ".utils.collation")
utils.collation.sparse_collate = sparse_collate
