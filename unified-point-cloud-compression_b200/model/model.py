"""UnifiedModel -- mirror of /root/reference/model/model.py:15-250 (forward / compress / decompress).

Same constructor config, sub-module names (g_a, g_s, entropy_model) and return conventions.  The
bitstream container, the ASCII-PLY + tmc3 (G-PCC) side channel for the stride-8 coordinates
(model.py:253-486) and Open3D are outside the graft: `compress(path=...)` / `decompress(path=...)`
raise, the component form (`path=None`) is the supported interface -- exactly what the reference
returns at model.py:186-187 and accepts at model.py:191.

`encode_block` / `decode_block` expose the GPU halves with the host range coder factored out, which is
what bench.py times (the coder stays on the host by design).
"""
import torch

from .. import MinkowskiEngine as ME
from .. import sparse_ops as so
from ..compressai.models.base import CompressionModel
from .entropy_models import MeanScaleHyperprior
from .transforms import AnalysisTransform, SparseSynthesisTransform


def partition_blocks(pointcloud, block_size):
    """Rows grouped by floor((xyz - min) / block_size), blocks in (bx,by,bz) order (model.py:121-127)."""
    xyz = pointcloud[:, :3]
    idx = ((xyz - xyz.min(dim=0)[0]) / block_size).floor().long()
    key = idx[:, 0] * 1_000_000 + idx[:, 1] * 1_000 + idx[:, 2]
    order = torch.argsort(key, stable=True)
    _, counts = torch.unique_consecutive(key[order], return_counts=True)
    return pointcloud[order], counts.tolist()


class UnifiedModel(CompressionModel):
    def __init__(self, config):
        super().__init__()
        self.g_a = AnalysisTransform(config["g_a"])
        self.g_s = SparseSynthesisTransform(config["g_s"])
        self.entropy_model = MeanScaleHyperprior(config["entropy_model"])

    def update(self):
        self.entropy_model.update(force=True)

    def aux_loss(self):
        return self.entropy_model.aux_loss()

    # -- training forward (model.py:45-90) -------------------------------------------------------
    def forward(self, x, q, Lambda):
        occ = torch.ones((x.C.shape[0], 1), device=x.device)
        x_in = ME.SparseTensor(torch.cat([occ, x.F], dim=1), _coord_set=x.coord_set)
        coords = ME.SparseTensor(occ, _coord_set=x.coord_set)
        y, k = self.g_a(x_in)
        y_hat, likelihoods = self.entropy_model(y, q)
        x_hat, points, predictions = self.g_s(y_hat, coords=coords, k=k)
        return {"prediction": x_hat, "points": points, "occ_predictions": predictions, "q_map": Lambda,
                "likelihoods": {"y": likelihoods[0], "z": likelihoods[1]}}

    # -- GPU halves --------------------------------------------------------------------------------
    def block_tensor(self, x_block):
        """(n,6) float [xyz|rgb] rows of one block -> deduplicated SparseTensor with the occupancy channel.

        Rows are put in Z-order first: MinkowskiEngine semantics do not depend on row order, and every derived
        coordinate set (strided / generated, first-occurrence order) inherits the spatial coherence, which is what
        lets a 128-row convolution tile skip the kernel offsets its surface patch does not use."""
        coords0, _ = so.gather_block(x_block)                      # integer coordinates, caller's row order (sort keys only)
        order = so.sort_rows_i32(coords0, morton=True)
        coords4, feats = so.gather_block(x_block, order)           # one pass: Z-ordered coordinates + [1, r, g, b] rows
        uniq, index, table = so.unique_first(coords4)
        if uniq.shape[0] != coords4.shape[0]:
            coords4, feats = uniq, feats[index.long()]
        cs = so.CoordSet(coords4, 1, table=table)
        cs.single_batch = True                                     # one block = one batch item: level sizes are row counts
        return ME.SparseTensor(feats, _coord_set=cs)

    def encode_block(self, x_block, q):
        """g_a + hyper-analysis + quantisation + CDF indexes on the GPU; returns device symbol buffers + k."""
        x = self.block_tensor(x_block)
        y, k = self.g_a(x)
        enc = self.entropy_model.encode_symbols(y, q)
        enc["k"] = k
        return enc

    def decode_block(self, y_points, z_symbols, y_symbol_source, k, q):
        """coords down x2 -> entropy decode -> g_s on the GPU; returns the SparseTensor x_hat."""
        y_points = y_points.to(torch.int32)
        y_set = so.coord_set_of(y_points, 8) or ME.SparseTensor(
            torch.empty((y_points.shape[0], 0), device=y_points.device), y_points, tensor_stride=8).coord_set
        z_set = SparseSynthesisTransform.coords_down(y_set, 2)
        y_hat = self.entropy_model.decode_symbols(y_set.coords, z_set.coords, z_symbols, y_symbol_source, q)
        return self.g_s(y_hat, k=k)

    # -- public API (model.py:94-250) ----------------------------------------------------------------
    def compress(self, pointcloud, q, path=None, block_size=1024, scaling_factor=1.0):
        if path:
            raise NotImplementedError("bitstream container + G-PCC side channel are outside the graft; use path=None")
        if scaling_factor != 1.0:
            pointcloud = pointcloud.clone()
            pointcloud[:, :3] = torch.round(pointcloud[:, :3] / scaling_factor).int()
        x_sorted, counts = partition_blocks(pointcloud, block_size)
        bitstreams, shapes, coords, q_vals, ks = [], [], [], [], []
        start = 0
        for count in counts:
            x = self.block_tensor(x_sorted[start:start + count])
            y, k = self.g_a(x)
            _, strings, shape = self.entropy_model.compress(y, q)
            bitstreams.append(strings); shapes.append(shape); coords.append(y.C); q_vals.append(q); ks.append(k)
            start += count
        return bitstreams, shapes, ks, coords, q_vals

    def decompress(self, path=None, coordinates=None, strings=None, shape=None, k=None, q_vals=None):
        if path:
            raise NotImplementedError("bitstream container + G-PCC side channel are outside the graft; pass the components")
        feats, pts = [], []
        for i, (b_strings, b_shape, b_coords, b_k) in enumerate(zip(strings, shape, coordinates, k)):
            b_coords = b_coords.to(torch.int32)
            y_set = so.coord_set_of(b_coords, 8) or ME.SparseTensor(
                torch.empty((b_coords.shape[0], 0), device=b_coords.device), b_coords, tensor_stride=8).coord_set
            z_set = SparseSynthesisTransform.coords_down(y_set, 2)
            y_hat = self.entropy_model.decompress([y_set.coords, z_set.coords], b_strings, b_shape, q_vals[i])
            x_hat = self.g_s(y_hat, k=b_k)
            feats.append(x_hat.F); pts.append(x_hat.C)
        return self.finalize(torch.cat(pts, 0), torch.cat(feats, 0))

    @staticmethod
    def finalize(coords, feats):
        """8-bit colour rounding + [xyz|rgb] rows (model.py:247-250)."""
        rgb = torch.clamp(torch.round(feats * 255), 0.0, 255.0) / 255
        return torch.cat([coords[:, 1:4].to(rgb.dtype), rgb], dim=1)
