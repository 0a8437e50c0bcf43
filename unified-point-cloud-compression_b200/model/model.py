"""UnifiedModel -- mirror of /root/reference/model/model.py:15-250 (forward / compress / decompress).

Same constructor config, sub-module names (g_a, g_s, entropy_model) and return conventions, both for the
component form (`path=None`: what the reference returns at model.py:186-187 and accepts at model.py:191) and
for the file form: `save_bitstream` / `load_bitstream` write the reference's container (model.py:253-385: same
fields in the same order, big-endian as the `bitstream` package lays them out).  The one payload that differs is
the stride-8 coordinate side channel: the reference shells out to the external G-PCC binary tmc3 through an ASCII
PLY (model.py:388-486), which is not part of its tree; `gpcc_encode` / `gpcc_decode` here are a native lossless
octree coder on the host (csrc/host_octree.cpp) with the same contract (a point SET in, bytes out, the set back).

`encode_block` / `decode_block` expose the GPU halves with the host range coder factored out, which is
what bench.py times (the coder stays on the host by design).
"""
import struct

import torch

from .. import MinkowskiEngine as ME
from .. import host_coder
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

    def prepare(self):
        """Fill every lazily built cache of the inference path -- packed convolution weights, the fused heads' host copies,
        the factorised prior's packed parameters, the host CDF tables -- and synchronise.  Call after the parameters are final
        (load_state_dict / update()) and before frames are coded concurrently; codec.FramePipeline does it for its model."""
        with torch.no_grad():
            for m in self.modules():
                if isinstance(m, ME.MinkowskiConvolutionBase):
                    m.prepare()
            for head in (self.g_s.predict_1, self.g_s.predict_2, self.g_s.predict_3):
                first, last = head[0], head[2]
                if first.kernel.is_cuda and so.head_fusable(first.kernel_size, first.kernel, last.kernel_size, last.kernel):
                    first._packed.get(first.kernel.detach())
                    so.head_host_params(first._packed, last.kernel, first.bias)
            self.entropy_model.prepare()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        return self

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
        order = so.sort_rows_i32(coords0, morton=True, bits=so.LOCALITY_BITS)
        coords4, feats = so.gather_block(x_block, order)           # one pass: Z-ordered coordinates + [1, r, g, b] rows
        uniq, index, table = so.unique_first(coords4)
        if uniq.shape[0] != coords4.shape[0]:
            coords4, feats = uniq, feats[index.long()]
        cs = so.CoordSet(coords4, 1, table=table)
        cs.single_batch = True                                     # one block = one batch item: level sizes are row counts
        return ME.SparseTensor(feats, _coord_set=cs)

    def blocks_tensor(self, x_sorted, counts):
        """All blocks of a frame as ONE batched SparseTensor: block i = batch item i (rows of `x_sorted` grouped by block,
        `counts[i]` rows each, model.py:121-137).  The reference codes the blocks one after the other, each as its own batch-0
        tensor; nothing in the codec crosses batch items (kernel maps, per-batch top-k, per-batch gains), so one batched pass
        computes every block's result with one set of launches instead of one per block."""
        dev = x_sorted.device
        batch = torch.repeat_interleave(torch.arange(len(counts), dtype=torch.int32, device=dev),
                                        torch.as_tensor(counts, device=dev), output_size=int(sum(counts)))
        coords0, _ = so.gather_block(x_sorted)
        order = so.sort_rows_i32(coords0, morton=True, bits=so.LOCALITY_BITS)      # blocks are disjoint cubes: global Z-order keeps them apart
        coords4, feats = so.gather_block(x_sorted, order)
        coords4[:, 0] = batch[order.long()]
        uniq, index, table = so.unique_first(coords4)
        if uniq.shape[0] != coords4.shape[0]:
            coords4, feats = uniq, feats[index.long()]
        return ME.SparseTensor(feats, _coord_set=so.CoordSet(coords4, 1, table=table))

    def encode_blocks(self, x_sorted, counts, q):
        """`encode_block` for all blocks of a frame at once (batched); q: (1,2) or (len(counts),2)."""
        x = self.blocks_tensor(x_sorted, counts)
        y, k = self.g_a(x)
        qb = q if q.shape[0] == len(counts) else q.expand(len(counts), 2)
        enc = self.entropy_model.encode_symbols(y, qb)
        enc["k"] = k
        enc["q"] = qb
        return enc

    def encode_block(self, x_block, q):
        """g_a + hyper-analysis + quantisation + CDF indexes on the GPU; returns device symbol buffers + k."""
        x = self.block_tensor(x_block)
        y, k = self.g_a(x)
        enc = self.entropy_model.encode_symbols(y, q)
        enc["k"] = k
        if getattr(self.g_s, "topk_from_encoder", False):          # measurement switch (bench.py --topk-from encoder)
            s1 = x.coord_set
            s2 = s1.strided(2)
            enc["keep_sets"] = [s2.strided(2), s2, s1]
        return enc

    def decode_block(self, y_points, z_symbols, y_symbol_source, k, q, keep_sets=None):
        """coords down x2 -> entropy decode -> g_s on the GPU; returns the SparseTensor x_hat."""
        y_points = y_points.to(torch.int32)
        y_set = so.coord_set_of(y_points, 8) or ME.SparseTensor(
            torch.empty((y_points.shape[0], 0), device=y_points.device), y_points, tensor_stride=8).coord_set
        z_set = SparseSynthesisTransform.coords_down(y_set, 2)
        y_hat = self.entropy_model.decode_symbols(y_set.coords, z_set.coords, z_symbols, y_symbol_source, q)
        return self.g_s(y_hat, k=k, keep_sets=keep_sets)

    # -- public API (model.py:94-250) ----------------------------------------------------------------
    def compress(self, pointcloud, q, path=None, block_size=1024, scaling_factor=1.0):
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
        if path:
            self.save_bitstream(path, coords, bitstreams, shapes, ks, q_vals)
            return None
        return bitstreams, shapes, ks, coords, q_vals

    def decompress(self, path=None, coordinates=None, strings=None, shape=None, k=None, q_vals=None):
        if path:
            coordinates, strings, shape, k, q_vals = self.load_bitstream(path)
            dev = next(self.parameters()).device
            coordinates = [torch.cat([torch.zeros((c.shape[0], 1), dtype=c.dtype), c], dim=1).to(dev) for c in coordinates]
            q_vals = [qv.to(dev) for qv in q_vals]
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

    # -- bitstream container (model.py:253-385) ------------------------------------------------------
    def save_bitstream(self, path, blocks_coordinates, blocks_strings, blocks_shapes, blocks_k, blocks_q):
        """int32 n_blocks; per block: shape, len(points), q0, q1 (float64), len(y string), len(z string), k1, k2, k3,
        then the three payloads -- the reference's field order (model.py:272-303)."""
        out = [struct.pack(">i", len(blocks_coordinates))]
        for points, strings, shape, k, q in zip(blocks_coordinates, blocks_strings, blocks_shapes, blocks_k, blocks_q):
            assert len(strings) == 2 and all(len(s) == 1 for s in strings) and len(k) == 3 and all(len(v) == 1 for v in k), \
                "one batch item per block (model.py:345-353 reads exactly two strings and three counts)"
            points_bitstream = self.gpcc_encode(points, path)
            out.append(b"".join(struct.pack(">i", int(v)) for v in shape))
            out.append(struct.pack(">i", len(points_bitstream)))
            out.append(struct.pack(">dd", float(q[0, 0]), float(q[0, 1])))
            out.extend(struct.pack(">i", len(s[0])) for s in strings)
            out.extend(struct.pack(">i", int(v[0])) for v in k)
            out.append(points_bitstream)
            out.extend(s[0] for s in strings)
        with open(path, "wb") as f:
            f.write(b"".join(out))

    def load_bitstream(self, path):
        """Inverse of `save_bitstream` (model.py:314-385); coordinates come back as (n,3) int32 CPU rows."""
        with open(path, "rb") as f:
            data = f.read()
        pos = 0

        def take(fmt):
            nonlocal pos
            vals = struct.unpack_from(fmt, data, pos)
            pos += struct.calcsize(fmt)
            return vals

        def take_bytes(n):
            nonlocal pos
            if n < 0 or pos + n > len(data):
                raise ValueError(f"{path}: truncated bitstream")
            b = data[pos:pos + n]
            pos += n
            return b

        try:
            (num_blocks,) = take(">i")
            coords, strings, shapes, ks, qs = [], [], [], [], []
            for _ in range(num_blocks):
                shape0, len_points = take(">II")
                q0, q1 = take(">dd")
                len_y, len_z, k1, k2, k3 = take(">IIIII")
                points_bitstream = take_bytes(len_points)
                strings.append([[take_bytes(len_y)], [take_bytes(len_z)]])
                coords.append(self.gpcc_decode(points_bitstream, path))
                shapes.append([int(shape0)])
                ks.append([[int(k1)], [int(k2)], [int(k3)]])
                qs.append(torch.tensor([q0, q1], dtype=torch.float64).reshape(1, 2).float())
        except struct.error:
            raise ValueError(f"{path}: truncated bitstream") from None
        return coords, strings, shapes, ks, qs

    def gpcc_encode(self, points, directory=None):
        """Lossless bytes of the latent coordinates (rows [b,x,y,z], b = 0, stride-8 lattice).  Native octree coder in
        place of the reference's tmc3 subprocess (model.py:388-431); `directory` is accepted and unused (no temp files)."""
        pts = points.detach().to("cpu", torch.int32).numpy()
        if pts.shape[0] and (pts[:, 0] != 0).any():
            raise ValueError("gpcc_encode: a block holds one batch item (model.py:147-161)")
        return host_coder.octree_encode(pts[:, 1:4], stride=8)

    def gpcc_decode(self, bin, directory=None):
        """Points of a `gpcc_encode` stream as (n,3) int32 xyz rows, like model.py:445-486 (the caller prepends the
        batch column, model.py:216-220); Morton order -- the decoder re-derives its canonical row order anyway."""
        return torch.from_numpy(host_coder.octree_decode(bytes(bin)))

    @staticmethod
    def finalize(coords, feats):
        """8-bit colour rounding + [xyz|rgb] rows (model.py:247-250)."""
        if feats.is_cuda and feats.dtype == torch.float32 and coords.dtype == torch.int32 and not (torch.is_grad_enabled() and feats.requires_grad):
            out = torch.empty((feats.shape[0], 3 + feats.shape[1]), dtype=torch.float32, device=feats.device)
            so._lib.call("pcg_finalize", so.ptr(coords.contiguous()), so.ptr(feats.contiguous()), feats.shape[1], feats.shape[0], so.ptr(out), so.stream())
            return out
        rgb = torch.clamp(torch.round(feats * 255), 0.0, 255.0) / 255
        return torch.cat([coords[:, 1:4].to(rgb.dtype), rgb], dim=1)
