"""AnalysisTransform / SparseSynthesisTransform -- mirror of /root/reference/model/transforms.py.

Same sub-module names (down_conv_1..3, up_1..3, predict_1..3, color_conv, prune, down_conv) so the
reference's checkpoints load.  Differences in *how* (not what):
  * point counts per level come from the coordinate sets (one bincount for batched input) instead of
    a python loop of `.item()` syncs (transforms.py:47-64);
  * the ReLU between the two occupancy-head convolutions is fused into the first conv's epilogue;
  * top-k is a device radix select (no python loop / torch.topk / nonzero, transforms.py:228-254) and
    the prune uses the mask directly -- the int64-key `torch.isin` of transforms.py:257-282 recomputes
    exactly that mask;
  * `down_conv` is only ever used for its output coordinates (transforms.py:219-220, model.py:227-229):
    `coords_down()` derives them with the stride-coordinates kernel, no convolution.
"""
import os

import torch
import torch.nn as nn

from .. import MinkowskiEngine as ME
from .. import sparse_ops as so
from .blocks import MinkowskiGDN


def _conv(cin, cout, k, s=1, bias=True):
    return ME.MinkowskiConvolution(in_channels=cin, out_channels=cout, kernel_size=k, stride=s, bias=bias, dimension=3)


def _gen(cin, cout, k=5, s=2):
    return ME.MinkowskiGenerativeConvolutionTranspose(in_channels=cin, out_channels=cout, kernel_size=k, stride=s,
                                                      bias=True, dimension=3)


def count_per_batch(x):
    """Number of rows per batch item, as a python list (transforms.py:47-64)."""
    b = x.C[:, 0]
    if x.C.shape[0] == 0:
        return []
    counts = torch.bincount(b.long())
    return [int(c) for c in counts.tolist() if c > 0]


class AnalysisTransform(nn.Module):
    def __init__(self, config):
        super().__init__()
        c_in, n1, n2, n3, n4 = (config[k] for k in ("C_in", "N1", "N2", "N3", "N4"))
        self.down_conv_1 = nn.Sequential(_conv(c_in, n1, 5, 2), MinkowskiGDN(n1))
        self.down_conv_2 = nn.Sequential(_conv(n1, n2, 5, 2), MinkowskiGDN(n2))
        self.down_conv_3 = nn.Sequential(_conv(n2, n3, 5, 2), MinkowskiGDN(n3), _conv(n3, n4, 5, 1))

    count_per_batch = staticmethod(count_per_batch)

    def forward(self, x):
        if getattr(x.coord_set, "single_batch", False):
            k = [[x.F.shape[0]]]                     # known single batch item (block entry): no bincount / sync
        else:
            k = [count_per_batch(x)]
        single = len(k[0]) == 1                      # one batch item: level sizes are just row counts
        x = self.down_conv_1(x)
        k.append([x.F.shape[0]] if single else count_per_batch(x))
        x = self.down_conv_2(x)
        k.append([x.F.shape[0]] if single else count_per_batch(x))
        x = self.down_conv_3(x)
        k.reverse()
        return x, k


class SparseSynthesisTransform(nn.Module):
    def __init__(self, config):
        super().__init__()
        c_out, n1, n2, n3, n4 = (config[k] for k in ("C_out", "N1", "N2", "N3", "N4"))
        self.up_1 = nn.Sequential(_conv(n4, n3, 5, 1), MinkowskiGDN(n3, inverse=True), _gen(n3, n2))
        self.up_2 = nn.Sequential(MinkowskiGDN(n2, inverse=True), _gen(n2, n1))
        self.up_3 = nn.Sequential(MinkowskiGDN(n1, inverse=True), _gen(n1, n1 // 4))
        self.color_conv = nn.Sequential(_conv(n1 // 4, c_out, 1, 1))
        self.predict_1 = nn.Sequential(_conv(n2, n2 // 2, 3), ME.MinkowskiReLU(inplace=False), _conv(n2 // 2, 1, 3))
        self.predict_2 = nn.Sequential(_conv(n1, n1 // 2, 3), ME.MinkowskiReLU(inplace=False), _conv(n1 // 2, 1, 3))
        self.predict_3 = nn.Sequential(_conv(n1 // 4, n4 // 8, 3), ME.MinkowskiReLU(inplace=False), _conv(n4 // 8, 1, 3))
        self.prune = ME.MinkowskiPruning()
        self.down_conv = ME.MinkowskiConvolution(in_channels=1, out_channels=1, kernel_size=3, stride=2, dimension=3)
        for up in (self.up_1, self.up_2, self.up_3):
            # the ~20x larger candidate set is written ONCE, as the bf16 operand image the occupancy head reads; the prune turns
            # the surviving rows back into fp32 (an fp32 copy of every candidate would double the up-sampling kernel's writes
            # for rows of which 80-90 % are dropped two kernels later).  PCG_CAND_F32=1 restores fp32 + bf16 ("both") for A/B runs.
            up[-1].emit = "both" if os.environ.get("PCG_CAND_F32") == "1" else "bf16"

    # -- helpers -----------------------------------------------------------------------------
    @staticmethod
    def _head(head, x):
        """conv -> ReLU -> conv with the ReLU folded into the first epilogue (inference path)."""
        first, last = head[0], head[2]
        if torch.is_grad_enabled() and x.F.requires_grad:
            return head(x)
        if so.head_fusable(first.kernel_size, first.kernel, last.kernel_size, last.kernel):
            o = so.occupancy_head(x.F, x.coord_set, first.kernel, first.bias, last.kernel, last.bias, packed=first._packed)
            return ME.SparseTensor(o, _coord_set=x.coord_set)
        h = so.sparse_conv(x.F, x.coord_set, x.coord_set, first.kernel, first.bias, first.kernel_size, False,
                           act=so.ACT_RELU, packed=first._packed, out="bf16")    # hidden rows only feed the next conv
        o = so.sparse_conv(h, x.coord_set, x.coord_set, last.kernel, last.bias, last.kernel_size, False,
                           packed=last._packed)
        return ME.SparseTensor(o, _coord_set=x.coord_set)

    @staticmethod
    def _topk_prediction(prediction, k):
        return so.topk_mask(prediction.F[:, 0], prediction.C, [int(v) for v in k])

    def _prune_tensor(self, x, mask, n_keep=None, reorder=False):
        """Rows under `mask`.  reorder=True additionally puts the survivors in Z-order (row order is free in ME
        semantics): they are the parents of the next generative level, whose candidates inherit the parents'
        order class by class -- without it spatial coherence (= L2 hit rate of every gather) decays level by level."""
        do_reorder = reorder and not os.environ.get("PCG_NO_REORDER")
        if torch.is_grad_enabled() and x.F.requires_grad:
            oc, _ = so.prune_rows(x.C, None, mask, n_keep=n_keep)
            of = x.F[mask]                              # differentiable row gather (training)
        elif n_keep is not None and (do_reorder or x.F.dtype == torch.bfloat16):
            oc, of = so.prune_rows_reordered(x.C, x.F, mask, n_keep, reorder=do_reorder)      # prune (+ Z-order) in one pass over the feature rows
            return ME.SparseTensor(of, _coord_set=so.CoordSet(oc, x.coord_set.ts))
        else:
            oc, of = so.prune_rows(x.C, x.F.float() if x.F.dtype == torch.bfloat16 else x.F, mask, n_keep=n_keep)
        if do_reorder and oc.shape[0] > 1:
            perm = so.sort_rows(oc, morton=True)
            oc, of = oc[perm].contiguous(), of[perm].contiguous()
        return ME.SparseTensor(of, _coord_set=so.CoordSet(oc, x.coord_set.ts))

    @staticmethod
    def coords_down(coord_set, times=1):
        """Output coordinate set of `down_conv` applied `times` times (a6)."""
        for _ in range(times):
            coord_set = coord_set.strided(2)
        return coord_set

    # -- forward -----------------------------------------------------------------------------
    def forward(self, y, coords=None, k=None, keep_sets=None):
        """keep_sets (measurement only, bench.py --topk-from encoder): the true coordinate sets at strides 4, 2, 1.  The head and
        the top-k still run, but each level is pruned with the TRUE occupancy -- what a trained model's top-k selects -- so the
        next level sees the candidate count a deployment sees instead of the scattered set random-init logits keep."""
        x = y
        predictions = []
        for lvl, (up, head, k_level) in enumerate(((self.up_1, self.predict_1, k[0]), (self.up_2, self.predict_2, k[1]),
                                                   (self.up_3, self.predict_3, k[2]))):
            x = up(x)
            pred = self._head(head, x)
            predictions.append(pred)
            mask = self._topk_prediction(pred, k_level)
            if keep_sets is not None:
                mask = keep_sets[lvl].lookup(x.C) >= 0
            n_avail = [x.F.shape[0]] if len(k_level) == 1 else count_per_batch(x)
            n_keep = sum(min(int(kk), na) for kk, na in zip(k_level, n_avail)) if len(n_avail) == len(k_level) else None
            x = self._prune_tensor(x, mask, n_keep, reorder=head is not self.predict_3)
        x = self.color_conv(x)
        if coords is None:
            return x
        with torch.no_grad():
            p1 = self.coords_down(coords.coord_set, 1)
            p2 = self.coords_down(p1, 1)
            ones = lambda cs: ME.SparseTensor(torch.ones((cs.n, 1), device=cs.coords.device), _coord_set=cs)
        return x, [ones(p2), ones(p1), coords], predictions
