"""MinkowskiGDN -- mirror of /root/reference/model/blocks.py:8-57 (GDN1: |x|, no square / sqrt).

The reference transposes F, runs `F.conv1d` with gamma as a 1x1 kernel and re-wraps the result in a
brand-new SparseTensor (new coordinate manager => every kernel map is rebuilt, blocks.py:53-56).
Here one fused kernel computes  x * (1 / (beta + |x| @ gamma^T))  (or x * norm for the inverse) and
the output keeps the input's coordinate set, so cached hash tables / kernel maps survive.
"""
import torch

from .. import sparse_ops as so
from ..compressai.layers import GDN
from ..MinkowskiEngine import SparseTensor


class MinkowskiGDN(GDN):
    def __init__(self, in_channels, inverse=False, beta_min=1e-6, gamma_init=0.1, kernel_size=1):
        super().__init__(in_channels, inverse, beta_min, gamma_init)
        self.kernel_size = kernel_size

    def forward(self, x):
        beta = self.beta_reparam(self.beta)
        gamma = self.gamma_reparam(self.gamma)
        if torch.is_grad_enabled() and (x.F.requires_grad or self.beta.requires_grad and self.training):
            norm = beta + torch.abs(x.F) @ gamma.t()          # differentiable device expression (training)
            feats = x.F * norm if self.inverse else x.F * (1.0 / norm)
        else:
            feats = so.gdn(x.F, beta, gamma, self.inverse)
        return SparseTensor(feats, _coord_set=x.coord_set)
