"""compressai.layers.GDN: parameter holder the reference's MinkowskiGDN subclasses (blocks.py:8-24)."""
import torch
import torch.nn as nn

from .ops.parametrizers import NonNegativeParametrizer


class GDN(nn.Module):
    def __init__(self, in_channels, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        beta_min, gamma_init = float(beta_min), float(gamma_init)
        self.inverse = bool(inverse)
        self.beta_reparam = NonNegativeParametrizer(minimum=beta_min)
        self.beta = nn.Parameter(self.beta_reparam.init(torch.ones(in_channels)))
        self.gamma_reparam = NonNegativeParametrizer()
        self.gamma = nn.Parameter(self.gamma_reparam.init(gamma_init * torch.eye(in_channels)))

    def forward(self, x):
        raise NotImplementedError("dense (B,C,H,W) GDN is not on the reference's path; see model.blocks.MinkowskiGDN")
