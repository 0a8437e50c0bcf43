"""compressai.ops.ops (entropy_models.py:285,310 use `ops.quantize_ste`)."""
import torch


def quantize_ste(x):
    """Rounding with a straight-through gradient: (round(x) - x).detach() + x."""
    return (torch.round(x) - x).detach() + x
