"""ColorJitter / RandomRotate -- mirror of /root/reference/data/transform.py:9-124 on CUDA tensors.

Same classes, constructor arguments, `sample` dict convention ("points", "colors", optional "cubes") and the SAME random
draws from torch's CPU generator in the same order (so a seeded reference run and a seeded run here use identical jitter
factors and angles); the arithmetic runs in csrc/pcg_data.cu.  Samples must live on the GPU: there is no CPU path.
"""
import math

import torch

from .. import sparse_ops as so
from .._lib import call, load, ptr, stream


def build_transforms(config):
    """transform.py:9-31: transforms in sorted-key order of the yaml section."""
    transforms = []
    if not config:
        return transforms
    for _, setting in sorted(config.items()):
        key = setting["key"]
        if key == "RandomRotate":
            transforms.append(RandomRotate(setting["block_size"]))
        elif key == "ColorJitter":
            transforms.append(ColorJitter())
        else:
            raise ValueError(f"Transform {key} not defined.")
    return transforms


def _need_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must live on a CUDA device: the graft has no CPU path")


class _PerCube:
    def __call__(self, sample):
        if "cubes" in sample:
            for idx, cube in enumerate(sample["cubes"]):
                sample["cubes"][idx] = self.transform(cube)
            return sample
        return self.transform(sample)


class ColorJitter(_PerCube):
    """transform.py:33-55: torchvision.transforms.ColorJitter(brightness=0.3, contrast=0.3, saturation=0.3, hue=0.3) over the
    colours of a sample seen as one 3 x N x 1 image."""

    def __init__(self, brightness=0.3, contrast=0.3, saturation=0.3, hue=0.3):
        self.brightness = (max(0.0, 1 - brightness), 1 + brightness)
        self.contrast = (max(0.0, 1 - contrast), 1 + contrast)
        self.saturation = (max(0.0, 1 - saturation), 1 + saturation)
        self.hue = (-hue, hue)

    def get_params(self):
        """The draws of torchvision's ColorJitter.get_params, in its order: randperm(4), then one uniform_ per factor."""
        order = torch.randperm(4).tolist()
        factors = [float(torch.empty(1).uniform_(lo, hi)) for lo, hi in (self.brightness, self.contrast, self.saturation, self.hue)]
        return order, factors

    def transform(self, sample):
        colors = sample["colors"]
        _need_cuda(colors, "sample['colors']")
        order, factors = self.get_params()
        sample["colors"] = self.apply(colors, order, factors)
        return sample

    @staticmethod
    def apply(colors, order, factors):
        c = colors.to(torch.float32).contiguous()
        out = torch.empty_like(c)
        ws = torch.empty(load().pcg_color_jitter_scratch_bytes(), dtype=torch.uint8, device=c.device)
        call("pcg_color_jitter", ptr(c), c.shape[0], *[int(o) for o in order], *[float(f) for f in factors], ptr(out), ptr(ws), stream())
        return out


class RandomRotate(_PerCube):
    """transform.py:57-124: rotate about the cube centre by a random roll and pitch, (optionally crop to the cube), floor and
    keep the first point of every voxel.  "points" come back as int32 voxel coordinates like ME.utils.sparse_quantize's."""

    def __init__(self, block_size, crop=False):
        self.block_size = block_size
        self.crop = crop

    @staticmethod
    def rotation_matrix_3d(phi, theta):
        """transform.py:105-124: R = R_y(theta) @ R_x(phi), built in fp32 on the host like the reference."""
        phi, theta = torch.as_tensor(phi, dtype=torch.float32).reshape(1), torch.as_tensor(theta, dtype=torch.float32).reshape(1)
        r_x = torch.tensor([[1, 0, 0], [0, torch.cos(phi), -torch.sin(phi)], [0, torch.sin(phi), torch.cos(phi)]])
        r_y = torch.tensor([[torch.cos(theta), 0, torch.sin(theta)], [0, 1, 0], [-torch.sin(theta), 0, torch.cos(theta)]])
        return torch.mm(r_y, r_x)

    def transform(self, sample):
        phi = torch.rand(1) * 2 * 3.141592653589793
        theta = torch.rand(1) * 2 * 3.141592653589793
        points, colors = self.rotate(sample["points"], sample["colors"], self.rotation_matrix_3d(phi, theta))
        sample["points"], sample["colors"] = points, colors
        return sample

    def rotate(self, points, colors, rotation):
        _need_cuda(points, "sample['points']")
        p = points.to(torch.float32).contiguous()
        n = p.shape[0]
        rot = rotation.to(torch.float32).contiguous().cpu().numpy()
        coords = torch.empty((max(n, 1), 4), dtype=torch.int32, device=p.device)
        keep = torch.empty(max(n, 1), dtype=torch.uint8, device=p.device) if self.crop else None
        call("pcg_rotate_floor", ptr(p), n, rot.ctypes.data, float(self.block_size) / 2, float(self.block_size) if self.crop else 0.0,
             ptr(coords), ptr(keep), stream())
        coords = coords[:n]
        if self.crop:
            sel = keep[:n].bool()
            coords, colors = coords[sel].contiguous(), colors[sel]
        if coords.shape[0] == 0:
            return coords[:, 1:4], colors[:0]
        uniq, index, _ = so.unique_first(coords)
        return uniq[:, 1:4].contiguous(), colors[index.long()]
