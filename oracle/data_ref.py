"""CPU restatement of the reference's tensor-level data path (TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py).

  slice_into_cubes   /root/reference/data/dataloader.py:167-208
  random_rotate      /root/reference/data/transform.py:78-103 (RandomRotate.transform, for a given rotation matrix)
  rotation_matrix_3d /root/reference/data/transform.py:105-124
  color_jitter       /root/reference/data/transform.py:33-55 = torchvision.transforms.ColorJitter.forward over the colours
                     as a 3 x N x 1 image; the third-party dependency itself (torchvision, importable here) is called, with
                     the op order / factors passed in instead of drawn.
Parity unpinned by the reference (it has no tests); pinned here by calling the dependency it wraps.
"""
import torch

from . import me_ref


def slice_into_cubes(points_tensor, colors_tensor, cube_size=64):
    min_boundary = torch.tensor([0, 0, 0])
    cube_indices = ((points_tensor - min_boundary) / cube_size).floor().long()
    unique_cube_indices, inverse_indices = torch.unique(cube_indices, dim=0, return_inverse=True)
    cubes = []
    for idx in range(unique_cube_indices.size(0)):
        mask = inverse_indices == idx
        shift = unique_cube_indices[idx] * cube_size
        cube = {"points": points_tensor[mask] - shift, "colors": colors_tensor[mask], "offset": shift,
                "num_points": torch.tensor(int(mask.sum()))}
        if cube["num_points"] > 0:
            cubes.append(cube)
    return cubes


def rotation_matrix_3d(phi, theta):
    r_x = torch.tensor([[1, 0, 0], [0, torch.cos(phi), -torch.sin(phi)], [0, torch.sin(phi), torch.cos(phi)]])
    r_y = torch.tensor([[torch.cos(theta), 0, torch.sin(theta)], [0, 1, 0], [-torch.sin(theta), 0, torch.cos(theta)]])
    return torch.mm(r_y, r_x)


def random_rotate(points, colors, rotation, block_size, crop=False):
    rotated = torch.mm(points - block_size / 2, rotation.T) + block_size / 2
    if crop:
        valid = ((rotated >= 0) & (rotated < block_size)).all(dim=1)
        rotated, colors = rotated[valid], colors[valid]
    points_q, colors_q = me_ref.sparse_quantize(rotated, colors, quantization_size=1.0)
    return points_q, colors_q, rotated


def color_jitter(colors, order, factors):
    import torchvision.transforms.functional as F
    img = colors.T.unsqueeze(-1)
    b, c, s, h = factors
    for fn_id in order:
        if fn_id == 0:
            img = F.adjust_brightness(img, b)
        elif fn_id == 1:
            img = F.adjust_contrast(img, c)
        elif fn_id == 2:
            img = F.adjust_saturation(img, s)
        else:
            img = F.adjust_hue(img, h)
    return img.squeeze(-1).T
