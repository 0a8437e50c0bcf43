"""Cube slicing -- mirror of /root/reference/data/dataloader.py:167-208 (`slice_into_cubes`).

The reference builds one boolean mask per cube (cubes x N work, done once per split "on cuda for faster slicing",
dataloader.py:129-134).  Here the rows are sorted once by cube index (stable radix sort, csrc/pcg_coords.cu) and cut at the
segment boundaries: same cubes in the same order (lexicographic cube index, like torch.unique(dim=0)), same row order inside
a cube (original order)."""
import torch

from .. import sparse_ops as so


def slice_into_cubes(points_tensor, colors_tensor, cube_size=64):
    """list of {"points": rows - cube origin, "colors", "offset": cube origin, "num_points"} (dataloader.py:167-208)."""
    if not points_tensor.is_cuda:
        raise RuntimeError("points must live on a CUDA device: the graft has no CPU path")
    n = points_tensor.shape[0]
    if n == 0:
        return []
    cube_idx = (points_tensor / cube_size).floor().to(torch.int32)                 # min_boundary = 0 (dataloader.py:184)
    rows = torch.cat([torch.zeros((n, 1), dtype=torch.int32, device=cube_idx.device), cube_idx], dim=1).contiguous()
    perm = so.sort_rows(rows).long()                                               # lexicographic, stable
    sorted_idx = cube_idx[perm]
    uniq, counts = torch.unique_consecutive(sorted_idx, dim=0, return_counts=True)
    counts = counts.tolist()                                                       # the one host sync: number of cubes
    pts, cols = points_tensor[perm], colors_tensor[perm]
    cubes, start = [], 0
    for c, count in enumerate(counts):
        shift = uniq[c].long() * cube_size
        cubes.append({"points": (pts[start:start + count] - shift).detach(), "colors": cols[start:start + count].detach(),
                      "offset": shift.detach(), "num_points": torch.tensor(count)})
        start += count
    return cubes
