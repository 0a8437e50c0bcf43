"""GPU data path of the training step -- mirror of /root/reference/data/{transform,dataloader}.py for the pieces that touch
tensors: cube slicing, RandomRotate + de-duplication, ColorJitter.  File loading (Open3D), dataset caching
and the torch DataLoader plumbing stay with the caller."""
from .dataloader import slice_into_cubes  # noqa: F401
from .transform import ColorJitter, RandomRotate, build_transforms  # noqa: F401
