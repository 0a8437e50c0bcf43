"""Frame / block sharding driver: one process per GPU, no data-path collective.

The codec is intra-only and the reference already codes the blocks of a frame independently and serially
(/root/reference/model/model.py:137-176, :225-238, no halo), so a sequence shards by frame and a frame by block.
Units are assigned largest-first (LPT) from their point counts -- a deterministic function of (sizes, world), so
every rank computes the same schedule without talking to the others.  Results (symbol buffers, reconstructions,
timings) are gathered on the host with `torch.distributed.gather_object`; nothing crosses NVLink.
"""
import torch
import torch.distributed as dist


def lpt_schedule(sizes, world):
    """Largest-processing-time-first assignment of units to `world` ranks.  Returns owner[i] for every unit.
    Ties are broken by unit index, loads by rank index: the same answer on every rank."""
    order = sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))
    load = [0] * world
    owner = [0] * len(sizes)
    for i in order:
        r = min(range(world), key=lambda j: (load[j], j))
        owner[i] = r
        load[r] += int(sizes[i])
    return owner


def my_units(sizes, world, rank):
    owner = lpt_schedule(sizes, world)
    return [i for i, r in enumerate(owner) if r == rank]


def gather_results(local, world, rank, dst=0):
    """local: {unit index: picklable result}.  Returns the merged dict on `dst`, None elsewhere."""
    if world == 1:
        return dict(local)
    bucket = [None] * world if rank == dst else None
    dist.gather_object(local, bucket, dst=dst)
    if rank != dst:
        return None
    merged = {}
    for part in bucket:
        merged.update(part)
    return merged


class StreamCodec:
    """Codes the frames of a sequence that fall to this rank.  `net` is a UnifiedModel on this rank's GPU."""

    def __init__(self, net, q, block_size=1024, rank=0, world=1):
        self.net, self.q, self.block_size = net, q, block_size
        self.rank, self.world = rank, world

    def code_frame(self, cloud):
        """compress + decompress one frame (component form, model.py:94-250) -> (N,6) reconstruction."""
        with torch.no_grad():
            strings, shapes, ks, coords, qs = self.net.compress(cloud, self.q, block_size=self.block_size)
            return self.net.decompress(coordinates=coords, strings=strings, shape=shapes, k=ks, q_vals=qs)

    def run(self, frames, sizes=None, keep=lambda rec: int(rec.shape[0])):
        """frames: sequence of (N,6) tensors or zero-argument loaders; sizes: point counts used for the schedule
        (default: len of each frame).  Returns {frame index: keep(reconstruction)} gathered on rank 0."""
        if sizes is None:
            sizes = [int(f.shape[0]) for f in frames]
        local = {}
        for i in my_units(sizes, self.world, self.rank):
            f = frames[i]() if callable(frames[i]) else frames[i]
            local[i] = keep(self.code_frame(f))
        return gather_results(local, self.world, self.rank)
