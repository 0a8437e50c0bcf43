"""Frame sharding host logic (codec.py) incl. a world_size-2 gloo run on CPU: the schedule is identical on every
rank, covers every unit exactly once, and the gathered result equals the single-rank result."""
import importlib
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _codec():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    return importlib.import_module("unified-point-cloud-compression_b200.codec")


def test_lpt_schedule_is_balanced_and_complete():
    codec = _codec()
    sizes = [1_100_000, 26_000, 640_000, 410_000, 300_000, 35_000]      # 11-bit frame: six 512-blocks
    for world in (1, 2, 4, 8):
        owner = codec.lpt_schedule(sizes, world)
        assert sorted(i for r in range(world) for i in codec.my_units(sizes, world, r)) == list(range(len(sizes)))
        loads = [sum(s for s, o in zip(sizes, owner) if o == r) for r in range(world)]
        assert max(loads) <= max(max(sizes), 1.34 * sum(sizes) / world)
    assert codec.lpt_schedule([5, 5, 5, 5], 2) == [0, 1, 0, 1]


class _FakeNet:
    """Stands in for UnifiedModel on CPU: 'codes' a frame by returning it (the GPU path has no CPU fallback)."""

    def compress(self, cloud, q, block_size=1024):
        return [cloud], None, None, None, None

    def decompress(self, coordinates=None, strings=None, shape=None, k=None, q_vals=None):
        return strings[0]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    codec = _codec()
    frames = [torch.full((100 + 37 * i, 6), float(i)) for i in range(9)]
    sc = codec.StreamCodec(_FakeNet(), None, rank=rank, world=world)
    res = sc.run(frames, keep=lambda rec: (int(rec.shape[0]), float(rec.sum())))
    if rank == 0:
        torch.save(res, out)
    dist.destroy_process_group()


def test_two_rank_gloo_run_matches_single_rank(tmp_path):
    codec = _codec()
    frames = [torch.full((100 + 37 * i, 6), float(i)) for i in range(9)]
    single = codec.StreamCodec(_FakeNet(), None).run(frames, keep=lambda rec: (int(rec.shape[0]), float(rec.sum())))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert torch.load(out) == single
