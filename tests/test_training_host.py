"""Host logic of the data-parallel training step on CPU: bucketed gradient all-reduce over a world_size-2 gloo group
equals the single-process mean, and the q -> Lambda map follows data/q_func.py."""
import importlib
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _training():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    return importlib.import_module("unified-point-cloud-compression_b200.training")


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    ps = [torch.nn.Parameter(torch.zeros(s)) for s in ((7, 5), (300,), (2, 3, 4), (1,))]
    for p in ps:
        p.grad = torch.randn(p.shape, generator=g)
    return ps


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ps = _params(rank)
    calls = _training().allreduce_mean_(ps, world, bucket_bytes=1024)       # small buckets: several all-reduces
    if rank == 0:
        torch.save({"grads": [p.grad for p in ps], "calls": calls}, out)
    dist.destroy_process_group()


def test_bucketed_allreduce_is_the_mean(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "g.pt")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = torch.load(out)
    want = [(a.grad + b.grad) / 2 for a, b in zip(_params(0), _params(1))]
    assert got["calls"] >= 2
    for g, w in zip(got["grads"], want):
        assert torch.allclose(g, w, atol=1e-6)


def test_q_to_lambda_map():
    tr = _training()
    q = torch.tensor([[0.0, 1.0], [0.5, 0.25]])
    lam = tr.scale_q_vals(q, "quadratic", a_G=3.0, b_G=0.5, a_A=2.0, b_A=0.1)
    assert torch.allclose(lam, torch.tensor([[0.5, 2.1], [1.25, 0.225]]))
    lam = tr.scale_q_vals(q, "exponential", a_G=2.0, b_G=0.0, a_A=1.0, b_A=1.0)
    assert torch.allclose(lam, torch.tensor([[1.0, 3.0], [2.0, 2 ** 0.25 + 1.0]]))
