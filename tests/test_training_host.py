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


def _hook_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(6, 40), torch.nn.Tanh(), torch.nn.Linear(40, 40), torch.nn.Tanh(), torch.nn.Linear(40, 3))
    unused = torch.nn.Parameter(torch.zeros(5))                     # a parameter the loss does not reach (e.g. rescale_nn)
    params = list(net.parameters()) + [unused]
    buckets = _training().GradBuckets(params, world, bucket_bytes=1000)       # several buckets
    x = torch.randn(16, 6, generator=torch.Generator().manual_seed(100 + rank))
    for _ in range(2):                                              # hooks re-arm every step
        for p in params:
            p.grad = None
        buckets.start()
        net(x).pow(2).sum().backward()                              # all-reduces are launched from the hooks during this call
        n = buckets.finish()
    if rank == 0:
        torch.save({"grads": [p.grad for p in net.parameters()], "n": n, "n_buckets": len(buckets.buckets), "unused": unused.grad}, out)
    dist.destroy_process_group()


def test_overlapped_gradient_buckets_equal_the_mean(tmp_path):
    """GradBuckets (post-accumulate-grad hooks, asynchronous bucket all-reduce launched during backward) over a world_size-2
    gloo group: every gradient ends as the mean over the ranks; a parameter without gradient does not stall its bucket."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "h.pt")
    mp.spawn(_hook_worker, args=(2, port, out), nprocs=2, join=True)
    got = torch.load(out)
    assert got["n_buckets"] >= 2 and got["n"] >= 2 and got["unused"] is None
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(6, 40), torch.nn.Tanh(), torch.nn.Linear(40, 40), torch.nn.Tanh(), torch.nn.Linear(40, 3))
    want = None
    for rank in range(2):
        for p in net.parameters():
            p.grad = None
        x = torch.randn(16, 6, generator=torch.Generator().manual_seed(100 + rank))
        net(x).pow(2).sum().backward()
        g = [p.grad.clone() for p in net.parameters()]
        want = g if want is None else [a + b for a, b in zip(want, g)]
    for g, w in zip(got["grads"], want):
        assert torch.allclose(g, w / 2, atol=1e-6)
