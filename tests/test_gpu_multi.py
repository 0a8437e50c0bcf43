"""Multi-GPU paths on real devices (skipped with fewer than 2 GPUs; `gpurun --gpus 2`): one process per GPU over NCCL.
  * training (BASELINE configs[3]): the data-parallel step's averaged gradients equal the mean of the per-shard gradients
    computed in one process -- with the all-reduce launched in buckets during backward and with the single flat all-reduce;
  * the sharded stream (configs[2]): codec.StreamCodec over 2 ranks reproduces the 1-rank run bit for bit.
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "unified-point-cloud-compression_b200"

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")]


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _setup(rank, world, port):
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    return importlib.import_module(PKG), dev


def _shard(shard, dev):
    from oracle import synth
    clouds = [synth.small_cloud(seed=10 * shard + s, radius=9.0 + s, centre=(18, 18, 18), n_samples=6000) for s in (1, 2)]
    coords = torch.cat([torch.cat([torch.full((len(p), 1), b, dtype=torch.int32), torch.from_numpy(p)], 1) for b, (p, _) in enumerate(clouds)]).to(dev)
    feats = torch.cat([torch.from_numpy(c.astype(np.float32)) for _, c in clouds]).to(dev)
    q = torch.tensor([[0.5, 0.5], [0.2, 0.8]], device=dev)
    return coords, feats, q


def _model(pkg, dev, width=32):
    from tests.helpers import model_config
    mm = importlib.import_module(pkg.__name__ + ".model")
    torch.manual_seed(0)
    return mm.UnifiedModel(model_config(width=width)).to(dev)


def _grads(pkg, dev, shard, world, overlap):
    training = importlib.import_module(pkg.__name__ + ".training")
    net = _model(pkg, dev)
    tr = training.Trainer(net, world=world, overlap=overlap)
    torch.manual_seed(1000 + shard)                      # train-mode quantisation noise: the same draw per shard everywhere
    tr.forward_backward(*_shard(shard, dev))
    return {n: p.grad.detach().cpu() for n, p in net.named_parameters() if p.grad is not None}


def _train_worker(rank, world, port, out):
    pkg, dev = _setup(rank, world, port)
    res = {ov: _grads(pkg, dev, rank, world, ov) for ov in (True, False)}
    if rank == 0:
        torch.save(res, out)
    dist.destroy_process_group()


def test_data_parallel_gradients_are_the_mean_over_ranks(tmp_path):
    out = str(tmp_path / "g.pt")
    mp.spawn(_train_worker, args=(2, _port(), out), nprocs=2, join=True)
    got = torch.load(out)
    pkg, dev = _setup(0, 1, 0)
    parts = [_grads(pkg, dev, shard, 1, False) for shard in (0, 1)]
    assert set(parts[0]) == set(parts[1]) == set(got[True]) == set(got[False])
    for name in parts[0]:
        want = (parts[0][name] + parts[1][name]) / 2
        scale = float(want.abs().max()) + 1e-12
        for ov in (True, False):
            err = float((got[ov][name] - want).abs().max()) / scale
            assert err < 1e-5, f"{name} overlap={ov}: {err:.2e}"


def _frames(dev):
    from oracle import synth
    out = []
    for seed in range(6):
        pts, rgb = synth.small_cloud(seed=seed + 1, radius=8.0 + 2 * (seed % 3), centre=(24, 24, 24), n_samples=5000 + 1500 * seed)
        out.append(torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb], axis=1)).to(dev))
    return out


def _stream(pkg, dev, rank, world, coder):
    from tests.helpers import spread_weights
    codec = importlib.import_module(pkg.__name__ + ".codec")
    net = _model(pkg, torch.device("cpu"))
    spread_weights(net)
    net = net.to(dev).eval()
    net.update()
    frames = _frames(dev)
    sc = codec.StreamCodec(net, torch.tensor([[0.5, 0.5]], device=dev), block_size=24, rank=rank, world=world, coder=coder)
    pipe = codec.FramePipeline(dev, 2, model=net)
    try:
        return sc.run(frames, keep=lambda rec: rec.cpu(), pipeline=pipe)
    finally:
        pipe.close()


def _stream_worker(rank, world, port, out):
    pkg, dev = _setup(rank, world, port)
    res = {coder: _stream(pkg, dev, rank, world, coder) for coder in ("bypass", "host")}
    if rank == 0:
        torch.save(res, out)
    dist.destroy_process_group()


def test_sharded_stream_equals_single_rank_run(tmp_path):
    out = str(tmp_path / "s.pt")
    mp.spawn(_stream_worker, args=(2, _port(), out), nprocs=2, join=True)
    got = torch.load(out)
    pkg, dev = _setup(0, 1, 0)
    for coder in ("bypass", "host"):
        one = _stream(pkg, dev, 0, 1, coder)
        assert sorted(got[coder]) == sorted(one) == list(range(6))
        for i in one:
            assert torch.equal(got[coder][i], one[i]), f"{coder}: frame {i} differs between the 2-rank and the 1-rank run"
    for i in range(6):                                   # the host range coder is lossless: same reconstruction as the bypass
        assert torch.equal(got["host"][i], got["bypass"][i])
