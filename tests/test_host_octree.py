"""Latent-coordinate side channel and bitstream container -- host code, no GPU needed.

The reference codes the stride-8 coordinates with the external tmc3 binary (model/model.py:388-486) and has no tests or
vectors for it; the contract checked here is the one its call sites rely on: a point SET survives the round trip exactly,
and the container holds the fields of model/model.py:272-303 in that order."""
import ctypes
import importlib
import os
import re
import struct

import numpy as np
import pytest
import torch

from oracle import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def hc():
    import __graft_entry__ as g
    g.build()
    return importlib.import_module("unified-point-cloud-compression_b200.host_coder")


def as_set(P):
    return set(map(tuple, np.asarray(P, dtype=np.int64).reshape(-1, 3).tolist()))


def test_host_library_exports_every_declared_symbol(hc):
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "pcg_host.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(pcg_[a-z0-9_]+)\s*\(", text)))
    assert len(names) == 6
    lib = ctypes.CDLL(hc.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in pcg_host.h but not exported"


@pytest.mark.parametrize("case", ["empty", "single", "duplicates", "negative", "dense", "random", "surface8", "wide"])
def test_octree_round_trip_is_lossless(hc, case):
    rng = np.random.default_rng(7)
    stride = 1
    if case == "empty":
        P = np.zeros((0, 3), np.int32)
    elif case == "single":
        P = np.array([[5, -3, 7]], np.int32)
    elif case == "duplicates":
        P = np.array([[0, 0, 0], [1, 1, 1], [0, 0, 0], [1, 1, 1], [9, 0, 2]], np.int32)
    elif case == "negative":
        P = rng.integers(-1000, 1000, (5000, 3)).astype(np.int32)
    elif case == "dense":
        g = np.arange(16)
        P = np.stack(np.meshgrid(g, g, g, indexing="ij"), axis=-1).reshape(-1, 3).astype(np.int32)
    elif case == "random":
        P = rng.integers(0, 1 << 10, (20000, 3)).astype(np.int32)
    elif case == "surface8":                      # what the codec feeds it: stride-8 parents of a 10-bit surface frame
        pts, _ = synth.frame10(seed=0)
        P, stride = np.unique(pts // 8 * 8, axis=0).astype(np.int32), 8
    else:                                          # 21 bits per axis
        P = np.array([[0, 0, 0], [(1 << 21) - 1, 5, 0], [7, (1 << 21) - 1, (1 << 21) - 1]], np.int32)
    data = hc.octree_encode(P, stride)
    back = hc.octree_decode(data)
    assert back.dtype == np.int32 and back.shape[1] == 3
    assert as_set(back) == as_set(P) and len(back) == len(as_set(P))
    assert hc.octree_encode(back, stride) == data                     # canonical: the bytes depend on the set only
    assert hc.octree_encode(P[rng.permutation(len(P))], stride) == data
    if case == "surface8":
        assert 8 * len(data) / len(P) < 3.0, "a surface's parents should cost < 3 bits per point"


def test_octree_rejects_bad_input_and_corrupt_streams(hc):
    with pytest.raises(ValueError, match="lattice"):
        hc.octree_encode(np.array([[0, 0, 0], [8, 8, 4]]), 8)
    with pytest.raises(ValueError, match="lattice|extent"):
        hc.octree_encode(np.array([[0, 0, 0], [1 << 21, 0, 0]]), 1)
    P = np.random.default_rng(1).integers(0, 256, (3000, 3)).astype(np.int32)
    data = hc.octree_encode(P)
    with pytest.raises(ValueError):
        hc.octree_decode(b"nope" + data[4:])
    with pytest.raises(ValueError):
        hc.octree_decode(data[:10])
    with pytest.raises(ValueError):
        hc.octree_decode(data[: len(data) // 2])                      # payload overrun / count mismatch
    bad = bytearray(data)
    bad[8:12] = struct.pack("<I", len(as_set(P)) + 5)                 # header count disagrees with the tree
    with pytest.raises(ValueError):
        hc.octree_decode(bytes(bad))


def test_container_fields_follow_the_reference_layout(hc, tmp_path):
    """save_bitstream / load_bitstream: parsed here field by field in the order of model/model.py:272-303."""
    from tests.helpers import model_config
    pkg = importlib.import_module("unified-point-cloud-compression_b200")
    model_mod = importlib.import_module(pkg.__name__ + ".model")
    net = model_mod.UnifiedModel(model_config(width=32))             # parameter container only: nothing runs on a device
    rng = np.random.default_rng(3)
    blocks = []
    for b in range(3):
        xyz = np.unique(rng.integers(0, 40, (500 + 100 * b, 3)), axis=0).astype(np.int32) * 8
        coords = torch.from_numpy(np.concatenate([np.zeros((len(xyz), 1), np.int32), xyz], axis=1))
        strings = [[rng.bytes(100 + b)], [rng.bytes(7 + b)]]
        blocks.append((coords, strings, [11 + b], [[1000 + b], [300 + b], [70 + b]], torch.tensor([[0.25 * b, 1.0 - 0.25 * b]])))
    path = str(tmp_path / "frame.bin")
    net.save_bitstream(path, *[[blk[i] for blk in blocks] for i in range(5)])
    data = open(path, "rb").read()
    pos = 4
    assert struct.unpack(">i", data[:4])[0] == 3
    for coords, strings, shape, k, q in blocks:
        shape0, len_pts = struct.unpack_from(">ii", data, pos); pos += 8
        q0, q1 = struct.unpack_from(">dd", data, pos); pos += 16
        len_y, len_z, k1, k2, k3 = struct.unpack_from(">iiiii", data, pos); pos += 20
        assert [shape0] == shape and [[k1], [k2], [k3]] == k and (q0, q1) == (float(q[0, 0]), float(q[0, 1]))
        assert as_set(hc.octree_decode(data[pos:pos + len_pts])) == as_set(coords[:, 1:].numpy()); pos += len_pts
        assert data[pos:pos + len_y] == strings[0][0]; pos += len_y
        assert data[pos:pos + len_z] == strings[1][0]; pos += len_z
    assert pos == len(data)
    c2, s2, sh2, k2_, q2 = net.load_bitstream(path)
    for (coords, strings, shape, k, q), c, s, sh, kk, qq in zip(blocks, c2, s2, sh2, k2_, q2):
        assert as_set(c.numpy()) == as_set(coords[:, 1:].numpy()) and s == strings and sh == shape and kk == k
        assert torch.allclose(qq, q)
    with open(path, "wb") as f:
        f.write(data[:-3])
    with pytest.raises(ValueError, match="truncated"):
        net.load_bitstream(path)
