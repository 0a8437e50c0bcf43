"""Pin the oracle's sparse-conv geometry and weight-index convention against dense torch convolutions
(independent of our own restatement of MinkowskiEngine; SURVEY.md §4 "oracle tier 0")."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import me_ref, native, synth


def densify(C, feats, size):
    g = torch.zeros((feats.shape[1], size, size, size), dtype=feats.dtype)
    g[:, C[:, 1], C[:, 2], C[:, 3]] = feats.t()
    return g


def dense_weight(W, k):
    """(K,Cin,Cout) with offset index ix + k*iy + k*k*iz -> conv3d weight (Cout,Cin,kx,ky,kz)."""
    K, cin, cout = W.shape
    return W.reshape(k, k, k, cin, cout).permute(4, 3, 2, 1, 0).contiguous()   # [iz,iy,ix] -> [.., ix, iy, iz]


@pytest.mark.parametrize("k", [3, 5])
def test_conv_stride1_matches_dense(k):
    torch.manual_seed(0)
    C = synth.random_cloud(0, 600, 12)
    feats = torch.randn(len(C), 3, dtype=torch.float64)
    W = torch.randn(k ** 3, 3, 4, dtype=torch.float64)
    maps = me_ref.kernel_map(C, C, k, 1)
    out = me_ref.sparse_conv_features(feats, maps, W, None, len(C))
    dense = F.conv3d(densify(C, feats, 12)[None], dense_weight(W, k), padding=k // 2)[0]
    ref = dense[:, C[:, 1], C[:, 2], C[:, 3]].t()
    assert torch.allclose(out, ref, atol=1e-10)


def test_conv_stride2_matches_dense():
    torch.manual_seed(1)
    C = synth.random_cloud(1, 500, 12)
    feats = torch.randn(len(C), 2, dtype=torch.float64)
    W = torch.randn(125, 2, 3, dtype=torch.float64)
    C_out = me_ref.stride_coords(C, 1, 2)
    assert (C_out[:, 1:] % 2 == 0).all() and len(np.unique(me_ref.pack_keys(C_out))) == len(C_out)
    maps = me_ref.kernel_map(C, C_out, 5, 1)
    out = me_ref.sparse_conv_features(feats, maps, W, None, len(C_out))
    dense = F.conv3d(densify(C, feats, 12)[None], dense_weight(W, 5), padding=2)[0]     # full-res, then sample even sites
    ref = dense[:, C_out[:, 1], C_out[:, 2], C_out[:, 3]].t()
    assert torch.allclose(out, ref, atol=1e-10)


def test_generative_transpose_k5_matches_dense():
    torch.manual_seed(2)
    Cf = synth.random_cloud(2, 120, 5)
    C = Cf.copy()
    C[:, 1:] = C[:, 1:] * 2 + 2                       # coarse grid (tensor stride 2), margin for the -2 offsets
    feats = torch.randn(len(C), 2, dtype=torch.float64)
    W = torch.randn(125, 2, 3, dtype=torch.float64)
    C_out = me_ref.generate_coords(C, 1, 5)
    maps = me_ref.kernel_map(C, C_out, 5, 1, transposed=True)
    assert sum(len(m[0]) for m in maps) == len(C) * 125          # every (input, offset) pair lands
    out = me_ref.sparse_conv_features(feats, maps, W, None, len(C_out))
    # dense: out[2i + (kidx - 2)] += x[i] * w[kidx]  == conv_transpose3d(stride 2, padding 2) on the coarse index grid
    coarse = torch.zeros((2, 7, 7, 7), dtype=torch.float64)
    ci = C[:, 1:] // 2
    coarse[:, ci[:, 0], ci[:, 1], ci[:, 2]] = feats.t()
    wt = W.reshape(5, 5, 5, 2, 3).permute(3, 4, 2, 1, 0).contiguous()     # (Cin,Cout,kx,ky,kz)
    dense = F.conv_transpose3d(coarse[None], wt, stride=2, padding=2)[0]      # size 2*7-1 = 13
    ref = dense[:, C_out[:, 1], C_out[:, 2], C_out[:, 3]].t()
    assert torch.allclose(out, ref, atol=1e-10)
    # and nothing outside the generated set
    mask = torch.ones(dense.shape[1:], dtype=torch.bool)
    mask[C_out[:, 1], C_out[:, 2], C_out[:, 3]] = False
    assert dense[:, mask].abs().max() < 1e-12


def test_generative_transpose_k2_children():
    C = synth.random_cloud(3, 50, 4)
    C[:, 1:] *= 2
    out = me_ref.generate_coords(C, 1, 2)
    assert len(out) == 8 * len(C)                      # the 8 octree children, disjoint per parent
    assert (np.floor_divide(out[:, 1:], 2) * 2 == np.repeat(C[:, 1:], 8, axis=0)).all()


def test_native_cpu_baseline_matches_python_oracle():
    torch.manual_seed(3)
    pts, _ = synth.small_cloud()
    C = np.concatenate([np.zeros((len(pts), 1), np.int32), pts], axis=1)
    feats = torch.randn(len(C), 8)
    for k, stride in ((3, 1), (5, 2)):
        W, b = torch.randn(k ** 3, 8, 16), torch.randn(1, 16)
        C_out = C if stride == 1 else me_ref.stride_coords(C, 1, stride)
        assert (C_out == (C if stride == 1 else native.stride_coords(C, stride))).all()
        nbr = native.kernel_map(C, C_out, k, 1)
        assert (nbr == me_ref.neighbour_table(C, C_out, k, 1)).all()
        ref = me_ref.sparse_conv_features(feats, me_ref.kernel_map(C, C_out, k, 1), W, b, len(C_out)).numpy()
        out = native.conv(feats.numpy(), nbr, W.numpy(), b.numpy())
        assert np.abs(out - ref).max() < 1e-3 * np.abs(ref).max()
    G = me_ref.generate_coords(C_out, 1, 5)
    assert (G == native.generate_coords(C_out, 5, 1)).all()


def test_interpolation_on_and_off_grid():
    C = torch.tensor([[0, 0, 0, 0], [0, 2, 0, 0], [0, 0, 2, 0]], dtype=torch.int32)
    Fm = torch.tensor([[1.0], [3.0], [5.0]])
    x = me_ref.SparseTensor(Fm, C, tensor_stride=2)
    on = x.features_at_coordinates(torch.tensor([[0, 2.0, 0, 0], [0, 4.0, 0, 0]]))
    assert on[0, 0] == 3.0 and on[1, 0] == 0.0
    mid = x.features_at_coordinates(torch.tensor([[0, 1.0, 0, 0], [0, 1.0, 1.0, 0.0]]))
    assert abs(mid[0, 0] - 2.0) < 1e-6                       # halfway between 1 and 3
    assert abs(mid[1, 0] - (0.25 * 1 + 0.25 * 3 + 0.25 * 5)) < 1e-6   # absent corner contributes 0


def test_sparse_tensor_dedupes_keeping_first():
    C = torch.tensor([[0, 1, 1, 1], [0, 2, 2, 2], [0, 1, 1, 1]], dtype=torch.int32)
    x = me_ref.SparseTensor(torch.tensor([[1.0], [2.0], [3.0]]), C)
    assert x.C.tolist() == [[0, 1, 1, 1], [0, 2, 2, 2]] and x.F[:, 0].tolist() == [1.0, 2.0]
    c, f = me_ref.sparse_quantize(torch.tensor([[0, 1.7, 1.2, 1.9], [0, 1.1, 1.0, 1.5]]), torch.tensor([[1.0], [2.0]]), quantization_size=1.0)
    assert c.tolist() == [[0, 1, 1, 1]] and f.tolist() == [[1.0]]


def test_head_second_layer_gemm_first_identity():
    """The occupancy head's second layer taken GEMM-first (sparse_ops.occupancy_head: s[k][p] = relu(h[p]) . w2[k] in the first layer's
    epilogue, then logit[o] = b2 + sum_k s[k][nbr[k][o]], offsets ascending) is the same function as conv(relu(conv(x)))
    (model/transforms.py:146-160): checked here on the CPU oracle, with integer operands bit for bit."""
    pts, _ = synth.small_cloud(seed=5)
    C = np.concatenate([np.zeros((len(pts), 1), np.int32), pts], axis=1)
    n = len(C)
    nbr = native.kernel_map(C, C, 3, 1, False)                      # [27][n], -1 where absent
    rng = np.random.default_rng(11)
    for integer in (False, True):
        if integer:
            X = rng.integers(-2, 3, (n, 32)).astype(np.float32)
            W1 = rng.integers(-1, 2, (27, 32, 16)).astype(np.float32)
            W2 = rng.integers(-1, 2, (27, 16, 1)).astype(np.float32)
            b1 = rng.integers(-3, 4, (1, 16)).astype(np.float32)
        else:
            X = rng.standard_normal((n, 32)).astype(np.float32)
            W1 = (rng.standard_normal((27, 32, 16)) * 0.1).astype(np.float32)
            W2 = (rng.standard_normal((27, 16, 1)) * 0.1).astype(np.float32)
            b1 = (rng.standard_normal((1, 16)) * 0.1).astype(np.float32)
        b2 = np.full((1, 1), 1.0, np.float32)
        h = native.conv(X, nbr, W1, b1, 1)                          # first layer + ReLU
        ref = native.conv(h, nbr, W2, b2, 0)[:, 0]
        planes = (h @ W2[:, :, 0].T).T.astype(np.float32)           # s[k][p]
        acc = np.zeros(n, np.float32)
        for k in range(27):
            idx = nbr[k][:n]
            acc += np.where(idx >= 0, planes[k][np.maximum(idx, 0)], 0.0).astype(np.float32)
        mine = acc + b2[0, 0]
        if integer:
            assert np.array_equal(mine, ref)
        else:
            assert np.abs(mine - ref).max() <= 1e-5 * max(1.0, np.abs(ref).max())
