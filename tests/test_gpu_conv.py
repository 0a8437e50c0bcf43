"""CUDA sparse convolution (fp32 SIMT and tcgen05 bf16) vs the fp32 oracle, on identical inputs.
Tolerances (north-star): fp32 path 1e-5 relative L2; bf16 tensor-core path 1e-2 relative L2."""
import importlib

import numpy as np
import pytest
import torch

from oracle import me_ref, native, synth
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu


def setup(dev, pkg, seed, ts=1, scale=1):
    pts, _ = synth.small_cloud(seed=seed, radius=9.0 * scale, centre=(16 * scale,) * 3, n_samples=6000 * scale * scale)
    C = np.concatenate([np.zeros((len(pts), 1), np.int32), pts * ts], axis=1)
    cs = pkg.sparse_ops.CoordSet(torch.from_numpy(C).to(dev), ts)
    return C, cs


def oracle_conv(C_in, C_out, Fm, W, b, ksize, ts, transposed, act=0):
    nbr = native.kernel_map(C_in, C_out, ksize, ts, transposed)
    return native.conv(Fm, nbr, W, b, act)


SHAPES = [  # (cin, cout, ksize, stride, transposed)
    (4, 128, 5, 2, False), (128, 128, 5, 2, False), (128, 128, 5, 1, False), (128, 64, 3, 1, False),
    (64, 1, 3, 1, False), (32, 16, 3, 1, False), (16, 1, 3, 1, False), (32, 3, 1, 1, False),
    (128, 192, 3, 1, False), (192, 192, 3, 2, False), (192, 256, 3, 1, False), (1, 1, 3, 2, False),
    (192, 192, 2, 2, True), (128, 128, 5, 2, True), (128, 32, 5, 2, True), (3, 5, 3, 1, False), (40, 48, 3, 1, False),
]


@pytest.mark.parametrize("cin,cout,ksize,stride,transposed", SHAPES)
def test_conv_matches_oracle(pkg, dev, cin, cout, ksize, stride, transposed):
    so = pkg.sparse_ops
    ts = 2 if transposed else 1
    C, cs = setup(dev, pkg, seed=cin + cout, ts=ts)
    out_set = cs.generated(ksize, stride) if transposed else cs.strided(stride)
    C_out = out_set.coords.cpu().numpy()
    rng = np.random.default_rng(cin * 7 + cout)
    Fm = rng.standard_normal((len(C), cin)).astype(np.float32)
    K = ksize ** 3
    W = (rng.standard_normal((K, cin, cout)) / np.sqrt(cin * K) * 3).astype(np.float32)
    b = rng.standard_normal((1, cout)).astype(np.float32)
    ref = oracle_conv(C, C_out, Fm, W, b, ksize, (ts // stride) if transposed else ts, transposed, act=1)
    Wt = torch.from_numpy(W if K > 1 else W[0]).to(dev)
    modes = [("f32", 2e-5)]
    if so.tc_eligible(K, cin, cout):
        modes.append(("tc", 1e-2))
        modes.append(("auto", 1e-2))          # narrow heads take the register-gather kernels (dot / mma32) here
    for mode, tol in modes:
        prev = so.set_conv_mode(mode)
        try:
            out = so.sparse_conv(torch.from_numpy(Fm).to(dev), cs, out_set, Wt, torch.from_numpy(b).to(dev), ksize,
                                 transposed, act=so.ACT_RELU)
        finally:
            so.set_conv_mode(prev)
        err = rel_l2(out.cpu().numpy(), ref)
        assert err < tol, f"{mode}: rel-L2 {err:.3e}"


def test_tc_path_is_exact_on_bf16_representable_inputs(pkg, dev):
    """Integer-valued operands are exact in bf16 and the sums exact in fp32: the tensor-core path must then
    agree with the fp32 oracle bit for bit (pins operand layout / descriptors, not just a tolerance)."""
    so = pkg.sparse_ops
    C, cs = setup(dev, pkg, seed=11)
    rng = np.random.default_rng(0)
    for cin, cout, ksize in ((64, 32, 3), (128, 128, 3), (192, 256, 3), (32, 16, 3), (16, 1, 3), (8, 24, 3), (4, 32, 5),
                            (40, 48, 3), (64, 1, 3), (3, 5, 3), (128, 1, 3), (24, 9, 3), (32, 1, 5), (20, 16, 5)):
        Fm = rng.integers(-4, 5, (len(C), cin)).astype(np.float32)
        W = rng.integers(-2, 3, (ksize ** 3, cin, cout)).astype(np.float32)
        b = rng.integers(-3, 4, (1, cout)).astype(np.float32)
        ref = oracle_conv(C, C, Fm, W, b, ksize, 1, False)
        for mode in ("tc", "auto"):
            prev = so.set_conv_mode(mode)
            try:
                out = so.sparse_conv(torch.from_numpy(Fm).to(dev), cs, cs, torch.from_numpy(W).to(dev),
                                     torch.from_numpy(b).to(dev), ksize, False)
            finally:
                so.set_conv_mode(prev)
            assert np.array_equal(out.cpu().numpy(), ref), f"{mode} cin={cin} cout={cout}: max |d| = {np.abs(out.cpu().numpy() - ref).max()}"


def test_two_tiles_per_cta_kernel_equals_single_tile_kernel(pkg, dev):
    """k_conv_tc_mt (two output tiles per CTA sharing each weight slice; taken by large grids) forced on small inputs:
    bit-exact on integer operands against the fp32 oracle, and bit-identical to the one-tile kernel on random operands
    (same per-element accumulation order), including an odd number of tiles and strided / transposed maps."""
    so = pkg.sparse_ops
    lib = pkg._lib.load()
    rng = np.random.default_rng(3)
    C, cs = setup(dev, pkg, seed=21, scale=2)
    assert (len(C) + 127) // 128 >= 9
    down = cs.strided(2)
    up = down.generated(2, 2)
    cases = [(cs, cs, 3, False, 128, 64), (cs, cs, 3, False, 192, 256), (cs, cs, 3, False, 64, 32), (cs, down, 5, False, 128, 128),
             (down, up, 2, True, 192, 192), (cs, cs, 1, False, 512, 128),
             (cs, cs, 3, False, 32, 16), (cs, cs, 3, False, 16, 1), (cs, cs, 3, False, 64, 1), (cs, down, 5, False, 4, 32)]   # 64 / 32-byte rows
    prev_mode = so.set_conv_mode("tc")
    try:
        for in_set, out_set, ksize, transposed, cin, cout in cases:
            K = ksize ** 3
            Fi = rng.integers(-4, 5, (in_set.n, cin)).astype(np.float32)
            Wi = rng.integers(-2, 3, (K, cin, cout)).astype(np.float32)
            bi = rng.integers(-3, 4, (1, cout)).astype(np.float32)
            Fr = torch.from_numpy(rng.standard_normal((in_set.n, cin)).astype(np.float32)).to(dev)
            Wr = torch.from_numpy((rng.standard_normal((K, cin, cout)) * 0.1).astype(np.float32)).to(dev)
            wi = torch.from_numpy(Wi if K > 1 else Wi[0]).to(dev)
            wr = Wr if K > 1 else Wr[0]
            res = {}
            for tiles in (0, 1):            # 0 = never (one tile per CTA), 1 = always
                prev = lib.pcg_tc_set_mt_min_tiles(tiles)
                try:
                    res[tiles] = (so.sparse_conv(torch.from_numpy(Fi).to(dev), in_set, out_set, wi, torch.from_numpy(bi).to(dev), ksize, transposed),
                                  so.sparse_conv(Fr, in_set, out_set, wr, None, ksize, transposed, act=so.ACT_RELU))
                finally:
                    lib.pcg_tc_set_mt_min_tiles(prev)
            ts_o = out_set.ts if transposed else in_set.ts
            ref = oracle_conv(in_set.coords.cpu().numpy(), out_set.coords.cpu().numpy(), Fi, Wi, bi, ksize, ts_o, transposed)
            tag = f"k{ksize} {cin}->{cout} transposed={transposed}"
            assert np.array_equal(res[1][0].cpu().numpy(), ref), tag
            assert torch.equal(res[0][0], res[1][0]), tag
            assert torch.equal(res[0][1], res[1][1]), tag
    finally:
        so.set_conv_mode(prev_mode)


@pytest.mark.parametrize("mode", ["f32", "tc"])
def test_conv_is_run_to_run_and_order_deterministic(pkg, dev, mode):
    """Fact 9: encoder and decoder must reproduce h_s bit-for-bit.  Same inputs -> same bits; and a row's result
    does not depend on which rows share its tile (permute the rows, compare coordinate-keyed)."""
    so = pkg.sparse_ops
    C, cs = setup(dev, pkg, seed=5)
    rng = np.random.default_rng(1)
    cin, cout = 64, 64
    Fm = torch.from_numpy(rng.standard_normal((len(C), cin)).astype(np.float32)).to(dev)
    W = torch.from_numpy((rng.standard_normal((27, cin, cout)) * 0.1).astype(np.float32)).to(dev)
    prev = so.set_conv_mode(mode)
    try:
        a = so.sparse_conv(Fm, cs, cs, W, None, 3, False)
        b = so.sparse_conv(Fm, cs, cs, W, None, 3, False)
        perm = torch.from_numpy(rng.permutation(len(C))).to(dev)
        cs2 = so.CoordSet(cs.coords[perm].contiguous(), 1)
        c = so.sparse_conv(Fm[perm].contiguous(), cs2, cs2, W, None, 3, False)
    finally:
        so.set_conv_mode(prev)
    assert torch.equal(a, b)
    assert torch.equal(a[perm], c)


def test_gdn_matches_reference_expression(pkg, dev):
    so = pkg.sparse_ops
    torch.manual_seed(0)
    for n, c in ((1000, 128), (77, 32), (5, 8)):
        x = torch.randn(n, c)
        beta = torch.rand(c) + 0.5
        gamma = torch.rand(c, c) * 0.1
        norm = torch.nn.functional.conv1d(torch.abs(x.T.unsqueeze(0)), gamma.reshape(c, c, 1), beta)[0].T
        for inverse in (False, True):
            ref = x * norm if inverse else x * (1.0 / norm)
            for mode, tol in (("f32", 1e-5), ("auto", 3e-3)):       # auto: the norm is one bf16 product (pcg_gdn_fwd_bf16)
                prev = so.set_conv_mode(mode)
                try:
                    o = so.gdn(x.to(dev), beta.to(dev), gamma.to(dev), inverse)
                finally:
                    so.set_conv_mode(prev)
                assert rel_l2(o.cpu().numpy(), ref.numpy()) < tol, (mode, n, c, inverse)
                shadow = getattr(o, "_pcg_bf16", None)
                if mode == "auto" and c in (32, 128):                # the bf16 operand image of the next layer comes out of the same kernel
                    assert shadow is not None and torch.equal(shadow[0], o.to(torch.bfloat16))


def test_channelwise_conv_k9(pkg, dev):
    so = pkg.sparse_ops
    C = synth.random_cloud(3, 700, 14)
    cs = so.CoordSet(torch.from_numpy(C).to(dev), 1)
    Fm = torch.randn(len(C), 4)
    w = torch.rand(729, 1)
    mod = me_ref.MinkowskiChannelwiseConvolution(4, kernel_size=9, dimension=3)
    mod.kernel = torch.nn.Parameter(w)
    ref = mod(me_ref.SparseTensor(Fm, torch.from_numpy(C))).F.detach()
    out = so.channelwise_conv(Fm.to(dev), cs, w.to(dev), 9).cpu()
    assert rel_l2(out.numpy(), ref.numpy()) < 1e-5


def test_topk_selection_consistent_with_fp32_logits(pkg, dev):
    """Occupancy head (conv k3 + ReLU + conv k3 -> 1 logit) on bf16 tensor cores, then top-k (transforms.py:146-160,
    228-254): the logits match the fp32 oracle to 1e-2 relative L2, and every selection difference lies inside the
    numerical error band around the oracle's k-th logit -- reduced precision can only flip near-ties."""
    so = pkg.sparse_ops
    C, cs = setup(dev, pkg, seed=21, ts=2)
    cand = cs.generated(5, 2)
    Cc = cand.coords.cpu().numpy()
    n = len(Cc)
    rng = np.random.default_rng(3)
    X = rng.standard_normal((n, 32)).astype(np.float32)
    W1 = (rng.standard_normal((27, 32, 16)) / np.sqrt(32 * 27) * 2).astype(np.float32)
    b1 = (rng.standard_normal((1, 16)) * 0.1).astype(np.float32)
    W2 = (rng.standard_normal((27, 16, 1)) / np.sqrt(16 * 27) * 2).astype(np.float32)
    b2 = np.zeros((1, 1), np.float32)
    nbr = native.kernel_map(Cc, Cc, 3, 1, False)
    ref = native.conv(native.conv(X, nbr, W1, b1, 1), nbr, W2, b2, 0)[:, 0]
    t = lambda a: torch.from_numpy(a).to(dev)
    prev = so.set_conv_mode("tc")
    try:
        h = so.sparse_conv(t(X), cand, cand, t(W1), t(b1), 3, False, act=so.ACT_RELU, out="bf16")
        assert h.dtype == torch.bfloat16
        logits = so.sparse_conv(h, cand, cand, t(W2), t(b2), 3, False)[:, 0]
    finally:
        so.set_conv_mode(prev)
    mine = logits.cpu().numpy()
    assert rel_l2(mine, ref) < 1e-2
    k = n // 5
    mask = so.topk_mask(logits, cand.coords, [k]).cpu().numpy()
    assert mask.sum() == k
    thr = np.sort(ref)[-k]
    band = 2.0 * np.abs(mine - ref).max()
    assert (ref[mask] >= thr - band).all() and (ref[~mask] <= thr + band).all()
    exact = np.zeros(n, bool)
    exact[np.argsort(-ref, kind="stable")[:k]] = True
    print(f"top-k agreement with the fp32 selection: {(mask & exact).sum() / k:.4f} (error band {band:.2e}, logit std {ref.std():.2e})")


@pytest.mark.parametrize("cin,cmid", [(32, 16), (128, 64), (64, 32)])
def test_fused_occupancy_head_matches_oracle_and_two_layer_path(pkg, dev, cin, cmid):
    """sparse_ops.occupancy_head (second head layer GEMM-first out of the first layer's tcgen05 epilogue + pcg_gather_sum) against the
    fp32 oracle (1e-2 relative L2, north-star) and against the two-convolution path it replaces (which rounds the hidden row to bf16:
    the fused form must be at least as close to the oracle); integer-valued operands must come out exactly; run-to-run identical."""
    so = pkg.sparse_ops
    C, cs = setup(dev, pkg, seed=23, ts=2)
    cand = cs.generated(5, 2)
    Cc = cand.coords.cpu().numpy()
    n = len(Cc)
    rng = np.random.default_rng(cin + cmid)
    X = rng.standard_normal((n, cin)).astype(np.float32)
    W1 = (rng.standard_normal((27, cin, cmid)) / np.sqrt(cin * 27) * 2).astype(np.float32)
    b1 = (rng.standard_normal((1, cmid)) * 0.1).astype(np.float32)
    W2 = (rng.standard_normal((27, cmid, 1)) / np.sqrt(cmid * 27) * 2).astype(np.float32)
    b2 = np.full((1, 1), 0.25, np.float32)
    nbr = native.kernel_map(Cc, Cc, 3, 1, False)
    ref = native.conv(native.conv(X, nbr, W1, b1, 1), nbr, W2, b2, 0)[:, 0]
    t = lambda a: torch.from_numpy(a).to(dev)
    assert so.head_fusable(3, t(W1), 3, t(W2))
    fused = so.occupancy_head(t(X), cand, t(W1), t(b1), t(W2), t(b2))
    assert fused.shape == (n, 1)
    again = so.occupancy_head(t(X), cand, t(W1), t(b1), t(W2), t(b2))
    assert torch.equal(fused, again)
    prev = so.set_conv_mode("tc")
    try:
        h = so.sparse_conv(t(X), cand, cand, t(W1), t(b1), 3, False, act=so.ACT_RELU, out="bf16")
        two = so.sparse_conv(h, cand, cand, t(W2), t(b2), 3, False)[:, 0].cpu().numpy()
    finally:
        so.set_conv_mode(prev)
    mine = fused[:, 0].cpu().numpy()
    e_fused, e_two = rel_l2(mine, ref), rel_l2(two, ref)
    print(f"fused head {cin}->{cmid}->1: rel L2 {e_fused:.2e} (two-layer bf16 path {e_two:.2e})")
    assert e_fused < 1e-2 and e_fused <= e_two * 1.05 + 1e-6
    # small integers: every product and partial sum is exact in bf16 x bf16 -> fp32, so the result must equal the oracle bit for bit
    Xi = rng.integers(-2, 3, (n, cin)).astype(np.float32)
    W1i = rng.integers(-1, 2, (27, cin, cmid)).astype(np.float32)
    W2i = rng.integers(-1, 2, (27, cmid, 1)).astype(np.float32)
    b1i = rng.integers(-3, 4, (1, cmid)).astype(np.float32)
    refi = native.conv(native.conv(Xi, nbr, W1i, b1i, 1), nbr, W2i, b2 * 4, 0)[:, 0]
    gi = so.occupancy_head(t(Xi), cand, t(W1i), t(b1i), t(W2i), t(b2 * 4))[:, 0].cpu().numpy()
    assert np.array_equal(gi, refi)


@pytest.mark.parametrize("cin,cout,kind", [(128, 128, "surface"), (128, 32, "surface"), (64, 32, "scattered"), (64, 64, "surface")])
def test_gen_conv_transpose_cells_matches_oracle(pkg, dev, cin, cout, kind):
    """Kernel-5/stride-2 generative transposed convolution on the octree-cell path (GEMM-first + gather-reduce) vs the
    fp32 oracle on the same coordinates, compared row by row."""
    so = pkg.sparse_ops
    ts = 2
    if kind == "surface":
        C, cs = setup(dev, pkg, seed=cin + cout, ts=ts)
    else:
        C = synth.random_cloud(3, 800, 50, batch=1)
        C[:, 1:] *= ts
        cs = so.CoordSet(torch.from_numpy(C).to(dev), ts)
    out_set = cs.generated_cells()
    C_out = out_set.coords.cpu().numpy()
    rng = np.random.default_rng(cin + 3 * cout)
    Fm = rng.standard_normal((len(C), cin)).astype(np.float32)
    W = (rng.standard_normal((125, cin, cout)) / np.sqrt(cin * 125) * 6).astype(np.float32)
    b = rng.standard_normal((1, cout)).astype(np.float32)
    ref = oracle_conv(C, C_out, Fm, W, b, 5, 1, True, act=1)
    out = so.gen_conv_transpose_cells(torch.from_numpy(Fm).to(dev), cs, out_set, torch.from_numpy(W).to(dev),
                                      torch.from_numpy(b).to(dev), act=so.ACT_RELU, out="both")
    err = rel_l2(out.cpu().numpy(), ref)
    assert err < 1e-2, f"rel-L2 {err:.3e}"
    shadow = out._pcg_bf16[0].float().cpu().numpy()
    assert rel_l2(shadow, ref) < 1.2e-2
    # run-to-run bit reproducibility (fact 9)
    again = so.gen_conv_transpose_cells(torch.from_numpy(Fm).to(dev), cs, out_set, torch.from_numpy(W).to(dev),
                                        torch.from_numpy(b).to(dev), act=so.ACT_RELU, out="f32")
    assert torch.equal(out, again)


@pytest.mark.parametrize("cin,cout,ksize,stride,transposed", [(8, 16, 3, 1, False), (16, 8, 5, 2, False), (8, 8, 2, 2, True),
                                                              (16, 16, 5, 2, True), (4, 24, 3, 2, False)])
def test_conv_gradients_match_oracle_autograd(pkg, dev, cin, cout, ksize, stride, transposed):
    """Training path: d(input), d(weight), d(bias) of the CUDA sparse convolution (fp32 mode) vs torch autograd through the
    CPU oracle's convolution on the same kernel map."""
    import importlib
    so = pkg.sparse_ops
    ag = importlib.import_module(pkg.__name__ + ".autograd")
    ts = 2 if transposed else 1
    C, cs = setup(dev, pkg, seed=3 * cin + cout, ts=ts)
    out_set = cs.generated(ksize, stride) if transposed else cs.strided(stride)
    C_out = out_set.coords.cpu().numpy()
    rng = np.random.default_rng(cin + cout)
    K = ksize ** 3
    Fm = torch.from_numpy(rng.standard_normal((len(C), cin)).astype(np.float32))
    W = torch.from_numpy((rng.standard_normal((K, cin, cout)) / np.sqrt(cin * K) * 3).astype(np.float32))
    b = torch.from_numpy(rng.standard_normal((1, cout)).astype(np.float32))
    G = torch.from_numpy(rng.standard_normal((len(C_out), cout)).astype(np.float32))
    # oracle: plain torch ops over the oracle's kernel map
    maps = me_ref.kernel_map(C, C_out, ksize, (ts // stride) if transposed else ts, transposed)
    Fr, Wr, br = Fm.clone().requires_grad_(True), W.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ref = torch.relu(me_ref.sparse_conv_features(Fr, maps, Wr, br, len(C_out)))
    (ref * G).sum().backward()
    prev = so.set_conv_mode("f32")
    try:
        Fd, Wd, bd = (t.clone().to(dev).requires_grad_(True) for t in (Fm, W, b))
        out = ag.sparse_conv_autograd(Fd, cs, out_set, Wd, bd, ksize, transposed, so.ACT_RELU)
        (out * G.to(dev)).sum().backward()
    finally:
        so.set_conv_mode(prev)
    assert rel_l2(out.detach().cpu().numpy(), ref.detach().numpy()) < 2e-5
    for name, got, want in (("d_input", Fd.grad, Fr.grad), ("d_weight", Wd.grad, Wr.grad), ("d_bias", bd.grad, br.grad)):
        err = rel_l2(got.cpu().numpy(), want.numpy())
        assert err < 5e-5, f"{name}: rel-L2 {err:.3e}"


@pytest.mark.parametrize("cin,cout,ksize,stride", [(128, 128, 5, 2), (192, 192, 2, 2), (24, 40, 3, 2), (16, 8, 3, 1)])
def test_conv_transpose_onto_given_coordinates_matches_oracle(pkg, dev, cin, cout, ksize, stride):
    """ME4: `MinkowskiConvolutionTranspose` (non-generative) -- the transposed convolution onto a coordinate set the CALLER gives
    (a SparseTensor or a bare (N,4) tensor), here a random half of the generative children plus voxels no parent reaches
    (their rows are bias only).  Same arithmetic as ME3 on another output set (no call site in the reference)."""
    ME = importlib.import_module(pkg.__name__ + ".MinkowskiEngine")
    so = pkg.sparse_ops
    ts_in = 2 * stride
    C, cs = setup(dev, pkg, seed=cin + ksize, ts=ts_in)
    ts_out = ts_in // stride
    rng = np.random.default_rng(cin + cout)
    children = me_ref.generate_coords(C, ts_out, ksize)
    keep = children[rng.random(len(children)) < 0.5]
    far = np.array([[0, 400, 400, 400], [0, 402 * ts_out, 0, 0]], np.int32) * np.array([1, ts_out, ts_out, ts_out], np.int32)
    C_out = np.concatenate([keep, far]).astype(np.int32)
    K = ksize ** 3
    Fm = rng.standard_normal((len(C), cin)).astype(np.float32)
    W = (rng.standard_normal((K, cin, cout)) / np.sqrt(cin * K) * 3).astype(np.float32)
    b = rng.standard_normal((1, cout)).astype(np.float32)
    ref = oracle_conv(C, C_out, Fm, W, b, ksize, ts_out, True)
    assert np.allclose(ref[-2:], b)                                     # unreachable voxels: bias only
    conv = ME.MinkowskiConvolutionTranspose(in_channels=cin, out_channels=cout, kernel_size=ksize, stride=stride, bias=True, dimension=3).to(dev)
    with torch.no_grad():
        conv.kernel.copy_(torch.from_numpy(W))
        conv.bias.copy_(torch.from_numpy(b))
        x = ME.SparseTensor(torch.from_numpy(Fm).to(dev), _coord_set=cs)
        target = ME.SparseTensor(features=torch.zeros((len(C_out), 1), device=dev), coordinates=torch.from_numpy(C_out).to(dev),
                                 tensor_stride=ts_out, device=dev)
        modes = [("f32", 2e-5)] + ([("auto", 1e-2)] if so.tc_eligible(K, cin, cout) else [])
        for mode, tol in modes:
            prev = so.set_conv_mode(mode)
            try:
                out_a = conv(x, target)                                  # coordinates as a SparseTensor
                out_b = conv(x, torch.from_numpy(C_out).to(dev))         # coordinates as an (N,4) tensor
            finally:
                so.set_conv_mode(prev)
            for out in (out_a, out_b):
                assert out.tensor_stride == [ts_out] * 3 and np.array_equal(out.C.cpu().numpy(), C_out)
                err = rel_l2(out.F.cpu().numpy(), ref)
                assert err < tol, f"{mode}: rel-L2 {err:.3e}"
        with pytest.raises(ValueError):
            conv(x)                                                      # the output set is not optional
        wrong = ME.SparseTensor(features=torch.zeros((len(C_out), 1), device=dev), coordinates=torch.from_numpy(C_out).to(dev),
                                tensor_stride=ts_in, device=dev)
        if stride > 1:
            with pytest.raises(RuntimeError):
                conv(x, wrong)                                           # wrong tensor stride


@pytest.mark.parametrize("cin,cout,ksize,stride,transposed", [(128, 128, 3, 1, False), (128, 128, 5, 2, False), (192, 256, 3, 1, False),
                                                              (192, 192, 2, 2, True), (128, 32, 5, 2, True), (64, 16, 3, 1, False),
                                                              (128, 64, 3, 1, False), (40, 48, 3, 1, False), (128, 192, 3, 1, False),
                                                              (64, 1, 3, 1, False), (32, 16, 3, 1, False), (16, 1, 3, 1, False),
                                                              (4, 128, 5, 2, False), (32, 3, 1, 1, False)])
def test_weight_gradient_on_tensor_cores(pkg, dev, cin, cout, ksize, stride, transposed):
    """pcg_conv_bwd_weight_tc (tcgen05, MN-major operands taken from the gathered row tiles, sliced + fixed-order reduction):
    bit-exact against the fp32 kernel on integer-valued operands (pins the MN-major descriptors / swizzle / TMEM lane mapping),
    <= 1e-2 relative L2 on random operands (bf16 rounding of F and G), and bit-reproducible run to run; then the same through
    autograd in "auto" mode (forward on the tensor cores, input gradient on the inverse map, weight gradient by this kernel)."""
    so = pkg.sparse_ops
    ag = importlib.import_module(pkg.__name__ + ".autograd")
    lib = pkg._lib.load()
    from importlib import import_module
    L = import_module(pkg.__name__ + "._lib")
    ts = 2 if transposed else 1
    C, cs = setup(dev, pkg, seed=cin + 3 * cout, ts=ts, scale=2)
    out_set = cs.generated(ksize, stride) if transposed else cs.strided(stride)
    K = ksize ** 3
    assert ag.bwd_weight_tc_eligible(K, cin, cout)
    nbr, tmask = cs.kernel_map(out_set, ksize, transposed)
    rng = np.random.default_rng(cin + cout)

    def both(Fm, G):
        Fd, Gd = torch.from_numpy(Fm).to(dev), torch.from_numpy(G).to(dev)
        ref = torch.empty((K, cin, cout), device=dev)
        L.call("pcg_conv_bwd_weight_f32", L.ptr(Fd), cin, L.ptr(nbr), nbr.shape[1], L.ptr(tmask), K, out_set.n, L.ptr(Gd), cout, L.ptr(ref), L.stream())
        fb, gb = ag._bf16_rows64(Fd), ag._bf16_rows64(Gd)
        ws = torch.empty(lib.pcg_conv_bwd_weight_tc_scratch_bytes(K, cin, cout, out_set.n), dtype=torch.uint8, device=dev)
        outs = []
        for _ in range(2):
            dw = torch.full((K, cin, cout), float("nan"), device=dev)
            L.call("pcg_conv_bwd_weight_tc", L.ptr(fb), cin, L.ptr(gb), cout, L.ptr(nbr), nbr.shape[1], L.ptr(tmask), K, out_set.n,
                   L.ptr(ws), ws.numel(), L.ptr(dw), L.stream())
            outs.append(dw)
        assert torch.equal(outs[0], outs[1]), "not reproducible"
        return ref, outs[0]

    ref, got = both(rng.integers(-3, 4, (cs.n, cin)).astype(np.float32), rng.integers(-2, 3, (out_set.n, cout)).astype(np.float32))
    assert torch.isfinite(got).all() and float(ref.abs().max()) > 0
    assert torch.equal(ref, got), f"integer operands: max |d| = {float((ref - got).abs().max())}"
    ref, got = both(rng.standard_normal((cs.n, cin)).astype(np.float32), rng.standard_normal((out_set.n, cout)).astype(np.float32))
    err = rel_l2(got.cpu().numpy(), ref.cpu().numpy())
    assert err < 1e-2, f"random operands: rel-L2 {err:.3e}"
    # through autograd, tensor-core mode end to end
    Fm = torch.from_numpy(rng.standard_normal((cs.n, cin)).astype(np.float32))
    W = torch.from_numpy((rng.standard_normal((K, cin, cout)) / np.sqrt(cin * K) * 3).astype(np.float32))
    G = torch.from_numpy(rng.standard_normal((out_set.n, cout)).astype(np.float32))
    grads = {}
    for mode in ("f32", "auto"):
        prev = so.set_conv_mode(mode)
        try:
            Fd, Wd = Fm.clone().to(dev).requires_grad_(True), W.clone().to(dev).requires_grad_(True)
            out = ag.sparse_conv_autograd(Fd, cs, out_set, Wd, None, ksize, transposed, so.ACT_NONE)      # (an activation's mask would flip with the forward rounding)
            (out * G.to(dev)).sum().backward()
            grads[mode] = (Fd.grad.cpu().numpy(), Wd.grad.cpu().numpy())
        finally:
            so.set_conv_mode(prev)
    assert rel_l2(grads["auto"][0], grads["f32"][0]) < 1.5e-2 and rel_l2(grads["auto"][1], grads["f32"][1]) < 1.5e-2


@pytest.mark.parametrize("cin,cout", [(32, 16), (16, 1), (64, 1), (128, 16)])
def test_weight_gradient_offsets_inner_variant(pkg, dev, cin, cout, monkeypatch):
    """k_conv_bwd_weight_tc_kin (narrow outputs on large candidate sets: offsets inner, the G tile gathered once per tile, one TMEM
    accumulator per offset), forced on a small input: bit-exact against the fp32 kernel on integer operands and bit-identical to the
    offsets-outer tensor-core kernel's result up to the summation grouping (compared through the fp32 kernel on integers)."""
    so = pkg.sparse_ops
    ag = importlib.import_module(pkg.__name__ + ".autograd")
    lib = pkg._lib.load()
    L = importlib.import_module(pkg.__name__ + "._lib")
    C, cs = setup(dev, pkg, seed=cin + cout, scale=2)
    K = 27
    nbr, tmask = cs.kernel_map(cs, 3, False)
    rng = np.random.default_rng(cin * 3 + cout)
    Fd = torch.from_numpy(rng.integers(-3, 4, (cs.n, cin)).astype(np.float32)).to(dev)
    Gd = torch.from_numpy(rng.integers(-2, 3, (cs.n, cout)).astype(np.float32)).to(dev)
    ref = torch.empty((K, cin, cout), device=dev)
    L.call("pcg_conv_bwd_weight_f32", L.ptr(Fd), cin, L.ptr(nbr), nbr.shape[1], L.ptr(tmask), K, cs.n, L.ptr(Gd), cout, L.ptr(ref), L.stream())
    fb, gb = ag._bf16_rows64(Fd), ag._bf16_rows64(Gd)
    monkeypatch.setenv("PCG_BWT_KIN_MIN_TILES", "1")
    ws = torch.empty(lib.pcg_conv_bwd_weight_tc_scratch_bytes(K, cin, cout, cs.n), dtype=torch.uint8, device=dev)
    outs = []
    for _ in range(2):
        dw = torch.full((K, cin, cout), float("nan"), device=dev)
        L.call("pcg_conv_bwd_weight_tc", L.ptr(fb), cin, L.ptr(gb), cout, L.ptr(nbr), nbr.shape[1], L.ptr(tmask), K, cs.n, L.ptr(ws), ws.numel(), L.ptr(dw), L.stream())
        outs.append(dw)
    torch.cuda.synchronize()
    assert ws.numel() >= 148 * K * cin * cout * 4            # the offsets-inner variant was selected (one slice per SM)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], ref), f"max |d| = {float((outs[0] - ref).abs().max())}"
