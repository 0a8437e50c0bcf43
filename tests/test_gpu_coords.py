"""CUDA coordinate kernels vs the oracle: bit-exact rows, order included (first-occurrence order == ME CPU)."""
import importlib

import numpy as np
import pytest
import torch

from oracle import me_ref, synth

pytestmark = pytest.mark.gpu


def cloud(seed=0, batch=1, n=3000, extent=40):
    return synth.random_cloud(seed, n, extent, batch=batch)


def to_dev(C, dev):
    return torch.from_numpy(np.ascontiguousarray(C, dtype=np.int32)).to(dev)


def test_hash_lookup_and_duplicates(pkg, dev):
    so = pkg.sparse_ops
    C = cloud(1, batch=2)
    dup = np.concatenate([C, C[::3]], axis=0)
    uniq, index, _ = so.unique_first(to_dev(dup, dev))
    first = me_ref.unique_first(dup)
    assert (uniq.cpu().numpy() == dup[first]).all() and (index.cpu().numpy() == first).all()
    cs = so.CoordSet(to_dev(C, dev), 1)
    q = np.concatenate([C[::7], C[::11] + np.array([0, 100, 0, 0])], axis=0)
    rows = cs.lookup(to_dev(q, dev)).cpu().numpy()
    assert (rows == me_ref._Lookup(C)(q)).all()


def test_empty_and_single(pkg, dev):
    so = pkg.sparse_ops
    e = so.CoordSet(torch.zeros((0, 4), dtype=torch.int32, device=dev), 1)
    assert e.strided(2).n == 0
    one = so.CoordSet(to_dev(np.array([[0, -3, 5, 7]]), dev), 1)
    assert one.strided(2).coords.cpu().tolist() == [[0, -4, 4, 6]]          # floor for negative coordinates
    assert one.generated(2, 1).n == 8


@pytest.mark.parametrize("ts,stride", [(1, 2), (2, 2), (8, 2), (1, 4)])
def test_stride_coords(pkg, dev, ts, stride):
    so = pkg.sparse_ops
    C = cloud(2, batch=2)
    C[:, 1:] = (C[:, 1:] - 7) * ts           # include negatives
    C = C[me_ref.unique_first(C)]
    out = so.CoordSet(to_dev(C, dev), ts).strided(stride)
    ref = me_ref.stride_coords(C, ts, stride)
    assert out.ts == ts * stride and (out.coords.cpu().numpy() == ref).all()
    # the derived table maps every output coordinate to its own row
    assert (out.lookup(out.coords).cpu().numpy() == np.arange(len(ref))).all()


@pytest.mark.parametrize("ksize,stride", [(2, 2), (5, 2), (3, 1)])
def test_generate_coords(pkg, dev, ksize, stride):
    so = pkg.sparse_ops
    pts, _ = synth.small_cloud(seed=3)
    C = np.concatenate([np.zeros((len(pts), 1), np.int32), pts * 2], axis=1)
    out = so.CoordSet(to_dev(C, dev), 2).generated(ksize, stride)
    ref = me_ref.generate_coords(C, 2 // stride, ksize)
    got = out.coords.cpu().numpy()
    # the SET is ME's; the row order is libpcg's own (grouped by child parity class, pcg.h)
    assert out.n == len(ref) and (np.sort(me_ref.pack_keys(got)) == np.sort(me_ref.pack_keys(ref))).all()
    assert len(np.unique(me_ref.pack_keys(got))) == len(got)
    ts_out = 2 // stride
    cls = ((got[:, 1] // ts_out) % stride * stride + (got[:, 2] // ts_out) % stride) * stride + (got[:, 3] // ts_out) % stride
    # classes are numbered by the offset residue; parents sit on multiples of `stride`, so the child's residue equals it
    assert (np.diff(cls) >= 0).all(), "rows must be grouped by parity class, classes ascending"
    assert (out.lookup(out.coords).cpu().numpy() == np.arange(len(ref))).all()


@pytest.mark.parametrize("ksize,ts,stride,transposed", [(3, 1, 1, False), (5, 1, 2, False), (5, 2, 2, False),
                                                        (2, 2, 2, True), (5, 2, 2, True), (1, 1, 1, False)])
def test_kernel_map(pkg, dev, ksize, ts, stride, transposed):
    so = pkg.sparse_ops
    pts, _ = synth.small_cloud(seed=4)
    C = np.concatenate([np.zeros((len(pts), 1), np.int32), pts * ts], axis=1)
    cs = so.CoordSet(to_dev(C, dev), ts)
    if transposed:
        out = cs.generated(ksize, stride)
        ref = me_ref.neighbour_table(C, out.coords.cpu().numpy(), ksize, ts // stride, True)
    else:
        out = cs.strided(stride)
        ref = me_ref.neighbour_table(C, out.coords.cpu().numpy(), ksize, ts, False)
    nbr, tmask = cs.kernel_map(out, ksize, transposed)
    n = ref.shape[1]
    assert nbr.shape[1] % 128 == 0 and nbr.shape[1] >= n
    assert (nbr[:, :n].cpu().numpy() == ref).all() and (nbr[:, n:] == -1).all()
    if transposed:
        assert (ref >= 0).sum() == len(C) * ksize ** 3          # every (input, offset) pair lands exactly once
    tm = tmask.cpu().numpy().view(np.uint32)
    for t in range(tm.shape[0]):
        any_k = (ref[:, t * 128:(t + 1) * 128] >= 0).any(axis=1)
        bits = np.array([(tm[t, k >> 5] >> (k & 31)) & 1 for k in range(ksize ** 3)], bool)
        assert (bits == any_k).all()


def test_kernel_map_k9_no_tile_mask(pkg, dev):
    so = pkg.sparse_ops
    C = cloud(5, n=500, extent=14)
    cs = so.CoordSet(to_dev(C, dev), 1)
    nbr, tmask = cs.kernel_map(cs, 9, False)
    assert tmask is None and (nbr[:, :len(C)].cpu().numpy() == me_ref.neighbour_table(C, C, 9, 1)).all()


def test_morton_sort(pkg, dev):
    so = pkg.sparse_ops
    C = cloud(9, batch=2)
    C[:, 1:] -= 13
    perm = so.sort_rows(to_dev(C, dev), morton=True).cpu().numpy()

    def spread(v):
        out = np.zeros(len(v), np.uint64)
        for b in range(16):
            out |= ((v.astype(np.uint64) >> np.uint64(b)) & np.uint64(1)) << np.uint64(3 * b)
        return out
    key = (C[:, 0].astype(np.uint64) << np.uint64(48)) | (spread(C[:, 1] + 32768) << np.uint64(2)) | (spread(C[:, 2] + 32768) << np.uint64(1)) | spread(C[:, 3] + 32768)
    assert (perm == np.argsort(key, kind="stable")).all()


def test_lexsort(pkg, dev):
    so = pkg.sparse_ops
    C = cloud(6, batch=3)
    C[:, 1:] -= 20
    C = C[np.random.default_rng(0).permutation(len(C))]
    perm = so.lexsort(to_dev(C, dev)).cpu().numpy()
    assert (perm == me_ref.lexsort_rows(C)).all()


@pytest.mark.parametrize("n,c", [(0, 4), (1, 3), (5000, 128), (4099, 7)])
def test_prune_rows(pkg, dev, n, c):
    so = pkg.sparse_ops
    rng = np.random.default_rng(n)
    C = rng.integers(0, 100, (n, 4)).astype(np.int32)
    Fm = rng.standard_normal((n, c)).astype(np.float32)
    mask = rng.random(n) < 0.3
    oc, of = so.prune_rows(to_dev(C, dev), torch.from_numpy(Fm).to(dev), torch.from_numpy(mask).to(dev))
    assert (oc.cpu().numpy() == C[mask]).all() and (of.cpu().numpy() == Fm[mask]).all()


@pytest.mark.parametrize("n,c", [(1, 3), (5000, 128), (4099, 7)])
def test_prune_rows_reordered_equals_prune_then_morton_sort(pkg, dev, n, c):
    """pcg_mask_to_rows + pcg_gather_rows: the fused prune + Z-order re-sort moves the same rows to the same places as
    the order-preserving prune followed by the Morton permutation."""
    so = pkg.sparse_ops
    rng = np.random.default_rng(n + c)
    C = np.unique(rng.integers(0, 60, (n, 4)).astype(np.int32), axis=0)
    C = C[rng.permutation(len(C))]
    C[:, 0] = 0
    C = np.unique(C, axis=0)[rng.permutation(len(np.unique(C, axis=0)))]
    n = len(C)
    Fm = rng.standard_normal((n, c)).astype(np.float32)
    mask = rng.random(n) < 0.4
    if n == 1:
        mask[:] = True
    Cd, Fd, md = to_dev(C, dev), torch.from_numpy(Fm).to(dev), torch.from_numpy(mask).to(dev)
    oc, of = so.prune_rows(Cd, Fd, md)
    perm = so.sort_rows(oc, morton=True)
    rc, rf = so.prune_rows_reordered(Cd, Fd, md, int(mask.sum()))
    assert torch.equal(rc, oc[perm]) and torch.equal(rf, of[perm])
    rows = torch.from_numpy(np.nonzero(mask)[0].astype(np.int32)).to(dev)
    gc, gf = so.gather_rows(Cd, Fd, rows)
    assert torch.equal(gc, oc) and torch.equal(gf, of)


@pytest.mark.parametrize("n,k", [(10, 3), (5000, 1234), (5000, 0), (5000, 5000), (5000, 7000), (70000, 35000), (3000000, 1000001)])
def test_topk_mask(pkg, dev, n, k):
    so = pkg.sparse_ops
    rng = np.random.default_rng(k + n)
    logits = rng.standard_normal(n).astype(np.float32)
    logits[::5] = np.round(logits[::5]) + 0.0        # plenty of exact ties (+0.0: no negative zeros)
    C = np.zeros((n, 4), np.int32)
    mask = so.topk_mask(torch.from_numpy(logits).to(dev), to_dev(C, dev), [k]).cpu().numpy()
    kk = min(k, n)
    assert mask.sum() == kk
    if 0 < kk < n:
        thr = np.sort(logits)[::-1][kk - 1]
        assert (logits[mask] >= thr).all() and (logits[~mask] <= thr).all()
        ties = np.nonzero(logits == thr)[0]
        taken = mask[ties]
        assert (taken[:taken.sum()]).all()           # ties resolved towards the smaller row index


def test_topk_mask_per_batch(pkg, dev):
    so = pkg.sparse_ops
    rng = np.random.default_rng(0)
    n = 6000
    logits = rng.standard_normal(n).astype(np.float32)
    C = np.zeros((n, 4), np.int32)
    C[:, 0] = rng.integers(0, 3, n)
    ks = [100, 0, 777]
    mask = so.topk_mask(torch.from_numpy(logits).to(dev), to_dev(C, dev), ks).cpu().numpy()
    for b, k in enumerate(ks):
        sel = C[:, 0] == b
        assert mask[sel].sum() == k
        if k:
            assert logits[sel][mask[sel]].min() >= np.sort(logits[sel])[::-1][k - 1]


@pytest.mark.parametrize("n,nb", [(50000, 8), (3000, 2), (400000, 32)])
def test_topk_mask_batched_equals_per_batch_selection(pkg, dev, n, nb):
    """pcg_topk_mask_batched (all batch items in one set of radix passes) selects exactly what the single-batch path selects for
    every item -- including exact ties at the threshold (smaller row first), k = 0, and k larger than the item."""
    so = pkg.sparse_ops
    rng = np.random.default_rng(n + nb)
    logits = rng.standard_normal(n).astype(np.float32)
    logits[::3] = np.round(logits[::3] * 2) / 2 + 0.0          # many exact ties
    C = np.zeros((n, 4), np.int32)
    C[:, 0] = rng.integers(0, nb, n)
    counts = np.bincount(C[:, 0], minlength=nb)
    ks = [int(rng.integers(0, c + 1)) for c in counts]
    ks[0], ks[-1] = 0, int(counts[-1]) + 5
    lg, Cd = torch.from_numpy(logits).to(dev), to_dev(C, dev)
    got = so.topk_mask(lg, Cd, ks).cpu().numpy()
    for b, k in enumerate(ks):
        sel = np.nonzero(C[:, 0] == b)[0]
        one = so.topk_mask(lg[torch.from_numpy(sel).to(dev)], to_dev(np.zeros((len(sel), 4), np.int32), dev), [k]).cpu().numpy()
        assert got[sel].sum() == min(k, len(sel)), (b, k)
        assert np.array_equal(got[sel], one), f"batch item {b}: batched and single-item selections differ"


def test_features_at(pkg, dev):
    so = pkg.sparse_ops
    C = cloud(7, n=800, extent=10)
    C[:, 1:] *= 2
    Fm = torch.randn(len(C), 5)
    x = me_ref.SparseTensor(Fm, torch.from_numpy(C), tensor_stride=2)
    rng = np.random.default_rng(1)
    q_on = np.concatenate([C[::3], C[::5] + np.array([0, 2, 0, 0])]).astype(np.float32)
    q_off = q_on + rng.uniform(0, 2, q_on.shape).astype(np.float32) * np.array([0, 1, 1, 1], np.float32)
    cs = so.CoordSet(to_dev(C, dev), 2)
    for q, tol in ((q_on, 0.0), (q_off, 1e-5)):
        out = so.features_at(Fm.to(dev), cs, torch.from_numpy(q).to(dev)).cpu()
        ref = x.features_at_coordinates(torch.from_numpy(q))
        assert (out - ref).abs().max() <= tol


# ---------------------------------------------------------------------------- octree-cell generative path
@pytest.mark.parametrize("ts,kind", [(2, "surface"), (4, "surface"), (8, "scattered"), (2, "batch2")])
def test_generated_cells_match_oracle_set_and_generic_maps(pkg, dev, ts, kind):
    """Cell-organised kernel-5/stride-2 candidates: the SET is ME's (oracle), rows are unique, the 3x3x3 map derived
    from the cells equals the hash-built map over the same rows, and the per-cell parent lists reproduce the
    transposed kernel map."""
    so = pkg.sparse_ops
    if kind == "surface":
        pts, _ = synth.small_cloud(seed=7)
        C = np.concatenate([np.zeros((len(pts), 1), np.int32), pts * ts], axis=1)
    elif kind == "scattered":
        C = cloud(9, n=700, extent=60)
        C[:, 1:] *= ts
    else:
        C = cloud(11, batch=2, n=900, extent=25)
        C[:, 1:] *= ts
        C[:5, 1:] = -C[:5, 1:]                      # negative coordinates: floor-to-cell must stay correct
        C = C[np.unique(me_ref.pack_keys(C), return_index=True)[1]]
    cs = so.CoordSet(to_dev(C, dev), ts)
    out = cs.generated_cells()
    ts_out = ts // 2
    ref = me_ref.generate_coords(C, ts_out, 5)
    got = out.coords.cpu().numpy()
    assert out.n == len(ref) and (np.sort(me_ref.pack_keys(got)) == np.sort(me_ref.pack_keys(ref))).all()
    assert len(np.unique(me_ref.pack_keys(got))) == len(got)
    # 3x3x3 map among the children: cells vs hash
    nbr_c, tm_c = out.kernel_map(out, 3, False)
    plain = so.CoordSet(out.coords.clone(), ts_out)
    nbr_h, tm_h = plain.kernel_map(plain, 3, False)
    assert torch.equal(nbr_c, nbr_h)
    assert torch.equal(tm_c, tm_h)
    # parent lists: child row r at cell c, bit b gets parent pnbr[e][c] through offset b - 2e
    ref_t = me_ref.neighbour_table(C, got, 5, ts_out, True)           # [125][n_out] parent rows
    pn = out.pnbr.cpu().numpy()
    cell = out.out_cell.cpu().numpy()[:out.n]
    cc = out.cells.coords.cpu().numpy()[cell]
    b = (got[:, 1:] - cc[:, 1:]) // ts_out
    assert ((b >= 0) & (b <= 1)).all()
    rebuilt = np.full_like(ref_t, -1)
    for ez in (-1, 0, 1):
        for ey in (-1, 0, 1):
            for ex in (-1, 0, 1):
                e = (ex + 1) + 3 * (ey + 1) + 9 * (ez + 1)
                off = b - 2 * np.array([ex, ey, ez])
                ok = (np.abs(off) <= 2).all(axis=1)
                k = (off[:, 0] + 2) + 5 * (off[:, 1] + 2) + 25 * (off[:, 2] + 2)
                rows = np.nonzero(ok)[0]
                rebuilt[k[rows], rows] = pn[e, cell[rows]]
    assert (rebuilt == ref_t).all()


def test_generated_cells_empty_and_single(pkg, dev):
    so = pkg.sparse_ops
    one = so.CoordSet(to_dev(np.array([[0, 4, 8, -4]]), dev), 4).generated_cells()
    assert one.n == 125 and one.ts == 2
    nbr, _ = one.kernel_map(one, 3, False)
    plain = so.CoordSet(one.coords.clone(), 2)
    assert torch.equal(nbr, plain.kernel_map(plain, 3, False)[0])
    empty = so.CoordSet(torch.zeros((0, 4), dtype=torch.int32, device=dev), 4).generated_cells()
    assert empty.n == 0 and empty.coords.shape == (0, 4)


@pytest.mark.parametrize("n", [1, 2, 1023, 1024, 1025, 4096, 100_003, 9_000_001])
@pytest.mark.parametrize("density", [0.0, 0.1, 0.93, 1.0])
def test_compaction_keeps_input_order_at_every_tile_boundary(pkg, dev, n, density):
    """Single-pass compaction (decoupled look-back over 1024-item tiles): the kept rows are exactly torch.nonzero's, in
    input order, for sizes around the tile boundary and for a level-3-sized mask (thousands of tiles in flight)."""
    so = pkg.sparse_ops
    g = torch.Generator(device="cpu").manual_seed(n % 1000 + int(density * 100))
    mask = (torch.rand(n, generator=g) < density).to(dev)
    coords = torch.stack([torch.zeros(n, dtype=torch.int32), torch.arange(n, dtype=torch.int32) % 30000,
                          torch.arange(n, dtype=torch.int32) // 30000, torch.zeros(n, dtype=torch.int32)], dim=1).to(dev)
    feats = torch.arange(n, dtype=torch.float32, device=dev).reshape(n, 1)
    oc, of = so.prune_rows(coords, feats, mask)
    want = torch.nonzero(mask).reshape(-1)
    assert oc.shape[0] == want.shape[0]
    assert torch.equal(of.reshape(-1).long(), want) and torch.equal(oc, coords[want])


def test_out_of_range_coordinates_raise_instead_of_wrapping(pkg, dev):
    """The hash packs (b, x, y, z) into 15 + 3 x 16 bits; a coordinate outside that range must raise (MinkowskiEngine takes any
    int32 -- here it would wrap onto another voxel's key and merge two voxels silently): ctor / unique_first, strided sets,
    and a table built lazily from caller-given rows (sticky status, read at the next size read-back)."""
    so = pkg.sparse_ops
    ME = importlib.import_module(pkg.__name__ + ".MinkowskiEngine")
    good = torch.tensor([[0, 1, 2, 3], [0, 4, 5, 6], [0, 32767, -32768, 0]], dtype=torch.int32, device=dev)
    t = ME.SparseTensor(features=torch.ones((3, 1), device=dev), coordinates=good, device=dev)
    assert t.C.shape[0] == 3
    for bad_row in ([0, 32768, 0, 0], [0, 0, -32769, 0], [32768, 0, 0, 0], [0, 1, 2, 3 + 65536]):
        bad = torch.cat([good, torch.tensor([bad_row], dtype=torch.int32, device=dev)])
        with pytest.raises(RuntimeError, match="packing range"):
            ME.SparseTensor(features=torch.ones((4, 1), device=dev), coordinates=bad, device=dev)
        with pytest.raises(RuntimeError, match="packing range"):
            so.unique_first(bad)
        cs = so.CoordSet(bad, 1)                     # trusted construction: no check yet
        with pytest.raises(RuntimeError, match="packing range"):
            cs.strided(2)
        cs = so.CoordSet(bad, 1)
        cs.table()                                   # lazily built table: the status is sticky ...
        with pytest.raises(RuntimeError, match="packing range"):
            so.check_status(dev)                     # ... and surfaces at the next read-back
        so.check_status(dev)                         # cleared after it was reported


@pytest.mark.parametrize("c", [32, 128])
def test_gather_rows_from_bf16_rows(pkg, dev, c):
    """pcg_gather_rows_bf16: survivors of a prune gathered from bf16 candidate rows come back as the exact fp32 images."""
    so = pkg.sparse_ops
    g = torch.Generator(device="cpu").manual_seed(c)
    feats = torch.randn((5000, c), generator=g).to(dev).to(torch.bfloat16)
    coords = torch.randint(0, 900, (5000, 4), generator=g, dtype=torch.int32).to(dev)
    rows = torch.randperm(5000, generator=g)[:1234].to(torch.int32).to(dev)
    oc, of = so.gather_rows(coords, feats, rows)
    assert of.dtype == torch.float32 and torch.equal(of, feats.float()[rows.long()]) and torch.equal(oc, coords[rows.long()])
    mask = torch.zeros(5000, dtype=torch.bool, device=dev)
    mask[rows.long()] = True
    oc2, of2 = so.prune_rows_reordered(coords, feats, mask, 1234, reorder=False)          # mask order = ascending source row
    keep = torch.sort(rows.long()).values
    assert torch.equal(of2, feats.float()[keep]) and torch.equal(oc2, coords[keep])


def test_locality_sort_equals_full_morton_sort_below_4096(pkg, dev):
    """Sorting on the low 36 Morton bits (12 per axis, 5 radix passes) gives the full (batch, Z-order) permutation for one batch
    item inside [0, 4096)^3 -- the 10/11-bit frames; outside, it is still a valid permutation (row order is free)."""
    so = pkg.sparse_ops
    g = torch.Generator(device="cpu").manual_seed(1)
    C = torch.cat([torch.zeros((20000, 1), dtype=torch.int32), torch.randint(0, 4096, (20000, 3), generator=g, dtype=torch.int32)], 1).to(dev)
    assert torch.equal(so.sort_rows_i32(C, morton=True), so.sort_rows_i32(C, morton=True, bits=so.LOCALITY_BITS))
    far = C.clone()
    far[:, 1:] += 20000
    perm = so.sort_rows_i32(far, morton=True, bits=so.LOCALITY_BITS)
    assert torch.equal(torch.sort(perm.long()).values, torch.arange(20000, device=dev))
