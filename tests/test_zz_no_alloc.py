"""pcg_mask_to_rows_static: the allocation-free variant of pcg_mask_to_rows (scratch in a file-scope device array of libpcg.so,
for callers that cannot call cudaMalloc: graph capture, serving loops).  No counterpart in the reference.
Called through the C ABI and compared, bit for bit (integer output), with the existing entry point on two
calls in a row (the shared array keeps the previous call's contents).  Collected last on purpose
(file name)."""
import ctypes
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ITEMS = 4 * 1024 * 1024


def _both(L, lib, mask, dev):
    """(rows, count) from the existing entry point (caller scratch) and from the static-scratch one."""
    n = int(mask.shape[0])
    st = L.stream()
    out = []
    ws = torch.empty(lib.pcg_scratch_bytes(max(n, 1)), dtype=torch.uint8, device=dev)
    for static in (False, True):
        rows = torch.full((max(n, 1),), -7, dtype=torch.int32, device=dev)
        cnt = torch.full((1,), -1, dtype=torch.int32, device=dev)
        if static:
            before = torch.cuda.memory_allocated(dev)
            L.call("pcg_mask_to_rows_static", L.ptr(mask), n, L.ptr(rows), L.ptr(cnt), st)
            assert torch.cuda.memory_allocated(dev) == before
        else:
            L.call("pcg_mask_to_rows", L.ptr(mask), n, L.ptr(rows), L.ptr(cnt), L.ptr(ws), ws.numel(), st)
        torch.cuda.synchronize()
        k = int(cnt.item())
        out.append((rows[:k].cpu().numpy(), k))
    return out


def test_static_scratch_entry_point_matches_the_scratch_one_twice_in_a_row(pkg, dev):
    L = importlib.import_module(pkg.__name__ + "._lib")
    lib = L.load()
    rng = np.random.default_rng(0)
    # two calls in a row with different sizes and densities (the second smaller: stale tile states of the first lie beyond it),
    # then the maximum size, an empty mask and n = 0
    for n, p in ((1_500_000, 0.3), (70_001, 0.9), (MAX_ITEMS, 0.05), (5000, 0.0), (0, 0.5)):
        m = torch.from_numpy((rng.random(n) < p).astype(np.uint8)).to(dev) if n else torch.zeros(0, dtype=torch.uint8, device=dev)
        (ref_rows, ref_k), (got_rows, got_k) = _both(L, lib, m, dev)
        want = np.nonzero(m.cpu().numpy())[0].astype(np.int32)
        assert ref_k == got_k == len(want), (n, ref_k, got_k, len(want))
        assert np.array_equal(got_rows, ref_rows) and np.array_equal(got_rows, want), n
    # the same mask twice in a row through the static entry point alone
    m = torch.from_numpy((rng.random(300_000) < 0.5).astype(np.uint8)).to(dev)
    a = _both(L, lib, m, dev)[1]
    b = _both(L, lib, m, dev)[1]
    assert a[1] == b[1] and np.array_equal(a[0], b[0])


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum(pkg, dev):
    L = importlib.import_module(pkg.__name__ + "._lib")
    lib = L.load()
    n = MAX_ITEMS + 1
    m = torch.zeros(n, dtype=torch.uint8, device=dev)
    rows = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = lib.pcg_mask_to_rows_static(L.ptr(m), n, L.ptr(rows), L.ptr(cnt), L.stream())
    assert rc != 0 and b"exceeds the built-in scratch" in lib.pcg_last_error()
    with pytest.raises(RuntimeError, match="exceeds the built-in scratch"):
        L.call("pcg_mask_to_rows_static", L.ptr(m), n, L.ptr(rows), L.ptr(cnt), L.stream())
