"""Functional layer over libpcg.so: coordinate sets, kernel maps, sparse convolution, pruning,
feature lookup.  Host-side mirror of what MinkowskiEngine's CoordinateManager + pybind functions
do for the reference; all arithmetic happens in the CUDA kernels.
"""
import contextlib
import ctypes
import os
import threading
import weakref

import torch

from . import _lib
from ._lib import call, ptr, stream

ACT_NONE, ACT_RELU, ACT_LEAKY = 0, 1, 2

_conv_mode = "auto"      # "auto" | "f32" | "tc"


def set_conv_mode(mode):
    """auto: tcgen05 bf16 where the shape allows, fp32 SIMT otherwise; f32 / tc force one path."""
    global _conv_mode
    assert mode in ("auto", "f32", "tc")
    prev, _conv_mode = _conv_mode, mode
    return prev


def get_conv_mode():
    return _conv_mode


# ----------------------------------------------------------------------------- scratch
_scratch = {}


def scratch(nbytes, device):
    """Per-device growable workspace (stream-ordered reuse on the current stream)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _lib.raw_stream())
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


_ctl = {}

RANGE_ERROR = ("coordinate outside the packing range of the GPU hash (batch in [0, 32767], x/y/z in [-32768, 32767]): "
               "MinkowskiEngine accepts any int32, this graft does not (include/pcg.h, conventions)")


def ctl(device):
    """Per-(device, stream) int32 control words, zero-initialised once: [0] = sticky status of pcg_hash_build (the one table
    builder without a size read-back of its own), [1:] = counters / status words of the call in flight.  Every C-ABI call
    that takes a counter clears it itself, and every reader synchronises right after its call, so the words are reused."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _lib.raw_stream())
    buf = _ctl.get(key)
    if buf is None:
        buf = _ctl[key] = torch.zeros(8, dtype=torch.int32, device=device)
    return buf


def read_ctl(c, n):
    """Words [1, 1 + n) of a control buffer as python ints -- ONE blocking read-back, which also carries the sticky
    coordinate-range status of every table built on this stream since the last read."""
    vals = c[:1 + n].tolist()
    if vals[0]:
        c[0:1].zero_()
        raise RuntimeError(RANGE_ERROR)
    return vals[1:]


def check_status(device):
    """Raise if any table built on this stream since the last read-back met a coordinate outside the packing range."""
    read_ctl(ctl(device), 0)


# Packed-operand caches (weights, tables) are filled lazily and shared by every host thread / CUDA stream that runs the model
# (codec.FramePipeline).  A miss is rare (first use, or a parameter changed): it is taken under this lock and the packing
# stream is synchronised BEFORE the new buffer is published, so no other stream can consume a buffer that is still being
# written; a hit takes no lock (one tuple read).  `UnifiedModel.prepare()` fills every cache up front.
pack_lock = threading.RLock()


def _need_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must live on a CUDA device: the graft has no CPU path")


# ----------------------------------------------------------------------------- coordinate sets
class CoordSet:
    """Unique (N,4) int32 coordinates + their GPU hash table + cached derived sets / kernel maps.

    Stands in for one ME coordinate-map key inside a CoordinateManager.
    """

    def __init__(self, coords, tensor_stride, table=None):
        _need_cuda(coords, "coordinates")
        assert coords.dtype == torch.int32 and coords.dim() == 2 and coords.shape[1] == 4
        self.coords = coords.contiguous()
        self.n = int(coords.shape[0])
        self.ts = int(tensor_stride)
        self._table = table          # (keys, vals, cap)
        self._strided = {}
        self._generated = {}
        self._kmaps = {}
        self.coords._pcg_coord_set = (weakref.ref(self), self.coords._version)

    # -- hash table --------------------------------------------------------------------------
    def table(self):
        if self._table is None:
            dev = self.coords.device
            cap = _lib.load().pcg_hash_capacity(max(self.n, 1))
            keys = torch.empty(cap, dtype=torch.int64, device=dev)
            vals = torch.empty(cap, dtype=torch.int32, device=dev)
            with hbm_span("hash_build", (self.n,), 16.0 * self.n + 12.0 * cap):      # coordinates read, table (key + row) written
                call("pcg_hash_build", ptr(self.coords), self.n, ptr(keys), ptr(vals), cap, ptr(ctl(dev)[0:1]), stream())
            self._table = (keys, vals, cap)
        return self._table

    def lookup(self, query):
        """Row of every query coordinate (int32 (M,4)), -1 when absent."""
        keys, vals, cap = self.table()
        query = query.to(torch.int32).contiguous()
        rows = torch.empty(query.shape[0], dtype=torch.int32, device=query.device)
        call("pcg_hash_lookup", ptr(keys), ptr(vals), cap, ptr(query), query.shape[0], ptr(rows), stream())
        return rows

    # -- derived sets ------------------------------------------------------------------------
    def strided(self, stride):
        """Coordinates of a stride-`stride` convolution output (first-occurrence order)."""
        stride = int(stride)
        if stride == 1:
            return self
        hit = self._strided.get(stride)
        if hit is None:
            dev = self.coords.device
            step = stride * self.ts
            cap = _lib.load().pcg_hash_capacity(max(self.n, 1))
            keys = torch.empty(cap, dtype=torch.int64, device=dev)
            vals = torch.empty(cap, dtype=torch.int32, device=dev)
            out = torch.empty((self.n, 4), dtype=torch.int32, device=dev)
            c = ctl(dev)
            nb = _lib.load().pcg_scratch_bytes(max(self.n, 1))
            ws = scratch(nb, dev)
            with hbm_span("stride_coords", (self.n,), 16.0 * self.n + 12.0 * cap):      # lower bound: the 16 * n_out written rows not counted
                call("pcg_stride_coords", ptr(self.coords), self.n, step, ptr(keys), ptr(vals), cap, ptr(out), ptr(c[1:2]), ptr(c[2:3]),
                     ptr(ws), ws.numel(), stream())
            n_out, status = read_ctl(c, 2)
            if status & 2:
                raise RuntimeError(RANGE_ERROR)
            hit = CoordSet(out[:n_out], step, table=(keys, vals, cap))
            self._strided[stride] = hit
        return hit

    def generated(self, ksize, stride, guess_per_row=40):
        """Coordinates of a generative transposed convolution (kernel ksize, up-sampling `stride`)."""
        key = (int(ksize), int(stride))
        hit = self._generated.get(key)
        if hit is None:
            assert self.ts % stride == 0 and self.ts // stride >= 1, "fractional output tensor stride"
            ts_out = self.ts // stride
            dev = self.coords.device
            K = ksize ** 3
            lib = _lib.load()
            n_cand = self.n * K
            # exact for disjoint children (even kernel == stride); a surface estimate otherwise, retried on overflow
            max_out = n_cand if (ksize == stride or self.n < 4096) else min(n_cand, guess_per_row * self.n)
            while True:
                cap = lib.pcg_hash_capacity(max(max_out, 1))
                keys = torch.empty(cap, dtype=torch.int64, device=dev)
                vals = torch.empty(cap, dtype=torch.int32, device=dev)
                out = torch.empty((max(max_out, 1), 4), dtype=torch.int32, device=dev)
                c = ctl(dev)
                ws = scratch(lib.pcg_scratch_bytes(max(n_cand, 1)), dev)
                with hbm_span("generate_coords", (self.n, K), 16.0 * self.n + 12.0 * cap):   # lower bound, as above
                    call("pcg_generate_coords", ptr(self.coords), self.n, int(ksize), int(stride), ts_out, ptr(keys), ptr(vals), cap,
                         ptr(out), max_out, ptr(c[1:2]), ptr(c[2:3]), ptr(ws), ws.numel(), stream())
                n_out, status = read_ctl(c, 2)
                if status & 2:
                    raise RuntimeError("generated " + RANGE_ERROR)
                if status == 0:
                    break
                if max_out >= n_cand:
                    raise RuntimeError("pcg_generate_coords: table overflow at full capacity")
                max_out = n_cand
            hit = CoordSet(out[:n_out], ts_out, table=(keys, vals, cap))
            self._generated[key] = hit
        return hit

    def generated_cells(self):
        """Children of the kernel-5 / stride-2 generative up-sampling, organised by octree cells (pcg_cells.cu).
        Same coordinate SET as generated(5, 2); rows are cell-major."""
        hit = self._generated.get("cells")
        if hit is None:
            assert self.ts % 2 == 0, "fractional output tensor stride"
            dev = self.coords.device
            lib = _lib.load()
            cells = self.generated(3, 1, guess_per_row=14)                # parents dilated by {-1,0,1}^3 * ts
            # pnbr[e][cell] = parent row at cell + e*ts, by scatter from the (parent, offset) pairs that generated the cells
            ckeys, cvals, ccap = cells.table()
            pitch = lib.pcg_nbr_pitch(cells.n)
            pnbr = torch.empty((27, pitch), dtype=torch.int32, device=dev)
            c = ctl(dev)
            call("pcg_cells_parent_map", ptr(self.coords), self.n, self.ts, ptr(ckeys), ptr(cvals), ccap, cells.n, ptr(pnbr), pitch,
                 ptr(c[2:3]), stream())
            mask = torch.empty(max(cells.n, 1), dtype=torch.uint8, device=dev)
            base = torch.empty(max(cells.n, 1), dtype=torch.int32, device=dev)
            ws = scratch(lib.pcg_cells_scratch_bytes(max(cells.n, 1)), dev)
            call("pcg_cells_info", ptr(pnbr), pnbr.shape[1], cells.n, ptr(mask), ptr(base), ptr(c[1:2]), ptr(ws), ws.numel(), stream())
            n_out, status = read_ctl(c, 2)
            if status:
                raise RuntimeError("pcg_cells_parent_map: a dilated parent position is missing from the cell table")
            out = torch.empty((n_out, 4), dtype=torch.int32, device=dev)
            out_cell = torch.empty(max(n_out, 1), dtype=torch.int32, device=dev)
            call("pcg_cells_emit", ptr(cells.coords), ptr(mask), ptr(base), cells.n, self.ts // 2, ptr(out), ptr(out_cell), stream())
            hit = CellCoordSet(out, self.ts // 2, self, cells, pnbr, mask, base, out_cell)
            self._generated["cells"] = hit
        return hit

    def identity_map(self):
        """nbr / tile_mask of the identity (row i <- row i): operand map of a dense product on the sparse-conv kernel."""
        hit = self._kmaps.get("identity")
        if hit is None:
            dev = self.coords.device
            pitch = _lib.load().pcg_nbr_pitch(self.n)
            nbr = torch.arange(pitch, dtype=torch.int32, device=dev)
            nbr[self.n:] = -1
            tmask = torch.zeros((max((self.n + 127) // 128, 1), 4), dtype=torch.int32, device=dev)
            tmask[:, 0] = 1
            hit = (nbr.view(1, pitch), tmask)
            self._kmaps["identity"] = hit
        return hit

    # -- kernel maps -------------------------------------------------------------------------
    def kernel_map(self, out_set, ksize, transposed):
        """nbr[K][pitch] (rows of THIS set feeding each output row; pitch = n_out rounded up to 128, padding -1)
        + per-128-row-tile offset mask."""
        key = (id(out_set), int(ksize), bool(transposed))
        hit = self._kmaps.get(key)
        if hit is None:
            keys, vals, cap = self.table()
            dev = self.coords.device
            K = ksize ** 3
            ts = out_set.ts if transposed else self.ts
            pitch = _lib.load().pcg_nbr_pitch(out_set.n)
            nbr = torch.empty((K, pitch), dtype=torch.int32, device=dev)
            tiles = (out_set.n + 127) // 128
            tmask = torch.empty((max(tiles, 1), 4), dtype=torch.int32, device=dev) if K <= 128 else None
            # output coordinates read once, the table read once (lower bound of K probes per row), neighbour table written
            with hbm_span("kernel_map", (K, out_set.n), 16.0 * out_set.n + 12.0 * cap + 4.0 * K * pitch):
                call("pcg_kernel_map", ptr(keys), ptr(vals), cap, ptr(out_set.coords), out_set.n, int(ksize), ts,
                     int(bool(transposed)), self.ts if transposed else 0, ptr(nbr), pitch, ptr(tmask), stream())
            hit = (nbr, tmask, None if out_set is self else out_set)   # keep out_set alive while cached (id() key)
            self._kmaps[key] = hit
        return hit[0], hit[1]


class CellCoordSet(CoordSet):
    """Candidate set of a kernel-5 / stride-2 generative up-sampling: children grouped by cell, with the cell table,
    the per-cell parent lists and child masks that replace a hash table over the children (pcg.h, octree-cell section)."""

    def __init__(self, coords, tensor_stride, parent, cells, pnbr, cell_mask, cell_base, out_cell):
        super().__init__(coords, tensor_stride)
        self._parent = weakref.ref(parent)      # weak: the parent caches this set (no reference cycle -> memory is freed by refcount)
        self.cells, self.pnbr = cells, pnbr
        self.cell_mask, self.cell_base, self.out_cell = cell_mask, cell_base, out_cell

    parent = property(lambda self: self._parent())

    def kernel_map(self, out_set, ksize, transposed):
        if out_set is not self or ksize != 3 or transposed:
            return super().kernel_map(out_set, ksize, transposed)
        key = (id(out_set), 3, False)
        hit = self._kmaps.get(key)
        if hit is None:
            dev = self.coords.device
            ckeys, cvals, ccap = self.cells.table()
            pitch = _lib.load().pcg_nbr_pitch(self.n)
            nbr = torch.empty((27, pitch), dtype=torch.int32, device=dev)
            tmask = torch.empty((max((self.n + 127) // 128, 1), 4), dtype=torch.int32, device=dev)
            with hbm_span("cells_kernel_map3", (27, self.n), 20.0 * self.n + 12.0 * ccap + 5.0 * self.cells.n + 4.0 * 27 * pitch):
                call("pcg_cells_kernel_map3", ptr(ckeys), ptr(cvals), ccap, ptr(self.coords), ptr(self.out_cell), self.n, self.ts,
                     ptr(self.cell_mask), ptr(self.cell_base), ptr(nbr), pitch, ptr(tmask), stream())
            hit = (nbr, tmask, None)
            self._kmaps[key] = hit
        return hit[0], hit[1]


def coord_set_of(coords, tensor_stride):
    """CoordSet already attached to this exact coordinate tensor, if still valid."""
    tag = getattr(coords, "_pcg_coord_set", None)
    if tag is None or tag[1] != coords._version:
        return None
    cs = tag[0]()
    if cs is not None and cs.ts == int(tensor_stride) and cs.coords is coords:
        return cs
    return None


def unique_first(coords):
    """(unique coords in first-occurrence order, source row index, table) -- ME ctor / sparse_quantize dedupe."""
    _need_cuda(coords, "coordinates")
    coords = coords.to(torch.int32).contiguous()
    n = coords.shape[0]
    dev = coords.device
    lib = _lib.load()
    cap = lib.pcg_hash_capacity(max(n, 1))
    keys = torch.empty(cap, dtype=torch.int64, device=dev)
    vals = torch.empty(cap, dtype=torch.int32, device=dev)
    out = torch.empty((max(n, 1), 4), dtype=torch.int32, device=dev)
    index = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    c = ctl(dev)
    ws = scratch(lib.pcg_scratch_bytes(max(n, 1)), dev)
    call("pcg_unique_first", ptr(coords), n, ptr(keys), ptr(vals), cap, ptr(out), ptr(index), ptr(c[1:2]), ptr(c[2:3]), ptr(ws),
         ws.numel(), stream())
    m, status = read_ctl(c, 2)
    if status & 2:
        raise RuntimeError(RANGE_ERROR)
    return out[:m], index[:m], (keys, vals, cap)


LOCALITY_BITS = 36      # Z-order on the low 12 bits per axis: enough for memory locality on 11-bit frames, 5 radix passes instead of 8


def sort_rows_i32(coords, morton=False, bits=None):
    """Row permutation (int32): lexicographic (b,x,y,z) (utils.py:142-187) or, with morton=True, (b, Z-order).
    bits: sort on the low `bits` key bits only (morton + LOCALITY_BITS: a coherence order for row layouts whose order is free)."""
    _need_cuda(coords, "coordinates")
    coords = coords.to(torch.int32).contiguous()
    n = coords.shape[0]
    perm = torch.empty(n, dtype=torch.int32, device=coords.device)
    if n:
        lib = _lib.load()
        ws = scratch(lib.pcg_sort_rows_scratch_bytes(n), coords.device)
        call("pcg_sort_rows_bits", ptr(coords), n, int(bool(morton)), 0, 64 if bits is None else int(bits), ptr(perm), ptr(ws), ws.numel(), stream())
    return perm


def sort_rows(coords, morton=False):
    return sort_rows_i32(coords, morton).long()


def gather_block(cloud, perm=None, batch=0):
    """(n,6) float rows [xyz|rgb] -> (coords (n,4) int32 = (batch, floor xyz), feats (n,4) = (1, rgb)), rows taken in `perm` order."""
    _need_cuda(cloud, "point cloud")
    c = cloud.to(torch.float32).contiguous()
    n = c.shape[0]
    coords = torch.empty((n, 4), dtype=torch.int32, device=c.device)
    feats = torch.empty((n, 4), dtype=torch.float32, device=c.device)
    call("pcg_gather_block", ptr(c), ptr(perm), n, int(batch), ptr(coords), ptr(feats), stream())
    return coords, feats


def lexsort(coords):
    return sort_rows(coords, morton=False)


# ----------------------------------------------------------------------------- features
def cin_padded(c):
    """Operand row width of the tensor-core path for c input channels (16, 32 or a multiple of 64)."""
    return 16 if c <= 16 else (32 if c <= 32 else (c + 63) // 64 * 64)


def to_bf16(feats):
    """bf16 operand rows of an fp32 feature matrix, zero-padded to pcg_tc_cin_padded(C) channels (cached on the tensor)."""
    if feats.dtype == torch.bfloat16:
        assert feats.shape[1] == cin_padded(feats.shape[1]) and feats.is_contiguous()
        return feats
    tag = getattr(feats, "_pcg_bf16", None)
    if tag is not None and tag[1] == feats._version:
        return tag[0]
    f = feats.contiguous()
    n, c = f.shape
    cp = cin_padded(c)
    out = torch.empty((n, cp), dtype=torch.bfloat16, device=f.device)
    call("pcg_f32_to_bf16_rows", ptr(f), n, c, cp, ptr(out), stream())
    feats._pcg_bf16 = (out, feats._version)
    return out


class PackedWeights:
    """bf16 operand image of a (K,cin,cout) kernel, rebuilt when the parameter changes.
    kind "umma": K-major swizzled tiles for tcgen05; kind "mma": mma.sync B fragments for the narrow-head kernel."""

    def __init__(self):
        self.entry = None                  # (key, buffer), published as ONE object
        self._proj_host = None

    def get(self, w3, kind="umma"):
        key = (w3.data_ptr(), w3._version, tuple(w3.shape), kind)
        hit = self.entry
        if hit is not None and hit[0] == key:
            return hit[1]
        with pack_lock:
            hit = self.entry
            if hit is not None and hit[0] == key:
                return hit[1]
            K, cin, cout = w3.shape
            lib = _lib.load()
            if kind == "im2col":        # (K,cin,cout) seen as one dense (1, roundup64(K*cin), cout) kernel
                rows = im2col_row_elems(K, cin)
                w2 = torch.zeros((1, rows, cout), dtype=torch.float32, device=w3.device)
                w2[0, :K * cin] = w3.reshape(K * cin, cout)
                w3, K, cin, kind = w2, 1, rows, "umma"
            nbytes = lib.pcg_packed_weight_bytes(K, cin, cout) if kind == "umma" else lib.pcg_mma_weight_bytes(K, cin, cout)
            buf = torch.empty(nbytes, dtype=torch.uint8, device=w3.device)
            call("pcg_pack_weights_bf16" if kind == "umma" else "pcg_pack_weights_mma", ptr(w3.contiguous()), K, cin, cout,
                 ptr(buf), stream())
            torch.cuda.current_stream().synchronize()
            self.entry = (key, buf)
        return buf


class ConvProfiler:
    """Optional per-launch CUDA-event timing of the convolution kernels (bench.py's roofline pass).
    Events are recorded on torch's current stream, which is the stream the kernels are launched on."""

    def __init__(self, detail=False):
        self.records = []
        self._pairs = {}
        self.detail = detail          # also record per-launch row counts and mean active offsets per tile
        self.launches = []
        self.impl_bytes = {}          # bytes of intermediates the implementation moves through HBM (not algorithmic), by name

    def pairs(self, nbr):
        key = id(nbr)                       # the tensor is kept alive below, so the id cannot be recycled
        if key not in self._pairs:
            self._pairs[key] = (int((nbr >= 0).sum().item()), nbr)
        return self._pairs[key][0]

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for kind, shape, flops, bytes_, e0, e1 in self.records:
            d = out.setdefault(kind, {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0, "by_shape": {}})
            ms = e0.elapsed_time(e1)
            d["launches"] += 1; d["ms"] += ms; d["flops"] += flops; d["bytes"] += bytes_
            b = d["by_shape"].setdefault(shape, [0, 0.0, 0.0])
            b[0] += 1; b[1] += ms; b[2] += flops
        return out


profiler = None      # set to a ConvProfiler() to time every convolution launch


@contextlib.contextmanager
def hbm_span(kind, shape, nbytes):
    """CUDA events around one HBM-bound C-ABI call (hash, kernel map, compaction, entropy model) when a profiler is set.
    nbytes = ALGORITHMIC bytes of the call: inputs read once + outputs written once."""
    prof = profiler
    if prof is None:
        yield
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    yield
    e1.record()
    prof.records.append((kind, shape, 0.0, float(nbytes), e0, e1))


def im2col_row_elems(K, cin):
    return (K * cin + 63) // 64 * 64


def im2col_eligible(K, cin, cout):
    """Tiny C_in with many offsets (the 4-channel input layer): dense col @ W instead of K row gathers per output."""
    return cin in (1, 2, 4, 8) and K >= 27 and im2col_row_elems(K, cin) <= 1024 and 16 <= cout <= 256


def tc_eligible(K, cin, cout):
    """Static shape rule (never by timing: determinism).  Everything with a reduction depth worth a
    tensor-core tile goes to tcgen05; the 1x1 colour head stays fp32 (its output is rounded to 8 bits)."""
    return K <= 128 and cout <= 256 and K * cin >= 32 and not (K == 1 and cout < 16)


def narrow_rule(K, cin, cout):
    """Static rule: layers that take the register-gather mma.sync kernel (k_conv_mma) in "auto" mode."""
    return _conv_mode == "auto" and cin <= 64 and cout <= 16 and K <= 128 and \
        (cin <= 16 or os.environ.get("PCG_NARROW") == "mma") and os.environ.get("PCG_NARROW") != "tc" and \
        _lib.load().pcg_mma_weight_bytes(K, cin, cout) <= 160 * 1024


def pack_kind(K, cin, cout, fp32_input=True):
    """Operand image the static rules of sparse_conv pick for a (K,cin,cout) kernel in the current mode; None = fp32 SIMT path."""
    if _conv_mode == "f32" or not tc_eligible(K, cin, cout):
        return None
    if _conv_mode == "auto" and fp32_input and im2col_eligible(K, cin, cout):
        return "im2col"
    return "mma" if narrow_rule(K, cin, cout) else "umma"


def sparse_conv(feats, in_set, out_set, weight, bias, ksize, transposed, act=ACT_NONE, packed=None, out="f32"):
    """out[o] = act(bias + sum_i F[nbr[i][o]] @ W[i]).

    feats: fp32 (N,cin), or bf16 (N,roundup8(cin)) as produced by a previous call with out="bf16".
    out: "f32" -> fp32 tensor; "both" -> fp32 tensor carrying a bf16 shadow for the next tensor-core layer;
    "bf16" -> bf16 tensor only (hidden activations that only feed another convolution).  On the fp32 SIMT path
    the result is always fp32.  weight: (K,cin,cout) or (cin,cout) for K == 1.
    """
    _need_cuda(feats, "features")
    w3 = weight if weight.dim() == 3 else weight.unsqueeze(0)
    K, cin, cout = w3.shape
    assert K == ksize ** 3, f"kernel {tuple(w3.shape)} vs kernel_size {ksize}"
    if feats.dtype == torch.bfloat16:
        assert feats.shape[1] == cin_padded(cin), f"kernel {tuple(w3.shape)} vs bf16 features {tuple(feats.shape)}"
    else:
        assert feats.shape[1] == cin, f"kernel {tuple(w3.shape)} vs features {tuple(feats.shape)}"
    nbr, tmask = in_set.kernel_map(out_set, ksize, transposed)
    n_out = out_set.n
    b = bias.reshape(-1).contiguous() if bias is not None else None
    mode = _conv_mode
    use_tc = mode != "f32" and tc_eligible(K, cin, cout) and tmask is not None and feats.shape[0] > 0
    if mode == "tc" and not use_tc:
        raise RuntimeError(f"conv shape K={K} cin={cin} cout={cout} is not tensor-core eligible")
    if not use_tc and feats.dtype == torch.bfloat16:
        feats = feats[:, :cin].float()          # a bf16 hidden activation feeding a layer too small for the tensor-core paths
    prof = profiler
    if prof is not None:
        pairs = prof.pairs(nbr)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if prof.detail:
            act_mean = float(sum(((tmask.view(-1).to(torch.int64) & 0xFFFFFFFF) >> i & 1).sum() for i in range(32)).item()) / max(tmask.shape[0], 1) if tmask is not None else float(K)
            prof.launches.append({"shape": (K, cin, cout), "n_in": int(feats.shape[0]), "n_out": int(n_out), "pairs": pairs,
                                  "active_per_tile": act_mean, "transposed": bool(transposed), "e": (e0, e1)})
    # "tc" forces the tcgen05 kernel (tests pin its layouts on these shapes too); "auto" sends the narrow heads to the
    # register-gather mma.sync kernel
    # (the LSU-gather tcgen05 kernel handles 32 / 64-byte rows too and is faster on the 32->16 and 64->1 heads -- level-3 candidates:
    # 1.20 against 1.47 ms, level 2: 0.46 against 0.60 -- while the 16->1 layer stays on k_conv_mma, 0.68 against 0.76 ms;
    # static shape rule, PCG_NARROW=mma sends every head layer to k_conv_mma for A/B runs)
    narrow = use_tc and narrow_rule(K, cin, cout)
    is_head = cin <= 64 and cout <= 16          # gather-bound occupancy-head layers: reported with the HBM/L2-bound kernels
    if use_tc and _conv_mode == "auto" and feats.dtype == torch.float32 and im2col_eligible(K, cin, cout):
        rows = im2col_row_elems(K, cin)
        col = torch.empty((n_out, rows), dtype=torch.bfloat16, device=feats.device)
        want_bf16 = out in ("bf16", "both") and cout == cin_padded(cout)
        want_f32 = out != "bf16" or not want_bf16
        of = torch.empty((n_out, cout), dtype=torch.float32, device=feats.device) if want_f32 else None
        ob = torch.empty((n_out, cout), dtype=torch.bfloat16, device=feats.device) if want_bf16 else None
        wp = (packed or PackedWeights()).get(w3.detach(), kind="im2col")
        inbr, itm = out_set.identity_map()
        if prof is not None:
            e0.record()
        call("pcg_im2col_bf16", ptr(feats.contiguous()), cin, ptr(nbr), nbr.shape[1], K, n_out, rows, ptr(col), stream())
        call("pcg_conv_fwd_tc", ptr(col), n_out, rows, ptr(inbr), inbr.shape[1], ptr(itm), 1, n_out, ptr(wp), ptr(b), cout,
             int(act), ptr(of), ptr(ob), stream())
        if prof is not None:
            e1.record()
            prof.records.append(("conv_tc", (K, cin, cout), 2.0 * pairs * cin * cout,
                                 4.0 * feats.shape[0] * cin + (4.0 if want_f32 else 2.0) * n_out * cout, e0, e1))
        if of is None:
            return ob
        if ob is not None:
            of._pcg_bf16 = (ob, of._version)
        return of
    if narrow:
        fb = to_bf16(feats)
        want_bf16 = out in ("bf16", "both")
        want_f32 = out != "bf16" or cout != 16          # the bf16 rows are 16 wide: alone they stand for the output only when cout == 16
        of = torch.empty((n_out, cout), dtype=torch.float32, device=feats.device) if want_f32 else None
        ob = torch.empty((n_out, 16), dtype=torch.bfloat16, device=feats.device) if want_bf16 else None
        if prof is not None:
            e0.record()
        wf = (packed or PackedWeights()).get(w3.detach(), kind="mma")
        call("pcg_conv_fwd_mma", ptr(fb), feats.shape[0], cin, ptr(nbr), nbr.shape[1], K, n_out, ptr(wf), ptr(b), cout,
             int(act), ptr(of), ptr(ob), stream())
        if prof is not None:
            e1.record()
            prof.records.append(("conv_narrow", (K, cin, cout), 2.0 * pairs * cin * cout,
                                 2.0 * feats.shape[0] * cin_padded(cin) + (4.0 if want_f32 else 2.0) * n_out * cout, e0, e1))
        if of is None:
            return ob
        if ob is not None:
            of._pcg_bf16 = (ob, of._version)
        return of
    if use_tc:
        fb = to_bf16(feats)
        wp = (packed or PackedWeights()).get(w3.detach())
        want_bf16 = out in ("bf16", "both") and cout == cin_padded(cout)
        want_f32 = out != "bf16" or not want_bf16
        of = torch.empty((n_out, cout), dtype=torch.float32, device=feats.device) if want_f32 else None
        ob = torch.empty((n_out, cout), dtype=torch.bfloat16, device=feats.device) if want_bf16 else None
        if prof is not None:
            e0.record()
        call("pcg_conv_fwd_tc", ptr(fb), feats.shape[0], cin, ptr(nbr), nbr.shape[1], ptr(tmask), K, n_out, ptr(wp), ptr(b), cout,
             int(act), ptr(of), ptr(ob), stream())
        if prof is not None:
            e1.record()
            if is_head:
                prof.records.append(("conv_narrow", (K, cin, cout), 2.0 * pairs * cin * cout,
                                     2.0 * feats.shape[0] * cin_padded(cin) + (4.0 if want_f32 else 2.0) * n_out * cout, e0, e1))
            else:
                prof.records.append(("conv_tc", (K, cin, cout), 2.0 * pairs * cin * cout,
                                     2.0 * feats.shape[0] * cin + (4.0 if want_f32 else 2.0) * n_out * cout, e0, e1))
        if of is None:
            return ob
        if ob is not None:
            of._pcg_bf16 = (ob, of._version)
        return of
    f = feats.contiguous()
    of = torch.empty((n_out, cout), dtype=torch.float32, device=feats.device)
    if prof is not None:
        e0.record()
    call("pcg_conv_fwd_f32", ptr(f), f.shape[0], cin, ptr(nbr), nbr.shape[1], K, n_out, ptr(w3.detach().contiguous()), ptr(b),
         cout, int(act), ptr(of), stream())
    if prof is not None:
        e1.record()
        prof.records.append(("conv_f32", (K, cin, cout), 2.0 * pairs * cin * cout,
                             4.0 * f.shape[0] * cin + 4.0 * n_out * cout, e0, e1))
    return of


def head_fusable(ksize_a, w_a, ksize_b, w_b):
    """Static rule for occupancy_head: conv k3 (C -> 16..64, tensor-core eligible, 64- or 128-byte operand rows), ReLU, conv k3 -> 1."""
    if _conv_mode == "f32" or os.environ.get("PCG_HEAD_FUSED") == "0" or ksize_a != 3 or ksize_b != 3 or w_a.dim() != 3 or w_b.dim() != 3:
        return False
    K, cin, cmid = w_a.shape
    # (64-byte operand rows, 17 <= cin <= 32, are covered by the parity tests with cmid = 16 only: wider hidden rows on them stay unfused)
    return (K == 27 and tuple(w_b.shape) == (27, cmid, 1) and cmid in (16, 32, 64) and (cin >= 33 or (cin >= 17 and cmid == 16))
            and tc_eligible(K, cin, cmid) and not im2col_eligible(K, cin, cmid))


def head_host_params(packed, w_b, b_a):
    """Host copies of the fused head's second-layer kernel (27, cmid) and first-layer bias: they travel in the launch's kernel
    parameters (constant bank).  Made once per parameter version and cached on the first layer's PackedWeights."""
    key = (w_b.data_ptr(), w_b._version, None if b_a is None else (b_a.data_ptr(), b_a._version))
    hit = getattr(packed, "_proj_host", None) if packed is not None else None
    if hit is None or hit[0] != key:
        with pack_lock:                    # .cpu() synchronises; the tuple is published as one object
            hit = (key, w_b.detach().reshape(27, -1).float().cpu().contiguous(),
                   None if b_a is None else b_a.detach().reshape(-1).float().cpu().contiguous())
            if packed is not None:
                packed._proj_host = hit
    return hit[1], hit[2]


def occupancy_head(feats, cset, w_a, b_a, w_b, b_b, packed=None):
    """logit = conv_b(relu(conv_a(x))) on one coordinate set (model/transforms.py:146-160), second layer GEMM-first:
    the first layer's epilogue (tcgen05 kernel, thread = output row, fp32 accumulator row in registers) also forms the 27 dot
    products s[k][p] = relu(h[p]) . w_b[k], stored offset-major; the second layer is then logit[o] = b_b + sum_k s[k][nbr[k][o]]
    -- one 4-byte read per pair from a plane in which neighbouring outputs hit neighbouring rows, instead of a 32- to 128-byte
    row gather per pair -- and the hidden rows are neither rounded to bf16 nor written.  Same fixed accumulation order (k ascending)."""
    _need_cuda(feats, "features")
    K, cin, cmid = w_a.shape
    nbr, tmask = cset.kernel_map(cset, 3, False)
    n = cset.n
    dev = feats.device
    out = torch.empty((n, 1), dtype=torch.float32, device=dev)
    if n == 0:
        return out
    pitch = nbr.shape[1]
    fb = to_bf16(feats)
    wp = (packed or PackedWeights()).get(w_a.detach())
    bb = b_b.reshape(-1).contiguous() if b_b is not None else None
    w2, ba_host = head_host_params(packed, w_b, b_a)
    planes = torch.empty((27, pitch), dtype=torch.float32, device=dev)
    prof = profiler
    if prof is not None:
        pairs = prof.pairs(nbr)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
    call("pcg_conv_fwd_tc_proj", ptr(fb), fb.shape[0], cin, ptr(nbr), pitch, ptr(tmask), K, n, ptr(wp),
         None if ba_host is None else ctypes.c_void_p(ba_host.data_ptr()), cmid,
         ACT_RELU, None, None, ctypes.c_void_p(w2.data_ptr()), 27, ptr(planes), pitch, stream())       # w2 / bias: the HOST pointers of the ABI
    if prof is not None:
        e1.record()
    call("pcg_gather_sum", ptr(planes), pitch, ptr(nbr), pitch, 27, n, ptr(bb), ACT_NONE, ptr(out), stream())
    if prof is not None:
        e2.record()
        # algorithmic bytes: inputs read once + outputs written once.  The projection planes and the neighbour
        # table are intermediates / implementation artefacts: they appear in the measured DRAM traffic, not here.  The pair
        # (first layer + gather-sum) is the unit with a defined input and output: rows in, one logit per row out.
        kind = "conv_narrow" if cin <= 64 and cmid <= 16 else "conv_tc"
        prof.records.append((kind, (K, cin, cmid), 2.0 * pairs * cin * cmid, 0.0, e0, e1))
        prof.records.append(("head_gather", (27, cmid, 1), 2.0 * pairs * cmid, 0.0, e1, e2))
        prof.records.append(("occupancy_head", (K, cin, cmid, 1), 2.0 * pairs * cmid * (cin + 1), 2.0 * fb.shape[0] * cin_padded(cin) + 4.0 * n, e0, e2))
    return out


class PackedGemmWeights:
    """bf16 operand image of a (K,cin,cout) kernel seen as the dense matrix [cin x K*cout] in column groups of 256."""

    def __init__(self):
        self.entry = None

    def get(self, w3):
        key = (w3.data_ptr(), w3._version, tuple(w3.shape))
        hit = self.entry
        if hit is not None and hit[0] == key:
            return hit[1]
        with pack_lock:
            hit = self.entry
            if hit is not None and hit[0] == key:
                return hit[1]
            K, cin, cout = w3.shape
            n_cols = K * cout
            G = (n_cols + 255) // 256
            wm = torch.zeros((cin, G * 256), dtype=torch.float32, device=w3.device)
            wm[:, :n_cols] = w3.permute(1, 0, 2).reshape(cin, n_cols)
            wg = wm.view(cin, G, 256).permute(1, 0, 2).contiguous()
            nbytes = _lib.load().pcg_packed_weight_bytes(G, cin, 256)
            buf = torch.empty(nbytes, dtype=torch.uint8, device=w3.device)
            call("pcg_pack_weights_bf16", ptr(wg), G, cin, 256, ptr(buf), stream())
            torch.cuda.current_stream().synchronize()
            self.entry = (key, buf)
        return buf


def cells_eligible(ksize, stride, cin, cout):
    """Static rule for the octree-cell / GEMM-first generative transposed convolution."""
    return ksize == 5 and stride == 2 and cout in (32, 64, 128, 256) and 33 <= cin <= 128 and _conv_mode != "f32"


def gen_conv_transpose_cells(feats, in_set, out_set, weight, bias, act=ACT_NONE, packed=None, out="f32"):
    """Kernel-5 / stride-2 generative transposed convolution, GEMM-first: T[p][k] = F[p] @ W[k] as one dense
    tcgen05 product over all parents (every (parent, offset) pair hits, so there is nothing to gather), then each
    child sums the <= 27 rows of T its parents hold for it (pcg_cells_reduce, fixed order)."""
    _need_cuda(feats, "features")
    K, cin, cout = weight.shape
    assert K == 125 and isinstance(out_set, CellCoordSet) and out_set.parent is in_set
    dev = feats.device
    n_par, n_out = in_set.n, out_set.n
    want_bf16 = out in ("bf16", "both")
    want_f32 = out != "bf16"
    of = torch.empty((n_out, cout), dtype=torch.float32, device=dev) if want_f32 else None
    ob = torch.empty((n_out, cout), dtype=torch.bfloat16, device=dev) if want_bf16 else None
    if n_out == 0 or n_par == 0:
        return of if of is not None else ob
    fb = to_bf16(feats)
    wp = (packed or PackedGemmWeights()).get(weight.detach())
    n_cols = K * cout
    T = torch.empty((n_par, n_cols), dtype=torch.bfloat16, device=dev)
    prof = profiler
    if prof is not None:
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
    call("pcg_gemm_bf16_grouped", ptr(fb), n_par, cin, ptr(wp), n_cols, ptr(T), stream())
    if prof is not None:
        e1.record()
    b = bias.reshape(-1).contiguous() if bias is not None else None
    call("pcg_cells_reduce", ptr(T), cout, ptr(out_set.pnbr), out_set.pnbr.shape[1], ptr(out_set.cell_mask), ptr(out_set.cell_base),
         out_set.cells.n, ptr(b), int(act), ptr(of), ptr(ob), stream())
    if prof is not None:
        e2.record()
        flops = 2.0 * n_par * K * cin * cout
        # T (n_par x 125 x cout bf16) is an intermediate: not algorithmic.  The pair is the unit: parent rows in, child rows out.
        prof.records.append(("gemm_first", (K, cin, cout), flops, 0.0, e0, e1))
        prof.records.append(("gent_reduce", (K, cin, cout), 0.0, 0.0, e1, e2))
        prof.records.append(("gen_transpose", (K, cin, cout), flops, 2.0 * n_par * cin + ((2.0 if want_bf16 else 0) + (4.0 if want_f32 else 0)) * n_out * cout, e0, e2))
        prof.impl_bytes["gen_transpose T (written + read back)"] = prof.impl_bytes.get("gen_transpose T (written + read back)", 0.0) + 4.0 * n_par * n_cols
    if of is None:
        return ob
    if ob is not None:
        of._pcg_bf16 = (ob, of._version)
    return of


def channelwise_conv(feats, cset, kernel, ksize):
    nbr, _ = cset.kernel_map(cset, ksize, False)
    f = feats.contiguous()
    out = torch.empty_like(f)
    k = kernel.detach().contiguous()
    call("pcg_channelwise_conv", ptr(f), f.shape[1], ptr(nbr), nbr.shape[1], ksize ** 3, cset.n, ptr(k), k.shape[1], ptr(out), stream())
    return out


def activation(feats, act):
    f = feats.contiguous()
    out = torch.empty_like(f)
    call("pcg_activation", ptr(f), f.numel(), int(act), ptr(out), stream())
    return out


def gdn(feats, beta, gamma, inverse):
    """x * (beta + |x| gamma^T)^-+1 (model/blocks.py:26-57).  "f32" mode: bf16x3 split products (fp32-grade).  Otherwise the norm
    is one bf16 product -- the precision of the convolutions on either side -- and the rows are also emitted as the bf16 operand
    image the next tensor-core layer reads (no separate conversion pass)."""
    f = feats.contiguous()
    out = torch.empty_like(f)
    c = f.shape[1]
    if _conv_mode != "f32" and c in (32, 64, 128) and os.environ.get("PCG_GDN_F32") != "1":
        ob = torch.empty((f.shape[0], c), dtype=torch.bfloat16, device=f.device) if c == cin_padded(c) else None
        call("pcg_gdn_fwd_bf16", ptr(f), f.shape[0], c, ptr(beta.detach().contiguous()), ptr(gamma.detach().contiguous()),
             int(bool(inverse)), ptr(out), ptr(ob), stream())
        if ob is not None:
            out._pcg_bf16 = (ob, out._version)
        return out
    call("pcg_gdn_fwd", ptr(f), f.shape[0], c, ptr(beta.detach().contiguous()), ptr(gamma.detach().contiguous()),
         int(bool(inverse)), ptr(out), stream())
    return out


def topk_mask(logits, coords, k_per_batch):
    """Bool mask of the k largest logits per batch item (transforms.py:228-254)."""
    lg = logits.reshape(-1).contiguous()
    n = lg.shape[0]
    mask = torch.zeros(n, dtype=torch.uint8, device=lg.device)
    if n == 0:
        return mask.view(torch.bool)
    lib = _lib.load()
    ws = scratch(lib.pcg_scratch_bytes(n), lg.device)
    # one entry in k = one batch item in the tensor (transforms.py:228-254 loops over the batch items): no batch filter, so the
    # radix passes read the logits only, not 16 bytes of coordinates per row
    single = len(k_per_batch) == 1
    if 1 < len(k_per_batch) <= 32:
        # all batch items in one set of radix passes (a frame coded as batched blocks, a training batch)
        ks = (ctypes.c_long * len(k_per_batch))(*[int(k) for k in k_per_batch])
        ws = scratch(lib.pcg_topk_mask_batched_scratch_bytes(n), lg.device)
        with hbm_span("topk", (n,), (4.0 + 4.0 + 1.0) * n):          # logits + batch index read once, mask written
            call("pcg_topk_mask_batched", ptr(lg), ptr(coords.contiguous()), n, len(k_per_batch), ctypes.cast(ks, ctypes.c_void_p), ptr(mask),
                 ptr(ws), ws.numel(), stream())
        return mask.view(torch.bool)
    for b, k in enumerate(k_per_batch):
        with hbm_span("topk", (n,), 4.0 * n + 1.0 * n):          # logits read once, mask written (the radix passes re-read: not algorithmic)
            call("pcg_topk_mask", ptr(lg), None if single else ptr(coords), n, b, int(k), ptr(mask), ptr(ws), ws.numel(), stream())
    return mask.view(torch.bool)          # the kernel writes 0 / 1


def prune_rows(coords, feats, mask, n_keep=None):
    """Order-preserving compaction of (coords, feats) under a bool mask (ME.MinkowskiPruning)."""
    n = mask.shape[0]
    dev = mask.device
    m8 = (mask.view(torch.uint8) if mask.dtype == torch.bool else mask.to(torch.uint8)).contiguous()
    lib = _lib.load()
    ws = scratch(lib.pcg_scratch_bytes(max(n, 1)) + 4 * max(n, 1), dev)
    cnt = ctl(dev)[3:4]                    # written by the compaction itself (never read back here)
    # size unknown: count first (one sync), then compact into exact buffers
    bound = int(m8.sum().item()) if n_keep is None else int(n_keep)
    oc = torch.empty((bound, 4), dtype=torch.int32, device=dev) if coords is not None else None
    of = torch.empty((bound, feats.shape[1]), dtype=torch.float32, device=dev) if feats is not None else None
    call("pcg_prune_rows", ptr(coords), ptr(feats.contiguous()) if feats is not None else None, n,
         feats.shape[1] if feats is not None else 0, ptr(m8), ptr(oc), ptr(of), ptr(cnt), ptr(ws), ws.numel(), stream())
    return oc, of


def gather_rows(coords, feats, rows):
    """(coords[rows], feats[rows]) for an int32 row list; either input may be None.  feats may be fp32 or bf16 (rows of a
    multiple of 8 channels, as the up-sampling kernels emit them); the gathered feature rows are always fp32."""
    n = int(rows.shape[0])
    dev = rows.device
    oc = torch.empty((n, 4), dtype=torch.int32, device=dev) if coords is not None else None
    of = torch.empty((n, feats.shape[1]), dtype=torch.float32, device=dev) if feats is not None else None
    src_bytes = 0 if feats is None else feats.element_size() * feats.shape[1]
    row_in = (16 if coords is not None else 0) + src_bytes
    row_out = (16 if coords is not None else 0) + (4 * feats.shape[1] if feats is not None else 0)
    with hbm_span("gather_rows", (n, row_in), (4.0 + row_in + row_out) * n):        # row list + rows read + rows written
        if feats is not None and feats.dtype == torch.bfloat16:
            call("pcg_gather_rows_bf16", ptr(coords), ptr(feats.contiguous()), feats.shape[1], ptr(rows), n, ptr(oc), ptr(of), stream())
        else:
            call("pcg_gather_rows", ptr(coords), ptr(feats.contiguous()) if feats is not None else None, feats.shape[1] if feats is not None else 0,
                 ptr(rows), n, ptr(oc), ptr(of), stream())
    return oc, of


def prune_rows_reordered(coords, feats, mask, n_keep, reorder=True):
    """Rows under `mask` in (batch, Z-order) order: the prune (ME.MinkowskiPruning) composed with the Z-order re-sort of the
    survivors.  The feature rows move once: kept-row list -> Morton permutation of the kept coordinates -> one gather.
    reorder=False: mask order kept (the plain prune, for bf16 candidate rows: the gather converts the survivors to fp32)."""
    n = mask.shape[0]
    dev = mask.device
    m8 = mask.view(torch.uint8) if mask.dtype == torch.bool else mask.to(torch.uint8)
    lib = _lib.load()
    n_keep = int(n_keep)
    rows = torch.empty(max(n_keep, 1), dtype=torch.int32, device=dev)
    cnt = ctl(dev)[3:4]
    ws = scratch(lib.pcg_scratch_bytes(max(n, 1)), dev)
    call("pcg_mask_to_rows", ptr(m8.contiguous()), n, ptr(rows), ptr(cnt), ptr(ws), ws.numel(), stream())
    rows = rows[:n_keep]
    if not reorder:
        return gather_rows(coords, feats, rows)
    oc, _ = gather_rows(coords, None, rows)
    if n_keep > 1:
        perm = sort_rows_i32(oc, morton=True, bits=LOCALITY_BITS)
        rows = torch.index_select(rows, 0, perm)
    return gather_rows(coords, feats, rows)


def features_at(feats, cset, query):
    """Trilinear feature lookup at float query coordinates (N,4) (ME features_at_coordinates)."""
    keys, vals, cap = cset.table()
    q = query.to(device=feats.device, dtype=torch.float32).contiguous()
    f = feats.contiguous()
    out = torch.empty((q.shape[0], f.shape[1]), dtype=torch.float32, device=f.device)
    with hbm_span("features_at", (q.shape[0], f.shape[1]), 16.0 * q.shape[0] + 8.0 * q.shape[0] * f.shape[1]):
        call("pcg_features_at", ptr(keys), ptr(vals), cap, ptr(f), f.shape[1], cset.ts, ptr(q), q.shape[0], ptr(out), stream())
    return out
