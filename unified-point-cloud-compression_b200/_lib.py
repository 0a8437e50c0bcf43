"""ctypes binding of libpcg.so (include/pcg.h) -- the only way the host code reaches the GPU kernels.

There is no CPU fallback: if the shared object is missing, or a call is made without CUDA
tensors, this raises.  PyTorch is used for device memory, streams and autograd plumbing only.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpcg.so")

_c = ctypes
_P, _L, _I, _F, _SZ = _c.c_void_p, _c.c_long, _c.c_int, _c.c_float, _c.c_size_t

# name -> (restype, argtypes); mirrors include/pcg.h one to one
_PROTOS = {
    "pcg_last_error": (_c.c_char_p, []),
    "pcg_version": (_I, []),
    "pcg_launch_count": (_L, []),
    "pcg_hash_capacity": (_L, [_L]),
    "pcg_nbr_pitch": (_L, [_L]),
    "pcg_tc_cin_padded": (_I, [_I]),
    "pcg_tc_set_mt_min_tiles": (_L, [_L]),
    "pcg_scratch_bytes": (_SZ, [_L]),
    "pcg_hash_build": (_I, [_P, _L, _P, _P, _L, _P, _P]),
    "pcg_hash_lookup": (_I, [_P, _P, _L, _P, _L, _P, _P]),
    "pcg_unique_first": (_I, [_P, _L, _P, _P, _L, _P, _P, _P, _P, _P, _SZ, _P]),
    "pcg_stride_coords": (_I, [_P, _L, _I, _P, _P, _L, _P, _P, _P, _P, _SZ, _P]),
    "pcg_generate_coords": (_I, [_P, _L, _I, _I, _I, _P, _P, _L, _P, _L, _P, _P, _P, _SZ, _P]),
    "pcg_kernel_map": (_I, [_P, _P, _L, _P, _L, _I, _I, _I, _I, _P, _L, _P, _P]),
    "pcg_sort_rows": (_I, [_P, _L, _I, _P, _P, _SZ, _P]),
    "pcg_sort_rows_bits": (_I, [_P, _L, _I, _I, _I, _P, _P, _SZ, _P]),
    "pcg_sort_rows_scratch_bytes": (_SZ, [_L]),
    "pcg_gather_block": (_I, [_P, _P, _L, _I, _P, _P, _P]),
    "pcg_conv_fwd_f32": (_I, [_P, _L, _I, _P, _L, _I, _L, _P, _P, _I, _I, _P, _P]),
    "pcg_conv_bwd_weight_f32": (_I, [_P, _I, _P, _L, _P, _I, _L, _P, _I, _P, _P]),
    "pcg_conv_bwd_weight_f32_scratch_bytes": (_SZ, [_I, _I, _I, _L]),
    "pcg_conv_bwd_weight_f32_sliced": (_I, [_P, _I, _P, _L, _P, _I, _L, _P, _I, _P, _SZ, _P, _P]),
    "pcg_conv_bwd_weight_tc_scratch_bytes": (_SZ, [_I, _I, _I, _L]),
    "pcg_conv_bwd_weight_tc": (_I, [_P, _I, _P, _I, _P, _L, _P, _I, _L, _P, _SZ, _P, _P]),
    "pcg_packed_weight_bytes": (_SZ, [_I, _I, _I]),
    "pcg_pack_weights_bf16": (_I, [_P, _I, _I, _I, _P, _P]),
    "pcg_conv_fwd_tc": (_I, [_P, _L, _I, _P, _L, _P, _I, _L, _P, _P, _I, _I, _P, _P, _P]),
    "pcg_conv_fwd_tc_proj": (_I, [_P, _L, _I, _P, _L, _P, _I, _L, _P, _P, _I, _I, _P, _P, _P, _I, _P, _L, _P]),
    "pcg_gather_sum": (_I, [_P, _L, _P, _L, _I, _L, _P, _I, _P, _P]),
    "pcg_im2col_bf16": (_I, [_P, _I, _P, _L, _I, _L, _I, _P, _P]),
    "pcg_mma_weight_bytes": (_SZ, [_I, _I, _I]),
    "pcg_pack_weights_mma": (_I, [_P, _I, _I, _I, _P, _P]),
    "pcg_conv_fwd_mma": (_I, [_P, _L, _I, _P, _L, _I, _L, _P, _P, _I, _I, _P, _P, _P]),
    "pcg_gemm_bf16_grouped": (_I, [_P, _L, _I, _P, _L, _P, _P]),
    "pcg_cells_parent_map": (_I, [_P, _L, _I, _P, _P, _L, _L, _P, _L, _P, _P]),
    "pcg_cells_scratch_bytes": (_SZ, [_L]),
    "pcg_cells_info": (_I, [_P, _L, _L, _P, _P, _P, _P, _SZ, _P]),
    "pcg_cells_emit": (_I, [_P, _P, _P, _L, _I, _P, _P, _P]),
    "pcg_cells_kernel_map3": (_I, [_P, _P, _L, _P, _P, _L, _I, _P, _P, _P, _L, _P, _P]),
    "pcg_cells_reduce": (_I, [_P, _I, _P, _L, _P, _P, _L, _P, _I, _P, _P, _P]),
    "pcg_f32_to_bf16": (_I, [_P, _P, _L, _P]),
    "pcg_f32_to_bf16_rows": (_I, [_P, _L, _I, _I, _P, _P]),
    "pcg_channelwise_conv": (_I, [_P, _I, _P, _L, _I, _L, _P, _I, _P, _P]),
    "pcg_activation": (_I, [_P, _L, _I, _P, _P]),
    "pcg_gdn_fwd": (_I, [_P, _L, _I, _P, _P, _I, _P, _P]),
    "pcg_gdn_fwd_bf16": (_I, [_P, _L, _I, _P, _P, _I, _P, _P, _P]),
    "pcg_finalize": (_I, [_P, _P, _I, _L, _P, _P]),
    "pcg_topk_mask": (_I, [_P, _P, _L, _I, _L, _P, _P, _SZ, _P]),
    "pcg_topk_mask_batched_scratch_bytes": (_SZ, [_L]),
    "pcg_topk_mask_batched": (_I, [_P, _P, _L, _I, _P, _P, _P, _SZ, _P]),
    "pcg_prune_rows": (_I, [_P, _P, _L, _I, _P, _P, _P, _P, _P, _SZ, _P]),
    "pcg_nn_search": (_I, [_P, _P, _L, _P, _P, _P, _L, _I, _P, _P, _P, _P, _P]),
    "pcg_mask_to_rows": (_I, [_P, _L, _P, _P, _P, _SZ, _P]),
    "pcg_mask_to_rows_static": (_I, [_P, _L, _P, _P, _P]),
    "pcg_gather_rows": (_I, [_P, _P, _I, _P, _L, _P, _P, _P]),
    "pcg_gather_rows_bf16": (_I, [_P, _P, _I, _P, _L, _P, _P, _P]),
    "pcg_color_jitter_scratch_bytes": (_SZ, []),
    "pcg_color_jitter": (_I, [_P, _L, _I, _I, _I, _I, _F, _F, _F, _F, _P, _P, _P]),
    "pcg_rotate_floor": (_I, [_P, _L, _P, _F, _F, _P, _P, _P]),
    "pcg_features_at": (_I, [_P, _P, _L, _P, _I, _I, _P, _L, _P, _P]),
    "pcg_eb_likelihood": (_I, [_P, _L, _I, _L, _L, _P, _F, _P, _P]),
    "pcg_quantize": (_I, [_P, _P, _L, _I, _L, _L, _L, _L, _I, _P, _P, _P]),
    "pcg_dequantize": (_I, [_P, _P, _L, _I, _L, _L, _L, _L, _P, _P]),
    "pcg_gc_likelihood": (_I, [_P, _P, _P, _L, _I, _L, _L, _F, _F, _P, _P]),
    "pcg_gc_build_indexes": (_I, [_P, _L, _I, _L, _L, _P, _I, _F, _P, _P]),
    "pcg_gc_dequant_offsets": (_I, [_P, _P, _P, _P, _P, _L, _I, _L, _L, _L, _L, _P, _F, _P, _P]),
}

_lib = None


def exported_symbols():
    return sorted(_PROTOS)


def load():
    """Load libpcg.so (built by __graft_entry__.build()).  Fails loudly when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback for the CUDA path)")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def launch_count():
    return int(load().pcg_launch_count())


def check(rc):
    if rc != 0:
        raise RuntimeError("libpcg: " + load().pcg_last_error().decode())


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libpcg needs CUDA tensors: the graft has no CPU path")
    return ctypes.c_void_p(t.data_ptr())


def raw_stream():
    """cudaStream_t of torch's current stream on the current device, as an int (the raw getter: ~0.3 us; building a
    torch.cuda.Stream object per C-ABI call cost several us of the ~15 us a launch takes on the host)."""
    try:
        return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())
    except AttributeError:                       # private getters renamed in some torch build: the public (slower) path
        return torch.cuda.current_stream().cuda_stream


def stream():
    return ctypes.c_void_p(raw_stream())


def call(name, *args):
    check(getattr(load(), name)(*args))
