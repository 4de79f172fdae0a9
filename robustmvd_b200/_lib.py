"""ctypes binding of the C ABI declared in ``include/rmvd_b200.h``.

The shared library is the product; this module only extracts pointers, sizes and the current CUDA stream from
torch tensors.  There is no CPU fallback: a missing library or a non-CUDA tensor raises.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p, POINTER

from . import _build

_lib = None
ABI_VERSION = 8

EXPORTS = (
    "rmvd_abi_version", "rmvd_status_string", "rmvd_psweep_workspace_bytes", "rmvd_psweep_corr_fwd",
    "rmvd_psweep_corr_fwd_peers", "rmvd_psweep_corr_fwd_static", "rmvd_copy_rows", "rmvd_psweep_corr_fwd_push",
    "rmvd_psweep_push_queue_bytes",
    "rmvd_psweep_warp_fwd", "rmvd_psweep_bwd_workspace_bytes", "rmvd_psweep_corr_bwd",
    "rmvd_corr_grid_fwd", "rmvd_warp_grid_fwd", "rmvd_sampling_points_fwd", "rmvd_mvs_variance_fwd", "rmvd_mvs_variance_fwd_ws",
    "rmvd_mvs_variance_workspace_bytes",
    "rmvd_homo_warp_fwd", "rmvd_mvs_regress_fwd", "rmvd_softargmax_partial", "rmvd_softargmax_combine",
    "rmvd_softargmax_confidence", "rmvd_softargmax_slab_fused", "rmvd_soft_argmin_fwd", "rmvd_volume_reduce_fwd", "rmvd_view_fusion_fwd", "rmvd_homography_groupcorr_fwd", "rmvd_homography_warp_fwd", "rmvd_vis_cost_volume_fwd", "rmvd_cvp_proj_cost_fwd", "rmvd_cvp_projections", "rmvd_debug_div_check",
    "rmvd_mask_pack_u8", "rmvd_host_expand_mask", "rmvd_host_expand_mask_async",
    "rmvd_mask_pack_bits", "rmvd_host_expand_mask_bits", "rmvd_host_expand_mask_bits_async", "rmvd_psweep_corr_fwd_act", "rmvd_psweep_warp_bwd", "rmvd_mvs_variance_bwd", "rmvd_mvs_regress_bwd",
)


class RmvdError(RuntimeError):
    pass


def _declare(lib):
    P = c_void_p
    PP = POINTER(c_void_p)
    PI = POINTER(c_int)
    PL = POINTER(c_int64)
    lib.rmvd_abi_version.restype = c_int
    lib.rmvd_status_string.restype = c_char_p
    lib.rmvd_status_string.argtypes = [c_int]
    lib.rmvd_psweep_workspace_bytes.restype = c_size_t
    lib.rmvd_psweep_workspace_bytes.argtypes = [c_int, c_int, c_int, c_int, c_int, PI, PI]
    lib.rmvd_psweep_corr_fwd.argtypes = [P, PP, P, PP, PP, P, PL] + [c_int] * 5 + [PI, PI] + [c_int] * 4 + \
        [P, P, c_int, c_int, P, P, P, P, c_size_t, P]
    lib.rmvd_psweep_corr_fwd_static.argtypes = [P, PP, P, PP, PP, P, PL] + [c_int] * 5 + [PI, PI] + [c_int] * 4 + \
        [P, P, c_int, c_int, P]
    lib.rmvd_psweep_corr_fwd_peers.argtypes = [P, PP, P, PP, PP, P, PL] + [c_int] * 5 + [PI, PI] + [c_int] * 8 + \
        [PP, PP, PP, ctypes.c_uint32, c_int, P, P, P, c_size_t, P]
    lib.rmvd_psweep_push_queue_bytes.restype = c_size_t
    lib.rmvd_psweep_push_queue_bytes.argtypes = [c_int] * 5
    lib.rmvd_psweep_corr_fwd_push.argtypes = [P, PP, P, PP, PP, P, PL] + [c_int] * 5 + [PI, PI] + [c_int] * 8 + \
        [PP, PP, PP, ctypes.c_uint32, P, c_size_t, c_int, P, c_size_t, P, P]
    lib.rmvd_copy_rows.argtypes = [P, P, c_size_t, c_size_t, c_int, P]
    lib.rmvd_psweep_bwd_workspace_bytes.restype = c_size_t
    lib.rmvd_psweep_bwd_workspace_bytes.argtypes = [c_int, c_int, c_int, c_int, c_int, PI, PI]
    lib.rmvd_psweep_corr_bwd.argtypes = [P, PP, P, PP, PP, P, PL] + [c_int] * 5 + [PI, PI] + [c_int] * 4 + \
        [P, c_int, c_int, P, PP, P, c_size_t, P]
    lib.rmvd_psweep_warp_fwd.argtypes = [P, P, P, P, P, PL] + [c_int] * 10 + [P, P, c_int, c_int, P]
    lib.rmvd_sampling_points_fwd.argtypes = [P, P, P, P, PL] + [c_int] * 6 + [P] * 7
    lib.rmvd_corr_grid_fwd.argtypes = [P, P, P, P] + [c_int] * 9 + [P, P, P, c_size_t, P]
    lib.rmvd_warp_grid_fwd.argtypes = [P, P] + [c_int] * 8 + [P, P, P]
    lib.rmvd_mvs_variance_fwd.argtypes = [P, P, P] + [c_int] * 8 + [P, c_int, c_int, P]
    lib.rmvd_mvs_variance_workspace_bytes.restype = c_size_t
    lib.rmvd_mvs_variance_workspace_bytes.argtypes = [c_int] * 5
    lib.rmvd_mvs_variance_fwd_ws.argtypes = [P, P, P] + [c_int] * 8 + [P, c_int, c_int, P, c_size_t, P]
    lib.rmvd_homo_warp_fwd.argtypes = [P, P, P, P] + [c_int] * 5 + [P, P, P, P]
    lib.rmvd_mvs_regress_fwd.argtypes = [P, P] + [c_int] * 4 + [P, P, P, P, P, P]
    lib.rmvd_softargmax_partial.argtypes = [P, P] + [c_int] * 5 + [P, P]
    lib.rmvd_softargmax_combine.argtypes = [P] + [c_int] * 4 + [P, P]
    lib.rmvd_softargmax_confidence.argtypes = [P, P] + [c_int] * 6 + [P, P]
    lib.rmvd_softargmax_slab_fused.argtypes = [PP, P] + [c_int] * 6 + [PP, PP, ctypes.c_uint32, P, P, P, P, P, P]
    lib.rmvd_soft_argmin_fwd.argtypes = [P] + [c_int] * 4 + [c_float, P, P, P, P, P]
    lib.rmvd_volume_reduce_fwd.argtypes = [P, P] + [c_int] * 5 + [P, P]
    lib.rmvd_view_fusion_fwd.argtypes = [PP, PP, PP] + [c_int] * 5 + [P, P, P, P]
    lib.rmvd_homography_groupcorr_fwd.argtypes = [P, P, P, PL] + [c_int] * 6 + [P, P]
    lib.rmvd_homography_warp_fwd.argtypes = [P, P, PL] + [c_int] * 5 + [P, P]
    lib.rmvd_vis_cost_volume_fwd.argtypes = [P, P, P, P, P, PL, P, PL] + [c_int] * 7 + [P, P, c_size_t, P]
    lib.rmvd_cvp_proj_cost_fwd.argtypes = [P, P, P] + [c_int] * 6 + [P, P]
    lib.rmvd_cvp_projections.argtypes = [P, P, P, P, c_int, c_int, P, P]
    lib.rmvd_debug_div_check.argtypes = [P, P, c_int, P, P, P]
    lib.rmvd_psweep_corr_fwd_act.argtypes = [P, PP, c_float, P, P, PP, PP, P, PL] + [c_int] * 5 + [PI, PI] + [c_int] * 4 + \
        [P, P, c_int, c_int, P, c_size_t, P]
    lib.rmvd_mvs_variance_bwd.argtypes = [P, P, P, P] + [c_int] * 6 + [P, P]
    lib.rmvd_mvs_regress_bwd.argtypes = [P, P, P] + [c_int] * 4 + [P, P]
    lib.rmvd_psweep_warp_bwd.argtypes = [P, P, P, P, PL] + [c_int] * 10 + [P, c_int, c_int, P, P]
    lib.rmvd_mask_pack_u8.argtypes = [P, P, c_size_t, P]
    lib.rmvd_host_expand_mask.argtypes = [P, P, c_size_t, c_int]
    lib.rmvd_host_expand_mask_async.argtypes = [P, P, c_size_t, c_int, P]
    lib.rmvd_mask_pack_bits.argtypes = [P, P, c_size_t, P]
    lib.rmvd_host_expand_mask_bits.argtypes = [P, P, c_size_t, c_int]
    lib.rmvd_host_expand_mask_bits_async.argtypes = [P, P, c_size_t, c_int, P]
    for name in EXPORTS:
        if name not in ("rmvd_abi_version", "rmvd_status_string", "rmvd_psweep_workspace_bytes",
                        "rmvd_psweep_bwd_workspace_bytes", "rmvd_mvs_variance_workspace_bytes", "rmvd_psweep_push_queue_bytes"):
            getattr(lib, name).restype = c_int


def load():
    """Load (once) ``csrc/librmvd_b200.so``.  Raises RmvdError when it has not been built."""
    global _lib
    if _lib is None:
        path = os.environ.get("RMVD_LIB_PATH") or _build.LIB    # RMVD_LIB_PATH: another build of the same sources (A/B timing only)
        if path == _build.LIB:
            if not os.path.exists(path):
                raise RmvdError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(robustmvd_b200 has no CPU or PyTorch fallback)")
            if _build.is_stale():
                # a source file is newer than the library: never run the old binary silently
                try:
                    _build.build()
                except RuntimeError as e:
                    raise RmvdError(f"{path} is older than its sources and could not be rebuilt: {e}") from e
        lib = ctypes.CDLL(path)
        _declare(lib)
        if lib.rmvd_abi_version() != ABI_VERSION:
            raise RmvdError(f"{path}: ABI version {lib.rmvd_abi_version()}, this binding expects {ABI_VERSION}")
        _lib = lib
    return _lib


def check(status, what):
    if status != 0:
        raise RmvdError(f"{what}: {load().rmvd_status_string(status).decode()} (status {status})")


def ptr(t):
    """Device pointer of a contiguous float32/int CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RmvdError("robustmvd_b200 kernels need CUDA tensors (there is no CPU path)")
    if not t.is_contiguous():
        raise RmvdError("tensor must be contiguous")
    return c_void_p(t.data_ptr())


def ptr_array(tensors):
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = ptr(t).value
    return arr


def raw_ptr_array(addresses):
    """Host array of device pointers given as integers (peer-mapped addresses of symmetric memory)."""
    arr = (c_void_p * len(addresses))()
    for i, a in enumerate(addresses):
        arr[i] = int(a)
    return arr


def ptr_array_opt(tensors):
    """Like ptr_array, with None entries -> NULL."""
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = None if t is None else ptr(t).value
    return arr


def int_array(vals):
    return (c_int * len(vals))(*vals)


def i64_array(vals):
    return (c_int64 * len(vals))(*vals)


def stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)
