"""ctypes binding of the in-tree C-ABI library `libdeltaconv_b200.so` (include/deltaconv_b200.h).

There is no CPU fallback: if the library is missing, or a tensor is not a contiguous CUDA fp32 /
int32 / int64 tensor as the entry point expects, the call raises.  PyTorch is used only for
device memory and the current stream.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libdeltaconv_b200.so')

_P, _I, _L, _F, _Z = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t

# name -> argtypes (restype is int unless listed in _RESTYPES)
_SIGNATURES = {
    'dc_abi_version': [],
    'dc_device_sm_count': [],
    'dc_launch_count': [],
    'dc_reset_launch_count': [],
    'dc_cloud_ptr': [_P, _I, _I, _P, _P],
    'dc_knn': [_P, _P, _I, _I, _I, _P, _P, _P, _P],
    'dc_morton_keys': [_P, _P, _I, _I, _P, _P, _Z, _P],
    'dc_knn_grid_ws_bytes': [_I, _I],
    'dc_geodesic_fps_ws_bytes': [_I],
    'dc_geodesic_fps': [_P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P, _Z, _P],
    'dc_set_knn_grid_density': [_F],
    'dc_set_knn_merge_min': [_I, _I],
    'dc_knn_grid': [_P, _P, _I, _I, _I, _P, _P, _P, _P, _Z, _P],
    'dc_ell_from_edge_index': [_P, _I, _I, _P, _P],
    'dc_edge_index_from_ell': [_P, _I, _I, _P, _P],
    'dc_ell_transpose_ws_bytes': [_I],
    'dc_ell_transpose': [_P, _I, _I, _P, _P, _P, _Z, _P],
    'dc_tangent_basis': [_P, _I, _P, _P, _P],
    'dc_estimate_basis': [_P, _P, _I, _I, _P, _P, _P, _P, _P],
    'dc_build_ws_bytes': [_I, _I],
    'dc_build_ws_min_bytes': [_I, _I],
    'dc_build_grad_div': [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _F, _F, _I, _I, _P, _P, _P, _P, _P, _P, _P, _Z, _P],
    'dc_set_build_chunk': [_I],
    'dc_build_static_max_points': [],
    'dc_build_static_max_clouds': [],
    'dc_build_grad_div_static': [_P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _F, _F, _I, _I, _P, _P, _P],
    'dc_set_spmm_staged': [_I],
    'dc_grad_fwd': [_P, _P, _I, _I, _I, _P, _L, _P, _L, _P, _P, _I, _I, _P],
    'dc_grad_bwd': [_P, _P, _P, _I, _I, _I, _P, _L, _P, _L, _P, _P, _I, _I, _P],
    'dc_divcurlnorm_fwd': [_P, _P, _I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _L, _P, _L, _P, _L, _P, _P, _I, _I, _P],
    'dc_divcurlnorm_bwd': [_P, _P, _P, _I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _L, _P, _L, _P, _P, _I, _I, _P],
    'dc_hodge_fwd': [_P, _P, _I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _L, _P, _L, _P, _P, _I, _I, _P],
    'dc_hodge_bwd': [_P, _P, _P, _I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _P, _I, _I, _P],
    'dc_gathermax_fwd': [_P, _I, _I, _I, _P, _L, _P, _L, _P, _P],
    'dc_gathermax_bwd': [_P, _P, _I, _I, _I, _P, _P, _L, _P, _L, _P],
    'dc_gathermax_bn_fwd': [_P, _I, _I, _I, _P, _L, _P, _P, _P, _P, _F, _P, _L, _P, _P],
    'dc_edge_diff_fwd': [_P, _I, _I, _I, _P, _L, _P, _P],
    'dc_edge_diff_bwd': [_P, _P, _I, _I, _I, _P, _P, _L, _P],
    'dc_segmax_fwd': [_P, _I, _I, _I, _P, _P, _L, _P, _P],
    'dc_segmax_bwd': [_P, _L, _I, _I, _I, _P, _P, _P],
    'dc_segmax_bn_fwd': [_P, _I, _I, _I, _P, _P, _P, _P, _P, _F, _P, _L, _P, _P],
    'dc_segmax_bn_bwd': [_P, _I, _I, _I, _P, _P, _L, _P, _P, _P, _P, _F, _I, _P, _P, _P, _P, _Z, _P],
    'dc_central_max_cin': [],
    'dc_central_ws_bytes': [_I, _I, _I, _I],
    'dc_central_stats': [_P, _I, _I, _I, _I, _P, _L, _P, _F, _F, _P, _P, _P, _P, _P, _P, _P, _Z, _P],
    'dc_central_fwd': [_P, _I, _I, _I, _I, _P, _L, _P, _P, _P, _P, _P, _F, _P, _L, _P, _P],
    'dc_central_bwd': [_P, _P, _P, _I, _I, _I, _I, _P, _L, _P, _P, _P, _P, _P, _P, _F, _I, _P, _P, _L, _P, _P, _P, _P, _L, _P, _Z, _P],
    'dc_central_edge_fwd': [_P, _I, _I, _I, _I, _P, _L, _P, _P, _P, _P, _P, _F, _P, _P],
    'dc_central_edge_bwd': [_P, _I, _I, _I, _I, _P, _L, _P, _P, _P, _P, _P, _P, _F, _I, _P, _L, _P, _P, _P, _P, _Z, _P],
    'dc_colreduce_ws_bytes': [_I],
    'dc_bn_stats': [_P, _L, _L, _I, _I, _F, _F, _P, _P, _P, _P, _P, _P, _Z, _P],
    'dc_bn_act_fwd': [_P, _L, _L, _I, _P, _P, _P, _P, _F, _P, _L, _P, _L, _P],
    'dc_bn_act_bwd': [_P, _L, _P, _L, _L, _I, _P, _P, _P, _P, _F, _I, _P, _P, _P, _L, _P, _Z, _P],
    'dc_col_sum': [_P, _L, _L, _I, _P, _P, _Z, _P],
    'dc_vecnonlin_fwd': [_P, _L, _L, _I, _P, _P, _P, _P, _P, _P, _L, _P],
    'dc_vecnonlin_bwd': [_P, _L, _P, _L, _L, _I, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P, _L, _P, _Z, _P],
    'dc_pool_fwd': [_P, _L, _P, _I, _I, _P, _L, _P, _L, _P, _P],
    'dc_pool_bwd': [_P, _I, _I, _I, _P, _P, _L, _P, _L, _P, _L, _P],
    'dc_bn_act_pool_fwd': [_P, _L, _P, _I, _I, _P, _P, _P, _P, _F, _P, _L, _P, _L, _P, _P],
    'dc_bn_act_pool_bwd': [_P, _L, _L, _I, _P, _I, _P, _P, _L, _P, _L, _P, _P, _P, _P, _F, _P, _P, _P, _L, _P, _Z, _P],
    'dc_sgd_step': [_P, _P, _P, _L, _P, _F, _F, _F, _P],
    'dc_adam_step': [_P, _P, _P, _P, _L, _P, _P, _F, _F, _F, _F, _F, _P],
    'dc_add_rows': [_P, _L, _P, _L, _L, _I, _P, _L, _P],
    'dc_bcast_rows': [_P, _L, _P, _I, _I, _I, _P, _L, _P],
    'dc_bcast_rows_add': [_P, _L, _P, _I, _I, _I, _P, _L, _P],
    'dc_set_gemm_mode': [_I],
    'dc_get_gemm_mode': [],
    'dc_set_gemm_no_ts': [_I],
    'dc_set_gemm_keep_raw_hi': [_I],
    'dc_set_gemm_tma_store': [_I],
    'dc_set_gemm_pair': [_I],
    'dc_set_gemm_trace': [_P],
    'dc_set_gemm_debug': [_I, _I],
    'dc_gemm_ws_bytes': [_L, _L, _L],
    'dc_gemm_tc_raw': [_P, _L, _I, _P, _L, _I, _P, _L, _I, _L, _P, _L, _P, _Z, _P],
    'dc_linear_fwd': [_P, _L, _P, _L, _P, _L, _I, _I, _P, _L, _P, _Z, _P],
    'dc_linear_bwd_input': [_P, _L, _P, _L, _L, _I, _I, _P, _L, _P, _Z, _P],
    'dc_linear_bwd_weight': [_P, _L, _P, _L, _L, _I, _I, _P, _L, _P, _Z, _P],
    'dc_expand_vec_weight': [_P, _I, _I, _I, _P, _P, _P],
    'dc_fold_vec_weight_grad': [_P, _I, _I, _I, _P, _P, _P],
}
_RESTYPES = {'dc_launch_count': ctypes.c_longlong, 'dc_reset_launch_count': None,
             'dc_ell_transpose_ws_bytes': _Z, 'dc_build_ws_bytes': _Z, 'dc_build_ws_min_bytes': _Z, 'dc_colreduce_ws_bytes': _Z,
             'dc_gemm_ws_bytes': _Z, 'dc_central_ws_bytes': _Z, 'dc_knn_grid_ws_bytes': _Z, 'dc_geodesic_fps_ws_bytes': _Z}

_lib = None
ABI_VERSION = 2          # dc_abi_version() of the library these signatures describe


class DeltaConvLibraryError(RuntimeError):
    pass


def load():
    """Load (once) and return the ctypes handle; raise loudly if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DeltaConvLibraryError(
            f"{LIB_PATH} not found — build it with `python __graft_entry__.py build` "
            f"(or `make -C deltaconv_b200/csrc`). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    lib.dc_last_error.restype = ctypes.c_char_p
    lib.dc_last_error.argtypes = []
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, _I)
    if lib.dc_abi_version() != ABI_VERSION:
        raise DeltaConvLibraryError(f"{LIB_PATH} has ABI version {lib.dc_abi_version()}, this package binds version {ABI_VERSION}: "
                                    f"rebuild it (`python __graft_entry__.py build`)")
    _lib = lib
    return lib


def exported_symbols():
    return ['dc_last_error'] + list(_SIGNATURES)


def stream_ptr():
    """Raw handle of the current CUDA stream of the current device (the fast private accessor when torch has it: an eager
    step makes ~200 library calls, each of which needs it)."""
    try:
        return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())
    except AttributeError:
        return torch.cuda.current_stream().cuda_stream


def ptr(t):
    """Device address of a tensor (None stays None).  A CPU tensor — e.g. a module parameter that was never moved to the
    device — would hand the kernels a host address: refused here, once, for every entry point."""
    if t is None:
        return None
    if not t.is_cuda:
        raise DeltaConvLibraryError(f"expected a CUDA tensor, got a {t.dtype} tensor of shape {tuple(t.shape)} on {t.device} "
                                    f"(deltaconv_b200 has no CPU path; move the module / data to the GPU)")
    return t.data_ptr()


def check(t, dtype, name):
    if not (torch.is_tensor(t) and t.is_cuda):
        raise DeltaConvLibraryError(f"{name}: expected a CUDA tensor (the CUDA library has no CPU path)")
    if t.dtype != dtype:
        raise DeltaConvLibraryError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise DeltaConvLibraryError(f"{name}: expected a contiguous tensor")
    return t


PROFILER = None      # optional object with .begin(name, args) -> token and .end(token); see bench.py


def call(name, *args):
    """Call an int-returning entry point on the current stream (appended as the last argument)."""
    lib = load()
    if PROFILER is not None:
        token = PROFILER.begin(name, args)
        rc = getattr(lib, name)(*args, stream_ptr())
        PROFILER.end(token)
    else:
        rc = getattr(lib, name)(*args, stream_ptr())
    if rc != 0:
        msg = lib.dc_last_error().decode(errors='replace')
        raise DeltaConvLibraryError(f"{name} failed (code {rc}): {msg}")


_query_cache = {}
_UNCACHED_QUERIES = {'dc_knn_grid_ws_bytes', 'dc_get_gemm_mode', 'dc_launch_count'}      # depend on run-time tuning switches / state


def query(name, *args):
    """Value of a pure size / limit query (`dc_*_ws_bytes`, ...), memoised: the answers depend on the arguments only."""
    if name in _UNCACHED_QUERIES:
        return getattr(load(), name)(*args)
    key = (name,) + args
    val = _query_cache.get(key)
    if val is None:
        val = getattr(load(), name)(*args)
        if len(_query_cache) < 65536:
            _query_cache[key] = val
    return val


def rows_ld(t):
    """(tensor, leading dimension) of a 2-D fp32 CUDA tensor whose rows are contiguous; column
    slices of a wider row-major buffer are accepted without a copy."""
    if not (t.is_cuda and t.dtype == torch.float32 and t.dim() == 2):
        raise DeltaConvLibraryError("expected a 2-D fp32 CUDA tensor")
    if t.size(1) > 1 and t.stride(1) != 1 or (t.size(0) > 1 and t.stride(0) < t.size(1)):
        t = t.contiguous()
    ld = t.stride(0) if t.size(0) > 1 else max(t.size(1), 1)
    if t.size(1) == 1 and t.size(0) > 1 and t.stride(0) < 1:
        t = t.contiguous()
        ld = 1
    return t, ld
