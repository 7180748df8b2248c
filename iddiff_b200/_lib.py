"""ctypes binding of libiddiff_b200.so (the C ABI declared in include/iddiff_b200.h).

There is NO fallback: if the library is missing or a call is rejected, an exception is raised.
The CPU implementation of this path is the reference itself (used only as oracle / baseline under
oracle/, never from here).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libiddiff_b200.so")

_lock = threading.Lock()
_lib = None

_vp = C.c_void_p
_i = C.c_int
_f = C.c_float
_d = C.c_double
_u64 = C.c_uint64
_ll = C.c_longlong

FMT_BF16, FMT_FP16 = 0, 1
SPECTRUM_JACOBI, SPECTRUM_TRIDIAG = 0, 1

# name -> (restype, argtypes); must list every symbol of include/iddiff_b200.h
SIGNATURES = {
    "idb_abi_version": (_i, []),
    "idb_last_error": (C.c_char_p, []),
    "idb_perturb": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _f, _f, _u64, _u64, _vp, _i, _f, _vp, _vp, _vp]),
    "idb_perturb_rows": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _f, _f, _u64, _u64, C.c_uint32, _vp, _i, _f, _vp, _vp, _vp]),
    "idb_pack_weight": (_i, [_vp, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp]),
    "idb_anchor_linear": (_i, [_vp, _i, _i, _f, _vp, _i, _vp, _vp, _i, _f, _vp, _i, _i, _i, _i, _vp]),
    "idb_fcn_hidden_layer": (_i, [_vp, _vp, _vp, _vp, _i, _f, _f, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "idb_fcn_output_layer": (_i, [_vp, _vp, _vp, _vp, _i, _f, _i, _i, _i, _vp, _vp]),
    "idb_fcn_anchor_layer": (_i, [_vp, _vp, _vp, _vp, _i, _f, _f, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "idb_fcn_scores": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp]),
    "idb_ksphere_gt_score": (_i, [_vp, _vp, _f, _i, _i, _i, _vp, _vp]),
    "idb_center_gram": (_i, [_vp, _ll, _i, _i, _i, _i, _d, _vp, _vp, _vp]),
    "idb_column_sums": (_i, [_vp, _ll, _i, _i, _i, _i, _vp, _vp]),
    "idb_gram_about": (_i, [_vp, _ll, _i, _i, _i, _i, _d, _vp, _vp, _vp]),
    "idb_dsm_sample": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _u64, _u64, _vp, _vp, _vp, _vp]),
    "idb_train_pack": (_i, [_vp, _i, _i, _i, _vp, _i, _i, _i, _i, _f, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "idb_dsm_loss": (_i, [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "idb_column_sums_f32": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "idb_grad_sumsq": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "idb_adam_step": (_i, [_vp, _i, _vp, _i, _vp, _vp, _i, _i, _f, _i, _f, _f, _f, _ll, _vp, _f, _vp, _vp]),
    "idb_spectrum_workspace_bytes": (C.c_size_t, [_i, _i]),
    "idb_spectrum": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "idb_id_from_spectrum": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "idb_upfirdn2d": (_i, [_vp, _vp] + [_i] * 14 + [_vp, _vp]),
    "idb_fused_bias_act": (_i, [_vp, _vp, _vp, _i, _i, _f, _f, _ll, _i, _i, _vp, _vp]),
    "idb_groupnorm_stats": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _d, _vp, _vp, _vp]),
    "idb_conv_prep": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _i, _f, _vp, _vp, _i, _vp, _vp]),
    "idb_conv_prep_fir": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _vp, _vp, _i, C.POINTER(C.c_float), _i, _i, _i, _i, _i, _i, _i, _f,
                               _vp, _vp, _i, _vp, _vp, _vp]),
    "idb_conv2d": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, C.POINTER(C.c_int), _i, _i, _f, _f, _vp, _vp, _i,
                        _vp, _i, _vp]),
    "idb_conv_stats_slots": (_i, [_i, _i]),
    "idb_conv2d_stats": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, C.POINTER(C.c_int), _i, _i, _f, _f, _vp,
                              _vp, _i, _vp, _i, _vp, _vp]),
    "idb_groupnorm_from_conv_stats": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _i, _i, _d, _vp, _vp]),
    "idb_row_gram": (_i, [_vp, _i, _i, _i, _d, _vp, _vp, _vp]),
    "idb_spectrum_large_workspace_bytes": (C.c_size_t, [_i]),
    "idb_spectrum_large": (_i, [_vp, _i, _i, _vp, _vp, _vp]),
    "idb_spectrum_large_variant": (_i, [_vp, _i, _i, _vp, _vp, _i, _vp]),
    "idb_spectrum_large_static_max_n": (_i, []),
    "idb_spectrum_large_static": (_i, [_vp, _i, _i, _vp, _vp]),
    "idb_attention_tc_workspace_bytes": (C.c_size_t, [_i, _i, _i]),
    "idb_attention_tc": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _i, _vp]),
    "idb_attention": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _i, _vp]),
}


class IdbError(RuntimeError):
    pass


def load():
    """Loads the shared library (once).  Raises IdbError when it has not been built."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise IdbError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU or PyTorch fallback for this path)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib


# CUDA kernels each entry point enqueues (for bench.py's `gpu_launches` claim)
KERNELS_PER_CALL = {"idb_perturb": 1, "idb_perturb_rows": 1, "idb_pack_weight": 1, "idb_anchor_linear": 1, "idb_fcn_hidden_layer": 1,
                    "idb_fcn_output_layer": 1, "idb_fcn_anchor_layer": 1, "idb_fcn_scores": 1, "idb_ksphere_gt_score": 1, "idb_dsm_sample": 1, "idb_train_pack": 1, "idb_dsm_loss": 2, "idb_column_sums_f32": 1, "idb_grad_sumsq": 1,
                    "idb_adam_step": 1, "idb_center_gram": 2, "idb_column_sums": 1, "idb_gram_about": 1,
                    "idb_spectrum": 2, "idb_id_from_spectrum": 1, "idb_upfirdn2d": 1, "idb_fused_bias_act": 1,
                    "idb_groupnorm_stats": 2, "idb_conv_prep": 1, "idb_conv_prep_fir": 1, "idb_conv2d": 1, "idb_conv2d_stats": 1, "idb_groupnorm_from_conv_stats": 1, "idb_attention": 1, "idb_attention_tc": 4,
                    "idb_row_gram": 3, "idb_spectrum_large": 2, "idb_spectrum_large_variant": 2,
                    "idb_spectrum_large_static": 2}
launch_count = 0


def call(name: str, *args):
    """Calls an int-returning entry point and raises IdbError with idb_last_error() on failure."""
    global launch_count
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise IdbError(f"{name} failed ({rc}): {lib.idb_last_error().decode()}")
    launch_count += KERNELS_PER_CALL.get(name, 0)


def ptr(t):
    """Device pointer of a CUDA tensor (or None)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise IdbError("iddiff_b200 kernels take CUDA tensors only (no CPU path)")
    return t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream
