"""ctypes binding of ``libcmf_b200.so`` -- the C ABI declared in ``include/cmf_b200.h``.

There is no CPU fallback: if the library is missing or a launch fails this raises.  The Python
layer above only ever passes raw device pointers, sizes and the current CUDA stream.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libcmf_b200.so")
ABI_VERSION = 1

_vp, _i, _ll, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t

# name -> (restype, argtypes); mirrors include/cmf_b200.h one to one
SIGNATURES = {
    "cmf_version": (_i, []),
    "cmf_last_error_string": (C.c_char_p, []),
    "cmf_launch_count": (C.c_ulonglong, []),
    "cmf_coupling_jvp_fwd": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_coupling_jvp_bwd_workspace": (_sz, [_i, _i, _i, _i]),
    "cmf_coupling_jvp_bwd": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                  _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "cmf_coupling_jvp_bwd_static_max_bytes": (_sz, []),
    "cmf_coupling_jvp_bwd_static": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                         _vp, _vp, _vp, _vp, _vp]),
    "cmf_gather_rows": (_i, [_ll, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "cmf_gram_fwd": (_i, [_i, _i, _i, _vp, _vp, _i, _vp]),
    "cmf_gram_fwd_path": (_i, [_i, _i, _i, _vp, _vp, _i, _i, _vp]),
    "cmf_gram_fwd2": (_i, [_i, _i, _i, _vp, _i, _vp, _vp, _i, _vp]),
    "cmf_gram_bwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp]),
    "cmf_gram_bwd_path": (_i, [_i, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "cmf_chol_logdet_fwd": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_chol_logdet_bwd": (_i, [_i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "cmf_rqs_coupling_jvp_fwd": (_i, [_i, _i, _i, _i, C.c_float] + [_vp] * 12),
    "cmf_rqs_coupling_jvp_bwd": (_i, [_i, _i, _i, _i, C.c_float] + [_vp] * 19),
    "cmf_preproc_logit_fwd": (_i, [_i, _i, _vp, _vp, C.c_float, C.c_float, _vp, _vp, _vp]),
    "cmf_coupling_x_to_z_fwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_coupling_x_to_z_bwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_sumsq_diff_fwd": (_i, [_i, _i, _vp, _vp, _vp, _vp]),
    "cmf_sumsq_diff_bwd": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp]),
    "cmf_flat_stack_usable": (_i, [_i, _i, _i, _i, _i, _i]),
    "cmf_flat_stack_fwd": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "cmf_flat_stack_bwd_partials": (_i, [_i]),
    "cmf_flat_stack_bwd": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_channel_sum_splits": (_i, [_i, _i]),
    "cmf_channel_sum": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp]),
    "cmf_convin_usable": (_i, [_i, _i, _i, _i]),
    "cmf_convin_wgrad_partials": (_i, [_i, _i, _i]),
    "cmf_convin_fwd": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "cmf_convin_dgrad": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "cmf_convin_wgrad": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "cmf_resnet_tail_usable": (_i, [_i, _i]),
    "cmf_resnet_tail_partial_floats": (_ll, [_i, _i, _i, _i, _i, _i]),
    "cmf_resnet_tail_fwd": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_resnet_tail_bwd": (_i, [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_mlp_usable": (_i, [_i, _vp, _i]),
    "cmf_mlp_pack_floats": (_ll, [_i]),
    "cmf_mlp_tiles": (_i, [_i, _i]),
    "cmf_mlp_save_floats": (_ll, [_i, _i, _i]),
    "cmf_mlp_partials": (_i, [_i, _i]),
    "cmf_mlp_partial_floats": (_ll, [_i, _i, _i]),
    "cmf_mlp_pack": (_i, [_i, _vp, _vp, _vp, _vp, _vp]),
    "cmf_mlp_jvp_fwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_mlp_jvp_bwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "cmf_conv3x3_c64_usable": (_i, [_i, _i, _i]),
    "cmf_conv3x3_c64_prep": (_i, [_vp, _i, _vp, _vp]),
    "cmf_conv3x3_c64_fwd": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "cmf_conv3x3_c64_fused": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "cmf_relu_mask_pack": (_i, [_i, _i, _vp, _vp, _vp]),
    "cmf_relu_and_pack": (_i, [_i, _i, _vp, _vp, _vp, _vp]),
    "cmf_conv3x3_c64_fused_bits": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "cmf_conv3x3_c64_joint": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "cmf_conv3x3_c64_wgrad_usable": (_i, [_i, _i, _i]),
    "cmf_conv3x3_c64_wgrad_workspace": (_sz, [_i, _i, _i]),
    "cmf_conv3x3_c64_wgrad": (_i, [_i, _i, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    "cmf_conv3x3_c64_wgrad_det": (_i, [_i, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _vp]),
}

_lib = None


class CmfB200Error(RuntimeError):
    pass


def load():
    """Load the shared library once; raise loudly if it is absent (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CmfB200Error(
            f"{LIB_PATH} not found: build it with `python -m cmf_b200.build` "
            "(nvcc, sm_100a). cmf_b200 has no CPU or eager fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here == header/library mismatch
        fn.restype, fn.argtypes = res, args
    if lib.cmf_version() != ABI_VERSION:
        raise CmfB200Error(f"libcmf_b200.so ABI {lib.cmf_version()} != expected {ABI_VERSION}")
    _lib = lib
    return lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = load().cmf_last_error_string().decode(errors="replace")
        kind = "bad argument" if code < 0 else f"CUDA error {code}"
        raise CmfB200Error(f"{what}: {kind}: {msg}")
