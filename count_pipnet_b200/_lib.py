"""ctypes binding of libcpnb200.so (include/cpn_b200.h). There is no CPU or PyTorch fallback behind this module:
if the library is missing or the device is not sm_100, the hot path raises."""
from __future__ import annotations

import ctypes as C
import os
import threading

_LIB = None
_LOCK = threading.Lock()
LIB_PATH = os.environ.get("CPN_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libcpnb200.so")  # env: A/B builds (dev)

_vp, _i, _f, _ll = C.c_void_p, C.c_int, C.c_float, C.c_longlong

# name -> argument types (all return int unless noted)
SIGNATURES = {
    "cpn_version": [],
    "cpn_check_device": [],
    "cpn_plan_create": [C.POINTER(_vp)],
    "cpn_plan_run": [_vp, _vp],
    "cpn_plan_capture": [_vp, _vp],
    "cpn_plan_has_graph": [_vp],
    "cpn_plan_num_ops": [_vp],
    "cpn_plan_run_timed": [_vp, _vp, C.POINTER(C.c_float), _i],
    "cpn_plan_op_info": [_vp, _i, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_double)],
    "cpn_plan_destroy": [_vp],
    "cpn_stem_ln": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp],
    "cpn_dwconv7_ln": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _vp],
    "cpn_dwconv7_stats": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_dwconv7_stats_split": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_gemm_ln_gelu_split": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp, _vp, _f, _vp, _vp, _vp],
    "cpn_mlp_fused": [_vp, _vp, _vp, _vp, _i, _i, _vp, _i, _vp, _vp, _vp, _vp, _f, _vp, _vp],
    "cpn_gemm_ln_gelu": [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp, _vp, _f, _vp, _vp],
    "cpn_ln_im2col2": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp],
    "cpn_gemm_bf16": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
    "cpn_addon_conv1x1": [_vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _vp],
    "cpn_addon_softmax_pool": [_vp, _vp, _vp, _vp, _vp, _f, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "cpn_softmax_pool": [_vp, _vp, _vp, _f, _i, _i, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_softmax_pool_arg": [_vp, _vp, _vp, _f, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_pool_argmax_reduce": [_vp, _vp, _vp, _i, _vp, _i, _i, _vp],
    "cpn_softmax_pool_static": [_vp, _i, _vp, _i, _i, _i, _vp],  # no plan, no allocation (static device scratch)
    "cpn_topk_update": [_vp, _vp, _vp, _i, _i, _i, _ll, _vp, _vp, _vp, _vp, _vp],
    "cpn_eval_batch": [_vp, _vp, _vp, _vp, _vp, _f, _f, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "cpn_head_tail": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_nhwc_to_nchw": [_vp, _vp, _vp, _i, _i, _i, _vp],
    "cpn_nchw_to_nhwc": [_vp, _vp, _vp, _i, _i, _i, _vp],
    "cpn_conv_bn_act": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp],
    "cpn_conv_bn_act_resb": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "cpn_im2col_bf16": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "cpn_im2col_nchw_f32": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp],
    "cpn_maxpool3s2": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_subsample2_bf16": [_vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_maxpool3s2_bf16": [_vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "cpn_preprocess_u8": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, C.POINTER(C.c_float), C.POINTER(C.c_float), _vp],
    # training side of the head (no plan argument)
    "cpn_pipnet_loss": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _f, _f, _f, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp],
    "cpn_scale_unless_one": [_vp, _ll, _vp, _vp],
    "cpn_softmax_pool_nchw": [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp],
    "cpn_softmax_pool_nchw_bwd": [_vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _vp],
}

ACT_NONE, ACT_GELU, ACT_RELU = 0, 1, 2
POOL_MAX, POOL_SUM = 0, 1
ENC_NONE, ENC_ONEHOT, ENC_LINEAR, ENC_LINEAR_FULL, ENC_BILINEAR, ENC_IDENTITY = 0, 1, 2, 3, 4, 5


# entry points that do not return an int status: name -> (argtypes, restype)
OTHER_SIGNATURES = {
    "cpn_last_error": ([], C.c_char_p),
    "cpn_plan_op_detail": ([_vp, _i], C.c_char_p),
    "cpn_conv3x3_tile_rows": ([_i, _i], C.c_int),  # a count, not a status
    "cpn_preprocess_ws_ints": ([_i, _i, _i, _i], C.c_longlong),  # a size, not a status
    "cpn_pipnet_loss_ws_floats": ([_i, _i, _i], C.c_longlong),
    "cpn_softmax_pool_nchw_ws_floats": ([_i, _i, _i], C.c_longlong),
}


class CpnError(RuntimeError):
    pass


def load():
    """Load the shared library once; raises if it has not been built (no fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        if not os.path.exists(LIB_PATH):
            raise CpnError(
                f"{LIB_PATH} not found: build it with `python -m count_pipnet_b200.build` "
                "(nvcc, sm_100a). The B200 forward path has no CPU/PyTorch fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = C.c_int
        for name, (args, res) in OTHER_SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = res
        _LIB = lib
    return _LIB


def check(rc: int) -> None:
    if rc != 0:
        msg = load().cpn_last_error()
        raise CpnError(msg.decode() if msg else f"libcpnb200 call failed with code {rc}")


def ptr(t):
    """Device pointer of a torch tensor (or None -> NULL)."""
    return None if t is None else t.data_ptr()


def call(name: str, *args) -> None:
    check(getattr(load(), name)(*args))
