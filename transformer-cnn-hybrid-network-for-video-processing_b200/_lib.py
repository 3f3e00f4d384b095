"""ctypes binding of libhybridnet_sm100.so (include/hybridnet.h).

The product path has NO fallback: if the shared object is missing, or a call
returns a non-zero status, a RuntimeError is raised.  Nothing here imports
``oracle/``.
"""
import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_float, c_int, c_longlong,
                    c_ulonglong, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhybridnet_sm100.so")


class HnConv(Structure):
    _fields_ = [(n, c_int) for n in (
        "B", "H", "W", "C1", "C2", "ld1", "ld2", "OH", "OW", "Cout", "ldy",
        "KH", "KW", "stride", "pad", "dil", "mode", "ups", "act", "pshuf",
        "ldr", "Kpad")]


ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID = 0, 1, 2, 3

_P, _I, _L, _F = c_void_p, c_int, c_longlong, c_float
_FP = c_void_p  # float* passed as raw address

# name -> argtypes (restype is always int unless listed in _SPECIAL)
SIGNATURES = {
    "hn_version": [],
    "hn_check_device": [],
    "hn_nchw_f32_to_nhwc_bf16": [_FP, _P, _I, _I, _I, _I, _I, _I, _P],
    "hn_nhwc_bf16_to_nchw_f32": [_P, _FP, _I, _I, _I, _I, _I, _P],
    "hn_avgpool2_fwd": [_P, _P, _I, _I, _I, _I, _I, _I, _P],
    "hn_head_fwd": [_P, _FP, _I, _I, _I, _I, _I, _I, _P],
    "hn_head_bwd": [_FP, _FP, _P, _I, _I, _I, _I, _I, _I, _P],
    "hn_conv_kpad": [_I, _I, _I],
    "hn_conv_npad": [_I],
    "hn_pack_weight": [_FP, _P, _I, _I, _I, _I, _I, _L, _L, _L, _L, _L, _L, _P],
    "hn_pack_weight_s2dgrad": [_FP, _P, _I, _I, _I, _P],
    "hn_pack_weight_multi": [_P, _I, _L, _P],
    "hn_conv2d_fwd": [POINTER(HnConv), _P, _P, _P, _FP, _P, _P, _FP, _FP, _P],
    "hn_conv2d_wgrad": [POINTER(HnConv), _P, _P, _P, _FP, _I, _L, _L, _L, _L, _P],
    "hn_act_bwd": [_P, _P, _P, _L, _I, _I, _I, _I, _I, _FP, _P],
    "hn_act_fwd": [_P, _P, _L, _I, _I, _I, _I, _P],
    "hn_colsum": [_P, _L, _I, _I, _FP, _P],
    "hn_add": [_P, _P, _P, _P, _L, _I, _I, _I, _I, _I, _P],
    "hn_gelu3_sum_fwd": [_P, _P, _P, _P, _L, _P],
    "hn_gelu3_sum_bwd": [_P, _P, _P, _P, _P, _P, _P, _L, _P],
    "hn_wf_combine_fwd": [_P, _P, _P, _FP, _FP, _FP, _P, _L, _I, _I, _I, _F, c_ulonglong, c_ulonglong, c_ulonglong, _P, _P],
    "hn_wf_combine_bwd": [_P, _P, _P, _P, _FP, _FP, _FP, _P, _P, _P, _L, _I, _I, _I, _I, _F, c_ulonglong, c_ulonglong,
                          c_ulonglong, _P, _FP, _FP, _FP, _P],
    "hn_wf_out_fwd": [_P, _FP, _P, _P, _L, _I, _I, _I, _I, _F, c_ulonglong, _P, _P],
    "hn_wf_out_bwd": [_P, _P, _FP, _P, _L, _I, _I, _I, _I, _F, c_ulonglong, _P, _FP, _P],
    "hn_maxpool2_fwd": [_P, _P, _I, _I, _I, _I, _I, _I, _P],
    "hn_maxpool2_bwd": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P],
    "hn_upsample2_bwd": [_P, _P, _I, _I, _I, _I, _I, _I, _P],
    "hn_dropout": [_P, _P, _I, _L, _I, _I, _I, _F, c_ulonglong, _P, _I, _P],
    "hn_layernorm_fwd": [_P, _P, _FP, _FP, _L, _I, _I, _I, _F, _P],
    "hn_layernorm_bwd": [_P, _P, _P, _FP, _FP, _FP, _L, _I, _I, _I, _I, _F, _P],
    "hn_bn_stats": [_P, _L, _I, _I, _FP, _FP, _P],
    "hn_bn_finalize": [_FP, _FP, _L, _I, _F, _F, _FP, _FP, _FP, _FP, _P, _P],
    "hn_bn_apply": [_P, _P, _P, _FP, _FP, _FP, _FP, _L, _I, _I, _I, _I, _I, _P],
    "hn_bn_apply_train": [_P, _P, _P, _FP, _FP, _L, _F, _F, _FP, _FP, _FP, _FP, _P, _FP, _FP, _L, _I, _I, _I, _I, _I, _P],
    "hn_bn_bwd_reduce": [_P, _P, _P, _P, _FP, _FP, _L, _I, _I, _I, _I, _I, _I, _FP, _FP, _P],
    "hn_bn_bwd_apply": [_P, _P, _P, _FP, _FP, _FP, _FP, _FP, _L, _I, _I, _I, _I, _I, _P],
    "hn_bn_relu_bwd_reduce": [_P, _P, _FP, _FP, _FP, _FP, _L, _I, _I, _I, _FP, _FP, _P],
    "hn_bn_relu_bwd_apply": [_P, _P, _P, _FP, _FP, _FP, _FP, _FP, _FP, _L, _I, _I, _I, _I, _I, _P],
    "hn_dwconv_relu_ln_fwd": [_P, _P, _P, _P, _FP, _FP, _FP, _FP, _I, _I, _I, _I, _I, _I, _F, _P],
    "hn_dwconv_relu_ln_bwd_pre": [_P, _P, _P, _P, _P, _P, _P, _FP, _FP, _FP, _FP, _FP, _FP,
                                  _I, _I, _I, _I, _I, _I, _F, _P],
    "hn_dwconv_bwd_data_weight": [_P, _P, _P, _P, _P, _FP, _FP, _I, _I, _I, _I, _I, _I, _I, _P],
    "hn_attn_fwd": [_P, _P, _P, _P, _FP, _I, _I, _I, _I, _I, _I, _I, _I, _F, _P],
    "hn_attn_bwd_workspace": [_I, _I, _I, _I],
    "hn_attn_tune": [_I],
    "hn_attn_bwd_tc_workspace": [_I, _I, _I, _I],
    "hn_attn_bwd_tc_supported": [_I, _I, _I, _I, _I, _I, _I, _I, _I, _I],
    "hn_attn_bwd_tc": [_P, _P, _P, _P, _P, _FP, _FP, _L, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _F, _I, _P],
    "hn_attn_bwd_tc_static": [_P, _P, _P, _P, _P, _FP, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _F, _I, _P],
    "hn_attn_fwd_tc_supported": [_I, _I, _I, _I, _I, _I],
    "hn_attn_fwd_tc": [_P, _P, _P, _P, _FP, _I, _I, _I, _I, _I, _I, _I, _I, _F, _I, _P],
    "hn_attn_tune_fwd": [_I],
    "hn_attn_bwd": [_P, _P, _P, _P, _P, _FP, _FP, _L, _P, _P, _P, _I, _I, _I, _I,
                    _I, _I, _I, _I, _I, _I, _I, _I, _F, _P],
    "hn_conv_tc_supported": [POINTER(HnConv)],
    "hn_conv_tc_halo_supported": [POINTER(HnConv)],
    "hn_conv_tc_fwd": [POINTER(HnConv), _P, _P, _P, _FP, _P, _FP, _FP, _P],
    "hn_conv_sc_supported": [POINTER(HnConv)],
    "hn_conv_sc_fwd": [POINTER(HnConv), _P, _P, _P, _FP, _P, _FP, _FP, _P],
    "hn_conv_sc_fwd_res": [POINTER(HnConv), _P, _P, _P, _FP, _P, _P, _FP, _FP, _P],
    "hn_conv_wgrad_sc_supported": [POINTER(HnConv)],
    "hn_conv_wgrad_sc": [POINTER(HnConv), _P, _P, _P, _FP, _I, _L, _L, _L, _L, _P],
    "hn_conv_wgrad_tc_supported": [POINTER(HnConv)],
    "hn_conv_wgrad_tc": [POINTER(HnConv), _P, _P, _P, _FP, _I, _L, _L, _L, _L, _FP, _L, _P],
    "hn_conv_wgrad_tc_workspace": [POINTER(HnConv), _L],
    "hn_adamw_step": [_FP, _FP, _FP, _FP, _L, _F, _F, _F, _F, _F, _I, _F, _P],
    "hn_adamw_step_dev": [_FP, _FP, _FP, _FP, _L, _F, _F, _F, _F, _F, _P, _F, _P],
    "hn_dice_fwd": [_FP, _FP, _I, _I, _I, _I, _F, _FP, _FP, _P],
    "hn_dice_bwd": [_FP, _FP, _FP, _FP, _I, _I, _I, _I, _F, _P],
    "hn_mse_fwd": [_FP, _FP, _L, _FP, _FP, _P],
    "hn_mse_bwd": [_FP, _FP, _L, _FP, _FP, _P],
    "hn_jaccard": [_FP, _FP, _I, _I, _I, _I, _FP, _FP, _P],
    "hn_augment_pack": [_P, _I, _P, _FP, _I, _I, _I, _I, _I, _F, c_ulonglong, _P, _I, POINTER(c_int), POINTER(c_int), _I, _P],
    "hn_avgpool2_bwd": [_P, _P, _I, _I, _I, _I, _I, _I, _P],
}

_LONG_RESULT = {"hn_conv_wgrad_tc_workspace", "hn_attn_bwd_workspace", "hn_attn_bwd_tc_workspace"}
_lib = None


def load():
    """Load the shared object once; raise loudly if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"hybridnet: {LIB_PATH} is missing -- build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` (or csrc/build.sh); "
            "there is no CPU/PyTorch fallback for the hot path")
    lib = ctypes.CDLL(LIB_PATH)
    lib.hn_last_error.restype = c_char_p
    lib.hn_last_error.argtypes = []
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)        # AttributeError => header/library mismatch
        fn.argtypes = args
        fn.restype = c_longlong if name in _LONG_RESULT else c_int
    _lib = lib
    return lib


# kernels launched per entry point (for the launch count bench.py reports)
LAUNCHES = {"hn_attn_bwd": 3, "hn_attn_bwd_tc": 3, "hn_attn_bwd_tc_static": 3}
stats = {"calls": 0, "launches": 0}
_profiler = None      # optional callable(name, args) -> context manager (bench.py roofline pass)


def set_profiler(p):
    global _profiler
    _profiler = p


def call(name, *args):
    lib = load()
    stats["calls"] += 1
    stats["launches"] += LAUNCHES.get(name, 1)
    if _profiler is not None:
        with _profiler(name, args):
            rc = getattr(lib, name)(*args)
    else:
        rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.hn_last_error().decode()}")


def exported_symbols():
    """Symbols declared in include/hybridnet.h (used by the CPU-side load test)."""
    return ["hn_last_error"] + list(SIGNATURES)
