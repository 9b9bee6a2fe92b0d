"""ctypes binding of libita_b200.so (C-ABI declared in include/ita_b200.h).

The shared object is built in-tree by ``__graft_entry__.build()`` (``make -C drone-oa-iree-vit-accelerator_b200/csrc``).
There is no CPU fallback: if the library is missing, or no CUDA device is present when a handle is created, the
call raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ITA_B200_LIB") or os.path.join(_HERE, "libita_b200.so")   # override: developer builds (tracing)

ABI_VERSION = 1
# enums of include/ita_b200.h
ROUND_HALF_UP, ROUND_FLOOR, ROUND_TORCH_RNE_F32 = 0, 1, 2
SOFTMAX_REPO_GLOBAL, SOFTMAX_STREAM16 = 0, 1
HEADS_PYITA, HEADS_TORCH = 0, 1
ST_Q, ST_K, ST_V, ST_QK, ST_AV, ST_O, ST_SUM, ST_FC1, ST_RELU, ST_FC2 = range(10)
NUM_STAGES = 10
STEP1_QP, STEP2_KP, STEP3_VP, STEP4_QK, STEP4_SOFTMAX, STEP5_AV, STEP6_O, STEP7_OSUM = 1, 2, 3, 4, 40, 5, 6, 7
PATH_AUTO, PATH_FUSED, PATH_STAGED = 0, 1, 2
LAYER_STATIC_SCRATCH_BYTES = 16 * 1024 * 1024   # ITA_LAYER_STATIC_SCRATCH_BYTES

_i8p, _u8p, _i32p, _f32p = C.POINTER(C.c_int8), C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float)


class MhaDesc(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in ("abi_version", "S", "E", "P", "H", "F", "round_mode", "softmax_mode",
                                          "head_mode", "device")] +
                [(n, _i8p) for n in ("Wq", "Wk", "Wv", "Wo")] + [(n, _i32p) for n in ("Bq", "Bk", "Bv", "Bo")] +
                [(n, _i8p) for n in ("Wff", "Wff2")] + [(n, _i32p) for n in ("Bff", "Bff2")] +
                [(n, _i32p) for n in ("mult", "shift", "add")] + [("eff", _f32p)])


class MhaStages(C.Structure):
    _fields_ = [("Qp", C.c_void_p), ("Kp", C.c_void_p), ("Vp", C.c_void_p), ("A", C.c_void_p), ("Asm", C.c_void_p),
                ("O", C.c_void_p), ("Out", C.c_void_p)]


class LayerParams(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("s_attn_in", "s_attn_out", "s_ffn_in", "s_ffn_out", "ln_eps")] + \
               [(n, C.c_void_p) for n in ("ln1_w", "ln1_b", "ln2_w", "ln2_b")]


# every symbol include/ita_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "ita_b200_abi_version": (C.c_int, []),
    "ita_last_error": (C.c_char_p, []),
    "ita_kernel_launch_count": (C.c_int64, []),
    "ita_mha_create": (C.c_int, [C.POINTER(MhaDesc), C.POINTER(C.c_void_p)]),
    "ita_mha_destroy": (None, [C.c_void_p]),
    "ita_mha_fused_supported": (C.c_int, [C.c_void_p]),
    "ita_mha_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "ita_mha_forward_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "ita_mha_forward_host_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "ita_mha_forward_host_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_int32, C.c_void_p]),
    "ita_mha_describe": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "ita_mha_forward_stages": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(MhaStages), C.c_void_p]),
    "ita_mha_forward_fused_debug": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(MhaStages), C.c_void_p]),
    "ita_mha_sync_check": (C.c_int, [C.c_void_p, C.c_void_p]),
    "ita_mha_step": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "ita_ffn_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "ita_ffn_forward_path": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "ita_ffn_fused_supported": (C.c_int, [C.c_void_p]),
    "ita_residual_layernorm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p,
                                         C.c_void_p, C.c_float, C.c_int64, C.c_int32, C.c_void_p]),
    "ita_encoder_layer_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "ita_encoder_layer_forward_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "ita_tokenizer_forward": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                                        C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]),
    "ita_fusion_conv_forward": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ita_lstm_cell_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "ita_quantize_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]),
    "ita_dequantize_s8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]),
    "ita_selftest": (C.c_int, [C.c_int32, C.c_char_p, C.c_int64]),
    "ita_microbench": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_double), C.c_int32]),
}

_lib = None


class ITAError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load libita_b200.so (once) and declare every prototype.  Raises if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ITAError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)          # AttributeError if the library does not export it
            fn.restype, fn.argtypes = res, args
        if lib.ita_b200_abi_version() != ABI_VERSION:
            raise ITAError(f"libita_b200.so ABI {lib.ita_b200_abi_version()} != binding ABI {ABI_VERSION}")
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().ita_last_error().decode("utf-8", "replace")
        raise ITAError(f"{what or 'libita_b200'} failed (code {rc}): {msg}")
