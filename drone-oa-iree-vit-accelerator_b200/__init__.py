"""B200-native ITA int8 multi-head attention (one hot path of OpenHardware-Initiative/Drone-OA-IREE-ViT-Accelerator).

    import importlib; ita = importlib.import_module("drone-oa-iree-vit-accelerator_b200")     # or: import ita_b200 as ita

* ``ITASelfAttention``   drop-in for the converted ``ITASelfAttention_QAT`` module        (attention.py)
* ``Transformer``        drop-in for the golden-model object of the reference harness      (golden.py)
* ``ITAEngine``          thin owner of a C-ABI handle (device tensors in / out)            (engine.py)
* ``params``             scale -> (mult, shift, add) derivation, reference names           (params.py)

All arithmetic lives in ``libita_b200.so`` (``csrc/``, hand-written sm_100a CUDA; C-ABI in ``include/ita_b200.h``).
Importing this package does not need a GPU; creating an engine does, and raises otherwise (no CPU fallback).
"""
from . import _lib
from ._lib import (HEADS_PYITA, HEADS_TORCH, PATH_AUTO, PATH_FUSED, PATH_STAGED, ROUND_FLOOR, ROUND_HALF_UP,
                   ROUND_TORCH_RNE_F32, SOFTMAX_REPO_GLOBAL, SOFTMAX_STREAM16, ITAError)
from .params import (ITAQuantParams, calculate_hw_params, dequantize_bias_to_int32, linear_eff_f32,
                     params_from_captured_block, translate_torch_scales_to_hw_params)
from .engine import ITAEngine, dequantize, microbench, quantize, residual_layernorm, selftest
from .attention import ITASelfAttention
from .golden import Transformer
from .model import ITAViTLSTM, layer_quant_params
from .sharding import all_shards, gather_frames, max_over_ranks, shard_range

__all__ = ["ITASelfAttention", "ITAViTLSTM", "Transformer", "ITAEngine", "ITAQuantParams", "ITAError", "calculate_hw_params",
           "dequantize_bias_to_int32", "translate_torch_scales_to_hw_params", "quantize", "dequantize", "selftest",
           "microbench"]
