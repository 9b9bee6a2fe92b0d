"""Host-side parameter derivation: float scales of a converted torch QAT block -> the golden model's per-stage
(mult, shift, add) and int32 biases.  Mirrors the reference harness functions by name and behaviour
(``tests/export_and_validation_W_B.py``):

* ``calculate_hw_params``                    :225-230
* ``dequantize_bias_to_int32``               :233-245
* ``translate_torch_scales_to_hw_params``    :247-305

plus ``ITAQuantParams``, the container the engine consumes (arrays indexed ``[stage, head]``).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, Optional

import numpy as np

from . import _lib as L

STAGE_KEYS = ["q_proj", "k_proj", "v_proj", "qk_matmul", "av_matmul", "out_proj", "head_sum", "fc1", "relu", "fc2"]


def calculate_hw_params(effective_scale: float):
    """float scale -> (mult, shift) with ``scale ~= mult / 2**shift`` and mult in [128, 255].

    Same arithmetic as the reference (``np.frexp``; ``round(mantissa * 2**8)``; ``shift = 8 - exponent``).  The
    reference can return mult == 256, which does not fit the accelerator's uint8 multiplier; (256, s) is
    renormalised to the identical rational (128, s - 1).
    """
    if effective_scale == 0:
        return 0, 0
    mantissa, exponent = np.frexp(effective_scale)
    mult = int(round(float(mantissa) * 2 ** 8))
    shift = int(8 - exponent)
    if mult == 256:
        mult, shift = 128, shift - 1
    return mult, shift


def dequantize_bias_to_int32(layer_params: dict) -> np.ndarray:
    """``B_int32 = round(B_float / (S_input * S_weight))`` on the reference's captured-layer dictionary."""
    bias = layer_params["bias"]
    if bias is None:
        return np.zeros(layer_params["weight"]["q_value"].shape[0], dtype=np.int32)
    s_acc = layer_params["input"][0]["scale"] * layer_params["weight"]["scale"]
    return np.round(np.asarray(bias) / s_acc).astype(np.int32)


def translate_torch_scales_to_hw_params(attn_block: dict, ffn_block: dict, H: int) -> Dict[str, dict]:
    """Per-stage ``{'mult': [..H], 'shift': [..H], 'add': [..H]}`` from the reference's organised capture dicts."""
    hw = {}
    for key, layer in {**attn_block, **ffn_block}.items():
        if layer.get("type") == "Linear":
            eff = (layer["input"][0]["scale"] * layer["weight"]["scale"]) / layer["output_scale"]
            m, s = calculate_hw_params(eff)
            n = 1 if key in ("fc1", "fc2") else H
            hw[key] = {"mult": [m] * n, "shift": [s] * n, "add": [layer["output_zp"]] * n, "eff": [eff] * n}
        elif layer.get("type") in ("MatMul_QK", "MatMul_AV"):
            so, zo = layer["output"]["scale"], layer["output"]["zp"]
            if key == "qk_matmul":
                s1, s2 = attn_block["q_proj"]["output"]["scale"], attn_block["k_proj"]["output"]["scale"]
            else:
                s1, s2 = attn_block["softmax"]["output"]["scale"], attn_block["v_proj"]["output"]["scale"]
            eff = (s1 * s2) / so
            m, s = calculate_hw_params(eff)
            hw[key] = {"mult": [m] * H, "shift": [s] * H, "add": [zo] * H, "eff": [eff] * H}
    if "relu" in ffn_block and "fc1" in ffn_block and "fc2" in ffn_block:
        s_in = ffn_block["fc1"]["output"]["scale"]
        s_out = ffn_block["fc2"]["input"][0]["scale"]
        eff = s_in / s_out if s_out > 0 else 0
        m, s = calculate_hw_params(eff)
        hw["relu"] = {"mult": m, "shift": s, "add": ffn_block["relu"]["output"]["zp"], "q_1": 0, "q_b": 0, "q_c": 0,
                      "eff": eff}
    return hw


@dataclasses.dataclass
class ITAQuantParams:
    """Everything ``ita_mha_create`` needs, as NumPy arrays in the golden model's layouts."""
    S: int
    E: int
    P: int
    H: int
    Wq: np.ndarray
    Wk: np.ndarray
    Wv: np.ndarray
    Wo: np.ndarray
    Bq: np.ndarray
    Bk: np.ndarray
    Bv: np.ndarray
    Bo: np.ndarray
    mult: np.ndarray
    shift: np.ndarray
    add: np.ndarray
    eff: Optional[np.ndarray] = None
    F: int = 0
    Wff: Optional[np.ndarray] = None
    Wff2: Optional[np.ndarray] = None
    Bff: Optional[np.ndarray] = None
    Bff2: Optional[np.ndarray] = None
    round_mode: int = L.ROUND_HALF_UP
    softmax_mode: int = L.SOFTMAX_REPO_GLOBAL
    head_mode: int = L.HEADS_PYITA

    @staticmethod
    def from_quant_dict(S, E, P, H, tensors: dict, quant_params: dict, F: int = 0, **modes) -> "ITAQuantParams":
        """Build from the golden-model constructor arguments (``Wq .. Bo`` [, ``Wff .. Bff2``] and the
        ``quant_params`` dictionary of ``translate_torch_scales_to_hw_params``)."""
        mult = np.zeros((L.NUM_STAGES, H), np.int32)
        shift = np.ones((L.NUM_STAGES, H), np.int32)
        add = np.zeros((L.NUM_STAGES, H), np.int32)
        eff = np.zeros((L.NUM_STAGES, H), np.float32)
        for st, key in enumerate(STAGE_KEYS):
            q = quant_params.get(key)
            if q is None:
                continue

            def vec(v):
                v = np.atleast_1d(np.asarray(v))
                return np.broadcast_to(v, (H,)) if v.size in (1, H) else v[:H]

            mult[st], shift[st], add[st] = vec(q["mult"]), vec(q["shift"]), vec(q["add"])
            if "eff" in q:
                eff[st] = vec(q["eff"]).astype(np.float32)
            else:
                eff[st] = (mult[st].astype(np.float64) / np.exp2(shift[st].astype(np.float64))).astype(np.float32)
        g = tensors.get
        return ITAQuantParams(S=S, E=E, P=P, H=H, Wq=g("Wq"), Wk=g("Wk"), Wv=g("Wv"), Wo=g("Wo"), Bq=g("Bq"),
                              Bk=g("Bk"), Bv=g("Bv"), Bo=g("Bo"), mult=mult, shift=shift, add=add, eff=eff, F=F,
                              Wff=g("Wff"), Wff2=g("Wff2"), Bff=g("Bff"), Bff2=g("Bff2"), **modes)
