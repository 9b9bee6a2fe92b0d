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


def linear_eff_f32(s_in: float, s_w: float, s_out: float) -> np.float32:
    """Requantisation scale of a converted ``quantized.Linear`` exactly as the reference's backend (qnnpack) forms it:
    ``(f32(s_in) * f32(s_w)) / f32(s_out)`` in float32 arithmetic.  (Computing it in double and casting differs by one
    ulp now and then, which flips near-tie outputs: 1 of 16384 in the fc2 fixture.)"""
    return np.float32(np.float32(np.float32(s_in) * np.float32(s_w)) / np.float32(s_out))


def bias_to_int32(bias, s_acc: float) -> np.ndarray:
    """``round(bias / s_acc)`` exactly as the reference evaluates it: a float32 bias array divided by a Python float stays
    float32 in NumPy (the scalar is cast to float32), then round-half-even (``W_B.py:233-245``)."""
    return np.round(np.asarray(bias, np.float32) / np.float32(s_acc)).astype(np.int32)


def dequantize_bias_to_int32(layer_params: dict) -> np.ndarray:
    """``B_int32 = round(B_float / (S_input * S_weight))`` on the reference's captured-layer dictionary."""
    bias = layer_params["bias"]
    if bias is None:
        return np.zeros(layer_params["weight"]["q_value"].shape[0], dtype=np.int32)
    s_acc = layer_params["input"][0]["scale"] * layer_params["weight"]["scale"]
    return bias_to_int32(bias, s_acc)


def translate_torch_scales_to_hw_params(attn_block: dict, ffn_block: dict, H: int) -> Dict[str, dict]:
    """Per-stage ``{'mult': [..H], 'shift': [..H], 'add': [..H]}`` from the reference's organised capture dicts."""
    hw = {}
    for key, layer in {**attn_block, **ffn_block}.items():
        if layer.get("type") == "Linear":
            eff = (layer["input"][0]["scale"] * layer["weight"]["scale"]) / layer["output_scale"]
            m, s = calculate_hw_params(eff)
            n = 1 if key in ("fc1", "fc2") else H
            e32 = float(linear_eff_f32(layer["input"][0]["scale"], layer["weight"]["scale"], layer["output_scale"]))
            hw[key] = {"mult": [m] * n, "shift": [s] * n, "add": [layer["output_zp"]] * n, "eff": [e32] * n}
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


def params_from_captured_block(q, prefix: str = "b0_", round_mode: int = L.ROUND_TORCH_RNE_F32,
                               softmax_mode: int = L.SOFTMAX_REPO_GLOBAL, with_ffn: bool = False) -> ITAQuantParams:
    """Parameter set of one converted reference block from its flattened capture (the keys the reference's
    ``PyTorchDataExtractor`` / ``organize_params_by_block`` produce, as stored in ``tests/golden/qat_*.npz``:
    ``<prefix>{q,k,v,out}_proj_{w,sw,bias,sin,sout}``, ``qk_sout``, ``softmax_sout``, ``av_sout`` [, ``fc1_*``, ``fc2_*``]).

    Follows the harness' recipe (``tests/export_and_validation_W_B.py:479-491,247-305``): ``W = weight.T[None]``,
    ``B = round(bias / (s_in * s_w))``, ``eff = s_in1 * s_in2 / s_out`` -> ``calculate_hw_params``; the fp32 requant scales
    are formed the way the backend forms them (``linear_eff_f32`` for Linear, double product cast to f32 for the matmuls).
    H = 1 for every shipped model; a capture that carries an ``H`` entry > 1 (``qat_block_H4_S256.npz``, the module
    ``ITASelfAttention_QAT(128, 256, 4)``) is split the way the module splits it (``models/ITA/QAT/layers.py:110-123``): heads
    are column slices of one projection, one out_proj over the concatenated contexts, every head shares the stage's scale."""
    g = lambda k: q[prefix + k]
    s = lambda k: float(q[prefix + k])
    P, E = g("q_proj_w").shape
    S = int(g("x").shape[-2]) if (prefix + "x") in q else 128
    mult = np.zeros((L.NUM_STAGES, 1), np.int32)
    shift = np.ones((L.NUM_STAGES, 1), np.int32)
    add = np.zeros((L.NUM_STAGES, 1), np.int32)
    eff = np.zeros((L.NUM_STAGES, 1), np.float32)
    lin = {L.ST_Q: "q_proj", L.ST_K: "k_proj", L.ST_V: "v_proj", L.ST_O: "out_proj"}
    if with_ffn:
        lin.update({L.ST_FC1: "fc1", L.ST_FC2: "fc2"})
    for st, n in lin.items():
        mult[st, 0], shift[st, 0] = calculate_hw_params(s(n + "_sin") * s(n + "_sw") / s(n + "_sout"))
        eff[st, 0] = linear_eff_f32(s(n + "_sin"), s(n + "_sw"), s(n + "_sout"))
    for st, e in ((L.ST_QK, s("q_proj_sout") * s("k_proj_sout") / s("qk_sout")),
                  (L.ST_AV, s("softmax_sout") * s("v_proj_sout") / s("av_sout"))):
        mult[st, 0], shift[st, 0] = calculate_hw_params(e)
        eff[st, 0] = np.float32(e)
    T = lambda k: np.ascontiguousarray(np.asarray(g(k + "_w")).T[None]).astype(np.int8)
    Bi = lambda k: bias_to_int32(g(k + "_bias"), s(k + "_sin") * s(k + "_sw"))[None]
    kw = {}
    if with_ffn:
        F = int(g("fc1_w").shape[0])
        # ReLU keeps the quantisation parameters (s_in == s_out): the golden model's relu requant is the identity
        mult[L.ST_RELU, 0] = 0
        kw = dict(F=F, Wff=T("fc1"), Wff2=T("fc2"), Bff=Bi("fc1"), Bff2=Bi("fc2"))
    H = int(q["H"]) if "H" in q else 1
    if H > 1:
        if with_ffn or P % H:
            raise ValueError("multi-head captures carry no FFN and need proj_dim divisible by num_heads")
        hd = int(P) // H
        split = lambda k: np.ascontiguousarray(np.asarray(g(k + "_w")).reshape(H, hd, E).transpose(0, 2, 1)).astype(np.int8)
        Wo = np.ascontiguousarray(np.asarray(g("out_proj_w")).T.reshape(H, hd, E)).astype(np.int8)
        Bo = np.zeros((H, int(E)), np.int32)
        Bo[0] = Bi("out_proj")[0]
        rep = lambda a: np.ascontiguousarray(np.repeat(a, H, axis=1))
        return ITAQuantParams(S=S, E=int(E), P=hd, H=H, Wq=split("q_proj"), Wk=split("k_proj"), Wv=split("v_proj"), Wo=Wo,
                              Bq=Bi("q_proj").reshape(H, hd), Bk=Bi("k_proj").reshape(H, hd), Bv=Bi("v_proj").reshape(H, hd), Bo=Bo,
                              mult=rep(mult), shift=rep(shift), add=rep(add), eff=rep(eff), round_mode=round_mode,
                              softmax_mode=softmax_mode, head_mode=L.HEADS_TORCH)
    return ITAQuantParams(S=S, E=int(E), P=int(P), H=1, Wq=T("q_proj"), Wk=T("k_proj"), Wv=T("v_proj"), Wo=T("out_proj"),
                          Bq=Bi("q_proj"), Bk=Bi("k_proj"), Bv=Bi("v_proj"), Bo=Bi("out_proj"), mult=mult, shift=shift, add=add,
                          eff=eff, round_mode=round_mode, softmax_mode=softmax_mode, head_mode=L.HEADS_PYITA, **kw)
