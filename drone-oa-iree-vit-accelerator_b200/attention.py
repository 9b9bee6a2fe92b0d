"""``ITASelfAttention`` -- the B200 drop-in for the reference operator ``ITASelfAttention_QAT`` *after conversion*
(reference ``models/ITA/QAT/layers.py:77-127``).

Same constructor (``embed_dim, proj_dim, num_heads``), same ``forward(x: f32[B,N,E]) -> f32[B,N,E]``, and the same
load-bearing submodule names (``q_proj, k_proj, v_proj, out_proj, custom_softmax, matmul1, matmul2, quant, dequant``) that
the reference's hooks and state-dict loading address (``tests/export_and_validation_W_B.py:86,167-179,409-411``).  The
submodules here are parameter holders: the arithmetic of the whole block is one call into libita_b200.so.

``from_converted(block)`` reads exactly the fields a converted block carries:
``{q,k,v,out}_proj.weight()/bias()/scale/zero_point``, ``matmul1/matmul2.scale``, ``quant.scale``.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L
from .engine import ITAEngine, dequantize, quantize
from .params import ITAQuantParams, bias_to_int32, calculate_hw_params, linear_eff_f32


class QuantizedLinearParams(nn.Module):
    """int8 weight [out,in] (torch layout), float bias, weight scale, output scale / zero point."""

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.register_buffer("weight_int8", torch.zeros(out_features, in_features, dtype=torch.int8))
        self.register_buffer("bias", torch.zeros(out_features, dtype=torch.float32))
        self.weight_scale = 1.0
        self.scale = 1.0          # output scale (same attribute name as torch.ao.nn.quantized.Linear)
        self.zero_point = 0


class ScaleHolder(nn.Module):
    def __init__(self, scale: float = 1.0, zero_point: int = 0):
        super().__init__()
        self.scale, self.zero_point = scale, zero_point


class ITASelfAttention(nn.Module):
    def __init__(self, embed_dim: int, proj_dim: int, num_heads: int, round_mode: int = L.ROUND_TORCH_RNE_F32,
                 softmax_mode: int = L.SOFTMAX_REPO_GLOBAL):
        super().__init__()
        if proj_dim % num_heads:
            raise ValueError("proj_dim must be divisible by num_heads")
        self.embed_dim, self.proj_dim = embed_dim, proj_dim
        self.num_heads = num_heads
        self.head_dim = proj_dim // num_heads
        self.q_proj = QuantizedLinearParams(embed_dim, proj_dim)
        self.k_proj = QuantizedLinearParams(embed_dim, proj_dim)
        self.v_proj = QuantizedLinearParams(embed_dim, proj_dim)
        self.out_proj = QuantizedLinearParams(proj_dim, embed_dim)
        self.custom_softmax = ScaleHolder(1.0 / 255.0, 0)       # u8 output, scale 1/255 (ITA_softmax.py:66-67)
        self.matmul1 = ScaleHolder()                              # logits scale
        self.matmul2 = ScaleHolder()                              # context scale
        self.quant = ScaleHolder()                                # input scale
        self.dequant = ScaleHolder()
        self.round_mode, self.softmax_mode = round_mode, softmax_mode
        self._engines: Dict[Tuple[int, int], ITAEngine] = {}

    # ------------------------------------------------------------------------------------------------ construction
    @classmethod
    def from_converted(cls, block, round_mode: int = L.ROUND_TORCH_RNE_F32,
                       softmax_mode: int = L.SOFTMAX_REPO_GLOBAL) -> "ITASelfAttention":
        """Build from a converted reference block (``torch.ao.quantization.convert`` output)."""
        qw = block.q_proj.weight()
        m = cls(qw.shape[1], qw.shape[0], int(block.num_heads), round_mode, softmax_mode)
        for name in ("q_proj", "k_proj", "v_proj", "out_proj"):
            src, dst = getattr(block, name), getattr(m, name)
            w = src.weight()
            dst.weight_int8.copy_(w.int_repr())
            dst.weight_scale = float(w.q_scale())
            b = src.bias()
            if b is not None:
                dst.bias.copy_(b.detach())
            dst.scale, dst.zero_point = float(src.scale), int(src.zero_point)
        m.matmul1.scale, m.matmul1.zero_point = float(block.matmul1.scale), int(block.matmul1.zero_point)
        m.matmul2.scale, m.matmul2.zero_point = float(block.matmul2.scale), int(block.matmul2.zero_point)
        m.quant.scale, m.quant.zero_point = float(block.quant.scale), int(block.quant.zero_point)
        return m

    def quant_params(self, seq_len: int) -> ITAQuantParams:
        """Scales -> golden-model parameter set, exactly as the reference harness derives it
        (``tests/export_and_validation_W_B.py:247-305,479-491``)."""
        H, hd, E = self.num_heads, self.head_dim, self.embed_dim
        s_x = float(self.quant.scale)
        s_q, s_k, s_v = float(self.q_proj.scale), float(self.k_proj.scale), float(self.v_proj.scale)
        s_a, s_sm, s_o, s_out = float(self.matmul1.scale), float(self.custom_softmax.scale), float(self.matmul2.scale), \
            float(self.out_proj.scale)
        effs = {
            L.ST_Q: s_x * self.q_proj.weight_scale / s_q, L.ST_K: s_x * self.k_proj.weight_scale / s_k,
            L.ST_V: s_x * self.v_proj.weight_scale / s_v, L.ST_QK: s_q * s_k / s_a, L.ST_AV: s_sm * s_v / s_o,
            L.ST_O: s_o * self.out_proj.weight_scale / s_out,
        }
        # The fp32 requant scale as the reference backend forms it: quantized Linear (qnnpack) computes
        # (f32(s_in) * f32(s_w)) / f32(s_out) in float32; the patched matmuls multiply by the double product cast to f32
        # (tests/export_and_validation_W_B.py:139-142).  The two differ by an ulp now and then, which flips near-ties.
        eff32 = {
            L.ST_Q: linear_eff_f32(s_x, self.q_proj.weight_scale, s_q), L.ST_K: linear_eff_f32(s_x, self.k_proj.weight_scale, s_k),
            L.ST_V: linear_eff_f32(s_x, self.v_proj.weight_scale, s_v), L.ST_QK: np.float32(effs[L.ST_QK]),
            L.ST_AV: np.float32(effs[L.ST_AV]), L.ST_O: linear_eff_f32(s_o, self.out_proj.weight_scale, s_out),
        }
        zps = {L.ST_Q: self.q_proj.zero_point, L.ST_K: self.k_proj.zero_point, L.ST_V: self.v_proj.zero_point,
               L.ST_QK: self.matmul1.zero_point, L.ST_AV: self.matmul2.zero_point, L.ST_O: self.out_proj.zero_point}
        mult = np.zeros((L.NUM_STAGES, H), np.int32)
        shift = np.ones((L.NUM_STAGES, H), np.int32)
        add = np.zeros((L.NUM_STAGES, H), np.int32)
        eff = np.zeros((L.NUM_STAGES, H), np.float32)
        for st, e in effs.items():
            mult[st], shift[st] = calculate_hw_params(e)
            add[st] = zps[st]
            eff[st] = eff32[st]

        def split_w(lin):   # torch [H*hd, E] -> golden [H, E, hd]
            w = lin.weight_int8.cpu().numpy()
            return np.ascontiguousarray(w.reshape(H, hd, E).transpose(0, 2, 1))

        def bias_i32(lin, s_in):
            return bias_to_int32(lin.bias.cpu().numpy(), s_in * lin.weight_scale)

        Wo = np.ascontiguousarray(self.out_proj.weight_int8.cpu().numpy().T.reshape(H, hd, E))   # [P_total,E] -> [H,hd,E]
        Bo = np.zeros((H, E), np.int32)
        Bo[0] = bias_i32(self.out_proj, s_o)
        return ITAQuantParams(S=seq_len, E=E, P=hd, H=H, Wq=split_w(self.q_proj), Wk=split_w(self.k_proj),
                              Wv=split_w(self.v_proj), Wo=Wo, Bq=bias_i32(self.q_proj, s_x).reshape(H, hd),
                              Bk=bias_i32(self.k_proj, s_x).reshape(H, hd), Bv=bias_i32(self.v_proj, s_x).reshape(H, hd),
                              Bo=Bo, mult=mult, shift=shift, add=add, eff=eff, round_mode=self.round_mode,
                              softmax_mode=self.softmax_mode,
                              head_mode=L.HEADS_TORCH if H > 1 else L.HEADS_PYITA)

    def engine(self, seq_len: int, device: torch.device) -> ITAEngine:
        key = (seq_len, device.index if device.index is not None else torch.cuda.current_device())
        if key not in self._engines:
            self._engines[key] = ITAEngine(self.quant_params(seq_len), torch.device("cuda", key[1]))
        return self._engines[key]

    # ------------------------------------------------------------------------------------------------ forward
    def forward_int8(self, x_q: torch.Tensor) -> torch.Tensor:
        """Pre-quantised entry: int8 cuda [B,N,E] -> int8 cuda [B,N,E] (skips the quant / dequant stubs)."""
        if x_q.dim() != 3 or x_q.shape[-1] != self.embed_dim:
            raise ValueError(f"expected [B,N,{self.embed_dim}], got {tuple(x_q.shape)}")
        if not x_q.is_cuda:
            raise ValueError("ITASelfAttention runs on CUDA tensors only (no CPU fallback)")
        return self.engine(int(x_q.shape[1]), x_q.device).forward(x_q.contiguous())

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """f32 [B,N,E] -> f32 [B,N,E]: quant -> int8 attention -> dequant (layers.py:101-127)."""
        if x.dtype == torch.int8:
            return self.forward_int8(x)
        if x.dim() != 3 or x.shape[-1] != self.embed_dim:
            raise ValueError(f"expected [B,N,{self.embed_dim}], got {tuple(x.shape)}")
        if not x.is_cuda:
            raise ValueError("ITASelfAttention runs on CUDA tensors only (no CPU fallback)")
        x_q = quantize(x.float(), float(self.quant.scale))
        y_q = self.forward_int8(x_q)
        return dequantize(y_q, float(self.out_proj.scale))
