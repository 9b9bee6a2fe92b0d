"""``ITASelfAttention`` -- the B200 drop-in for the reference operator ``ITASelfAttention_QAT`` *after conversion*
(reference ``models/ITA/QAT/layers.py:77-127``).

Same constructor (``embed_dim, proj_dim, num_heads``), same ``forward(x: f32[B,N,E]) -> f32[B,N,E]``, and the same
load-bearing submodule names (``q_proj, k_proj, v_proj, out_proj, custom_softmax, matmul1, matmul2, quant, dequant``) that
the reference's hooks and state-dict loading address (``tests/export_and_validation_W_B.py:86,167-179,409-411``;
``tests/inspect_quantized_model.py:138-146``).  The submodules are parameter holders -- the arithmetic of the whole block
is one call into libita_b200.so -- but they behave like the reference's where the reference's tools touch them:

* **state dict**: every scale / zero-point / int8 weight is a buffer, so ``state_dict()`` round-trips, and
  ``load_state_dict`` also accepts the *converted reference checkpoint* keys (``q_proj._packed_params._packed_params``,
  ``q_proj.scale``, ``q_proj.zero_point``, ``matmul1.scale`` ...; loaded ``strict=False`` at
  ``tests/export_and_validation_W_B.py:387-392``);
* **forward hooks** registered on a named submodule fire with quantised torch tensors (``int_repr()`` / ``q_scale()`` as
  the reference's capture hooks read them, ``W_B.py:30-47``): a forward with hooks present runs the per-stage kernels.
  (The reference's ``PyTorchDataExtractor`` selects modules by ``isinstance(m, nnq.Linear)``, which these holders are not:
  hook by name, as ``inspect_quantized_model.py`` does.)

``from_converted(block)`` reads exactly the fields a converted block carries.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L
from .engine import ITAEngine, dequantize, quantize
from .params import ITAQuantParams, bias_to_int32, calculate_hw_params, linear_eff_f32


class _ScalarBuffers(nn.Module):
    """Scalar quantisation parameters kept as float64 / int64 buffers that can still be assigned plain numbers."""
    _SCALARS: Dict[str, torch.dtype] = {}

    def __setattr__(self, name, value):
        if name in self._SCALARS and name in self._buffers and not isinstance(value, torch.Tensor):
            self._buffers[name] = torch.tensor(value, dtype=self._SCALARS[name])
            return
        super().__setattr__(name, value)

    def _load_scalars(self, state_dict, prefix):
        for name, dt in self._SCALARS.items():
            if prefix + name in state_dict:
                state_dict[prefix + name] = torch.as_tensor(state_dict[prefix + name]).reshape(()).to(dt)


class QuantizedLinearParams(_ScalarBuffers):
    """int8 weight [out,in] (torch layout), float bias, weight scale, output scale / zero point."""
    _SCALARS = {"weight_scale": torch.float64, "scale": torch.float64, "zero_point": torch.int64}

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.register_buffer("weight_int8", torch.zeros(out_features, in_features, dtype=torch.int8))
        self.register_buffer("bias", torch.zeros(out_features, dtype=torch.float32))
        self.register_buffer("weight_scale", torch.tensor(1.0, dtype=torch.float64))
        self.register_buffer("scale", torch.tensor(1.0, dtype=torch.float64))     # output scale, as torch.ao.nn.quantized.Linear
        self.register_buffer("zero_point", torch.tensor(0, dtype=torch.int64))

    def set_packed(self, packed) -> None:
        """(qweight, bias) tuple or LinearPackedParams ScriptObject of a converted ``quantized.Linear``."""
        if not isinstance(packed, (tuple, list)):
            packed = torch.ops.quantized.linear_unpack(packed)
        qw, b = packed
        if tuple(qw.shape) != (self.out_features, self.in_features):
            raise ValueError(f"packed weight has shape {tuple(qw.shape)}, expected {(self.out_features, self.in_features)}")
        if int(qw.q_zero_point()) != 0:
            raise ValueError("weight zero-point must be 0 (ita_symmetric_qconfig)")
        self.weight_int8.copy_(qw.int_repr())
        self.weight_scale = float(qw.q_scale())
        self.bias.copy_(torch.zeros_like(self.bias) if b is None else b.detach().float())

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        pk = prefix + "_packed_params._packed_params"
        if pk in state_dict:                        # converted reference checkpoint
            self.set_packed(state_dict.pop(pk))
            state_dict.pop(prefix + "_packed_params.dtype", None)
            for name in ("weight_int8", "bias", "weight_scale"):
                state_dict.setdefault(prefix + name, getattr(self, name))
        self._load_scalars(state_dict, prefix)
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)


class ScaleHolder(_ScalarBuffers):
    _SCALARS = {"scale": torch.float64, "zero_point": torch.int64}

    def __init__(self, scale: float = 1.0, zero_point: int = 0):
        super().__init__()
        self.register_buffer("scale", torch.tensor(float(scale), dtype=torch.float64))
        self.register_buffer("zero_point", torch.tensor(int(zero_point), dtype=torch.int64))

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        self._load_scalars(state_dict, prefix)
        for name in self._SCALARS:                  # FloatFunctional / DeQuantize / softmax carry no entry in a reference checkpoint
            state_dict.setdefault(prefix + name, getattr(self, name))
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)


class ITASelfAttention(nn.Module):
    _LINEARS = ("q_proj", "k_proj", "v_proj", "out_proj")

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
        self._engines: Dict[tuple, ITAEngine] = {}

    # ------------------------------------------------------------------------------------------------ construction
    @classmethod
    def from_converted(cls, block, round_mode: int = L.ROUND_TORCH_RNE_F32,
                       softmax_mode: int = L.SOFTMAX_REPO_GLOBAL) -> "ITASelfAttention":
        """Build from a converted reference block (``torch.ao.quantization.convert`` output)."""
        qw = block.q_proj.weight()
        m = cls(qw.shape[1], qw.shape[0], int(block.num_heads), round_mode, softmax_mode)
        for name in cls._LINEARS:
            src, dst = getattr(block, name), getattr(m, name)
            dst.set_packed((src.weight(), src.bias()))
            dst.scale, dst.zero_point = float(src.scale), int(src.zero_point)
        m.matmul1.scale, m.matmul1.zero_point = float(block.matmul1.scale), int(block.matmul1.zero_point)
        m.matmul2.scale, m.matmul2.zero_point = float(block.matmul2.scale), int(block.matmul2.zero_point)
        m.quant.scale, m.quant.zero_point = float(block.quant.scale), int(block.quant.zero_point)
        return m

    def _params_version(self) -> tuple:
        """Changes whenever a buffer is replaced or written in place: cached engines are keyed by it."""
        return tuple((id(b), b._version) for b in self.buffers())

    def quant_params(self, seq_len: int) -> ITAQuantParams:
        """Scales -> golden-model parameter set, exactly as the reference harness derives it
        (``tests/export_and_validation_W_B.py:247-305,479-491``)."""
        H, hd, E = self.num_heads, self.head_dim, self.embed_dim
        for name in self._LINEARS + ("matmul1", "matmul2", "quant", "custom_softmax"):
            zp = int(getattr(self, name).zero_point)
            if zp != 0:
                raise ValueError(f"{name}.zero_point = {zp}: the ITA attention path is symmetric (zero zero-points, "
                                 "ita_symmetric_qconfig, models/ITA/QAT/layers.py:10-21)")
        s_x = float(self.quant.scale)
        s_q, s_k, s_v = float(self.q_proj.scale), float(self.k_proj.scale), float(self.v_proj.scale)
        s_a, s_sm, s_o, s_out = float(self.matmul1.scale), float(self.custom_softmax.scale), float(self.matmul2.scale), \
            float(self.out_proj.scale)
        ws = {n: float(getattr(self, n).weight_scale) for n in self._LINEARS}
        effs = {
            L.ST_Q: s_x * ws["q_proj"] / s_q, L.ST_K: s_x * ws["k_proj"] / s_k, L.ST_V: s_x * ws["v_proj"] / s_v,
            L.ST_QK: s_q * s_k / s_a, L.ST_AV: s_sm * s_v / s_o, L.ST_O: s_o * ws["out_proj"] / s_out,
        }
        # The fp32 requant scale as the reference backend forms it: quantized Linear (qnnpack) computes
        # (f32(s_in) * f32(s_w)) / f32(s_out) in float32; the patched matmuls multiply by the double product cast to f32
        # (tests/export_and_validation_W_B.py:139-142).  The two differ by an ulp now and then, which flips near-ties.
        eff32 = {
            L.ST_Q: linear_eff_f32(s_x, ws["q_proj"], s_q), L.ST_K: linear_eff_f32(s_x, ws["k_proj"], s_k),
            L.ST_V: linear_eff_f32(s_x, ws["v_proj"], s_v), L.ST_QK: np.float32(effs[L.ST_QK]),
            L.ST_AV: np.float32(effs[L.ST_AV]), L.ST_O: linear_eff_f32(s_o, ws["out_proj"], s_out),
        }
        mult = np.zeros((L.NUM_STAGES, H), np.int32)
        shift = np.ones((L.NUM_STAGES, H), np.int32)
        add = np.zeros((L.NUM_STAGES, H), np.int32)
        eff = np.zeros((L.NUM_STAGES, H), np.float32)
        for st, e in effs.items():
            mult[st], shift[st] = calculate_hw_params(e)
            eff[st] = eff32[st]

        def split_w(lin):   # torch [H*hd, E] -> golden [H, E, hd]
            w = lin.weight_int8.cpu().numpy()
            return np.ascontiguousarray(w.reshape(H, hd, E).transpose(0, 2, 1))

        def bias_i32(lin, s_in, name):
            return bias_to_int32(lin.bias.cpu().numpy(), s_in * ws[name])

        Wo = np.ascontiguousarray(self.out_proj.weight_int8.cpu().numpy().T.reshape(H, hd, E))   # [P_total,E] -> [H,hd,E]
        Bo = np.zeros((H, E), np.int32)
        Bo[0] = bias_i32(self.out_proj, s_o, "out_proj")
        return ITAQuantParams(S=seq_len, E=E, P=hd, H=H, Wq=split_w(self.q_proj), Wk=split_w(self.k_proj),
                              Wv=split_w(self.v_proj), Wo=Wo, Bq=bias_i32(self.q_proj, s_x, "q_proj").reshape(H, hd),
                              Bk=bias_i32(self.k_proj, s_x, "k_proj").reshape(H, hd),
                              Bv=bias_i32(self.v_proj, s_x, "v_proj").reshape(H, hd),
                              Bo=Bo, mult=mult, shift=shift, add=add, eff=eff, round_mode=self.round_mode,
                              softmax_mode=self.softmax_mode,
                              head_mode=L.HEADS_TORCH if H > 1 else L.HEADS_PYITA)

    def engine(self, seq_len: int, device: torch.device) -> ITAEngine:
        dev = device.index if device.index is not None else torch.cuda.current_device()
        key = (seq_len, dev, self.round_mode, self.softmax_mode, self._params_version())
        if key not in self._engines:
            for e in self._engines.values():          # parameters changed: drop the stale handles
                e.close()
            self._engines.clear()
            self._engines[key] = ITAEngine(self.quant_params(seq_len), torch.device("cuda", dev))
        return self._engines[key]

    # ------------------------------------------------------------------------------------------------ forward
    def _has_hooks(self) -> bool:
        return any(len(m._forward_hooks) or len(m._forward_pre_hooks) for m in self.children())

    def _check(self, x: torch.Tensor) -> None:
        if x.dim() != 3 or x.shape[-1] != self.embed_dim:
            raise ValueError(f"expected [B,N,{self.embed_dim}], got {tuple(x.shape)}")
        if not x.is_cuda:
            raise ValueError("ITASelfAttention runs on CUDA tensors only (no CPU fallback)")

    def forward_int8(self, x_q: torch.Tensor) -> torch.Tensor:
        """Pre-quantised entry: int8 cuda [B,N,E] -> int8 cuda [B,N,E] (skips the quant / dequant stubs)."""
        self._check(x_q)
        return self.engine(int(x_q.shape[1]), x_q.device).forward(x_q.contiguous())

    def _forward_with_hooks(self, x: torch.Tensor) -> torch.Tensor:
        """Per-stage kernels, firing the hooks registered on the named submodules with quantised tensors shaped like the
        reference's (``layers.py:103-126``): Q/K/V [B,N,P], logits and probabilities [B,H,N,N], context [B,N,P]."""
        B, N = int(x.shape[0]), int(x.shape[1])
        H, hd = self.num_heads, self.head_dim

        def qt(t, scale):      # int8 -> qint8, uint8 -> quint8, on the host like the reference's CPU model
            return torch._make_per_tensor_quantized_tensor(t.cpu().contiguous(), float(scale), 0)

        def fire(mod, inputs, output):
            for hk in mod._forward_pre_hooks.values():
                hk(mod, inputs)
            for hk in mod._forward_hooks.values():
                hk(mod, inputs, output)
            return output

        merge = lambda t: t.permute(0, 2, 1, 3).reshape(B, N, H * hd)      # [B,H,N,hd] -> [B,N,P]
        x_q = quantize(x.float(), float(self.quant.scale))
        xq_t = fire(self.quant, (x.cpu(),), qt(x_q, self.quant.scale))
        st = self.engine(N, x.device).forward_stages(x_q)
        for name, key in (("q_proj", "Qp"), ("k_proj", "Kp"), ("v_proj", "Vp")):
            fire(getattr(self, name), (xq_t,), qt(merge(st[key]), getattr(self, name).scale))
        a_t = fire(self.matmul1, (), qt(st["A"], self.matmul1.scale))
        p_t = fire(self.custom_softmax, (a_t,), qt(st["Asm"], self.custom_softmax.scale))
        c_t = qt(merge(st["O"]), self.matmul2.scale)
        fire(self.matmul2, (p_t,), c_t)
        y_t = fire(self.out_proj, (c_t,), qt(st["Y"], self.out_proj.scale))
        y = dequantize(st["Y"], float(self.out_proj.scale))
        fire(self.dequant, (y_t,), y.cpu())
        return y

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """f32 [B,N,E] -> f32 [B,N,E]: quant -> int8 attention -> dequant (layers.py:101-127)."""
        if x.dtype == torch.int8:
            return self.forward_int8(x)
        self._check(x)
        if self._has_hooks():
            return self._forward_with_hooks(x)
        x_q = quantize(x.float(), float(self.quant.scale))
        y_q = self.forward_int8(x_q)
        return dequantize(y_q, float(self.out_proj.scale))
