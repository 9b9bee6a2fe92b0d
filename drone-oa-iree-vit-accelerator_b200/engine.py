"""ITAEngine -- owner of one ``ita_mha_handle`` (device copies of an attention layer's int8 weights, int32 biases and
per-stage requant parameters) and the Python face of the C-ABI forward calls.

torch is used for what the task allows it for: device memory, streams, and (in bench/tests) ``torch.distributed``.
The arithmetic is entirely inside libita_b200.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib as L
from .params import ITAQuantParams


def _np_ptr(a: Optional[np.ndarray], ty):
    return a.ctypes.data_as(ty) if a is not None else ty()


def _stream_ptr(stream: Optional[torch.cuda.Stream], device: torch.device):
    s = stream if stream is not None else torch.cuda.current_stream(device)
    return C.c_void_p(s.cuda_stream)


class ITAEngine:
    """One ITA attention layer on one GPU.

    ``forward(x)``          x int8 cuda [B,S,E] -> int8 cuda [B,S,E]   (asynchronous on the current stream)
    ``forward_host(x)``     x int8 host ndarray/tensor -> int8 host    (H2D + kernels + D2H, synchronous)
    ``forward_stages(x)``   every golden-model intermediate (staged kernels) as a dict of cuda tensors
    ``step(...)``           one golden-model step
    ``ffn(x)``              the FFN block (needs F > 0 parameters)
    """

    def __init__(self, params: ITAQuantParams, device: Optional[torch.device] = None):
        self.lib = L.load()
        if not torch.cuda.is_available():
            raise L.ITAError("ITAEngine needs a CUDA device: libita_b200 has no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        p = params
        self.S, self.E, self.P, self.H, self.F = p.S, p.E, p.P, p.H, p.F
        self.head_mode = p.head_mode
        keep = []

        def c(a, dt, shape=None):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=dt)
            if shape is not None and tuple(a.shape) != tuple(shape):
                raise ValueError(f"parameter has shape {a.shape}, expected {shape}")
            keep.append(a)
            return a

        H, E, P, F = p.H, p.E, p.P, p.F
        eff = p.eff if p.eff is not None else np.zeros((L.NUM_STAGES, H), np.float32)
        d = L.MhaDesc(L.ABI_VERSION, p.S, E, P, H, F, p.round_mode, p.softmax_mode, p.head_mode, self.device.index,
                      _np_ptr(c(p.Wq, np.int8, (H, E, P)), L._i8p), _np_ptr(c(p.Wk, np.int8, (H, E, P)), L._i8p),
                      _np_ptr(c(p.Wv, np.int8, (H, E, P)), L._i8p), _np_ptr(c(p.Wo, np.int8, (H, P, E)), L._i8p),
                      _np_ptr(c(p.Bq, np.int32, (H, P)), L._i32p), _np_ptr(c(p.Bk, np.int32, (H, P)), L._i32p),
                      _np_ptr(c(p.Bv, np.int32, (H, P)), L._i32p), _np_ptr(c(p.Bo, np.int32, (H, E)), L._i32p),
                      _np_ptr(c(p.Wff, np.int8, (1, E, F)) if F else None, L._i8p),
                      _np_ptr(c(p.Wff2, np.int8, (1, F, E)) if F else None, L._i8p),
                      _np_ptr(c(p.Bff, np.int32, (1, F)) if F else None, L._i32p),
                      _np_ptr(c(p.Bff2, np.int32, (1, E)) if F else None, L._i32p),
                      _np_ptr(c(p.mult, np.int32, (L.NUM_STAGES, H)), L._i32p),
                      _np_ptr(c(p.shift, np.int32, (L.NUM_STAGES, H)), L._i32p),
                      _np_ptr(c(p.add, np.int32, (L.NUM_STAGES, H)), L._i32p),
                      _np_ptr(c(eff, np.float32, (L.NUM_STAGES, H)), L._f32p))
        h = C.c_void_p()
        L.check(self.lib.ita_mha_create(C.byref(d), C.byref(h)), "ita_mha_create")
        self._h = h

    # -- lifetime
    def close(self):
        if getattr(self, "_h", None):
            self.lib.ita_mha_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def fused_supported(self) -> bool:
        return bool(self.lib.ita_mha_fused_supported(self._h))

    def fused_unsupported_reason(self) -> str:
        return "" if self.fused_supported else self.lib.ita_last_error().decode()

    def describe(self) -> str:
        """Which kernels / requant forms this engine runs (``ita_mha_describe``)."""
        buf = C.create_string_buffer(1024)
        L.check(self.lib.ita_mha_describe(self._h, buf, len(buf)), "ita_mha_describe")
        return buf.value.decode()

    # -- helpers
    def _check_x(self, x: torch.Tensor, last: int) -> int:
        if x.dtype not in (torch.int8, torch.uint8):
            raise TypeError(f"expected an int8 tensor, got {x.dtype}")
        if not x.is_cuda or x.device != self.device:
            raise ValueError(f"expected a tensor on {self.device}, got {x.device}")
        if x.dim() < 2 or x.shape[-1] != last:
            raise ValueError(f"bad shape {tuple(x.shape)}")
        if not x.is_contiguous():
            raise ValueError("expected a contiguous tensor")
        return 0

    def _batch_of(self, x: torch.Tensor) -> int:
        self._check_x(x, self.E)
        if x.dim() == 2:
            if x.shape[0] != self.S:
                raise ValueError(f"bad shape {tuple(x.shape)}, expected [{self.S},{self.E}]")
            return 1
        if x.dim() != 3 or x.shape[1] != self.S:
            raise ValueError(f"bad shape {tuple(x.shape)}, expected [B,{self.S},{self.E}]")
        return int(x.shape[0])

    # -- forward calls
    def forward(self, x: torch.Tensor, out: Optional[torch.Tensor] = None, path: int = L.PATH_AUTO,
                stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        B = self._batch_of(x)
        y = torch.empty_like(x) if out is None else out
        if out is not None:
            self._check_x(out, self.E)
            if out.shape != x.shape:
                raise ValueError("out has the wrong shape")
        L.check(self.lib.ita_mha_forward(self._h, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), B, path,
                                         _stream_ptr(stream, self.device)), "ita_mha_forward")
        return y

    def sync_check(self, stream: Optional[torch.cuda.Stream] = None) -> None:
        L.check(self.lib.ita_mha_sync_check(self._h, _stream_ptr(stream, self.device)), "ita_mha_sync_check")

    def forward_host(self, x, out=None, path: int = L.PATH_AUTO, stream: Optional[torch.cuda.Stream] = None,
                     sync: bool = True):
        """x: int8 host ``np.ndarray`` or CPU ``torch.Tensor`` [B,S,E] (pinned memory makes the copies asynchronous).
        ``sync=False`` returns as soon as the work is queued on ``stream``: the result is valid after ``sync_check(stream)``
        (two engines on two streams then overlap one call's H2D copies with the other's D2H copies)."""
        is_np = isinstance(x, np.ndarray)
        xt = torch.from_numpy(x) if is_np else x
        if xt.is_cuda or xt.dtype != torch.int8 or not xt.is_contiguous():
            raise ValueError("forward_host expects a contiguous int8 host tensor")
        if xt.dim() != 3 or tuple(xt.shape[1:]) != (self.S, self.E):
            raise ValueError(f"bad shape {tuple(xt.shape)}, expected [B,{self.S},{self.E}]")
        yt = torch.empty_like(xt) if out is None else (torch.from_numpy(out) if isinstance(out, np.ndarray) else out)
        with torch.cuda.device(self.device):
            fn = self.lib.ita_mha_forward_host if sync else self.lib.ita_mha_forward_host_async
            L.check(fn(self._h, C.c_void_p(xt.data_ptr()), C.c_void_p(yt.data_ptr()), int(xt.shape[0]), path,
                       _stream_ptr(stream, self.device)), "ita_mha_forward_host")
        return yt.numpy() if is_np and out is None else (out if out is not None else yt)

    def _alloc_stages(self, B: int) -> Dict[str, torch.Tensor]:
        S, E, P, H = self.S, self.E, self.P, self.H
        nout = H if self.head_mode == L.HEADS_PYITA else 1
        mk = lambda *shape, dt=torch.int8: torch.empty(shape, dtype=dt, device=self.device)
        return {"Qp": mk(B, H, S, P), "Kp": mk(B, H, S, P), "Vp": mk(B, H, S, P), "A": mk(B, H, S, S),
                "Asm": mk(B, H, S, S, dt=torch.uint8), "O": mk(B, H, S, P), "Out": mk(B, nout, S, E), "Y": mk(B, S, E)}

    def forward_stages(self, x: torch.Tensor, fused_debug: bool = False,
                       stream: Optional[torch.cuda.Stream] = None) -> Dict[str, torch.Tensor]:
        """All golden-model intermediates.  ``fused_debug=True`` runs the fused tcgen05 kernel with its on-chip
        intermediates dumped instead of the staged kernels (one head: ``Out`` is then a copy of ``Y``; the multi-head
        kernels dump ``Out`` themselves)."""
        B = self._batch_of(x)
        t = self._alloc_stages(B)
        st = L.MhaStages(*(C.c_void_p(t[k].data_ptr()) for k in ("Qp", "Kp", "Vp", "A", "Asm", "O", "Out")))
        fn = self.lib.ita_mha_forward_fused_debug if fused_debug else self.lib.ita_mha_forward_stages
        L.check(fn(self._h, C.c_void_p(x.data_ptr()), C.c_void_p(t["Y"].data_ptr()), B, C.byref(st),
                   _stream_ptr(stream, self.device)), fn.__name__)
        if fused_debug and self.H == 1:
            t["Out"] = t["Y"].reshape(B, 1, self.S, self.E).clone()
        return t

    def step(self, step: int, in0: torch.Tensor, in1: Optional[torch.Tensor], out: torch.Tensor, batch: int,
             stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        for t in (in0, in1, out):
            if t is not None and (not t.is_cuda or not t.is_contiguous()):
                raise ValueError("step() needs contiguous cuda tensors")
        L.check(self.lib.ita_mha_step(self._h, step, C.c_void_p(in0.data_ptr()),
                                      C.c_void_p(in1.data_ptr() if in1 is not None else None),
                                      C.c_void_p(out.data_ptr()), batch, _stream_ptr(stream, self.device)), "ita_mha_step")
        return out

    @property
    def ffn_fused_supported(self) -> bool:
        return bool(self.lib.ita_ffn_fused_supported(self._h))

    def ffn_fused_unsupported_reason(self) -> str:
        return "" if self.ffn_fused_supported else self.lib.ita_last_error().decode()

    def ffn(self, x: torch.Tensor, want_f1: bool = False, stream: Optional[torch.cuda.Stream] = None,
            path: int = L.PATH_AUTO, out: Optional[torch.Tensor] = None):
        """FFN block fc2(relu(fc1(x))), int8 in / int8 out.  ``path``: PATH_FUSED = the tcgen05 kernel (F1 stays on chip),
        PATH_STAGED = CUDA-core GEMMs, PATH_AUTO = fused when supported."""
        B = self._batch_of(x)
        f2 = torch.empty_like(x) if out is None else out
        f1 = torch.empty((B, self.S, self.F), dtype=torch.int8, device=self.device) if want_f1 else None
        L.check(self.lib.ita_ffn_forward_path(self._h, C.c_void_p(x.data_ptr()),
                                              C.c_void_p(f1.data_ptr() if f1 is not None else None),
                                              C.c_void_p(f2.data_ptr()), B, path, _stream_ptr(stream, self.device)),
                "ita_ffn_forward")
        return (f1, f2) if want_f1 else f2

    def layer_params(self, s_attn_in: float, s_attn_out: float, s_ffn_in: float, s_ffn_out: float, ln1_w, ln1_b, ln2_w, ln2_b,
                     eps: float = 1e-5):
        """Bundle the float side of an encoder layer (scales of the four stubs, the two LayerNorms) for ``layer()``."""
        t = [torch.as_tensor(a, dtype=torch.float32).to(self.device).contiguous() for a in (ln1_w, ln1_b, ln2_w, ln2_b)]
        for a in t:
            if a.numel() != self.E:
                raise ValueError(f"LayerNorm parameters must have {self.E} elements")
        lp = L.LayerParams(float(s_attn_in), float(s_attn_out), float(s_ffn_in), float(s_ffn_out), float(eps),
                           *(C.c_void_p(a.data_ptr()) for a in t))
        lp._keep = t
        return lp

    def layer(self, x: torch.Tensor, lp, out: Optional[torch.Tensor] = None, stream: Optional[torch.cuda.Stream] = None):
        """One encoder layer, f32 cuda [B,S,E] -> f32 cuda [B,S,E] (reference QAT/model.py:100-113)."""
        if x.dtype != torch.float32 or not x.is_cuda or not x.is_contiguous() or x.dim() != 3 or \
                tuple(x.shape[1:]) != (self.S, self.E):
            raise ValueError(f"layer() expects a contiguous float32 cuda tensor [B,{self.S},{self.E}]")
        y = torch.empty_like(x) if out is None else out
        L.check(self.lib.ita_encoder_layer_forward(self._h, C.byref(lp), C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                   int(x.shape[0]), _stream_ptr(stream, self.device)),
                "ita_encoder_layer_forward")
        return y


    @property
    def layer_static_max_batch(self) -> int:
        """Largest batch ``layer_static`` accepts for this handle's S and E (ITA_LAYER_STATIC_SCRATCH_BYTES / (8*S*E))."""
        return L.LAYER_STATIC_SCRATCH_BYTES // (8 * self.S * self.E)

    def layer_static(self, x: torch.Tensor, lp, out: torch.Tensor, stream: Optional[torch.cuda.Stream] = None):
        """``layer()`` for a caller that must not allocate device memory at call time (a real-time loop, a host whose memory
        pool is already handed out): the C call keeps its intermediates in a fixed 16 MiB array inside the library and
        performs no cudaMalloc / cudaFree, and this wrapper creates no tensor either, so ``out`` (f32 cuda [B,S,E]) is
        mandatory.  Bit-identical to ``layer()``.  Limits: ``B <= layer_static_max_batch``; single-head handles on the
        fused tcgen05 kernels only; one array per GPU shared by every call, so calls that could overlap on the GPU must be
        ordered (one stream); first call on a device belongs to start-up.  See include/ita_b200.h."""
        for t, name in ((x, "x"), (out, "out")):
            if t.dtype != torch.float32 or not t.is_cuda or not t.is_contiguous() or t.dim() != 3 or \
                    tuple(t.shape[1:]) != (self.S, self.E):
                raise ValueError(f"layer_static() expects {name} as a contiguous float32 cuda tensor [B,{self.S},{self.E}]")
        if out.shape[0] != x.shape[0]:
            raise ValueError("layer_static(): x and out differ in batch")
        L.check(self.lib.ita_encoder_layer_forward_static(self._h, C.byref(lp), C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                                          int(x.shape[0]), _stream_ptr(stream, self.device)),
                "ita_encoder_layer_forward_static")
        return out


def quantize(x: torch.Tensor, scale: float) -> torch.Tensor:
    """QuantStub of the operator boundary (reference models/ITA/QAT/layers.py:103): sat8(rne(x / scale))."""
    if not x.is_cuda or x.dtype != torch.float32:
        raise ValueError("quantize expects a float32 cuda tensor")
    x = x.contiguous()
    q = torch.empty(x.shape, dtype=torch.int8, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.load().ita_quantize_f32(C.c_void_p(x.data_ptr()), C.c_void_p(q.data_ptr()), x.numel(), float(scale),
                                          _stream_ptr(None, x.device)), "ita_quantize_f32")
    return q


def residual_layernorm(x: torch.Tensor, q: torch.Tensor, q_scale: float, weight: torch.Tensor, bias: torch.Tensor, eps: float = 1e-5,
                       quant_scale: Optional[float] = None):
    """``LayerNorm(x + q_scale * q)`` (dequant stub + ``self.add.add`` + LayerNorm of QAT/model.py:104-106) and, when
    ``quant_scale`` is given, the next block's quant stub on the result.  Returns ``y`` or ``(y, y_q)``."""
    if not (x.is_cuda and x.dtype == torch.float32 and q.dtype == torch.int8 and x.shape == q.shape):
        raise ValueError("residual_layernorm expects float32 x and int8 q of the same shape on a CUDA device")
    x, q = x.contiguous(), q.contiguous()
    E = int(x.shape[-1])
    w = weight.to(x.device, torch.float32).contiguous()
    b = bias.to(x.device, torch.float32).contiguous()
    y = torch.empty_like(x)
    yq = torch.empty(x.shape, dtype=torch.int8, device=x.device) if quant_scale is not None else None
    with torch.cuda.device(x.device):
        L.check(L.load().ita_residual_layernorm(C.c_void_p(x.data_ptr()), C.c_void_p(q.data_ptr()), float(q_scale),
                                                C.c_void_p(w.data_ptr()), C.c_void_p(b.data_ptr()), float(eps),
                                                C.c_void_p(y.data_ptr()), C.c_void_p(yq.data_ptr() if yq is not None else None),
                                                float(quant_scale or 0.0), x.numel() // E, E, _stream_ptr(None, x.device)),
                "ita_residual_layernorm")
    return y if yq is None else (y, yq)


def dequantize(q: torch.Tensor, scale: float) -> torch.Tensor:
    """DeQuantStub (reference models/ITA/QAT/layers.py:126): scale * q."""
    if not q.is_cuda or q.dtype != torch.int8:
        raise ValueError("dequantize expects an int8 cuda tensor")
    q = q.contiguous()
    y = torch.empty(q.shape, dtype=torch.float32, device=q.device)
    with torch.cuda.device(q.device):
        L.check(L.load().ita_dequantize_s8(C.c_void_p(q.data_ptr()), C.c_void_p(y.data_ptr()), q.numel(), float(scale),
                                           _stream_ptr(None, q.device)), "ita_dequantize_s8")
    return y


def selftest(device: int = 0) -> str:
    buf = C.create_string_buffer(1 << 14)
    rc = L.load().ita_selftest(device, buf, len(buf))
    rep = buf.value.decode()
    if rc != 0:
        raise L.ITAError(f"ita_selftest failed (code {rc}): {L.load().ita_last_error().decode()}\n{rep}")
    return rep


def microbench(kind: int, device: int = 0, arg: float = 0.0):
    """``ita_microbench`` (roofline denominators).  ``arg`` is the input some kinds read from out[0] (kind 3: seconds)."""
    out = (C.c_double * 8)()
    out[0] = float(arg)
    L.check(L.load().ita_microbench(device, kind, out, 8), "ita_microbench")
    return list(out)
