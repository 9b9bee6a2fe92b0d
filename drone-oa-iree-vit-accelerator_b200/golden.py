"""``Transformer`` -- drop-in for the golden-model object the reference harness drives
(``from third_party.ITA_FPGA.PyITA.ITA import Transformer``; call sites ``tests/export_and_validation_W_B.py:518-555``,
``tests/tuki.py:426-461``, ``tests/onnx_backup/onnx_ITA_verification.py:236-267``).

Same constructor keywords, same ``step1_Qp() .. step6_O()``, ``feedforward_layer()``, ``export_hwpe()``, same NumPy
result attributes (``Qp_requant, Kp_requant, Vp_requant, A_requant, A_partial_softmax`` (writable), ``O_soft_requant,
Out_soft_requant, FFp_requant, FF2p_requant``).  Every step runs on the GPU through ``ita_mha_step`` /
``ita_ffn_forward``; to run the reference harness unmodified alias this module:

    sys.modules["third_party.ITA_FPGA.PyITA.ITA"] = importlib.import_module("drone-oa-iree-vit-accelerator_b200.golden")

The upstream golden model is absent from the reference tree; arithmetic follows SURVEY.md App. A (``round_mode`` /
``softmax_mode`` keywords select the variants; defaults: round-half-up requant, the repo's ITAmax).
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch

from . import _lib as L
from .engine import ITAEngine
from .params import ITAQuantParams


class Transformer:
    def __init__(self, ITA_N: int = 16, path: Optional[str] = None, activation: str = "relu", S: int = 64, P: int = 64,
                 E: int = 64, F: int = 64, H: int = 1, Q=None, K=None, V=None, Wq=None, Wk=None, Wv=None, Wo=None,
                 Bq=None, Bk=None, Bv=None, Bo=None, FF_in=None, Wff=None, Wff2=None, Bff=None, Bff2=None,
                 quant_params: Optional[dict] = None, round_mode: int = L.ROUND_HALF_UP,
                 softmax_mode: int = L.SOFTMAX_REPO_GLOBAL, device=None, **_ignored):
        if activation not in ("relu", "identity"):
            raise ValueError(f"activation {activation!r} is not supported (relu | identity)")
        if quant_params is None:
            raise ValueError("quant_params is required (per-stage mult/shift/add)")
        self.ITA_N, self.path, self.activation = ITA_N, path, activation
        self.S, self.P, self.E, self.F, self.H = S, P, E, F, H
        i8 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.int8)
        i32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.int32)
        self.Q, self.K, self.V = i8(Q), i8(K), i8(V)
        if self.Q is None or self.Q.shape != (S, E):
            raise ValueError(f"Q must be int8 [{S},{E}]")
        self.FF_in = i8(FF_in)
        tensors = dict(Wq=i8(Wq), Wk=i8(Wk), Wv=i8(Wv), Wo=i8(Wo), Bq=i32(Bq), Bk=i32(Bk), Bv=i32(Bv), Bo=i32(Bo),
                       Wff=i8(Wff), Wff2=i8(Wff2), Bff=i32(Bff), Bff2=i32(Bff2))
        has_ffn = F > 0 and all(tensors[k] is not None for k in ("Wff", "Wff2", "Bff", "Bff2"))
        qp = dict(quant_params)
        if activation == "identity":
            qp.pop("relu", None)
        self.params = ITAQuantParams.from_quant_dict(S, E, P, H, tensors, qp, F=F if has_ffn else 0,
                                                     round_mode=round_mode, softmax_mode=softmax_mode,
                                                     head_mode=L.HEADS_PYITA)
        if "relu" not in qp:
            self.params.mult[L.ST_RELU] = 0       # plain ReLU, no second requant
        self._eng = ITAEngine(self.params, device)
        self._dev = self._eng.device
        self.Qp_requant = self.Kp_requant = self.Vp_requant = None
        self.A_requant = self.A_partial_softmax = None
        self.O_soft_requant = self.Out_soft_requant = self.Out_sum_requant = None
        self.FFp_requant = self.FF2p_requant = None

    # ---- plumbing
    def _run(self, step, in0, in1, shape, dt=torch.int8) -> np.ndarray:
        out = torch.empty(shape, dtype=dt, device=self._dev)
        t0 = torch.from_numpy(np.ascontiguousarray(in0)).to(self._dev)
        t1 = torch.from_numpy(np.ascontiguousarray(in1)).to(self._dev) if in1 is not None else None
        self._eng.step(step, t0, t1, out, 1)
        return out.cpu().numpy()

    # ---- golden-model steps
    def step1_Qp(self):
        self.Qp_requant = self._run(L.STEP1_QP, self.Q, None, (self.H, self.S, self.P))

    def step2_Kp(self):
        self.Kp_requant = self._run(L.STEP2_KP, self.K if self.K is not None else self.Q, None, (self.H, self.S, self.P))

    def step3_Vp(self):
        self.Vp_requant = self._run(L.STEP3_VP, self.V if self.V is not None else self.Q, None, (self.H, self.S, self.P))

    def step4_QK(self, no_partial_softmax: bool = False):
        self.A_requant = self._run(L.STEP4_QK, self.Qp_requant, self.Kp_requant, (self.H, self.S, self.S))
        self.A_partial_softmax = self._run(L.STEP4_SOFTMAX, self.A_requant, None, (self.H, self.S, self.S), torch.uint8)

    def step5_AV(self):
        asm = np.ascontiguousarray(self.A_partial_softmax).astype(np.uint8)
        self.O_soft_requant = self._run(L.STEP5_AV, asm, self.Vp_requant, (self.H, self.S, self.P))

    def step6_O(self):
        self.Out_soft_requant = self._run(L.STEP6_O, self.O_soft_requant, None, (self.H, self.S, self.E))

    def step7_Osum(self):
        if self.H > 1:
            self.Out_sum_requant = self._run(L.STEP7_OSUM, self.Out_soft_requant, None, (1, self.S, self.E))
        else:
            self.Out_sum_requant = self.Out_soft_requant.copy()

    def feedforward_layer(self):
        if self.params.F == 0 or self.FF_in is None:
            raise RuntimeError("Transformer was built without FFN tensors")
        x = torch.from_numpy(self.FF_in.reshape(1, self.S, self.E)).to(self._dev)
        f1, f2 = self._eng.ffn(x, want_f1=True)
        self.FFp_requant = f1.cpu().numpy()          # [1,S,F], post-activation (the harness compares it to the ReLU output)
        self.FF2p_requant = f2.cpu().numpy()         # [1,S,E]

    def run_all(self):
        self.step1_Qp(); self.step2_Kp(); self.step3_Vp(); self.step4_QK(); self.step5_AV(); self.step6_O(); self.step7_Osum()
        return self.Out_sum_requant

    def export_hwpe(self):
        """Writes every tensor of the run as ``<path>/<name>.npy`` plus a ``manifest.txt``.  The upstream golden model
        writes ``mem.txt`` etc. in the RTL test-bench's format; that format is unpinned here (absent code), so this is
        a plain, documented container rather than a guess at it."""
        if not self.path:
            return
        os.makedirs(self.path, exist_ok=True)
        names = ["Q", "K", "V", "Qp_requant", "Kp_requant", "Vp_requant", "A_requant", "A_partial_softmax",
                 "O_soft_requant", "Out_soft_requant", "Out_sum_requant", "FFp_requant", "FF2p_requant"]
        with open(os.path.join(self.path, "manifest.txt"), "w") as mf:
            mf.write(f"S={self.S} E={self.E} P={self.P} F={self.F} H={self.H} activation={self.activation}\n")
            for n in names:
                a = getattr(self, n, None)
                if a is None:
                    continue
                np.save(os.path.join(self.path, n + ".npy"), a)
                mf.write(f"{n} {a.dtype} {list(a.shape)}\n")
            for k in ("mult", "shift", "add"):
                np.save(os.path.join(self.path, f"requant_{k}.npy"), getattr(self.params, k))
