"""ctypes binding of the C oracle (oracle/ita_ref.c).  TEST INFRASTRUCTURE ONLY -- see oracle/ita_ref.py header.

``load(native=False)`` builds the shared object on demand with the Makefile next to this file (gcc is part of the
image both here and on the GPU box) and returns a thin wrapper whose ``mha`` / ``ffn`` take an
``ita_ref.ITAParams`` and NumPy arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_i8p, _u8p, _i32p, _f32p = C.POINTER(C.c_int8), C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float)


class _Params(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("S", "E", "P", "H", "F", "round_mode", "softmax_mode", "head_mode")] + \
               [(n, _i8p) for n in ("Wq", "Wk", "Wv", "Wo")] + [(n, _i32p) for n in ("Bq", "Bk", "Bv", "Bo")] + \
               [(n, _i8p) for n in ("Wff", "Wff2")] + [(n, _i32p) for n in ("Bff", "Bff2")] + \
               [(n, _i32p) for n in ("mult", "shift", "add")] + [("eff", _f32p)]


def _ptr(a: Optional[np.ndarray], ty):
    return a.ctypes.data_as(ty) if a is not None else ty()


class RefLib:
    def __init__(self, path: str):
        self.path = path
        self.lib = C.CDLL(path)
        self.lib.ita_ref_mha.restype = C.c_int
        self.lib.ita_ref_mha.argtypes = [C.POINTER(_Params), _i8p, C.c_int64, _i8p, _i8p, _i8p, _i8p, _i8p, _u8p,
                                         _i8p, _i8p, C.c_int]
        self.lib.ita_ref_ffn.restype = C.c_int
        self.lib.ita_ref_ffn.argtypes = [C.POINTER(_Params), _i8p, C.c_int64, _i8p, _i8p, C.c_int]
        self.lib.ita_ref_itamax.restype = None
        self.lib.ita_ref_itamax.argtypes = [_i8p, _u8p, C.c_int64, C.c_int, C.c_int]
        self.lib.ita_ref_requant.restype = None
        self.lib.ita_ref_requant.argtypes = [_i32p, _i8p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
        self.lib.ita_ref_max_threads.restype = C.c_int

    def max_threads(self) -> int:
        return int(self.lib.ita_ref_max_threads())

    @staticmethod
    def _pack(p):
        keep = []

        def c(a, dt):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=dt)
            keep.append(a)
            return a

        eff = p.eff if p.eff is not None else np.zeros_like(p.mult, dtype=np.float32)
        cp = _Params(p.S, p.E, p.P, p.H, p.F, p.round_mode, p.softmax_mode, p.head_mode,
                     _ptr(c(p.Wq, np.int8), _i8p), _ptr(c(p.Wk, np.int8), _i8p), _ptr(c(p.Wv, np.int8), _i8p),
                     _ptr(c(p.Wo, np.int8), _i8p),
                     _ptr(c(p.Bq, np.int32), _i32p), _ptr(c(p.Bk, np.int32), _i32p), _ptr(c(p.Bv, np.int32), _i32p),
                     _ptr(c(p.Bo, np.int32), _i32p),
                     _ptr(c(p.Wff, np.int8), _i8p), _ptr(c(p.Wff2, np.int8), _i8p),
                     _ptr(c(p.Bff, np.int32), _i32p), _ptr(c(p.Bff2, np.int32), _i32p),
                     _ptr(c(p.mult, np.int32), _i32p), _ptr(c(p.shift, np.int32), _i32p),
                     _ptr(c(p.add, np.int32), _i32p), _ptr(c(eff, np.float32), _f32p))
        return cp, keep

    def mha(self, p, X: np.ndarray, intermediates: bool = True, nthreads: int = 0) -> Dict[str, np.ndarray]:
        X = np.ascontiguousarray(X, dtype=np.int8)
        if X.ndim == 2:
            X = X[None]
        B = X.shape[0]
        cp, keep = self._pack(p)
        nout = p.H if p.head_mode == 0 else 1
        out = {"Y": np.empty((B, p.S, p.E), np.int8)}
        if intermediates:
            out.update(Qp=np.empty((B, p.H, p.S, p.P), np.int8), Kp=np.empty((B, p.H, p.S, p.P), np.int8),
                       Vp=np.empty((B, p.H, p.S, p.P), np.int8), A=np.empty((B, p.H, p.S, p.S), np.int8),
                       Asm=np.empty((B, p.H, p.S, p.S), np.uint8), O=np.empty((B, p.H, p.S, p.P), np.int8),
                       Out=np.empty((B, nout, p.S, p.E), np.int8))
        g = out.get
        rc = self.lib.ita_ref_mha(C.byref(cp), _ptr(X, _i8p), B, _ptr(out["Y"], _i8p), _ptr(g("Qp"), _i8p),
                                  _ptr(g("Kp"), _i8p), _ptr(g("Vp"), _i8p), _ptr(g("A"), _i8p), _ptr(g("Asm"), _u8p),
                                  _ptr(g("O"), _i8p), _ptr(g("Out"), _i8p), nthreads)
        if rc != 0:
            raise RuntimeError(f"ita_ref_mha failed: {rc}")
        return out

    def ffn(self, p, X: np.ndarray, nthreads: int = 0) -> Dict[str, np.ndarray]:
        X = np.ascontiguousarray(X, dtype=np.int8)
        if X.ndim == 2:
            X = X[None]
        B = X.shape[0]
        cp, keep = self._pack(p)
        F1, F2 = np.empty((B, p.S, p.F), np.int8), np.empty((B, p.S, p.E), np.int8)
        rc = self.lib.ita_ref_ffn(C.byref(cp), _ptr(X, _i8p), B, _ptr(F1, _i8p), _ptr(F2, _i8p), nthreads)
        if rc != 0:
            raise RuntimeError(f"ita_ref_ffn failed: {rc}")
        return {"FF1": F1, "FF2": F2}

    def itamax(self, a: np.ndarray, mode: int = 0) -> np.ndarray:
        a = np.ascontiguousarray(a, dtype=np.int8)
        out = np.empty(a.shape, np.uint8)
        self.lib.ita_ref_itamax(_ptr(a, _i8p), _ptr(out, _u8p), a.size // a.shape[-1], a.shape[-1], mode)
        return out

    def requant(self, acc: np.ndarray, mult: int, shift: int, add: int, mode: int = 0, eff: float = 0.0) -> np.ndarray:
        acc = np.ascontiguousarray(acc, dtype=np.int32)
        out = np.empty(acc.shape, np.int8)
        self.lib.ita_ref_requant(_ptr(acc, _i32p), _ptr(out, _i8p), acc.size, mult, shift, add, mode, eff)
        return out


def build(native: bool = False) -> str:
    target = "libita_ref_native.so" if native else "libita_ref.so"
    subprocess.run(["make", "-s", "-C", _HERE, "native" if native else "all"], check=True)
    return os.path.join(_HERE, target)


_cache: Dict[bool, RefLib] = {}


def load(native: bool = False) -> RefLib:
    if native not in _cache:
        path = os.path.join(_HERE, "libita_ref_native.so" if native else "libita_ref.so")
        try:
            if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "ita_ref.c")):
                path = build(native)
            _cache[native] = RefLib(path)
        except OSError:
            _cache[native] = RefLib(build(native))
    return _cache[native]
