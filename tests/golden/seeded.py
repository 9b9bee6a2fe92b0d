"""Deterministic float parameters for the model-level fixture (tests/golden/qat_model_E64.npz).

The full ITAViTLSTM has ~3 M float parameters (decoder 4608x512, LSTM) -- too many to commit as a fixture.  They are
instead generated from their NAME with a seeded NumPy generator, identically by ``make_golden.py`` (build container, where
the reference model is importable) and by the tests (GPU box, where it is not).  torch-like scales: weights uniform in
+-1/sqrt(fan_in), biases small, LayerNorm gains near one."""
import zlib

import numpy as np


def float_param(name: str, shape) -> np.ndarray:
    shape = tuple(int(s) for s in shape)
    rng = np.random.default_rng(zlib.crc32(name.encode()))
    is_norm = ".norm." in name or name.startswith("norms")
    if name.endswith("weight") and is_norm:
        return (1.0 + 0.1 * rng.standard_normal(shape)).astype(np.float32)
    if "bias" in name.rsplit(".", 1)[-1]:
        return (0.05 * rng.standard_normal(shape)).astype(np.float32)
    fan_in = int(np.prod(shape[1:])) if len(shape) > 1 else shape[0]
    return (rng.uniform(-1.0, 1.0, size=shape) / np.sqrt(fan_in)).astype(np.float32)


def is_float_param(name: str) -> bool:
    """Parameters outside the int8 blocks (those are captured quantised in the fixture itself)."""
    return not (name.startswith("attention_blocks.") or name.startswith("ffn_blocks."))
