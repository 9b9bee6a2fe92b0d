"""GPU test of ita_encoder_layer_forward_static, the encoder-layer entry point for callers that cannot allocate.

Who those callers are: the per-packet inference loop on the drone (a cudaMalloc or cudaFree inside it stalls the whole GPU
and lands in the control period), a host runtime that has already handed out all device memory from its own pool, and code
that runs inside another component's stream-ordered region.  For them the layer's intermediates live in a fixed 16 MiB
array inside libita_b200.so (include/ita_b200.h, csrc/layer_static.cu) and the call performs no cudaMalloc / cudaFree.

What is checked, through the C-ABI (ctypes), against the allocating entry point ita_encoder_layer_forward:
  * two calls in a row (a full-size batch, then a smaller batch of other data, so the array still holds the first call's
    bytes) each give the same output as the existing entry point.  Both run the same five kernels on the same values, so
    the tolerance is zero: the f32 outputs must be equal bit for bit;
  * the driver's free-memory figure does not drop across a call;
  * a batch beyond the array, and a handle whose kernels need allocated scratch, are refused with a clear message.

This file is named test_zz_* so that pytest collects it last: `pytest -x` stops at the first failure, and a failure here
must not hide the parity tests.  It passes alone in a fresh process and after the other test files (the array is shared by
the whole process, but every intermediate in it is rewritten before it is read).
"""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import GOLDEN, to_engine_params

import sys
sys.path.insert(0, GOLDEN)

torch = pytest.importorskip("torch")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")]


def _layer_fixture(ita):
    import seeded
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    p = ita.params_from_captured_block(q, "b0_", with_ffn=True)
    eng = ita.ITAEngine(p, torch.device("cuda:0"))
    ln = [seeded.float_param(k, (64,)) for k in ("norms1.0.weight", "norms1.0.bias", "norms2.0.weight", "norms2.0.bias")]
    lp = eng.layer_params(float(q["b0_s_x"]), float(q["b0_out_proj_sout"]), float(q["b0_fc1_sin"]), float(q["b0_fc2_sout"]), *ln)
    return q, eng, lp


def _call_static(eng, lp, x, y):
    """The raw C call, the way a C host would make it: caller-owned input and output, current stream."""
    stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    return eng.lib.ita_encoder_layer_forward_static(eng._h, C.byref(lp), C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                    int(x.shape[0]), stream)


def test_static_layer_equals_allocating_layer_on_two_calls_in_a_row(ita):
    q, eng, lp = _layer_fixture(ita)
    assert eng.fused_supported and eng.ffn_fused_supported
    dev = torch.device("cuda:0")
    max_b = eng.layer_static_max_batch
    assert max_b == ita._lib.LAYER_STATIC_SCRATCH_BYTES // (8 * 128 * 64) == 256
    g = torch.Generator().manual_seed(7)
    amp = float(np.abs(q["attn_in"]).max())
    # call 1 fills the whole array; call 2 is smaller and of other data, so stale bytes of call 1 lie behind and inside it
    xs = [(torch.rand((max_b, 128, 64), generator=g) * 2 - 1) * amp, torch.cat([torch.from_numpy(q["attn_in"]),
                                                                                  (torch.rand((35, 128, 64), generator=g) * 2 - 1) * amp])]
    xs = [x.to(dev).contiguous() for x in xs]
    outs = [torch.full_like(x, float("nan")) for x in xs]            # everything allocated before the calls
    want = [eng.layer(x, lp) for x in xs]                            # the existing entry point (allocates its scratch)
    eng.sync_check()
    free_before = None
    for i, (x, y) in enumerate(zip(xs, outs)):
        if i == 1:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
        rc = _call_static(eng, lp, x, y)
        assert rc == 0, eng.lib.ita_last_error().decode()
        eng.sync_check()
        if i == 1:
            assert torch.cuda.mem_get_info(dev)[0] >= free_before, "device memory was consumed by a call that must not allocate"
        got, ref = y.cpu().numpy(), want[i].cpu().numpy()
        assert np.isfinite(got).all()
        assert np.array_equal(got, ref), (i, float(np.abs(got - ref).max()))
    # and the reference-generated trace, with the tolerance test_encoder_layer_vs_model_fixture uses for the float LayerNorm
    n = q["attn_in"].shape[0]
    d = np.abs(outs[1][:n].cpu().numpy() - q["norm2_out"])
    assert (d > 1e-4).mean() <= 0.01 and d.max() <= 0.25, ((d > 1e-4).mean(), d.max())
    # the wrapper gives the same bytes and allocates no tensor of its own
    y2 = torch.empty_like(xs[1])
    assert eng.layer_static(xs[1], lp, y2) is y2
    eng.sync_check()
    assert torch.equal(y2, want[1])
    eng.close()


def test_static_layer_refuses_what_it_cannot_hold(ita, R):
    q, eng, lp = _layer_fixture(ita)
    dev = torch.device("cuda:0")
    x = torch.zeros((eng.layer_static_max_batch + 1, 128, 64), dtype=torch.float32, device=dev)
    y = torch.full_like(x, 3.0)
    assert _call_static(eng, lp, x, y) == 1                          # ITA_ERR_INVALID
    msg = eng.lib.ita_last_error().decode()
    assert "at most 256 frames" in msg and "ita_encoder_layer_forward" in msg, msg
    torch.cuda.synchronize()
    assert bool((y == 3.0).all())                                    # nothing was launched
    assert _call_static(eng, lp, x[:0], y[:0]) == 0                  # an empty batch is a no-op
    with pytest.raises(ita.ITAError, match="static scratch holds"):
        eng.layer_static(x, lp, y)
    eng.close()
    # a multi-head handle runs kernels that keep an allocated context buffer: not available without allocation
    p = R.make_synthetic(256, 128, 64, 4, seed=0, head_mode=1, F=256)
    mh = ita.ITAEngine(to_engine_params(ita, p), dev)
    ones = [np.ones(128, np.float32), np.zeros(128, np.float32)] * 2
    lp4 = mh.layer_params(0.05, 0.05, 0.05, 0.05, *ones)
    x4 = torch.zeros((2, 256, 128), dtype=torch.float32, device=dev)
    assert _call_static(mh, lp4, x4, torch.empty_like(x4)) == 2      # ITA_ERR_UNSUPPORTED
    assert "allocat" in mh.lib.ita_last_error().decode()
    mh.close()
