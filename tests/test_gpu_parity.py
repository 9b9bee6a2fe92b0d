"""GPU parity tests (run on the B200 box with ``-m gpu``): every CUDA path of libita_b200.so, called through the C-ABI
(ctypes -> ``ITAEngine``), against the CPU oracle on identical seeded inputs, against the committed golden fixtures that
were generated from the importable reference, and -- at BASELINE.json's full sizes -- through size-independent
properties (fused == staged, batch-permutation equivariance, duplicate frames).

Bar: bit-exact (integer path).  Nothing here reads /root/reference.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, to_engine_params

import sys
sys.path.insert(0, GOLDEN)

torch = pytest.importorskip("torch")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")]

STAGES = ("Qp", "Kp", "Vp", "A", "Asm", "O", "Y")


def _dev():
    return torch.device("cuda", 0)


def _engine(ita, p):
    return ita.ITAEngine(to_engine_params(ita, p), _dev())


def _cmp(got, want, what):
    got = got.cpu().numpy() if hasattr(got, "cpu") else got
    want = want.reshape(got.shape)
    bad = int((got != want).sum())
    assert bad == 0, f"{what}: {bad}/{got.size} mismatches, first at {np.argwhere(got != want)[:3].tolist()}"


# ----------------------------------------------------------------------------------------------------- building blocks
def test_tcgen05_selftest(ita):
    """Descriptor geometry (K-major / MN-major int8 operands, u8 x s8), TMEM load mapping and 2-D TMA panel loads."""
    rep = ita.selftest(0)
    assert "FAIL" not in rep, rep


# ----------------------------------------------------------------------------------------------------- staged kernels
@pytest.mark.parametrize("dims", [(64, 64, 64, 1), (128, 64, 192, 1), (128, 128, 192, 1), (256, 128, 64, 4), (32, 32, 32, 2),
                                  (48, 40, 24, 3)])
@pytest.mark.parametrize("round_mode", [0, 1, 2])
@pytest.mark.parametrize("softmax_mode", [0, 1])
def test_staged_all_stages_vs_oracle(ita, R, dims, round_mode, softmax_mode):
    S, E, P, H = dims
    for head_mode in ((0, 1) if H > 1 else (0,)):
        p = R.make_synthetic(S, E, P, H, seed=11, rand_add=True, round_mode=round_mode, softmax_mode=softmax_mode,
                             head_mode=head_mode)
        X = R.make_input(3, S, E, seed=21)
        ref = R.mha_forward(p, X)
        eng = _engine(ita, p)
        st = eng.forward_stages(torch.from_numpy(X).to(_dev()))
        for k in STAGES + ("Out",):
            _cmp(st[k], ref[k], f"staged {k} dims={dims} round={round_mode} softmax={softmax_mode} heads={head_mode}")
        y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_STAGED)
        _cmp(y, ref["Y"], "staged forward Y")
        eng.close()


# ----------------------------------------------------------------------------------------------------- fused kernel
FUSED_SHAPES = [(64, 192), (64, 64), (64, 128), (128, 64), (128, 192), (128, 128)]


@pytest.mark.parametrize("EP", FUSED_SHAPES)
@pytest.mark.parametrize("round_mode", [0, 1, 2])
@pytest.mark.parametrize("softmax_mode", [0, 1])
def test_fused_every_intermediate_vs_oracle(ita, R, EP, round_mode, softmax_mode):
    """The fused tcgen05 kernel with its on-chip intermediates dumped: every stage bit-exact, ragged batch (7 frames)."""
    E, P = EP
    p = R.make_synthetic(128, E, P, 1, seed=5, rand_add=True, round_mode=round_mode, softmax_mode=softmax_mode)
    eng = _engine(ita, p)
    if not eng.fused_supported:
        pytest.skip("fused kernel not instantiated for this shape: " + eng.fused_unsupported_reason())
    X = R.make_input(7, 128, E, seed=31)
    ref = R.mha_forward(p, X)
    st = eng.forward_stages(torch.from_numpy(X).to(_dev()), fused_debug=True)
    for k in STAGES:
        _cmp(st[k], ref[k], f"fused {k} E={E} P={P} round={round_mode} softmax={softmax_mode}")
    y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
    eng.sync_check()
    _cmp(y, ref["Y"], "fused forward Y")
    eng.close()


@pytest.mark.parametrize("EP", [(64, 192), (64, 64)])
@pytest.mark.parametrize("fp", ["1", "0"])
def test_fused_torch_form_zero_zero_points(ita, R, EP, fp, monkeypatch):
    """fp32 requant form (round_mode torch_rne_f32) with the reference's zero zero-points: the fused kernel then runs it on
    offset accumulators with three FFMAs per element (RQM_TORCH_FP; ITA_FUSED_FP=0: the I2F/FMUL/F2I form).  Every
    intermediate (dump kernel) and the production kernel's output bit-exact vs the oracle, incl. saturating frames."""
    monkeypatch.setenv("ITA_FUSED_FP", fp)
    E, P = EP
    p = R.make_synthetic(128, E, P, 1, seed=11, rand_add=False, round_mode=2, softmax_mode=0)
    eng = _engine(ita, p)
    if not eng.fused_supported:
        pytest.skip(eng.fused_unsupported_reason())
    X = R.make_input(9, 128, E, seed=12)
    X[0] = 127
    X[1] = -128
    ref = R.mha_forward(p, X)
    st = eng.forward_stages(torch.from_numpy(X).to(_dev()), fused_debug=True)
    for k in STAGES:
        _cmp(st[k], ref[k], f"fused torch form {k} E={E} P={P} ITA_FUSED_FP={fp}")
    y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
    eng.sync_check()
    _cmp(y, ref["Y"], f"fused torch form forward Y ITA_FUSED_FP={fp}")
    eng.close()


@pytest.mark.parametrize("batch", [0, 1, 2, 3, 295, 296, 297, 1000])
def test_fused_batch_sizes(ita, R, RC, batch):
    """Empty, single, odd, exactly one wave (2 x 148), one more, and multi-wave batches."""
    p = R.make_synthetic(128, 64, 192, 1, seed=0)
    eng = _engine(ita, p)
    assert eng.fused_supported, eng.fused_unsupported_reason()
    X = R.make_input(batch, 128, 64, seed=batch)
    y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
    eng.sync_check()
    assert tuple(y.shape) == (batch, 128, 64)
    if batch:
        _cmp(y, RC.mha(p, X, intermediates=False)["Y"], f"fused batch={batch}")
    eng.close()


def test_fused_general_requant_ranges(ita, R, RC):
    """Parameter sets outside the fast `hi32` requant form (small shifts, large multipliers) take the general form."""
    p = R.make_synthetic(128, 64, 192, 1, seed=9)
    p.mult = p.mult * 2
    p.shift = p.shift + 1          # same rational, mult > 255 => general path
    p.mult[R.ST_QK] = 3
    p.shift[R.ST_QK] = 5
    eng = _engine(ita, p)
    X = R.make_input(5, 128, 64, seed=2)
    ref = RC.mha(p, X)
    if eng.fused_supported:
        st = eng.forward_stages(torch.from_numpy(X).to(_dev()), fused_debug=True)
        for k in STAGES:
            _cmp(st[k], ref[k], f"fused(general requant) {k}")
    st = eng.forward_stages(torch.from_numpy(X).to(_dev()))
    for k in STAGES:
        _cmp(st[k], ref[k], f"staged(general requant) {k}")
    eng.close()


def test_edge_inputs(ita, R, RC):
    """Saturating inputs, all-equal logit rows, one-hot rows: all-zero / all-extreme frames and weights."""
    p = R.make_synthetic(128, 64, 192, 1, seed=4)
    X = np.zeros((6, 128, 64), np.int8)
    X[1] = 127
    X[2] = -128
    X[3, ::2] = 127
    X[3, 1::2] = -128
    X[4, 0] = 127                      # one token dominates
    X[5] = R.make_input(1, 128, 64, seed=8)[0]
    eng = _engine(ita, p)
    ref = RC.mha(p, X)
    for fused in ((True, False) if eng.fused_supported else (False,)):
        st = eng.forward_stages(torch.from_numpy(X).to(_dev()), fused_debug=fused)
        for k in STAGES:
            _cmp(st[k], ref[k], f"edge inputs {'fused' if fused else 'staged'} {k}")
    eng.close()
    # extreme weights: every accumulator at its bound, requant saturates both ways
    p2 = R.make_synthetic(128, 64, 192, 1, seed=4)
    p2.Wq[:] = 127
    p2.Wk[:] = -128
    p2.Wv[:, ::2] = 127
    eng = _engine(ita, p2)
    ref = RC.mha(p2, X)
    for fused in ((True, False) if eng.fused_supported else (False,)):
        st = eng.forward_stages(torch.from_numpy(X).to(_dev()), fused_debug=fused)
        for k in STAGES:
            _cmp(st[k], ref[k], f"extreme weights {'fused' if fused else 'staged'} {k}")
    eng.close()


@pytest.mark.parametrize("fp", ["1", "0"])
def test_fused_requant_forms_at_accumulator_bounds(ita, R, RC, fp, monkeypatch):
    """The production (non-dump) fused kernel in both requant forms -- FFMA.RM on offset accumulators (ITA_FUSED_FP=1,
    default) and the integer IMAD forms (ITA_FUSED_FP=0) -- on inputs and weights that drive every accumulator to its
    bound in both directions, plus ordinary frames: output bit-exact vs the oracle."""
    monkeypatch.setenv("ITA_FUSED_FP", fp)
    X = np.zeros((8, 128, 64), np.int8)
    X[1] = 127
    X[2] = -128
    X[3, ::2] = 127
    X[3, 1::2] = -128
    X[4, :, ::2] = -128
    X[4, :, 1::2] = 127
    X[5:] = R.make_input(3, 128, 64, seed=9)
    for case in range(4):
        p = R.make_synthetic(128, 64, 192, 1, seed=6 + case, rand_add=(case % 2 == 1))
        if case >= 1:
            p.Wq[:] = 127 if case != 2 else -128
            p.Wk[:] = -128
            p.Wv[:, ::2] = 127
            p.Wv[:, 1::2] = -128
        if case == 3:
            p.Wo[:] = 127
        eng = _engine(ita, p)
        if not eng.fused_supported:
            pytest.skip(eng.fused_unsupported_reason())
        ref = RC.mha(p, X)
        y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
        eng.sync_check()
        _cmp(y, ref["Y"], f"fused forward at accumulator bounds, ITA_FUSED_FP={fp}, case {case}")
        eng.close()


# ----------------------------------------------------------------------------------------------------- golden fixtures
def _fixture_params(R, q, pre, round_mode):
    from test_oracle import _block_params
    return _block_params(R, q, pre, round_mode)


@pytest.mark.parametrize("variant,blocks", [("E64", 1), ("E128", 2)])
def test_reference_fixture_bit_exact_on_gpu(ita, R, variant, blocks):
    """The converted reference block captured in tests/golden/qat_block_*.npz (generated by importing the reference,
    tests/golden/make_golden.py): in torch-rounding mode the CUDA paths reproduce EVERY stage bit-exactly."""
    from test_oracle import STAGE_FIX
    q = np.load(os.path.join(GOLDEN, f"qat_block_{variant}.npz"))
    for bi in range(blocks):
        pre = f"b{bi}_"
        p = _fixture_params(R, q, pre, R.ROUND_TORCH_RNE_F32)
        eng = _engine(ita, p)
        x = torch.from_numpy(q[pre + "x"]).to(_dev())
        for fused in ((True, False) if eng.fused_supported else (False,)):
            st = eng.forward_stages(x, fused_debug=fused)
            for k, fk in STAGE_FIX.items():
                _cmp(st[k], q[pre + fk], f"fixture {variant} block {bi} {'fused' if fused else 'staged'} {k}")
        eng.close()


def test_reference_multi_head_fixture_bit_exact_on_gpu(ita, R):
    """BASELINE config 3 pinned to the reference: the converted ``ITASelfAttention_QAT(128, 256, 4)`` on 256 tokens captured in
    tests/golden/qat_block_H4_S256.npz (generated by importing the reference).  The multi-head tcgen05 kernels (k_ita_mh +
    k_ita_mh_out, torch head convention) and the staged kernels reproduce EVERY stage of the reference bit-exactly, and the
    module built from the reference's converted state_dict returns the reference's float output."""
    from test_oracle import STAGE_FIX, _heads_fixture_stage
    q = np.load(os.path.join(GOLDEN, "qat_block_H4_S256.npz"))
    pre, H = "b0_", int(q["H"])
    eng = ita.ITAEngine(ita.params_from_captured_block(q, pre), _dev())
    assert eng.fused_supported and "k_ita_mh" in eng.describe(), eng.describe()
    x = torch.from_numpy(q[pre + "x"]).to(_dev())
    for fused in (True, False):
        st = eng.forward_stages(x, fused_debug=fused)
        for k, fk in STAGE_FIX.items():
            _cmp(st[k], _heads_fixture_stage(q, pre, fk, H), f"H=4 fixture {'tcgen05' if fused else 'staged'} {k}")
    _cmp(eng.forward(x), q[pre + "out_proj_out"], "H=4 fixture forward")
    eng.sync_check()
    eng.close()
    m = ita.ITASelfAttention(128, 256, 4)
    m.load_state_dict(torch.load(os.path.join(GOLDEN, "qat_state_H4_S256.pt"), weights_only=False), strict=False)
    y = m(torch.from_numpy(q["x_float"]).to(_dev()))
    assert np.array_equal(y.cpu().numpy(), q["y_float"])


def test_itamax_fixture_on_gpu(ita, R):
    """Reference IntegerApproximatedSoftmaxFn vectors through the CUDA ITAmax step (ita_mha_step STEP4_SOFTMAX)."""
    g = np.load(os.path.join(GOLDEN, "itamax_repo.npz"))
    for k in g.files:
        if not k.startswith("x"):
            continue
        x = g[k]
        rows, S = x.shape
        p = R.make_synthetic(S, 32, 32, 1, seed=0)
        eng = _engine(ita, p)
        # rows are independent: present them as `rows` frames' first logit row, padded to [rows,1,S,S]
        a = np.zeros((rows, 1, S, S), np.int8)
        a[:, 0, 0, :] = x
        out = torch.empty((rows, 1, S, S), dtype=torch.uint8, device=_dev())
        eng.step(ita._lib.STEP4_SOFTMAX, torch.from_numpy(a).to(_dev()), None, out, rows)
        _cmp(out[:, 0, 0, :], g["y" + k[1:]], f"itamax fixture {k}")
        eng.close()


@pytest.mark.parametrize("fixture", ["qat_block_E64.npz", "qat_block_E128.npz", "qat_model_E64.npz"])
def test_ffn_fixture_on_gpu(ita, R, fixture):
    """The converted reference FFN block (fc1 -> ReLU -> fc2, layers.py:61-75) through ``ita_ffn_forward``: the CUDA result
    equals the reference's captured ``relu_out`` / ``fc2_out`` bit-exactly (torch_rne_f32, requant scale formed in float32
    like the backend)."""
    from test_oracle import _ffn_params
    q = np.load(os.path.join(GOLDEN, fixture))
    p = _ffn_params(R, q)
    eng = _engine(ita, p)
    assert eng.ffn_fused_supported, eng.ffn_fused_unsupported_reason()
    x = torch.from_numpy(q["b0_ffn_x"]).to(_dev())
    for path in (ita.PATH_FUSED, ita.PATH_STAGED):
        f1, f2 = eng.ffn(x, want_f1=True, path=path)
        eng.sync_check()
        _cmp(f1, q["b0_relu_out"], f"ffn fixture {fixture} relu_out path={path}")
        _cmp(f2, q["b0_fc2_out"], f"ffn fixture {fixture} fc2_out path={path}")
    _cmp(eng.ffn(x, path=ita.PATH_FUSED), q["b0_fc2_out"], f"ffn fixture {fixture} fused, F1 on chip only")
    eng.close()


@pytest.mark.parametrize("E", [64, 128])
@pytest.mark.parametrize("fp", ["2", "1", "0"])
def test_ffn_vs_oracle_all_roundings(ita, R, RC, E, fp, monkeypatch):
    """Fused tcgen05 FFN kernel and the staged kernels vs the oracle: 3 roundings, random zero-points, every requant form
    (ITA_FUSED_FP = 2: FFMA2 forms, 1: three-FFMA torch form, 0: integer / conversion forms), ragged batches."""
    monkeypatch.setenv("ITA_FUSED_FP", fp)
    for rm in (0, 1, 2):
        p = R.make_synthetic(128, E, 64, 1, seed=1, F=256, rand_add=(rm != 2), round_mode=rm)
        p.mult[R.ST_RELU] = 0
        eng = _engine(ita, p)
        assert eng.ffn_fused_supported, eng.ffn_fused_unsupported_reason()
        for B in (1, 7, 600):
            X = R.make_input(B, 128, E, seed=3 + B)
            X[0] = 127
            ref = RC.ffn(p, X)
            for path in (ita.PATH_FUSED, ita.PATH_STAGED):
                f1, f2 = eng.ffn(torch.from_numpy(X).to(_dev()), want_f1=True, path=path)
                eng.sync_check()
                _cmp(f1, ref["FF1"], f"ffn FF1 E={E} round={rm} B={B} path={path} fp={fp}")
                _cmp(f2, ref["FF2"], f"ffn FF2 E={E} round={rm} B={B} path={path} fp={fp}")
            _cmp(eng.ffn(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED), ref["FF2"], "ffn fused without the F1 copy")
        eng.close()


def test_ffn_relu_requant_routes_to_staged(ita, R, RC):
    """A non-identity ReLU requant (golden-model option, W_B.py:278-301) is outside the fused FFN kernel: AUTO takes the
    staged kernels and stays bit-exact."""
    p = R.make_synthetic(128, 64, 64, 1, seed=1, F=256)
    assert int(p.mult[R.ST_RELU, 0]) != 0
    eng = _engine(ita, p)
    X = R.make_input(3, 128, 64, seed=2)
    _cmp(eng.ffn(torch.from_numpy(X).to(_dev())), RC.ffn(p, X)["FF2"], "ffn with relu requant (auto path)")
    eng.close()


def test_encoder_layer_vs_model_fixture(ita, R):
    """One encoder layer f32 -> f32 (quant, int8 attention, dequant + residual + LayerNorm, quant, int8 FFN, dequant +
    residual + LayerNorm; reference QAT/model.py:100-113) against the full converted-model trace.  The int8 stages are
    bit-exact; the float LayerNorm is compared with atol 2e-5 (different summation order than ATen's); a LayerNorm output
    within that distance of a rounding boundary may quantise one LSB apart, which the final comparison allows for on
    at most 0.2 % of the elements."""
    import seeded
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    p = ita.params_from_captured_block(q, "b0_", with_ffn=True)
    eng = ita.ITAEngine(p, _dev())
    assert eng.fused_supported and eng.ffn_fused_supported
    ln = {k: seeded.float_param(k, (64,)) for k in ("norms1.0.weight", "norms1.0.bias", "norms2.0.weight", "norms2.0.bias")}
    x = torch.from_numpy(q["attn_in"]).to(_dev())
    # stage by stage
    yq = eng.forward(ita.quantize(x, float(q["b0_s_x"])))
    x1, x1q = ita.residual_layernorm(x, yq, float(q["b0_out_proj_sout"]), torch.from_numpy(ln["norms1.0.weight"]),
                                     torch.from_numpy(ln["norms1.0.bias"]), 1e-5, quant_scale=float(q["b0_fc1_sin"]))
    assert np.abs(x1.cpu().numpy() - q["norm1_out"]).max() <= 2e-5
    bad = int((x1q.cpu().numpy() != q["b0_ffn_x"]).sum())
    assert bad <= 0.002 * x1q.numel(), bad
    # LayerNorm on the reference's own inputs: residual add is bit-exact, so feed the captured pre-norm tensor through q = 0
    z = torch.zeros_like(yq)
    y1 = ita.residual_layernorm(torch.from_numpy(q["norm1_in"]).to(_dev()), z, 1.0, torch.from_numpy(ln["norms1.0.weight"]),
                                torch.from_numpy(ln["norms1.0.bias"]))
    assert np.abs(y1.cpu().numpy() - q["norm1_out"]).max() <= 2e-5
    # the layer as one call
    lp = eng.layer_params(float(q["b0_s_x"]), float(q["b0_out_proj_sout"]), float(q["b0_fc1_sin"]), float(q["b0_fc2_sout"]),
                          ln["norms1.0.weight"], ln["norms1.0.bias"], ln["norms2.0.weight"], ln["norms2.0.bias"])
    y = eng.layer(x, lp).cpu().numpy()
    eng.sync_check()
    d = np.abs(y - q["norm2_out"])
    assert (d > 1e-4).mean() <= 0.01 and d.max() <= 0.25, ((d > 1e-4).mean(), d.max())
    eng.close()


def test_quantize_fixture_on_gpu(ita):
    """QuantStub (layers.py:103): the CUDA quantiser equals torch.quantize_per_tensor on the reference-generated vectors
    (random floats, exact ties, saturation, awkward scales); dequantise is scale * q exactly."""
    g = np.load(os.path.join(GOLDEN, "quantize.npz"))
    for s, x, q in zip(g["scales"], g["x"], g["q"]):
        got = ita.quantize(torch.from_numpy(x).to(_dev()), float(s))
        _cmp(got, q, f"quantize scale={float(s)}")
        back = ita.dequantize(got, float(s)).cpu().numpy()
        assert np.array_equal(back, q.astype(np.float32) * np.float32(s))


def test_model_fixture_block_on_gpu(ita, R):
    """The attention block inside the full converted model trace (qat_model_E64.npz): f32 in -> f32 out through the module
    surface equals the reference's captured block output bit-exactly, every int8 stage included."""
    from test_oracle import STAGE_FIX
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    p = ita.params_from_captured_block(q, "b0_")
    eng = ita.ITAEngine(p, _dev())
    assert eng.fused_supported, eng.fused_unsupported_reason()
    xq = ita.quantize(torch.from_numpy(q["attn_in"]).to(_dev()), float(q["b0_s_x"]))
    _cmp(xq, q["b0_x"], "model fixture: quant stub")
    for fused in (True, False):
        st = eng.forward_stages(xq, fused_debug=fused)
        for k, fk in STAGE_FIX.items():
            _cmp(st[k], q["b0_" + fk], f"model fixture {'fused' if fused else 'staged'} {k}")
    y = ita.dequantize(eng.forward(xq, path=ita.PATH_FUSED), float(q["b0_out_proj_sout"]))
    assert np.array_equal(y.cpu().numpy(), q["attn_out"])
    eng.close()


# ----------------------------------------------------------------------------------------------------- host surfaces
def test_forward_host_and_chunked_pipeline(ita, R, RC):
    p = R.make_synthetic(128, 64, 192, 1, seed=6)
    eng = _engine(ita, p)
    eng2 = _engine(ita, p)
    for B in (3, 1500, 4100):                # >= 1024: 4-chunk copy/compute pipeline, >= 4096: 8 chunks (ragged last chunk)
        X = R.make_input(B, 128, 64, seed=B)
        want = RC.mha(p, X, intermediates=False)["Y"]
        y = eng.forward_host(X)
        _cmp(y, want, f"forward_host pageable B={B}")
        xp = torch.from_numpy(X).pin_memory()
        yp = torch.empty_like(xp).pin_memory()
        eng.forward_host(xp, out=yp)
        _cmp(yp.numpy(), want, f"forward_host pinned B={B}")
        # asynchronous entry: two handles on two streams, results valid after sync_check
        outs = [torch.zeros_like(xp).pin_memory() for _ in range(2)]
        streams = [torch.cuda.Stream(_dev()) for _ in range(2)]
        for i, e in enumerate((eng, eng2)):
            e.forward_host(xp, out=outs[i], stream=streams[i], sync=False)
        for i, e in enumerate((eng, eng2)):
            e.sync_check(streams[i])
            _cmp(outs[i].numpy(), want, f"forward_host async handle {i} B={B}")
    eng.close()
    eng2.close()


def test_golden_facade_steps(ita, R):
    """`Transformer` facade driven exactly like the reference harness (W_B.py:518-542), incl. overwriting
    A_partial_softmax before step5_AV (:531-540)."""
    for (S, E, P, H, F) in [(64, 64, 64, 1, 64), (128, 64, 192, 1, 256), (64, 64, 32, 2, 0)]:
        p = R.make_synthetic(S, E, P, H, seed=13, F=F)
        X = R.make_input(1, S, E, seed=14)[0]
        names = ["q_proj", "k_proj", "v_proj", "qk_matmul", "av_matmul", "out_proj", "head_sum", "fc1", "relu", "fc2"]
        qp = {n: {"mult": p.mult[i].tolist(), "shift": p.shift[i].tolist(), "add": p.add[i].tolist()} for i, n in enumerate(names)}
        t = ita.Transformer(ITA_N=16, path=None, activation="relu", S=S, P=P, E=E, F=F, H=H, Q=X, K=X, V=X, Wq=p.Wq, Wk=p.Wk,
                            Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv, Bo=p.Bo, FF_in=X, Wff=p.Wff, Wff2=p.Wff2, Bff=p.Bff,
                            Bff2=p.Bff2, quant_params=qp)
        ref = R.mha_forward(p, X)
        t.step1_Qp(); t.step2_Kp(); t.step3_Vp(); t.step4_QK(no_partial_softmax=False)
        _cmp(t.Qp_requant, ref["Qp"], "facade Qp"); _cmp(t.Kp_requant, ref["Kp"], "facade Kp")
        _cmp(t.Vp_requant, ref["Vp"], "facade Vp"); _cmp(t.A_requant, ref["A"], "facade A")
        _cmp(t.A_partial_softmax, ref["Asm"], "facade Asm")
        t.step5_AV(); t.step6_O(); t.step7_Osum()
        _cmp(t.O_soft_requant, ref["O"], "facade O"); _cmp(t.Out_soft_requant, ref["Out"], "facade Out")
        _cmp(t.Out_sum_requant, ref["Y"], "facade Y")
        # inject a different softmax (the harness' --isolate_softmax pattern) and re-run step 5
        inj = np.roll(ref["Asm"][0], 1, axis=-1)
        t.A_partial_softmax = inj
        t.step5_AV()
        want = np.stack([p.rq(R._mm(inj[h], ref["Vp"][0, h]), R.ST_AV, h) for h in range(H)])
        _cmp(t.O_soft_requant, want, "facade injected softmax -> O")
        if F:
            t.feedforward_layer()
            rf = R.ffn_forward(p, X)
            _cmp(t.FFp_requant, rf["FF1"], "facade FF1"); _cmp(t.FF2p_requant, rf["FF2"], "facade FF2")


def test_module_surface(ita, R):
    """ITASelfAttention: same ctor / forward(f32[B,N,E]) -> f32[B,N,E] as ITASelfAttention_QAT (layers.py:79,101-127);
    quant / dequant stubs bit-exact vs their definition."""
    q = np.load(os.path.join(GOLDEN, "qat_block_E64.npz"))
    pre = "b0_"
    m = ita.ITASelfAttention(64, 192, 1)
    for name in ("q_proj", "k_proj", "v_proj", "out_proj"):
        lin = getattr(m, name)
        lin.weight_int8.copy_(torch.from_numpy(q[pre + name + "_w"]))
        lin.weight_scale = float(q[pre + name + "_sw"])
        lin.scale = float(q[pre + name + "_sout"])
        s_in = float(q[pre + name + "_sin"])
        lin.bias.copy_(torch.from_numpy(q[pre + name + "_bias_i32"].astype(np.float64) * s_in * float(lin.weight_scale)).float())
    m.quant.scale = float(q[pre + "q_proj_sin"])
    m.matmul1.scale = float(q[pre + "qk_sout"])
    m.matmul2.scale = float(q[pre + "av_sout"])
    xq = q[pre + "x"]
    x = torch.from_numpy(xq.astype(np.float32) * np.float32(float(m.quant.scale))).to(_dev())
    y = m(x)
    assert y.dtype == torch.float32 and tuple(y.shape) == tuple(x.shape)
    want = q[pre + "out_proj_out"].reshape(xq.shape).astype(np.float32) * np.float32(float(m.out_proj.scale))
    assert np.array_equal(y.cpu().numpy(), want)
    yq = m.forward_int8(torch.from_numpy(xq).to(_dev()))
    _cmp(yq, q[pre + "out_proj_out"], "module forward_int8 vs fixture")
    with pytest.raises(ValueError):
        m(torch.zeros(1, 128, 32, device=_dev()))
    with pytest.raises(ValueError):
        m(torch.zeros(1, 128, 64))            # CPU tensor: no CPU fallback
    # quantize: clamp(rne(x/s)); ties to even
    s = 0.5
    v = torch.tensor([0.25, 0.75, -0.25, -0.75, 1.25, 1000.0, -1000.0, 0.0], device=_dev())
    assert ita.quantize(v, s).cpu().tolist() == [0, 2, 0, -2, 2, 127, -128, 0]


def test_module_multi_head_runs_the_tcgen05_kernels(ita, R):
    """``ITASelfAttention(embed_dim=128, proj_dim=256, num_heads=4)`` on 256 tokens (the module form of BASELINE config 3: heads
    are column slices of one projection, contexts concatenated into one out_proj, reference layers.py:82,106-123) takes the
    multi-head tcgen05 kernels and equals the oracle's torch head convention bit-exactly."""
    import bench
    g = np.random.default_rng(17)
    m = ita.ITASelfAttention(128, 256, 4)
    for name, (n_out, n_in) in (("q_proj", (256, 128)), ("k_proj", (256, 128)), ("v_proj", (256, 128)), ("out_proj", (128, 256))):
        lin = getattr(m, name)
        lin.weight_int8.copy_(torch.from_numpy(np.clip(np.rint(g.normal(0, 32, (n_out, n_in))), -128, 127).astype(np.int8)))
        lin.weight_scale = 0.004
        lin.bias.copy_(torch.from_numpy(g.normal(0, 0.3, n_out).astype(np.float32)))
    m.quant.scale, m.q_proj.scale, m.k_proj.scale, m.v_proj.scale = 0.03, 0.06, 0.06, 0.05
    m.matmul1.scale, m.matmul2.scale, m.out_proj.scale = 0.9, 0.02, 0.05
    p = m.quant_params(256)
    assert p.H == 4 and p.P == 64 and p.head_mode == ita._lib.HEADS_TORCH
    eng = m.engine(256, _dev())
    assert eng.fused_supported and "k_ita_mh" in eng.describe(), eng.describe()
    X = R.make_input(3, 256, 128, seed=8)
    y = m.forward_int8(torch.from_numpy(X).to(_dev()))
    eng.sync_check()
    _cmp(y, R.mha_forward(bench.to_oracle(R, p), X)["Y"], "multi-head module vs oracle (torch head convention)")
    yf = m(torch.from_numpy(X.astype(np.float32) * np.float32(0.03)).to(_dev()))
    assert np.array_equal(yf.cpu().numpy(), y.cpu().numpy().astype(np.float32) * np.float32(0.05))


def test_module_hooks_fire_with_quantised_stage_tensors(ita, R):
    """Forward hooks on the named submodules (the way tests/export_and_validation_W_B.py:186-205 captures a block) receive
    quantised tensors with the reference's shapes, scales and integer values; registering hooks does not change the output."""
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    sd_ref = torch.load(os.path.join(GOLDEN, "qat_state_E64.pt"), weights_only=False)
    sub = {k[len("attention_blocks.0."):]: v for k, v in sd_ref.items() if k.startswith("attention_blocks.0.")}
    m = ita.ITASelfAttention(64, 192, 1)
    m.load_state_dict(sub, strict=False)
    x = torch.from_numpy(q["attn_in"]).to(_dev())
    y_plain = m(x)
    got = {}
    handles = [getattr(m, n).register_forward_hook(lambda mod, i, o, n=n: got.__setitem__(n, o))
               for n in ("quant", "q_proj", "k_proj", "v_proj", "matmul1", "custom_softmax", "matmul2", "out_proj", "dequant")]
    y_hook = m(x)
    for h in handles:
        h.remove()
    assert torch.equal(y_plain, y_hook)
    assert np.array_equal(y_plain.cpu().numpy(), q["attn_out"])
    want = {"quant": "b0_x", "q_proj": "b0_q_proj_out", "k_proj": "b0_k_proj_out", "v_proj": "b0_v_proj_out", "matmul1": "b0_qk_out",
            "custom_softmax": "b0_softmax_out", "matmul2": "b0_av_out", "out_proj": "b0_out_proj_out"}
    for name, key in want.items():
        t = got[name]
        assert t.is_quantized and tuple(t.shape) == tuple(q[key].shape), (name, tuple(t.shape), q[key].shape)
        assert np.array_equal(t.int_repr().numpy(), q[key]), name
    assert abs(got["q_proj"].q_scale() - float(q["b0_q_proj_sout"])) <= 1e-12 * float(q["b0_q_proj_sout"]) + 1e-15
    assert got["custom_softmax"].dtype == torch.quint8 and got["matmul1"].dtype == torch.qint8
    assert np.array_equal(got["dequant"].numpy(), q["attn_out"])
    assert not m._has_hooks()


def test_full_model_forward_vs_reference_trace(ita, R):
    """``ITAViTLSTM.forward`` ([img, velocity, quaternion] -> command, LSTM state; reference QAT/model.py:88-133) built from
    the converted reference checkpoint, against the trace of the reference model on the same inputs
    (tests/golden/make_golden.py --new).  Float stages run in torch on the GPU, the encoder layer in libita_b200.
    Tolerances: tokens 1e-5 (cuDNN conv vs the CPU conv of the trace); encoder output like test_encoder_layer (a token within
    float noise of a rounding boundary may quantise one LSB apart); command and LSTM state 5e-3 absolute."""
    import importlib
    import seeded
    M = importlib.import_module("drone-oa-iree-vit-accelerator_b200.model")
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    sd = dict(torch.load(os.path.join(GOLDEN, "qat_state_E64.pt"), weights_only=False))
    for name, shape in M.reference_float_shapes(1, 64).items():
        sd[name] = torch.from_numpy(seeded.float_param(name, shape))
    m = M.ITAViTLSTM.from_state_dict(sd).to(_dev())
    dev = lambda k: torch.from_numpy(q[k]).to(_dev())
    tok = m.tokens(dev("img"))                      # the CUDA tokenizer kernel
    assert np.abs(tok.cpu().numpy() - q["tokenizer_out"]).max() <= 2e-5, np.abs(tok.cpu().numpy() - q["tokenizer_out"]).max()
    assert np.abs(m.tokens_torch(dev("img")).cpu().numpy() - q["tokenizer_out"]).max() <= 2e-5
    enc = m.encode(dev("tokenizer_out")).cpu().numpy()
    d = np.abs(enc - q["norm2_out"])
    assert (d > 1e-4).mean() <= 0.01 and d.max() <= 0.25, ((d > 1e-4).mean(), d.max())
    out, (hn, cn) = m([dev("img"), dev("vel"), dev("quat")])
    assert tuple(out.shape) == (3, 3) and tuple(hn.shape) == (3, 3, 128)
    assert np.abs(out.cpu().numpy() - q["out"]).max() <= 5e-3, np.abs(out.cpu().numpy() - q["out"]).max()
    assert np.abs(hn.cpu().numpy() - q["hn"]).max() <= 5e-3 and np.abs(cn.cpu().numpy() - q["cn"]).max() <= 5e-3
    # second step: the recurrent state goes back in (the trace's second call flipped the batch)
    out2, (hn2, _) = m([torch.flip(dev("img"), dims=[0]), dev("vel"), dev("quat"), (hn, cn)])
    assert np.abs(out2.cpu().numpy() - q["out2"]).max() <= 5e-3 and np.abs(hn2.cpu().numpy() - q["hn2"]).max() <= 5e-3


@pytest.mark.parametrize("E", [64, 128])
@pytest.mark.parametrize("u8", [False, True])
def test_tokenizer_kernel_vs_torch_fp32(ita, E, u8):
    """``ita_tokenizer_forward`` (conv 7x7 s2 p3 + bilinear 30x45 -> 8x16 + LayerNorm, reference models/ITA/QAT/layers.py:25-45;
    uint8 frames scaled by 1/255 as main.cpp:168) against the same stage in plain PyTorch fp32 (TF32 off).  The kernel blends
    the four bilinear taps before the convolution (both are linear), so the sums are reordered: tolerance 2e-5 absolute on
    O(1) LayerNorm outputs.  The fused quant stub must equal quantising the kernel's own float tokens exactly, and sit within
    one LSB of the torch reference's on every element (a token within float noise of a rounding boundary may flip)."""
    import ctypes as C
    import torch.nn.functional as F
    L = ita._lib
    g = torch.Generator().manual_seed(5 + E)
    B = 37
    img8 = torch.randint(0, 256, (B, 60, 90), generator=g, dtype=torch.uint8)
    img8[0] = 0
    img8[1] = 255
    imgf = torch.rand((B, 60, 90), generator=g) * 3.0 - 1.0
    w = (torch.randn((E, 1, 7, 7), generator=g) * 0.2).to(_dev())
    b = (torch.randn((E,), generator=g) * 0.1).to(_dev())
    lw = (1.0 + 0.1 * torch.randn((E,), generator=g)).to(_dev())
    lb = (0.1 * torch.randn((E,), generator=g)).to(_dev())
    img = (img8 if u8 else imgf).to(_dev()).contiguous()
    x = (img.float() / 255.0 if u8 else img).unsqueeze(1)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        ref = F.conv2d(x, w, b, stride=2, padding=3)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert tuple(ref.shape) == (B, E, 30, 45)
    ref = F.interpolate(ref, size=(8, 16), mode="bilinear", align_corners=False)
    ref = F.layer_norm(ref.flatten(2).transpose(1, 2), (E,), lw, lb, 1e-5).contiguous()
    tok = torch.empty((B, 128, E), dtype=torch.float32, device=_dev())
    q = torch.empty((B, 128, E), dtype=torch.int8, device=_dev())
    s_q = 0.0371
    L.check(L.load().ita_tokenizer_forward(img.data_ptr(), int(u8), B, E, w.data_ptr(), b.data_ptr(), lw.data_ptr(), lb.data_ptr(), 1e-5,
                                           tok.data_ptr(), q.data_ptr(), s_q, torch.cuda.current_stream().cuda_stream), "ita_tokenizer_forward")
    torch.cuda.synchronize()
    d = (tok - ref).abs().max().item()
    assert d <= 2e-5, d
    inv = np.float32(1.0) / np.float32(s_q)
    own = np.clip(np.rint(tok.cpu().numpy() * inv), -128, 127).astype(np.int8)
    assert np.array_equal(q.cpu().numpy(), own)
    refq = np.clip(np.rint(ref.cpu().numpy() * inv), -128, 127).astype(np.int32)
    dq = np.abs(q.cpu().numpy().astype(np.int32) - refq)
    assert dq.max() <= 1 and (dq != 0).mean() <= 1e-3, (dq.max(), (dq != 0).mean())
    # float-only / int8-only outputs
    tok2 = torch.empty_like(tok)
    L.check(L.load().ita_tokenizer_forward(img.data_ptr(), int(u8), B, E, w.data_ptr(), b.data_ptr(), lw.data_ptr(), lb.data_ptr(), 1e-5,
                                           tok2.data_ptr(), None, 0.0, torch.cuda.current_stream().cuda_stream), "ita_tokenizer_forward")
    assert torch.equal(tok, tok2)


@pytest.mark.parametrize("E", [64, 128])
def test_fusion_conv_kernel_vs_torch_fp32(ita, E):
    """``ita_fusion_conv_forward`` (PixelShuffle(2) | bilinear Upsample(16x32, align_corners=True) -> cat -> Conv2d 3x3 padding 1,
    reference models/ITA_single_layer_upsample_shuffle/QAT/model.py:116-121) against the same ops in plain PyTorch fp32 (TF32
    off).  fp32 sums in a different order: tolerance 5e-5 absolute on O(1) outputs."""
    import torch.nn.functional as F
    L = ita._lib
    g = torch.Generator().manual_seed(9 + E)
    B = 19
    x = torch.randn((B, 128, E), generator=g).to(_dev())
    cin = E // 4 + E
    w = (torch.randn((9, cin, 3, 3), generator=g) * 0.05).to(_dev())
    b = (torch.randn((9,), generator=g) * 0.1).to(_dev())
    x2d = x.transpose(1, 2).reshape(B, E, 8, 16)
    fused = torch.cat([F.pixel_shuffle(x2d, 2), F.interpolate(x2d, size=(16, 32), mode="bilinear", align_corners=True)], dim=1)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        ref = F.conv2d(fused, w, b, padding=1).flatten(1)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    out = torch.empty((B, 9 * 16 * 32), dtype=torch.float32, device=_dev())
    L.check(L.load().ita_fusion_conv_forward(x.data_ptr(), B, E, w.data_ptr(), b.data_ptr(), out.data_ptr(),
                                             torch.cuda.current_stream().cuda_stream), "ita_fusion_conv_forward")
    torch.cuda.synchronize()
    d = (out - ref).abs().max().item()
    assert d <= 5e-5, d


def test_lstm_cell_kernel_vs_torch_lstm(ita):
    """``ita_lstm_cell_forward`` behind two cuBLAS gate GEMMs == one time step of ``torch.nn.LSTM`` (3 layers, 517 -> 128, the
    model's head, reference QAT/model.py:68,128) computed on the CPU in fp32; tolerance 1e-5 (GEMM summation order, expf / tanhf vs
    ATen's kernels)."""
    L = ita._lib
    torch.manual_seed(3)
    B, I, Hd, NL = 53, 517, 128, 3
    lstm = torch.nn.LSTM(input_size=I, hidden_size=Hd, num_layers=NL).to(_dev()).eval()
    x = torch.randn(1, B, I, device=_dev())
    h0, c0 = torch.randn(NL, B, Hd, device=_dev()) * 0.5, torch.randn(NL, B, Hd, device=_dev()) * 0.5
    with torch.no_grad():
        # reference on the CPU in fp32 (cuDNN's RNN kernels use TF32 tensor ops by default: 1e-3 off)
        ref, (hr, cr) = lstm.cpu()(x.cpu(), (h0.cpu(), c0.cpu()))
        ref, hr, cr = ref.to(_dev()), hr.to(_dev()), cr.to(_dev())
        lstm.to(_dev())
        xs = x[0]
        hn, cn = torch.empty_like(h0), torch.empty_like(c0)
        for l in range(NL):
            gates = torch.addmm(getattr(lstm, f"bias_ih_l{l}") + getattr(lstm, f"bias_hh_l{l}"), xs, getattr(lstm, f"weight_ih_l{l}").t())
            gates.addmm_(h0[l], getattr(lstm, f"weight_hh_l{l}").t())
            L.check(L.load().ita_lstm_cell_forward(gates.data_ptr(), c0[l].contiguous().data_ptr(), hn[l].data_ptr(), cn[l].data_ptr(), B, Hd,
                                                   torch.cuda.current_stream().cuda_stream), "ita_lstm_cell_forward")
            xs = hn[l]
    torch.cuda.synchronize()
    assert (hn - hr).abs().max().item() <= 1e-5 and (cn - cr).abs().max().item() <= 1e-5, ((hn - hr).abs().max().item(), (cn - cr).abs().max().item())
    assert (xs - ref[0]).abs().max().item() <= 1e-5


def test_graphed_model_step_equals_eager(ita, R):
    """``GraphedStep`` (one ITAViTLSTM step captured in a CUDA graph, LSTM state carried inside) reproduces the eager forward
    over three consecutive steps of a 5-drone batch, bit for bit (the graph replays the same kernels)."""
    import importlib
    import seeded
    M = importlib.import_module("drone-oa-iree-vit-accelerator_b200.model")
    sd = dict(torch.load(os.path.join(GOLDEN, "qat_state_E64.pt"), weights_only=False))
    for name, shape in M.reference_float_shapes(1, 64).items():
        sd[name] = torch.from_numpy(seeded.float_param(name, shape))
    m = M.ITAViTLSTM.from_state_dict(sd).to(_dev())
    g = torch.Generator().manual_seed(4)
    B = 5
    step = M.GraphedStep(m, B, _dev())
    hidden = None
    for k in range(3):
        img = torch.randint(0, 256, (B, 60, 90), generator=g, dtype=torch.uint8).to(_dev())
        vel = (torch.rand((B, 1), generator=g) * 5).to(_dev())
        quat = torch.rand((B, 4), generator=g).to(_dev())
        want, hidden = m([img, vel, quat] + ([hidden] if hidden is not None else []))
        got = step(img, vel, quat)
        torch.cuda.synchronize()
        assert torch.equal(got, want), (k, (got - want).abs().max().item())
    assert torch.equal(step.h, hidden[0]) and torch.equal(step.c, hidden[1])
    step.reset_state()
    assert float(step.h.abs().max()) == 0.0
    # An eager call at a LARGER batch on the same model grows (frees and reallocates) the scratch of the model's shared handles.
    # The graph replays on the step's private handles, so it must neither fault nor change its answer.
    big = 8 * B + 3
    m([torch.randint(0, 256, (big, 60, 90), generator=g, dtype=torch.uint8).to(_dev()), torch.rand((big, 1), generator=g).to(_dev()),
       torch.rand((big, 4), generator=g).to(_dev())])
    junk = torch.full((64 << 20,), 0x55, dtype=torch.uint8, device=_dev())       # recycle whatever the reallocation released
    img = torch.randint(0, 256, (B, 60, 90), generator=g, dtype=torch.uint8).to(_dev())
    vel, quat = (torch.rand((B, 1), generator=g) * 5).to(_dev()), torch.rand((B, 4), generator=g).to(_dev())
    want, _ = m([img, vel, quat])
    got = step(img, vel, quat)
    torch.cuda.synchronize()
    assert torch.equal(got, want), (got - want).abs().max().item()
    assert all(e is not s_ for (e, _), (s_, _) in zip(m._layer_engines(_dev()), step._engines))
    del junk
    step.close()


# ----------------------------------------------------------------------------------------------------- full-size properties
def test_full_size_properties(ita, R, RC):
    """BASELINE configs[1] size (4096 frames): fused == staged on every frame, permutation equivariance, duplicate
    frames give duplicate results, and a 256-frame sample equals the oracle."""
    p = R.make_synthetic(128, 64, 192, 1, seed=0)
    eng = _engine(ita, p)
    assert eng.fused_supported, eng.fused_unsupported_reason()
    B = 4096
    X = R.make_input(B, 128, 64, seed=77)
    X[17] = X[4000]
    x = torch.from_numpy(X).to(_dev())
    y = eng.forward(x, path=ita.PATH_FUSED)
    eng.sync_check()
    ys = eng.forward(x, path=ita.PATH_STAGED)
    assert torch.equal(y, ys)
    assert torch.equal(y[17], y[4000])
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(1)).to(_dev())
    yp = eng.forward(x[perm].contiguous(), path=ita.PATH_FUSED)
    eng.sync_check()
    assert torch.equal(yp, y[perm])
    idx = np.arange(0, B, 16)
    _cmp(y[torch.from_numpy(idx).to(_dev())], RC.mha(p, X[idx], intermediates=False)["Y"], "full-size sample vs oracle")
    eng.close()


def test_multi_head_full_config3_sample(ita, R, RC):
    """BASELINE configs[2] dims (H=4, S=256, E=128, P=64) on a 64-frame sample, both ITAmax variants, on the tcgen05
    multi-head kernels (k_ita_mh + k_ita_mh_out)."""
    for sm in (0, 1):
        p = R.make_synthetic(256, 128, 64, 4, seed=2, softmax_mode=sm)
        X = R.make_input(64, 256, 128, seed=5)
        eng = _engine(ita, p)
        assert eng.fused_supported, eng.fused_unsupported_reason()
        assert "k_ita_mh" in eng.describe()
        y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
        eng.sync_check()
        _cmp(y, RC.mha(p, X, intermediates=False)["Y"], f"config 3 sample softmax={sm}")
        eng.close()


@pytest.mark.parametrize("round_mode", [0, 1, 2])
@pytest.mark.parametrize("softmax_mode", [0, 1])
@pytest.mark.parametrize("head_mode", [0, 1])
@pytest.mark.parametrize("H", [4, 1, 3])
def test_multi_head_every_intermediate_vs_oracle(ita, R, round_mode, softmax_mode, head_mode, H):
    """Config 3 dims (S=256, E=128, P=64) on the multi-head tcgen05 kernels with every on-chip intermediate dumped (Qp, Kp,
    Vp, A, Asm, O per head, Out, Y): bit-exact vs the oracle for 3 roundings x 2 ITAmax variants x both head conventions
    (per-head step6_O + step7_Osum / the module's concatenated projection), ragged batch (5 frames, fewer than the SMs)."""
    if H == 1 and head_mode == 1:
        pytest.skip("one head: the conventions coincide")
    p = R.make_synthetic(256, 128, 64, H, seed=7 + H, rand_add=True, round_mode=round_mode, softmax_mode=softmax_mode,
                         head_mode=head_mode)
    eng = _engine(ita, p)
    assert eng.fused_supported, eng.fused_unsupported_reason()
    X = R.make_input(5, 256, 128, seed=41)
    X[0] = 127
    X[1] = -128
    ref = R.mha_forward(p, X)
    st = eng.forward_stages(torch.from_numpy(X).to(_dev()), fused_debug=True)
    for k in STAGES + ("Out",):
        _cmp(st[k], ref[k], f"multi-head {k} H={H} round={round_mode} softmax={softmax_mode} heads={head_mode}")
    y = eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
    eng.sync_check()
    _cmp(y, ref["Y"], "multi-head forward Y")
    eng.close()


def test_multi_head_limits_and_fallbacks(ita, R, RC):
    """Edges of the multi-head tcgen05 path: an empty batch is a no-op; 8 heads exceed what k_ita_mh_out keeps on chip (three
    context tiles of H * P columns in shared memory; per-head accumulators H * E / 2 > 256 tensor-memory columns) and take the
    staged kernels -- reported, and still bit-exact; E != 128 at S = 256 is staged."""
    X = R.make_input(2, 256, 128, seed=3)
    p4 = R.make_synthetic(256, 128, 64, 4, seed=21, head_mode=1)
    eng = _engine(ita, p4)
    assert eng.fused_supported and "k_ita_mh" in eng.describe()
    assert tuple(eng.forward(torch.empty((0, 256, 128), dtype=torch.int8, device=_dev())).shape) == (0, 256, 128)
    eng.close()
    p8 = R.make_synthetic(256, 128, 64, 8, seed=21, head_mode=1)
    eng = _engine(ita, p8)
    assert not eng.fused_supported and "shared memory" in eng.fused_unsupported_reason(), eng.fused_unsupported_reason()
    y = eng.forward(torch.from_numpy(X).to(_dev()))                     # PATH_AUTO -> staged (says so once on stderr)
    _cmp(y, RC.mha(p8, X, intermediates=False)["Y"], "8 heads, concatenated (staged)")
    eng.close()
    # 5 and 6 heads would fit the shared memory of the concatenated form, but the pipelines are budgeted and checked up to 4
    p5 = R.make_synthetic(256, 128, 64, 5, seed=22, head_mode=1)
    eng = _engine(ita, p5)
    assert not eng.fused_supported and "H <= 4" in eng.fused_unsupported_reason(), eng.fused_unsupported_reason()
    _cmp(eng.forward(torch.from_numpy(X).to(_dev())), RC.mha(p5, X, intermediates=False)["Y"], "5 heads, concatenated (staged)")
    eng.close()
    p8p = R.make_synthetic(256, 128, 64, 8, seed=21, head_mode=0)
    eng = _engine(ita, p8p)
    assert not eng.fused_supported and "tensor-memory" in eng.fused_unsupported_reason()
    y = eng.forward(torch.from_numpy(X).to(_dev()))                     # PATH_AUTO -> staged (says so once on stderr)
    _cmp(y, RC.mha(p8p, X, intermediates=False)["Y"], "8 heads, per-head convention (staged)")
    with pytest.raises(ita.ITAError):
        eng.forward(torch.from_numpy(X).to(_dev()), path=ita.PATH_FUSED)
    eng.close()
    p64 = R.make_synthetic(256, 64, 64, 2, seed=5, head_mode=1)
    eng = _engine(ita, p64)
    assert not eng.fused_supported
    X64 = R.make_input(2, 256, 64, seed=4)
    _cmp(eng.forward(torch.from_numpy(X64).to(_dev())), RC.mha(p64, X64, intermediates=False)["Y"], "S=256, E=64 (staged)")
    eng.close()


@pytest.mark.parametrize("fp", ["2", "1", "0"])
def test_multi_head_requant_forms_and_batches(ita, R, RC, fp, monkeypatch):
    """The multi-head kernels in every requant form (ITA_FUSED_FP: 2 = two-FFMA / one-FFMA offset forms, 1 = three-FFMA, 0 =
    conversion / 64-bit integer forms), batches around the SM count (1 frame; 149 and 300: several frames per CTA, ragged),
    against the C oracle and the staged kernels."""
    monkeypatch.setenv("ITA_FUSED_FP", fp)
    for rm in (0, 2):
        p = R.make_synthetic(256, 128, 64, 4, seed=3, rand_add=(rm != 2), round_mode=rm, head_mode=1)
        eng = _engine(ita, p)
        assert eng.fused_supported, eng.fused_unsupported_reason()
        for B in (1, 149, 300):
            X = R.make_input(B, 256, 128, seed=50 + B)
            x = torch.from_numpy(X).to(_dev())
            y = eng.forward(x, path=ita.PATH_FUSED)
            eng.sync_check()
            _cmp(y, RC.mha(p, X, intermediates=False)["Y"], f"multi-head B={B} round={rm} ITA_FUSED_FP={fp} [{eng.describe()}]")
        assert torch.equal(eng.forward(x, path=ita.PATH_FUSED), eng.forward(x, path=ita.PATH_STAGED))
        eng.close()


# ----------------------------------------------------------------------------------------------------- IREE dispatch symbol
@pytest.mark.parametrize("f16", [False, True])
def test_dispatch_shim_computes_the_operator(ita, f16):
    """``ITASelfAttention_workgroup(in, out)`` (reference plugin/ITA_dispatch.c:17-27; expanded-memref form :31-56) backed by
    libita_b200: host [1,128,E] buffers in, the module's forward out -- equal to ``ITASelfAttention.forward`` on the same
    tokens (bit-exact for f32 elements; for float16_t elements equal to the f32 result computed from the f16-rounded input
    and rounded to f16)."""
    import ctypes as C
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    p = ita.params_from_captured_block(q, "b0_")
    eng = ita.ITAEngine(p, _dev())
    s_in, s_out = float(q["b0_s_x"]), float(q["b0_out_proj_sout"])
    pkg = os.path.dirname(ita._lib.LIB_PATH)
    shim = C.CDLL(os.path.join(pkg, "libita_dispatch_shim_f16.so" if f16 else "libita_dispatch_shim.so"))
    shim.ita_dispatch_bind.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int32, C.c_int32]
    assert shim.ita_dispatch_bind(eng._h, s_in, s_out, 128, 64) == 0
    x32 = np.ascontiguousarray(q["attn_in"][1])                      # [128, 64] f32
    dt = np.float16 if f16 else np.float32
    x = x32.astype(dt)
    want = ita.dequantize(eng.forward(ita.quantize(torch.from_numpy(x.astype(np.float32)[None]).to(_dev()), s_in)), s_out)
    want = want.cpu().numpy()[0].astype(dt)
    if not f16:
        assert np.array_equal(want, q["attn_out"][1])               # and that is the reference's captured block output
    out = np.full_like(x, 7)
    shim.ITASelfAttention_workgroup.argtypes = [C.c_void_p, C.c_void_p]
    shim.ITASelfAttention_workgroup(x.ctypes.data, out.ctypes.data)
    assert shim.ita_dispatch_last_status() == 0
    assert np.array_equal(out, want)
    out2 = np.full_like(x, 9)
    shim.ITASelfAttention_workgroup_expanded.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t] * 2
    shim.ITASelfAttention_workgroup_expanded(x.ctypes.data, x.ctypes.data, 0, x.size, 1, out2.ctypes.data, out2.ctypes.data, 0, out2.size, 1)
    assert shim.ita_dispatch_last_status() == 0 and np.array_equal(out2, want)
    eng.close()
