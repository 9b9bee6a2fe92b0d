"""CPU tests: the oracle (oracle/ita_ref.py + oracle/ita_ref.c) against the golden fixtures generated from the importable
reference (tests/golden/make_golden.py), its two implementations against each other, and the domain's edge cases."""
import os

import numpy as np
import pytest

from conftest import GOLDEN


def test_itamax_repo_bit_exact_vs_reference(R, RC):
    """IntegerApproximatedSoftmaxFn.forward (reference models/ITA/QAT/ITA_softmax.py:19-77): 0 mismatches."""
    g = np.load(os.path.join(GOLDEN, "itamax_repo.npz"))
    n = 0
    for k in g.files:
        if not k.startswith("x"):
            continue
        want = g["y" + k[1:]]
        assert np.array_equal(R.itamax_repo_global(g[k]), want), k
        assert np.array_equal(RC.itamax(g[k], R.SOFTMAX_REPO_GLOBAL), want), k
        n += want.size
    assert n > 50000


def test_hw_params_vs_reference(R):
    """calculate_hw_params / dequantize_bias_to_int32 (reference tests/export_and_validation_W_B.py:225-245)."""
    h = np.load(os.path.join(GOLDEN, "hw_params.npz"))
    for s, m, sh in zip(h["scales"], h["mult"], h["shift"]):
        m, sh = int(m), int(sh)
        if m == 256:                      # the reference's uint8 overflow case, renormalised to the same rational
            m, sh = 128, sh - 1
        assert R.calculate_hw_params(float(s)) == (m, sh)
        if s > 0:
            assert 128 <= m <= 255 and abs(m / 2.0 ** sh - s) <= s * 2 ** -8
    for i in range(len(h["bias"])):
        assert np.array_equal(R.bias_to_int32(h["bias"][i], float(h["s_act"][i]), float(h["s_w"][i])), h["bias_i32"][i])


def _block_params(R, q, pre, round_mode):
    """Oracle parameter set for one captured reference block (the harness' recipe, W_B.py:479-491)."""
    wq = q[pre + "q_proj_w"]
    P, E = wq.shape
    S = q[pre + "x"].shape[1]
    mult = np.zeros((R.NUM_STAGES, 1), np.int32); shift = np.ones((R.NUM_STAGES, 1), np.int32)
    add = np.zeros((R.NUM_STAGES, 1), np.int32); eff = np.zeros((R.NUM_STAGES, 1), np.float32)
    s = lambda k: float(q[pre + k])
    effs = {R.ST_Q: s("q_proj_sin") * s("q_proj_sw") / s("q_proj_sout"), R.ST_K: s("k_proj_sin") * s("k_proj_sw") / s("k_proj_sout"),
            R.ST_V: s("v_proj_sin") * s("v_proj_sw") / s("v_proj_sout"), R.ST_QK: s("q_proj_sout") * s("k_proj_sout") / s("qk_sout"),
            R.ST_AV: s("softmax_sout") * s("v_proj_sout") / s("av_sout"), R.ST_O: s("out_proj_sin") * s("out_proj_sw") / s("out_proj_sout")}
    names = {R.ST_Q: "q_proj", R.ST_K: "k_proj", R.ST_V: "v_proj", R.ST_QK: "qk_matmul", R.ST_AV: "av_matmul", R.ST_O: "out_proj"}
    f32 = np.float32
    lin32 = lambda n: f32(f32(f32(s(n + "_sin")) * f32(s(n + "_sw"))) / f32(s(n + "_sout")))   # qnnpack: float32 arithmetic
    eff32 = {R.ST_Q: lin32("q_proj"), R.ST_K: lin32("k_proj"), R.ST_V: lin32("v_proj"), R.ST_O: lin32("out_proj"),
             R.ST_QK: f32(effs[R.ST_QK]), R.ST_AV: f32(effs[R.ST_AV])}     # patched matmuls: double product cast to f32
    for st, e in effs.items():
        mult[st, 0], shift[st, 0], add[st, 0] = [int(v) for v in q[pre + "hw_" + names[st]]]
        if mult[st, 0] == 256:
            mult[st, 0], shift[st, 0] = 128, shift[st, 0] - 1
        eff[st, 0] = eff32[st]
        assert (int(mult[st, 0]), int(shift[st, 0])) == R.calculate_hw_params(e)
    T = lambda k: np.ascontiguousarray(q[pre + k + "_w"].T[None])
    B = lambda k: q[pre + k + "_bias_i32"][None].astype(np.int32)
    return R.ITAParams(S=S, E=E, P=P, H=1, Wq=T("q_proj"), Wk=T("k_proj"), Wv=T("v_proj"), Wo=T("out_proj"), Bq=B("q_proj"),
                       Bk=B("k_proj"), Bv=B("v_proj"), Bo=B("out_proj"), mult=mult, shift=shift, add=add, eff=eff,
                       round_mode=round_mode)


def _block_params_heads(R, q, pre, round_mode):
    """Oracle parameter set of the captured MULTI-HEAD reference block (qat_block_H4_S256.npz): heads are column slices of
    one projection and one out_proj runs over the concatenated contexts (layers.py:110-123) -> HEADS_TORCH; every head shares
    the stage's scale."""
    H = int(q["H"])
    p1 = _block_params(R, q, pre, round_mode)                      # H = 1 view: [1,E,H*hd] weights, per-stage constants
    PT, E = q[pre + "q_proj_w"].shape
    hd = PT // H
    split = lambda k: np.ascontiguousarray(q[pre + k + "_w"].reshape(H, hd, E).transpose(0, 2, 1))      # [H*hd,E] -> [H,E,hd]
    bsplit = lambda k: q[pre + k + "_bias_i32"].astype(np.int32).reshape(H, hd)
    Wo = np.ascontiguousarray(q[pre + "out_proj_w"].T.reshape(H, hd, E))                               # [E,H*hd] -> [H,hd,E]
    Bo = np.zeros((H, E), np.int32)
    Bo[0] = q[pre + "out_proj_bias_i32"]
    rep = lambda a: np.ascontiguousarray(np.repeat(a, H, axis=1))
    return R.ITAParams(S=p1.S, E=E, P=hd, H=H, Wq=split("q_proj"), Wk=split("k_proj"), Wv=split("v_proj"), Wo=Wo, Bq=bsplit("q_proj"),
                       Bk=bsplit("k_proj"), Bv=bsplit("v_proj"), Bo=Bo, mult=rep(p1.mult), shift=rep(p1.shift), add=rep(p1.add),
                       eff=rep(p1.eff), round_mode=round_mode, head_mode=R.HEADS_TORCH)


def _heads_fixture_stage(q, pre, key, H):
    """A stage of qat_block_H4_S256.npz in the oracle's layout: projections [B,N,H*hd] -> [B,H,N,hd]; the rest as stored."""
    a = q[pre + key]
    if key in ("q_proj_out", "k_proj_out", "v_proj_out"):
        B, N, PT = a.shape
        return np.ascontiguousarray(a.reshape(B, N, H, PT // H).transpose(0, 2, 1, 3))
    return a


def test_converted_multi_head_block_torch_rounding_bit_exact(R, RC):
    """BASELINE config 3 as the reference module, ``ITASelfAttention_QAT(128, 256, 4)`` on 256 tokens, converted and captured by
    tests/golden/make_golden.py: the oracle (torch head convention, torch rounding) reproduces EVERY stage bit-exactly, which
    pins the head split / context merge / single out_proj of layers.py:110-123 to reference output."""
    q = np.load(os.path.join(GOLDEN, "qat_block_H4_S256.npz"))
    pre, H = "b0_", int(q["H"])
    assert H == 4 and q[pre + "x"].shape == (2, 256, 128)
    p = _block_params_heads(R, q, pre, R.ROUND_TORCH_RNE_F32)
    assert (p.S, p.E, p.P, p.H) == (256, 128, 64, 4)
    for impl in (lambda: R.mha_forward(p, q[pre + "x"]), lambda: RC.mha(p, q[pre + "x"])):
        out = impl()
        for k, fk in STAGE_FIX.items():
            want = _heads_fixture_stage(q, pre, fk, H).reshape(out[k].shape)
            assert np.array_equal(out[k], want), (k, int((out[k] != want).sum()))
    # the float boundary: quant stub of the input, dequant of the output (layers.py:103,126)
    xq = np.clip(np.rint(q["x_float"] / np.float32(q["quant_scale"])), -128, 127).astype(np.int8)
    assert np.array_equal(xq, q[pre + "x"])
    assert np.array_equal(q[pre + "out_proj_out"].astype(np.float32) * np.float32(q["out_scale"]), q["y_float"])


STAGE_FIX = {"Qp": "q_proj_out", "Kp": "k_proj_out", "Vp": "v_proj_out", "A": "qk_out", "Asm": "softmax_out", "O": "av_out",
             "Y": "out_proj_out"}


def _fixture_file(variant):
    return os.path.join(GOLDEN, f"qat_model_{variant[6:]}.npz" if variant.startswith("model_") else f"qat_block_{variant}.npz")


@pytest.mark.parametrize("variant,blocks", [("E64", 1), ("E128", 2), ("model_E64", 1)])
def test_converted_block_torch_rounding_bit_exact(R, RC, variant, blocks):
    """Whole converted ITASelfAttention_QAT block (qnnpack, matmul2 patched as W_B.py:408-423): in torch-rounding mode the
    oracle reproduces EVERY stage of the reference bit-exactly, stage by stage AND end to end."""
    q = np.load(_fixture_file(variant))
    for bi in range(blocks):
        pre = f"b{bi}_"
        p = _block_params(R, q, pre, R.ROUND_TORCH_RNE_F32)
        x = q[pre + "x"]
        for impl in (lambda: R.mha_forward(p, x), lambda: RC.mha(p, x)):
            out = impl()
            for k, fk in STAGE_FIX.items():
                want = q[pre + fk].reshape(out[k].shape)
                assert np.array_equal(out[k], want), (variant, bi, k, int((out[k] != want).sum()))


@pytest.mark.parametrize("variant,blocks", [("E64", 1), ("E128", 2)])
def test_converted_block_golden_rounding_within_reference_tolerance(R, variant, blocks):
    """With the golden model's mult/shift round-half-up requant the reference's own acceptance rule must hold per
    stage: mean(|delta|) <= 1.0 LSB (verify_step, W_B.py:308-334); we also require max |delta| <= 1 on linear stages."""
    q = np.load(os.path.join(GOLDEN, f"qat_block_{variant}.npz"))
    for bi in range(blocks):
        pre = f"b{bi}_"
        p = _block_params(R, q, pre, R.ROUND_HALF_UP)
        out = R.mha_forward(p, q[pre + "x"])
        for k, fk in STAGE_FIX.items():
            want = q[pre + fk].reshape(out[k].shape).astype(np.int32)
            mae = np.abs(out[k].astype(np.int32) - want).mean()
            assert mae <= 1.0, (variant, bi, k, mae)
        for k in ("Qp", "Kp", "Vp"):
            want = q[pre + STAGE_FIX[k]].reshape(out[k].shape).astype(np.int32)
            assert np.abs(out[k].astype(np.int32) - want).max() <= 1


def _ffn_params(R, q, pre="b0_"):
    """Oracle parameter set of the captured converted FFN block (fc1 -> ReLU -> fc2)."""
    x = q[pre + "ffn_x"]
    S, E = x.shape[1:]
    F = q[pre + "fc1_w"].shape[0]
    mult = np.zeros((R.NUM_STAGES, 1), np.int32); shift = np.ones((R.NUM_STAGES, 1), np.int32)
    add = np.zeros((R.NUM_STAGES, 1), np.int32); eff = np.zeros((R.NUM_STAGES, 1), np.float32)
    s = lambda k: float(q[pre + k])
    f32 = np.float32          # the backend (qnnpack) forms the requant scale in float32 arithmetic
    eff[R.ST_FC1, 0] = f32(f32(f32(s("fc1_sin")) * f32(s("fc1_sw"))) / f32(s("fc1_sout")))
    eff[R.ST_FC2, 0] = f32(f32(f32(s("fc2_sin")) * f32(s("fc2_sw"))) / f32(s("fc2_sout")))
    z = np.zeros((1, E, 1), np.int8)
    p = R.ITAParams(S=S, E=E, P=1, H=1, Wq=z, Wk=z, Wv=z, Wo=np.zeros((1, 1, E), np.int8), Bq=np.zeros((1, 1), np.int32),
                    Bk=np.zeros((1, 1), np.int32), Bv=np.zeros((1, 1), np.int32), Bo=np.zeros((1, E), np.int32), mult=mult,
                    shift=shift, add=add, eff=eff, F=F, Wff=np.ascontiguousarray(q[pre + "fc1_w"].T[None]),
                    Wff2=np.ascontiguousarray(q[pre + "fc2_w"].T[None]), Bff=q[pre + "fc1_bias_i32"][None].astype(np.int32),
                    Bff2=q[pre + "fc2_bias_i32"][None].astype(np.int32), round_mode=R.ROUND_TORCH_RNE_F32)
    p.mult[R.ST_RELU, 0] = 0          # quantised ReLU keeps the scale: no second requant
    return p


@pytest.mark.parametrize("fixture", ["qat_block_E64.npz", "qat_block_E128.npz", "qat_model_E64.npz"])
def test_ffn_fixture(R, RC, fixture):
    """FFN block of the converted model: fc1 -> ReLU -> fc2 (reference models/ITA/QAT/layers.py:61-75), bit-exact."""
    q = np.load(os.path.join(GOLDEN, fixture))
    pre = "b0_"
    p = _ffn_params(R, q, pre)
    x = q[pre + "ffn_x"]
    for out in (R.ffn_forward(p, x), RC.ffn(p, x)):
        assert np.array_equal(out["FF1"], q[pre + "relu_out"])
        assert np.array_equal(out["FF2"], q[pre + "fc2_out"])   # bit-exact once eff is formed in float32 like the backend


@pytest.mark.parametrize("dims", [(64, 64, 64, 1, 0), (128, 64, 192, 1, 256), (256, 128, 64, 4, 0), (32, 32, 32, 2, 64)])
@pytest.mark.parametrize("round_mode", [0, 1, 2])
@pytest.mark.parametrize("softmax_mode", [0, 1])
def test_c_oracle_equals_numpy_oracle(R, RC, dims, round_mode, softmax_mode):
    S, E, P, H, F = dims
    for head_mode in (0, 1):
        p = R.make_synthetic(S, E, P, H, seed=3, F=F, rand_add=True, round_mode=round_mode, softmax_mode=softmax_mode,
                             head_mode=head_mode)
        X = R.make_input(3, S, E, seed=9)
        a, b = R.mha_forward(p, X), RC.mha(p, X)
        for k in a:
            assert np.array_equal(a[k], b[k]), (k, dims, round_mode, softmax_mode, head_mode)
        if F:
            a, b = R.ffn_forward(p, X), RC.ffn(p, X)
            for k in a:
                assert np.array_equal(a[k], b[k]), k


def test_requant_edges(R, RC):
    acc = np.array([-2 ** 25, -70000, -32769, -129, -128, -1, 0, 1, 127, 128, 255, 256, 32767, 65535, 2 ** 25 - 1,
                    3 * 2 ** 13, 3 * 2 ** 13 - 1, -3 * 2 ** 13, -3 * 2 ** 13 - 1], np.int32)
    # round-half-up: exact ties go towards +inf for both signs
    assert list(R.requant(np.array([3, -3, 1, -1]), 1, 1, 0)) == [2, -1, 1, 0]
    assert list(R.requant(np.array([3, -3, 1, -1]), 1, 1, 0, R.ROUND_FLOOR)) == [1, -2, 0, -1]
    assert list(R.requant(np.array([5, 7, -5, -7]), 0, 0, 0, R.ROUND_TORCH_RNE_F32, 0.5)) == [2, 4, -2, -4]   # half-even
    for mult, shift, add in [(128, 7, 0), (255, 14, 3), (64, 9, -8), (1, 0, 0), (200, 31, 0), (129, 24, -128)]:
        for mode in (0, 1):
            want = R.requant(acc, mult, shift, add, mode)
            assert np.array_equal(RC.requant(acc, mult, shift, add, mode), want)
            assert want.min() >= -128 and want.max() <= 127
    eff = np.float32(0.0037)
    assert np.array_equal(RC.requant(acc, 0, 0, 2, 2, float(eff)), R.requant(acc, 0, 0, 2, 2, eff))


def test_itamax_edge_rows(R, RC):
    for S in (16, 64, 256):
        eq = np.full((1, S), 7, np.int8)                       # all equal: sum = 256*S
        y = R.itamax_repo_global(eq)
        assert (y == (256 * ((255 * 65536) // (256 * S))) >> 16).all()
        one = np.full((1, S), -128, np.int8); one[0, 5] = 127   # one-hot: sum = 256 -> 255
        y = R.itamax_repo_global(one)
        assert y[0, 5] == 255 and y.sum() == 255
        for mode in (0, 1):
            x = np.random.default_rng(S + mode).integers(-128, 128, size=(33, S)).astype(np.int8)
            assert np.array_equal(R.itamax(x, mode), RC.itamax(x, mode))
            yy = R.itamax(x, mode).astype(np.int64)
            assert yy.max() <= 255 and yy.sum(axis=-1).max() <= 256 + S      # rows sum to <= ~255 (floor rounding)
    # streaming variant: the running max only grows, and a late large value rescales the sum
    x = np.zeros((1, 32), np.int8); x[0, 31] = 100
    y = R.itamax_stream16(x)
    assert y[0, 31] > y[0, 0]


def test_empty_and_single_frame(R, RC):
    p = R.make_synthetic(64, 64, 64, 1, seed=0)
    X = R.make_input(1, 64, 64)
    a = R.mha_forward(p, X[0])          # 2-D input = one frame
    b = RC.mha(p, X)
    assert np.array_equal(a["Y"], b["Y"])
    e = RC.mha(p, np.zeros((0, 64, 64), np.int8))
    assert e["Y"].shape == (0, 64, 64)


def test_properties_batch_independence(R):
    """Frames are independent: permuting / duplicating the batch permutes / duplicates the result."""
    p = R.make_synthetic(128, 64, 192, 1, seed=2)
    X = R.make_input(6, 128, 64, seed=4)
    y = R.mha_forward(p, X)["Y"]
    perm = np.array([3, 0, 5, 1, 4, 2])
    assert np.array_equal(R.mha_forward(p, X[perm])["Y"], y[perm])
    assert np.array_equal(R.mha_forward(p, np.concatenate([X[:1], X[:1]]))["Y"], np.stack([y[0], y[0]]))


def test_ops_per_frame(R):
    assert R.ops_per_frame(64, 64, 64, 1) == 3145728            # 3.15 Mop   
    assert R.ops_per_frame(128, 64, 192, 1) == 25165824         # 25.17 Mop
    assert R.ops_per_frame(128, 64, 192, 1, 256) == 33554432    # 33.55 Mop with FFN
    assert R.ops_per_frame(256, 128, 64, 4) == 134217728        # 134.2 Mop


def test_package_block_loader_matches_harness_recipe(R, ita):
    """params_from_captured_block (product side) == the harness' recipe as restated for the oracle (_block_params)."""
    for f, pre in (("qat_block_E64.npz", "b0_"), ("qat_block_E128.npz", "b0_"), ("qat_block_E128.npz", "b1_"), ("qat_model_E64.npz", "b0_")):
        q = np.load(os.path.join(GOLDEN, f))
        a, b = _block_params(R, q, pre, 2), ita.params_from_captured_block(q, pre, with_ffn=True)
        for k in ("Wq", "Wk", "Wv", "Wo", "Bq", "Bk", "Bv", "Bo"):
            assert np.array_equal(getattr(a, k), getattr(b, k)), (f, k)
        for st in range(6):
            assert (a.mult[st, 0], a.shift[st, 0], a.eff[st, 0]) == (b.mult[st, 0], b.shift[st, 0], b.eff[st, 0]), (f, st)
        fp = _ffn_params(R, q, pre)
        assert np.array_equal(fp.Wff, b.Wff) and np.array_equal(fp.Bff2, b.Bff2) and np.array_equal(fp.Bff, b.Bff)
        assert fp.eff[R.ST_FC1, 0] == b.eff[R.ST_FC1, 0] and fp.eff[R.ST_FC2, 0] == b.eff[R.ST_FC2, 0]


def test_quantize_fixture_definition():
    """torch.quantize_per_tensor (the QuantStub of layers.py:103) == clamp(rne(x * f32(1/scale))) -- the form the CUDA quant
    stub uses -- on the reference-generated vectors (ties, saturation, awkward scales)."""
    g = np.load(os.path.join(GOLDEN, "quantize.npz"))
    for s, x, q in zip(g["scales"], g["x"], g["q"]):
        inv = np.float32(1.0) / np.float32(s)
        with np.errstate(over="ignore", invalid="ignore"):
            y = np.clip(np.rint(x * inv), -128, 127).astype(np.int8)
        assert np.array_equal(y, q), float(s)


def test_layer_oracle_vs_model_fixture(R, ita):
    """Encoder layer (quant, int8 attention, residual + LayerNorm, quant, int8 FFN, residual + LayerNorm; reference
    QAT/model.py:100-113) restated in the oracle vs the full converted-model trace: int8 stages bit-exact, float stages
    within 2e-6 (LayerNorm summation order)."""
    import sys
    sys.path.insert(0, GOLDEN)
    import seeded
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    p = _block_params(R, q, "b0_", R.ROUND_TORCH_RNE_F32)
    f = _ffn_params(R, q)
    p.F, p.Wff, p.Wff2, p.Bff, p.Bff2 = f.F, f.Wff, f.Wff2, f.Bff, f.Bff2
    p.eff[R.ST_FC1], p.eff[R.ST_FC2], p.mult[R.ST_RELU] = f.eff[R.ST_FC1], f.eff[R.ST_FC2], 0
    ln = [seeded.float_param(k, (64,)) for k in ("norms1.0.weight", "norms1.0.bias", "norms2.0.weight", "norms2.0.bias")]
    o = R.layer_forward(p, q["attn_in"], float(q["b0_s_x"]), float(q["b0_out_proj_sout"]), float(q["b0_fc1_sin"]),
                        float(q["b0_fc2_sout"]), *ln)
    assert np.array_equal(o["xq"], q["b0_x"]) and np.array_equal(o["yq"], q["b0_out_proj_out"].reshape(o["yq"].shape))
    assert np.abs(o["x1"] - q["norm1_out"]).max() <= 2e-6
    assert np.array_equal(o["x1q"], q["b0_ffn_x"]) and np.array_equal(o["f2q"], q["b0_fc2_out"])
    assert np.abs(o["y"] - q["norm2_out"]).max() <= 2e-6
