"""CPU tests: the C-ABI library loads and exports every symbol include/ita_b200.h declares; the host-side mirror of the
reference interface (parameter derivation, module surface, sharding) behaves like the reference's."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT, to_engine_params


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "ita_b200.h")).read()
    return sorted(set(re.findall(r"^ITA_API\s+[\w\s\*]+?\b(ita_\w+)\s*\(", src, flags=re.M)))


def test_library_exports_every_declared_symbol(ita):
    names = _declared_symbols()
    assert len(names) >= 17, names
    assert sorted(ita._lib.SYMBOLS) == names          # the binding declares exactly the header's symbols
    lib = ctypes.CDLL(ita._lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), n
    assert ita._lib.load().ita_b200_abi_version() == 1
    assert ita._lib.load().ita_kernel_launch_count() >= 0


def test_enums_match_header_and_oracle(ita, R):
    src = open(os.path.join(ROOT, "include", "ita_b200.h")).read()
    for name, val in [("ITA_ROUND_TORCH_RNE_F32", 2), ("ITA_SOFTMAX_STREAM16", 1), ("ITA_HEADS_TORCH", 1), ("ITA_STEP4_SOFTMAX", 40),
                      ("ITA_PATH_STAGED", 2)]:
        assert re.search(rf"{name} = {val}\b", src), name
    L = ita._lib
    assert (L.ROUND_HALF_UP, L.ROUND_FLOOR, L.ROUND_TORCH_RNE_F32) == (R.ROUND_HALF_UP, R.ROUND_FLOOR, R.ROUND_TORCH_RNE_F32)
    assert (L.ST_Q, L.ST_QK, L.ST_AV, L.ST_O, L.ST_SUM, L.ST_FC2, L.NUM_STAGES) == (R.ST_Q, R.ST_QK, R.ST_AV, R.ST_O, R.ST_SUM,
                                                                                    R.ST_FC2, R.NUM_STAGES)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(ita, R):
    """Without a GPU the product path must fail loudly, not compute on the CPU."""
    p = R.make_synthetic(64, 64, 64, 1)
    with pytest.raises(ita.ITAError):
        ita.ITAEngine(to_engine_params(ita, p))
    with pytest.raises(ita.ITAError):
        ita.selftest()
    m = ita.ITASelfAttention(64, 64, 1)
    with pytest.raises(ValueError):
        m(torch.zeros(1, 64, 64))
    # the C entry point itself reports the error through its return code / ita_last_error
    lib = ita._lib.load()
    assert lib.ita_mha_create(None, None) != 0
    assert b"null" in lib.ita_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "drone-oa-iree-vit-accelerator_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                s = open(os.path.join(dp, f)).read()
                assert "ita_ref" not in s and "oracle" not in s.lower().replace("no oracle", ""), os.path.join(dp, f)


def test_hw_param_derivation_matches_reference(ita):
    h = np.load(os.path.join(GOLDEN, "hw_params.npz"))
    for s, m, sh in zip(h["scales"], h["mult"], h["shift"]):
        m, sh = (128, int(sh) - 1) if m == 256 else (int(m), int(sh))
        assert ita.calculate_hw_params(float(s)) == (m, sh)
    for i in range(len(h["bias"])):
        lp = {"bias": h["bias"][i], "input": [{"scale": float(h["s_act"][i])}],
              "weight": {"scale": float(h["s_w"][i]), "q_value": np.zeros((64, 1))}}
        assert np.array_equal(ita.dequantize_bias_to_int32(lp), h["bias_i32"][i])
    assert np.array_equal(ita.dequantize_bias_to_int32({"bias": None, "weight": {"q_value": np.zeros((5, 3))}}), np.zeros(5, np.int32))


def _module_from_fixture(ita, q, pre):
    P, E = q[pre + "q_proj_w"].shape
    m = ita.ITASelfAttention(E, P, 1)
    for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
        lin = getattr(m, nm)
        lin.weight_int8.copy_(torch.from_numpy(q[pre + nm + "_w"]))
        lin.bias.copy_(torch.from_numpy(q[pre + nm + "_bias"]))
        lin.weight_scale = float(q[pre + nm + "_sw"])
        lin.scale = float(q[pre + nm + "_sout"])
    m.quant.scale = float(q[pre + "s_x"])
    m.matmul1.scale = float(q[pre + "qk_sout"])
    m.matmul2.scale = float(q[pre + "av_sout"])
    return m


@pytest.mark.parametrize("variant", ["E64", "E128"])
def test_module_quant_params_match_reference_harness(ita, variant):
    """ITASelfAttention.quant_params == what the reference harness derives for the golden model
    (translate_torch_scales_to_hw_params / dequantize_bias_to_int32 / weight.T[None], W_B.py:247-305,479-491)."""
    q = np.load(os.path.join(GOLDEN, f"qat_block_{variant}.npz"))
    pre = "b0_"
    m = _module_from_fixture(ita, q, pre)
    p = m.quant_params(128)
    L = ita._lib
    for st, nm in [(L.ST_Q, "q_proj"), (L.ST_K, "k_proj"), (L.ST_V, "v_proj"), (L.ST_QK, "qk_matmul"), (L.ST_AV, "av_matmul"),
                   (L.ST_O, "out_proj")]:
        mult, shift, add = [int(v) for v in q[pre + "hw_" + nm]]
        if mult == 256:
            mult, shift = 128, shift - 1
        assert (int(p.mult[st, 0]), int(p.shift[st, 0]), int(p.add[st, 0])) == (mult, shift, add), nm
    assert np.array_equal(p.Wq[0], q[pre + "q_proj_w"].T) and np.array_equal(p.Wo[0], q[pre + "out_proj_w"].T)
    for nm, b in [("q_proj", p.Bq), ("k_proj", p.Bk), ("v_proj", p.Bv), ("out_proj", p.Bo)]:
        assert np.array_equal(b[0], q[pre + nm + "_bias_i32"]), nm
    assert p.round_mode == L.ROUND_TORCH_RNE_F32 and p.head_mode == L.HEADS_PYITA


def test_module_surface(ita):
    m = ita.ITASelfAttention(embed_dim=64, proj_dim=192, num_heads=1)
    for name in ("q_proj", "k_proj", "v_proj", "out_proj", "custom_softmax", "matmul1", "matmul2", "quant", "dequant"):
        assert hasattr(m, name), name                      # names the reference's hooks / state dicts address
    assert m.head_dim == 192 and m.custom_softmax.scale == 1.0 / 255.0
    with pytest.raises(ValueError):
        ita.ITASelfAttention(64, 100, 3)
    # multi-head torch convention: heads are column slices of one projection (layers.py:110-113)
    m4 = ita.ITASelfAttention(32, 64, 4)
    m4.q_proj.weight_int8.copy_(torch.arange(64 * 32).reshape(64, 32).to(torch.int8))
    p = m4.quant_params(16)
    assert p.P == 16 and p.H == 4 and p.head_mode == ita.HEADS_TORCH
    w = m4.q_proj.weight_int8.numpy()
    assert np.array_equal(p.Wq[2], w[32:48, :].T)


def test_quant_dict_container(ita, R):
    p = R.make_synthetic(64, 64, 64, 2, seed=1, F=32)
    qd = {k: {"mult": list(p.mult[i]), "shift": list(p.shift[i]), "add": list(p.add[i])} for i, k in enumerate(R.STAGE_NAMES)
          if k != "relu"}
    qd["relu"] = {"mult": int(p.mult[R.ST_RELU, 0]), "shift": int(p.shift[R.ST_RELU, 0]), "add": 0, "q_1": 0, "q_b": 0, "q_c": 0}
    t = dict(Wq=p.Wq, Wk=p.Wk, Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv, Bo=p.Bo, Wff=p.Wff, Wff2=p.Wff2, Bff=p.Bff, Bff2=p.Bff2)
    e = ita.ITAQuantParams.from_quant_dict(64, 64, 64, 2, t, qd, F=32)
    assert np.array_equal(e.mult[:7], p.mult[:7]) and np.array_equal(e.shift[:7], p.shift[:7])
    assert e.mult[R.ST_RELU, 1] == p.mult[R.ST_RELU, 0]      # scalar relu params broadcast over heads


def test_shard_range(ita):
    for total in (0, 1, 2, 7, 8, 4096, 4097, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            sh = ita.all_shards(total, world)
            assert sh[0][0] == 0 and sh[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
            sizes = [e - b for b, e in sh]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        ita.shard_range(4, 4, 4)


def test_dispatch_shim_exports_reference_symbols(ita):
    """libita_dispatch_shim*.so export the C symbol the reference's compiled model imports
    (samples/inference_udp_FPGA_custom_dispatch/plugin/ITA_dispatch.c:17-27 and :31-56), f32 and float16_t element builds."""
    pkg = os.path.dirname(ita._lib.LIB_PATH)
    for name, size in (("libita_dispatch_shim.so", 4), ("libita_dispatch_shim_f16.so", 2)):
        lib = ctypes.CDLL(os.path.join(pkg, name))
        for sym in ("ITASelfAttention_workgroup", "ITASelfAttention_workgroup_expanded", "ita_dispatch_bind", "ita_dispatch_last_status"):
            assert hasattr(lib, sym), (name, sym)
        assert lib.ita_dispatch_elem_size() == size
        # unbound: the call must not crash, reports an error status
        buf_in = (ctypes.c_float * 16)()
        buf_out = (ctypes.c_float * 16)()
        lib.ITASelfAttention_workgroup(buf_in, buf_out)
        assert lib.ita_dispatch_last_status() != 0
        assert lib.ita_dispatch_bind(None, ctypes.c_float(1.0), ctypes.c_float(1.0), 128, 64) != 0


def _same_params(a, b):
    for k in ("Wq", "Wk", "Wv", "Wo", "Bq", "Bk", "Bv", "Bo", "mult", "shift", "add", "eff"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k


def test_module_state_dict_round_trip_and_reference_checkpoint(ita):
    """``state_dict()`` carries every scale / zero-point / weight (round trip), and ``load_state_dict`` accepts the keys of
    a CONVERTED REFERENCE checkpoint (``attention_blocks.N.q_proj._packed_params._packed_params``, ``.scale``,
    ``.zero_point``, ``matmul1.scale`` ..., loaded strict=False at tests/export_and_validation_W_B.py:387-392)."""
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    want = ita.params_from_captured_block(q, "b0_")
    sd_ref = torch.load(os.path.join(GOLDEN, "qat_state_E64.pt"), weights_only=False)
    sub = {k[len("attention_blocks.0."):]: v for k, v in sd_ref.items() if k.startswith("attention_blocks.0.")}
    m = ita.ITASelfAttention(64, 192, 1)
    res = m.load_state_dict(sub, strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    _same_params(m.quant_params(128), want)
    # round trip through our own state dict
    sd = m.state_dict()
    for k in ("q_proj.scale", "q_proj.zero_point", "q_proj.weight_scale", "q_proj.weight_int8", "matmul1.scale", "quant.scale"):
        assert k in sd, k
    m2 = ita.ITASelfAttention(64, 192, 1)
    m2.load_state_dict(sd)
    _same_params(m2.quant_params(128), want)
    # a parameter change invalidates cached engines (the cache key follows the buffers)
    v0 = m2._params_version()
    m2.quant.scale = 0.5
    assert m2._params_version() != v0
    m2.q_proj.weight_int8.add_(1)
    assert m2._params_version() != v0
    # non-zero zero-points are refused, not silently ignored
    m2.matmul1.zero_point = 3
    with pytest.raises(ValueError):
        m2.quant_params(128)


def test_multi_head_reference_checkpoint_and_capture_agree(ita, R):
    """The converted ``ITASelfAttention_QAT(128, 256, 4)`` of tests/golden/make_golden.py: its state_dict loaded into
    ``ITASelfAttention``, its flattened capture through ``params_from_captured_block`` and the oracle-side recipe of
    tests/test_oracle.py give the same multi-head parameter set (torch head convention, layers.py:110-123)."""
    from test_oracle import _block_params_heads
    q = np.load(os.path.join(GOLDEN, "qat_block_H4_S256.npz"))
    want = ita.params_from_captured_block(q, "b0_")
    assert (want.S, want.E, want.P, want.H, want.head_mode) == (256, 128, 64, 4, ita._lib.HEADS_TORCH)
    o = _block_params_heads(R, q, "b0_", R.ROUND_TORCH_RNE_F32)
    for k in ("Wq", "Wk", "Wv", "Wo", "Bq", "Bk", "Bv", "Bo", "eff"):
        assert np.array_equal(getattr(want, k), getattr(o, k)), k
    for st in (R.ST_Q, R.ST_K, R.ST_V, R.ST_QK, R.ST_AV, R.ST_O):
        assert np.array_equal(want.mult[st], o.mult[st]) and np.array_equal(want.shift[st], o.shift[st]), st
    m = ita.ITASelfAttention(128, 256, 4)
    res = m.load_state_dict(torch.load(os.path.join(GOLDEN, "qat_state_H4_S256.pt"), weights_only=False), strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    _same_params(m.quant_params(256), want)


def test_full_model_state_dict_reader_and_float_stages(ita):
    """``ITAViTLSTM.from_state_dict`` on a converted reference checkpoint: the int8 layer parameters equal the harness'
    recipe on the captured block, and the float pre/post stages (tokenizer; fusion, decoder, LSTM, head) reproduce the
    reference's captured tensors (CPU float ops, exact)."""
    import importlib
    import sys
    sys.path.insert(0, GOLDEN)
    import seeded
    M = importlib.import_module("drone-oa-iree-vit-accelerator_b200.model")
    q = np.load(os.path.join(GOLDEN, "qat_model_E64.npz"))
    sd = dict(torch.load(os.path.join(GOLDEN, "qat_state_E64.pt"), weights_only=False))
    p, sc = M.layer_quant_params(sd, 0)
    want = ita.params_from_captured_block(q, "b0_", with_ffn=True)
    _same_params(p, want)
    for k in ("Wff", "Wff2", "Bff", "Bff2"):
        assert np.array_equal(getattr(p, k), getattr(want, k)), k
    assert sc == {"s_attn_in": float(q["b0_s_x"]), "s_attn_out": float(q["b0_out_proj_sout"]), "s_ffn_in": float(q["b0_fc1_sin"]),
                  "s_ffn_out": float(q["b0_fc2_sout"])}
    with pytest.raises(KeyError):
        M.ITAViTLSTM.from_state_dict(sd)                 # float tensors missing
    for name, shape in M.reference_float_shapes(1, 64).items():
        sd[name] = torch.from_numpy(seeded.float_param(name, shape))
    m = M.ITAViTLSTM.from_state_dict(sd)
    with torch.no_grad():
        tok = m.tokens_torch(torch.from_numpy(q["img"]))     # torch twin of the CUDA tokenizer kernel (the checker on the GPU box)
        assert torch.equal(tok, torch.from_numpy(q["tokenizer_out"]))
        with pytest.raises(ValueError):
            m.tokens(torch.from_numpy(q["img"]))             # the product stage is the CUDA kernel: no CPU fallback
        out, (h, c) = m.head(torch.from_numpy(q["norm2_out"]), torch.from_numpy(q["vel"]), torch.from_numpy(q["quat"]))
        assert np.abs(out.numpy() - q["out"]).max() <= 1e-6 and np.abs(h.numpy() - q["hn"]).max() <= 1e-6
    with pytest.raises(ValueError):
        m([torch.zeros(1, 1, 60, 90), torch.zeros(1, 1), torch.zeros(1, 4)])     # CPU tensors: no CPU fallback


def test_committed_bench_lines_follow_the_contract():
    """The bench lines committed under profiles/ (taken on the B200 box with the final build) carry every key of the bench
    contract: the driver parses the same line from ``python bench.py``."""
    import json
    prof = os.path.join(ROOT, "profiles")
    for name in ("bench_r2_final_c2.json", "bench_r2_final_c3.json", "bench_r2_final_c4.json"):
        d = json.load(open(os.path.join(prof, name)))
        for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                  "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks"):
            assert k in d, (name, k)
        assert d["unit"] == "frames/s" and d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
        assert "workload" in d["config"] and d["config"]["rounding"] == "torch_rne_f32"
        for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
            assert k in d["roofline"], (name, k)
        for k in ("value", "unit", "cores", "kind", "sample"):
            assert k in d["cpu_baseline"], (name, k)
        for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
            assert k in d["e2e"], (name, k)
        assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["gpu_launches"] > 0 and d["parity_spot_check"] is True
        assert d["clocks"]["reasons"] == [] or d["clocks"]["reasons"] == ["sw_power_cap"]
    ref = json.load(open(os.path.join(prof, "bench_r2_final_c2_reference_arm.json")))
    assert ref["impl"] == "reference" and ref["e2e"]["h2d_bytes_per_step"] == 0 and ref["cpu_baseline"]["kind"] == "port"
    assert ref["config"]["frames_per_gpu_per_step"] == json.load(open(os.path.join(prof, "bench_r2_final_c2.json")))["config"]["frames_per_gpu_per_step"]
