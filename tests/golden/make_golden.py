#!/usr/bin/env python
"""Generates the committed golden fixtures under tests/golden/ by IMPORTING the reference from /root/reference.

Run in the build container only (the reference tree does not exist on the GPU box):

    python tests/golden/make_golden.py

What is produced (all small .npz files, seeded):

* itamax_repo.npz     inputs/outputs of the reference's ``IntegerApproximatedSoftmaxFn.forward`` on qint8 tensors
                      (reference models/ITA/QAT/ITA_softmax.py:19-77) incl. edge rows.
* hw_params.npz       ``calculate_hw_params`` and ``dequantize_bias_to_int32`` of the reference harness
                      (tests/export_and_validation_W_B.py:225-245), imported with the absent PyITA module stubbed.
* qat_block_E64.npz   one converted (prepare_qat -> convert, qnnpack) ``ITALSTMNetVIT_QAT`` layer
  qat_block_E128.npz  (models/ITA_single_layer_upsample_shuffle/QAT/model.py, models/ITA/QAT/model.py), matmul2
                      patched exactly as tests/export_and_validation_W_B.py:408-423, captured with the reference's own
                      ``PyTorchDataExtractor`` / ``organize_params_by_block`` / ``translate_torch_scales_to_hw_params``:
                      int8 input, int8 weights, float biases, every scale, every per-stage int8/u8 output, and the
                      (mult, shift, add) dictionary the harness would hand to the golden model.
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)


def _import_harness():
    """tests/export_and_validation_W_B.py imports the absent PyITA at line 19; stub it."""
    for name in ("third_party", "third_party.ITA_FPGA", "third_party.ITA_FPGA.PyITA", "third_party.ITA_FPGA.PyITA.ITA"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["third_party.ITA_FPGA.PyITA.ITA"].Transformer = object
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_harness", os.path.join(REF, "tests", "export_and_validation_W_B.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def gen_itamax():
    from models.ITA.QAT.ITA_softmax import IntegerApproximatedSoftmax
    sm = IntegerApproximatedSoftmax(dim=-1)
    rng = np.random.default_rng(0)
    out = {}
    idx = 0
    for S in (16, 64, 128, 256):
        for sigma in (3, 10, 20, 60, 200):
            x = np.clip(np.rint(rng.normal(0, sigma, size=(24, S))), -128, 127).astype(np.int8)
            out[f"x{idx}"] = x
            idx += 1
    # edge rows: all equal, one +127 among -128s, all -128, all 127, ramps, sum == 256*S, max shift boundaries
    for S in (64, 128, 256):
        e = np.full((8, S), -128, np.int8)
        e[0, :] = 5
        e[1, 3] = 127
        e[2, :] = -128
        e[3, :] = 127
        e[4, :] = (np.arange(S) % 256 - 128).astype(np.int8)
        e[5, :] = np.where(np.arange(S) % 2 == 0, 10, 2).astype(np.int8)       # shift exactly 8
        e[6, :] = np.where(np.arange(S) % 2 == 0, 10, 1).astype(np.int8)       # shift exactly 9 -> 0
        e[7, :] = np.where(np.arange(S) < 2, 100, 100 - 7).astype(np.int8)
        out[f"x{idx}"] = e
        idx += 1
    res = {}
    for k, x in out.items():
        q = torch._make_per_tensor_quantized_tensor(torch.from_numpy(x), scale=0.05, zero_point=0)
        y = sm(q)
        assert y.dtype == torch.quint8 and abs(y.q_scale() - 1.0 / 255.0) < 1e-12 and y.q_zero_point() == 0
        res[k] = x
        res["y" + k[1:]] = y.int_repr().numpy().astype(np.uint8)
    np.savez_compressed(os.path.join(HERE, "itamax_repo.npz"), **res)
    print("itamax_repo.npz:", idx, "tensors")


def gen_hw_params(h):
    rng = np.random.default_rng(1)
    scales = np.concatenate([np.exp(rng.uniform(np.log(1e-6), np.log(4.0), size=400)),
                             [0.5, 0.25, 1.0, 2.0 ** -10, 0.99999999 * 2.0 ** -7, 255.9 / 256 / 64, 0.0]])
    ms = np.array([h.calculate_hw_params(float(s)) for s in scales], dtype=np.int64)
    bias = rng.normal(0, 0.3, size=(5, 64)).astype(np.float32)
    s_act = np.array([0.02, 0.013, 0.2, 0.0031, 0.05])
    s_w = np.array([0.004, 0.0021, 0.01, 0.0007, 0.003])
    b32 = np.stack([h.dequantize_bias_to_int32({"bias": bias[i], "input": [{"scale": float(s_act[i])}],
                                                "weight": {"scale": float(s_w[i]), "q_value": np.zeros((64, 1))}})
                    for i in range(5)])
    np.savez_compressed(os.path.join(HERE, "hw_params.npz"), scales=scales, mult=ms[:, 0], shift=ms[:, 1],
                        bias=bias, s_act=s_act, s_w=s_w, bias_i32=b32)
    print("hw_params.npz:", len(scales), "scales; mult==256 cases:", int((ms[:, 0] == 256).sum()))


def gen_qat_block(h, variant: str):
    torch.manual_seed(0)
    torch.backends.quantized.engine = "qnnpack"
    from models.ITA.QAT.layers import ita_symmetric_qconfig
    if variant == "E64":
        from models.ITA_single_layer_upsample_shuffle.QAT.model import ITALSTMNetVIT_QAT
        m = ITALSTMNetVIT_QAT(num_layers=1)
    else:
        from models.ITA.QAT.model import ITALSTMNetVIT_QAT
        m = ITALSTMNetVIT_QAT()
    # random-init weights are tiny; scale them up so that the int8 ranges are exercised like a trained model
    with torch.no_grad():
        for blk in list(m.attention_blocks) + list(m.ffn_blocks):
            for lin in blk.modules():
                if isinstance(lin, torch.nn.Linear):
                    lin.weight.mul_(3.0)
                    lin.bias.normal_(0, 0.3)
    m.attention_blocks.qconfig = ita_symmetric_qconfig
    m.ffn_blocks.qconfig = ita_symmetric_qconfig
    mp = torch.ao.quantization.prepare_qat(m.train())
    for _ in range(3):
        mp([torch.rand(2, 1, 60, 90), torch.rand(2, 1), torch.rand(2, 4), None])
    mc = torch.ao.quantization.convert(mp.eval())
    for block in mc.attention_blocks:                       # tests/export_and_validation_W_B.py:408-423
        mm2 = block.matmul2
        block.matmul2.matmul = (lambda s, z: (lambda a, b: h.custom_quantized_matmul(a, b, output_scale=s,
                                                                                     output_zero_point=z)))(mm2.scale, mm2.zero_point)
    B = 2
    dummy = [torch.rand(B, 1, 60, 90), torch.randn(B, 1), torch.randn(B, 4), None]
    ext = h.PyTorchDataExtractor(mc)
    # matmul1's raw output is needed as well (the harness reads it as the softmax hook's input)
    flat = ext.extract_and_capture(dummy)
    org = h.organize_params_by_block(flat)
    n_blocks = len(org["attention"])
    save = {"n_blocks": n_blocks, "B": B}
    for bi in sorted(org["attention"].keys()):
        at, ff = org["attention"][bi], org["ffn"].get(bi)
        hw = h.translate_torch_scales_to_hw_params(at, ff if ff is not None else {}, H=1)
        pre = f"b{bi}_"
        save[pre + "x"] = at["q_proj"]["input"][0]["q_value"].astype(np.int8)
        save[pre + "s_x"] = at["q_proj"]["input"][0]["scale"]
        for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
            save[pre + nm + "_w"] = at[nm]["weight"]["q_value"].astype(np.int8)      # torch layout [out,in]
            save[pre + nm + "_sw"] = at[nm]["weight"]["scale"]
            save[pre + nm + "_bias"] = at[nm]["bias"].astype(np.float32)
            save[pre + nm + "_sin"] = at[nm]["input"][0]["scale"]
            save[pre + nm + "_sout"] = at[nm]["output_scale"]
            save[pre + nm + "_out"] = at[nm]["output"]["q_value"].astype(np.int8)
            save[pre + nm + "_bias_i32"] = h.dequantize_bias_to_int32(at[nm])
        save[pre + "qk_out"] = at["qk_matmul"]["output"]["q_value"].astype(np.int8)
        save[pre + "qk_sout"] = at["qk_matmul"]["output"]["scale"]
        save[pre + "softmax_out"] = at["softmax"]["output"]["q_value"].astype(np.uint8)
        save[pre + "softmax_sout"] = at["softmax"]["output"]["scale"]
        save[pre + "av_out"] = at["av_matmul"]["output"]["q_value"].astype(np.int8)
        save[pre + "av_sout"] = at["av_matmul"]["output"]["scale"]
        for st in ("q_proj", "k_proj", "v_proj", "qk_matmul", "av_matmul", "out_proj"):
            save[pre + "hw_" + st] = np.array([hw[st]["mult"][0], hw[st]["shift"][0], hw[st]["add"][0]], np.int64)
        if ff is not None and "fc1" in ff:
            save[pre + "ffn_x"] = ff["fc1"]["input"][0]["q_value"].astype(np.int8)
            for nm in ("fc1", "fc2"):
                save[pre + nm + "_w"] = ff[nm]["weight"]["q_value"].astype(np.int8)
                save[pre + nm + "_sw"] = ff[nm]["weight"]["scale"]
                save[pre + nm + "_bias"] = ff[nm]["bias"].astype(np.float32)
                save[pre + nm + "_sin"] = ff[nm]["input"][0]["scale"]
                save[pre + nm + "_sout"] = ff[nm]["output_scale"]
                save[pre + nm + "_out"] = ff[nm]["output"]["q_value"].astype(np.int8)
                save[pre + nm + "_bias_i32"] = h.dequantize_bias_to_int32(ff[nm])
                save[pre + "hw_" + nm] = np.array([hw[nm]["mult"][0], hw[nm]["shift"][0], hw[nm]["add"][0]], np.int64)
            if "relu" in ff and "output" in ff["relu"]:
                save[pre + "relu_out"] = ff["relu"]["output"]["q_value"].astype(np.int8)
                save[pre + "relu_sout"] = ff["relu"]["output"]["scale"]
            if "relu" in hw:
                save[pre + "hw_relu"] = np.array([hw["relu"]["mult"], hw["relu"]["shift"], hw["relu"]["add"]], np.int64)
            save[pre + "ffn_modules"] = np.array([f"{k}:{v.get('type')}" for k, v in ff.items()])
    # float boundary of the operator (reference models/ITA/QAT/layers.py:103,126): input float -> quant, output dequant
    blk0 = mc.attention_blocks[0]
    save["quant_scale"] = float(blk0.quant.scale)
    save["out_scale"] = float(blk0.out_proj.scale)
    np.savez_compressed(os.path.join(HERE, f"qat_block_{variant}.npz"), **save)
    print(f"qat_block_{variant}.npz: blocks={n_blocks}",
          {k: (v.shape if hasattr(v, 'shape') and getattr(v, 'ndim', 0) else v) for k, v in save.items() if k.startswith('b0_hw')})


def gen_qat_block_h4(h):
    """BASELINE config 3 as the reference module: ``ITASelfAttention_QAT(embed_dim=128, proj_dim=256, num_heads=4)``
    (models/ITA/QAT/layers.py:79-127) on 256 tokens, standalone: prepare_qat -> calibrate -> convert (qnnpack), matmul2 patched as
    tests/export_and_validation_W_B.py:408-423, every submodule output captured by forward hooks.  Pins the head split
    (reshape(B,N,H,hd).permute, layers.py:110-113), the context merge and the single out_proj (layers.py:120-123).
    Keys follow qat_block_*.npz (prefix b0_), stage outputs in the module's own layouts: q/k/v_proj_out [B,N,H*hd],
    qk_out / softmax_out [B,H,N,N], av_out [B,H,N,hd], out_proj_out [B,N,E]."""
    torch.manual_seed(4)
    torch.backends.quantized.engine = "qnnpack"
    from models.ITA.QAT.layers import ITASelfAttention_QAT, ita_symmetric_qconfig
    E, PT, H, N, B = 128, 256, 4, 256, 2
    m = ITASelfAttention_QAT(E, PT, H)
    with torch.no_grad():
        for lin in (m.q_proj, m.k_proj, m.v_proj, m.out_proj):
            lin.weight.mul_(3.0)
            lin.bias.normal_(0, 0.3)
    m.qconfig = ita_symmetric_qconfig
    mp = torch.ao.quantization.prepare_qat(m.train())
    g = torch.Generator().manual_seed(9)
    for _ in range(3):
        mp(torch.randn(B, N, E, generator=g))
    mc = torch.ao.quantization.convert(mp.eval())
    mm2 = mc.matmul2
    mc.matmul2.matmul = (lambda s, z: (lambda a, b: h.custom_quantized_matmul(a, b, output_scale=s, output_zero_point=z)))(mm2.scale,
                                                                                                                         mm2.zero_point)
    assert mc.num_heads == H and mc.head_dim == PT // H
    cap = {}

    def hook(name):
        def f(mod, inp, out):
            cap[name + "_in"] = inp[0].detach().clone()
            cap[name + "_out"] = out.detach().clone()
        return f

    hs = [getattr(mc, n).register_forward_hook(hook(n)) for n in ("quant", "q_proj", "k_proj", "v_proj", "out_proj", "custom_softmax")]
    x = torch.randn(B, N, E, generator=g)
    with torch.no_grad():
        y = mc(x)
    for k in hs:
        k.remove()
    for n in ("quant", "q_proj", "k_proj", "v_proj", "out_proj", "custom_softmax"):
        assert cap[n + "_out"].is_quantized and cap[n + "_out"].q_zero_point() == 0, n
    pre = "b0_"
    xq = cap["quant_out"]
    save = {"H": H, "B": B, "x_float": x.numpy(), "y_float": y.numpy(), pre + "x": xq.int_repr().numpy(), pre + "s_x": xq.q_scale()}
    for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
        lin = getattr(mc, nm)
        w = lin.weight()
        assert w.qscheme() == torch.per_tensor_symmetric or w.q_zero_point() == 0
        entry = {"bias": lin.bias().detach().numpy(), "input": [{"scale": cap[nm + "_in"].q_scale()}],
                 "weight": {"scale": w.q_scale(), "q_value": w.int_repr().numpy()}}
        save[pre + nm + "_w"] = w.int_repr().numpy().astype(np.int8)           # torch layout [out,in]
        save[pre + nm + "_sw"] = w.q_scale()
        save[pre + nm + "_bias"] = lin.bias().detach().numpy().astype(np.float32)
        save[pre + nm + "_sin"] = cap[nm + "_in"].q_scale()
        save[pre + nm + "_sout"] = float(lin.scale)
        save[pre + nm + "_out"] = cap[nm + "_out"].int_repr().numpy().astype(np.int8)
        save[pre + nm + "_bias_i32"] = h.dequantize_bias_to_int32(entry)
        save[pre + "hw_" + nm] = np.array([*h.calculate_hw_params(cap[nm + "_in"].q_scale() * w.q_scale() / float(lin.scale)), 0], np.int64)
    logits, probs, ctx = cap["custom_softmax_in"], cap["custom_softmax_out"], cap["out_proj_in"]
    assert probs.dtype == torch.quint8 and logits.dtype == torch.qint8
    save[pre + "qk_out"] = logits.int_repr().numpy().astype(np.int8)                         # [B,H,N,N]
    save[pre + "qk_sout"] = logits.q_scale()
    save[pre + "softmax_out"] = probs.int_repr().numpy().astype(np.uint8)
    save[pre + "softmax_sout"] = probs.q_scale()
    # context as out_proj sees it: [B,N,H*hd] (layers.py:121) -> per head [B,H,N,hd]
    save[pre + "av_out"] = np.ascontiguousarray(ctx.int_repr().numpy().reshape(B, N, H, PT // H).transpose(0, 2, 1, 3)).astype(np.int8)
    save[pre + "av_sout"] = ctx.q_scale()
    save[pre + "hw_qk_matmul"] = np.array([*h.calculate_hw_params(float(mc.q_proj.scale) * float(mc.k_proj.scale) / logits.q_scale()), 0], np.int64)
    save[pre + "hw_av_matmul"] = np.array([*h.calculate_hw_params(probs.q_scale() * float(mc.v_proj.scale) / ctx.q_scale()), 0], np.int64)
    save["quant_scale"] = float(mc.quant.scale)
    save["out_scale"] = float(mc.out_proj.scale)
    assert abs(save[pre + "qk_sout"] - float(mc.matmul1.scale)) < 1e-12 and abs(save[pre + "av_sout"] - float(mc.matmul2.scale)) < 1e-12
    np.savez_compressed(os.path.join(HERE, "qat_block_H4_S256.npz"), **save)
    torch.save(mc.state_dict(), os.path.join(HERE, "qat_state_H4_S256.pt"))
    sat = {k: float((np.abs(save[pre + k].astype(np.int32)) >= 127).mean()) for k in ("q_proj_out", "qk_out", "av_out", "out_proj_out")}
    print("qat_block_H4_S256.npz: scales", {k: round(float(save[pre + k]), 5) for k in ("s_x", "q_proj_sout", "qk_sout", "av_sout", "out_proj_sout")},
          "saturated share", sat, "softmax nonzero", float((save[pre + "softmax_out"] > 0).mean()))


def gen_quantize():
    """QuantStub at the operator's float boundary (reference models/ITA/QAT/layers.py:103): torch.quantize_per_tensor over
    random floats and awkward scales (ties, saturation, denormal-ish scales)."""
    rng = np.random.default_rng(7)
    scales = np.array([0.5, 0.1, 0.05, 1.0 / 255.0, 0.0123456789, 3.3e-3, 7.77e-5, 1.9, 0.024999999, 2.0 ** -7], np.float64)
    xs, qs = [], []
    for s in scales:
        x = (rng.standard_normal(4096) * s * 60).astype(np.float32)
        x[:64] = (np.arange(-32, 32) + 0.5).astype(np.float32) * np.float32(s)          # exact ties
        x[64:72] = np.array([1e30, -1e30, 127.5 * s, -128.5 * s, 127.49 * s, -128.51 * s, 0.0, -0.0], np.float32)
        q = torch.quantize_per_tensor(torch.from_numpy(x), float(s), 0, torch.qint8).int_repr().numpy()
        xs.append(x); qs.append(q)
    np.savez_compressed(os.path.join(HERE, "quantize.npz"), scales=scales, x=np.stack(xs), q=np.stack(qs))
    print("quantize.npz:", len(scales), "scales x 4096 floats")


def gen_model(h):
    """Full converted ITALSTMNetVIT_QAT (E=64, one layer) forward, every block boundary captured: tokenizer output,
    attention in/out, residual + LayerNorm, FFN in/out, fusion, decoder, LSTM, head (reference
    models/ITA_single_layer_upsample_shuffle/QAT/model.py:93-132).  Float parameters outside the int8 blocks come from
    tests/golden/seeded.py (regenerated by name on the test side); the converted blocks' state-dict entries are saved with
    torch.save exactly as a converted checkpoint holds them (tests/export_and_validation_W_B.py:387-392)."""
    import seeded
    torch.manual_seed(1)
    torch.backends.quantized.engine = "qnnpack"
    from models.ITA.QAT.layers import ita_symmetric_qconfig
    from models.ITA_single_layer_upsample_shuffle.QAT.model import ITALSTMNetVIT_QAT
    m = ITALSTMNetVIT_QAT(num_layers=1)
    with torch.no_grad():
        for name, prm in m.named_parameters():
            if seeded.is_float_param(name):
                prm.copy_(torch.from_numpy(seeded.float_param(name, prm.shape)))
        for blk in list(m.attention_blocks) + list(m.ffn_blocks):
            for lin in blk.modules():
                if isinstance(lin, torch.nn.Linear):
                    lin.weight.mul_(3.0)
                    lin.bias.normal_(0, 0.3)
    m.attention_blocks.qconfig = ita_symmetric_qconfig
    m.ffn_blocks.qconfig = ita_symmetric_qconfig
    mp = torch.ao.quantization.prepare_qat(m.train())
    g = torch.Generator().manual_seed(5)
    for _ in range(4):
        mp([torch.rand(4, 1, 60, 90, generator=g), torch.rand(4, 1, generator=g), torch.rand(4, 4, generator=g), None])
    mc = torch.ao.quantization.convert(mp.eval())
    for block in mc.attention_blocks:
        mm2 = block.matmul2
        block.matmul2.matmul = (lambda s, z: (lambda a, b: h.custom_quantized_matmul(a, b, output_scale=s,
                                                                                     output_zero_point=z)))(mm2.scale, mm2.zero_point)
    B = 3
    img = torch.rand(B, 1, 60, 90, generator=g)
    vel = torch.rand(B, 1, generator=g) * 10
    quat = torch.randn(B, 4, generator=g)
    cap = {}

    def hook(name):
        def f(mod, inp, out):
            o = out[0] if isinstance(out, tuple) else out
            cap[name + "_in"] = inp[0].detach().clone() if isinstance(inp[0], torch.Tensor) else None
            cap[name + "_out"] = o.detach().clone() if isinstance(o, torch.Tensor) else o
        return f

    hs = [mc.tokenizer.register_forward_hook(hook("tokenizer")), mc.attention_blocks[0].register_forward_hook(hook("attn")),
          mc.norms1[0].register_forward_hook(hook("norm1")), mc.ffn_blocks[0].register_forward_hook(hook("ffn")),
          mc.norms2[0].register_forward_hook(hook("norm2")), mc.down_sample.register_forward_hook(hook("down")),
          mc.decoder.register_forward_hook(hook("decoder")), mc.nn_fc2.register_forward_hook(hook("fc"))]
    with torch.no_grad():
        out, (hn, cn) = mc([img.clone(), vel.clone(), quat.clone(), None])
        for x in hs:
            x.remove()
        # second step carrying the LSTM state, as the deployed loop does (samples/.../main.cpp:217-224)
        out2, (hn2, cn2) = mc([img.flip(0).clone(), vel.clone(), quat.clone(), (hn, cn)])
    # per-block capture in the same key format as qat_block_*.npz
    ext = h.PyTorchDataExtractor(mc)
    flat = ext.extract_and_capture([img.clone(), vel.clone(), quat.clone(), None])
    org = h.organize_params_by_block(flat)
    at, ff = org["attention"][0], org["ffn"][0]
    hw = h.translate_torch_scales_to_hw_params(at, ff, H=1)
    save = {"img": img.numpy(), "vel": vel.numpy(), "quat": quat.numpy(), "out": out.numpy(), "hn": hn.numpy(), "cn": cn.numpy(),
            "out2": out2.numpy(), "hn2": hn2.numpy(), "cn2": cn2.numpy()}
    for k in ("tokenizer_out", "attn_in", "attn_out", "norm1_in", "norm1_out", "ffn_in", "ffn_out", "norm2_in", "norm2_out",
              "down_in", "down_out", "decoder_out", "fc_out"):
        save[k] = cap[k].numpy().astype(np.float32)
    pre = "b0_"
    save[pre + "x"] = at["q_proj"]["input"][0]["q_value"].astype(np.int8)
    save[pre + "s_x"] = at["q_proj"]["input"][0]["scale"]
    for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
        save[pre + nm + "_w"] = at[nm]["weight"]["q_value"].astype(np.int8)
        save[pre + nm + "_sw"] = at[nm]["weight"]["scale"]
        save[pre + nm + "_bias"] = at[nm]["bias"].astype(np.float32)
        save[pre + nm + "_sin"] = at[nm]["input"][0]["scale"]
        save[pre + nm + "_sout"] = at[nm]["output_scale"]
        save[pre + nm + "_out"] = at[nm]["output"]["q_value"].astype(np.int8)
        save[pre + nm + "_bias_i32"] = h.dequantize_bias_to_int32(at[nm])
    save[pre + "qk_out"] = at["qk_matmul"]["output"]["q_value"].astype(np.int8)
    save[pre + "qk_sout"] = at["qk_matmul"]["output"]["scale"]
    save[pre + "softmax_out"] = at["softmax"]["output"]["q_value"].astype(np.uint8)
    save[pre + "softmax_sout"] = at["softmax"]["output"]["scale"]
    save[pre + "av_out"] = at["av_matmul"]["output"]["q_value"].astype(np.int8)
    save[pre + "av_sout"] = at["av_matmul"]["output"]["scale"]
    for st in ("q_proj", "k_proj", "v_proj", "qk_matmul", "av_matmul", "out_proj"):
        save[pre + "hw_" + st] = np.array([hw[st]["mult"][0], hw[st]["shift"][0], hw[st]["add"][0]], np.int64)
    save[pre + "ffn_x"] = ff["fc1"]["input"][0]["q_value"].astype(np.int8)
    for nm in ("fc1", "fc2"):
        save[pre + nm + "_w"] = ff[nm]["weight"]["q_value"].astype(np.int8)
        save[pre + nm + "_sw"] = ff[nm]["weight"]["scale"]
        save[pre + nm + "_bias"] = ff[nm]["bias"].astype(np.float32)
        save[pre + nm + "_sin"] = ff[nm]["input"][0]["scale"]
        save[pre + nm + "_sout"] = ff[nm]["output_scale"]
        save[pre + nm + "_out"] = ff[nm]["output"]["q_value"].astype(np.int8)
        save[pre + nm + "_bias_i32"] = h.dequantize_bias_to_int32(ff[nm])
        save[pre + "hw_" + nm] = np.array([hw[nm]["mult"][0], hw[nm]["shift"][0], hw[nm]["add"][0]], np.int64)
    save[pre + "relu_out"] = ff["relu"]["output"]["q_value"].astype(np.int8)
    np.savez_compressed(os.path.join(HERE, "qat_model_E64.npz"), **save)
    # converted checkpoint entries of the int8 blocks (what torch.save(model.state_dict()) holds for them)
    sd = {k: v for k, v in mc.state_dict().items() if not seeded.is_float_param(k)}
    torch.save(sd, os.path.join(HERE, "qat_state_E64.pt"))
    print("qat_model_E64.npz: out", out.numpy().round(4).tolist(), "| qat_state_E64.pt keys:", len(sd))
    print("   state keys sample:", list(sd.keys())[:12])


if __name__ == "__main__":
    h = _import_harness()
    sys.path.insert(0, HERE)
    if len(sys.argv) > 1 and sys.argv[1] == "--h4":      # the multi-head fixture added in round 3 (the others stay as committed)
        gen_qat_block_h4(h)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "--new":     # only the fixtures added in round 2 (the others stay as committed)
        gen_quantize()
        gen_model(h)
        sys.exit(0)
    gen_itamax()
    gen_hw_params(h)
    gen_qat_block(h, "E64")
    gen_qat_block(h, "E128")
    gen_qat_block_h4(h)
