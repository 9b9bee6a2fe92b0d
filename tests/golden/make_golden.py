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


if __name__ == "__main__":
    h = _import_harness()
    gen_itamax()
    gen_hw_params(h)
    gen_qat_block(h, "E64")
    gen_qat_block(h, "E128")
