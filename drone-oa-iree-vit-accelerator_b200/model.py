"""``ITAViTLSTM`` -- the full obstacle-avoidance model around the ITA hot path (reference
``models/ITA_single_layer_upsample_shuffle/QAT/model.py:36-132``, converted int8 form), B200 side:

    depth frame [B,1,60,90] -> tokenizer (conv 7x7 s2 + bilinear to 8x16 + LayerNorm; ``models/ITA/QAT/layers.py:25-45``)
      -> N x encoder layer: int8 ITA attention + int8 FFN with the float residual / LayerNorm stages between them
         (one C-ABI call per layer, ``ita_encoder_layer_forward``: tcgen05 kernels for the two int8 blocks)
      -> pixel-shuffle / bilinear-upsample fusion + conv 3x3 (``model.py:116-121``) -> decoder 4608 -> 512
      -> LSTM(517 -> 128, 3 layers, one step, state carried by the caller) -> fc 128 -> 3     (``model.py:124-130``)

The float pre/post stages use torch (cuDNN / cuBLAS) on the GPU -- they are the steps either side of the hot path, not the
hot path; everything between the tokenizer's LayerNorm and the fusion is libita_b200.

``from_state_dict`` reads a *converted* reference checkpoint as ``torch.save(model.state_dict())`` writes it and as the
reference harness loads it (``tests/export_and_validation_W_B.py:387-392``): ``attention_blocks.N.{q,k,v,out}_proj.
{scale,zero_point,_packed_params._packed_params}``, ``.matmul1/.matmul2.{scale,zero_point}``, ``.quant.{scale,zero_point}``,
``ffn_blocks.N.{fc1,fc2}.*``, ``ffn_blocks.N.quant.*`` plus the float tensors (tokenizer, norms, down_sample, decoder, lstm,
nn_fc2).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L
from .engine import ITAEngine
from .params import ITAQuantParams, bias_to_int32, calculate_hw_params, linear_eff_f32


def unpack_linear(sd: dict, prefix: str):
    """(int8 weight [out,in], weight scale, float bias [out], output scale, output zero-point) of one converted
    ``quantized.Linear`` from its state-dict entries.  ``_packed_params._packed_params`` is the (qweight, bias) tuple in
    a saved checkpoint, or the packed ScriptObject of a live module."""
    pp = sd[prefix + "._packed_params._packed_params"]
    if not isinstance(pp, (tuple, list)):
        pp = torch.ops.quantized.linear_unpack(pp)
    qw, bias = pp
    if qw.qscheme() not in (torch.per_tensor_affine, torch.per_tensor_symmetric):
        raise ValueError(f"{prefix}: per-tensor weight quantisation expected (ita_symmetric_qconfig), got {qw.qscheme()}")
    if int(qw.q_zero_point()) != 0:
        raise ValueError(f"{prefix}: weight zero-point must be 0 (symmetric quantisation)")
    w = qw.int_repr().cpu().numpy().astype(np.int8)
    b = np.zeros(w.shape[0], np.float32) if bias is None else bias.detach().cpu().numpy().astype(np.float32)
    return w, float(qw.q_scale()), b, float(sd[prefix + ".scale"]), int(sd[prefix + ".zero_point"])


def layer_quant_params(sd: dict, i: int, num_heads: int = 1, seq_len: int = 128, round_mode: int = L.ROUND_TORCH_RNE_F32,
                       softmax_mode: int = L.SOFTMAX_REPO_GLOBAL) -> Tuple[ITAQuantParams, Dict[str, float]]:
    """Encoder layer ``i`` of a converted checkpoint -> (int8 parameter set incl. FFN, the four stub scales).

    The harness' recipe (``tests/export_and_validation_W_B.py:247-305,479-491``) applied to state-dict entries; the fp32
    requant scales are formed as the backend forms them (``params.linear_eff_f32``)."""
    a, f = f"attention_blocks.{i}", f"ffn_blocks.{i}"
    lin = {n: unpack_linear(sd, f"{a}.{n}") for n in ("q_proj", "k_proj", "v_proj", "out_proj")}
    ffn = {n: unpack_linear(sd, f"{f}.{n}") for n in ("fc1", "fc2")} if f"{f}.fc1.scale" in sd else None
    s_x = float(sd[f"{a}.quant.scale"])
    s_a, s_o = float(sd[f"{a}.matmul1.scale"]), float(sd[f"{a}.matmul2.scale"])
    for k in (f"{a}.quant.zero_point", f"{a}.matmul1.zero_point", f"{a}.matmul2.zero_point"):
        if int(sd[k]) != 0:
            raise ValueError(f"{k} = {int(sd[k])}: the ITA path needs zero zero-points (ita_symmetric_qconfig)")
    s_q, s_k, s_v, s_out = lin["q_proj"][3], lin["k_proj"][3], lin["v_proj"][3], lin["out_proj"][3]
    P_total, E = lin["q_proj"][0].shape
    H, hd = num_heads, P_total // num_heads
    mult = np.zeros((L.NUM_STAGES, H), np.int32)
    shift = np.ones((L.NUM_STAGES, H), np.int32)
    add = np.zeros((L.NUM_STAGES, H), np.int32)
    eff = np.zeros((L.NUM_STAGES, H), np.float32)

    def put(st, e64, e32, zp):
        mult[st], shift[st] = calculate_hw_params(e64)
        eff[st], add[st] = e32, zp

    for st, n, s_in in ((L.ST_Q, "q_proj", s_x), (L.ST_K, "k_proj", s_x), (L.ST_V, "v_proj", s_x), (L.ST_O, "out_proj", s_o)):
        w, sw, b, so, zp = lin[n]
        put(st, s_in * sw / so, linear_eff_f32(s_in, sw, so), zp)
    put(L.ST_QK, s_q * s_k / s_a, np.float32(s_q * s_k / s_a), 0)
    put(L.ST_AV, (1.0 / 255.0) * s_v / s_o, np.float32((1.0 / 255.0) * s_v / s_o), 0)       # softmax scale 1/255 (ITA_softmax.py:66)

    def split_w(w):      # torch [H*hd, E] -> golden [H, E, hd]
        return np.ascontiguousarray(w.reshape(H, hd, E).transpose(0, 2, 1))

    Bo = np.zeros((H, E), np.int32)
    Bo[0] = bias_to_int32(lin["out_proj"][2], s_o * lin["out_proj"][1])
    kw, scales = {}, {"s_attn_in": s_x, "s_attn_out": s_out}
    if ffn is not None:
        s_f = float(sd[f"{f}.quant.scale"])
        for st, n, s_in in ((L.ST_FC1, "fc1", s_f), (L.ST_FC2, "fc2", ffn["fc1"][3])):
            w, sw, b, so, zp = ffn[n]
            put(st, s_in * sw / so, linear_eff_f32(s_in, sw, so), zp)
        mult[L.ST_RELU] = 0       # the quantised ReLU keeps fc1's output scale
        kw = dict(F=int(ffn["fc1"][0].shape[0]), Wff=np.ascontiguousarray(ffn["fc1"][0].T[None]),
                  Wff2=np.ascontiguousarray(ffn["fc2"][0].T[None]), Bff=bias_to_int32(ffn["fc1"][2], s_f * ffn["fc1"][1])[None],
                  Bff2=bias_to_int32(ffn["fc2"][2], ffn["fc1"][3] * ffn["fc2"][1])[None])
        scales.update({"s_ffn_in": s_f, "s_ffn_out": ffn["fc2"][3]})
    p = ITAQuantParams(S=seq_len, E=int(E), P=int(hd), H=H, Wq=split_w(lin["q_proj"][0]), Wk=split_w(lin["k_proj"][0]),
                       Wv=split_w(lin["v_proj"][0]), Wo=np.ascontiguousarray(lin["out_proj"][0].T.reshape(H, hd, E)),
                       Bq=bias_to_int32(lin["q_proj"][2], s_x * lin["q_proj"][1]).reshape(H, hd),
                       Bk=bias_to_int32(lin["k_proj"][2], s_x * lin["k_proj"][1]).reshape(H, hd),
                       Bv=bias_to_int32(lin["v_proj"][2], s_x * lin["v_proj"][1]).reshape(H, hd), Bo=Bo, mult=mult, shift=shift,
                       add=add, eff=eff, round_mode=round_mode, softmax_mode=softmax_mode,
                       head_mode=L.HEADS_TORCH if H > 1 else L.HEADS_PYITA, **kw)
    return p, scales


def reference_float_shapes(num_layers: int = 1, embed_dim: int = 64) -> Dict[str, tuple]:
    """Names (as in the reference checkpoint) and shapes of the float tensors of the model."""
    E = embed_dim
    d = {"tokenizer.conv.weight": (E, 1, 7, 7), "tokenizer.conv.bias": (E,), "tokenizer.norm.weight": (E,), "tokenizer.norm.bias": (E,),
         "down_sample.weight": (9, E // 4 + E, 3, 3), "down_sample.bias": (9,), "decoder.weight": (512, 4608), "decoder.bias": (512,),
         "nn_fc2.weight": (3, 128), "nn_fc2.bias": (3,)}
    for i in range(num_layers):
        for n in ("norms1", "norms2"):
            d[f"{n}.{i}.weight"], d[f"{n}.{i}.bias"] = (E,), (E,)
    for l in range(3):
        d[f"lstm.weight_ih_l{l}"] = (512, 517 if l == 0 else 128)
        d[f"lstm.weight_hh_l{l}"] = (512, 128)
        d[f"lstm.bias_ih_l{l}"], d[f"lstm.bias_hh_l{l}"] = (512,), (512,)
    return d


def refine_inputs(X: list) -> list:
    """Reference ``refine_inputs`` (``QAT/model.py:22-31``): default quaternion, resize the frame to 60x90."""
    if len(X) < 3 or X[2] is None:
        quat = torch.zeros((X[0].shape[0], 4), dtype=torch.float32, device=X[0].device)
        quat[:, 0] = 1
        if len(X) < 3:
            X.append(quat)
        else:
            X[2] = quat
    if X[0].shape[-2:] != (60, 90):
        X[0] = F.interpolate(X[0], size=(60, 90), mode="bilinear", align_corners=False)
    return X


class ITAViTLSTM(nn.Module):
    """Same ``forward(X) -> (out [B,3], (h, c))`` as the reference model, ``X = [img [B,1,60,90], vel [B,1], quat [B,4],
    (h, c) | None]``; CUDA only (the int8 blocks have no CPU fallback)."""

    def __init__(self, num_layers: int = 1, embed_dim: int = 64, proj_dim: int = 192, ffn_dim: int = 256, num_heads: int = 1):
        super().__init__()
        self.num_layers, self.E, self.S, self.P, self.F, self.H = num_layers, embed_dim, 128, proj_dim, ffn_dim, num_heads
        E = embed_dim
        self.tokenizer_conv = nn.Conv2d(1, E, kernel_size=7, stride=2, padding=3)
        self.tokenizer_norm = nn.LayerNorm(E)
        self.norms1 = nn.ModuleList([nn.LayerNorm(E) for _ in range(num_layers)])
        self.norms2 = nn.ModuleList([nn.LayerNorm(E) for _ in range(num_layers)])
        self.pxShuffle = nn.PixelShuffle(upscale_factor=2)
        self.up_sample = nn.Upsample(size=(16, 32), mode="bilinear", align_corners=True)
        self.down_sample = nn.Conv2d(E // 4 + E, 9, kernel_size=3, padding=1)
        self.decoder = nn.Linear(4608, 512)
        self.lstm = nn.LSTM(input_size=517, hidden_size=128, num_layers=3, dropout=0.1)
        self.nn_fc2 = nn.Linear(128, 3)
        self.layer_params: List[Optional[ITAQuantParams]] = [None] * num_layers
        self.layer_scales: List[Dict[str, float]] = [{} for _ in range(num_layers)]
        self._engines: Dict[int, list] = {}
        self._lstm_bias = None                       # b_ih + b_hh per LSTM layer, formed once per device
        self.eval()

    # ------------------------------------------------------------------------------------------------ checkpoint
    _FLOAT_MAP = {"tokenizer.conv.": "tokenizer_conv.", "tokenizer.norm.": "tokenizer_norm."}

    @classmethod
    def from_state_dict(cls, sd: dict, num_layers: Optional[int] = None, num_heads: int = 1) -> "ITAViTLSTM":
        if num_layers is None:
            num_layers = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("attention_blocks."))
        qw = sd["attention_blocks.0.q_proj._packed_params._packed_params"]
        qw = qw[0] if isinstance(qw, (tuple, list)) else torch.ops.quantized.linear_unpack(qw)[0]
        P, E = qw.shape
        Fd = sd["ffn_blocks.0.fc1._packed_params._packed_params"]
        Fd = (Fd[0] if isinstance(Fd, (tuple, list)) else torch.ops.quantized.linear_unpack(Fd)[0]).shape[0]
        m = cls(num_layers, int(E), int(P), int(Fd), num_heads)
        m.load_converted_state_dict(sd)
        return m

    def load_converted_state_dict(self, sd: dict) -> None:
        own = dict(self.named_parameters())
        missing = []
        for name, prm in own.items():
            src = name
            for ref, mine in self._FLOAT_MAP.items():
                if name.startswith(mine):
                    src = ref + name[len(mine):]
            if src not in sd:
                missing.append(src)
                continue
            with torch.no_grad():
                prm.copy_(torch.as_tensor(sd[src]).to(prm.dtype))
        if missing:
            raise KeyError(f"converted state dict lacks float tensors: {missing[:6]}{' ...' if len(missing) > 6 else ''}")
        for i in range(self.num_layers):
            self.layer_params[i], self.layer_scales[i] = layer_quant_params(sd, i, self.H, self.S)
        self._engines.clear()
        self._lstm_bias = None

    # ------------------------------------------------------------------------------------------------ forward
    def _layer_engines(self, device: torch.device):
        key = device.index if device.index is not None else torch.cuda.current_device()
        if key not in self._engines:
            self._engines[key] = self._build_engines(key)
        return self._engines[key]

    def _build_engines(self, key: int) -> list:
        """A fresh set of per-layer handles on device ``key`` (each owns its device scratch).  ``_layer_engines`` caches one set
        per device for eager calls; ``GraphedStep`` takes a private set, because a captured graph holds the raw addresses of the
        handles' scratch buffers and an eager call at a larger batch reallocates those of the shared set."""
        engs = []
        for i in range(self.num_layers):
            if self.layer_params[i] is None:
                raise RuntimeError("ITAViTLSTM has no int8 parameters: load a converted state dict first")
            eng = ITAEngine(self.layer_params[i], torch.device("cuda", key))
            sc = self.layer_scales[i]
            lp = eng.layer_params(sc["s_attn_in"], sc["s_attn_out"], sc["s_ffn_in"], sc["s_ffn_out"], self.norms1[i].weight.detach(),
                                  self.norms1[i].bias.detach(), self.norms2[i].weight.detach(), self.norms2[i].bias.detach(),
                                  self.norms1[i].eps)
            engs.append((eng, lp))
        return engs

    @torch.no_grad()
    def tokens(self, img: torch.Tensor, want_q: bool = False):
        """OverlapPatchMerging (``models/ITA/QAT/layers.py:25-45``): [B,1,60,90] (or [B,60,90]) -> [B,128,E] f32, as ONE
        libita_b200 kernel (conv 7x7 s2 + bilinear to 8x16 + LayerNorm).  ``img`` float32, or uint8 as it comes off the wire
        (scaled by 1/255 in the kernel, ``main.cpp:168``).  ``want_q=True`` also returns the int8 tokens of the first encoder
        layer's quant stub."""
        if not img.is_cuda:
            raise ValueError("ITAViTLSTM runs on CUDA tensors only (no CPU fallback)")
        if img.dtype not in (torch.uint8, torch.float32):
            img = img.float()
        img = img.reshape(-1, 60, 90).contiguous()
        B = img.shape[0]
        tok = torch.empty((B, 128, self.E), dtype=torch.float32, device=img.device)
        q, s_q = None, 0.0
        if want_q:
            q = torch.empty((B, 128, self.E), dtype=torch.int8, device=img.device)
            s_q = float(self._layer_engines(img.device)[0][1].s_attn_in)
        cw, cb = self.tokenizer_conv.weight.contiguous(), self.tokenizer_conv.bias.contiguous()
        with torch.cuda.device(img.device):
            L.check(L.load().ita_tokenizer_forward(img.data_ptr(), int(img.dtype == torch.uint8), B, self.E, cw.data_ptr(), cb.data_ptr(),
                                                   self.tokenizer_norm.weight.data_ptr(), self.tokenizer_norm.bias.data_ptr(),
                                                   float(self.tokenizer_norm.eps), tok.data_ptr(), q.data_ptr() if q is not None else None,
                                                   s_q, torch.cuda.current_stream(img.device).cuda_stream), "ita_tokenizer_forward")
        return (tok, q) if want_q else tok

    @torch.no_grad()
    def tokens_torch(self, img: torch.Tensor) -> torch.Tensor:
        """The same stage with torch ops (cuDNN conv + ATen resize + LayerNorm): the fp32 reference the tokenizer kernel is
        tested against, not used by ``forward``."""
        x = self.tokenizer_conv(img)
        x = F.interpolate(x, size=(8, 16), mode="bilinear", align_corners=False)
        return self.tokenizer_norm(x.flatten(2).transpose(1, 2)).contiguous()

    @torch.no_grad()
    def encode(self, x: torch.Tensor, engines: Optional[list] = None) -> torch.Tensor:
        """The encoder stack on f32 tokens [B,128,E] (``QAT/model.py:100-113``): libita_b200 only.  ``engines``: a private set
        from ``_build_engines`` instead of the model's cached one (``GraphedStep``)."""
        for eng, lp in (engines if engines is not None else self._layer_engines(x.device)):
            x = eng.layer(x, lp)
        return x

    @torch.no_grad()
    def head(self, x: torch.Tensor, vel: torch.Tensor, quat: torch.Tensor, hidden=None):
        """Fusion, decoder, LSTM step, output layer (``QAT/model.py:116-132``)."""
        B = x.shape[0]
        if x.is_cuda:
            # pixel-shuffle | bilinear upsample -> cat -> conv 3x3 as ONE libita_b200 kernel (the fused map never reaches HBM)
            xc = x.contiguous()
            down = torch.empty((B, 9 * 16 * 32), dtype=torch.float32, device=x.device)
            with torch.cuda.device(x.device):
                L.check(L.load().ita_fusion_conv_forward(xc.data_ptr(), B, self.E, self.down_sample.weight.contiguous().data_ptr(),
                                                         self.down_sample.bias.data_ptr(), down.data_ptr(),
                                                         torch.cuda.current_stream(x.device).cuda_stream), "ita_fusion_conv_forward")
            out = self.decoder(down)
        else:                                              # CPU (tests of the float stages): the stock modules
            x_2d = x.transpose(1, 2).reshape(B, self.E, 8, 16)
            x_fused = torch.cat([self.pxShuffle(x_2d), self.up_sample(x_2d)], dim=1)
            out = self.decoder(self.down_sample(x_fused).flatten(1))
        out_cat = torch.cat([out, vel / 10.0, quat], dim=1).unsqueeze(0)
        if not out_cat.is_cuda:
            out_lstm, h = self.lstm(out_cat, hidden)          # CPU (tests of the float stages): the stock module
            return self.nn_fc2(out_lstm.squeeze(0)), h
        # One time step of the 3-layer LSTM as GEMMs over the batch (gate order i, f, g, o as nn.LSTM; dropout is inactive in
        # eval): cuDNN's persistent RNN kernel is built for long sequences and takes ~1 ms per layer for a [1, 4096, .] step.
        xs = out_cat.squeeze(0)
        L_, Hd = self.lstm.num_layers, self.lstm.hidden_size
        if hidden is None:
            h0 = c0 = xs.new_zeros((L_, B, Hd))
        else:
            h0, c0 = hidden
        hn = xs.new_empty((L_, B, Hd))
        cn = xs.new_empty((L_, B, Hd))
        c0 = c0.contiguous()
        lib, st = L.load(), torch.cuda.current_stream(xs.device).cuda_stream
        if self._lstm_bias is None or self._lstm_bias[0].device != xs.device:
            self._lstm_bias = [(getattr(self.lstm, f"bias_ih_l{l}") + getattr(self.lstm, f"bias_hh_l{l}")).detach() for l in range(L_)]
        for l in range(L_):
            w_ih, w_hh = getattr(self.lstm, f"weight_ih_l{l}"), getattr(self.lstm, f"weight_hh_l{l}")
            gates = torch.addmm(self._lstm_bias[l], xs, w_ih.t())
            gates.addmm_(h0[l], w_hh.t())
            with torch.cuda.device(xs.device):
                L.check(lib.ita_lstm_cell_forward(gates.data_ptr(), c0[l].data_ptr(), hn[l].data_ptr(), cn[l].data_ptr(), B, Hd, st),
                        "ita_lstm_cell_forward")
            xs = hn[l]
        return self.nn_fc2(xs), (hn, cn)

    @torch.no_grad()
    def forward(self, X: list, _engines: Optional[list] = None):
        X = refine_inputs(list(X))
        img, vel, quat = X[0], X[1], X[2]
        if not img.is_cuda:
            raise ValueError("ITAViTLSTM runs on CUDA tensors only (no CPU fallback)")
        hidden = X[3] if len(X) > 3 else None
        return self.head(self.encode(self.tokens(img), _engines), vel.float(), quat.float(), hidden)


class GraphedStep:
    """One fixed-batch ``ITAViTLSTM`` step captured in a CUDA graph: the ~35 launches of a forward (tokenizer, encoder layer,
    fusion conv, decoder / LSTM GEMMs, cell kernels) become one graph launch, which is what bounds the latency of the small
    batches a serving loop sees (batch 1: ~270 us eager).  Inputs are copied into static device buffers, the LSTM state lives in
    the step's own buffers between calls (``reset_state`` zeroes it), outputs are views of static buffers (copy before the next
    call if they must survive it).  The graph holds raw device addresses of the encoder handles' scratch buffers, so the step
    owns a PRIVATE set of handles, used at this one batch size only: eager calls on the same model (``DroneServer.infer`` at any
    batch) can grow and reallocate the model's shared handles without touching what the graph replays.  ``close()`` frees them.

        step = GraphedStep(model, batch=8, device=torch.device("cuda", 0))
        cmd = step(img_u8, vel, quat)            # [8, 3]; state carried inside
    """

    def __init__(self, model: "ITAViTLSTM", batch: int, device: torch.device, image_dtype: torch.dtype = torch.uint8):
        if batch < 1:
            raise ValueError("batch must be positive")
        self.model, self.batch, self.device = model, int(batch), device
        L_, Hd = model.lstm.num_layers, model.lstm.hidden_size
        self.img = torch.zeros((batch, 60, 90), dtype=image_dtype, device=device)
        self.vel = torch.zeros((batch, 1), dtype=torch.float32, device=device)
        self.quat = torch.zeros((batch, 4), dtype=torch.float32, device=device)
        self.h = torch.zeros((L_, batch, Hd), dtype=torch.float32, device=device)
        self.c = torch.zeros_like(self.h)
        self._engines = model._build_engines(device.index if device.index is not None else torch.cuda.current_device())
        side = torch.cuda.Stream(device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):                     # warm-up off the capture: handles, scratch buffers, cuBLAS workspaces
            for _ in range(3):
                model([self.img, self.vel, self.quat, (self.h, self.c)], _engines=self._engines)
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            out, (hn, cn) = model([self.img, self.vel, self.quat, (self.h, self.c)], _engines=self._engines)
            self.h.copy_(hn)
            self.c.copy_(cn)
            self.out = out
        self.reset_state()

    def reset_state(self) -> None:
        self.h.zero_()
        self.c.zero_()

    def close(self) -> None:
        """Drop the graph, then free the private handles it replays on."""
        self.graph = None
        for eng, _ in self._engines:
            eng.close()
        self._engines = []

    @torch.no_grad()
    def __call__(self, img: torch.Tensor, vel: torch.Tensor, quat: torch.Tensor) -> torch.Tensor:
        self.img.copy_(img.reshape(self.img.shape), non_blocking=True)
        self.vel.copy_(vel.reshape(self.vel.shape), non_blocking=True)
        self.quat.copy_(quat.reshape(self.quat.shape), non_blocking=True)
        self.graph.replay()
        return self.out
