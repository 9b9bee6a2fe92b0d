"""Developer tool: kernel-only frames/s of the fused kernel per rounding mode / softmax variant (C2 shape, 4096 frames),
on the seeded synthetic parameters and on the converted reference block (tests/golden/qat_block_E64.npz)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch
import ita_b200 as ita
import ita_ref as R
from bench import to_oracle
x = torch.from_numpy(R.make_input(4096, 128, 64, seed=1)).cuda()
xs = [x.clone() for _ in range(8)]
q = np.load(os.path.join(ROOT, "tests", "golden", "qat_block_E64.npz"))
cases = [("synthetic", rm, sm) for rm in (0, 1, 2) for sm in (0, 1)] + [("fixture", 0, 0), ("fixture", 2, 0)]
if len(sys.argv) > 1 and sys.argv[1] == "--quick":
    cases = [("synthetic", 0, 0), ("synthetic", 2, 0), ("fixture", 0, 0), ("fixture", 2, 0)]
for src, rm, sm in cases:
    if src == "synthetic":
        p = R.make_synthetic(128, 64, 192, 1, seed=0, round_mode=rm, softmax_mode=sm)
        params = ita.ITAQuantParams(S=p.S, E=p.E, P=p.P, H=p.H, Wq=p.Wq, Wk=p.Wk, Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv, Bo=p.Bo,
                                    mult=p.mult, shift=p.shift, add=p.add, eff=p.eff, round_mode=p.round_mode,
                                    softmax_mode=p.softmax_mode, head_mode=p.head_mode)
    else:
        params = ita.params_from_captured_block(q, "b0_", round_mode=rm, softmax_mode=sm)
        p = to_oracle(R, params)
    eng = ita.ITAEngine(params, torch.device("cuda", 0))
    for i in range(5):
        y = eng.forward(xs[i % 8], path=ita.PATH_FUSED)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40):
        y = eng.forward(xs[i % 8], path=ita.PATH_FUSED)
    e1.record(); torch.cuda.synchronize()
    ref = R.mha_forward(p, x[:8].cpu().numpy())["Y"]
    ok = np.array_equal(eng.forward(x[:8].contiguous(), path=ita.PATH_FUSED).cpu().numpy(), ref.reshape(8, 128, 64))
    print(f"{src:9s} round_mode {rm} softmax {sm}: {4096 * 40 / (e0.elapsed_time(e1) * 1e-3) / 1e6:.2f} M frames/s  bit-exact {ok}  [{eng.describe().split(';')[0]}]")
    eng.close()
