"""Developer tool: run one fused launch of the bench workload with the -DITA_TRACE build and save CTA 0's clock stamps."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch
import ita_b200 as ita
import ita_ref as R
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "trace.bin")
p = R.make_synthetic(128, 64, 192, 1, seed=0)
params = ita.ITAQuantParams(S=p.S, E=p.E, P=p.P, H=p.H, Wq=p.Wq, Wk=p.Wk, Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv, Bo=p.Bo,
                            mult=p.mult, shift=p.shift, add=p.add, eff=p.eff, round_mode=p.round_mode,
                            softmax_mode=p.softmax_mode, head_mode=p.head_mode)
eng = ita.ITAEngine(params, torch.device("cuda", 0))
x = torch.from_numpy(R.make_input(4096, 128, 64, seed=1)).cuda()
for _ in range(3):
    eng.forward(x, path=ita.PATH_FUSED)
torch.cuda.synchronize()
os.environ["ITA_TRACE_FILE"] = out
eng.forward(x, path=ita.PATH_FUSED)
torch.cuda.synchronize()
print("trace written to", out)
