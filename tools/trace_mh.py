"""Developer tool: run one k_ita_mh launch (config 3) with the -DITA_TRACE build and print CTA 0's per-position timeline.
usage: ITA_B200_LIB=build_tools/trace/libita_b200.so python tools/trace_mh.py [round_mode] [iteration]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch
import ita_b200 as ita
import ita_ref as R
sys.path.insert(0, os.path.join(ROOT, 'tests'))
from conftest import to_engine_params
rm = int(sys.argv[1]) if len(sys.argv) > 1 else 2
it = int(sys.argv[2]) if len(sys.argv) > 2 else 5
out = os.path.join(ROOT, "gpurun_out", "trace_mh.bin")
p = R.make_synthetic(256, 128, 64, 4, seed=0, round_mode=rm, head_mode=1)
eng = ita.ITAEngine(to_engine_params(ita, p), torch.device("cuda", 0))
print(eng.describe())
x = torch.from_numpy(R.make_input(2048, 256, 128, seed=1)).cuda()
for _ in range(2):
    eng.forward(x, path=ita.PATH_FUSED)
torch.cuda.synchronize()
os.environ["ITA_TRACE_FILE"] = out
eng.forward(x, path=ita.PATH_FUSED)
torch.cuda.synchronize()
t = np.fromfile(out, dtype=np.int64).reshape(-1, 9, 32, 4)
names = ["S t0", "P' t0", "AV t0", "S t1", "P' t1", "AV t1"]
mma = 16
base = t[mma, it, 0, 0]
print(f"iteration {it}: period (MMA pos 0 start -> next) = {t[mma, it + 1, 0, 0] - base} cycles")
print("pos     | MMA: start gotEmpty issued committed | EPI warp 0 (part 0): wait0 full released done | busy   | EPI warp 12 (part 3): full done")
for c, nm in enumerate(names):
    m = t[mma, it, c] - base
    e = t[0, it, c] - base
    e3 = t[12, it, c] - base
    print(f"{nm:7s} | {m[0]:6d} {m[1]:6d} {m[2]:6d} {m[3]:6d} | {e[0]:6d} {e[1]:6d} {e[2]:6d} {e[3]:6d} | {e[3] - e[1]:5d} | {e3[1]:6d} {e3[3]:6d}")
