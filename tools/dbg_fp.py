import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch
import ita_b200 as ita
import ita_ref as R
p = R.make_synthetic(128, 64, 192, 1, seed=5, rand_add=True)
params = ita.ITAQuantParams(S=p.S, E=p.E, P=p.P, H=p.H, Wq=p.Wq, Wk=p.Wk, Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv, Bo=p.Bo,
                            mult=p.mult, shift=p.shift, add=p.add, eff=p.eff, round_mode=p.round_mode,
                            softmax_mode=p.softmax_mode, head_mode=p.head_mode)
eng = ita.ITAEngine(params, torch.device("cuda", 0))
for nb in (1, 7, 600):
    X = R.make_input(nb, 128, 64, seed=31)
    ref = R.mha_forward(p, X)
    x = torch.from_numpy(X).cuda()
    for rep in range(2):
        y = eng.forward(x, path=ita.PATH_FUSED).cpu().numpy()
        bad = (y != ref["Y"].reshape(y.shape))
        print(f"batch {nb} rep {rep}: {bad.sum()}/{bad.size} mismatches; bad frames {np.unique(np.argwhere(bad)[:,0])[:20].tolist()}")
        if bad.any():
            f = np.argwhere(bad)[0][0]
            print("  rows with errors in first bad frame:", np.unique(np.argwhere(bad[f])[:,0])[:40].tolist())
            print("  got", y[f, 0, :8], "want", ref["Y"].reshape(y.shape)[f, 0, :8])
