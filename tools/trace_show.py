"""Developer tool: print the per-chunk timeline recorded by tools/trace_run.py (cycles relative to the iteration start)."""
import sys
import numpy as np
t = np.fromfile(sys.argv[1], dtype=np.int64).reshape(-1, 9, 32, 4)
NQ = 3
names = ["QK0", "QK1", "QK2", "V0", "S", "V1", "AV0", "AV1", "O"]   # schedule positions of Cfg<64,192>
it = int(sys.argv[2]) if len(sys.argv) > 2 else 4
WG = int(sys.argv[3]) if len(sys.argv) > 3 else 0
for slot in (0, 1):
    mma = 16 + slot
    base = t[mma, it, 0, 0]
    print(f"==== slot {slot}: base - slot0 base = {base - t[16, it, 0, 0]}")
    print(f"==== slot {slot} iteration {it}: frame period (MMA Q0 start -> next Q0 start) = {t[mma, it + 1, 0, 0] - base} cycles")
    print("chunk  | MMA: wait0  gotEmpty  gotDeps  committed | EPI(warp q0 of owner): wait0 full released done | dur")
    for c, nm in enumerate(names):
        m = t[mma, it, c] - base
        row = None
        for w in range(slot * 8, slot * 8 + 8):
            if t[w, it, c, 1] != 0 and (w & 3) == 0 and (w >> 2) % 2 == WG and row is None:
                row = (w, t[w, it, c] - base)
        if row is None:
            print(f"{nm:6s} | {m[0]:6d} {m[1]:6d} {m[2]:6d} {m[3]:6d} |   (S1 handled with S0)")
        else:
            w, e = row
            print(f"{nm:6s} | {m[0]:6d} {m[1]:6d} {m[2]:6d} {m[3]:6d} | w{w:2d} {e[0]:6d} {e[1]:6d} {e[2]:6d} {e[3]:6d} | {e[3] - e[1]:5d}")
