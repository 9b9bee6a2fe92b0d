"""Developer tool: per-opcode executed warp-instruction counts and stall samples from `ncu --page source --csv` (SASS view)."""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
frames = float(sys.argv[2]) if len(sys.argv) > 2 else 4096.0
hdr = rows[1]
ci, ce, cs = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = {h: i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h}
ops = collections.Counter(); samp = collections.Counter(); st = collections.Counter()
tot = 0
for r in rows[2:]:
    if len(r) <= ce or not r[ce]:
        continue
    src = r[ci].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+(\.[A-Z0-9_]+)?)", src)
    op = m.group(2) if m else src.split()[0]
    if not op.startswith(("IMAD.HI", "IMAD.MOV", "IMAD.WIDE", "LDS", "STS", "STG", "SYNCS", "BAR")):
        op = op.split(".")[0]
    n = int(r[ce]); ops[op] += n; tot += n
    s = int(r[cs] or 0); samp[op] += s
    for h, i in stall_cols.items():
        if r[i]:
            st[h] += int(r[i])
print(f"total warp-instr {tot}  per frame {tot / frames:.0f}")
print("op            per-frame   samples%")
ts = sum(samp.values())
for op, n in ops.most_common(28):
    print(f"{op:12s} {n / frames:9.1f}  {100.0 * samp[op] / ts:6.1f}")
print("stall samples:", {k: round(100.0 * v / sum(st.values()), 1) for k, v in st.most_common(10)})
