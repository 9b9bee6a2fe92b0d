"""Developer tool: per-kernel SASS opcode histogram of the built objects (cuobjdump -sass), with the Blackwell-native opcodes
(UTCIMMA = tcgen05.mma kind::i8, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor, UBLKCP = cp.async.bulk, UTCBAR =
tcgen05.commit, SYNCS = mbarrier) called out.   usage: python tools/sass_hist.py <object.o> <kernel-name regex> [...]"""
import collections, re, subprocess, sys
obj, pats = sys.argv[1], sys.argv[2:]
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
fn, hist = None, collections.defaultdict(collections.Counter)
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", ln)
    if m and fn:
        op = m.group(1)
        base = op.split(".")[0]
        hist[fn][base if base not in ("UTCIMMA", "LDTM", "UTMALDG", "UBLKCP", "UTCBAR", "SYNCS", "FFMA2", "FFMA") else op] += 1
KEY = ("UTCIMMA", "LDTM", "UTMALDG", "UBLKCP", "UTCBAR", "SYNCS", "UTCATOMSWS", "FFMA2", "I2IP", "STS", "FENCE")
for fn in sorted(hist):
    dem = subprocess.run(["cu++filt", fn], capture_output=True, text=True).stdout.strip() or fn
    if not any(re.search(p, dem) for p in pats):
        continue
    h = hist[fn]
    print(f"== {dem[:150]}\n   {sum(h.values())} SASS instructions")
    key = {k: v for k, v in h.items() if k.startswith(KEY)}
    print("   tcgen05 / TMA / mbarrier / epilogue opcodes:", ", ".join(f"{k} x{v}" for k, v in sorted(key.items())))
    print("   top opcodes:", ", ".join(f"{k} {v}" for k, v in h.most_common(14)))
