"""Developer tool: DRAM bytes per launch of the dominant kernel from an `ncu --set full` capture of THIS build, written to
profiles/roofline_traffic.json keyed by bench config and stamped with the build id (bench.py reports `roofline.traffic` only
when the stamp matches the running build).

    ncu -i capture.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_traffic.py <config> raw.csv <kernel substring> <frames per launch> [more kernel substrings ...]
"""
import csv, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
cfg, path, frames = sys.argv[1], sys.argv[2], int(sys.argv[4])
names = [sys.argv[3]] + sys.argv[5:]
rows = list(csv.reader(open(path)))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
units = rows[1]
def val(r, name):
    v = float(r[col[name]].replace(",", ""))
    u = units[col[name]].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)
out = {"build_id": bench.build_id(), "frames_per_launch": frames, "kernels": {}, "source": f"{os.path.basename(path)} (ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum)"}
tot = 0
for nm in names:
    sel = [r for r in rows[2:] if nm in r[col["Kernel Name"]]]
    if not sel:
        raise SystemExit(f"no launch of {nm} in {path}")
    r = sel[-1]          # the last captured launch (steady state)
    rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
    out["kernels"][r[col["Kernel Name"]][:80]] = {"dram_bytes_read": rd, "dram_bytes_write": wr, "duration_us": val(r, "gpu__time_duration.sum") if "gpu__time_duration.sum" in col else None}
    tot += rd + wr
out["dram_bytes_per_launch"] = tot
p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
try:
    doc = json.load(open(p))
    if "dram_bytes_per_launch" in doc:       # round-1 single-entry format
        doc = {"r1_c2": doc}
except Exception:
    doc = {}
doc[cfg] = out
json.dump(doc, open(p, "w"), indent=1)
print(cfg, json.dumps(out)[:400])
