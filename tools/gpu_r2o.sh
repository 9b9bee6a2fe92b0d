#!/bin/bash
mkdir -p gpurun_out
for e in 0 1; do CUBLAS_EMULATE_SINGLE_PRECISION=$e timeout 300 python bench.py --config c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2o_c4_emu$e.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/r2o_c4_emu$e.json')); print('emu$e', d['value']/1e6, 'Mfps', d['ms_per_step'], 'ms', d['parity_spot_check'], 'e2e', d['e2e']['value']/1e6)"; done
CUBLAS_EMULATE_SINGLE_PRECISION=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file gpurun_out/r2o_launches_emu.csv python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(l for l in open('gpurun_out/r2o_launches_emu.csv') if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.OrderedDict()
for r in rows[1:][-50:]:
    agg.setdefault(r[ki][:80],[]).append(float(r[vi].replace(',','')))
for k,v in agg.items():
    if sum(v)/len(v) > 20000: print(f"{sum(v)/len(v)/1e3:9.1f} us x{len(v):3d}  {k}")
PY
