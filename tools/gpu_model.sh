#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "tokenizer or fusion or full_model or serve" > gpurun_out/model_pytest.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/model_pytest.log
timeout 300 python bench.py --config c4 --steps 10 --warmup 3 > gpurun_out/model_bench_c4.json 2> gpurun_out/model_bench_c4.err; echo "bench exit $?"; tail -3 gpurun_out/model_bench_c4.err; python -c "
import json; d=json.load(open('gpurun_out/model_bench_c4.json')); print('c4', d['value']/1e6, 'Mfps', d['ms_per_step'], 'ms', d['parity_spot_check'], 'e2e', d['e2e']['value']/1e6, d['e2e']['parity_spot_check'], 'launches', d['gpu_launches'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/model_launches_c4.csv python bench.py --config c4 --steps 2 --warmup 3 > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(l for l in open('gpurun_out/model_launches_c4.csv') if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.OrderedDict()
for r in rows[1:][-60:]:
    agg.setdefault(r[ki][:70],[]).append(float(r[vi].replace(',','')))
for k,v in agg.items(): print(f"{sum(v)/len(v)/1e3:9.1f} us x{len(v):3d}  {k}")
PY
