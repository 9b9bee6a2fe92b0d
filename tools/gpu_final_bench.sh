#!/bin/bash
# final build: bench lines of every config on one GPU (+ the reference arms of c2 and c4)
mkdir -p gpurun_out
python bench.py > gpurun_out/final_bench_c2.json 2> gpurun_out/final_bench_c2.err; echo "c2 default exit $?"
for c in c2_e128 c2_ffn c2_stack c3 c3_pyita c4 c1; do
  timeout 600 python bench.py --config $c --steps 20 --warmup 5 > gpurun_out/final_bench_$c.json 2> gpurun_out/final_bench_$c.err; echo "$c exit $?"
done
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/final_bench_c2_reference_arm.json 2>/dev/null; echo "ref c2 exit $?"
timeout 600 python bench.py --impl reference --config c4 --steps 20 --warmup 3 > gpurun_out/final_bench_c4_reference_arm.json 2>/dev/null; echo "ref c4 exit $?"
timeout 600 python bench.py --impl reference --config c3 --steps 20 --warmup 3 > gpurun_out/final_bench_c3_reference_arm.json 2>/dev/null; echo "ref c3 exit $?"
for f in gpurun_out/final_bench_*.json; do python - "$f" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
r=d.get('roofline',{}); e=d.get('e2e',{}); s=d.get('sustained') or {}; c=d.get('cpu_baseline') or {}
print(sys.argv[1].split('final_bench_')[1][:-5], 'value %.3fM'%(d['value']/1e6), 'frac', round(r.get('frac',0),4) if r else None, 'bound', r.get('bound'), 'traffic', r.get('traffic'), 'e2e %.3fM'%((e.get('value') or 0)/1e6), 'sust %.2fM'%((s.get('value') or 0)/1e6), 'cpu', c.get('value'), 'parity', d.get('parity_spot_check'), e.get('parity_spot_check'), 'clk', (d.get('clocks') or {}).get('sm_mhz'), (d.get('clocks') or {}).get('reasons'))
PY
done
