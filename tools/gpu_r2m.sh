#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_head or fused or fixture or edge" 2>&1 | tail -3
timeout 300 python tools/mode_bench.py --quick 2>&1 | cut -c1-90
for c in c3; do timeout 300 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --no-e2e > gpurun_out/r2m_bench_$c.json 2> gpurun_out/r2m_bench_$c.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r2m_bench_$c.json')); print('$c', d['value']/1e6, 'Mfps', d['ms_per_step'], 'ms frac', d['roofline']['frac'], d['parity_spot_check'], d['modes']['ita_half_up']['value']/1e6)"; done
