#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_head" 2>&1 | tail -3
ITA_B200_LIB=$PWD/build_tools/trace/libita_b200.so timeout 120 python tools/trace_mh.py 2 5 2>&1 | tee gpurun_out/r2l_trace_mh.txt
for c in c3 c3_pyita; do timeout 300 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --no-e2e > gpurun_out/r2l_bench_$c.json 2> gpurun_out/r2l_bench_$c.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r2l_bench_$c.json')); print('$c', d['value']/1e6, 'Mfps', d['ms_per_step'], 'ms frac', d['roofline']['frac'], d['parity_spot_check'], d['modes']['ita_half_up']['value']/1e6)"; done
for v in base hint500 hint2000; do echo "== variant $v"; ITA_B200_LIB=$PWD/build_tools/$v/libita_b200.so timeout 300 python tools/mode_bench.py --quick 2>&1 | cut -c1-90; done
