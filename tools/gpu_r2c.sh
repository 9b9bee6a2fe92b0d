#!/bin/bash
# round 2, GPU call C: new chunk schedule (V chunk before the logits, per-position dependencies), spin-loop unroll
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2c_pytest.log
tail -5 gpurun_out/r2c_pytest.log
echo "== modes, this build"; timeout 300 python tools/mode_bench.py 2>&1 | tee gpurun_out/r2c_modes.log
ITA_B200_LIB=$PWD/build_tools/trace/libita_b200.so timeout 120 python tools/trace_run.py gpurun_out/r2c_trace.bin && python tools/trace_show.py gpurun_out/r2c_trace.bin 4 > gpurun_out/r2c_timeline.txt; cat gpurun_out/r2c_timeline.txt
timeout 300 python bench.py --config c2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2c_bench_c2.json 2> gpurun_out/r2c_bench_c2.err; echo "bench exit $?"; cut -c1-200 gpurun_out/r2c_bench_c2.json
timeout 300 python bench.py --config c2_e128 --steps 20 --warmup 5 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2c_bench_c2_e128.json 2> gpurun_out/r2c_bench_c2_e128.err; cut -c1-200 gpurun_out/r2c_bench_c2_e128.json
