#!/bin/bash
# round 2, GPU call D: multi-head S=256 tcgen05 kernels, first run
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_head" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest exit $?"
tail -25 gpurun_out/r2d_pytest.log
timeout 300 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2d_bench_c3.json 2> gpurun_out/r2d_bench_c3.err; echo "bench exit $?"; cut -c1-400 gpurun_out/r2d_bench_c3.json; tail -5 gpurun_out/r2d_bench_c3.err
