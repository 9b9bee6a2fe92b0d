#!/bin/bash
# round 2, GPU call I: batch sweep (config 5) on one GPU, ncu full captures of the c2 and c3 kernels (traffic + source view)
mkdir -p gpurun_out
timeout 900 python tools/sweep.py --config c2 --out gpurun_out/r2i_sweep_c2_n1.json 2>&1 | tee gpurun_out/r2i_sweep_c2_n1.txt | tail -25
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ita_fused -s 6 -c 1 -o gpurun_out/r2i_c2 -f python bench.py --config c2 --steps 3 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --no-e2e > gpurun_out/r2i_ncu_c2.log 2>&1; echo "ncu c2 exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ita_mh -s 6 -c 2 -o gpurun_out/r2i_c3 -f python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --no-e2e > gpurun_out/r2i_ncu_c3.log 2>&1; echo "ncu c3 exit $?"
ls -la gpurun_out/*.ncu-rep
