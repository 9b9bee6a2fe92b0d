#!/bin/bash
# round 2, GPU call E: launch list of the c3 bench (k_ita_mh vs k_ita_mh_out split)
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2e_launches_c3.csv python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --sustained-seconds 0 --no-e2e > gpurun_out/r2e_ncu.log 2>&1
grep -E "k_ita_mh" gpurun_out/r2e_launches_c3.csv | awk -F'","' '{print $5, $NF}' | tail -12
