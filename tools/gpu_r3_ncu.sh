#!/bin/bash
# round 3: ncu launch list of the default bench command, full captures of the dominant kernels of c2 / c3 (traffic stamp)
mkdir -p gpurun_out
F="--no-cpu-baseline --sustained-seconds 0 --no-other-configs"
python bench.py --steps 2 --warmup 3 $F > gpurun_out/r3_plain_c2.json 2>/dev/null; echo "plain exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3_launches_c2.csv python bench.py --steps 2 --warmup 3 $F > /dev/null 2>&1; echo "launch list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ita_fused -s 3 -c 1 -o gpurun_out/r3_c2 -f python bench.py --config c2 --steps 3 --warmup 3 $F --no-e2e > gpurun_out/r3_ncu_c2.log 2>&1; echo "ncu c2 exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ita_mh -s 6 -c 2 -o gpurun_out/r3_c3 -f python bench.py --config c3 --steps 3 --warmup 3 $F --no-e2e > gpurun_out/r3_ncu_c3.log 2>&1; echo "ncu c3 exit $?"
ls -la gpurun_out/r3_*.ncu-rep
