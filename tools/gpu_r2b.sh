#!/bin/bash
# round 2, GPU call B: parity suite, modes on synthetic + fixture params, wait-hint variants, bench lines, ncu launch list + full capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2b_pytest.log
tail -4 gpurun_out/r2b_pytest.log
echo "== modes, this build"; python tools/mode_bench.py 2>&1 | tee gpurun_out/r2b_modes.log
for v in 2000 20000; do echo "== wait hint $v ns"; ITA_B200_LIB=$PWD/build_tools/csrc_hint$v/libita_b200.so python tools/mode_bench.py --quick 2>&1 | tee gpurun_out/r2b_modes_hint$v.log; done
for c in c2 c2_e128 c2_ffn c2_stack; do
  python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2b_bench_$c.json 2> gpurun_out/r2b_bench_$c.err; echo "bench $c exit $?"; cut -c1-250 gpurun_out/r2b_bench_$c.json
done
# ncu: launch list of the default bench command, then one full capture of the headline kernel
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2b_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2b_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2b_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ita_fused -s 4 -c 1 -o gpurun_out/r2b_prof python bench.py --steps 5 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2b_ncu2.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/r2b_prof.ncu-rep
