#!/bin/bash
# round 2, GPU call A: parity suite, FFMA2 exactness + issue rates, mode timings (f32x2 on/off), bench lines
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
./build_tools/fp_requant_check > gpurun_out/r2a_fpcheck.log 2>&1; tail -3 gpurun_out/r2a_fpcheck.log
./build_tools/pipe_rate_check > gpurun_out/r2a_piperate.log 2>&1; cat gpurun_out/r2a_piperate.log
echo "== modes, this build (f32x2)"; python tools/mode_bench.py 2>&1 | tee gpurun_out/r2a_modes_f32x2.log
echo "== modes, this build, ITA_FUSED_FP=1 (three-FFMA torch form)"; ITA_FUSED_FP=1 python tools/mode_bench.py 2>&1 | tee gpurun_out/r2a_modes_fp1.log
echo "== modes, -DITA_NO_F32X2"; ITA_B200_LIB=$PWD/build_tools/libita_b200.so python tools/mode_bench.py 2>&1 | tee gpurun_out/r2a_modes_scalar.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench exit $?"; cat gpurun_out/r2a_bench.json
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; cat gpurun_out/r2a_bench_ref.json
python bench.py --config c2_e128 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2a_bench_e128.json 2> gpurun_out/r2a_bench_e128.err; cat gpurun_out/r2a_bench_e128.json
python bench.py --config c2_ffn --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2a_bench_ffn.json 2> gpurun_out/r2a_bench_ffn.err; cat gpurun_out/r2a_bench_ffn.json
python bench.py --config c2_stack --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2a_bench_stack.json 2> gpurun_out/r2a_bench_stack.err; cat gpurun_out/r2a_bench_stack.json
tail -3 gpurun_out/r2a_bench.err gpurun_out/r2a_bench_e128.err gpurun_out/r2a_bench_ffn.err gpurun_out/r2a_bench_stack.err
