#!/bin/bash
# round 3 (4 GPUs): BASELINE config 5 batch sweep at 2 and 4 GPUs, bench lines of c2 at 1/2/4 GPUs, the two-GPU NCCL test
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python -m pytest tests/test_multirank_nccl.py -m gpu -x -q 2>&1 | tail -2
for n in 2 4; do
  timeout 400 $TR --nproc-per-node $n --master-port 2951$n tools/sweep.py --config c2 --max-log2 17 --no-cpu --out gpurun_out/r3_sweep_c2_n$n.json 2>/dev/null > gpurun_out/r3_sweep_c2_n$n.txt; echo "sweep n=$n exit $?"
done
python bench.py --steps 20 --warmup 5 --no-other-configs > gpurun_out/r3_bench_c_n1.json 2>/dev/null; echo "bench n=1 exit $?"
for n in 2 4; do
  timeout 400 $TR --nproc-per-node $n --master-port 2952$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r3_bench_c_n$n.json 2>/dev/null; echo "bench n=$n exit $?"
done
