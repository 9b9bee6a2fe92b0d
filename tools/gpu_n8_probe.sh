#!/bin/bash
# round 2, GPU call K (8 GPUs): per-rank PCIe bandwidth vs ranks active, 8-GPU bench lines, 8-GPU batch sweep
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r2k_topo.txt 2>&1
$TR --nproc-per-node 8 --master-port 29511 tools/pcie_probe.py --bind --out gpurun_out/r2k_pcie_n8_bound.json 2>/dev/null | tee gpurun_out/r2k_pcie_n8_bound.txt
$TR --nproc-per-node 8 --master-port 29512 tools/pcie_probe.py --out gpurun_out/r2k_pcie_n8.json 2>/dev/null | tee gpurun_out/r2k_pcie_n8.txt
for c in c2 c3 c4; do
  $TR --nproc-per-node 8 --master-port 29513 bench.py --gpus 8 --config $c --steps 20 --warmup 5 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2k_bench_${c}_n8.json 2> gpurun_out/r2k_bench_${c}_n8.err; echo "bench $c n8 exit $?"; cut -c1-220 gpurun_out/r2k_bench_${c}_n8.json
done
for n in 2 4; do
  $TR --nproc-per-node $n --master-port 29514 bench.py --gpus $n --config c2 --steps 20 --warmup 5 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/r2k_bench_c2_n$n.json 2> gpurun_out/r2k_bench_c2_n$n.err; echo "bench c2 n$n exit $?"; cut -c1-220 gpurun_out/r2k_bench_c2_n$n.json
done
timeout 400 $TR --nproc-per-node 8 --master-port 29515 tools/sweep.py --config c2 --max-log2 17 --no-cpu --out gpurun_out/r2k_sweep_c2_n8.json 2>/dev/null | tee gpurun_out/r2k_sweep_c2_n8.txt | tail -22
