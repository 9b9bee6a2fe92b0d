#!/bin/bash
# final build on 8 GPUs: bench lines of c2 (2, 4, 8 GPUs), c3 and c4 (8 GPUs)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for c in c2 c3 c4; do
  $TR --nproc-per-node 8 --master-port 29513 bench.py --gpus 8 --config $c --steps 20 --warmup 5 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/final8_bench_${c}_n8.json 2> gpurun_out/final8_bench_${c}_n8.err; echo "bench $c n8 exit $?"
done
for n in 2 4; do
  $TR --nproc-per-node $n --master-port 29514 bench.py --gpus $n --config c2 --steps 20 --warmup 5 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/final8_bench_c2_n$n.json 2> gpurun_out/final8_bench_c2_n$n.err; echo "bench c2 n$n exit $?"
done
for f in gpurun_out/final8_bench_*.json; do python -c "
import json,sys; d=json.load(open('$f')); print('$f'.split('final8_bench_')[1][:-5], 'value %.1fM'%(d['value']/1e6), 'e2e %.2fM'%(d['e2e']['value']/1e6), d['parity_spot_check'], d['e2e']['parity_spot_check'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; done
