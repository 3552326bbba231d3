#!/bin/bash
# round 2, 8-GPU call (gpurun --gpus 8): 2-rank exchange / sampler test, strong-scaling bench lines of configs 4 and 5,
# the sharded device-side sampler on configs 5 and 4
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider -rfs --timeout 500 -s > gpurun_out/r2_pytest_multi.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_multi.txt; tail -6 gpurun_out/r2_pytest_multi.txt
for n in 8 ${EXTRA_N:-}; do
  for c in c4 c5; do
    timeout 500 $TR --nproc-per-node $n --master-port $((29500 + n)) bench.py --gpus $n --config $c --steps 20 --warmup 3 \
      > gpurun_out/r2_bench_${c}_n$n.json 2> gpurun_out/r2_bench_${c}_n$n.err
    python -c "import sys,json; l=json.load(open('gpurun_out/r2_bench_${c}_n$n.json')); print('$c n=$n', round(l['value'],1), 'ms', round(l['ms_per_step'],4), 'e2e', round(l['e2e']['value'],1), 'sync', round(l['e2e']['sync_per_step']['value'],1), 'frac', round(l['roofline']['frac'],4))" || tail -5 gpurun_out/r2_bench_${c}_n$n.err
  done
done
timeout 600 $TR --nproc-per-node 8 --master-port 29611 tools/nuts_bench.py --config c5 --chains 2 --warmup ${W5:-300} --samples ${S5:-200} \
  --out gpurun_out/r2_nuts_c5_n8.json > gpurun_out/r2_nuts_c5_n8.log 2>&1; tail -c 900 gpurun_out/r2_nuts_c5_n8.log
timeout 600 $TR --nproc-per-node 8 --master-port 29612 tools/nuts_bench.py --config c4 --chains 2 --warmup ${W4:-100} --samples ${S4:-100} \
  --out gpurun_out/r2_nuts_c4_n8.json > gpurun_out/r2_nuts_c4_n8.log 2>&1; tail -c 900 gpurun_out/r2_nuts_c4_n8.log
