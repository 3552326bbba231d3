#!/bin/bash
# round 2, 2-GPU check (gpurun --gpus 2) of what the round-end scaling run launches: bench.py with its defaults
# (interleaved end-to-end chains over the in-kernel exchange) in both arms, and the 2-rank exchange test
mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node $N --master-port 29502 bench.py --gpus $N --steps 20 --warmup 3 \
  > gpurun_out/r2_bench_c4_n$N.json 2> gpurun_out/r2_bench_c4_n$N.err
echo "bench exit $?"
python -c "import sys,json; l=json.load(open('gpurun_out/r2_bench_c4_n$N.json')); print('c4 n=$N', round(l['value'],1), 'ms', round(l['ms_per_step'],4), 'e2e', round(l['e2e']['value'],1), 'chains', l['e2e']['chains'], 'single', round(l['e2e']['single_chain_queued']['value'],1), 'sync', round(l['e2e']['sync_per_step']['value'],1))" || tail -8 gpurun_out/r2_bench_c4_n$N.err
if [ -z "$SKIP_REST" ]; then
timeout 300 $TR --nproc-per-node $N --master-port 29503 bench.py --impl reference --gpus $N --steps 3 --warmup 1 \
  > gpurun_out/r2_bench_ref_c4_n$N.json 2> gpurun_out/r2_bench_ref_c4_n$N.err
echo "reference arm exit $?"; tail -c 600 gpurun_out/r2_bench_ref_c4_n$N.json
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider -rfs --timeout 350 > gpurun_out/r2_pytest_multi.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_multi.txt; tail -4 gpurun_out/r2_pytest_multi.txt
fi
