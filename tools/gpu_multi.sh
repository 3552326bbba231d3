#!/bin/bash
# bench.py on N GPUs of one box (run under `gpurun --gpus N`): tools/gpu_multi.sh N
N=$1; mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "exit $?"; tail -1 gpurun_out/bench_n$N.json | python -c "
import sys,json; l=json.loads(sys.stdin.readline()); print(l['n_gpus'], round(l['value'],1), 'ms/step', round(l['ms_per_step'],4), 'kernel', round(l['roofline']['kernel_ms'],4), l['config']['parallelism'], 'e2e', round(l['e2e']['value'],1))"
tail -2 gpurun_out/bench_n$N.err
