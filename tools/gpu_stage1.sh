#!/bin/bash
# Stage-1 GPU bring-up: hardware probe, FFMA-engine parity, then tcgen05 engine under short timeouts.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
for v in 4 0 1 2 3; do
  timeout 30 ./scalable_stellar_inference_b200/csi_probe $v >> gpurun_out/probe.log 2>&1
  echo "variant $v exit $?" >> gpurun_out/probe.log
done
cat gpurun_out/probe.log
timeout 900 python -m pytest tests -m gpu -q -x -k "ffma or operators" > gpurun_out/ffma.log 2>&1
echo "ffma exit $?" >> gpurun_out/ffma.log
tail -15 gpurun_out/ffma.log
timeout 180 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "test_ragged_sizes and tcgen05" > gpurun_out/tc_small.log 2>&1
echo "tc_small exit $?" >> gpurun_out/tc_small.log
tail -30 gpurun_out/tc_small.log
