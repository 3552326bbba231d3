#!/bin/bash
# gated host-buffer path: e2e number for CSI_HOST_GATE = 0 (kernel reads PCIe) / 1 / 2 and the timing experiments
mkdir -p gpurun_out; rm -f gpurun_out/gate.txt
for rep in 1 2; do
for cfg in "0 0" "1 0" "1 1" "1 2" "1 3" "2 1" "2 3"; do
  set -- $cfg
  CSI_HOST_GATE=$1 CSI_HOST_GATE_EXP=$2 CSI_SKIP_E2E_CHECK=1 timeout 200 python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>&1 | grep '^{' | \
    python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('gate $1 exp $2 rep $rep e2e', round(l['e2e']['value'],1), 'value', round(l['value'],1))" \
    >> gpurun_out/gate.txt 2>&1
done
done
cat gpurun_out/gate.txt
