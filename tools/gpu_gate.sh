#!/bin/bash
# gated host-buffer path: parity test, then the e2e number for CSI_HOST_GATE = 0 (kernel reads PCIe) against the
# hybrid modes 3 / 4 with CSI_HOST_GATE_DIRECT rounds read by the kernel itself
mkdir -p gpurun_out; rm -f gpurun_out/gate.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "gated or zero_copy" > gpurun_out/gate_pytest.txt 2>&1
tail -3 gpurun_out/gate_pytest.txt
for rep in 1 2 3; do
for cfg in ${CFGS:-"0 0" "4 6" "4 10" "4 14" "4 20"}; do
  set -- $cfg
  CSI_HOST_GATE=$1 CSI_HOST_GATE_DIRECT=$2 timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu-baseline 2>&1 | grep '^{' | \
    python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('gate $1 direct $2 rep $rep e2e', round(l['e2e']['value'],1), 'value', round(l['value'],1))" \
    >> gpurun_out/gate.txt 2>&1
done
done
grep "^gate" gpurun_out/gate.txt
