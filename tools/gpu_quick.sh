#!/bin/bash
# Quick GPU check of a kernel change: parity tests under a short timeout, bench (gate on / off), clock trace.
mkdir -p gpurun_out; rm -f gpurun_out/quick_*.txt
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/quick_pytest.txt 2>&1
echo "pytest exit $?" >> gpurun_out/quick_pytest.txt; tail -4 gpurun_out/quick_pytest.txt
for d in 0 16; do
  CSI_TC_DEBUG=$d timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | \
    python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('debug $d kernel_ms', round(l['roofline']['kernel_ms'],4), 'e2e', round(l['e2e']['value'],1))" \
    >> gpurun_out/quick_bench.txt 2>&1
done
cat gpurun_out/quick_bench.txt
CSI_TC_DEBUG=8 timeout 300 python tools/trace_tc.py > gpurun_out/trace.txt 2>&1; tail -1 gpurun_out/trace.txt
