#!/bin/bash
# bench + clock trace of CTA 0 (no pytest): for timing experiments on a kernel change
mkdir -p gpurun_out; rm -f gpurun_out/quick_bench.txt
for d in ${DEBUGS:-0}; do
  CSI_TC_DEBUG=$d timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | \
    python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('debug $d kernel_ms', round(l['roofline']['kernel_ms'],4), 'e2e', round(l['e2e']['value'],1))" \
    >> gpurun_out/quick_bench.txt 2>&1
done
cat gpurun_out/quick_bench.txt
CSI_TC_DEBUG=8 timeout 300 python tools/trace_tc.py > gpurun_out/trace.txt 2>&1; tail -1 gpurun_out/trace.txt
