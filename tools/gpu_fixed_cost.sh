#!/bin/bash
# fixed cost of one launch: kernel time against the number of full rounds (CSI_TC_DEBUG bits via DEBUGS)
mkdir -p gpurun_out; rm -f gpurun_out/fixed_cost.txt
for d in ${DEBUGS:-0 16}; do
for u in ${UNITS:-1 2 3 4 6 8 27}; do
  n=$((37888 * u)); if [ $u = 27 ]; then n=1000000; fi
  CSI_TC_DEBUG=$d timeout 200 python bench.py --stars $n --steps 30 --warmup 3 --no-cpu-baseline --e2e-blocks 1 2>/dev/null | \
    python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('debug $d units $u stars $n kernel_ms', round(l['roofline']['kernel_ms'],4))" \
    >> gpurun_out/fixed_cost.txt 2>&1
done
done
cat gpurun_out/fixed_cost.txt
