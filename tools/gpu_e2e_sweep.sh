#!/bin/bash
# end-to-end (host buffers) throughput against the block schedule / issue mode of the pipelined host path
mkdir -p gpurun_out; rm -f gpurun_out/e2e_sweep.txt
for rep in 1 2; do
for nb in ${BLOCKS:-6}; do
  for mode in "--e2e-equal --e2e-no-graph" "--e2e-no-graph" "--e2e-equal" ""; do
    timeout 200 python bench.py --steps ${STEPS:-40} --warmup 3 --no-cpu-baseline --e2e-blocks $nb $mode 2>gpurun_out/e2e_err.txt | \
      python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('rep $rep blocks $nb [$mode] e2e', round(l['e2e']['value'],1), 'value', round(l['value'],1), l['e2e'].get('cuda_graph'), l['e2e'].get('cuda_graph_error'))" \
      >> gpurun_out/e2e_sweep.txt 2>&1
  done
done
done
cat gpurun_out/e2e_sweep.txt
