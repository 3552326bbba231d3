#!/bin/bash
# A/B timing of two builds of libcsi.so on the SAME box: ab/libcsi_a.so vs ab/libcsi_b.so, alternating
mkdir -p gpurun_out; rm -f gpurun_out/ab.txt
for rep in 1 2 3; do
  for v in a b; do
    cp ab/libcsi_$v.so scalable_stellar_inference_b200/libcsi.so
    timeout 200 python bench.py --stars ${STARS:-1000000} --steps ${STEPS:-20} --warmup 3 --no-cpu-baseline 2>/dev/null | \
      python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('$v rep $rep kernel_ms', round(l['roofline']['kernel_ms'],4), 'clk', l['clocks'].get('sm_mhz'))" \
      >> gpurun_out/ab.txt 2>&1
  done
done
cat gpurun_out/ab.txt
