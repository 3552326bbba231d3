#!/bin/bash
# Timing decomposition of the tcgen05 kernel with the CSI_TC_DEBUG experiment bits (results are wrong when set;
# only the times matter): 1 = no weight streaming, 2 = no star_pre/star_mid, 4 = half of the epilogue chunks.
mkdir -p gpurun_out
for d in 0 1 2 4 3 6 7; do
  CSI_TC_DEBUG=$d timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | \
    python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('debug $d kernel_ms', round(l['roofline']['kernel_ms'],4))" \
    >> gpurun_out/experiments.txt 2>&1
done
cat gpurun_out/experiments.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:potential_tc_kernel -s 16 -c 1 \
  -o gpurun_out/prof_pot -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_pot.log 2>&1
tail -2 gpurun_out/ncu_full_pot.log | cut -c1-300
