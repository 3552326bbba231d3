#!/bin/bash
# ncu launch list + one full capture of the potential kernel (after the plain bench has exited 0)
mkdir -p gpurun_out
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_plain.json 2>/dev/null || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:potential_tc_kernel -s 16 -c 1 \
  -o gpurun_out/prof_pot -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_pot.log 2>&1
ls -la gpurun_out/prof_pot.ncu-rep
