#!/bin/bash
# round 2, fifth GPU call (1 GPU): GPU suite with per-test time-outs, bench lines of the four configurations and the
# reference arm, operator bench, ncu launch list + full capture of the potential kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -rfs --timeout 500 --durations=8 > gpurun_out/r2_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_gpu.txt; tail -25 gpurun_out/r2_pytest_gpu.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1; tail -1 gpurun_out/r2_smoke.txt
for c in c4 c3 c5 c2; do
  extra="--no-cpu-baseline"; [ $c = c4 ] && extra=""
  timeout 400 python bench.py --config $c --steps 20 --warmup 3 $extra > gpurun_out/r2_bench_$c.json 2> gpurun_out/r2_bench_$c.err
  python -c "import sys,json; l=json.load(open('gpurun_out/r2_bench_$c.json')); print('$c', round(l['value'],1), 'e2e', round(l['e2e']['value'],1), 'frac', round(l['roofline']['frac'],4))"
done
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_bench_ref_c4.json 2>/dev/null; cut -c1-300 gpurun_out/r2_bench_ref_c4.json
timeout 300 python tools/op_bench.py > gpurun_out/r2_operator_bench.jsonl 2> gpurun_out/r2_operator_bench.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_plain.json 2>/dev/null && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r2_launches.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:potential_tc_kernel -s 16 -c 1 \
  -o gpurun_out/r2_prof_pot -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_pot.log 2>&1
ls -la gpurun_out/r2_prof_pot.ncu-rep
