#!/bin/bash
# round 2, fourth GPU call: the 8-warp-epilogue kernel with register budgets inside the launch pool (library B = the
# tree's) against the round-1 kernel (A); BC emulator operator on the tensor cores; adapter tests
mkdir -p gpurun_out
cp ab/libcsi_b.so scalable_stellar_inference_b200/libcsi.so
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_b.txt 2>&1
rc=$?; echo "smoke(B) rc=$rc"; tail -2 gpurun_out/r2_smoke_b.txt
if [ $rc -ne 0 ]; then
  echo "new kernel failed its smoke test: library A for the rest of this call"
  cp ab/libcsi_a.so scalable_stellar_inference_b200/libcsi.so
else
  STEPS=20 bash tools/gpu_ab.sh > gpurun_out/r2_ab_8warp.txt 2>&1
  for c in c3 c5; do
    for v in a b; do
      cp ab/libcsi_$v.so scalable_stellar_inference_b200/libcsi.so
      timeout 200 python bench.py --config $c --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | \
        python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('$c $v kernel_ms', round(l['roofline']['kernel_ms'],4), 'value', round(l['value'],1))" >> gpurun_out/r2_ab_8warp.txt 2>&1
    done
  done
  cat gpurun_out/r2_ab_8warp.txt
  cp ab/libcsi_b.so scalable_stellar_inference_b200/libcsi.so
  timeout 300 python tools/op_bench.py > gpurun_out/r2_operator_bench.jsonl 2> gpurun_out/r2_operator_bench.err; cat gpurun_out/r2_operator_bench.jsonl | cut -c1-160
fi
timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider -rfs 2>&1 | tail -120 > gpurun_out/r2_pytest_gpu.txt
tail -12 gpurun_out/r2_pytest_gpu.txt
timeout 600 python tools/nuts_bench.py --config c3 --chains 1 --warmup 150 --samples 50 --out gpurun_out/r2_nuts_c3_quick_b.json > gpurun_out/r2_nuts_c3_quick_b.log 2>&1
tail -c 700 gpurun_out/r2_nuts_c3_quick_b.log
