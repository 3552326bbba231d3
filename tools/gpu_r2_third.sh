#!/bin/bash
# round 2, third GPU call: 8-warp-epilogue kernel (ab/libcsi_b.so = the tree's library) against the previous one
# (ab/libcsi_a.so): smoke with a short timeout first (a synchronisation bug would hang), then same-box A/B, the GPU
# suite, and the sampler run that replicates the reference's recorded 1000-star hierarchical run
mkdir -p gpurun_out
cp ab/libcsi_b.so scalable_stellar_inference_b200/libcsi.so
timeout 180 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_b.txt 2>&1
rc=$?; echo "smoke(B) rc=$rc"; tail -2 gpurun_out/r2_smoke_b.txt
if [ $rc -ne 0 ]; then
  echo "new kernel failed its smoke test: falling back to library A for the rest of this call"
  cp ab/libcsi_a.so scalable_stellar_inference_b200/libcsi.so
else
  STEPS=20 bash tools/gpu_ab.sh > gpurun_out/r2_ab_8warp.txt 2>&1; cat gpurun_out/r2_ab_8warp.txt
  cp ab/libcsi_b.so scalable_stellar_inference_b200/libcsi.so
  for c in c3 c5; do
    for v in a b; do
      cp ab/libcsi_$v.so scalable_stellar_inference_b200/libcsi.so
      timeout 200 python bench.py --config $c --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | \
        python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('$c $v kernel_ms', round(l['roofline']['kernel_ms'],4), 'value', round(l['value'],1))" >> gpurun_out/r2_ab_8warp.txt 2>&1
    done
  done
  tail -4 gpurun_out/r2_ab_8warp.txt
  cp ab/libcsi_b.so scalable_stellar_inference_b200/libcsi.so
fi
timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider -rfs 2>&1 | tail -40 > gpurun_out/r2_pytest_gpu.txt
tail -12 gpurun_out/r2_pytest_gpu.txt
timeout 1500 python tools/nuts_bench.py --config r1k --chains 4 --warmup ${WARM:-4000} --samples ${SAMP:-2000} \
  --out gpurun_out/r2_nuts_r1k.json > gpurun_out/r2_nuts_r1k.log 2>&1
tail -c 1800 gpurun_out/r2_nuts_r1k.log
