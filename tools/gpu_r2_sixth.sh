#!/bin/bash
# round 2, final single-GPU call: the whole GPU suite on the final library, smoke, then the like-for-like replica of the
# reference's recorded hierarchical run (1000 stars, 4 chains, 16000 warm-up + 2000 samples)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider -rfs --timeout 500 --durations=5 > gpurun_out/r2_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_gpu.txt; tail -14 gpurun_out/r2_pytest_gpu.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1; tail -1 gpurun_out/r2_smoke.txt
timeout 1500 python tools/nuts_bench.py --config r1k --chains 4 --warmup ${WARM:-16000} --samples ${SAMP:-2000} \
  --out gpurun_out/r2_nuts_r1k_16000.json > gpurun_out/r2_nuts_r1k_16000.log 2>&1
tail -c 1500 gpurun_out/r2_nuts_r1k_16000.log
