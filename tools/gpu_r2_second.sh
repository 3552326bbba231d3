#!/bin/bash
# round 2, second GPU call: the whole GPU suite (no -x), then the real ESS/s run of config 3 (4 chains side by side)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider -rs 2>&1 | tail -60 > gpurun_out/r2_pytest_gpu.txt
tail -8 gpurun_out/r2_pytest_gpu.txt
timeout 1500 python tools/nuts_bench.py --config c3 --chains 4 --warmup ${WARM:-2000} --samples ${SAMP:-2000} \
  --out gpurun_out/r2_nuts_c3.json > gpurun_out/r2_nuts_c3.log 2>&1
tail -c 1500 gpurun_out/r2_nuts_c3.log
