#!/bin/bash
# round 2, closing single-GPU call: the allocation-free entry point's test alone in a fresh process, the whole GPU
# suite (which ends with it), smoke, and bench.py with the interleaved (default) and the single-chain end-to-end mode
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider -rfs --timeout 300 > gpurun_out/r2_pytest_no_alloc_alone.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_no_alloc_alone.txt; tail -5 gpurun_out/r2_pytest_no_alloc_alone.txt
timeout 1300 python -m pytest tests -m gpu -q -p no:cacheprovider -rfs --timeout 500 --durations=5 > gpurun_out/r2_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_gpu.txt; tail -14 gpurun_out/r2_pytest_gpu.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1; tail -1 gpurun_out/r2_smoke.txt
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err
echo "bench exit $?"; tail -c 3000 gpurun_out/r2_bench_c4_n1.json; tail -c 600 gpurun_out/r2_bench_c4_n1.err
timeout 400 python bench.py --steps 20 --warmup 3 --e2e-chains 1 > gpurun_out/r2_bench_c4_n1_chain1.json 2> gpurun_out/r2_bench_c4_n1_chain1.err
echo "bench exit $?"; tail -c 1500 gpurun_out/r2_bench_c4_n1_chain1.json
