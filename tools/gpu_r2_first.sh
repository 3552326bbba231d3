#!/bin/bash
# round 2, first GPU call: the whole GPU suite on the refactored library, bench lines of the four configurations,
# a first device-sampler timing, topology probe for the NUMA binding
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2_gpu.txt 2>&1
python - > gpurun_out/r2_numa_probe.txt 2>&1 <<'PY'
import os, glob, torch
p = torch.cuda.get_device_properties(0)
bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
print("bus", bus, "exists", os.path.exists(f"/sys/bus/pci/devices/{bus}"))
for f in ("local_cpulist", "numa_node"):
    try: print(f, open(f"/sys/bus/pci/devices/{bus}/{f}").read().strip())
    except Exception as e: print(f, "ERR", e)
print("nodes", glob.glob("/sys/devices/system/node/node*")[:8])
print("affinity", len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:4], "...")
import sys; sys.path.insert(0, ".")
from scalable_stellar_inference_b200.parallel import gpu_local_cpus, bind_host_to_device
c, how = gpu_local_cpus(0); print("gpu_local_cpus", None if c is None else len(c), how)
print({k: (v if k != "previous" else (None if v is None else len(v))) for k, v in bind_host_to_device(0).items()})
PY
timeout 2400 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -40 > gpurun_out/r2_pytest_gpu.txt
tail -5 gpurun_out/r2_pytest_gpu.txt
for c in c4 c3 c5 c2; do
  extra="--no-cpu-baseline"; [ $c = c4 ] && extra=""
  timeout 600 python bench.py --config $c --steps 20 --warmup 3 $extra > gpurun_out/r2_bench_$c.json 2> gpurun_out/r2_bench_$c.err
  tail -c 600 gpurun_out/r2_bench_$c.json
done
timeout 900 python tools/nuts_bench.py --config c3 --chains 1 --warmup 150 --samples 50 --out gpurun_out/r2_nuts_c3_quick.json > gpurun_out/r2_nuts_c3_quick.log 2>&1
tail -3 gpurun_out/r2_nuts_c3_quick.log
timeout 900 python tools/nuts_bench.py --config c3 --chains 4 --warmup 150 --samples 50 --out gpurun_out/r2_nuts_c3_quick4.json > gpurun_out/r2_nuts_c3_quick4.log 2>&1
tail -3 gpurun_out/r2_nuts_c3_quick4.log
