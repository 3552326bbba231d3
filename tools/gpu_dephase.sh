#!/bin/bash
# does de-phasing the CTAs (CSI_TC_DEBUG=32: four start phases 8 us apart) help the kernel when it reads / writes host memory?
mkdir -p gpurun_out; rm -f gpurun_out/dephase.txt
for rep in 1 2; do for d in 0 32; do
  echo "== CSI_TC_DEBUG=$d rep $rep" >> gpurun_out/dephase.txt
  CSI_TC_DEBUG=$d timeout 200 python tools/zc_probe.py --steps 20 2>&1 | grep "flush 0\|wall" >> gpurun_out/dephase.txt
done; done
cat gpurun_out/dephase.txt
