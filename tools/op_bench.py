#!/usr/bin/env python
"""Roofline numbers of the standalone operator kernels (plug point 1): BC table interpolation and its
VJP (HBM-bound), emulator / BC-net forward and VJP (tensor pipe above 2048 rows).  CUDA events on the launching
stream, inputs larger than L2, 3 warm-ups, mean of 10."""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from scalable_stellar_inference_b200 import synthetic
from scalable_stellar_inference_b200.emulator import DeviceOps, load_emulator_weights, load_bc_weights

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
dev = torch.device("cuda", 0)
grid = synthetic.synthetic_bc_grid(bands=3)
ops = DeviceOps(load_emulator_weights(), load_bc_weights(), grid=grid, bands=3)
lib, h = ops.lib, ops.h
rng = np.random.default_rng(0)
pts = grid[0]
xg = np.stack([rng.uniform(p[1], p[-2], n) for p in pts], 1).astype(np.float32)
x4 = torch.from_numpy(xg).to(dev)
x4n = torch.from_numpy(np.stack([rng.uniform(3.45, 4.3, n), rng.uniform(2.2, 5.8, n), rng.uniform(-0.95, 0.45, n),
                                 rng.uniform(0, 3.5, n)], 1).astype(np.float32)).to(dev)   # BC emulator inputs: [log_Teff, log_g, M_H, Av]
gy3 = torch.randn(n, 3, device=dev)
y3 = torch.empty(n, 3, device=dev)
gx4 = torch.empty(n, 4, device=dev)
x5 = torch.from_numpy(np.stack([rng.uniform(0.05, 0.9, n), rng.uniform(0.8, 2.0, n), rng.uniform(-0.8, 0.4, n),
                                rng.uniform(0.24, 0.32, n), rng.uniform(1.5, 2.5, n)], 1).astype(np.float32)).to(dev)
y4 = torch.empty(n, 4, device=dev)
gy4 = torch.randn(n, 4, device=dev)
gx5 = torch.empty(n, 5, device=dev)
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0}
st = torch.cuda.current_stream(dev).cuda_stream

def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

p = lambda t: C.c_void_p(t.data_ptr())
rows = []
def add(name, ms, bytes_per, flop_per):
    rows.append({"op": name, "n": n, "ms": ms, "GB/s": bytes_per * n / ms / 1e6, "frac_hbm": bytes_per * n / ms / 1e6 / peaks["hbm_gbs"],
                 "TFLOP/s": flop_per * n / ms / 1e9, "algorithmic_bytes_per_point": bytes_per})
add("csi_grid_forward (4-D table, 3 bands)", timeit(lambda: lib.csi_grid_forward(h, p(x4), p(y3), n, C.c_void_p(st))), 16 + 12, 200)
add("csi_grid_vjp", timeit(lambda: lib.csi_grid_vjp(h, p(x4), p(gy3), p(gx4), n, C.c_void_p(st))), 16 + 12 + 16, 300)
add("csi_net_forward (emulator 5-128x6-4)", timeit(lambda: lib.csi_net_forward(h, 0, p(x5), p(y4), n, C.c_void_p(st))), 20 + 16, 166144)
add("csi_net_vjp (emulator)", timeit(lambda: lib.csi_net_vjp(h, 0, p(x5), p(gy4), p(gx5), n, C.c_void_p(st))), 20 + 16 + 20, 332288)
add("csi_net_forward (BC net 4-64x6-3)", timeit(lambda: lib.csi_net_forward(h, 1, p(x4n), p(y3), n, C.c_void_p(st))), 16 + 12, 41856)
add("csi_net_vjp (BC net)", timeit(lambda: lib.csi_net_vjp(h, 1, p(x4n), p(gy3), p(gx4), n, C.c_void_p(st))), 16 + 12 + 16, 83712)
for r in rows:
    print(json.dumps(r))
