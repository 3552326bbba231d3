"""Where does the host-buffer step lose time against the device-resident one?  Kernel time (CUDA events on the
launching stream, L2 flushed between launches) of csi_potential for the four placements of (z, gradient):
device / device, host / device, device / host, host / host (page-locked host buffers, read / written by the
kernel itself over PCIe), then the wall clock of the synchronous csi_potential_host call.
Run on the GPU box: python tools/zc_probe.py [--stars N] [--steps K]"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stars", type=int, default=1_000_000)
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    import torch
    import scalable_stellar_inference_b200 as ssi
    from scalable_stellar_inference_b200 import synthetic
    from bench import BANDS, OBSERVABLES

    dev = torch.device("cuda", 0)
    grid = synthetic.synthetic_bc_grid(bands=3)
    model = ssi.HierarchicalStarModel(args.stars, list(OBSERVABLES), bands=BANDS, bc="grid", bc_grid=grid, device=0)
    z0, obs, diag, ml, ms, _ = synthetic.make_problem(model, args.stars)
    model.condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    n = model.flat_size
    z_host = torch.from_numpy(model.flatten(z0)).pin_memory()
    g_host = torch.zeros(n, dtype=torch.float32).pin_memory()
    z_dev = z_host.to(dev)
    g_dev = torch.zeros(n, dtype=torch.float32, device=dev)
    out_dev = torch.zeros(8, dtype=torch.float64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    K = args.steps
    for zname, z in (("dev", z_dev), ("host", z_host)):
        for gname, g in (("dev", g_dev), ("host", g_host)):
            for fl in (True, False):
                ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
                for k in range(-3, K):
                    if fl:
                        flush.fill_(k & 0xFF)
                    if k >= 0:
                        ev[k][0].record()
                    model.potential_device(z.data_ptr(), out_dev.data_ptr(), g.data_ptr(), stream)
                    if k >= 0:
                        ev[k][1].record()
                torch.cuda.synchronize(dev)
                ms_ = np.array([a.elapsed_time(b) for a, b in ev])
                print(f"z {zname:4s} grad {gname:4s} flush {int(fl)} kernel_ms median {np.median(ms_):.4f} "
                      f"min {ms_.min():.4f} max {ms_.max():.4f} U {float(out_dev[0].item()):.6e}", flush=True)
    out = np.zeros(8, np.float64)
    for _ in range(3):
        model.potential_host(z_host.data_ptr(), out.ctypes.data, g_host.data_ptr())
    t = []
    for k in range(K):
        t0 = time.perf_counter()
        model.potential_host(z_host.data_ptr(), out.ctypes.data, g_host.data_ptr())
        t.append((time.perf_counter() - t0) * 1e3)
    t = np.array(t)
    print(f"csi_potential_host wall ms median {np.median(t):.4f} min {t.min():.4f} max {t.max():.4f} U {out[0]:.6e}")


if __name__ == "__main__":
    main()
