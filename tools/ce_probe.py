"""Copy-engine rates for the gated host-buffer path: 1-D and 2-D (site x star-block) host -> device copies from
page-locked memory, at 16-byte-aligned and at 20-byte offsets, with and without the 8-byte progress copies."""
import ctypes
import numpy as np
import torch

rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemcpy2DAsync.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t,
                                 ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
n, sites = 1_000_000, 7
dev = torch.device("cuda", 0)
h = torch.zeros(n * sites + 16, dtype=torch.float32).pin_memory()
d = torch.zeros(n * sites + 16, dtype=torch.float32, device=dev)
hf = torch.zeros(64, dtype=torch.int64).pin_memory()
df = torch.zeros(1, dtype=torch.int64, device=dev)
st = torch.cuda.Stream(dev)
round_ = 128 * 148
bounds = [0]
for r in (2, 4, 8, 12, 12, 12, 12):
    bounds.append(min(n, bounds[-1] + r * round_))
    if bounds[-1] == n:
        break


def run(label, fn, reps=10):
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record(st)
            fn()
            e1.record(st)
        torch.cuda.synchronize(dev)
        ts.append(e0.elapsed_time(e1))
    ts = np.array(ts)
    print(f"{label:60s} median {np.median(ts):.4f} ms  min {ts.min():.4f}  -> {n * sites * 4 / np.median(ts) / 1e6:.1f} GB/s")


def one_d(off):
    rt.cudaMemcpyAsync(d.data_ptr() + off, h.data_ptr() + off, n * sites * 4, 1, st.cuda_stream)


def two_d(off, flags, per_site=False):
    for b in range(len(bounds) - 1):
        lo, hi = bounds[b], bounds[b + 1]
        if per_site:
            for s in range(sites):
                rt.cudaMemcpyAsync(d.data_ptr() + off + (s * n + lo) * 4, h.data_ptr() + off + (s * n + lo) * 4,
                                   (hi - lo) * 4, 1, st.cuda_stream)
        else:
            rt.cudaMemcpy2DAsync(d.data_ptr() + off + lo * 4, n * 4, h.data_ptr() + off + lo * 4, n * 4, (hi - lo) * 4,
                                 sites, 1, st.cuda_stream)
        if flags:
            rt.cudaMemcpyAsync(df.data_ptr(), hf.data_ptr() + 8 * b, 8, 1, st.cuda_stream)


for off in (0, 20):
    run(f"1-D 28 MB, offset {off}", lambda: one_d(off))
    run(f"2-D blocks {len(bounds) - 1}, offset {off}, no progress copies", lambda: two_d(off, False))
    run(f"2-D blocks {len(bounds) - 1}, offset {off}, progress copies", lambda: two_d(off, True))
    run(f"1-D per site x block, offset {off}, progress copies", lambda: two_d(off, True, True))
