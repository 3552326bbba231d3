"""Pinned-host <-> device copy rates of this box (what bounds the host-buffer path): H2D, D2H, both at once."""
import json, time, torch
dev = torch.device("cuda", 0)
n = 28_000_000 // 4
hp, hp2 = torch.zeros(n).pin_memory(), torch.zeros(n).pin_memory()
d, d2 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
def run(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t) / reps
def h2d():
    with torch.cuda.stream(s1): d.copy_(hp, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): hp2.copy_(d2, non_blocking=True)
def both():
    h2d(); d2h()
gb = n * 4 / 1e9
out = {"bytes": n * 4, "h2d_GBps": gb / run(h2d), "d2h_GBps": gb / run(d2h), "both_ms": run(both) * 1e3}
out["both_GBps_each"] = gb / (out["both_ms"] * 1e-3)
print(json.dumps(out))
