"""Prints the clock64 trace of CTA 0 of the tcgen05 kernel (CSI_TC_DEBUG=8 [+ other experiment bits])."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import parity_util as PU
n = int(os.environ.get("TRACE_STARS", 148 * 256 * 3)) // 2048 * 2048
name = "hier_phot_grid"
spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, 2048)
rep = n // 2048
tile = lambda a: np.tile(np.asarray(a), (rep,) + (1,) * (np.ndim(a) - 1))
zb = {k: (v if np.ndim(v) == 0 else tile(v)) for k, v in z0.items()}
model = PU.product_model(name, n, "tcgen05").condition(tile(obs), tile(diag), M_H_loc=tile(ml), M_H_scale=tile(ms))
zf = model.flatten(zb)
for _ in range(3):
    model.potential_and_grad_flat(zf)
buf = np.zeros(8 * 1024, np.int64)
lib = model._lib
lib.csi_tc_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
assert lib.csi_tc_trace(model._h, buf.ctypes.data, buf.size) == 0
names = ["w0(tileA)", "w3(tileA)", "w4(tileB)", "w7(tileB)", "mma", "cta"]
ev = []
for who in range(6):
    reg = buf[who * 1024:(who + 1) * 1024].reshape(-1, 2)
    for tag, clk in reg:
        if tag:
            ev.append((clk, who, tag))
ev.sort()
t0 = ev[0][0]
# one pair only: print events of the 2nd pair region
for clk, who, tag in ev:
    print(f"{clk - t0:8d} {names[who]:10s} {tag}")
