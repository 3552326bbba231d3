import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, parity_util as PU
from oracle import model as M
np.set_printoptions(precision=3, linewidth=220)
for name in ["hier_diag_seis", "hier_phot_grid"]:
    n = 20000
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    u_ref, g_ref, terms = M.potential_and_grad(spec, z0, obs, diag, ml, ms, return_terms=True)
    for eng in ["ffma", "tcgen05"]:
        model = PU.product_model(name, n, eng).condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
        d = model.predict(z0)
        mean = np.stack([d[nm] for nm in model.output_names], -1).astype(np.float64)
        err = mean - terms["mean"]
        print(name, eng, "mean err: bias", err.mean(0), "\n      rms", err.std(0), "\n      max", np.abs(err).max(0))
        u, g = model.potential_and_grad(z0)
        print("   U err", u - u_ref, "rel", (u - u_ref) / abs(u_ref))
