#!/usr/bin/env python
"""Writes tests/golden/posterior_hier64_seis.json: posterior summary of a hierarchical model under the fp64 ORACLE
potential (oracle.model.potential_and_grad), sampled on the CPU with the host-recursive NUTS (nuts.NUTS, numpyro's
settings: target 0.8, diagonal mass matrix, max depth 10).  tests/test_gpu_nuts.py compares the device-side sampler
over the CUDA potential against it (north_star's posterior criterion).  Chains run as separate processes:

    python tools/gen_posterior_fixture.py --stars 128 --warmup 500 --samples 400 --chains 4
"""
import argparse
import json
import math
import os
import sys
import time
from concurrent.futures import ProcessPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("OMP_NUM_THREADS", "1")

import numpy as np  # noqa: E402


def chain(args):
    name, n, nw, ns, seed = args
    import torch
    import parity_util as PU
    from oracle import model as M
    from scalable_stellar_inference_b200.nuts import NUTS, VectorSpace
    torch.set_num_threads(1)
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    hs, ss = spec.hyper_sites(), spec.star_sites()
    nh = len(hs)

    def flat(d):
        return np.concatenate([np.atleast_1d(np.asarray(d[k], float)) for k in hs] + [np.asarray(d[k], float) for k in ss])

    def unflat(v):
        d = {k: v[i] for i, k in enumerate(hs)}
        for j, k in enumerate(ss):
            d[k] = v[nh + j * n: nh + (j + 1) * n]
        return d

    def pot(z):
        u, g = M.potential_and_grad(spec, unflat(z.numpy()), obs, diag, ml, ms)
        return float(u), torch.as_tensor(flat(g))

    t0 = time.time()
    out = NUTS(pot, VectorSpace(nh), seed=seed).run(torch.as_tensor(flat(z0)), nw, ns,
                                                     collect=lambda z: z.numpy().copy())
    return dict(samples=np.array(out["samples"]), depth=float(np.mean(out["depth"])), accept=float(np.mean(out["accept"])),
                div=int(sum(out["diverging"])), step=out["step_size"], seconds=time.time() - t0, n_evals=out["n_evals"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="hier_diag")
    ap.add_argument("--stars", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=500)
    ap.add_argument("--samples", type=int, default=400)
    ap.add_argument("--chains", type=int, default=4)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "posterior_hier64_seis.json"))
    a = ap.parse_args()
    from scalable_stellar_inference_b200.nuts import effective_sample_size, split_rhat
    with ProcessPoolExecutor(a.chains) as ex:
        res = list(ex.map(chain, [(a.case, a.stars, a.warmup, a.samples, 500 + c) for c in range(a.chains)]))
    x = np.stack([r["samples"] for r in res])            # [chain, draw, coordinate]
    import parity_util as PU
    spec = PU.oracle_spec(a.case, a.stars)
    nh, ns = len(spec.hyper_sites()), len(spec.star_sites())
    # coordinates kept: every hyper site + the sites of 6 stars spread over the population
    stars = np.linspace(0, a.stars - 1, 6).astype(int)
    index = list(range(nh)) + [nh + s * a.stars + int(i) for s in range(ns) for i in stars]
    xs = x[:, :, index]
    ess = [effective_sample_size(xs[:, :, k]) for k in range(len(index))]
    fx = dict(case=a.case, stars=a.stars, num_warmup=a.warmup, num_samples=a.samples, chains=a.chains,
              sites=spec.hyper_sites() + spec.star_sites(), index=index,
              mean=xs.mean(axis=(0, 1)).tolist(), sd=xs.std(axis=(0, 1)).tolist(), ess=ess,
              rhat=[split_rhat(xs[:, :, k]) for k in range(len(index))],
              mean_depth=[r["depth"] for r in res], mean_accept=[r["accept"] for r in res],
              divergences=[r["div"] for r in res], step_size=[r["step"] for r in res],
              cpu_seconds=[r["seconds"] for r in res], n_evals=[r["n_evals"] for r in res],
              generator="tools/gen_posterior_fixture.py: nuts.NUTS over oracle.model.potential_and_grad (fp64)")
    with open(a.out, "w") as fh:
        json.dump(fx, fh, indent=1)
    print(json.dumps({k: fx[k] for k in ("ess", "rhat", "mean_depth", "cpu_seconds")}))


if __name__ == "__main__":
    main()
