#!/usr/bin/env python
"""NUTS ESS/s on the B200 path (north_star metric): HierarchicalStarModel, observables
[log_Teff, log_g] (BASELINE config 3 at --stars 10000), numpyro's settings (scripts/model_star.py:18-22).
ESS = the smallest bulk ESS over the 5 hyper-parameters; time = sampling wall time (warm-up reported apart)."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import scalable_stellar_inference_b200 as ssi
from scalable_stellar_inference_b200 import synthetic
from scalable_stellar_inference_b200.nuts import sample_model, effective_sample_size

ap = argparse.ArgumentParser()
ap.add_argument("--stars", type=int, default=10000)
ap.add_argument("--warmup", type=int, default=150)
ap.add_argument("--samples", type=int, default=150)
ap.add_argument("--max-tree-depth", type=int, default=10)
a = ap.parse_args()
obs_names = ["log_Teff", "log_g"]
model = ssi.HierarchicalStarModel(a.stars, obs_names)
z0, obs, diag, ml, ms, _ = synthetic.make_problem(model, a.stars)
model.condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
zf = model.flatten(z0)
marks = {}
def progress(it, warm, step, depth, acc):
    if it + 1 == a.warmup:
        torch.cuda.synchronize(); marks["t_warm"] = time.time()
torch.cuda.synchronize(); t0 = time.time()
from scalable_stellar_inference_b200.nuts import NUTS, VectorSpace, DevicePotential
pot = DevicePotential(model)
nuts = NUTS(pot, VectorSpace(pot.nh), seed=0, max_tree_depth=a.max_tree_depth)
out = nuts.run(torch.as_tensor(zf, device=torch.device("cuda", model.device)), a.warmup, a.samples, progress=progress)
torch.cuda.synchronize(); t1 = time.time()
s = np.array(out["samples"])
ess = [effective_sample_size(s[:, k]) for k in range(s.shape[1])]
t_samp = t1 - marks.get("t_warm", t0)
ev_s = out["n_evals"] - getattr(nuts, "n_evals_warmup", 0)
print(json.dumps({"metric": "NUTS ESS/s (min bulk ESS over the 5 hyper-parameters / sampling wall time)",
                  "value": min(ess) / t_samp, "unit": "ESS/s", "stars": a.stars, "observables": obs_names,
                  "num_warmup": a.warmup, "num_samples": a.samples, "max_tree_depth": a.max_tree_depth,
                  "ess_hyper": [round(e, 1) for e in ess], "sampling_s": t_samp, "warmup_s": marks.get("t_warm", t0) - t0,
                  "evals_sampling": ev_s, "evals_per_s_in_sampler": ev_s / t_samp,
                  "mean_tree_depth": float(np.mean(out["depth"])), "mean_accept": float(np.mean(out["accept"])),
                  "divergences": int(sum(out["diverging"])), "step_size": out["step_size"], "n_gpus": 1}))
