#!/usr/bin/env python
"""NUTS ESS/s on the B200 path (the second half of BASELINE.json's metric): the device-side sampler (csi_nuts_*)
over one of the BASELINE configurations, numpyro's settings as the reference uses them (scripts/model_star.py:18-24:
target_accept 0.8, heuristic initial step, regularised diagonal mass matrix, max tree depth 10, init at the median
-- here the jittered truth -- and 2000 + 2000 draws unless told otherwise).

    python tools/nuts_bench.py --config c3 --chains 4 --warmup 2000 --samples 2000          # 1 GPU, chains side by side
    torchrun --nproc-per-node 8 ... tools/nuts_bench.py --config c4 --chains 4 ...           # star-sharded, chains in turn

ESS = bulk effective sample size over all chains (split chains, Geyer's initial monotone sequence), min over the 5
hyper-parameters; ESS/s = that divided by the SAMPLING wall time (warm-up reported apart).  On one GPU the chains run
concurrently, one host thread + stream + handle each, the potential limited to SMs / chains CTAs (csi_set_grid_limit);
sharded over N GPUs every chain uses all GPUs and the chains run one after the other.
"""
import argparse
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench as B  # noqa: E402  (configuration table only)
import scalable_stellar_inference_b200 as ssi  # noqa: E402
from scalable_stellar_inference_b200 import _lib, synthetic  # noqa: E402
from scalable_stellar_inference_b200.nuts import DeviceNUTS, effective_sample_size, split_rhat  # noqa: E402
from scalable_stellar_inference_b200.parallel import setup_peer_exchange, shard_bounds, shard_problem  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3", choices=sorted(B.CONFIGS))
    ap.add_argument("--stars", type=int, default=0)
    ap.add_argument("--chains", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=2000)
    ap.add_argument("--samples", type=int, default=2000)
    ap.add_argument("--max-tree-depth", type=int, default=10)
    ap.add_argument("--serial", action="store_true", help="one chain after the other on the whole GPU")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    cfg = B.CONFIGS[a.config]
    n = a.stars or cfg["stars"]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    grid = synthetic.synthetic_bc_grid(bands=3) if cfg["phot"] else None

    def make_model():
        kw = dict(device=local)
        if cfg["phot"]:
            kw.update(bands=B.BANDS, bc="grid", bc_grid=grid)
        cls = ssi.HierarchicalStarModel if cfg["level"] == "hierarchical" else ssi.MultiStarModel
        return cls(n, list(cfg["obs"]), **kw)

    gen = make_model()
    z0, obs, diag, ml, ms, _ = synthetic.make_problem(gen, n)
    del gen
    lo, hi = shard_bounds(n, rank, world)
    zl, ol, dl, mll, msl = shard_problem(z0, obs, diag, ml, ms, rank, world)
    concurrent = world == 1 and a.chains > 1 and not a.serial
    sms = torch.cuda.get_device_properties(local).multi_processor_count
    models, keep = [], []
    for c in range(a.chains if concurrent else 1):
        m = make_model().set_shard(rank, world)
        m.condition(ol, dl, M_H_loc=mll, M_H_scale=msl)
        if concurrent:
            _lib.check(m._lib.csi_set_grid_limit(m._h, max(1, sms // a.chains)))
        if world > 1:
            ex = setup_peer_exchange(m)
            if ex is None:
                raise SystemExit("the sharded sampler needs the in-kernel peer exchange")
            keep.append(ex)
        models.append(m)
    nh = len(models[0].hyper_sites)
    zf = models[0].flatten(zl)
    results = [None] * a.chains

    def run_chain(c, model):
        torch.cuda.set_device(local)
        marks = {}

        def progress(it, warm, step, depth, acc):
            if it + 1 == a.warmup:
                marks["t_warm"] = time.time()

        t0 = time.time()
        nuts = DeviceNUTS(model, zf, seed=1000 + c, max_tree_depth=a.max_tree_depth, star_offset=lo, n_total=n)
        out = nuts.run(a.warmup, a.samples, progress=progress)
        t1 = time.time()
        out.update(t_warm=marks.get("t_warm", t0) - t0, t_samp=t1 - marks.get("t_warm", t0))
        results[c] = out

    t_all = time.time()
    if concurrent:
        th = [threading.Thread(target=run_chain, args=(c, models[c])) for c in range(a.chains)]
        [t.start() for t in th]
        [t.join() for t in th]
    else:
        for c in range(a.chains):
            run_chain(c, models[0])
    torch.cuda.synchronize()
    wall = time.time() - t_all
    if rank == 0:
        x = np.stack([np.array(r["samples"]) for r in results])          # [chain, draw, hyper]
        ess = [effective_sample_size(x[:, :, k]) for k in range(x.shape[2])]
        rhat = [split_rhat(x[:, :, k]) for k in range(x.shape[2])]
        # chains side by side: sampling time = the slowest chain's; in turn: the sum
        t_samp = max(r["t_samp"] for r in results) if concurrent else sum(r["t_samp"] for r in results)
        t_warm = max(r["t_warm"] for r in results) if concurrent else sum(r["t_warm"] for r in results)
        ev_s = sum(r["n_evals"] - r["n_evals_warmup"] for r in results)
        line = {"metric": "NUTS ESS/s (min bulk ESS over the hyper-parameters, all chains / sampling wall time)",
                "value": min(ess) / t_samp, "unit": "ESS/s", "config": a.config, "workload": cfg["text"], "stars": n,
                "n_gpus": world, "chains": a.chains, "chains_concurrent": concurrent,
                "num_warmup": a.warmup, "num_samples": a.samples, "max_tree_depth": a.max_tree_depth,
                "sites": models[0].hyper_sites if nh else models[0].star_sites[:1],
                "ess": [round(e, 1) for e in ess], "rhat": [round(r, 4) for r in rhat],
                "posterior_mean": x.mean(axis=(0, 1)).tolist(), "posterior_sd": x.std(axis=(0, 1)).tolist(),
                "sampling_s": t_samp, "warmup_s": t_warm, "wall_s": wall,
                "evals_sampling": int(ev_s), "evals_per_s_in_sampler": ev_s / t_samp,
                "mean_tree_depth": float(np.mean([np.mean(r["depth"]) for r in results])),
                "mean_accept": float(np.mean([np.mean(r["accept"]) for r in results])),
                "divergences": int(sum(sum(r["diverging"]) for r in results)),
                "step_size": [r["step_size"] for r in results],
                "host_syncs_per_leapfrog": sum(r["host_syncs"] for r in results) / max(1, sum(r["n_evals"] for r in results)),
                "launches_per_leapfrog": sum(r["launches"] for r in results) / max(1, sum(r["n_evals"] for r in results))}
        s = json.dumps(line)
        print(s)
        if a.out:
            with open(a.out, "w") as fh:
                fh.write(s + "\n")
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
