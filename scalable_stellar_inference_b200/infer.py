"""The driver objects of the reference's sampling script with numpyro's names and call shapes, over the CUDA path.

The reference runs (scripts/model_star.py:132-196)

    sampler = NUTS(model, init_strategy=init_to_median, target_accept_prob=0.8, find_heuristic_step_size=True,
                   dense_mass=False, regularize_mass_matrix=True)
    mcmc = MCMC(sampler, num_warmup=2000, num_samples=2000, num_chains=10)
    mcmc.run(key, obs=y, diag=diag);  mcmc.print_summary()
    posterior = mcmc.get_samples(group_by_chain=True);  sample_stats = mcmc.get_extra_fields(group_by_chain=True)
    y = Predictive(model, posterior_samples=posterior, return_sites=["y"], batch_ndims=2)(key)["y"]
    az.from_dict(...).to_netcdf(...)

numpyro / jax / arviz are third-party and absent here; these classes keep that script's shape with the potential, the
leapfrogs and the tree on the B200 (``nuts.DeviceNUTS``), the forward pass of ``Predictive`` on the tensor-core
predict kernel (``csi_predict``) and the results written as NetCDF-3 files (``scipy.io.netcdf_file``; one file per
arviz group, since NetCDF-3 has no groups).  ``rng_key`` is an integer seed.
"""
from __future__ import annotations

import json
import math
import os

import numpy as np

from .nuts import DeviceNUTS, effective_sample_size, split_rhat


def init_to_median(model, num_samples=15, seed=0):
    """numpyro.infer.init_to_median: the median of ``num_samples`` prior draws per site, in unconstrained space."""
    from . import synthetic
    rng = np.random.default_rng(seed)
    n = max(model.n_local, 1)
    draws = [synthetic.truths_to_z(model, synthetic.draw_star_truths(model, n, rng)) for _ in range(num_samples)]
    return {k: np.median(np.stack([np.asarray(d[k], np.float64) for d in draws]), axis=0) for k in draws[0]}


class NUTS:
    """numpyro.infer.NUTS(model, ...) with the arguments the reference sets (scripts/model_star.py:18-24,134-136)."""

    def __init__(self, model, init_strategy=init_to_median, target_accept_prob=0.8, find_heuristic_step_size=True,
                 dense_mass=False, regularize_mass_matrix=True, max_tree_depth=10, step_size=1.0, adapt_mass_matrix=True):
        if dense_mass:
            raise NotImplementedError("dense_mass=True: the reference samples with a diagonal mass matrix "
                                      "(scripts/model_star.py:22); only that is implemented")
        if not regularize_mass_matrix:
            raise NotImplementedError("regularize_mass_matrix=False is not implemented (the reference sets True)")
        self.model, self.init_strategy = model, init_strategy
        self.target_accept_prob, self.find_heuristic_step_size = target_accept_prob, find_heuristic_step_size
        self.max_tree_depth, self.step_size, self.adapt_mass_matrix = max_tree_depth, step_size, adapt_mass_matrix


class MCMC:
    """numpyro.infer.MCMC(sampler, num_warmup=, num_samples=, num_chains=): chains run one after the other on the
    model's GPU (``tools/nuts_bench.py`` runs them side by side with one handle per chain)."""

    def __init__(self, sampler: NUTS, num_warmup: int, num_samples: int, num_chains: int = 1, progress_bar: bool = False):
        self.sampler, self.num_warmup, self.num_samples, self.num_chains = sampler, num_warmup, num_samples, num_chains
        self._z = self._stats = None
        self._kwargs = {}

    def run(self, rng_key=0, *args, init_params=None, **kwargs):
        """``kwargs`` are the model's call arguments (``obs=``, ``diag=``, and here ``M_H_loc=`` / ``M_H_scale=``)."""
        s, model = self.sampler, self.sampler.model
        if args or kwargs:
            model(*args, **kwargs)               # conditions the model: the obs= / diag= of model.__call__
        if model.n_local < 1:
            raise ValueError("MCMC.run needs observations: pass obs= (and diag=)")
        self._kwargs = {k: np.asarray(v) for k, v in kwargs.items() if v is not None}
        seed = int(rng_key)
        zs, stats = [], []
        for c in range(self.num_chains):
            z0 = init_params if init_params is not None else s.init_strategy(model, seed=seed + c)
            nuts = DeviceNUTS(model, model.flatten(z0), seed=seed * 1000 + c, target_accept=s.target_accept_prob,
                              max_tree_depth=s.max_tree_depth, adapt_mass=s.adapt_mass_matrix)
            out = nuts.run(self.num_warmup, self.num_samples, collect=lambda z: z.detach().cpu().numpy().copy(),
                           step_size=None if s.find_heuristic_step_size else s.step_size)
            zs.append(np.array(out["samples"]))
            stats.append(dict(diverging=np.array(out["diverging"]), accept_prob=np.array(out["accept"]),
                              tree_depth=np.array(out["depth"]), num_steps=np.array(out["n_leapfrog"]),
                              potential_energy=np.array(out["potential"]),
                              step_size=np.full(self.num_samples, out["step_size"])))
            del nuts
        self._z = np.stack(zs)                    # [chain, draw, flat unconstrained vector]
        self._stats = {k: np.stack([st[k] for st in stats]) for k in stats[0]}
        return self

    # ---- results ------------------------------------------------------------------------------------------
    def get_samples(self, group_by_chain=False):
        """Constrained site values (+ the ``numpyro.deterministic`` sites ``evol``, ``mass``) per draw."""
        model = self.sampler.model
        c, d, L = self._z.shape
        n = model.n_local
        out = {}
        hyper = {}
        flat = self._z.reshape(c * d, L)
        for i in range(c * d):
            p = model.predict(model.unflatten(flat[i]), constrained=True)
            for k, v in p.items():
                out.setdefault(k, []).append(v)
        res = {k: np.stack(v) for k, v in out.items()}
        if model.num_stars == 1:
            res = {k: v.reshape(c * d, *v.shape[2:]) if v.ndim > 1 and v.shape[1] == 1 else v for k, v in res.items()}
        res.update(hyper)
        if group_by_chain:
            res = {k: v.reshape((c, d) + v.shape[1:]) for k, v in res.items()}
        return res

    def get_extra_fields(self, group_by_chain=False):
        return {k: (v if group_by_chain else v.reshape((-1,) + v.shape[2:])) for k, v in self._stats.items()}

    def summary(self, prob=0.9):
        """Per-site mean / std / median / 5 % / 95 % / n_eff / r_hat, as numpyro's print_summary computes them."""
        samples = self.get_samples(group_by_chain=True)
        rows = {}
        lo, hi = (1 - prob) / 2, 1 - (1 - prob) / 2
        for k, v in samples.items():
            v = v.reshape(v.shape[0], v.shape[1], -1)
            for j in range(v.shape[2]):
                x = v[:, :, j]
                name = k if v.shape[2] == 1 else f"{k}[{j}]"
                rows[name] = dict(mean=float(x.mean()), std=float(x.std(ddof=1)), median=float(np.median(x)),
                                  q_lo=float(np.quantile(x, lo)), q_hi=float(np.quantile(x, hi)),
                                  n_eff=effective_sample_size(x), r_hat=split_rhat(x) if x.shape[1] >= 4 else float("nan"))
        return rows

    def print_summary(self, prob=0.9, exclude_deterministic=True):
        rows = self.summary(prob)
        print(f"\n{'':>20s} {'mean':>9s} {'std':>9s} {'median':>9s} {100 * (1 - prob) / 2:>8.1f}% {100 * (1 + prob) / 2:>8.1f}% {'n_eff':>9s} {'r_hat':>9s}")
        for k, r in rows.items():
            print(f"{k:>20s} {r['mean']:9.2f} {r['std']:9.2f} {r['median']:9.2f} {r['q_lo']:9.2f} {r['q_hi']:9.2f} {r['n_eff']:9.2f} {r['r_hat']:9.2f}")
        print(f"\nNumber of divergences: {int(self._stats['diverging'].sum())}")


class Predictive:
    """numpyro.infer.Predictive(model, posterior_samples=..., num_samples=...): samples ``y`` -- all 7 outputs (+ the
    photometric ones), ``Normal(loc, sqrt(variance / scaled_precision))`` as ``likelihood`` does when ``obs is None``
    (numpyro_models.py:223-235).  The forward pass runs on the GPU (csi_predict), the Gaussian noise on the host."""

    def __init__(self, model, posterior_samples=None, num_samples=None, return_sites=None, parallel=False, batch_ndims=1):
        self.model, self.posterior, self.num_samples = model, posterior_samples, num_samples
        self.return_sites, self.batch_ndims = return_sites, batch_ndims

    def __call__(self, rng_key=0, *args, **kwargs):
        from . import synthetic
        model = self.model
        rng = np.random.default_rng(int(rng_key))
        nm = len(model.output_names)
        if model.n_local < 1:   # unconditioned model: predict needs the star count only
            model.condition(np.zeros((model.num_stars, len(model.observable_indices))), 1.0)
        n = model.n_local
        if self.posterior is None:
            zs = [synthetic.truths_to_z(model, synthetic.draw_star_truths(model, n, rng)) for _ in range(self.num_samples or 1)]
            shape = (len(zs),)
        else:
            uz = self.posterior["_unconstrained"] if "_unconstrained" in self.posterior else None
            if uz is None:
                raise ValueError("posterior_samples must come from MCMC.get_unconstrained() (key '_unconstrained')")
            shape = uz.shape[:-1]
            zs = [model.unflatten(v) for v in uz.reshape(-1, uz.shape[-1])]
        ys, extra = [], {}
        var = np.concatenate([model.variance, np.zeros(nm - 7)])
        for z in zs:
            p = model.predict(z, constrained=True)
            mean = np.stack([p[k] for k in model.output_names], axis=-1)
            sd = np.sqrt(var[None, :] / p["scaled_precision"][:, None])
            ys.append(mean + sd * rng.standard_normal(mean.shape))
            for k, v in p.items():
                extra.setdefault(k, []).append(v)
        y = np.stack(ys).reshape(shape + (n, nm))
        if model.num_stars == 1:
            y = y[..., 0, :]
        out = {"y": y}
        out.update({k: np.stack(v).reshape(shape + (n,)) for k, v in extra.items()})
        if self.return_sites:
            out = {k: out[k] for k in self.return_sites}
        return out


def get_unconstrained(mcmc: MCMC, group_by_chain=True):
    """The raw draws for ``Predictive(posterior_samples=...)``: {'_unconstrained': [chain, draw, L]}."""
    z = mcmc._z if group_by_chain else mcmc._z.reshape(-1, mcmc._z.shape[-1])
    return {"_unconstrained": z}


def save_inference_data(path, posterior=None, sample_stats=None, constant_data=None, attrs=None, **groups):
    """Results writer (scripts/model_star.py:157-196 builds an arviz InferenceData and calls ``to_netcdf``): one
    NetCDF-3 file per group under ``path`` -- ``posterior.nc`` with dims (chain, draw, ...), ``sample_stats.nc``,
    ``constant_data.nc`` -- readable by xarray / netCDF4 / scipy; ``attrs.json`` holds the attributes."""
    from scipy.io import netcdf_file
    os.makedirs(path, exist_ok=True)
    groups = dict(groups)
    if posterior is not None:
        groups["posterior"] = posterior
    if sample_stats is not None:
        groups["sample_stats"] = sample_stats
    if constant_data is not None:
        groups["constant_data"] = constant_data
    written = []
    for gname, data in groups.items():
        fn = os.path.join(path, f"{gname}.nc")
        with netcdf_file(fn, "w") as f:
            dims = {}

            def dim(name, size):
                if name not in dims:
                    f.createDimension(name, size)
                    dims[name] = size
                return name
            for k, v in data.items():
                if k.startswith("_"):
                    continue
                v = np.asarray(v)
                if v.dtype == bool:
                    v = v.astype(np.int8)
                if v.dtype == np.int64:
                    v = v.astype(np.int32)
                if v.dtype.kind not in "fi":
                    continue
                if v.ndim == 0:          # scalars as length-1 variables (NetCDF-3 scalars are awkward in scipy)
                    v = v.reshape(1)
                if gname == "constant_data":
                    names = [dim(f"{k}_dim_{i}", s) for i, s in enumerate(v.shape)]
                else:
                    names = [dim("chain", v.shape[0]), dim("draw", v.shape[1])] if v.ndim >= 2 else [dim(f"{k}_dim_0", v.shape[0])]
                    names += [dim(f"{k}_dim_{i}", s) for i, s in enumerate(v.shape[2:])]
                var = f.createVariable(k, v.dtype.str[1:], tuple(names))
                var[:] = v
            for ak, av in (attrs or {}).items():
                if isinstance(av, (int, float, str)):
                    setattr(f, ak, av)
        written.append(fn)
    with open(os.path.join(path, "attrs.json"), "w") as fh:
        json.dump({k: (v if isinstance(v, (int, float, str, list)) else str(v)) for k, v in (attrs or {}).items()}, fh)
    return written


def flatten_dict(dictionary, parent_key="", separator="_"):
    """scripts/model_star.py:32-41: flattens the nested ``const`` dict for the constant_data group."""
    items = []
    for key, value in dictionary.items():
        new_key = parent_key + separator + key if parent_key else key
        if isinstance(value, dict):
            items.extend(flatten_dict(value, new_key, separator=separator).items())
        else:
            items.append((new_key, value))
    return dict(items)
