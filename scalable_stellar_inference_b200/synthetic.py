"""Synthetic stellar populations for benchmarks and sampler checks (no external datasets).

Follows the reference's recipe: hyper-truths of ``scripts/sample_truths.py:17-23``; per-star truths drawn from the
model's own priors (``src/celestify/numpyro_models.py:120-125,301-331``); Gaussian observation noise of
``scripts/observations.py:15-25`` (+ the photometric / seismic scales printed in the single-star notebooks); the
observed metallicity enters as a per-star M_H prior (``scripts/model_star.py:126-128``); stars hotter than 7500 K
or older than 14 Gyr are redrawn (``scripts/clean_nearest_neighbour_truths.py:5-6,33``).  Where the reference
snaps draws to an external MESA grid, the emulator itself (csi_predict on the GPU) supplies the "true" observables.
"""
from __future__ import annotations

from math import inf, log

import numpy as np
from scipy import special as sp

HYPER_TRUTH = dict(Y_0=0.247, dY_dZ=1.5, sigma_Y=0.005, mu_a=2.0, sigma_a=0.05)
NOISE = dict(log_age=0.05, log_Teff=0.02 / log(10), log_R=0.01, log_L=0.02 / log(10), log_Dnu=0.01 / log(10),
             log_numax=0.05, log_g=0.1, mag=2e-3, plx=0.05, M_H=0.1)


def truncnorm(rng, loc, scale, low, high, size):
    """Inverse-CDF draws from Normal(loc, scale) truncated to [low, high] (arrays broadcast)."""
    a = sp.ndtr((low - loc) / scale) if np.all(np.isfinite(low)) else 0.0
    b = sp.ndtr((high - loc) / scale) if np.all(np.isfinite(high)) else 1.0
    u = rng.uniform(size=size)
    return loc + scale * sp.ndtri(np.clip(a + u * (b - a), 1e-300, 1 - 1e-16))


def synthetic_bc_grid(bands=3, seed=3, shape=(70, 26, 18, 13)):
    """A smooth stand-in for the MIST bolometric-correction table (same rank and extent, non-uniform axes);
    the real table needs the third-party `isochrones` package and a download (bolometric_corrections.py:28-45)."""
    rng = np.random.default_rng(seed)

    def axis(lo, hi, n, geometric=False):
        u = np.linspace(0.0, 1.0, n)
        u[1:-1] += rng.uniform(-0.3, 0.3, n - 2) / (n - 1)
        return lo * (hi / lo) ** u if geometric else lo + (hi - lo) * u

    pts = [axis(2500.0, 50000.0, shape[0], True), axis(-1.0, 6.0, shape[1]), axis(-2.0, 0.75, shape[2]),
           axis(0.0, 4.0, shape[3])]
    T, G, F, A = np.meshgrid(*pts, indexing="ij")
    lt = np.log10(T) - 3.76
    vals = np.empty(tuple(shape) + (bands,))
    for b in range(bands):
        c = rng.normal(size=8)
        vals[..., b] = (0.1 * c[0] - (2.0 + 0.5 * c[1]) * lt ** 2 + (0.8 * b - 0.6 + 0.1 * c[2]) * lt
                        + 0.02 * c[3] * (G - 4.4) + 0.05 * c[4] * F + 0.03 * c[5] * F * lt
                        - (0.8 + 0.1 * b + 0.02 * c[6]) * A + 0.01 * c[7] * A * lt)
    return pts, vals


def _logit(p):
    return np.log(p) - np.log1p(-p)


def _unconstrain(v, low=-inf, high=inf):
    lo, hi = np.isfinite(low), np.isfinite(high)
    if lo and hi:
        return _logit((v - low) / (high - low))
    if hi:
        return np.log(high - v)
    if lo:
        return np.log(v - low)
    return v


def draw_star_truths(model, n, rng):
    c = model.const
    h = HYPER_TRUTH
    b = lambda d, k, default: d.get(k, default)  # noqa: E731
    le = truncnorm(rng, c["log_evol"]["loc"], c["log_evol"]["scale"], b(c["log_evol"], "low", -inf), b(c["log_evol"], "high", inf), n)
    lm = truncnorm(rng, c["log_mass"]["loc"], c["log_mass"]["scale"], c["log_mass"]["low"], c["log_mass"]["high"], n)
    mh = truncnorm(rng, 0.0, 0.5, c["M_H"]["low"], c["M_H"]["high"], n)
    f = h["dY_dZ"] / (10 ** -(mh + model.log_zx_sun) + 1)
    mu_y = (h["Y_0"] + f) / (1 + f)
    y = truncnorm(rng, mu_y, h["sigma_Y"], c["Y"]["low"], c["Y"]["high"], n)
    a = truncnorm(rng, h["mu_a"], h["sigma_a"], c["a_MLT"]["low"], c["a_MLT"]["high"], n)
    df = c["delta"]["df"]
    tau = rng.gamma(df / 2, 2 / df, n)
    sd = rng.beta(c["scaled_distance"]["concentration1"], c["scaled_distance"]["concentration0"], n)
    return dict(log_evol=le, log_mass=lm, M_H=mh, Y=y, a_MLT=a, scaled_precision=tau, scaled_distance=sd)


def truths_to_z(model, t):
    """Constrained truths -> the model's unconstrained site dict (hyper sites at HYPER_TRUTH)."""
    c, h = model.const, HYPER_TRUTH
    ys = _logit((t["Y"] - c["Y"]["low"]) / (c["Y"]["high"] - c["Y"]["low"]))
    as_ = _logit((t["a_MLT"] - c["a_MLT"]["low"]) / (c["a_MLT"]["high"] - c["a_MLT"]["low"]))
    n = len(ys)
    z = dict(scaled_precision=np.log(t["scaled_precision"]),
             log_evol=_unconstrain(t["log_evol"], c["log_evol"].get("low", -inf), c["log_evol"].get("high", inf)),
             log_mass=_unconstrain(t["log_mass"], c["log_mass"]["low"], c["log_mass"]["high"]),
             M_H=_unconstrain(t["M_H"], c["M_H"]["low"], c["M_H"]["high"]),
             Y=ys, a_MLT=as_, Y_decentered=ys, a_decentered=as_,
             scaled_distance=_logit(t["scaled_distance"]),
             Y_0=np.float64(h["Y_0"]),
             dY_dZ=np.float64(_unconstrain(h["dY_dZ"], c["dY_dZ"]["low"], c["dY_dZ"]["high"])),
             mu_a=np.float64(_unconstrain(h["mu_a"], c["mu_a"]["low"], c["mu_a"]["high"])))
    if model.hyper_sites:
        z["precision_Y"], z["precision_a"] = np.float64(np.log(h["sigma_Y"] ** -2)), np.float64(np.log(h["sigma_a"] ** -2))
    else:
        z["precision_Y"], z["precision_a"] = np.full(n, np.log(h["sigma_Y"] ** -2)), np.full(n, np.log(h["sigma_a"] ** -2))
    return {k: z[k] for k in model.site_names}


def make_problem(model, n, seed=0, noise_seed=101, point_seed=7, jitter=0.1, max_rounds=40):
    """Truths -> (GPU forward) -> noisy observations.  Returns (z0, obs, diag, M_H_loc, M_H_scale, truths) for the
    FULL population of n stars; shard it with ``parallel.shard_problem``.  Leaves ``model`` conditioned on the
    full population when it is not sharded."""
    rng = np.random.default_rng(seed)
    k = len(model.observable_indices)
    t = draw_star_truths(model, n, rng)
    todo = np.arange(n)
    for _ in range(max_rounds):
        model.condition(np.zeros((n, k)), np.ones((n, k)))
        d = model.predict(truths_to_z(model, t))
        bad = ~((d["log_Teff"] < np.log10(7500.0)) & (d["log_age"] < np.log10(14.0)))
        if model.bands:
            bad |= ~np.isfinite(d[model.output_names[7]])
        todo = np.nonzero(bad)[0]
        if todo.size == 0:
            break
        new = draw_star_truths(model, todo.size, rng)
        for key in t:
            t[key][todo] = new[key]
    else:
        raise RuntimeError("rejection sampling of synthetic stars did not converge")
    z = truths_to_z(model, t)
    nrng = np.random.default_rng(noise_seed)
    mh_loc = np.clip(t["M_H"] + NOISE["M_H"] * nrng.standard_normal(n), -0.85, 0.45)
    mh_scale = np.full(n, NOISE["M_H"])
    mean = np.stack([d[model.output_names[i]] for i in model.observable_indices], axis=1).astype(np.float64)
    sig = np.array([NOISE["mag" if model.output_names[i].startswith("mag_") else model.output_names[i]]
                    for i in model.observable_indices])
    obs = mean + sig * nrng.standard_normal(mean.shape)
    diag = np.broadcast_to(sig ** 2, mean.shape).copy()
    prng = np.random.default_rng(point_seed)
    z0 = {key: np.asarray(v) + jitter * prng.standard_normal(np.shape(v)) for key, v in z.items()}
    return z0, obs, diag, mh_loc, mh_scale, t
