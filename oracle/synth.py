"""Deterministic synthetic stars / observations for tests and benches (oracle side; tests only).

Recipe from SURVEY 8d: hyper-truths of scripts/sample_truths.py:17-23, per-star draws from the
priors (src/celestify/numpyro_models.py:120-125,301-331), observation noise of
scripts/observations.py:15-25, per-star M_H prior loc = noisy M_H, scale 0.1
(scripts/model_star.py:126-128), rejection of Teff >= 7500 K or age >= 14 Gyr
(scripts/clean_nearest_neighbour_truths.py:5-6,33 -- applied by redrawing)."""
from __future__ import annotations

import numpy as np
from scipy import stats, special as sp

from . import model as M

HYPER_TRUTH = dict(Y_0=0.247, dY_dZ=1.5, sigma_Y=0.005, mu_a=2.0, sigma_a=0.05)
NOISE = dict(log_age=0.05, log_Teff=0.02 / np.log(10), log_R=0.01, log_L=0.02 / np.log(10), log_Dnu=0.01 / np.log(10),
             log_numax=0.05, log_g=0.1, mag=2e-3, plx=0.05, M_H=0.1)


def _tn(rng, loc, scale, low, high, size):
    a, b = (low - loc) / scale, (high - loc) / scale
    return stats.truncnorm.rvs(a, b, loc=loc, scale=scale, size=size, random_state=rng)


def draw_truths(n, seed=0, phot=False):
    rng = np.random.default_rng(seed)
    emu = M.load_emulator(np.float64)
    h = HYPER_TRUTH
    df = M.emulator_precision()["df"]
    out = {}
    keep = np.zeros(n, bool)
    cols = {k: np.empty(n) for k in ["log_evol", "log_mass", "M_H", "Y", "a_MLT", "scaled_precision", "scaled_distance"]}
    todo = np.arange(n)
    while todo.size:
        m = todo.size
        le = _tn(rng, -0.7, 0.4, -np.inf, 0.0, m)
        lm = _tn(rng, 0.0, 0.3, np.log10(0.7), np.log10(2.3), m)
        mh = _tn(rng, 0.0, 0.5, -0.9, 0.5, m)
        f = h["dY_dZ"] / (10 ** -(mh + M.LOG_ZX_SUN) + 1)
        mu_y = (h["Y_0"] + f) / (1 + f)
        y = _tn(rng, mu_y, h["sigma_Y"], 0.24, 0.32, m)
        a = _tn(rng, h["mu_a"], h["sigma_a"], 1.5, 2.5, m)
        tau = rng.gamma(df / 2, 2 / df, m)
        sd = rng.beta(3.0, 1.0, m)
        o = emu(np.stack([10 ** le, 10 ** lm, mh, y, a], -1))
        ok = (o[:, 1] < np.log10(7500.0)) & (o[:, 0] - 9 < np.log10(14.0))
        for k_, v in zip(cols, [le, lm, mh, y, a, tau, sd]):
            cols[k_][todo[ok]] = v[ok]
        todo = todo[~ok]
    return cols


def unconstrain_truths(spec: M.ModelSpec, t: dict):
    """Map constrained truths to the unconstrained site vector of ``spec``."""
    c = spec.const
    logit = lambda p: np.log(p) - np.log1p(-p)
    z = {}

    def tn_u(name, v):
        low, high = M._bounds(c[name])
        if np.isfinite(low) and np.isfinite(high):
            return logit((v - low) / (high - low))
        if np.isfinite(high):
            return np.log(high - v)
        if np.isfinite(low):
            return np.log(v - low)
        return v
    z["scaled_precision"] = np.log(t["scaled_precision"])
    z["log_evol"] = tn_u("log_evol", t["log_evol"])
    z["log_mass"] = tn_u("log_mass", t["log_mass"])
    z["M_H"] = tn_u("M_H", t["M_H"])
    ys = logit((t["Y"] - c["Y"]["low"]) / (c["Y"]["high"] - c["Y"]["low"]))
    as_ = logit((t["a_MLT"] - c["a_MLT"]["low"]) / (c["a_MLT"]["high"] - c["a_MLT"]["low"]))
    h = HYPER_TRUTH
    if spec.level == "hierarchical":
        z["Y_decentered"], z["a_decentered"] = ys, as_
        z["Y_0"] = np.float64(h["Y_0"])
        z["dY_dZ"] = np.float64(tn_u("dY_dZ", h["dY_dZ"]))
        z["precision_Y"] = np.float64(np.log(h["sigma_Y"] ** -2))
        z["mu_a"] = np.float64(tn_u("mu_a", h["mu_a"]))
        z["precision_a"] = np.float64(np.log(h["sigma_a"] ** -2))
    else:
        z["Y"], z["a_MLT"] = ys, as_
        if spec.prior == "informative":
            n = spec.num_stars
            z["precision_Y"] = np.full(n, np.log(h["sigma_Y"] ** -2))
            z["precision_a"] = np.full(n, np.log(h["sigma_a"] ** -2))
    if spec.phot:
        z["scaled_distance"] = logit(t["scaled_distance"])
    return {k: z[k] for k in spec.hyper_sites() + spec.star_sites()}


def make_problem(spec: M.ModelSpec, seed=0, noise_seed=101, point_seed=7, jitter=0.1):
    """Truths -> noisy observations -> (z0, obs, diag, mh_loc, mh_scale)."""
    n = spec.num_stars
    t = draw_truths(n, seed, spec.phot)
    rng = np.random.default_rng(noise_seed)
    mh_loc = np.clip(t["M_H"] + NOISE["M_H"] * rng.standard_normal(n), -0.85, 0.45)
    mh_scale = np.full(n, NOISE["M_H"])
    z = unconstrain_truths(spec, t)
    _, _, terms = M.potential_and_grad(spec, z, np.zeros((n, len(spec.obs_idx))), 1.0, mh_loc, mh_scale, return_terms=True)
    mean = terms["mean"][:, spec.obs_idx]
    sig = np.array([NOISE["mag" if nm.startswith("mag_") else nm] for nm in (spec.names[i] for i in spec.obs_idx)])
    obs = mean + sig * rng.standard_normal(mean.shape)
    diag = np.broadcast_to(sig ** 2, mean.shape).copy()
    prng = np.random.default_rng(point_seed)
    z0 = {k: np.asarray(v) + jitter * prng.standard_normal(np.shape(v)) for k, v in z.items()}
    return z0, obs, diag, mh_loc, mh_scale
