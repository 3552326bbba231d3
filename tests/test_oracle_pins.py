"""Further pins of the oracle (the numpyro boundary holds no exact reference vector, so everything that CAN
be checked against something other than the oracle's own code is checked here):

  * oracle/dists.py against scipy.stats (a second independent implementation besides torch.distributions);
  * a complete single-star potential re-assembled from scipy.stats pieces;
  * the reference's own posterior summary of a fully specified single-star problem
    (notebook/single-star-spec.ipynb cells [17]-[19], [23]): observations, const and posterior mean / sd printed by
    the reference's numpyro run.  That run used the older prior variant with log_sigma_Y / log_sigma_a sites
    (Normal(-2.3, 0.2) / Normal(-1.3, 0.2) on log10 sigma) where HEAD has Gamma precisions; Y's and a_MLT's prior
    widths are dominated by the enrichment-law / mu_a uncertainty terms (numpyro_models.py:167-193), so the
    summaries pin the likelihood + star-prior block to the printed 2-3 digits, not the whole graph;
  * the shipped weight files against the reference's HDF5 files read with the in-repo reader (skipped where
    /root/reference is absent).
"""
import math
import os

import numpy as np
import pytest
import torch
from scipy import stats

from oracle import dists as D, model as M
from oracle.nets import load_emulator
from scalable_stellar_inference_b200.nuts import NUTS, VectorSpace, effective_sample_size

REF = "/root/reference/src/celestify/data"


def test_dists_against_scipy_stats():
    rng = np.random.default_rng(0)
    x = rng.uniform(-0.8, 0.4, 50)
    for loc, scale, low, high in [(0.0, 0.5, -0.9, 0.5), (-0.7, 0.4, -np.inf, 0.0), (0.3, 0.005, 0.24, 0.32), (2.0, 0.05, 1.5, 2.5)]:
        xs = np.clip(x * 0.1 + loc, max(low, loc - 3 * scale) + 1e-9, min(high, loc + 3 * scale) - 1e-9)
        a, b = (low - loc) / scale, (high - loc) / scale
        want = stats.truncnorm.logpdf(xs, a, b, loc=loc, scale=scale)
        got = D.normal_logpdf(xs, loc, scale) - D.truncnorm_const_logz(loc, scale, low, high)
        np.testing.assert_allclose(got, want, rtol=1e-10, atol=1e-10)
    xp = rng.uniform(0.05, 4.0, 50)
    np.testing.assert_allclose(D.gamma_logpdf(xp, 2.09, 2.09), stats.gamma.logpdf(xp, 2.09, scale=1 / 2.09), rtol=1e-12)
    np.testing.assert_allclose(D.gamma_logpdf(xp * 1e4, 3.0, 1e-4), stats.gamma.logpdf(xp * 1e4, 3.0, scale=1e4), rtol=1e-12)
    xb = rng.uniform(0.01, 0.99, 50)
    np.testing.assert_allclose(D.beta_logpdf(xb, 3.0, 1.0), stats.beta.logpdf(xb, 3.0, 1.0), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(D.normal_logpdf(x, 0.247, 0.001), stats.norm.logpdf(x, 0.247, 0.001), rtol=1e-12)
    # far tails: log(Phi(b) - Phi(a)) where a naive difference underflows
    for a, b in [(-40.0, -30.0), (30.0, 40.0), (-12.0, 4.0), (-np.inf, -25.0)]:
        want = stats.norm.logcdf(b) + np.log1p(-np.exp(stats.norm.logcdf(a) - stats.norm.logcdf(b))) if b <= 0 else \
            stats.norm.logsf(a) + np.log1p(-np.exp(stats.norm.logsf(b) - stats.norm.logsf(a)))
        assert D.log_ndtr_diff(a, b)[0] == pytest.approx(want, rel=1e-11)
    # log|J| of the bijections = log |d value / du| (central differences)
    u = rng.normal(size=20)
    for low, high in [(-0.9, 0.5), (-np.inf, 0.0), (0.0, np.inf)]:
        v, dv, ld, dld = D.bounded_transform(u, low, high)
        h = 1e-6
        num = (D.bounded_transform(u + h, low, high)[0] - D.bounded_transform(u - h, low, high)[0]) / (2 * h)
        np.testing.assert_allclose(dv, num, rtol=1e-7)
        np.testing.assert_allclose(ld, np.log(np.abs(dv)), rtol=1e-10, atol=1e-12)


def _single_spec(kind="diag", prior="informative", obs=("log_Teff", "log_L", "log_Dnu"), const=None):
    return M.ModelSpec("single", 1, list(obs), kind=kind, prior=prior, const=const or {})


@pytest.mark.parametrize("kind,prior", [("diag", "flat"), ("full", "flat"), ("diag", "informative")])
def test_single_star_potential_reassembled_from_scipy(kind, prior):
    """-log joint of SingleStarModel (numpyro_models.py:148-245) written out with scipy.stats distributions and the
    oracle's emulator forward only; must equal oracle.model.potential_and_grad's U."""
    spec = _single_spec(kind, prior)
    c = spec.const
    rng = np.random.default_rng(5)
    z = {k: rng.normal(scale=0.7, size=1) for k in spec.star_sites()}
    if prior == "informative":
        z["precision_Y"] = np.log(rng.gamma(3.0, 1e4, 1))
        z["precision_a"] = np.log(rng.gamma(3.0, 1e2, 1))
    obs = np.array([[3.70, -0.5, 2.28]])
    diag = np.array([[7.5e-5, 7.5e-5, 1.9e-5]])
    u, _ = M.potential_and_grad(spec, z, obs, diag)

    def tn(name, uu):
        cc = c[name]
        low, high = cc.get("low", -np.inf), cc.get("high", np.inf)
        v, dv, ld, _ = D.bounded_transform(uu, low, high)
        a, b = (low - cc["loc"]) / cc["scale"], (high - cc["loc"]) / cc["scale"]
        return v, stats.truncnorm.logpdf(v, a, b, loc=cc["loc"], scale=cc["scale"]) + ld

    lp = 0.0
    log_evol, l = tn("log_evol", z["log_evol"]); lp += l
    log_mass, l = tn("log_mass", z["log_mass"]); lp += l
    mh, l = tn("M_H", z["M_H"]); lp += l
    ylo, yhi, alo, ahi = c["Y"]["low"], c["Y"]["high"], c["a_MLT"]["low"], c["a_MLT"]["high"]
    y, _, ldy, _ = D.bounded_transform(z["Y"], ylo, yhi)
    a_mlt, _, lda, _ = D.bounded_transform(z["a_MLT"], alo, ahi)
    if prior == "flat":
        lp += stats.uniform.logpdf(y, ylo, yhi - ylo) + ldy + stats.uniform.logpdf(a_mlt, alo, ahi - alo) + lda
    else:
        py = np.exp(z["precision_Y"])
        lp += stats.gamma.logpdf(py, 3.0, scale=1e4) + z["precision_Y"]
        zxp1 = 10.0 ** -(mh + M.LOG_ZX_SUN) + 1
        f = c["dY_dZ"]["loc"] / zxp1
        mu_y = (c["Y_0"]["loc"] + f) / (1 + f)
        fp1 = c["dY_dZ"]["loc"] / zxp1 + 1                       # numpyro_models.py:140-146
        sig = np.sqrt(1 / py + ((1 - c["Y_0"]["loc"]) / fp1 ** 2 / zxp1 * c["dY_dZ"]["scale"]) ** 2
                      + (1 / fp1 * c["Y_0"]["scale"]) ** 2)
        lp += stats.truncnorm.logpdf(y, (ylo - mu_y) / sig, (yhi - mu_y) / sig, loc=mu_y, scale=sig) + ldy
        pa = np.exp(z["precision_a"])
        lp += stats.gamma.logpdf(pa, 3.0, scale=1e2) + z["precision_a"]
        sig = np.sqrt(1 / pa + c["mu_a"]["scale"] ** 2)
        mu = c["mu_a"]["loc"]
        lp += stats.truncnorm.logpdf(a_mlt, (alo - mu) / sig, (ahi - mu) / sig, loc=mu, scale=sig) + lda
    df = c["delta"]["df"]
    tau = np.exp(z["scaled_precision"])
    lp += stats.gamma.logpdf(tau, df / 2, scale=2 / df) + z["scaled_precision"]
    x = np.stack([10.0 ** log_evol, 10.0 ** log_mass, mh, y, a_mlt], -1)
    out = load_emulator(np.float64)(x) + c["delta"]["loc"]
    mean = out @ M.A_MAT.T + M.C_VEC
    mean[:, 5] += log_mass
    mean[:, 6] += log_mass
    m = mean[0, spec.obs_idx]
    if kind == "diag":
        lp += stats.norm.logpdf(obs[0], m, np.sqrt(spec.var_ext[spec.obs_idx] / tau + diag[0])).sum()
    else:
        cov = spec.cov_ext[np.ix_(spec.obs_idx, spec.obs_idx)] / tau + np.diag(diag[0])
        lp += stats.multivariate_normal.logpdf(obs[0], m, cov)
    assert float(u) == pytest.approx(-float(np.sum(lp)), rel=1e-10, abs=1e-9)


@pytest.mark.timeout(900)
def test_posterior_matches_the_reference_notebook_summary(kat):
    """notebook/single-star-spec.ipynb: star 10, observables [log_Teff, log_L, log_Dnu] (cell [17]: y and diag),
    M_H prior from the observed metallicity (cell [6]), 10 chains x (2000 + 2000), x64 -> print_summary (cell [19])
    and az.summary (cell [23]).  NUTS over the fp64 oracle potential must land on the same posterior."""
    k = kat["posterior_single_star_spec"]
    const = {"M_H": dict(loc=k["M_H_loc"], scale=k["M_H_scale"], low=-0.9, high=0.5)}
    spec = _single_spec("diag", "informative", k["observables"], const)
    obs, diag = np.array([k["y"]]), np.array([k["diag"]])
    sites = spec.star_sites()

    def pot(z):
        zd = {s: z[i:i + 1].numpy() for i, s in enumerate(sites)}
        u, g = M.potential_and_grad(spec, zd, obs, diag)
        return float(u), torch.as_tensor(np.array([float(np.asarray(g[s]).reshape(-1)[0]) for s in sites]))

    z0 = torch.zeros(len(sites), dtype=torch.float64)
    z0[sites.index("precision_Y")] = math.log(3e4)
    z0[sites.index("precision_a")] = math.log(3e2)
    chains = []
    for c in range(3):
        out = NUTS(pot, VectorSpace(), seed=20 + c).run(z0.clone(), 300, 700, collect=lambda z: z.numpy().copy())
        chains.append(np.array(out["samples"]))
    x = np.stack(chains)
    c = spec.const

    def col(name):
        return x[:, :, sites.index(name)]

    sig = lambda u: 1 / (1 + np.exp(-u))  # noqa: E731
    post = {
        "log_evol": c["log_evol"]["high"] - np.exp(col("log_evol")),
        "log_mass": c["log_mass"]["low"] + (c["log_mass"]["high"] - c["log_mass"]["low"]) * sig(col("log_mass")),
        "M_H": -0.9 + 1.4 * sig(col("M_H")),
        "Y": 0.24 + 0.08 * sig(col("Y")),
        "a_MLT": 1.5 + sig(col("a_MLT")),
        "scaled_precision": np.exp(col("scaled_precision")),
    }
    post["evol"] = 10.0 ** post["log_evol"]
    for name, (mean, sd) in k["summary"].items():
        v = post[name]
        ess = effective_sample_size(v)
        mcse = v.std() / math.sqrt(max(ess, 10.0))
        half_digit = 0.5 * 10.0 ** -k["decimals"][name]          # the reference prints this many decimals
        assert abs(v.mean() - mean) <= half_digit + 4 * mcse + 0.04 * sd, (name, v.mean(), mean, mcse)
        assert abs(v.std() / sd - 1) <= 0.12 + half_digit / sd, (name, v.std(), sd)


@pytest.mark.skipif(not os.path.isdir(REF), reason="/root/reference is not present on this machine")
def test_shipped_weights_equal_the_reference_hdf5_files():
    """The .npz files the package ships == the datasets of the reference's HDF5 files, read with the in-repo HDF5
    reader (scalable_stellar_inference_b200/h5lite.py; emulator.py:12-19, bolometric_corrections.py:16-23)."""
    from scalable_stellar_inference_b200 import h5lite
    from scalable_stellar_inference_b200.emulator import DATA_DIR
    pairs = [("emulator_weights.h5", "emulator_weights.npz"), ("bc_weights.h5", "bc_weights.npz")]
    checked = 0
    for h5, npz in pairs:
        src, dst = os.path.join(REF, h5), os.path.join(DATA_DIR, npz)
        if not (os.path.exists(src) and os.path.exists(dst)):
            continue
        f = h5lite.H5File(src)
        got = np.load(dst)
        keys = list(f.keys())
        assert len(keys) >= 10 and sorted(keys) == sorted(got.files), (h5, keys[:4], got.files[:4])
        for name in keys:
            arr = np.asarray(f[name])
            assert arr.dtype == got[name].dtype and np.array_equal(arr, got[name]), (h5, name)
            checked += 1
    assert checked >= 20
    import json
    a = json.load(open(os.path.join(REF, "emulator_error.json")))
    b = json.load(open(os.path.join(DATA_DIR, "emulator_error.json")))
    assert a == b
