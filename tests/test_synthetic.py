"""The synthetic-population generator (scripts/sample_truths.py + scripts/observations.py of the
reference): host-side pieces on the CPU, the whole recipe through the GPU forward pass under -m gpu."""
from math import log10

import numpy as np
import pytest
from scipy import stats

from scalable_stellar_inference_b200 import synthetic


class _FakeModel:
    """What the host-side helpers read from a model: const, site names, log_zx_sun."""
    log_zx_sun = log10(0.0181)

    def __init__(self, hier=True, phot=False):
        from scalable_stellar_inference_b200.numpyro_models import _default_const
        self.const = _default_const()
        self.hyper_sites = ["Y_0", "dY_dZ", "precision_Y", "mu_a", "precision_a"] if hier else []
        star = (["scaled_precision", "log_evol", "log_mass", "M_H", "Y_decentered", "a_decentered"] if hier else
                ["log_evol", "log_mass", "M_H", "precision_Y", "Y", "precision_a", "a_MLT", "scaled_precision"])
        self.star_sites = star + (["scaled_distance"] if phot else [])
        self.site_names = self.hyper_sites + self.star_sites


def test_truncnorm_draws_follow_the_truncated_normal():
    rng = np.random.default_rng(0)
    x = synthetic.truncnorm(rng, 0.0, 0.5, -0.9, 0.5, 40000)
    assert x.min() >= -0.9 and x.max() <= 0.5
    ks = stats.kstest(x, stats.truncnorm(-0.9 / 0.5, 0.5 / 0.5, loc=0.0, scale=0.5).cdf)
    assert ks.pvalue > 1e-3
    y = synthetic.truncnorm(rng, -0.7, 0.4, -np.inf, 0.0, 40000)          # one-sided (log_evol)
    assert y.max() <= 0.0 and stats.kstest(y, stats.truncnorm(-np.inf, 0.7 / 0.4, loc=-0.7, scale=0.4).cdf).pvalue > 1e-3


def test_truths_follow_the_hyper_truths_and_round_trip_through_z():
    """scripts/sample_truths.py:17-23: mu_a 2.0, sigma_a 0.05, Y_0 0.247, dY_dZ 1.5, sigma_Y 0.005."""
    m = _FakeModel(hier=True, phot=True)
    rng = np.random.default_rng(1)
    t = synthetic.draw_star_truths(m, 20000, rng)
    assert abs(t["a_MLT"].mean() - 2.0) < 2e-3 and abs(t["a_MLT"].std() - 0.05) < 2e-3
    f = 1.5 / (10 ** -(t["M_H"] + m.log_zx_sun) + 1)
    resid = t["Y"] - (0.247 + f) / (1 + f)
    assert abs(resid.std() - 0.005) < 5e-4 and abs(resid.mean()) < 3e-4           # Y ~ TN(mu_Y(M_H), sigma_Y) (truncation at 0.24 bites a little)
    assert abs(t["scaled_precision"].mean() - 1.0) < 0.03                          # Gamma(df/2, df/2) has mean 1
    assert abs(t["scaled_distance"].mean() - 0.75) < 0.01                          # Beta(3, 1)
    z = synthetic.truths_to_z(m, t)
    assert list(z) == m.site_names
    sig = lambda u: 1 / (1 + np.exp(-u))  # noqa: E731
    c = m.const
    np.testing.assert_allclose(c["log_mass"]["low"] + (c["log_mass"]["high"] - c["log_mass"]["low"]) * sig(z["log_mass"]),
                               t["log_mass"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(0.0 - np.exp(z["log_evol"]), t["log_evol"], atol=1e-12)
    np.testing.assert_allclose(0.24 + 0.08 * sig(z["Y_decentered"]), t["Y"], atol=1e-12)
    np.testing.assert_allclose(np.exp(z["scaled_precision"]), t["scaled_precision"], rtol=1e-12)
    assert float(np.exp(z["precision_Y"])) == pytest.approx(0.005 ** -2) and float(3 * sig(z["dY_dZ"])) == pytest.approx(1.5)
    m2 = _FakeModel(hier=False)
    z2 = synthetic.truths_to_z(m2, synthetic.draw_star_truths(m2, 50, rng))
    assert list(z2) == m2.site_names and z2["precision_Y"].shape == (50,)


def test_synthetic_bc_grid_has_the_mist_table_shape():
    pts, vals = synthetic.synthetic_bc_grid(bands=3)
    assert [len(p) for p in pts] == [70, 26, 18, 13] and vals.shape == (70, 26, 18, 13, 3)
    assert all(np.all(np.diff(p) > 0) for p in pts) and np.all(np.isfinite(vals))
    pts2, vals2 = synthetic.synthetic_bc_grid(bands=3)
    assert all(np.array_equal(a, b) for a, b in zip(pts, pts2)) and np.array_equal(vals, vals2)   # deterministic


@pytest.mark.gpu
def test_make_problem_on_the_gpu():
    """The full recipe: prior draws -> csi_predict forward -> rejection of Teff >= 7500 K / age >= 14 Gyr
    (scripts/clean_nearest_neighbour_truths.py:5-6,33) -> Gaussian noise of scripts/observations.py:15-25."""
    import scalable_stellar_inference_b200 as ssi
    n = 30000
    model = ssi.HierarchicalStarModel(n, ["log_Teff", "log_L", "log_Dnu"])
    z0, obs, diag, ml, ms, t = synthetic.make_problem(model, n)
    z1, obs1, *_ = synthetic.make_problem(ssi.HierarchicalStarModel(n, ["log_Teff", "log_L", "log_Dnu"]), n)
    assert np.array_equal(obs, obs1) and all(np.array_equal(z0[k], z1[k]) for k in z0)           # seeds 0 / 101 / 7
    model.condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    d = model.predict(synthetic.truths_to_z(model, t))
    assert np.all(d["log_Teff"] < np.log10(7500.0)) and np.all(d["log_age"] < np.log10(14.0))
    sig = np.array([synthetic.NOISE[k] for k in ("log_Teff", "log_L", "log_Dnu")])
    zs = (obs - np.stack([d["log_Teff"], d["log_L"], d["log_Dnu"]], 1)) / sig
    assert np.all(np.abs(zs.mean(0)) < 0.03) and np.all(np.abs(zs.std(0) - 1) < 0.03)
    np.testing.assert_allclose(diag, np.broadcast_to(sig ** 2, diag.shape))
    assert abs((ml - t["M_H"]).std() - 0.1) < 0.01 and np.all(ms == 0.1)
    # the evaluation point is the unconstrained truth + 0.1 N(0,1) jitter
    zt = synthetic.truths_to_z(model, t)
    assert abs(np.std(z0["log_mass"] - zt["log_mass"]) - 0.1) < 0.005
