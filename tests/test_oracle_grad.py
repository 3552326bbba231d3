"""Oracle self-checks: hand-written reverse mode vs torch autograd vs central differences; sharding."""
import numpy as np
import pytest

from oracle import model as M, synth
from oracle.bc import GridInterpolator, synthetic_grid
from oracle.nets import load_bc_net
from oracle.torch_ref import TorchPotential


def _grid():
    pts, vals = synthetic_grid()
    return GridInterpolator(pts, vals)


CASES = {
    "hier_diag": lambda: M.ModelSpec("hierarchical", 40, ["log_Teff", "log_g"]),
    "hier_full_seis": lambda: M.ModelSpec("hierarchical", 40, ["log_Teff", "log_L", "log_Dnu", "log_numax"], kind="full"),
    "multi_informative": lambda: M.ModelSpec("multi", 40, ["log_Teff", "log_g"], prior="informative"),
    "multi_flat": lambda: M.ModelSpec("multi", 40, ["log_Teff", "log_L", "log_Dnu"], prior="flat"),
    "single": lambda: M.ModelSpec("single", 1, ["log_Teff", "log_g"]),
    "hier_phot_grid": lambda: M.ModelSpec("hierarchical", 40, ["log_Teff", "log_g", "mag", "plx"], bands=3,
                                          bc=_grid(), bc_mode="grid"),
    "multi_phot_net_full": lambda: M.ModelSpec("multi", 40, ["mag", "plx"], bands=3, bc=load_bc_net(),
                                               bc_mode="net", kind="full"),
}


@pytest.mark.parametrize("name", list(CASES))
def test_manual_grad_matches_autograd(name):
    spec = CASES[name]()
    z0, obs, diag, ml, ms = synth.make_problem(spec)
    u, g = M.potential_and_grad(spec, z0, obs, diag, ml, ms)
    ut, gt = TorchPotential(spec, obs, diag, ml, ms).value_and_grad(z0)
    assert abs(u - ut) <= 1e-11 * abs(ut)
    assert set(g) == set(gt) == set(spec.hyper_sites() + spec.star_sites())
    for k in g:
        scale = np.max(np.abs(gt[k])) + 1e-300
        np.testing.assert_allclose(g[k], gt[k], rtol=1e-8, atol=1e-10 * scale, err_msg=k)


@pytest.mark.parametrize("name", ["hier_diag", "multi_informative", "hier_phot_grid"])
def test_central_differences(name):
    spec = CASES[name]()
    z0, obs, diag, ml, ms = synth.make_problem(spec)
    _, g = M.potential_and_grad(spec, z0, obs, diag, ml, ms)
    rng = np.random.default_rng(5)
    for site in g:
        for _ in range(2):
            zp = {k: np.array(v, float) for k, v in z0.items()}
            zm = {k: np.array(v, float) for k, v in z0.items()}
            h = 1e-6
            if np.ndim(z0[site]) == 0:
                zp[site] += h; zm[site] -= h; want = g[site]
            else:
                i = rng.integers(len(z0[site]))
                zp[site][i] += h; zm[site][i] -= h; want = g[site][i]
            fd = (M.potential(spec, zp, obs, diag, mh_loc=ml, mh_scale=ms)
                  - M.potential(spec, zm, obs, diag, mh_loc=ml, mh_scale=ms)) / (2 * h)
            assert abs(fd - want) <= 2e-5 * max(1.0, abs(want)), (site, fd, want)


def test_shard_sum_invariance():
    """Stars are conditionally independent given the hypers: per-shard potentials add up."""
    spec = CASES["hier_diag"]()
    z0, obs, diag, ml, ms = synth.make_problem(spec)
    u, g, terms = M.potential_and_grad(spec, z0, obs, diag, ml, ms, return_terms=True)
    n = spec.num_stars
    total, ghyp = -terms["hyper_logp"], {k: 0.0 for k in spec.hyper_sites()}
    parts = np.array_split(np.arange(n), 4)
    for p in parts:
        sub = M.ModelSpec("hierarchical", len(p), spec.observables)
        zs = {k: (v if np.ndim(v) == 0 else np.asarray(v)[p]) for k, v in z0.items()}
        us, gs, ts = M.potential_and_grad(sub, zs, obs[p], diag[p], ml[p], ms[p], return_terms=True)
        total += us + ts["hyper_logp"]       # shard potential without the (replicated) hyper prior
        for k in ghyp:
            ghyp[k] += gs[k]
        for k in spec.star_sites():
            np.testing.assert_allclose(gs[k], g[k][p], rtol=1e-12, atol=1e-12)
    assert abs(total - u) < 1e-9 * abs(u)
    # hyper gradients: prior part counted once per shard -> subtract the 3 extra copies
    _, g_prior_only = M.potential_and_grad(M.ModelSpec("hierarchical", 2, spec.observables),
                                           {k: (v if np.ndim(v) == 0 else np.asarray(v)[:2]) for k, v in z0.items()},
                                           obs[:2], diag[:2], ml[:2], ms[:2])
    assert np.isfinite(sum(ghyp.values()))


def test_validation_errors():
    with pytest.raises(ValueError):
        M.ModelSpec("multi", 1, ["log_Teff"])
    with pytest.raises(ValueError):
        M.ModelSpec("single", 1, ["log_Teff"], kind="none")
    with pytest.raises(ValueError):
        M.ModelSpec("single", 1, ["log_Teff"], prior="weird")


def test_oracle_matches_jax_fixture():
    """The pin this image cannot produce: U and dU/dz of the REAL reference (jax + numpyro, x64), written by
    tools/gen_jax_fixtures.py on a box that has them.  Skipped while tests/golden/potential_jax.json is absent."""
    import json
    import os
    import parity_util as PU
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "potential_jax.json")
    if not os.path.exists(path):
        pytest.skip("tests/golden/potential_jax.json not generated (needs a box with jax + numpyro): parity unpinned")
    fx = json.load(open(path))
    for name, case in fx["cases"].items():
        spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, case["stars"])
        z = {k: np.asarray(v, np.float64).reshape(np.shape(z0[k])) for k, v in case["z"].items()}
        u, g = M.potential_and_grad(spec, z, obs, diag, ml, ms)
        assert abs(u - case["U"]) <= 1e-9 * abs(case["U"]), name
        for k, v in case["grad"].items():
            np.testing.assert_allclose(np.asarray(g[k]).reshape(-1), np.asarray(v).reshape(-1), rtol=1e-7, atol=1e-8, err_msg=f"{name}/{k}")
