"""The adapters and driver mirrors over the C ABI on the GPU: torch.autograd.Function, the
model callable, MCMC / Predictive with numpyro's call shapes, and the reference notebook's single-star posterior
through the CUDA path (BASELINE config 1).  -m gpu."""
import math

import numpy as np
import pytest
import torch

import parity_util as PU
from oracle import model as M
from oracle.nets import load_emulator

pytestmark = pytest.mark.gpu


def test_torch_potential_function_gradients():
    """U = potential_function(model)(z): value and autograd gradient == the C ABI's (value, gradient) == the oracle's
    (standard parity bounds); the chain rule through torch ops composes (d(2U + sum z)/dz = 2 g + 1)."""
    from scalable_stellar_inference_b200.torch_adapter import potential_fn, potential_function
    name, n = "hier_phot_grid", 1500   # the size of the main parity test (the L2 criterion sits at the fp32 floor)
    spec, problem = PU.make_problem(name, n)
    z0, obs, diag, ml, ms = problem
    model = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    zf = torch.tensor(model.flatten(z0), device="cuda", requires_grad=True)
    u = potential_function(model)(zf)
    assert u.dtype == torch.float64
    (2.0 * u + zf.double().sum()).backward()
    out, grad = model.potential_and_grad_flat(model.flatten(z0))
    nh = len(model.hyper_sites)
    g = np.array(grad, np.float64)
    g[:nh] = out[1:1 + nh]
    assert float(u.detach()) == pytest.approx(out[0], rel=1e-12)
    np.testing.assert_allclose(zf.grad.cpu().numpy(), 2.0 * g.astype(np.float32) + 1.0, rtol=1e-5, atol=1e-5)
    # dict interface, as numpyro's potential_fn takes it, against the oracle
    zd = {k: torch.tensor(np.asarray(v, np.float32), device="cuda", requires_grad=True) for k, v in z0.items()}
    u2 = potential_fn(model)(zd)
    u2.backward()
    assert float(u2.detach()) == pytest.approx(out[0], rel=1e-12)
    u_ref, g_ref = M.potential_and_grad(spec, z0, obs, diag, ml, ms)
    gd = {k: v.grad.cpu().numpy().astype(np.float64).reshape(np.shape(g_ref[k])) for k, v in zd.items()}
    PU.check_parity(name, n, float(u2), gd, u_ref, g_ref, problem, spec.hyper_sites(), spec.star_sites())


def test_torch_emulator_function_input_vjp():
    """emulator_function: forward == oracle net, autograd's input gradient == the oracle's input-VJP (both engines of
    the operator: CUDA cores below 2048 rows, tcgen05 above)."""
    import scalable_stellar_inference_b200 as ssi
    from scalable_stellar_inference_b200.torch_adapter import emulator_function
    emu = ssi.Emulator("cuda")
    ref = load_emulator(np.float64)
    f = emulator_function(emu)
    rng = np.random.default_rng(3)
    for n in (300, 5000):
        x = np.stack([rng.uniform(0.01, 1.0, n), rng.uniform(0.7, 2.3, n), rng.uniform(-0.9, 0.5, n),
                      rng.uniform(0.24, 0.32, n), rng.uniform(1.5, 2.5, n)], -1).astype(np.float32)
        w = rng.standard_normal((n, 4)).astype(np.float32)
        xt = torch.tensor(x, device="cuda", requires_grad=True)
        y = f(xt)
        (y * torch.tensor(w, device="cuda")).sum().backward()
        y_ref, masks = ref.forward(x.astype(np.float64), cache=True)
        gx_ref = ref.vjp(masks, w.astype(np.float64))
        np.testing.assert_allclose(y.detach().cpu().numpy(), y_ref, rtol=0, atol=1e-5)   # log_age ~ 10: 1 ulp = 1e-6
        gx = xt.grad.cpu().numpy()
        assert np.linalg.norm(gx - gx_ref) / np.linalg.norm(gx_ref) < 2e-3
        r = np.abs(gx - gx_ref) / (np.abs(gx_ref) + np.sqrt(np.mean(gx_ref ** 2)))
        assert np.quantile(r, 0.99) < 1e-4


def test_model_callable_and_predictive():
    """model() with obs=None is a prior-predictive trace (tests/test_numpyro_models.py:8-17 of the reference: one star,
    one trace); model(obs=, diag=) conditions; Predictive samples y around predict's means with variance / tau."""
    import scalable_stellar_inference_b200 as ssi
    from scalable_stellar_inference_b200.infer import Predictive
    model = ssi.SingleStarModel(["log_Teff", "log_g"])
    tr = model(obs=None, rng_key=1)
    assert tr["y"].shape == (7,) and np.all(np.isfinite(tr["y"]))
    for k in ("evol", "mass", "M_H", "Y", "a_MLT", "scaled_precision", "log_Teff"):
        assert k in tr
    assert abs(tr["log_g"] - (-2 * tr["log_R"] + np.log10(tr["mass"]) + 4.438067655544355)) < 1e-4
    multi = ssi.MultiStarModel(64, ["log_Teff", "log_L"])
    prior = Predictive(multi, num_samples=40)(7)
    assert prior["y"].shape == (40, 64, 7)
    resid = (prior["y"] - np.stack([prior[k] for k in multi.output_names], -1)) * np.sqrt(prior["scaled_precision"][..., None] / multi.variance)
    assert abs(resid.std() - 1.0) < 0.05 and abs(resid.mean()) < 0.05
    assert multi(obs=np.zeros((64, 2)) + [3.7, 0.0], diag=1e-4) is multi and multi.n_local == 64
    with pytest.raises(ValueError, match="zero total variance"):
        ssi.MultiStarModel(8, ["plx"], bands=["G", "BP", "RP"], bc="grid", bc_grid=PU.grid()).condition(np.ones((8, 1)))


def test_mcmc_single_star_reproduces_the_reference_notebook(kat, tmp_path):
    """BASELINE config 1 through the CUDA path with the reference's driver shape: NUTS(model, ...) -> MCMC(...).run(key,
    obs=y, diag=diag) -> print_summary / get_samples / Predictive / results files, on the fully specified problem of
    notebook/single-star-spec.ipynb (cells [5]-[6], [17]-[19], [23]); the posterior means and spreads the reference
    printed must come out (same tolerances as the oracle-side test, tests/test_oracle_pins.py)."""
    import scalable_stellar_inference_b200 as ssi
    from scalable_stellar_inference_b200 import infer
    k = kat["posterior_single_star_spec"]
    const = {"M_H": dict(loc=k["M_H_loc"], scale=k["M_H_scale"], low=-0.9, high=0.5)}
    model = ssi.SingleStarModel(k["observables"], const=const)
    sampler = infer.NUTS(model, target_accept_prob=0.8, find_heuristic_step_size=True)
    mcmc = infer.MCMC(sampler, num_warmup=500, num_samples=1000, num_chains=4)
    mcmc.run(3, obs=np.array(k["y"]), diag=np.array(k["diag"]))
    mcmc.print_summary()
    post = mcmc.get_samples(group_by_chain=True)
    stats = mcmc.get_extra_fields(group_by_chain=True)
    assert post["log_mass"].shape == (4, 1000) and stats["diverging"].shape == (4, 1000)
    assert stats["diverging"].mean() < 0.05
    rows = mcmc.summary()
    for name, (mean, sd) in k["summary"].items():
        v = post[name]
        mcse = v.std() / math.sqrt(max(rows[name]["n_eff"], 10.0))
        half_digit = 0.5 * 10.0 ** -k["decimals"][name]
        assert abs(v.mean() - mean) <= half_digit + 4 * mcse + 0.04 * sd, (name, v.mean(), mean, mcse)
        assert abs(v.std() / sd - 1) <= 0.12 + half_digit / sd, (name, v.std(), sd)
        assert rows[name]["r_hat"] < 1.05
    y = infer.Predictive(model, posterior_samples=infer.get_unconstrained(mcmc), return_sites=["y"], batch_ndims=2)(4)["y"]
    assert y.shape == (4, 1000, 7)
    # posterior-predictive log_Teff / log_L / log_Dnu against the notebook's az.summary (cell [23]: 3.703, -0.5163, 2.278)
    for j, want, sd in ((1, 3.703, 0.003654), (4, -0.5163, 0.01097), (3, 2.278, 0.005571)):
        assert abs(y[..., j].mean() - want) < 0.0006 + 0.15 * sd and abs(y[..., j].std() / sd - 1) < 0.2
    files = infer.save_inference_data(str(tmp_path / "run"), posterior=dict(post, y=y), sample_stats=stats,
                                      constant_data=infer.flatten_dict(model.const) | mcmc._kwargs,
                                      attrs=dict(star=10, observables=",".join(model.observables)))
    assert len(files) == 3
