"""NUTS driver (the caller of the hot path; numpyro's role, scripts/model_star.py:132-155): CPU checks of the
sampler itself, and on the GPU posterior agreement between the CUDA potential and the oracle potential."""
import math

import numpy as np
import pytest
import torch

import parity_util as PU
from oracle import model as M
from scalable_stellar_inference_b200.nuts import NUTS, VectorSpace, effective_sample_size


def test_adaptation_windows_follow_stan_schedule():
    assert NUTS.adaptation_windows(1000) == [(75, 100), (100, 150), (150, 250), (250, 450), (450, 950)]
    assert NUTS.adaptation_windows(10) == []
    w = NUTS.adaptation_windows(100)           # 15 % / 75 % / 10 % rule
    assert w[0][0] == 15 and w[-1][1] == 90


def test_gaussian_moments_and_ess():
    d = 12
    sig = torch.linspace(0.2, 2.5, d, dtype=torch.float64)

    def pot(z):
        return float(0.5 * (z ** 2 / sig ** 2).sum()), z / sig ** 2

    out = NUTS(pot, VectorSpace(), seed=3).run(torch.full((d,), 0.3, dtype=torch.float64), 300, 800,
                                                 collect=lambda z: z.numpy().copy())
    s = np.array(out["samples"])
    ess = np.array([effective_sample_size(s[:, k]) for k in range(d)])
    assert np.all(np.abs(s.mean(0)) < 5 * sig.numpy() / np.sqrt(ess))
    assert np.all(np.abs(s.std(0) / sig.numpy() - 1) < 0.15)
    assert sum(out["diverging"]) == 0 and 0.7 < np.mean(out["accept"]) <= 1.0


def test_ess_of_independent_and_correlated_draws():
    rng = np.random.default_rng(0)
    x = rng.standard_normal(4000)
    assert 3000 < effective_sample_size(x) < 5500
    y = np.zeros(4000)
    for i in range(1, 4000):                    # AR(1), rho = 0.9 -> ESS ~ n (1 - rho) / (1 + rho) ~ 210
        y[i] = 0.9 * y[i - 1] + rng.standard_normal()
    assert 100 < effective_sample_size(y) < 450


def test_sharded_inner_products_match_unsharded():
    # [hyper | stars] split over two "ranks" by hand: hyper counted once, star parts summed
    rng = np.random.default_rng(1)
    nh, n = 3, 10
    a, b = torch.tensor(rng.standard_normal(nh + n)), torch.tensor(rng.standard_normal(nh + n))
    full = VectorSpace(nh).dots([(a, b)])[0]
    parts = 0.0
    for lo, hi in ((0, 4), (4, 10)):
        sl = torch.cat([a[:nh], a[nh + lo: nh + hi]]), torch.cat([b[:nh], b[nh + lo: nh + hi]])
        parts += VectorSpace(0).dots([(sl[0][nh:], sl[1][nh:])])[0]
    parts += float(torch.dot(a[:nh], b[:nh]))
    assert math.isclose(full, parts, rel_tol=1e-12)


@pytest.mark.gpu
def test_posterior_of_cuda_potential_matches_oracle_potential():
    """Posterior means / spreads of every site: NUTS over libcsi vs the same NUTS over the fp64 oracle on the same
    4 synthetic stars (MultiStarModel, informative prior); agreement within Monte-Carlo standard error
    (north_star's posterior criterion).  The hierarchical posterior needs tree depth 9-10 at this size -- minutes of
    oracle time -- so the cheap well-conditioned model carries this test."""
    from scalable_stellar_inference_b200.nuts import sample_model
    name, n = "multi_informative", 4
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    model = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    zf = model.flatten(z0)
    nw, ns = 150, 250
    keep = lambda z: z.detach().double().cpu().numpy().copy()  # noqa: E731
    out_gpu = sample_model(model, zf, nw, ns, seed=5, collect=keep)

    def oracle_pot(z):
        zn = z.numpy()
        if not np.all(np.abs(zn) < 1e30):   # the step-size search probes absurd steps: a non-finite energy, as on the GPU
            return math.inf, torch.zeros_like(z)
        zd = model.unflatten(zn.astype(np.float32))
        with np.errstate(all="ignore"):
            u, g = M.potential_and_grad(spec, zd, obs, diag, ml, ms)
        gf = model.flatten(g).astype(np.float64)
        if not (np.isfinite(u) and np.all(np.isfinite(gf))):
            return math.inf, torch.zeros_like(z)
        return float(u), torch.as_tensor(gf)

    out_cpu = NUTS(oracle_pot, VectorSpace(len(model.hyper_sites)), seed=6).run(
        torch.as_tensor(zf.astype(np.float64)), nw, ns, collect=keep)
    a, b = np.array(out_gpu["samples"]), np.array(out_cpu["samples"])
    bad = 0
    for k in range(a.shape[1]):
        ea, eb = effective_sample_size(a[:, k]), effective_sample_size(b[:, k])
        mcse = math.sqrt(a[:, k].var() / ea + b[:, k].var() / eb)
        bad += abs(a[:, k].mean() - b[:, k].mean()) > 4.0 * mcse
        assert abs(a[:, k].mean() - b[:, k].mean()) < 6.0 * mcse, (k, a[:, k].mean(), b[:, k].mean(), mcse)
        assert 0.4 < a[:, k].std() / b[:, k].std() < 2.5
    assert bad <= 1
    assert sum(out_gpu["diverging"]) <= 0.1 * ns


@pytest.mark.gpu
@pytest.mark.parametrize("name,n", [("hier_phot_grid", 3000), ("multi_informative", 64)])
def test_fused_leapfrog_and_uturn_kernels_match_the_vector_ops(name, n):
    """csi_leapfrog_pre / csi_leapfrog_post / csi_uturn == the same step written with torch ops around the potential
    (velocity-Verlet and the generalised U-turn dots), for a model with and one without hyper-parameters."""
    from scalable_stellar_inference_b200.nuts import DevicePotential
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    model = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    dev = torch.device("cuda", model.device)
    pot = DevicePotential(model)
    nh = pot.nh
    gen = torch.Generator(device=dev).manual_seed(3)
    z = torch.as_tensor(model.flatten(z0), device=dev)
    r = torch.randn(z.shape, generator=gen, device=dev)
    im = torch.rand(z.shape, generator=gen, device=dev) + 0.5
    u0, g0 = pot(z)
    for eps in (1e-3, -2e-3):
        # reference: torch ops
        rh = r - (0.5 * eps) * g0
        z1_ref = z + eps * (im * rh)
        u1_ref, g1_ref = pot(z1_ref)
        r1_ref = rh - (0.5 * eps) * g1_ref
        k_ref = 0.5 * float(torch.dot((im * r1_ref).double(), r1_ref.double()))
        z1, r1, g1, u1, k1 = pot.leapfrog(z, r, g0, eps, im)
        assert torch.allclose(z1, z1_ref, rtol=0, atol=1e-6 * float(z1_ref.abs().max()))
        assert u1 == pytest.approx(u1_ref, rel=1e-6)   # z1 differs by an ulp (fused multiply-add), U is steep
        assert torch.allclose(g1, g1_ref, rtol=1e-5, atol=1e-5 * float(g1_ref.abs().max()))
        assert torch.allclose(r1, r1_ref, rtol=1e-5, atol=1e-5 * float(r1_ref.abs().max()))
        assert k1 == pytest.approx(k_ref, rel=1e-6)
        if nh:   # hyper slots of the returned gradient are the reduced sums, not the fp32 local partials
            assert torch.equal(g1[:nh], pot.sp.out_dev[1:1 + nh].float())
    r_left, r_right = r, torch.randn(z.shape, generator=gen, device=dev)
    r_sum = r_left + r_right + torch.randn(z.shape, generator=gen, device=dev)
    rc = (r_sum - 0.5 * (r_left + r_right)).double()
    a_ref = float(torch.dot((im * r_left).double(), rc)); b_ref = float(torch.dot((im * r_right).double(), rc))
    a, b = pot.uturn(r_left, r_right, r_sum, im)
    assert a == pytest.approx(a_ref, rel=1e-6, abs=1e-6 * n) and b == pytest.approx(b_ref, rel=1e-6, abs=1e-6 * n)


def test_hash_mirror_matches_libcsi():
    """nuts.hash_u64 (Python) == csi::nuts::hash_u64 (the stream behind every device-side decision)."""
    from scalable_stellar_inference_b200 import _lib, nuts
    lib = _lib.load()
    rng = np.random.default_rng(0)
    for _ in range(200):
        a = [int(v) for v in rng.integers(0, 2 ** 63, 4)]
        a[2] = a[2] % 5
        assert lib.csi_nuts_hash(*a) == nuts.hash_u64(*a)
    u = [nuts.hash_uniform(7, d, nuts.K_LEAF, 5) for d in range(4000)]
    assert 0.0 <= min(u) and max(u) < 1.0 and abs(np.mean(u) - 0.5) < 0.02


def test_leaf_to_checkpoint_indices():
    from scalable_stellar_inference_b200.nuts import leaf_to_ckpt
    # numpyro hmc_util._leaf_idx_to_ckpt_idxs docstring examples: 6 -> idx_max 2, no test; 7 -> 3 sub-trees; 13 -> 1
    assert leaf_to_ckpt(6) == (3, 2) and leaf_to_ckpt(7) == (0, 2) and leaf_to_ckpt(13) == (2, 2)
    assert leaf_to_ckpt(0) == (1, 0) and leaf_to_ckpt(1) == (0, 0)


def test_iterative_nuts_samples_a_gaussian():
    """The host restatement of the device sampler (IterativeNUTS) targets the right distribution: fixed step size,
    correlated scales, moments within Monte-Carlo error; trajectories stop by U-turn, not by depth."""
    from scalable_stellar_inference_b200.nuts import IterativeNUTS
    d = 10
    sig = np.linspace(0.3, 2.0, d)

    def pot(z):
        z = np.asarray(z, np.float64)
        return float(0.5 * np.sum(z ** 2 / sig ** 2)), z / sig ** 2

    rng = np.random.default_rng(4)
    nuts = IterativeNUTS(pot, seed=9, max_tree_depth=8, dtype=np.float64)
    z = np.full(d, 0.2)
    u, g = pot(z)
    im = np.ones(d)
    draws, depth, acc = [], [], []
    for t in range(1500):
        out = nuts.transition(z, u, g, rng.standard_normal(d), im, 0.25, t)
        z, u, g = out["z"], out["u"], out["g"]
        draws.append(z.copy()); depth.append(out["depth"]); acc.append(out["accept"])
    s = np.array(draws[200:])
    ess = np.array([effective_sample_size(s[:, k]) for k in range(d)])
    assert np.all(np.abs(s.mean(0)) < 5 * sig / np.sqrt(ess))
    assert np.all(np.abs(s.std(0) / sig - 1) < 0.15)
    assert 2 <= np.mean(depth) < 7.5 and np.mean(acc) > 0.7
