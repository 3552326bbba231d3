"""The device-side sampler (csi_nuts_*, csrc/csi_nuts.cuh) against its host restatement and against the oracle's
posterior.  -m gpu."""
import json
import math
import os

import numpy as np
import pytest
import torch

import parity_util as PU
from scalable_stellar_inference_b200.nuts import (DeviceNUTS, IterativeNUTS, effective_sample_size, sample_model,
                                                  split_rhat)

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _model(name, n, engine="tcgen05", seed=0):
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n, seed)
    model = PU.product_model(name, spec.num_stars, engine).condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    return spec, model, model.flatten(z0)


def _host_potential(model):
    nh = len(model.hyper_sites)

    def pot(z):
        out, grad = model.potential_and_grad_flat(np.asarray(z, np.float32))
        g = np.array(grad, np.float32)
        g[:nh] = out[1:1 + nh].astype(np.float32)   # the reduced sums, as the device sampler sees them
        return float(out[0]), g
    return pot


@pytest.mark.parametrize("name,n,engine", [("hier_diag", 300, "tcgen05"), ("multi_informative", 40, "ffma"),
                                           ("hier_phot_grid", 700, "tcgen05")])
def test_device_transitions_match_the_host_restatement(name, n, engine):
    """csi_nuts_transition == nuts.IterativeNUTS (the same algorithm written with numpy around the SAME potential
    kernel) transition by transition: same momentum (read back from the device), same hash stream, so the trees must
    have the same depth, leaf count and proposal.  fp32 rounding differs between the fused device updates and numpy,
    so a decision that sits within ~1e-6 of its threshold may flip: require agreement on >= 85 % of the transitions
    and compare the states of those."""
    spec, model, zf = _model(name, n, engine)
    pot = _host_potential(model)
    nuts = DeviceNUTS(model, zf, seed=11, max_tree_depth=6)
    ref = IterativeNUTS(pot, seed=11, max_tree_depth=6, dtype=np.float32)
    L = zf.size
    step = (0.12 if name == "hier_phot_grid" else 0.5) * nuts.find_heuristic_step_size()   # several doublings per transition
    nuts.inv_mass.copy_(torch.rand(L, device=nuts.dev) * 0.5 + 0.75)   # exercise a non-trivial mass matrix
    torch.cuda.synchronize(nuts.dev)   # the sampler runs on its own stream
    agree, depths, total = 0, [], 12
    for t in range(total):
        z0 = nuts.z.cpu().numpy().copy()
        g0 = nuts.g.cpu().numpy().copy()
        u0 = nuts.u
        u_chk, g_chk = pot(z0)
        assert u_chk == pytest.approx(u0, rel=1e-12) and np.array_equal(g_chk, g0)   # the state is self-consistent
        draw = nuts.draw
        from scalable_stellar_inference_b200 import _lib
        r0 = np.empty(L, np.float32)
        with torch.cuda.device(nuts.dev):
            _lib.check(nuts.lib.csi_nuts_begin_only(nuts._s, step, draw))
            _lib.check(nuts.lib.csi_nuts_peek(nuts._s, 1, r0.ctypes.data))
        st = nuts.transition(step)
        assert nuts.draw == draw + 1
        out = ref.transition(z0, u0, g0, r0, nuts.inv_mass.cpu().numpy(), step, draw)
        assert st.energy0 == pytest.approx(out["h0"], rel=1e-6)
        depths.append(st.depth)
        same = (st.depth == out["depth"] and st.num_leapfrog == out["n_leapfrog"]
                and bool(st.diverging) == out["diverging"] and bool(st.turning) == out["turning"])
        if same and np.allclose(nuts.z.cpu().numpy(), out["z"], rtol=0, atol=2e-4 * (1 + np.abs(out["z"]).max())):
            agree += 1
            assert st.potential == pytest.approx(out["u"], rel=1e-5, abs=1e-3)
            assert st.accept_prob == pytest.approx(out["accept"], rel=1e-3, abs=1e-4)
        assert st.host_syncs <= st.depth and st.launches >= 3 * st.num_leapfrog
    print(name, "depths", depths, "agree", agree, "of", total)
    assert agree >= math.ceil(0.85 * total)
    assert max(depths) >= 3, "the comparison must exercise sub-trees with checkpointed U-turn tests"


def test_device_sampler_is_bit_reproducible_and_independent_of_graphs(monkeypatch):
    """Same seed -> identical chains (every sum is a fixed-order reduction, every decision a counter-based hash), with
    CUDA graphs and with plain stream launches (CSI_NUTS_NO_GRAPH=1)."""
    spec, model, zf = _model("hier_diag", 500)
    keep = lambda z: z.detach().cpu().numpy().copy()  # noqa: E731
    runs = []
    for no_graph in ("0", "0", "1"):
        monkeypatch.setenv("CSI_NUTS_NO_GRAPH", no_graph)
        out = sample_model(model, zf, 8, 6, seed=3, collect=keep, max_tree_depth=5)
        runs.append((np.array(out["samples"]), out["n_evals"], out["step_size"]))
    for a in runs[1:]:
        assert np.array_equal(a[0], runs[0][0]) and a[1] == runs[0][1] and a[2] == runs[0][2]


def test_posterior_of_device_sampler_matches_oracle_posterior():
    """north_star's posterior criterion on a HIERARCHICAL model: NUTS over the CUDA potential (device-side tree, 4
    chains) against NUTS over the fp64 oracle potential on the same 64 synthetic stars with the seismic observables
    (hier_diag_seis).  The oracle chains cost 4 x 40 min of CPU at this size (trees of depth 9-10, 2.6 ms per
    evaluation -- a 256-star fixture would take a working day), so their summary is a committed fixture
    (tests/golden/posterior_hier64_seis.json: mean, sd and ESS of the 5 hyper-parameters and of the sites of 6 stars,
    written by tools/gen_posterior_fixture.py with the host-recursive nuts.NUTS over oracle.model.potential_and_grad).
    Means must agree within Monte-Carlo standard error (both chains' ESS), spreads within 25 % where both runs have
    ESS >= 200 for the coordinate (for the decentred sites: for the hyper-parameters as well)."""
    path = os.path.join(GOLDEN, "posterior_hier64_seis.json")
    if not os.path.exists(path):
        pytest.skip("fixture not generated yet")
    fx = json.load(open(path))
    name, n = fx["case"], fx["stars"]
    spec, model, zf = _model(name, n)
    nh = len(model.hyper_sites)
    idx = np.array(fx["index"])
    keep = lambda z: z.detach().double().cpu().numpy()[idx].copy()  # noqa: E731
    chains = []
    for c in range(4):
        out = sample_model(model, zf, fx["num_warmup"], fx["num_samples"], seed=100 + c, collect=keep)
        chains.append(np.array(out["samples"]))
        assert sum(out["diverging"]) <= 0.12 * fx["num_samples"]   # the oracle chains: 11-28 of 500 (funnel)
    x = np.stack(chains)                                   # [chain, draw, coordinate]
    mean_ref, sd_ref, ess_ref = np.array(fx["mean"]), np.array(fx["sd"]), np.array(fx["ess"])
    bad, fails, hyper_resolved = 0, [], float("inf")
    for k in range(x.shape[2]):
        ess = effective_sample_size(x[:, :, k])
        mean, sd = x[:, :, k].mean(), x[:, :, k].std()
        mcse = math.sqrt(sd ** 2 / max(ess, 4.0) + sd_ref[k] ** 2 / max(ess_ref[k], 4.0))
        dev = abs(mean - mean_ref[k]) / mcse
        bad += dev > 3.5
        if not dev < 6.0:
            fails.append(("mean", k, mean, mean_ref[k], mcse))
        # a sample sd from ESS effective draws is itself uncertain by ~ 1/sqrt(2 ESS) (x2-3 for the heavy-tailed
        # funnel coordinates): 25 % only where both runs resolved the coordinate (the fixture's ESS of precision_a
        # is 38, of mu_a 20), a gross-failure guard elsewhere
        resolved = min(ess, ess_ref[k])
        if k < nh:
            hyper_resolved = min(hyper_resolved, resolved)
        elif "decentered" in model.star_sites[(idx[k] - nh) // n]:
            # Y_decentered / a_decentered are the funnel partners of the hyper-parameters: their spread follows the
            # range of precision_Y / precision_a the chains visited, whatever their own autocorrelation says
            resolved = min(resolved, hyper_resolved)
        lo_r, hi_r = (0.75, 1.33) if resolved >= 200 else (0.4, 2.5)
        if not lo_r < sd / sd_ref[k] < hi_r:
            fails.append(("sd", k, sd, sd_ref[k], ess, ess_ref[k]))
        # r-hat of the hyper-parameters: the oracle's own chains reach 1.16 on mu_a at this length, and the
        # prior-dominated neck of the funnel (precision_a, ESS < 50 in both runs) does not mix in 500 draws in
        # either implementation (the reference's own 1000-star run ends at r-hat 1.4 after 16 000 warm-up draws,
        # DESIGN section 5): a gross-failure guard, tighter where the coordinate is resolved
        if k < nh:
            rhat = split_rhat(x[:, :, k])
            if not rhat < (1.35 if resolved >= 50 else 2.5):
                fails.append(("rhat", k, rhat, ess, ess_ref[k]))
    assert not fails, fails
    assert bad <= max(1, x.shape[2] // 20)
