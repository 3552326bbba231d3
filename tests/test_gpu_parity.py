"""Parity of the CUDA potential (through the C ABI / Python host layer) with the fp64 oracle.  -m gpu."""
import numpy as np
import pytest

import parity_util as PU
from oracle import model as M

pytestmark = pytest.mark.gpu

ENGINES = ["ffma", "tcgen05"]


def _eval(name, n, engine, seed=0):
    spec, problem = PU.make_problem(name, n, seed)
    z0, obs, diag, ml, ms = problem
    u_ref, g_ref = M.potential_and_grad(spec, z0, obs, diag, ml, ms)
    model = PU.product_model(name, spec.num_stars, engine)
    assert model.hyper_sites == spec.hyper_sites() and model.star_sites == spec.star_sites()
    model.condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    u, g = model.potential_and_grad(z0)
    return spec, problem, u, g, u_ref, g_ref, model


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("name", list(PU.CASES))
def test_potential_parity(name, engine):
    if not PU.engine_supports(name, engine):
        pytest.skip("BC-emulator net inside the potential: FFMA engine only this round")
    n = 1500
    spec, problem, u, g, u_ref, g_ref, model = _eval(name, n, engine)
    rep = PU.check_parity(name, n, u, g, u_ref, g_ref, problem, spec.hyper_sites(), spec.star_sites())
    print(name, engine, {k: (v if not isinstance(v, tuple) else tuple(f"{x:.1e}" for x in v)) for k, v in rep.items()})
    assert model.info()["launches"] == 1


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("n", [2, 127, 128, 129, 257, 1000])
def test_ragged_sizes(n, engine):
    """Tile tails: star counts around the 128-star tile / 256-star tile pair boundaries."""
    name = "hier_diag"
    spec, problem, u, g, u_ref, g_ref, _ = _eval(name, n, engine, seed=n)
    PU.check_parity(name, n, u, g, u_ref, g_ref, problem, spec.hyper_sites(), spec.star_sites())


@pytest.mark.parametrize("engine", ENGINES)
def test_three_points_config3(engine):
    """BASELINE config 3 (hierarchical spectroscopic, 10k stars) at 3 fixed points."""
    name, n = "hier_diag", 10000
    spec, problem = PU.make_problem(name, n)
    z0, obs, diag, ml, ms = problem
    model = PU.product_model(name, n, engine).condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    rng = np.random.default_rng(11)
    for p in range(3):
        z = PU.f32_point({k: np.asarray(v) + (0.05 * p) * rng.standard_normal(np.shape(v)) for k, v in z0.items()})
        u_ref, g_ref = M.potential_and_grad(spec, z, obs, diag, ml, ms)
        u, g = model.potential_and_grad(z)
        PU.check_parity(name, n, u, g, u_ref, g_ref, (z, obs, diag, ml, ms), spec.hyper_sites(), spec.star_sites())


def test_no_diag_no_mh_arrays():
    """diag=None (= 0, numpyro_models.py:229-230) and the constant M_H prior."""
    name, n = "multi_flat", 300
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    u_ref, g_ref = M.potential_and_grad(spec, z0, obs, None, None, None)
    for engine in ENGINES:
        model = PU.product_model(name, n, engine).condition(obs)
        u, g = model.potential_and_grad(z0)
        assert abs(u - u_ref) <= 1e-5 * abs(u_ref)
        for k in spec.star_sites():
            l2 = np.linalg.norm(g[k] - g_ref[k]) / np.linalg.norm(g_ref[k])
            assert l2 < 2e-3, (k, l2)


def test_engines_agree_and_deterministic():
    name, n = "hier_diag_seis", 5000
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    res = {}
    for engine in ENGINES:
        model = PU.product_model(name, n, engine).condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
        zf = model.flatten(z0)
        o1, g1 = model.potential_and_grad_flat(zf)
        o2, g2 = model.potential_and_grad_flat(zf)
        assert np.array_equal(o1, o2) and np.array_equal(g1, g2), "same launch configuration must be bit-reproducible"
        res[engine] = (o1, g1)
    oa, ga = res["ffma"]
    ob, gb = res["tcgen05"]
    assert abs(oa[0] - ob[0]) <= 5e-6 * abs(oa[0]), (oa[0], ob[0])
    np.testing.assert_allclose(oa[1:], ob[1:], rtol=1e-5, atol=1e-5 * np.abs(oa[1:]).max())
    assert np.linalg.norm(ga - gb) <= 1e-3 * np.linalg.norm(ga)


def test_shard_sum_invariance():
    """Two shards with csi_set_shard reproduce the single-handle result after a SUM of `out`."""
    name, n = "hier_diag", 3001
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    full = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    o_full, g_full = full.potential_and_grad_flat(full.flatten(z0))
    cut = 1400
    tot = np.zeros(6)
    pieces = {k: [] for k in spec.star_sites()}
    for r, sl in enumerate([slice(0, cut), slice(cut, n)]):
        m = PU.product_model(name, n, "tcgen05").set_shard(r, 2)
        m.condition(obs[sl], diag[sl], M_H_loc=ml[sl], M_H_scale=ms[sl])
        zl = {k: (v if np.ndim(v) == 0 else np.asarray(v)[sl]) for k, v in z0.items()}
        o, g = m.potential_and_grad_flat(m.flatten(zl))
        tot += o
        for k, v in m.unflatten(g).items():
            if k in pieces:
                pieces[k].append(v)
    np.testing.assert_allclose(tot, o_full, rtol=1e-9, atol=1e-9 * np.abs(o_full).max())
    gf = full.unflatten(g_full)
    for k in spec.star_sites():
        np.testing.assert_array_equal(np.concatenate(pieces[k]), gf[k])


def test_out_of_grid_is_nan_not_error():
    """Outside the BC table the reference's interpolator returns NaN: U becomes NaN (a divergence), no error."""
    name, n = "multi_phot_grid", 200
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    pts, vals = PU.grid()
    pts2 = [p.copy() for p in pts]
    pts2[0] = pts2[0] * 0.02  # Teff axis now ends at 1000 K, below every star
    import scalable_stellar_inference_b200 as ssi
    model = ssi.MultiStarModel(n, ["mag", "plx"], bands=PU.BANDS, bc="grid", bc_grid=(pts2, vals), engine="tcgen05")
    model.condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    u, _ = model.potential_and_grad(z0)
    assert np.isnan(u)


@pytest.mark.parametrize("n", [100_000, 1_000_000])
def test_full_size_properties(n):
    """BASELINE sizes (configs 4/5): finite, deterministic, and additive over star blocks.  The oracle is only
    run on a 2000-star slice of the same problem."""
    import scalable_stellar_inference_b200 as ssi
    rng = np.random.default_rng(3)
    name = "hier_diag_seis"
    m0 = 2000
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, m0)
    reps = n // m0
    tile = lambda a: np.tile(np.asarray(a), (reps,) + (1,) * (np.ndim(a) - 1))  # noqa: E731
    zb = {k: (v if np.ndim(v) == 0 else tile(v)) for k, v in z0.items()}
    model = ssi.HierarchicalStarModel(n, PU.CASES[name]["obs"]).condition(tile(obs), tile(diag), M_H_loc=tile(ml), M_H_scale=tile(ms))
    zf = model.flatten(zb)
    o1, g1 = model.potential_and_grad_flat(zf)
    o2, g2 = model.potential_and_grad_flat(zf)
    assert np.all(np.isfinite(o1)) and np.all(np.isfinite(g1))
    assert np.array_equal(o1, o2) and np.array_equal(g1, g2)
    # additivity: the tiled problem is `reps` copies of the slice -> star terms scale linearly
    u_ref, g_ref, terms = M.potential_and_grad(spec, z0, obs, diag, ml, ms, return_terms=True)
    u_expect = -(terms["hyper_logp"] + reps * float(np.sum(terms["per_star"])))
    assert abs(o1[0] - u_expect) <= 1e-5 * abs(u_expect)
    gd = model.unflatten(g1)
    for k in spec.star_sites():
        blocks = gd[k].reshape(reps, m0)
        assert np.array_equal(blocks[0], blocks[-1]) and np.array_equal(blocks[0], blocks[reps // 2])
        l2 = np.linalg.norm(blocks[0] - g_ref[k]) / np.linalg.norm(g_ref[k])
        assert l2 <= 2e-3, (k, l2)


@pytest.mark.gpu
def test_pipelined_host_blocks_reproduce_single_handle():
    """parallel.PipelinedHostPotential (the end-to-end host-buffer API: blocks on streams) == one handle."""
    import scalable_stellar_inference_b200 as ssi
    from scalable_stellar_inference_b200.parallel import PipelinedHostPotential, shard_problem
    name, n, nb = "hier_phot_grid", 1000, 3
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    full = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    out, grad = full.potential_and_grad_flat(full.flatten(z0))
    g_full = full.unflatten(grad)
    blocks, zb = [], []
    for b in range(nb):
        m = PU.product_model(name, n, "tcgen05")
        m.set_shard(b, nb)
        zl, o, d, a, s = shard_problem(z0, obs, diag, ml, ms, b, nb)
        m.condition(o, d, M_H_loc=a, M_H_scale=s)
        blocks.append(m); zb.append(m.flatten(zl))
    pipe = PipelinedHostPotential(blocks)
    pipe.load_blocks(zb)
    for _ in range(2):
        u, gh, gviews = pipe.value_and_grad_host()
    assert abs(u - out[0]) <= 1e-9 * abs(out[0])
    assert np.allclose(gh, out[1:6], rtol=1e-9, atol=1e-9 * np.abs(out[1:6]).max())
    for site in spec.star_sites():
        cat = np.concatenate([gv[site] for gv in gviews])
        assert np.array_equal(cat, g_full[site]), site


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["tcgen05", "ffma"])
def test_zero_copy_host_potential_is_bit_identical_to_device_buffers(engine):
    """parallel.ZeroCopyHostPotential: the kernel reading z from / writing the gradient to pinned host memory gives
    exactly what it gives with device buffers (several evaluations: the staging column is reused across tiles)."""
    from scalable_stellar_inference_b200.parallel import ZeroCopyHostPotential
    name, n = "hier_phot_grid", 40_000 if engine == "tcgen05" else 3000
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, 1000)
    rep = n // 1000
    tile = lambda a: np.tile(np.asarray(a), (rep,) + (1,) * (np.ndim(a) - 1))  # noqa: E731
    zb = {k: (v if np.ndim(v) == 0 else tile(v)) for k, v in z0.items()}
    model = PU.product_model(name, n, engine).condition(tile(obs), tile(diag), M_H_loc=tile(ml), M_H_scale=tile(ms))
    zf = model.flatten(zb)
    out, grad = model.potential_and_grad_flat(zf)
    zc = ZeroCopyHostPotential(model)
    zc.load(zf)
    for _ in range(3):
        zc.grad_pin.zero_()
        u, gh, gview = zc.value_and_grad_host()
        assert u == out[0] and np.array_equal(gh, out[1:6])
        assert np.array_equal(zc.grad_pin.numpy()[5:], np.asarray(grad)[5:])
    assert zc.h2d_bytes == zf.size * 4 and zc.launches == 3


@pytest.mark.gpu
@pytest.mark.parametrize("gate", ["1", "2", "3", "4"])
def test_gated_host_sites_copy_is_bit_identical_to_device_buffers(gate, monkeypatch):
    """csi_potential_host on page-locked buffers with CSI_HOST_GATE = 1 / 2 (opt-in; launches of >= 8 tile rounds): the copy engine brings the sites in
    star blocks while the kernel runs and the kernel stages a tile only once its block has been published
    (Launch::z_ready).  A different z every call, so a tile staged before its block landed would read the previous
    call's sites; potential, hyper-gradient and per-star gradient must equal the device-buffer evaluation bit for bit."""
    from scalable_stellar_inference_b200.parallel import ZeroCopyHostPotential
    monkeypatch.setenv("CSI_HOST_GATE", gate)
    monkeypatch.setenv("CSI_HOST_GATE_DIRECT", "3")  # mode 3: the first 3 rounds straight from the caller's buffer
    name, n = "hier_phot_grid", 230_000
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, 1000)
    rep = n // 1000
    tile = lambda a: np.tile(np.asarray(a), (rep,) + (1,) * (np.ndim(a) - 1))  # noqa: E731
    zb = {k: (v if np.ndim(v) == 0 else tile(v)) for k, v in z0.items()}
    model = PU.product_model(name, n, "tcgen05").condition(tile(obs), tile(diag), M_H_loc=tile(ml), M_H_scale=tile(ms))
    assert n >= 8 * 128 * model.info()["sm_count"], "the launch must be long enough for the gated path"
    zf = np.asarray(model.flatten(zb), np.float32)
    zc = ZeroCopyHostPotential(model)
    wiggle = np.sin(np.arange(zf.size - 5, dtype=np.float64)).astype(np.float32)
    for k in range(4):
        zk = zf.copy()
        zk[5:] += np.float32(2e-3 * k) * wiggle
        out, grad = model.potential_and_grad_flat(zk)
        zc.load(zk)
        zc.grad_pin.zero_()
        u, gh, gview = zc.value_and_grad_host()
        assert np.isfinite(u) and u == out[0] and np.array_equal(gh, out[1:6])
        assert np.array_equal(zc.grad_pin.numpy()[5:], np.asarray(grad)[5:], equal_nan=True)
