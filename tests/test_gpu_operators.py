"""Operator-level parity (Emulator / BolometricCorrections / Star plug points) on the GPU.  -m gpu."""
import numpy as np
import pytest

import parity_util as PU
from oracle.bc import GridInterpolator
from oracle.nets import load_bc_net, load_emulator
from oracle.star import star_model

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ssi():
    import scalable_stellar_inference_b200 as m
    return m


def _rand_inputs(n, seed=0):
    rng = np.random.default_rng(seed)
    return np.stack([rng.uniform(0.01, 1.0, n), rng.uniform(0.7, 2.3, n), rng.uniform(-0.9, 0.5, n),
                     rng.uniform(0.24, 0.32, n), rng.uniform(1.5, 2.5, n)], -1).astype(np.float32)


def test_emulator_backend_names(ssi):
    # tests/test_emulator.py:9-12 of the reference: the factory accepts backend spellings / rejects unknown ones
    for b in ["cuda", "CUDA", "gpu"]:
        assert ssi.Emulator(b) is not None
    with pytest.raises(ValueError):
        ssi.Emulator("numpy")


def test_emulator_kat5_zero(ssi, kat):
    # tests/test_emulator.py:14-18 (call at 5*[0.0]) + the fp32 JAX-path answer
    out = ssi.Emulator("cuda")(np.array(5 * [0.0]))
    np.testing.assert_allclose(out, kat["kat5_zero"]["outputs"], rtol=5e-7)


def test_emulator_forward_vjp_vs_oracle(ssi):
    emu = ssi.Emulator("cuda")
    x = _rand_inputs(4099, 1)
    ref = load_emulator(np.float64)
    y_ref, masks = ref.forward(x.astype(np.float64), cache=True)
    y = emu(x)
    assert y.shape == (4099, 4)
    np.testing.assert_allclose(y, y_ref, rtol=0, atol=5e-6)
    g = np.random.default_rng(2).standard_normal((4099, 4)).astype(np.float32)
    gx_ref = ref.vjp(masks, g.astype(np.float64))
    gx = emu.vjp(x, g)
    l2 = np.linalg.norm(gx - gx_ref) / np.linalg.norm(gx_ref)
    assert l2 < 2e-3, l2   # dominated by the handful of rows sitting on a ReLU kink
    r = np.abs(gx - gx_ref) / (np.abs(gx_ref) + np.sqrt(np.mean(gx_ref ** 2)))
    assert np.quantile(r, 0.99) < 1e-4
    assert emu(np.zeros((0, 5), np.float32)).shape == (0, 4)   # empty batch


def test_bc_net_forward_vjp(ssi):
    bc = ssi.BolometricCorrections(["G", "BP", "RP"], "cuda")
    rng = np.random.default_rng(4)
    x = np.stack([rng.uniform(3.5, 4.0, 999), rng.uniform(2.5, 5.0, 999), rng.uniform(-0.9, 0.5, 999),
                  rng.uniform(0, 1, 999)], -1).astype(np.float32)
    ref = load_bc_net(np.float64)
    y_ref, masks = ref.forward(x.astype(np.float64), cache=True)
    np.testing.assert_allclose(bc(x), y_ref, rtol=0, atol=5e-6)
    # tests/test_bolometric_corrections.py:9-14: call at [5772, 4.44, 0, 0] (smoke)
    assert np.all(np.isfinite(bc([5772, 4.44, 0.0, 0.0])))
    g = rng.standard_normal((999, 3)).astype(np.float32)
    gx_ref = ref.vjp(masks, g.astype(np.float64))
    gx = bc.model_vjp(x, g)
    assert np.linalg.norm(gx - gx_ref) / np.linalg.norm(gx_ref) < 2e-3
    with pytest.raises(RuntimeError):
        bc.interpolate(x)   # no table without `isochrones`


def test_bc_grid_interpolate_and_vjp(ssi):
    pts, vals = PU.grid()
    bc = ssi.BolometricCorrections(["G", "BP", "RP"], "cuda", grid=(pts, vals))
    ref = GridInterpolator(pts, vals)
    rng = np.random.default_rng(5)
    n = 5000
    x = np.stack([rng.uniform(2600, 40000, n), rng.uniform(-0.9, 5.9, n), rng.uniform(-1.9, 0.7, n),
                  rng.uniform(0, 3.9, n)], -1)
    x[:5] = [[pts[0][0], pts[1][0], pts[2][0], pts[3][0]],            # lower corner (on the boundary: in bounds)
             [pts[0][-1], pts[1][-1], pts[2][-1], pts[3][-1]],        # upper corner
             [pts[0][3], pts[1][4], pts[2][5], pts[3][6]],            # exactly on nodes
             [1000.0, 4.0, 0.0, 0.0],                                 # below the Teff axis -> NaN
             [5000.0, 7.0, 0.0, 0.0]]                                 # above log g -> NaN
    x32 = x.astype(np.float32)
    y_ref, cache = ref.forward(x32.astype(np.float64), cache=True)
    y = bc.interpolate(x32)
    assert np.array_equal(np.isnan(y), np.isnan(y_ref))
    assert np.isnan(y[3]).all() and np.isnan(y[4]).all() and np.isfinite(y[:3]).all()
    ok = ~np.isnan(y_ref[:, 0])
    np.testing.assert_allclose(y[ok], y_ref[ok], rtol=0, atol=2e-5)
    g = rng.standard_normal((n, 3)).astype(np.float32)
    gx_ref = ref.vjp(cache, g.astype(np.float64))
    gx = bc.interpolate_vjp(x32, g)
    assert np.all(gx[~ok] == 0)
    scale = np.abs(gx_ref[ok]).max(0)
    assert np.all(np.abs(gx[ok] - gx_ref[ok]) <= 2e-3 * scale + 1e-7)


def test_star_kats(ssi, kat):
    # tests/test_star.py:4-24 of the reference (construct + call) with the notebook outputs as known answers
    k = kat["kat1_seis"]
    d = ssi.Star()(dict(k["inputs"]))
    for name, want in k["outputs"].items():
        assert abs(float(d[name]) - want) <= 3e-6 * max(1.0, abs(want)), name
    k = kat["kat2_phot"]
    d = ssi.Star(bands=["G", "BP", "RP"], bc="net")(dict(k["inputs"]))
    for name, want in k["outputs"].items():
        np.testing.assert_allclose(np.asarray(d[name], float), want, rtol=0, atol=4e-6 * max(1.0, np.max(np.abs(want))),
                                   err_msg=name)
    # batch == the oracle's star model on a synthetic table
    pts, vals = PU.grid()
    star = ssi.Star(bands=["G", "BP", "RP"], grid=(pts, vals), bc="grid")
    x = _rand_inputs(300, 9)
    p = dict(evol=x[:, 0], log_mass=np.log10(x[:, 1]), M_H=x[:, 2], Y=x[:, 3], a_MLT=x[:, 4],
             distance=np.full(300, 100.0), Av=0.0)
    d = star(p)
    ref = star_model(load_emulator(np.float64), p, bc=GridInterpolator(pts, vals), bc_mode="grid")
    for name in ("log_Teff", "log_g", "log_L", "log_numax", "mag", "plx"):
        both = np.isfinite(ref[name]) & np.isfinite(d[name])
        assert np.array_equal(np.isfinite(ref[name]), np.isfinite(d[name])), name
        np.testing.assert_allclose(np.asarray(d[name])[both], np.asarray(ref[name])[both], rtol=0, atol=5e-5, err_msg=name)


def test_predict_matches_oracle_means(ssi):
    name, n = "hier_phot_grid", 700
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    from oracle import model as M
    _, _, terms = M.potential_and_grad(spec, z0, obs, diag, ml, ms, return_terms=True)
    model = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    d = model.predict(z0)
    for i, nm in enumerate(model.output_names):
        np.testing.assert_allclose(d[nm], terms["mean"][:, i], rtol=0, atol=3e-5, err_msg=nm)
    np.testing.assert_allclose(d["evol"], terms["x"][:, 0], rtol=2e-6)


def test_errors_mirror_reference(ssi):
    # numpyro_models.py:58-59, 100-102, 251-252
    with pytest.raises(ValueError):
        ssi.SingleStarModel(["log_Teff"], kind="bogus")
    with pytest.raises(ValueError):
        ssi.SingleStarModel(["log_Teff"], prior="weird")
    with pytest.raises(ValueError):
        ssi.MultiStarModel(1, ["log_Teff"])
    with pytest.raises(ValueError):
        ssi.SingleStarModel(["not_an_output"])
    m = ssi.HierarchicalStarModel(10, ["log_Teff"])
    with pytest.raises(Exception):
        m.potential_and_grad_flat(np.zeros(3, np.float32))   # before condition()


def test_bc_net_on_the_tensor_cores(ssi):
    """Batches >= 2048 rows of the BC emulator (4 -> 64 x 6 -> 3) run on the tcgen05 kernel with zero-padded operand
    images (N = 64 MMAs over the first K-half): same answers as the oracle net and as the CUDA-core path (< 2048 rows),
    forward and input-VJP (src/celestify/_jax/bolometric_corrections.py:17-25)."""
    bc = ssi.BolometricCorrections(["G", "BP", "RP"], "cuda")
    rng = np.random.default_rng(8)
    n = 9001   # not a multiple of the 128-row tile
    x = np.stack([rng.uniform(3.45, 4.3, n), rng.uniform(2.2, 5.8, n), rng.uniform(-0.95, 0.45, n),
                  rng.uniform(0, 3.5, n)], -1).astype(np.float32)
    ref = load_bc_net(np.float64)
    y_ref, masks = ref.forward(x.astype(np.float64), cache=True)
    y = bc(x)
    np.testing.assert_allclose(y, y_ref, rtol=0, atol=1.5e-5)   # |y| ~ 3: 6e-6 = 25 fp32 ulps over 7 layers (measured max 6.0e-6)
    np.testing.assert_allclose(y[:1500], bc(x[:1500]), rtol=0, atol=1.5e-5)          # tcgen05 vs CUDA-core operator
    g = rng.standard_normal((n, 3)).astype(np.float32)
    gx_ref = ref.vjp(masks, g.astype(np.float64))
    gx = bc.model_vjp(x, g)
    assert np.linalg.norm(gx - gx_ref) / np.linalg.norm(gx_ref) < 2e-3
    r = np.abs(gx - gx_ref) / (np.abs(gx_ref) + np.sqrt(np.mean(gx_ref ** 2)))
    assert np.quantile(r, 0.99) < 1e-4
