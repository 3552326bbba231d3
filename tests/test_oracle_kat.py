"""Pin the oracle against the reference's own executed-notebook outputs (KAT-1..5)."""
import numpy as np
import pytest

from oracle import model as M
from oracle.nets import load_emulator, load_bc_net
from oracle.star import star_model


def test_kat5_zero_input(kat):
    k = kat["kat5_zero"]
    out = load_emulator(np.float32)(np.array(k["inputs"], np.float32))[0]
    np.testing.assert_allclose(out, k["outputs"], rtol=3e-7, atol=0)


def test_kat1_seis_fp32(kat):
    k = kat["kat1_seis"]
    p = dict(k["inputs"])
    d = star_model(load_emulator(np.float32), p)
    for name, want in k["outputs"].items():
        assert abs(float(d[name][0]) - want) <= 2e-6 * max(1.0, abs(want)), name


def test_kat2_phot_bcnet_fp32(kat):
    k = kat["kat2_phot"]
    d = star_model(load_emulator(np.float32), dict(k["inputs"]), bc=load_bc_net(np.float32), bc_mode="net")
    for name, want in k["outputs"].items():
        got = np.asarray(d[name][0], np.float64)
        np.testing.assert_allclose(got, want, rtol=0, atol=3e-6 * max(1.0, np.max(np.abs(want))), err_msg=name)


def test_kat3_track(kat):
    k = kat["kat3_track7000"]
    p = dict(k["inputs"]); p["log_mass"] = np.log10(p["mass"])
    d = star_model(load_emulator(np.float32), p)
    o = k["outputs"]
    assert abs(d["log_age"][0] + 9 - o["log_age_yr"]) < k["tol"]
    for name in ("log_Teff", "log_R", "log_Dnu", "log_g", "log_numax", "log_L"):
        assert abs(d[name][0] - o[name]) < k["tol"], name


def test_kat4_constants(kat):
    k = kat["kat4_const"]
    c = M.default_const()
    assert abs(c["delta"]["df"] - k["df"]) < 1e-8
    np.testing.assert_allclose(c["delta"]["loc"], k["loc"], rtol=1e-12)
    np.testing.assert_allclose(c["delta"]["cov"], k["cov"], rtol=2e-8)
    spec = M.ModelSpec("single", 1, ["log_Teff"])
    np.testing.assert_allclose(spec.var_ext, k["diag_A_cov_At"], rtol=5e-6)  # fixture printed to 6 digits
    assert abs(M.LOG_G_SUN - k["log_g_sun"]) < 1e-12
    assert abs(M.C_VEC[6] - k["c6"]) < 1e-12
    assert abs(c["log_mass"]["low"] - k["log_mass_low"]) < 1e-15
    assert abs(c["log_mass"]["high"] - k["log_mass_high"]) < 1e-15
