"""Deterministic star model restated (oracle; tests only): src/celestify/_jax/star.py:7-66.

Inputs: evol, log_mass, M_H, Y, a_MLT (+ distance, Av when a BC source is given).  The reference
uses the raw emulator outputs here (no emulator-error offset; that is added only in
numpyro_models.predict).  ``bc_mode``: 'grid' feeds [Teff, log_g, M_H, Av] to the table
(:40-44), 'net' feeds [log_Teff, log_g, M_H, Av] to the BC emulator (:48-52)."""
from __future__ import annotations

import numpy as np
from . import model as M


def star_model(emulator, params: dict, bc=None, bc_mode="none"):
    f = emulator.dtype
    g = lambda k: np.atleast_1d(np.asarray(params[k], f))
    log_mass, mh = g("log_mass"), g("M_H")
    d = {}
    d["mass"] = mass = 10 ** log_mass
    x = np.stack([g("evol"), mass, mh, g("Y"), g("a_MLT")], axis=-1)
    o = emulator(x)
    d["log_age"] = log_age = o[:, 0] - f.type(9)
    d["log_Teff"] = log_teff = o[:, 1]
    d["log_R"] = log_rad = o[:, 2]
    d["log_Dnu"] = log_dnu = o[:, 3]
    d["log_g"] = log_g = log_mass - 2 * log_rad + f.type(M.LOG_G_SUN)
    d["log_L"] = log_lum = 2 * log_rad + 4 * (log_teff - f.type(M.LOG_TEFF_SUN))
    d["Teff"] = teff = 10 ** log_teff
    d["R"] = 10 ** log_rad
    d["age"] = 10 ** log_age
    d["Dnu"] = 10 ** log_dnu
    d["log_numax"] = f.type(M.LOG_NUMAX_SUN) + log_g - f.type(M.LOG_G_SUN) - f.type(0.5) * (log_teff - f.type(M.LOG_TEFF_SUN))
    d["numax"] = 10 ** d["log_numax"]
    if bc is None:
        return d
    av, dist = g("Av") + 0 * mh, g("distance")
    first = teff if bc_mode == "grid" else log_teff
    bcv = bc(np.stack([first, log_g, mh, av], axis=-1).astype(bc.dtype)).astype(f)
    d["bol_mag"] = bol = f.type(M.BOL_MAG_SUN) - f.type(2.5) * log_lum
    d["abs_mag"] = abs_mag = bol[:, None] - bcv
    d["mag"] = abs_mag - 5 * (1 - np.log10(dist))[:, None]
    d["plx"] = f.type(1e3) / dist
    d["G"] = d["mag"][:, 0]
    d["BP_RP"] = bcv[:, 2] - bcv[:, 1]
    return d
