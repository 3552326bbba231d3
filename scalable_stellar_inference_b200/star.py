"""Deterministic star model (params -> observables), CUDA backend for the emulator / BC evaluations.

Mirrors ``src/celestify/star.py`` (constants ``:10-17``, factory ``:33-39``) and ``src/celestify/_jax/star.py:7-66``.
``bc="grid"`` feeds ``[Teff, log_g, M_H, Av]`` to the table (``:40-44``, what HEAD does); ``bc="net"`` feeds
``[log_Teff, log_g, M_H, Av]`` to the BC emulator (``:48-52``, what the photometric notebook used).  Inputs may be
scalars (one star, as in the reference) or arrays (a batch, what ``vmap(star)`` does in the notebooks).
"""
from __future__ import annotations

from math import log10
from typing import Optional

import numpy as np

from .bolometric_corrections import BolometricCorrections
from .emulator import Emulator

grav_constant = 6.6743e-8  # in cgs units


class StarBase:
    log_m_sun = log10(1.988410) + 33
    log_r_sun = log10(6.957) + 10
    log_g_sun = log10(grav_constant) + log_m_sun - 2 * log_r_sun
    log_l_sun = log10(3.828) + 33
    log_teff_sun = log10(5772.0)
    bol_mag_sun = 4.75
    log_zx_sun = log10(0.0181)
    log_numax_sun = log10(3090.0)

    def __init__(self, bands: Optional[list] = None, backend: str = "cuda", grid=None, bc: str = "grid"):
        self.emulator = Emulator(backend=backend)
        self.bolometric_corrections = None
        if bands is not None:
            self.bolometric_corrections = BolometricCorrections(bands, backend=backend, grid=grid)
        self.bc = bc

    def model(self, params: dict) -> dict:
        raise NotImplementedError()

    def __call__(self, params: dict):
        return self.model(params)


class CudaStar(StarBase):
    def model(self, params: dict) -> dict:
        f = np.float32
        g = lambda k: np.atleast_1d(np.asarray(params[k], f))  # noqa: E731
        scalar = np.ndim(params["log_mass"]) == 0
        d = {}
        log_mass, mh = g("log_mass"), g("M_H")
        d["mass"] = mass = f(10) ** log_mass
        x = np.stack(np.broadcast_arrays(g("evol"), mass, mh, g("Y"), g("a_MLT")), axis=-1)
        o = self.emulator(x)
        d["log_age"] = log_age = o[:, 0] - f(9)
        d["log_Teff"] = log_teff = o[:, 1]
        d["log_R"] = log_rad = o[:, 2]
        d["log_Dnu"] = o[:, 3]
        d["log_g"] = log_g = log_mass - 2 * log_rad + f(self.log_g_sun)
        d["log_L"] = log_lum = 2 * log_rad + 4 * (log_teff - f(self.log_teff_sun))
        d["Teff"] = teff = f(10) ** log_teff
        d["R"] = f(10) ** log_rad
        d["age"] = f(10) ** log_age
        d["Dnu"] = f(10) ** d["log_Dnu"]
        d["log_numax"] = f(self.log_numax_sun) + log_g - f(self.log_g_sun) - f(0.5) * (log_teff - f(self.log_teff_sun))
        d["numax"] = f(10) ** d["log_numax"]
        if self.bolometric_corrections is not None:
            av = g("Av") + 0 * mh
            dist = g("distance")
            if self.bc == "grid":
                bc = self.bolometric_corrections.interpolate(np.stack([teff, log_g, mh, av], axis=-1))
            else:
                bc = self.bolometric_corrections(np.stack([log_teff, log_g, mh, av], axis=-1))
            d["bol_mag"] = bol = f(self.bol_mag_sun) - f(2.5) * log_lum
            d["abs_mag"] = abs_mag = bol[:, None] - bc
            d["mag"] = abs_mag - 5 * (1 - np.log10(dist))[:, None]
            d["plx"] = f(1e3) / dist
            d["G"] = d["mag"][:, 0]
            d["BP_RP"] = bc[:, 2] - bc[:, 1] if bc.shape[1] > 2 else np.full_like(log_teff, np.nan)
        if scalar:
            d = {k: v[0] for k, v in d.items()}
        return d


def Star(bands: Optional[list] = None, backend: str = "cuda", grid=None, bc: str = "grid"):
    """Factory with the reference's signature (src/celestify/star.py:33-39)."""
    if backend in ("cuda", "b200", "gpu"):
        return CudaStar(bands, backend=backend, grid=grid, bc=bc)
    raise ValueError(f"Backend value {backend!r} is invalid.")
