"""Bolometric corrections: 4-D table interpolation and the BC emulator net, CUDA backend.

Mirrors ``src/celestify/bolometric_corrections.py`` (base ``:8-51``, factory ``:54-60``) and
``src/celestify/_jax/bolometric_corrections.py`` (``interpolate :13-15``, ``model :17-25``).

The reference builds its table from ``isochrones.mist.bc.MISTBolometricCorrectionGrid`` (third party, downloads
data at run time -- unavailable offline), so here the table is an argument: ``grid=(points, values)`` with
``points`` four increasing axes ``[Teff, logg, [Fe/H], Av]`` and ``values[nT, ng, nF, nA, len(bands)]``, the same
arrays ``_load_grid`` (``:28-45``) produces.  Without ``grid`` only the BC emulator (``model``) is available.
"""
from __future__ import annotations

import os

import numpy as np

from .emulator import DATA_DIR, DeviceOps, load_bc_weights, load_emulator_weights


class BolometricCorrectionsBase:
    _filename = os.path.join(DATA_DIR, "bc_weights.npz")

    def __init__(self, bands: list, grid=None):
        self.bands = list(bands)
        self.points, self.values = (None, None) if grid is None else self._check_grid(grid)
        w = load_bc_weights(self._filename)
        self._w = w
        self.weights, self.bias = w.weights, w.bias
        self.minval, self.maxval = w.in_loc, w.in_loc + w.in_scale
        self.loc, self.scale = w.out_loc, w.out_scale

    def _check_grid(self, grid):
        points, values = grid
        points = [np.asarray(p, np.float64) for p in points]
        values = np.asarray(values, np.float64)
        if len(points) != 4 or values.ndim != 5 or values.shape[:4] != tuple(len(p) for p in points):
            raise ValueError("grid must be (points[4], values[nT, ng, nF, nA, nbands])")
        if values.shape[-1] != len(self.bands):
            raise ValueError("values.shape[-1] must equal len(bands)")
        if values.shape[-1] > 4:
            raise ValueError("at most 4 bands are supported")
        return points, values

    def model(self, x):
        raise NotImplementedError()

    def __call__(self, x):
        return self.model(x)


class CudaBolometricCorrections(BolometricCorrectionsBase):
    def __init__(self, bands: list, grid=None, device=0):
        super().__init__(bands, grid)
        g = None if self.points is None else (self.points, self.values)
        self._ops = DeviceOps(load_emulator_weights(), bc_net=self._w, grid=g, bands=len(self.bands), device=device)

    def interpolate(self, x):
        """Multilinear table interpolation at x[..., 4] = [Teff, log_g, M_H, Av]; NaN outside the grid."""
        if self.points is None:
            raise RuntimeError("no BC table: the MIST grid needs `isochrones` (unavailable offline); pass grid=")
        x = np.asarray(x, np.float32)
        y = self._ops.grid_forward(x.reshape(-1, 4))
        return y.reshape(x.shape[:-1] + (len(self.bands),))

    def interpolate_vjp(self, x, g_out):
        x = np.asarray(x, np.float32)
        g = self._ops.grid_vjp(x.reshape(-1, 4), np.asarray(g_out, np.float32).reshape(-1, len(self.bands)))
        return g.reshape(x.shape)

    def model(self, x):
        """BC emulator at x[..., 4] = [log_Teff, log_g, M_H, Av] -> [G, BP, RP] corrections."""
        x = np.atleast_2d(np.asarray(x, np.float32))
        return self._ops.net_forward(1, x.reshape(-1, 4), 4, 3).reshape(x.shape[:-1] + (3,))

    def model_vjp(self, x, g_out):
        x = np.atleast_2d(np.asarray(x, np.float32))
        return self._ops.net_vjp(1, x.reshape(-1, 4), np.asarray(g_out, np.float32).reshape(-1, 3), 4, 3).reshape(x.shape)


def BolometricCorrections(bands: list, backend: str = "cuda", grid=None, device=0):
    """Factory with the reference's signature (src/celestify/bolometric_corrections.py:54-60)."""
    if backend.lower() in ("cuda", "b200", "gpu"):
        return CudaBolometricCorrections(bands, grid=grid, device=device)
    raise ValueError(f"Backend value {backend!r} is invalid.")
