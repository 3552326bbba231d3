"""Bolometric-correction table interpolation restated (oracle; tests only).

The reference builds ``jax.scipy.interpolate.RegularGridInterpolator(points, values)`` over the MIST
grid ``values[nTeff, nlogg, nFeH, nAv, nbands]`` (``src/celestify/_jax/bolometric_corrections.py:11-15``,
grid assembly ``src/celestify/bolometric_corrections.py:28-45``) and calls it with
``x = [Teff, log_g, M_H, Av]`` (``src/celestify/_jax/star.py:40-44``).  Third-party algorithm (jax,
unpinned), restated from its published source: per axis ``i = clip(searchsorted(g, x, 'left') - 1,
0, n-2)``, ``t = (x - g[i]) / (g[i+1] - g[i])``, value = sum over the 2^d corners of
``values[corner] * prod(t or 1-t)``, NaN where any coordinate is outside ``[g[0], g[-1]]``.
The MIST table itself (third-party ``isochrones``, downloaded at run time) is unavailable:
PARITY UNPINNED for table *data*; ``synthetic_grid`` makes a smooth stand-in of the same rank.
"""
from __future__ import annotations

from itertools import product
import numpy as np


class GridInterpolator:
    def __init__(self, points, values, dtype=np.float64):
        self.dtype = np.dtype(dtype)
        self.points = [np.asarray(p, dtype) for p in points]
        self.values = np.asarray(values, dtype)
        assert self.values.ndim == len(self.points) + 1
        for p, n in zip(self.points, self.values.shape):
            assert p.ndim == 1 and p.shape[0] == n and n >= 2 and np.all(np.diff(p) > 0)

    def astype(self, dtype):
        return GridInterpolator(self.points, self.values, dtype)

    def _locate(self, x):
        idx, t, dt, oob = [], [], [], np.zeros(x.shape[0], bool)
        for d, g in enumerate(self.points):
            xd = x[:, d]
            i = np.searchsorted(g, xd, side="left") - 1
            i = np.clip(i, 0, g.size - 2)
            w = g[i + 1] - g[i]
            idx.append(i)
            t.append((xd - g[i]) / w)
            dt.append(1.0 / w)
            oob |= (xd < g[0]) | (xd > g[-1])
        return idx, t, dt, oob

    def forward(self, x, cache=False):
        x = np.atleast_2d(np.asarray(x, self.dtype))
        idx, t, dt, oob = self._locate(x)
        out = np.zeros((x.shape[0], self.values.shape[-1]), self.dtype)
        for corner in product(*[(0, 1)] * len(idx)):
            w = np.ones(x.shape[0], self.dtype)
            for c, ti in zip(corner, t):
                w = w * (ti if c else 1 - ti)
            out += self.values[tuple(i + c for i, c in zip(idx, corner))] * w[:, None]
        out = np.where(oob[:, None], np.nan, out)
        return (out, (idx, t, dt, oob)) if cache else out

    __call__ = forward

    def vjp(self, cache, g_out):
        """d(sum(g_out * value))/dx, shape [N, ndim]; zero where out of bounds (jnp.where semantics)."""
        idx, t, dt, oob = cache
        nd = len(idx)
        g_out = np.asarray(g_out, self.dtype)
        gx = np.zeros((g_out.shape[0], nd), self.dtype)
        for corner in product(*[(0, 1)] * nd):
            v = np.sum(self.values[tuple(i + c for i, c in zip(idx, corner))] * g_out, axis=1)
            for d in range(nd):
                w = np.ones_like(v)
                for e, (c, ti) in enumerate(zip(corner, t)):
                    if e == d:
                        w = w * (dt[d] if c else -dt[d])
                    else:
                        w = w * (ti if c else 1 - ti)
                gx[:, d] += v * w
        return np.where(oob[:, None], 0.0, gx)


def synthetic_grid(bands=3, seed=3, shape=(70, 26, 18, 13), dtype=np.float64):
    """Smooth synthetic BC table on a non-uniform grid of MIST-like extent.

    Axes: Teff [2500, 50000] K (log-spaced with jitter), log g [-1, 6], [Fe/H] [-2, 0.75], Av [0, 4].
    Values are a smooth blackbody-ish colour law plus low-order cross terms; seeded, deterministic."""
    rng = np.random.default_rng(seed)

    def axis(lo, hi, n, log=False):
        u = np.linspace(0.0, 1.0, n)
        u[1:-1] += rng.uniform(-0.3, 0.3, n - 2) / (n - 1)
        g = lo * (hi / lo) ** u if log else lo + (hi - lo) * u
        return np.asarray(g, np.float64)

    pts = [axis(2500.0, 50000.0, shape[0], log=True), axis(-1.0, 6.0, shape[1]),
           axis(-2.0, 0.75, shape[2]), axis(0.0, 4.0, shape[3])]
    T, G, F, A = np.meshgrid(*pts, indexing="ij")
    lt = np.log10(T) - 3.76
    vals = np.empty(shape + (bands,), np.float64)
    for b in range(bands):
        c = rng.normal(size=8)
        vals[..., b] = (0.1 * c[0] - (2.0 + 0.5 * c[1]) * lt ** 2 + (0.8 * b - 0.6 + 0.1 * c[2]) * lt
                        + 0.02 * c[3] * (G - 4.4) + 0.05 * c[4] * F + 0.03 * c[5] * F * lt
                        - (0.8 + 0.1 * b + 0.02 * c[6]) * A + 0.01 * c[7] * A * lt)
    return [p.astype(dtype) for p in pts], vals.astype(dtype)
