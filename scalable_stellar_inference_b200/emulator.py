"""Stellar-evolution emulator (MLP 5 -> 128x6 ReLU -> 4) with a CUDA backend.

Mirrors ``src/celestify/emulator.py``: ``EmulatorBase`` loads ``dense_7..13`` kernels/biases, the input
normalisation (``normalization/mean:0``, ``variance:0 ** 0.5``) and the output rescaling from the weight
file (``:8-21``); ``Emulator(backend)`` is the factory (``:39-51``).  The only backend here is ``"cuda"``
(the reference's numpy / pytensor / tensorflow backends are stale at HEAD and north_star forbids
multi-backend dispatch).  ``model(x)`` follows ``src/celestify/_jax/emulator.py:44-50``; ``vjp`` is the
input cotangent JAX autodiff would produce.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib

DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def _read_arrays(path):
    """Weight archive: the converted ``.npz`` shipped with the package, or the reference's HDF5 file."""
    if path.endswith(".npz"):
        with np.load(path) as z:
            return {k: z[k] for k in z.files}
    from .h5lite import H5File
    f = H5File(path)
    return {k: f[k] for k in f.keys()}


class DenseWeights:
    """Host copy of a dense ReLU net in the reference's conventions + its ctypes view (csi_net)."""

    def __init__(self, weights, bias, in_loc, in_scale, out_loc, out_scale):
        self.weights = [np.ascontiguousarray(w, np.float32) for w in weights]
        self.bias = [np.ascontiguousarray(b, np.float32) for b in bias]
        self.in_loc = np.ascontiguousarray(in_loc, np.float64)
        self.in_scale = np.ascontiguousarray(in_scale, np.float64)
        self.out_loc = np.ascontiguousarray(out_loc, np.float64)
        self.out_scale = np.ascontiguousarray(out_scale, np.float64)
        self.n_in = self.weights[0].shape[0]
        self.n_hidden = self.weights[0].shape[1]
        self.n_out = self.weights[-1].shape[1]

    def as_struct(self) -> _lib.Net:
        s = _lib.Net()
        s.n_layers, s.n_in, s.n_hidden, s.n_out = len(self.weights), self.n_in, self.n_hidden, self.n_out
        fp, dp = C.POINTER(C.c_float), C.POINTER(C.c_double)
        for i, (w, b) in enumerate(zip(self.weights, self.bias)):
            s.weights[i] = w.ctypes.data_as(fp)
            s.bias[i] = b.ctypes.data_as(fp)
        s.in_loc, s.in_scale = self.in_loc.ctypes.data_as(dp), self.in_scale.ctypes.data_as(dp)
        s.out_loc, s.out_scale = self.out_loc.ctypes.data_as(dp), self.out_scale.ctypes.data_as(dp)
        return s


def load_emulator_weights(path=None) -> DenseWeights:
    """src/celestify/emulator.py:8-21."""
    a = _read_arrays(path or os.path.join(DATA_DIR, "emulator_weights.npz"))
    keys = [f"dense_{i}" for i in range(7, 14)]
    return DenseWeights([a[f"{k}/{k}/kernel:0"] for k in keys], [a[f"{k}/{k}/bias:0"] for k in keys],
                        a["normalization/mean:0"], a["normalization/variance:0"] ** 0.5,
                        a["rescaling/offset"], a["rescaling/scale"])


def load_bc_weights(path=None) -> DenseWeights:
    """src/celestify/bolometric_corrections.py:15-26."""
    a = _read_arrays(path or os.path.join(DATA_DIR, "bc_weights.npz"))
    keys = ["dense"] + [f"dense_{i}" for i in range(1, 7)]
    lo, hi = a["input_scaling/minval"], a["input_scaling/maxval"]
    return DenseWeights([a[f"{k}/{k}/kernel:0"] for k in keys], [a[f"{k}/{k}/bias:0"] for k in keys],
                        lo, hi - lo, a["output_rescaling/loc"], a["output_rescaling/scale"])


class DeviceOps:
    """A libcsi handle used only for the standalone operators (net forward/VJP, BC table)."""

    def __init__(self, emulator: DenseWeights, bc_net: DenseWeights | None = None, grid=None, bands=0, device=0):
        from .numpyro_models import _base_config
        self._keep = [emulator, bc_net, grid]
        cfg = _base_config()
        cfg.level, cfg.prior, cfg.kind = _lib.LEVEL_MULTI, _lib.PRIOR_FLAT, _lib.KIND_DIAG
        cfg.n_obs = 1
        cfg.obs_index[0] = 1
        cfg.bands, cfg.bc_mode = 0, _lib.BC_NONE
        cfg.emulator = emulator.as_struct()
        if bc_net is not None:
            cfg.bc_net = bc_net.as_struct()
        if grid is not None:
            pts, vals = grid
            self._pts = [np.ascontiguousarray(p, np.float64) for p in pts]
            self._vals = np.ascontiguousarray(vals, np.float64)
            g = _lib.Grid()
            for d in range(4):
                g.shape[d] = len(self._pts[d])
                g.points[d] = self._pts[d].ctypes.data_as(C.POINTER(C.c_double))
            g.values = self._vals.ctypes.data_as(C.POINTER(C.c_double))
            cfg.bc_grid = g
            cfg.bands = int(self._vals.shape[-1])
            cfg.bc_mode = _lib.BC_GRID
            cfg.n_obs = 1
        cfg.device, cfg.engine = device, _lib.ENGINE_FFMA
        self.bands = int(cfg.bands) if grid is not None else int(bands)
        self.lib = _lib.load()
        h = C.c_void_p()
        _lib.check(self.lib.csi_create(C.byref(cfg), C.byref(h)))
        self.h = h

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.csi_destroy(self.h)
            self.h = None

    # -- helpers -------------------------------------------------------------------------------------
    def _run(self, fn, inputs, out_cols, n):
        lib, h = self.lib, self.h
        ptrs = []
        try:
            for a in inputs:
                p = C.c_void_p()
                _lib.check(lib.csi_dev_alloc(h, a.nbytes, C.byref(p)))
                ptrs.append(p)
                _lib.check(lib.csi_memcpy_h2d(h, p, a.ctypes.data, a.nbytes))
            out = np.empty((n, out_cols), np.float32)
            po = C.c_void_p()
            _lib.check(lib.csi_dev_alloc(h, max(out.nbytes, 16), C.byref(po)))
            ptrs.append(po)
            _lib.check(fn(*ptrs))
            _lib.check(lib.csi_synchronize(h))
            _lib.check(lib.csi_memcpy_d2h(h, out.ctypes.data, po, out.nbytes))
            return out
        finally:
            for p in ptrs:
                lib.csi_dev_free(h, p)

    def net_forward(self, which, x, n_in, n_out):
        x = np.ascontiguousarray(np.atleast_2d(np.asarray(x, np.float32)))
        assert x.shape[1] == n_in, f"expected {n_in} inputs, got {x.shape[1]}"
        n = x.shape[0]
        return self._run(lambda px, py: self.lib.csi_net_forward(self.h, which, px, py, n, None), [x], n_out, n)

    def net_vjp(self, which, x, gy, n_in, n_out):
        x = np.ascontiguousarray(np.atleast_2d(np.asarray(x, np.float32)))
        gy = np.ascontiguousarray(np.atleast_2d(np.asarray(gy, np.float32)))
        assert x.shape[1] == n_in and gy.shape == (x.shape[0], n_out)
        n = x.shape[0]
        return self._run(lambda px, pg, po: self.lib.csi_net_vjp(self.h, which, px, pg, po, n, None), [x, gy], n_in, n)

    def grid_forward(self, x):
        x = np.ascontiguousarray(np.atleast_2d(np.asarray(x, np.float32)))
        n = x.shape[0]
        return self._run(lambda px, py: self.lib.csi_grid_forward(self.h, px, py, n, None), [x], self.bands, n)

    def grid_vjp(self, x, gy):
        x = np.ascontiguousarray(np.atleast_2d(np.asarray(x, np.float32)))
        gy = np.ascontiguousarray(np.atleast_2d(np.asarray(gy, np.float32)))
        n = x.shape[0]
        return self._run(lambda px, pg, po: self.lib.csi_grid_vjp(self.h, px, pg, po, n, None), [x, gy], 4, n)


class EmulatorBase:
    """Attribute-compatible with the reference's EmulatorBase (src/celestify/emulator.py:5-36)."""
    PATH = os.path.join(DATA_DIR, "emulator_weights.npz")

    def __init__(self):
        w = load_emulator_weights(self.PATH)
        self._w = w
        self.weights, self.bias = w.weights, w.bias
        self.input_loc, self.input_scale = w.in_loc.astype(np.float32), w.in_scale.astype(np.float32)
        self.output_loc, self.output_scale = w.out_loc, w.out_scale

    def model(self, x):
        raise NotImplementedError()

    def __call__(self, x):
        return self.model(x)


class CudaEmulator(EmulatorBase):
    def __init__(self, device=0):
        super().__init__()
        self._ops = DeviceOps(self._w, device=device)

    def model(self, x):
        """Emulator model; x[..., 5] -> y[..., 4] (src/celestify/_jax/emulator.py:35-50)."""
        x = np.asarray(x, np.float32)
        y = self._ops.net_forward(0, x.reshape(-1, 5), 5, 4)
        return y.reshape(x.shape[:-1] + (4,))

    def vjp(self, x, g_out):
        """Cotangent of the inputs for output cotangent ``g_out`` (what jax.vjp(model, x)[1](g_out) gives)."""
        x = np.asarray(x, np.float32)
        g = self._ops.net_vjp(0, x.reshape(-1, 5), np.asarray(g_out, np.float32).reshape(-1, 4), 5, 4)
        return g.reshape(x.shape)


def Emulator(backend="cuda", device=0):
    """Factory with the reference's signature (src/celestify/emulator.py:39-51); only ``"cuda"`` exists here."""
    if backend.lower() in ("cuda", "b200", "gpu"):
        return CudaEmulator(device=device)
    raise ValueError(f"Backend value {backend!r} is invalid.")
