"""celestify's probabilistic models with the NUTS potential evaluated by libcsi (hand-written sm_100a CUDA).

Mirrors ``src/celestify/numpyro_models.py``: ``SingleStarModel(observables, const, kind, prior)`` (``:54``),
``MultiStarModel(num_stars, observables, const, kind)`` (``:249``), ``HierarchicalStarModel(num_stars, ...)``
(``:263``), the ``const`` dict with ``setdefault`` semantics (``:117-134``) and the attributes ``A, B, c,
covariance, variance, observables, observable_indices`` (``:94-103``).

numpyro itself (third party) is not part of the hot path any more: where the reference hands the model to
``NUTS(model)`` and numpyro builds ``potential_fn`` by tracing it, these classes expose that function directly:

    model.condition(obs, diag, M_H_loc=..., M_H_scale=...)   # the obs= / diag= arguments of model.__call__
    U = model.potential_fn(z)                                 # z: dict site -> unconstrained array (numpyro's)
    U, grad = model.potential_and_grad(z)                     # jax.value_and_grad(potential_fn)(z)

so ``numpyro.infer.NUTS(potential_fn=model.potential_fn)`` drives it when numpyro/jax are installed (through
the jax.ffi adapter described in INTEGRATION.md), and ``scalable_stellar_inference_b200.nuts.NUTS`` drives it
natively on the device otherwise.  Site names, order and transforms are the reference's.

Photometric models (``bands=``, observables ``"mag"`` / ``"plx"``) reconstruct the model of
``notebook/single-star-phot.ipynb`` (absent from the tree at HEAD; forward math from ``_jax/star.py:39-65``).
"""
from __future__ import annotations

import ctypes as C
import json
import os
from math import inf, log10
from typing import Optional

import numpy as np

from . import _lib
from .emulator import DATA_DIR, load_bc_weights, load_emulator_weights

grav_constant = 6.6743e-8  # in cgs units


def _emulator_precision() -> dict:
    """src/celestify/numpyro_models.py:107-115."""
    with open(os.path.join(DATA_DIR, "emulator_error.json"), "r") as file:
        params = json.loads(file.read())
    scale_tril = np.sqrt(np.array(params["theta"]))[:, None] * np.array(params["L_omega"])
    return dict(df=float(params["df"]), loc=np.array(params["mu"]), cov=scale_tril @ scale_tril.T)


def _default_const(const: Optional[dict] = None) -> dict:
    """src/celestify/numpyro_models.py:117-134 (+ the reconstructed photometric constants)."""
    const = {} if const is None else const
    const.setdefault("delta", _emulator_precision())
    const.setdefault("log_evol", dict(loc=-0.7, scale=0.4, high=0.0))
    const.setdefault("log_mass", dict(loc=0.0, scale=0.3, low=log10(0.7), high=log10(2.3)))
    const.setdefault("M_H", dict(loc=0.0, scale=0.5, low=-0.9, high=0.5))
    const.setdefault("Y", dict(low=0.24, high=0.32))
    const.setdefault("a_MLT", dict(low=1.5, high=2.5))
    const.setdefault("Y_0", dict(loc=0.247, scale=0.001))
    const.setdefault("dY_dZ", dict(loc=1.5, scale=1.0, low=0.0, high=3.0))
    const.setdefault("precision_Y", dict(concentration=3.0, rate=1e-4))
    const.setdefault("log_sigma_Y", dict(loc=-2.3, scale=0.2))
    const.setdefault("mu_a", dict(loc=2.0, scale=0.2, low=1.5, high=2.5))
    const.setdefault("precision_a", dict(concentration=3.0, rate=1e-2))
    const.setdefault("log_sigma_a", dict(loc=-1.3, scale=0.2))
    const.setdefault("scaled_distance", dict(concentration1=3.0, concentration0=1.0))
    const.setdefault("max_distance", 500.0)
    const.setdefault("Av", 0.0)
    return const


def _tn(d, name="") -> _lib.Tn:
    """Truncated-normal constants of one site.  The reference's ``TruncatedNormal(**const[...])`` also accepts per-star
    ARRAYS for loc / scale (scripts/model_star.py:126-128 does that for M_H); the C ABI carries per-star arrays for
    M_H only (``condition(M_H_loc=, M_H_scale=)``), so arrays anywhere else are an error, not silently averaged."""
    for k in ("loc", "scale"):
        if np.ndim(d[k]) != 0 and np.size(d[k]) != 1:
            raise ValueError(f"const[{name!r}][{k!r}] is an array: per-star prior parameters are supported for M_H only, "
                             "through condition(M_H_loc=..., M_H_scale=...)")
    return _lib.Tn(float(np.reshape(d["loc"], -1)[0]), float(np.reshape(d["scale"], -1)[0]), float(d.get("low", -inf)),
                   float(d.get("high", inf)))


def _base_config(const: Optional[dict] = None) -> _lib.Config:
    c = _default_const(dict(const) if const else None)
    cfg = _lib.Config()
    cfg.log_evol, cfg.log_mass = _tn(c["log_evol"], "log_evol"), _tn(c["log_mass"], "log_mass")
    mh = dict(c["M_H"])
    if np.size(mh["loc"]) > 1 or np.size(mh["scale"]) > 1:   # per-star arrays travel through condition() (see there)
        mh["loc"], mh["scale"] = float(np.mean(mh["loc"])), float(np.mean(mh["scale"]))
    cfg.m_h = _tn(mh, "M_H")
    cfg.y_low, cfg.y_high = c["Y"]["low"], c["Y"]["high"]
    cfg.a_low, cfg.a_high = c["a_MLT"]["low"], c["a_MLT"]["high"]
    cfg.y0_loc, cfg.y0_scale = c["Y_0"]["loc"], c["Y_0"]["scale"]
    cfg.dydz, cfg.mu_a = _tn(c["dY_dZ"], "dY_dZ"), _tn(c["mu_a"], "mu_a")
    cfg.py_conc, cfg.py_rate = c["precision_Y"]["concentration"], c["precision_Y"]["rate"]
    cfg.pa_conc, cfg.pa_rate = c["precision_a"]["concentration"], c["precision_a"]["rate"]
    cfg.delta_df = float(c["delta"]["df"])
    for i in range(4):
        cfg.delta_loc[i] = float(c["delta"]["loc"][i])
    cov = np.asarray(c["delta"]["cov"], np.float64)
    for i in range(16):
        cfg.delta_cov[i] = float(cov.flat[i])
    cfg.sd_c1 = c["scaled_distance"]["concentration1"]
    cfg.sd_c0 = c["scaled_distance"]["concentration0"]
    cfg.max_distance, cfg.av = float(c["max_distance"]), float(c["Av"])
    return cfg


class SingleStarModel:
    log_m_sun = log10(1.988410) + 33
    log_r_sun = log10(6.957) + 10
    log_g_sun = log10(grav_constant) + log_m_sun - 2 * log_r_sun
    log_l_sun = log10(3.828) + 33
    log_teff_sun = log10(5772.0)
    log_zx_sun = log10(0.0181)
    log_numax_sun = log10(3090.0)
    outputs = ["log_age", "log_Teff", "log_R", "log_Dnu", "log_L", "log_g", "log_numax"]
    _level = _lib.LEVEL_SINGLE
    num_stars = 1

    def __init__(self, observables: list, const: Optional[dict] = None, kind: str = "diag",
                 prior: str = "informative", *, bands: Optional[list] = None, bc: str = "grid", bc_grid=None,
                 engine: str = "tcgen05", device: int = 0):
        const = _default_const(const)
        if kind not in ["none", "diag", "full"]:
            raise ValueError("Kind must be one of 'none', 'diag', or 'full'.")
        if kind == "none":
            # numpyro_models.py:211-212 returns a 1-tuple that __call__ (:244) cannot unpack
            raise ValueError("Kind 'none' cannot be evaluated in the reference either; use 'diag' or 'full'.")
        prior = prior.lower()
        priors = ["flat", "informative"]
        if prior not in priors:
            raise ValueError(f"Variable 'prior' must be one of {priors}.")
        if engine not in ("tcgen05", "ffma"):
            raise ValueError("engine must be 'tcgen05' or 'ffma'")
        self.bands = list(bands) if bands else []
        if self.bands and bc not in ("grid", "net"):
            raise ValueError("bc must be 'grid' or 'net'")
        if self.bands and bc == "grid" and bc_grid is None:
            raise ValueError("bc='grid' needs bc_grid=(points, values): the MIST table is not shipped")
        nb = len(self.bands)
        A = np.vstack([np.eye(4), [0.0, 4.0, 2.0, 0.0], [0.0, 0.0, -2.0, 0.0], [0.0, -0.5, -2.0, 0.0]])
        B = np.vstack([np.zeros((5, 7)), [0.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0], [0.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0]])
        c = np.array([-9.0, 0.0, 0.0, 0.0, -4.0 * self.log_teff_sun, self.log_g_sun,
                      self.log_numax_sun + 0.5 * self.log_teff_sun])
        covariance = A @ const["delta"]["cov"] @ A.T
        names = list(self.outputs) + [f"mag_{b}" for b in self.bands] + (["plx"] if nb else [])
        idx = []
        for o in observables:
            if o == "mag":
                if not nb:
                    raise ValueError("observable 'mag' needs bands=")
                idx += [7 + b for b in range(nb)]
            else:
                idx.append(names.index(o))  # ValueError for unknown names, like list.index in the reference
        self.emulator_weights = load_emulator_weights()
        self.const, self.kind, self._prior = const, kind, prior
        self.A, self.B, self.c = A, B, c
        self.covariance, self.variance = covariance, np.diag(covariance)
        self.observables = list(observables)
        self.observable_indices = np.array(idx, int)
        self.output_names = names
        self.bc, self.engine, self.device = bc, engine, device

        cfg = _base_config(const)
        cfg.level = self._level
        cfg.prior = _lib.PRIOR_FLAT if prior == "flat" else _lib.PRIOR_INFORMATIVE
        cfg.kind = _lib.KIND_DIAG if kind == "diag" else _lib.KIND_FULL
        if len(idx) > _lib.MAX_OBS:
            raise ValueError("too many observables")
        cfg.n_obs = len(idx)
        for j, i in enumerate(idx):
            cfg.obs_index[j] = i
        cfg.bands = nb
        cfg.bc_mode = _lib.BC_NONE if not nb else (_lib.BC_GRID if bc == "grid" else _lib.BC_NET)
        self._keep = [self.emulator_weights]
        cfg.emulator = self.emulator_weights.as_struct()
        if nb and bc == "net":
            w = load_bc_weights()
            self._keep.append(w)
            cfg.bc_net = w.as_struct()
        if nb and bc == "grid":
            pts = [np.ascontiguousarray(p, np.float64) for p in bc_grid[0]]
            vals = np.ascontiguousarray(bc_grid[1], np.float64)
            if vals.shape != tuple(len(p) for p in pts) + (nb,):
                raise ValueError("bc_grid values must have shape [nT, ng, nF, nA, len(bands)]")
            self._keep += [pts, vals]
            g = _lib.Grid()
            for d in range(4):
                g.shape[d] = len(pts[d])
                g.points[d] = pts[d].ctypes.data_as(C.POINTER(C.c_double))
            g.values = vals.ctypes.data_as(C.POINTER(C.c_double))
            cfg.bc_grid = g
        cfg.device = device
        cfg.engine = _lib.ENGINE_TCGEN05 if engine == "tcgen05" else _lib.ENGINE_FFMA
        self._lib = _lib.load()  # raises when the CUDA library is missing: no CPU fallback
        h = C.c_void_p()
        _lib.check(self._lib.csi_create(C.byref(cfg), C.byref(h)))
        self._h = h
        nh, ns = C.c_int32(), C.c_int32()
        names_buf = (C.c_char_p * (_lib.MAX_HYPER + _lib.MAX_STAR_SITES))()
        _lib.check(self._lib.csi_layout(h, C.byref(nh), C.byref(ns), names_buf))
        self.hyper_sites = [names_buf[i].decode() for i in range(nh.value)]
        self.star_sites = [names_buf[nh.value + i].decode() for i in range(ns.value)]
        self.n_local = 0
        self._rank, self._world = 0, 1

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.csi_destroy(self._h)
            self._h = None

    # ---- data -----------------------------------------------------------------------------------------
    def set_shard(self, rank: int, world: int):
        """Star-axis data parallelism: this process holds block ``rank`` of ``world`` (call before condition)."""
        _lib.check(self._lib.csi_set_shard(self._h, rank, world))
        self._rank, self._world = rank, world
        return self

    def set_peer_exchange(self, peer_ptrs):
        """Fuse the allreduce of the 6 sums into the kernel: ``peer_ptrs[r]`` = device pointer (valid in this
        process) to rank r's exchange buffer (csi_set_peer_exchange); ``None`` switches it off."""
        if not peer_ptrs:
            _lib.check(self._lib.csi_set_peer_exchange(self._h, 1, 0, None))
            self._peer_exchange_on = False
            return self
        arr = (C.c_void_p * len(peer_ptrs))(*[C.c_void_p(int(p)) for p in peer_ptrs])
        _lib.check(self._lib.csi_set_peer_exchange(self._h, len(peer_ptrs), self._rank, arr))
        self._peer_exchange_on = len(peer_ptrs) > 1
        return self

    def condition(self, obs, diag=None, M_H_loc=None, M_H_scale=None):
        """Load the observations of the LOCAL stars: ``obs`` / ``diag`` are [n, k] with k observed columns
        (``"mag"`` expands to one column per band); ``diag=None`` is 0 as in ``likelihood`` (``:229-230``)."""
        k = len(self.observable_indices)
        obs = np.ascontiguousarray(np.asarray(obs, np.float64).reshape(-1, k))
        n = obs.shape[0]
        pd = C.POINTER(C.c_double)
        dg = None
        if diag is not None:
            dg = np.ascontiguousarray(np.broadcast_to(np.asarray(diag, np.float64), obs.shape))
        # an observable without emulator-error variance (the parallax row) needs diag > 0: Normal(mean, sqrt(0 + 0))
        var_obs = np.concatenate([self.variance, np.zeros(len(self.output_names) - 7)])[self.observable_indices]
        if self.kind == "diag":
            total = var_obs[None, :] + (0.0 if dg is None else dg)
            if np.any(np.broadcast_to(total, obs.shape) <= 0.0):
                bad = [self.output_names[i] for i, v in zip(self.observable_indices, np.min(np.broadcast_to(total, obs.shape), axis=0)) if v <= 0]
                raise ValueError(f"observables {bad} have zero total variance: pass diag > 0 for them")
        ml = ms = None
        if M_H_loc is None and (np.size(self.const["M_H"]["loc"]) > 1 or np.size(self.const["M_H"]["scale"]) > 1):
            M_H_loc, M_H_scale = self.const["M_H"]["loc"], self.const["M_H"]["scale"]   # const arrays, as the reference takes them
        if M_H_loc is not None:
            ml = np.ascontiguousarray(np.broadcast_to(np.asarray(M_H_loc, np.float64), (n,)))
            ms = np.ascontiguousarray(np.broadcast_to(np.asarray(M_H_scale, np.float64), (n,)))
        p = lambda a: None if a is None else a.ctypes.data_as(pd)  # noqa: E731
        _lib.check(self._lib.csi_set_data(self._h, n, p(obs), p(dg), p(ml), p(ms)))
        self.n_local = n
        return self

    def __call__(self, obs=None, diag=None, *, M_H_loc=None, M_H_scale=None, rng_key=0):
        """The model callable of the reference (``numpyro_models.py:242-245,255-259,341-346``).  numpyro runs it
        under handlers; here it has the two effects those runs have on the hot path: with ``obs`` it CONDITIONS the
        model (what ``mcmc.run(key, obs=y, diag=diag)`` does -- afterwards ``potential_fn`` is the conditioned
        potential) and returns the model; with ``obs=None`` it is a prior-predictive trace
        (``handlers.trace(handlers.seed(model, key)).get_trace(obs=None)``, tests/test_numpyro_models.py:8-17): one
        draw of every site from the priors, the deterministic outputs through the GPU forward pass, and ``y`` sampled
        for all outputs -- returned as a dict."""
        if obs is not None:
            return self.condition(obs, diag, M_H_loc=M_H_loc, M_H_scale=M_H_scale)
        from .infer import Predictive
        return {k: (v[0] if np.ndim(v) else v) for k, v in Predictive(self, num_samples=1)(rng_key).items()}

    # ---- layout ---------------------------------------------------------------------------------------
    @property
    def site_names(self):
        return self.hyper_sites + self.star_sites

    @property
    def flat_size(self):
        return len(self.hyper_sites) + len(self.star_sites) * self.n_local

    def flatten(self, z: dict) -> np.ndarray:
        parts = [np.asarray(z[s], np.float32).reshape(1) for s in self.hyper_sites]
        parts += [np.broadcast_to(np.asarray(z[s], np.float32), (self.n_local,)) for s in self.star_sites]
        return np.ascontiguousarray(np.concatenate(parts), np.float32)

    def unflatten(self, v: np.ndarray) -> dict:
        nh, n = len(self.hyper_sites), self.n_local
        d = {s: v[i] for i, s in enumerate(self.hyper_sites)}
        for j, s in enumerate(self.star_sites):
            d[s] = v[nh + j * n: nh + (j + 1) * n]
        return d

    # ---- the hot path -----------------------------------------------------------------------------------
    def potential_and_grad_flat(self, z: np.ndarray):
        """Host buffers in, host buffers out (csi_potential_host): (U_local, dU/dz_local, hyper partials)."""
        z = np.ascontiguousarray(z, np.float32)
        if z.shape != (self.flat_size,):
            raise ValueError(f"z must have shape ({self.flat_size},)")
        out = np.zeros(6, np.float64)
        grad = np.empty_like(z)
        _lib.check(self._lib.csi_potential_host(self._h, z.ctypes.data, out.ctypes.data, grad.ctypes.data))
        return out, grad

    def potential_and_grad(self, z: dict):
        out, grad = self.potential_and_grad_flat(self.flatten(z))
        g = self.unflatten(grad)
        for i, s in enumerate(self.hyper_sites):
            g[s] = out[1 + i]
        return float(out[0]), g

    def potential_fn(self, z: dict) -> float:
        return self.potential_and_grad(z)[0]

    def potential_device(self, z_ptr: int, out_ptr: int, grad_ptr: int, stream: int = 0):
        """Device pointers (e.g. ``torch.Tensor.data_ptr()``), stream-ordered, no synchronisation."""
        _lib.check(self._lib.csi_potential(self._h, z_ptr, out_ptr, grad_ptr, stream or None))

    def potential_host(self, z_ptr: int, out_ptr: int, grad_ptr: int):
        """Host pointers, synchronous (csi_potential_host).  Page-locked z / gradient buffers (e.g. pinned torch
        tensors) are read and written by the kernel itself -- no copies; pageable ones are staged."""
        _lib.check(self._lib.csi_potential_host(self._h, z_ptr, out_ptr, grad_ptr))

    def time_potential(self, z_ptr: int, out_ptr: int, grad_ptr: int, iters: int) -> float:
        ms = C.c_float()
        _lib.check(self._lib.csi_time_potential(self._h, z_ptr, out_ptr, grad_ptr, iters, C.byref(ms)))
        return float(ms.value)

    def info(self) -> dict:
        a = (C.c_int64 * 8)()
        _lib.check(self._lib.csi_info(self._h, a))
        return dict(sm_count=a[0], engine="tcgen05" if a[1] == 0 else "ffma", launches=a[2], n_local=a[3],
                    smem_bytes=a[4], grid=a[5], rank=a[6], world=a[7])

    # ---- forward-only (Predictive) ------------------------------------------------------------------------
    def predict(self, z: dict, constrained: bool = False) -> dict:
        """Deterministic sites and the predicted means at ``z`` (what Predictive / numpyro.deterministic expose).
        ``constrained=True`` adds the constrained values of the remaining sampled sites (``log_evol``, the
        precisions, the hyper-parameters and ``sigma_Y`` / ``sigma_a``), i.e. everything ``mcmc.get_samples()`` holds."""
        lib, h, n = self._lib, self._h, self.n_local
        zf = self.flatten(z)
        ptrs = []
        try:
            def alloc(nbytes):
                p = C.c_void_p()
                _lib.check(lib.csi_dev_alloc(h, nbytes, C.byref(p)))
                ptrs.append(p)
                return p
            pz, pm, pp = alloc(zf.nbytes), alloc(n * _lib.MAX_MEAN * 4), alloc(n * 8 * 4)
            _lib.check(lib.csi_memcpy_h2d(h, pz, zf.ctypes.data, zf.nbytes))
            _lib.check(lib.csi_predict(h, pz, pm, pp, None))
            _lib.check(lib.csi_synchronize(h))
            mean = np.empty((n, _lib.MAX_MEAN), np.float32)
            par = np.empty((n, 8), np.float32)
            _lib.check(lib.csi_memcpy_d2h(h, mean.ctypes.data, pm, mean.nbytes))
            _lib.check(lib.csi_memcpy_d2h(h, par.ctypes.data, pp, par.nbytes))
        finally:
            for p in ptrs:
                lib.csi_dev_free(h, p)
        self._last_mean = mean
        d = {nm: mean[:, i] for i, nm in enumerate(self.output_names)}
        for i, nm in enumerate(["evol", "mass", "M_H", "Y", "a_MLT", "log_mass", "scaled_precision"]):
            d[nm] = par[:, i]
        if self.bands:
            d["distance"] = par[:, 7]
        if constrained:
            c = self.const
            sig = lambda u: 1.0 / (1.0 + np.exp(-np.asarray(u, np.float64)))  # noqa: E731
            d["log_evol"] = np.log10(d["evol"])
            if self.hyper_sites:
                d["Y_0"] = np.float64(z["Y_0"])
                d["dY_dZ"] = c["dY_dZ"]["low"] + (c["dY_dZ"]["high"] - c["dY_dZ"]["low"]) * sig(z["dY_dZ"])
                d["mu_a"] = c["mu_a"]["low"] + (c["mu_a"]["high"] - c["mu_a"]["low"]) * sig(z["mu_a"])
            for nm in ("precision_Y", "precision_a"):
                if nm in z:
                    d[nm] = np.exp(np.asarray(z[nm], np.float64))
                    d["sigma_" + nm[-1]] = d[nm] ** -0.5
            if self.bands:
                d["scaled_distance"] = d["distance"] / c["max_distance"]
        return d


class MultiStarModel(SingleStarModel):
    _level = _lib.LEVEL_MULTI

    def __init__(self, num_stars: int, observables: list, const: Optional[dict] = None, kind: str = "diag",
                 prior: str = "informative", **kw):
        if num_stars < 2:
            raise ValueError("Variable num_stars must be greater than 1.")
        super().__init__(observables, const=const, kind=kind, prior=prior, **kw)
        self.num_stars = num_stars


class HierarchicalStarModel(MultiStarModel):
    _level = _lib.LEVEL_HIERARCHICAL

    def __init__(self, num_stars: int, observables: list, const: Optional[dict] = None, kind: str = "diag", **kw):
        super().__init__(num_stars, observables, const=const, kind=kind, **kw)
