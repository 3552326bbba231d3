"""numpyro ``potential_fn`` (= -log joint in unconstrained space) and its gradient, restated
in numpy with hand-written reverse mode (oracle; tests only).

Follows ``src/celestify/numpyro_models.py``: constants ``:42-48,61-84,107-134``; non-hierarchical
priors ``:148-201`` (+ ``enrichment_law :136-138``, ``_enrichment_unc_fac :140-146``); hierarchical
hyper-priors ``:277-297`` and per-star priors ``:299-339``; ``predict :203-218``; ``likelihood
:220-240``; orchestration ``:242-245,255-259,341-346``.  The photometric extension follows
``src/celestify/_jax/star.py:39-65`` (the numpyro photometric model is not in the tree at HEAD --
see SURVEY 0; sites/priors reconstructed from notebook/single-star-phot.ipynb cells [5],[7]).

Known reference defects that are NOT reproduced (documented in DESIGN.md):
  * kind="full" at HEAD adds ``diag[..., None] * identity(mean.shape[0])`` (``:239``), which uses the
    star count for batched means; here each star gets its own k x k ``cov_sub/tau + diag(diag_i)``;
  * kind="none" cannot be unpacked by ``__call__`` (``:211-212,244``) and is rejected.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field
from math import log10

import numpy as np
from scipy import special as sp

from . import dists as D
from .nets import DATA_DIR, load_emulator

LN10 = D.LN10

# src/celestify/numpyro_models.py:38-48
GRAV_CONSTANT = 6.6743e-8
LOG_M_SUN = log10(1.988410) + 33
LOG_R_SUN = log10(6.957) + 10
LOG_G_SUN = log10(GRAV_CONSTANT) + LOG_M_SUN - 2 * LOG_R_SUN
LOG_TEFF_SUN = log10(5772.0)
LOG_ZX_SUN = log10(0.0181)
LOG_NUMAX_SUN = log10(3090.0)
BOL_MAG_SUN = 4.75  # src/celestify/star.py:15
OUTPUTS = ["log_age", "log_Teff", "log_R", "log_Dnu", "log_L", "log_g", "log_numax"]  # :49-52

# :61-81
A_MAT = np.vstack([np.eye(4), [0.0, 4.0, 2.0, 0.0], [0.0, 0.0, -2.0, 0.0], [0.0, -0.5, -2.0, 0.0]])
C_VEC = np.array([-9.0, 0.0, 0.0, 0.0, -4.0 * LOG_TEFF_SUN, LOG_G_SUN, LOG_NUMAX_SUN + 0.5 * LOG_TEFF_SUN])
# B @ params adds params[5] = log_mass to rows 5 (log_g) and 6 (log_numax)   (:70-76)


def emulator_precision():
    """:107-115 -- Student-t fit of the emulator error: df, loc, cov = L L^T, L = sqrt(theta)[:,None]*L_omega."""
    with open(os.path.join(DATA_DIR, "emulator_error.json")) as fh:
        p = json.load(fh)
    scale_tril = np.sqrt(np.array(p["theta"]))[:, None] * np.array(p["L_omega"])
    return dict(df=float(p["df"]), loc=np.array(p["mu"]), cov=scale_tril @ scale_tril.T)


def default_const(const=None):
    """:117-134 (setdefault semantics), plus the reconstructed photometric constants."""
    const = dict(const or {})
    const.setdefault("delta", emulator_precision())
    const.setdefault("log_evol", dict(loc=-0.7, scale=0.4, high=0.0))
    const.setdefault("log_mass", dict(loc=0.0, scale=0.3, low=log10(0.7), high=log10(2.3)))
    const.setdefault("M_H", dict(loc=0.0, scale=0.5, low=-0.9, high=0.5))
    const.setdefault("Y", dict(low=0.24, high=0.32))
    const.setdefault("a_MLT", dict(low=1.5, high=2.5))
    const.setdefault("Y_0", dict(loc=0.247, scale=0.001))
    const.setdefault("dY_dZ", dict(loc=1.5, scale=1.0, low=0.0, high=3.0))
    const.setdefault("precision_Y", dict(concentration=3.0, rate=1e-4))
    const.setdefault("mu_a", dict(loc=2.0, scale=0.2, low=1.5, high=2.5))
    const.setdefault("precision_a", dict(concentration=3.0, rate=1e-2))
    # photometric reconstruction (notebook/single-star-phot.ipynb cell [7] output)
    const.setdefault("scaled_distance", dict(concentration1=3.0, concentration0=1.0))
    const.setdefault("max_distance", 500.0)
    const.setdefault("Av", 0.0)
    return const


@dataclass
class ModelSpec:
    """level: 'single' | 'multi' | 'hierarchical'; prior: 'flat' | 'informative' (non-hierarchical only)."""
    level: str
    num_stars: int
    observables: list
    kind: str = "diag"
    prior: str = "informative"
    const: dict = field(default_factory=dict)
    bands: int = 0               # number of photometric bands (0 = no photometry)
    bc: object = None            # GridInterpolator or DenseNet (BC net) or None
    bc_mode: str = "none"        # 'none' | 'grid' | 'net'
    emulator: object = None

    def __post_init__(self):
        if self.kind not in ("diag", "full"):
            raise ValueError("Kind must be one of 'diag' or 'full' ('none' is unusable in the reference).")
        if self.prior not in ("flat", "informative"):
            raise ValueError("Variable 'prior' must be one of ['flat', 'informative'].")
        if self.level in ("multi", "hierarchical") and self.num_stars < 2:
            raise ValueError("Variable num_stars must be greater than 1.")
        self.const = default_const(self.const)
        if self.emulator is None:
            self.emulator = load_emulator(np.float64)
        self.phot = self.bands > 0
        names = list(OUTPUTS) + [f"mag_{b}" for b in range(self.bands)] + (["plx"] if self.phot else [])
        self.names = names
        idx = []
        for o in self.observables:
            if o == "mag":
                idx += [names.index(f"mag_{b}") for b in range(self.bands)]
            else:
                idx.append(names.index(o))
        self.obs_idx = np.array(idx, int)
        a_ext = [A_MAT] + [-2.5 * A_MAT[4:5]] * self.bands + ([np.zeros((1, 4))] if self.phot else [])
        self.a_ext = np.vstack(a_ext)
        self.cov_ext = self.a_ext @ self.const["delta"]["cov"] @ self.a_ext.T     # :83
        self.var_ext = np.diag(self.cov_ext).copy()                               # :84

    # ---- site bookkeeping -----------------------------------------------------------------
    def hyper_sites(self):
        return ["Y_0", "dY_dZ", "precision_Y", "mu_a", "precision_a"] if self.level == "hierarchical" else []

    def star_sites(self):
        if self.level == "hierarchical":
            s = ["scaled_precision", "log_evol", "log_mass", "M_H", "Y_decentered", "a_decentered"]
        elif self.prior == "flat":
            s = ["log_evol", "log_mass", "M_H", "Y", "a_MLT", "scaled_precision"]
        else:
            s = ["log_evol", "log_mass", "M_H", "precision_Y", "Y", "precision_a", "a_MLT", "scaled_precision"]
        return s + (["scaled_distance"] if self.phot else [])


def _bounds(c):
    return float(c.get("low", -np.inf)), float(c.get("high", np.inf))


def _tn_site(u, c, loc=None, scale=None):
    """Constant-bound TruncatedNormal site in unconstrained space.

    Returns value, dvalue/du, log-term (log_prob + log|J|), dlog-term/du."""
    loc = c["loc"] if loc is None else loc
    scale = c["scale"] if scale is None else scale
    low, high = _bounds(c)
    v, dv, ld, dld = D.bounded_transform(u, low, high)
    alpha = (low - loc) / scale if np.isfinite(low) else -np.inf
    beta = (high - loc) / scale if np.isfinite(high) else np.inf
    logz = D.log_ndtr_diff(alpha, beta)[0]
    lp = D.normal_logpdf(v, loc, scale) - logz
    dlp = -(v - loc) / scale ** 2 * dv
    return v, dv, lp + ld, dlp + dld


def _tn_logpdf_grads(x, mu, sigma, low, high):
    """log TN(x; mu, sigma, low, high) and its partials wrt x, mu, sigma (finite bounds)."""
    alpha, beta = (low - mu) / sigma, (high - mu) / sigma
    logz, ra, rb = D.log_ndtr_diff(alpha, beta)
    r = (x - mu) / sigma
    lp = -0.5 * r * r - np.log(sigma) - 0.5 * D.LOG_2PI - logz
    dlogz_dmu = -(rb - ra) / sigma
    dlogz_dsig = -(beta * rb - alpha * ra) / sigma
    return lp, -r / sigma, r / sigma - dlogz_dmu, (r * r - 1.0) / sigma - dlogz_dsig


def potential_and_grad(spec: ModelSpec, z: dict, obs, diag=None, mh_loc=None, mh_scale=None,
                       return_terms=False, dtype=np.float64):
    """U(z) = -(sum of site log-probs + log-Jacobians) and dU/dz for every site in ``z``.

    z: unconstrained values, scalars for hyper sites, shape [N] for per-star sites.
    obs, diag: [N, k] (k = number of observable columns; 'mag' expands to one column per band).
    mh_loc / mh_scale: optional per-star M_H prior arrays (scripts/model_star.py:126-128).
    dtype: np.float64 (the oracle proper) or np.float32 -- every array operation in single precision, i.e. the
    reference's DEFAULT precision (jax without --x64, scripts/model_star.py:73,86); used only to calibrate what
    round-off an fp32 evaluation of this graph carries (tests/test_oracle_fp32_floor.py)."""
    c = spec.const
    emu = spec.emulator
    n = spec.num_stars
    f64 = np.dtype(dtype).type
    zz = {k: np.asarray(v, f64) for k, v in z.items()}
    obs = np.asarray(obs, f64).reshape(n, -1)
    diag = np.zeros_like(obs) if diag is None else np.broadcast_to(np.asarray(diag, f64), obs.shape)
    logp = 0.0
    g = {}

    # ---- hyper-parameters (:277-297) ------------------------------------------------------
    hier = spec.level == "hierarchical"
    if hier:
        y0 = zz["Y_0"]
        lp = D.normal_logpdf(y0, c["Y_0"]["loc"], c["Y_0"]["scale"])
        g_y0 = -(y0 - c["Y_0"]["loc"]) / c["Y_0"]["scale"] ** 2
        logp += lp
        dydz, ddydz, lt, g_udydz = _tn_site(zz["dY_dZ"], c["dY_dZ"])
        logp += lt
        prec_y = np.exp(zz["precision_Y"])
        cy = c["precision_Y"]
        logp += D.gamma_logpdf(prec_y, cy["concentration"], cy["rate"]) + zz["precision_Y"]
        g_upy = ((cy["concentration"] - 1) / prec_y - cy["rate"]) * prec_y + 1.0
        sigma_y = prec_y ** -0.5
        mu_a, dmu_a, lt, g_umua = _tn_site(zz["mu_a"], c["mu_a"])
        logp += lt
        prec_a = np.exp(zz["precision_a"])
        ca = c["precision_a"]
        logp += D.gamma_logpdf(prec_a, ca["concentration"], ca["rate"]) + zz["precision_a"]
        g_upa = ((ca["concentration"] - 1) / prec_a - ca["rate"]) * prec_a + 1.0
        sigma_a = prec_a ** -0.5

    # ---- per-star priors (:148-201 / :299-339) ---------------------------------------------
    df = c["delta"]["df"]
    u_tau = zz["scaled_precision"]
    tau = np.exp(u_tau)
    lp_star = D.gamma_logpdf(tau, df / 2, df / 2) + u_tau
    g_utau = ((df / 2 - 1) / tau - df / 2) * tau + 1.0

    log_evol, dle, lt, g_uevol = _tn_site(zz["log_evol"], c["log_evol"])
    lp_star = lp_star + lt
    evol = 10.0 ** log_evol
    log_mass, dlm, lt, g_umass = _tn_site(zz["log_mass"], c["log_mass"])
    lp_star = lp_star + lt
    mass = 10.0 ** log_mass
    loc_mh = c["M_H"]["loc"] if mh_loc is None else np.asarray(mh_loc, f64)
    scale_mh = c["M_H"]["scale"] if mh_scale is None else np.asarray(mh_scale, f64)
    mh, dmh, lt, g_umh = _tn_site(zz["M_H"], c["M_H"], loc_mh, scale_mh)
    lp_star = lp_star + lt

    ylo, yhi = c["Y"]["low"], c["Y"]["high"]
    alo, ahi = c["a_MLT"]["low"], c["a_MLT"]["high"]
    t = 10.0 ** -(mh + LOG_ZX_SUN)
    q = 1.0 / (1.0 + t)                       # 1 / zxp1
    dq_dmh = LN10 * q * (1.0 - q)
    g_mh_prior = 0.0                           # d(log prior)/d mh through mu_Y etc.
    extra = {}
    extra_terms = {}
    if hier:
        # :318-333 -- decentred, bounds depend on (hyper, M_H); Y = mu + sigma*Y_dec = ylo + w*s
        f = dydz * q
        mu_y = (y0 + f) / (1.0 + f)
        dmu_dy0 = 1.0 / (1.0 + f)
        dmu_df = (1.0 - y0) / (1.0 + f) ** 2
        sy = D.sigmoid(zz["Y_decentered"])
        y = ylo + (yhi - ylo) * sy
        dy = (yhi - ylo) * sy * (1 - sy)
        # log N(Y_dec;0,1) - log Z + log(beta-alpha) + log s(1-s), with Y_dec = (y-mu)/sigma
        lpy, dl_dy, dl_dmu, dl_dsig = _tn_logpdf_grads(y, mu_y, sigma_y, ylo, yhi)
        lp_star = lp_star + lpy + math.log(yhi - ylo) + D.log_sigmoid_prod(zz["Y_decentered"])
        g_uy = dl_dy * dy + (1 - 2 * sy)
        g_mh_prior = g_mh_prior + dl_dmu * dmu_df * dydz * dq_dmh
        g_y0 = g_y0 + np.sum(dl_dmu * dmu_dy0)
        g_dydz = np.sum(dl_dmu * dmu_df * q)
        g_sigy = np.sum(dl_dsig)
        sa = D.sigmoid(zz["a_decentered"])
        a_mlt = alo + (ahi - alo) * sa
        da = (ahi - alo) * sa * (1 - sa)
        lpa, dl_da, dl_dmua, dl_dsiga = _tn_logpdf_grads(a_mlt, mu_a, sigma_a, alo, ahi)
        lp_star = lp_star + lpa + math.log(ahi - alo) + D.log_sigmoid_prod(zz["a_decentered"])
        g_ua = dl_da * da + (1 - 2 * sa)
        g_mua = np.sum(dl_dmua)
        g_siga = np.sum(dl_dsiga)
        # conditioning of the five hyper-gradient sums: sum_i |contribution_i| in unconstrained units
        extra_terms = dict(hyper_abs=np.array([
            np.sum(np.abs(dl_dmu * dmu_dy0)), np.sum(np.abs(dl_dmu * dmu_df * q)) * abs(ddydz),
            np.sum(np.abs(dl_dsig)) * 0.5 * sigma_y, np.sum(np.abs(dl_dmua)) * abs(dmu_a),
            np.sum(np.abs(dl_dsiga)) * 0.5 * sigma_a], float))
    elif spec.prior == "flat":
        # :163-165 Uniform priors, interval transform
        y, dy, ldy, dldy = D.bounded_transform(zz["Y"], ylo, yhi)
        lp_star = lp_star - math.log(yhi - ylo) + ldy
        g_uy = dldy
        a_mlt, da, lda, dlda = D.bounded_transform(zz["a_MLT"], alo, ahi)
        lp_star = lp_star - math.log(ahi - alo) + lda
        g_ua = dlda
    else:
        # :167-193 informative: per-star precision sites; uncertainty of the enrichment law folded in
        cy, ca = c["precision_Y"], c["precision_a"]
        py = np.exp(zz["precision_Y"])
        lp_star = lp_star + D.gamma_logpdf(py, cy["concentration"], cy["rate"]) + zz["precision_Y"]
        g_upy_s = ((cy["concentration"] - 1) / py - cy["rate"]) * py + 1.0
        sig_y0 = py ** -0.5
        dydz_c, y0_c = c["dY_dZ"]["loc"], c["Y_0"]["loc"]
        zxp1 = t + 1.0
        fp1 = dydz_c / zxp1 + 1.0
        mu_y = (y0_c + dydz_c / zxp1) / fp1                       # enrichment_law :136-138
        f_dydz = (1.0 - y0_c) / fp1 ** 2 / zxp1                    # :140-146
        f_y0 = 1.0 / fp1
        s1, s0 = c["dY_dZ"]["scale"], c["Y_0"]["scale"]
        var_y = sig_y0 ** 2 + (f_dydz * s1) ** 2 + (f_y0 * s0) ** 2
        sig_yt = np.sqrt(var_y)
        y, dy, ldy, dldy = D.bounded_transform(zz["Y"], ylo, yhi)
        lpy, dl_dy, dl_dmu, dl_dsig = _tn_logpdf_grads(y, mu_y, sig_yt, ylo, yhi)
        lp_star = lp_star + lpy + ldy
        g_uy = dl_dy * dy + dldy
        # d/d mh of mu_y, f_dydz, f_y0   (zxp1' = -ln10 * t)
        dz = -LN10 * t
        dfp1 = -dydz_c / zxp1 ** 2 * dz
        dmu = (1.0 - y0_c) / fp1 ** 2 * (-dfp1) * -1.0            # mu = 1 - (1-y0)/fp1
        dfd = (1.0 - y0_c) * (-2.0 * dfp1 / fp1 ** 3 / zxp1 - dz / fp1 ** 2 / zxp1 ** 2)
        dfy = -dfp1 / fp1 ** 2
        dsig = (f_dydz * s1 ** 2 * dfd + f_y0 * s0 ** 2 * dfy) / sig_yt
        g_mh_prior = g_mh_prior + dl_dmu * dmu + dl_dsig * dsig
        # d sig_yt / d u_precision_Y = sig_y0/sig_yt * (-sig_y0/2)
        g_upy_s = g_upy_s + dl_dsig * (-0.5 * sig_y0 ** 2 / sig_yt)
        pa = np.exp(zz["precision_a"])
        lp_star = lp_star + D.gamma_logpdf(pa, ca["concentration"], ca["rate"]) + zz["precision_a"]
        g_upa_s = ((ca["concentration"] - 1) / pa - ca["rate"]) * pa + 1.0
        sig_a0 = pa ** -0.5
        sig_at = np.sqrt(sig_a0 ** 2 + c["mu_a"]["scale"] ** 2)
        a_mlt, da, lda, dlda = D.bounded_transform(zz["a_MLT"], alo, ahi)
        lpa, dl_da, _, dl_dsiga = _tn_logpdf_grads(a_mlt, c["mu_a"]["loc"], sig_at, alo, ahi)
        lp_star = lp_star + lpa + lda
        g_ua = dl_da * da + dlda
        g_upa_s = g_upa_s + dl_dsiga * (-0.5 * sig_a0 ** 2 / sig_at)
        extra["precision_Y"], extra["precision_a"] = g_upy_s, g_upa_s

    if spec.phot:
        cd = c["scaled_distance"]
        sd = D.sigmoid(zz["scaled_distance"])
        c1, c0 = cd["concentration1"], cd["concentration0"]
        lp_star = lp_star + D.beta_logpdf(sd, c1, c0) + D.log_sigmoid_prod(zz["scaled_distance"])
        g_usd = c1 * (1 - sd) - c0 * sd
        dist = c["max_distance"] * sd

    # ---- predict (:203-218) -----------------------------------------------------------------
    x = np.stack([evol, mass, mh, y, a_mlt], axis=-1).astype(emu.dtype)
    out, masks = emu.forward(x, cache=True)
    out = out.astype(f64) + np.asarray(c["delta"]["loc"], f64)
    mean = out @ A_MAT.T.astype(f64) + C_VEC.astype(f64)
    mean[:, 5] += log_mass
    mean[:, 6] += log_mass
    if spec.phot:
        av = np.full(n, float(c["Av"]), f64)
        log_teff, log_g, log_l = mean[:, 1], mean[:, 5], mean[:, 4]
        if spec.bc_mode == "grid":      # _jax/star.py:40-44
            teff = 10.0 ** log_teff
            bc_in = np.stack([teff, log_g, mh, av], axis=-1)
        else:                           # _jax/star.py:48-52
            bc_in = np.stack([log_teff, log_g, mh, av], axis=-1)
        bcv, bc_cache = spec.bc.forward(bc_in.astype(spec.bc.dtype), cache=True)
        bcv = bcv.astype(f64)
        bol = BOL_MAG_SUN - 2.5 * log_l                                   # :54
        mag = (bol[:, None] - bcv) - 5.0 * (1.0 - np.log10(dist))[:, None]  # :55,61
        plx = 1e3 / dist                                                  # :62
        mean = np.concatenate([mean, mag, plx[:, None]], axis=1)

    # ---- likelihood (:220-240) ----------------------------------------------------------------
    idx = spec.obs_idx
    m = mean[:, idx]
    r = obs - m
    g_mean = np.zeros_like(mean)
    if spec.kind == "diag":
        var_ext = spec.var_ext.astype(f64)
        s2 = var_ext[idx] / tau[:, None] + diag
        ll = -0.5 * r * r / s2 - 0.5 * np.log(s2) - 0.5 * D.LOG_2PI
        g_m = r / s2
        dl_ds2 = 0.5 * r * r / s2 ** 2 - 0.5 / s2
        g_tau = np.sum(dl_ds2 * (-var_ext[idx] / tau[:, None] ** 2), axis=1)
        ll = ll.sum(axis=1)
    else:
        csub = spec.cov_ext[np.ix_(idx, idx)].astype(f64)
        k = len(idx)
        S = csub[None] / tau[:, None, None] + diag[:, :, None] * np.eye(k, dtype=f64)[None]
        L = np.linalg.cholesky(S)
        yv = np.linalg.solve(S, r[:, :, None])[:, :, 0]
        logdet = 2.0 * np.sum(np.log(np.diagonal(L, axis1=1, axis2=2)), axis=1)
        ll = -0.5 * np.sum(r * yv, axis=1) - 0.5 * logdet - 0.5 * k * D.LOG_2PI
        g_m = yv
        sinv_c = np.linalg.solve(S, np.broadcast_to(csub, S.shape))
        tr = np.trace(sinv_c, axis1=1, axis2=2)
        quad = np.einsum("ni,ij,nj->n", yv, csub, yv)
        g_tau = 0.5 / tau ** 2 * (tr - quad)
    g_mean[:, idx] = g_m
    loglik = ll

    # ---- reverse mode -------------------------------------------------------------------------
    g_logmass = g_mean[:, 5] + g_mean[:, 6]
    g_mh = np.zeros(n, f64)
    if spec.phot:
        g_mag = g_mean[:, 7:7 + spec.bands]
        g_plx = g_mean[:, 7 + spec.bands]
        g_bcin = spec.bc.vjp(bc_cache, -g_mag).astype(f64)
        g_mean[:, 4] += -2.5 * g_mag.sum(axis=1)
        g_mean[:, 1] += g_bcin[:, 0] * (LN10 * teff if spec.bc_mode == "grid" else 1.0)
        g_mean[:, 5] += g_bcin[:, 1]
        g_logmass = g_logmass + g_bcin[:, 1]
        g_mh = g_mh + g_bcin[:, 2]
        g_dist = g_mag.sum(axis=1) * 5.0 / (dist * LN10) - g_plx * 1e3 / dist ** 2
        g_usd = g_usd + g_dist * c["max_distance"] * sd * (1 - sd)
    g_out = g_mean[:, :7] @ A_MAT.astype(f64)
    g_x = emu.vjp(masks, g_out.astype(emu.dtype)).astype(f64)

    grad = {}
    grad["scaled_precision"] = -(g_utau + g_tau * tau)
    grad["log_evol"] = -(g_uevol + g_x[:, 0] * LN10 * evol * dle)
    grad["log_mass"] = -(g_umass + (g_x[:, 1] * LN10 * mass + g_logmass) * dlm)
    grad["M_H"] = -(g_umh + (g_x[:, 2] + g_mh + g_mh_prior) * dmh)
    ykey, akey = ("Y_decentered", "a_decentered") if hier else ("Y", "a_MLT")
    grad[ykey] = -(g_uy + g_x[:, 3] * dy)
    grad[akey] = -(g_ua + g_x[:, 4] * da)
    for k_, v_ in extra.items():
        grad[k_] = -v_
    if spec.phot:
        grad["scaled_distance"] = -g_usd
    if hier:
        grad["Y_0"] = -g_y0
        grad["dY_dZ"] = -(g_udydz + g_dydz * ddydz)
        grad["precision_Y"] = -(g_upy + g_sigy * (-0.5 * sigma_y))
        grad["mu_a"] = -(g_umua + g_mua * dmu_a)
        grad["precision_a"] = -(g_upa + g_siga * (-0.5 * sigma_a))

    per_star = lp_star + loglik
    u = -(logp + math.fsum(per_star.tolist()))
    if return_terms:
        return u, grad, dict(mean=mean, per_star=per_star, hyper_logp=logp, x=x, out=out, **extra_terms)
    return u, grad


def potential(spec, z, obs, diag=None, **kw):
    return potential_and_grad(spec, z, obs, diag, **kw)[0]
