"""Independent torch-autograd restatement of the same potential (oracle cross-check; tests only).

Built from torch.distributions (Normal, Gamma, Beta, Uniform, MultivariateNormal, and
``biject_to``-style transforms) with reverse-mode autodiff, so neither the log-densities nor the
gradients share code with ``oracle/model.py``.  Also used as the multi-threaded fp32 CPU baseline
("torch-CPU autograd").  Reference: src/celestify/numpyro_models.py:148-346.
"""
from __future__ import annotations

import math
import numpy as np
import torch
import torch.distributions as td
from torch.distributions import constraints as tc

from . import model as M


def _log_ndtr_diff(a, b):
    """log(Phi(b) - Phi(a)); mirror to the lower tail when both are positive."""
    flip = a > 0
    a2 = torch.where(flip, -b, a)
    b2 = torch.where(flip, -a, b)
    lb = torch.special.log_ndtr(b2)
    la = torch.special.log_ndtr(a2)
    return lb + torch.log1p(-torch.exp(la - lb))


def _tn_logprob(x, loc, scale, low, high):
    dt = x.dtype
    base = td.Normal(torch.as_tensor(loc, dtype=dt), torch.as_tensor(scale, dtype=dt)).log_prob(x)
    lo_f, hi_f = math.isfinite(low), math.isfinite(high)
    if lo_f and hi_f:
        return base - _log_ndtr_diff((low - loc) / scale + 0 * x, (high - loc) / scale + 0 * x)
    if hi_f:
        return base - torch.special.log_ndtr(torch.as_tensor((high - loc) / scale, dtype=dt))
    if lo_f:
        return base - torch.special.log_ndtr(-torch.as_tensor((low - loc) / scale, dtype=dt))
    return base


def _constrain(u, low, high):
    lo_f, hi_f = math.isfinite(low), math.isfinite(high)
    if lo_f and hi_f:
        t = td.transform_to(tc.interval(low, high))
    elif hi_f:
        t = td.transform_to(tc.less_than(high))
    elif lo_f:
        t = td.transform_to(tc.greater_than(low))
    else:
        return u, torch.zeros_like(u)
    v = t(u)
    return v, t.log_abs_det_jacobian(u, v)


class TorchNet(torch.nn.Module):
    def __init__(self, net, dtype):
        super().__init__()
        self.w = [torch.as_tensor(np.asarray(w), dtype=dtype) for w in net.weights]
        self.b = [torch.as_tensor(np.asarray(b), dtype=dtype) for b in net.bias]
        self.il, self.isc = torch.as_tensor(net.in_loc, dtype=dtype), torch.as_tensor(net.in_scale, dtype=dtype)
        self.ol, self.osc = torch.as_tensor(net.out_loc, dtype=dtype), torch.as_tensor(net.out_scale, dtype=dtype)

    def forward(self, x):
        h = (x - self.il) / self.isc
        for w, b in zip(self.w[:-1], self.b[:-1]):
            h = torch.relu(h @ w + b)
        return self.ol + self.osc * (h @ self.w[-1] + self.b[-1])


def _grid_interp(points, values, x):
    """Multilinear interpolation, torch version (autograd supplies d/dx)."""
    nd = len(points)
    idx, tt = [], []
    oob = torch.zeros(x.shape[0], dtype=torch.bool)
    for d in range(nd):
        g = points[d]
        xd = x[:, d]
        i = torch.clamp(torch.searchsorted(g, xd.detach().contiguous(), right=False) - 1, 0, g.numel() - 2)
        idx.append(i)
        tt.append((xd - g[i]) / (g[i + 1] - g[i]))
        oob |= (xd < g[0]) | (xd > g[-1])
    out = 0
    for corner in range(1 << nd):
        w = 1
        sel = []
        for d in range(nd):
            cbit = (corner >> d) & 1
            w = w * (tt[d] if cbit else 1 - tt[d])
            sel.append(idx[d] + cbit)
        out = out + values[tuple(sel)] * w[:, None]
    return torch.where(oob[:, None], torch.full_like(out, float("nan")), out)


class TorchPotential:
    """potential(z) with autograd; dtype float64 (cross-check) or float32 (CPU baseline)."""

    def __init__(self, spec: M.ModelSpec, obs, diag=None, mh_loc=None, mh_scale=None, dtype=torch.float64):
        self.spec, self.dtype = spec, dtype
        t = lambda a: torch.as_tensor(np.array(a, np.float64), dtype=dtype)
        n = spec.num_stars
        self.obs = t(obs).reshape(n, -1)
        self.diag = torch.zeros_like(self.obs) if diag is None else t(np.broadcast_to(diag, self.obs.shape))
        c = spec.const
        self.mh_loc = t(c["M_H"]["loc"] if mh_loc is None else mh_loc)
        self.mh_scale = t(c["M_H"]["scale"] if mh_scale is None else mh_scale)
        self.emu = TorchNet(spec.emulator, dtype)
        self.A = t(M.A_MAT)
        self.cvec = t(M.C_VEC)
        self.dloc = t(c["delta"]["loc"])
        self.var = t(spec.var_ext)
        self.cov = t(spec.cov_ext)
        self.idx = torch.as_tensor(spec.obs_idx)
        if spec.bc_mode == "grid":
            self.bc_points = [t(p) for p in spec.bc.points]
            self.bc_values = t(spec.bc.values)
        elif spec.bc_mode == "net":
            self.bc_net = TorchNet(spec.bc, dtype)

    def log_joint(self, z):
        s, c, dt = self.spec, self.spec.const, self.dtype
        lp = torch.zeros((), dtype=dt)
        hier = s.level == "hierarchical"
        T = lambda v: torch.as_tensor(v, dtype=dt)

        def tn(name, u, loc=None, scale=None):
            cc = c[name]
            low, high = M._bounds(cc)
            v, ld = _constrain(u, low, high)
            loc = cc["loc"] if loc is None else loc
            scale = cc["scale"] if scale is None else scale
            return v, _tn_logprob(v, loc, scale, low, high) + ld

        def gamma(u, conc, rate):
            v = torch.exp(u)
            return v, td.Gamma(torch.as_tensor(conc, dtype=dt), torch.as_tensor(rate, dtype=dt)).log_prob(v) + u

        if hier:
            y0 = z["Y_0"]
            lp = lp + td.Normal(T(c["Y_0"]["loc"]), T(c["Y_0"]["scale"])).log_prob(y0)
            dydz, l = tn("dY_dZ", z["dY_dZ"]); lp = lp + l
            py, l = gamma(z["precision_Y"], c["precision_Y"]["concentration"], c["precision_Y"]["rate"]); lp = lp + l
            sigma_y = py ** -0.5
            mu_a, l = tn("mu_a", z["mu_a"]); lp = lp + l
            pa, l = gamma(z["precision_a"], c["precision_a"]["concentration"], c["precision_a"]["rate"]); lp = lp + l
            sigma_a = pa ** -0.5
        df = c["delta"]["df"]
        tau, l = gamma(z["scaled_precision"], df / 2, df / 2); lp = lp + l.sum()
        log_evol, l = tn("log_evol", z["log_evol"]); lp = lp + l.sum()
        log_mass, l = tn("log_mass", z["log_mass"]); lp = lp + l.sum()
        mh, l = tn("M_H", z["M_H"], self.mh_loc, self.mh_scale); lp = lp + l.sum()
        evol, mass = 10.0 ** log_evol, 10.0 ** log_mass
        ylo, yhi, alo, ahi = c["Y"]["low"], c["Y"]["high"], c["a_MLT"]["low"], c["a_MLT"]["high"]
        if hier:
            f = dydz / (10.0 ** -(mh + M.LOG_ZX_SUN) + 1)
            mu_y = (y0 + f) / (1 + f)
            low, high = (ylo - mu_y) / sigma_y, (yhi - mu_y) / sigma_y
            s01 = torch.sigmoid(z["Y_decentered"])
            y_dec = low + (high - low) * s01
            ld = torch.log(high - low) + torch.nn.functional.logsigmoid(z["Y_decentered"]) \
                + torch.nn.functional.logsigmoid(-z["Y_decentered"])
            lp = lp + (td.Normal(T(0.0), T(1.0)).log_prob(y_dec) - _log_ndtr_diff(low, high) + ld).sum()
            y = mu_y + sigma_y * y_dec
            low, high = (alo - mu_a) / sigma_a, (ahi - mu_a) / sigma_a
            s01 = torch.sigmoid(z["a_decentered"])
            a_dec = low + (high - low) * s01
            ld = torch.log(high - low) + torch.nn.functional.logsigmoid(z["a_decentered"]) \
                + torch.nn.functional.logsigmoid(-z["a_decentered"])
            lp = lp + (td.Normal(T(0.0), T(1.0)).log_prob(a_dec) - _log_ndtr_diff(low + 0 * a_dec, high + 0 * a_dec) + ld).sum()
            a_mlt = mu_a + sigma_a * a_dec
        elif s.prior == "flat":
            y, ld = _constrain(z["Y"], ylo, yhi)
            lp = lp + (td.Uniform(T(ylo), T(yhi)).log_prob(y.clamp(ylo + 1e-12, yhi - 1e-12)) + ld).sum()
            a_mlt, ld = _constrain(z["a_MLT"], alo, ahi)
            lp = lp + (td.Uniform(T(alo), T(ahi)).log_prob(a_mlt.clamp(alo + 1e-12, ahi - 1e-12)) + ld).sum()
        else:
            py, l = gamma(z["precision_Y"], c["precision_Y"]["concentration"], c["precision_Y"]["rate"]); lp = lp + l.sum()
            sigma_y = py ** -0.5
            dydz, y0 = c["dY_dZ"]["loc"], c["Y_0"]["loc"]
            zxp1 = 10.0 ** -(mh + M.LOG_ZX_SUN) + 1
            f = dydz / zxp1
            mu_y = (y0 + f) / (1 + f)
            fp1 = dydz / zxp1 + 1
            f_dydz, f_y0 = (1 - y0) / fp1 ** 2 / zxp1, 1 / fp1
            sig = torch.sqrt(sigma_y ** 2 + (f_dydz * c["dY_dZ"]["scale"]) ** 2 + (f_y0 * c["Y_0"]["scale"]) ** 2)
            y, ld = _constrain(z["Y"], ylo, yhi)
            lp = lp + (td.Normal(mu_y, sig).log_prob(y) - _log_ndtr_diff((ylo - mu_y) / sig, (yhi - mu_y) / sig) + ld).sum()
            pa, l = gamma(z["precision_a"], c["precision_a"]["concentration"], c["precision_a"]["rate"]); lp = lp + l.sum()
            sig = torch.sqrt(pa ** -1.0 + c["mu_a"]["scale"] ** 2)
            mu_a = c["mu_a"]["loc"]
            a_mlt, ld = _constrain(z["a_MLT"], alo, ahi)
            lp = lp + (td.Normal(T(mu_a), sig).log_prob(a_mlt) - _log_ndtr_diff((alo - mu_a) / sig, (ahi - mu_a) / sig) + ld).sum()
        x = torch.stack([evol, mass, mh, y, a_mlt], dim=-1)
        out = self.emu(x) + self.dloc
        mean = out @ self.A.T + self.cvec
        add = torch.zeros_like(mean)
        add[:, 5] = log_mass
        add[:, 6] = log_mass
        mean = mean + add
        if s.phot:
            cd = c["scaled_distance"]
            sd = torch.sigmoid(z["scaled_distance"])
            ld = torch.nn.functional.logsigmoid(z["scaled_distance"]) + torch.nn.functional.logsigmoid(-z["scaled_distance"])
            lp = lp + (td.Beta(torch.as_tensor(cd["concentration1"], dtype=dt),
                               torch.as_tensor(cd["concentration0"], dtype=dt)).log_prob(sd) + ld).sum()
            dist = c["max_distance"] * sd
            av = torch.full_like(mh, float(c["Av"]))
            if s.bc_mode == "grid":
                bc = _grid_interp(self.bc_points, self.bc_values, torch.stack([10.0 ** mean[:, 1], mean[:, 5], mh, av], -1))
            else:
                bc = self.bc_net(torch.stack([mean[:, 1], mean[:, 5], mh, av], -1))
            bol = M.BOL_MAG_SUN - 2.5 * mean[:, 4]
            mag = (bol[:, None] - bc) - 5 * (1 - torch.log10(dist))[:, None]
            mean = torch.cat([mean, mag, (1e3 / dist)[:, None]], dim=1)
        m = mean[:, self.idx]
        if s.kind == "diag":
            sc = torch.sqrt(self.var[self.idx] / tau[:, None] + self.diag)
            lp = lp + td.Normal(m, sc).log_prob(self.obs).sum()
        else:
            k = len(self.idx)
            cov = self.cov[self.idx][:, self.idx][None] / tau[:, None, None] + torch.diag_embed(self.diag)
            lp = lp + td.MultivariateNormal(m, covariance_matrix=cov).log_prob(self.obs).sum()
        return lp

    def value_and_grad(self, z):
        zt = {k: torch.as_tensor(np.asarray(v, np.float64), dtype=self.dtype).requires_grad_(True) for k, v in z.items()}
        u = -self.log_joint(zt)
        u.backward()
        return float(u.detach()), {k: v.grad.numpy().astype(np.float64) for k, v in zt.items()}
