"""numpyro transforms / log-densities restated with analytic derivatives (oracle; tests only).

Third-party arithmetic: numpyro (0.13.2 / 0.14.0 observed in the reference's notebooks; unpinned
in pyproject.toml:9; not importable here) -- PARITY UNPINNED, closed forms restated:

  biject_to(positive)        = exp(u)                         log|J| = u
  biject_to(interval(a,b))   = a + (b-a)*sigmoid(u)           log|J| = log(b-a) - softplus(u) - softplus(-u)
  biject_to(less_than(b))    = b - exp(u)                     log|J| = u
  biject_to(greater_than(a)) = a + exp(u)                     log|J| = u
  Normal.log_prob            = -((x-mu)/s)^2/2 - log s - log(2 pi)/2
  Gamma(c, rate).log_prob    = c log rate + (c-1) log x - rate x - lgamma(c)
  Beta(c1, c0).log_prob      = (c1-1) log x + (c0-1) log(1-x) - betaln(c1, c0)
  Uniform(a, b).log_prob     = -log(b-a)
  TruncatedNormal            = Normal.log_prob - log(Phi(beta) - Phi(alpha))   (one side -> +-inf)

Call sites in the reference: src/celestify/numpyro_models.py:149-196, 279-288, 302-331.
numpyro evaluates Phi(beta)-Phi(alpha) on the tail nearer to loc; here log-space forms
(scipy.special.log_ndtr) are used, which agree to rounding and do not underflow.
"""
from __future__ import annotations

import math
import numpy as np
from scipy import special as sp

LOG_2PI = math.log(2.0 * math.pi)
LN10 = math.log(10.0)


def sigmoid(u):
    return sp.expit(u)


def log_sigmoid_prod(u):
    """log(s (1-s)) = -softplus(u) - softplus(-u)."""
    return -(np.logaddexp(0.0, u) + np.logaddexp(0.0, -u))


def normal_logpdf(x, loc, scale):
    r = (x - loc) / scale
    return -0.5 * r * r - np.log(scale) - 0.5 * LOG_2PI


def phi(x):
    return np.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)


def log_ndtr_diff(alpha, beta):
    """log(Phi(beta) - Phi(alpha)) for alpha < beta, stable in both tails; +-inf allowed.

    Returns (logZ, phi(alpha)/Z, phi(beta)/Z)."""
    dt = np.result_type(np.asarray(alpha).dtype, np.asarray(beta).dtype)
    dt = dt if dt == np.float32 else np.float64
    alpha, beta = np.broadcast_arrays(np.asarray(alpha, dt), np.asarray(beta, dt))
    # mirror so that the mass sits in the lower tail whenever both bounds are positive
    flip = alpha > 0
    a = np.where(flip, -beta, alpha)
    b = np.where(flip, -alpha, beta)
    lb = sp.log_ndtr(b)
    la = sp.log_ndtr(a)  # -inf when a = -inf
    with np.errstate(divide="ignore", invalid="ignore"):
        logz = lb + np.log1p(-np.exp(la - lb))
    with np.errstate(over="ignore", invalid="ignore"):
        ra = np.where(np.isfinite(alpha), np.exp(-0.5 * alpha * alpha - 0.5 * LOG_2PI - logz), 0.0)
        rb = np.where(np.isfinite(beta), np.exp(-0.5 * beta * beta - 0.5 * LOG_2PI - logz), 0.0)
    return logz, ra, rb


def bounded_transform(u, low, high):
    """numpyro biject_to for real / greater_than / less_than / interval supports.

    Returns value, dvalue/du, log|J|, dlog|J|/du."""
    lo_f, hi_f = np.isfinite(low), np.isfinite(high)
    if lo_f and hi_f:
        s = sigmoid(u)
        w = high - low
        return low + w * s, w * s * (1 - s), math.log(w) + log_sigmoid_prod(u), 1 - 2 * s
    if hi_f:
        e = np.exp(u)
        return high - e, -e, u, np.ones_like(u)
    if lo_f:
        e = np.exp(u)
        return low + e, e, u, np.ones_like(u)
    return u, np.ones_like(u), np.zeros_like(u), np.zeros_like(u)


def truncnorm_const_logz(loc, scale, low, high):
    alpha = (low - loc) / scale if np.isfinite(low) else -np.inf
    beta = (high - loc) / scale if np.isfinite(high) else np.inf
    return log_ndtr_diff(alpha, beta)[0]


def gamma_logpdf(x, conc, rate):
    return conc * np.log(rate) + (conc - 1) * np.log(x) - rate * x - sp.gammaln(conc)


def beta_logpdf(x, c1, c0):
    return (c1 - 1) * np.log(x) + (c0 - 1) * np.log1p(-x) - sp.betaln(c1, c0)
