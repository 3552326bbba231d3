"""Star-sharded No-U-Turn sampler: the caller of the hot path (numpyro's role in the reference).

The reference drives its models with ``numpyro.infer.NUTS`` / ``MCMC`` (scripts/model_star.py:132-155) using
``init_to_median``, ``target_accept_prob=0.8``, ``find_heuristic_step_size=True``, a regularised DIAGONAL mass
matrix and ``max_tree_depth=10`` (scripts/model_star.py:18-22).  numpyro is third-party and absent here, so this
module restates the same algorithm (multinomial NUTS with the generalised U-turn criterion, dual-averaging step
size, Stan-style windowed variance adaptation -- Hoffman & Gelman 2014, Betancourt 2017) around ONE primitive:

    potential(z) -> (U, dU/dz)      one evaluation of the hot path (libcsi on the GPU, the oracle in the CPU tests)

Vectors are torch tensors (CPU or CUDA: PyTorch is plumbing here).  With the star axis sharded over ranks a vector
is ``[hyper (replicated) | local stars]``: inner products add the hyper part once and SUM-allreduce the star part,
momenta of the hyper sites and every accept / direction decision come from a generator seeded identically on all
ranks, so that every rank builds the same tree.  Energies are kept in float64 (|U| ~ 1e7 at 1 M stars).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch


def _check(rc):
    from . import _lib
    _lib.check(rc)


class VectorSpace:
    """Inner products of ``[hyper | local stars]`` vectors; ``group`` = torch.distributed group or None."""

    def __init__(self, n_hyper: int = 0, group=None, sharded: bool = False):
        self.n_hyper, self.group, self.sharded = n_hyper, group, sharded

    def dots(self, pairs):
        """Several inner products with ONE collective: returns a list of python floats."""
        nh = self.n_hyper
        loc = torch.stack([torch.dot(a[nh:].double(), b[nh:].double()) for a, b in pairs])
        if self.sharded:
            import torch.distributed as dist
            dist.all_reduce(loc, group=self.group)
        if nh:
            loc = loc + torch.stack([torch.dot(a[:nh].double(), b[:nh].double()) for a, b in pairs])
        return loc.tolist()


@dataclass
class _Tree:
    z_left: torch.Tensor
    r_left: torch.Tensor
    g_left: torch.Tensor
    z_right: torch.Tensor
    r_right: torch.Tensor
    g_right: torch.Tensor
    z_prop: torch.Tensor
    u_prop: float
    g_prop: torch.Tensor
    log_w: float            # log of the summed multinomial weights exp(-H)
    r_sum: torch.Tensor
    turning: bool
    diverging: bool
    sum_accept: float
    n_prop: int


def _logaddexp(a, b):
    m = max(a, b)
    return m if m == -math.inf else m + math.log(math.exp(a - m) + math.exp(b - m))


class NUTS:
    """One chain.  ``potential(z) -> (U: float, grad: tensor)`` must return a tensor it will not overwrite."""

    def __init__(self, potential, space: VectorSpace | None = None, target_accept: float = 0.8,
                 max_tree_depth: int = 10, seed: int = 0, rank: int = 0, adapt_mass: bool = True,
                 max_delta_energy: float = 1000.0):
        self.potential, self.space = potential, space or VectorSpace()
        self.target, self.max_depth, self.max_de = target_accept, max_tree_depth, max_delta_energy
        self.adapt_mass = adapt_mass
        self.rng = np.random.default_rng(seed)                    # shared decisions: same seed on every rank
        self.seed, self.rank = seed, rank
        self.gen_shared = self.gen_local = None
        self.n_evals = 0

    # ---- primitives ------------------------------------------------------------------------------------------
    def _pot(self, z):
        self.n_evals += 1
        return self.potential(z)

    def _momentum(self, like):
        nh = self.space.n_hyper
        if self.gen_shared is None:
            self.gen_shared = torch.Generator(device=like.device).manual_seed(1234567 + self.seed)
            self.gen_local = torch.Generator(device=like.device).manual_seed(7654321 + 1000 * self.seed + self.rank)
        r = torch.empty_like(like)
        if nh:
            r[:nh] = torch.randn(nh, generator=self.gen_shared, device=like.device, dtype=like.dtype)
        r[nh:] = torch.randn(like.numel() - nh, generator=self.gen_local, device=like.device, dtype=like.dtype)
        return r * self.inv_mass.rsqrt()                          # r ~ N(0, M), M = 1 / inv_mass

    def _kinetic(self, r):
        return 0.5 * self.space.dots([(r, self.inv_mass * r)])[0]

    def _leapfrog(self, z, r, g, eps):
        r = r - (0.5 * eps) * g
        z = z + eps * (self.inv_mass * r)
        u, g = self._pot(z)
        r = r - (0.5 * eps) * g
        return z, r, u, g

    def _is_turning(self, r_left, r_right, r_sum):
        # numpyro / Stan generalised criterion on the summed momentum (end points counted half)
        if hasattr(self.potential, "uturn"):
            a, b = self.potential.uturn(r_left, r_right, r_sum, self.inv_mass)
            return a <= 0.0 or b <= 0.0
        rc = r_sum - 0.5 * (r_left + r_right)
        a, b = self.space.dots([(self.inv_mass * r_left, rc), (self.inv_mass * r_right, rc)])
        return a <= 0.0 or b <= 0.0

    # ---- tree ------------------------------------------------------------------------------------------------
    def _base(self, z, r, g, eps, h0):
        if hasattr(self.potential, "leapfrog"):   # fused vector kernels around the potential: one host sync
            self.n_evals += 1
            z1, r1, g1, u1, k1 = self.potential.leapfrog(z, r, g, eps, self.inv_mass)
            h1 = u1 + k1
        else:
            z1, r1, u1, g1 = self._leapfrog(z, r, g, eps)
            h1 = u1 + self._kinetic(r1)
        de = h1 - h0
        if not math.isfinite(de):
            de = math.inf
        return _Tree(z1, r1, g1, z1, r1, g1, z1, u1, g1, -de, r1, False, de > self.max_de,
                     math.exp(min(-de, 0.0)) if de < math.inf else 0.0, 1)

    def _combine(self, cur: _Tree, new: _Tree, going_right: bool, biased: bool):
        log_w = _logaddexp(cur.log_w, new.log_w)
        p = math.exp(min(0.0, new.log_w - cur.log_w)) if biased else math.exp(new.log_w - log_w)
        if new.turning or new.diverging:
            p = 0.0  # no proposal is taken from a sub-tree that turned or diverged
        take = self.rng.random() < p
        left, right = (cur, new) if going_right else (new, cur)
        r_sum = cur.r_sum + new.r_sum
        src = new if take else cur
        return _Tree(left.z_left, left.r_left, left.g_left, right.z_right, right.r_right, right.g_right,
                     src.z_prop, src.u_prop, src.g_prop, log_w, r_sum, False, False,
                     cur.sum_accept + new.sum_accept, cur.n_prop + new.n_prop)

    def _build(self, z, r, g, eps, depth, h0):
        if depth == 0:
            return self._base(z, r, g, eps, h0)
        half = self._build(z, r, g, eps, depth - 1, h0)
        if half.turning or half.diverging:
            return half
        going_right = eps > 0
        if going_right:
            other = self._build(half.z_right, half.r_right, half.g_right, eps, depth - 1, h0)
        else:
            other = self._build(half.z_left, half.r_left, half.g_left, eps, depth - 1, h0)
        tree = self._combine(half, other, going_right, biased=False)
        tree.diverging = other.diverging
        tree.turning = other.turning or (not other.diverging and self._is_turning(tree.r_left, tree.r_right, tree.r_sum))
        return tree

    def transition(self, z, u, g, step_size):
        r0 = self._momentum(z)
        h0 = u + self._kinetic(r0)
        tree = _Tree(z, r0, g, z, r0, g, z, u, g, 0.0, r0, False, False, 0.0, 0)
        depth, diverging = 0, False
        while depth < self.max_depth:
            going_right = self.rng.random() < 0.5
            if going_right:
                new = self._build(tree.z_right, tree.r_right, tree.g_right, step_size, depth, h0)
            else:
                new = self._build(tree.z_left, tree.r_left, tree.g_left, -step_size, depth, h0)
            depth += 1
            if new.turning or new.diverging:
                tree.sum_accept += new.sum_accept
                tree.n_prop += new.n_prop
                diverging = new.diverging
                break
            tree = self._combine(tree, new, going_right, biased=True)
            if self._is_turning(tree.r_left, tree.r_right, tree.r_sum):
                break
        accept = tree.sum_accept / max(tree.n_prop, 1)
        return tree.z_prop, tree.u_prop, tree.g_prop, accept, depth, diverging

    # ---- step size and mass matrix adaptation -------------------------------------------------------------------
    def find_heuristic_step_size(self, z, u, g, step_size=1.0):
        """Double / halve the step until the one-step acceptance probability crosses the target (numpyro
        ``find_reasonable_step_size``)."""
        log_target = math.log(self.target)
        direction = 0
        for _ in range(60):
            r = self._momentum(z)
            h0 = u + self._kinetic(r)
            _, r1, u1, _ = self._leapfrog(z, r, g, step_size)
            de = (u1 + self._kinetic(r1)) - h0
            new_dir = 1 if (math.isfinite(de) and -de > log_target) else -1
            if direction and new_dir != direction:
                break
            direction = new_dir
            step_size *= 2.0 ** direction
            if not (1e-10 < step_size < 1e10):
                break
        return step_size

    @staticmethod
    def adaptation_windows(num_warmup):
        """(start, end) of the slow (mass-matrix) windows, Stan / numpyro schedule."""
        if num_warmup < 20:
            return []
        init, term, base = 75, 50, 25
        if init + base + term > num_warmup:
            init, term = int(0.15 * num_warmup), int(0.1 * num_warmup)
            base = num_warmup - init - term
        wins, start, size = [], init, base
        end_slow = num_warmup - term
        while start < end_slow:
            end = start + size
            if end + 2 * size > end_slow:
                end = end_slow
            wins.append((start, end))
            start, size = end, 2 * size
        return wins

    def run(self, z0, num_warmup, num_samples, collect=None, progress=None):
        """Returns dict(samples=[collect(z) per draw], accept, depth, diverging, step_size, n_evals, potential).
        ``collect`` maps the state vector to what is kept (default: the first ``n_hyper`` entries, on the host)."""
        nh = self.space.n_hyper
        collect = collect or (lambda z: z[:max(nh, 1)].detach().double().cpu().numpy().copy())
        z = z0.clone()
        self.inv_mass = torch.ones_like(z)
        u, g = self._pot(z)
        if not math.isfinite(u):
            raise ValueError("non-finite potential at the initial point")
        step = self.find_heuristic_step_size(z, u, g)
        # dual averaging (Nesterov / Hoffman-Gelman): log step
        t0, kappa, gamma = 10.0, 0.75, 0.05
        da = dict(mu=math.log(10.0 * step), x_bar=0.0, h_bar=0.0, t=0)
        windows = self.adaptation_windows(num_warmup) if self.adapt_mass else []
        ends = {e: s for s, e in windows}
        in_window = lambda i: any(s <= i < e for s, e in windows)  # noqa: E731
        w_n, w_mean, w_m2 = 0, torch.zeros_like(z, dtype=torch.float64), torch.zeros_like(z, dtype=torch.float64)
        out = dict(samples=[], accept=[], depth=[], diverging=[], potential=[])
        for it in range(num_warmup + num_samples):
            warm = it < num_warmup
            z, u, g, acc, depth, div = self.transition(z, u, g, step)
            if warm:
                da["t"] += 1
                t = da["t"]
                da["h_bar"] = (1 - 1 / (t + t0)) * da["h_bar"] + (self.target - acc) / (t + t0)
                x = da["mu"] - math.sqrt(t) / gamma * da["h_bar"]
                eta = t ** (-kappa)
                da["x_bar"] = eta * x + (1 - eta) * da["x_bar"]
                step = math.exp(x)
                if in_window(it):
                    w_n += 1
                    d = z.double() - w_mean
                    w_mean += d / w_n
                    w_m2 += d * (z.double() - w_mean)
                if (it + 1) in ends and w_n > 1:
                    var = w_m2 / (w_n - 1)
                    var = var * (w_n / (w_n + 5.0)) + 1e-3 * (5.0 / (w_n + 5.0))      # Stan's regularisation
                    self.inv_mass = var.to(z.dtype)
                    w_n = 0; w_mean.zero_(); w_m2.zero_()
                    step = self.find_heuristic_step_size(z, u, g, math.exp(da["x_bar"]))
                    da = dict(mu=math.log(10.0 * step), x_bar=0.0, h_bar=0.0, t=0)
                if it + 1 == num_warmup:
                    step = math.exp(da["x_bar"]) if da["t"] > 0 else step
                    self.n_evals_warmup = self.n_evals
            else:
                out["samples"].append(collect(z))
                out["accept"].append(acc); out["depth"].append(depth); out["diverging"].append(div)
                out["potential"].append(u)
            if progress:
                progress(it, warm, step, depth, acc)
        out.update(step_size=step, n_evals=self.n_evals, z_last=z)
        return out


# ---------------------------------------------------------------------------------------------------------------
# counter-based random numbers: the Python mirror of csi::nuts::hash_u64 (csrc/csi_nuts.cuh)
# ---------------------------------------------------------------------------------------------------------------
_M64 = (1 << 64) - 1
K_MOMENTUM, K_DIRECTION, K_LEAF, K_MERGE = 1, 2, 3, 4


def _mix64(x):
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & _M64
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & _M64
    x ^= x >> 31
    return x


def hash_u64(seed, draw, kind, idx):
    g = 0x9E3779B97F4A7C15
    h = _mix64((seed + g) & _M64)
    h = _mix64(h ^ ((draw + g) & _M64))
    h = _mix64(h ^ ((kind * 0xD1B54A32D192ED03) & _M64))
    return _mix64(h ^ ((idx + g) & _M64))


def hash_uniform(seed, draw, kind, idx):
    return (hash_u64(seed, draw, kind, idx) >> 11) * (1.0 / 9007199254740992.0)


def leaf_to_ckpt(n):
    """numpyro hmc_util._leaf_idx_to_ckpt_idxs: (idx_min, idx_max) of the checkpoints leaf n is tested against."""
    idx_max = bin(n >> 1).count("1")
    trailing = 0
    while n & 1:
        trailing += 1
        n >>= 1
    return idx_max - trailing + 1, idx_max


class IterativeNUTS:
    """Host restatement of the DEVICE sampler (csrc/csi_nuts.cuh), decision for decision: iterative sub-tree building
    with checkpointed U-turn tests, in-place leapfrogs on the growing end, lazily copied sub-tree proposal, the same
    hash stream for directions / leaf accepts / merges.  ``potential(z) -> (U, grad)`` on numpy vectors.  It is the
    reference the GPU tests compare ``csi_nuts_transition`` with; momenta are passed in (the device draws its own)."""

    def __init__(self, potential, seed=0, max_tree_depth=10, max_delta_energy=1000.0, dtype=np.float32):
        self.potential, self.seed, self.max_depth, self.max_de = potential, int(seed), max_tree_depth, max_delta_energy
        self.dtype = dtype

    def transition(self, z, u, g, r0, inv_mass, eps, draw):
        f = self.dtype
        im = np.asarray(inv_mass, f)
        dot = lambda a, b: float(np.dot(a.astype(np.float64), b.astype(np.float64)))  # noqa: E731
        z_prop, g_prop, u_prop = np.array(z, f), np.array(g, f), float(u)
        end = {+1: [z_prop.copy(), np.array(r0, f), g_prop.copy()], -1: [z_prop.copy(), np.array(r0, f), g_prop.copy()]}
        r_sum = np.array(r0, f)
        h0 = u_prop + 0.5 * dot(im * end[1][1], end[1][1])
        logw_tree, sum_acc, n_prop, depth, turning, diverging = 0.0, 0.0, 0, 0, False, False
        while depth < self.max_depth:
            sign = +1 if (hash_u64(self.seed, draw, K_DIRECTION, depth) >> 63) else -1
            e = f(sign * eps)
            zc, rc, gc = end[sign]
            ck_r, ck_rs = {}, {}
            logw_sub, u_sub, z_sub, g_sub, r_sum_sub = -math.inf, u_prop, None, None, None
            pending, stop, sub_turning = False, False, False
            for leaf in range(1 << depth):
                if pending:
                    z_sub, g_sub = zc.copy(), gc.copy()
                rc -= f(0.5) * e * gc
                zc += e * (im * rc)
                u1, g1 = self.potential(zc)
                gc[:] = np.asarray(g1, f)
                rc -= f(0.5) * e * gc
                kin = 0.5 * dot(im * rc, rc)
                r_sum_sub = rc.copy() if leaf == 0 else r_sum_sub + rc
                idx_min, idx_max = leaf_to_ckpt(leaf)
                if leaf % 2 == 0:
                    ck_r[idx_max], ck_rs[idx_max] = rc.copy(), r_sum_sub.copy()
                de = (u1 + kin) - h0
                if not math.isfinite(de):
                    de = math.inf
                w = -de
                div = de > self.max_de
                sum_acc += math.exp(min(-de, 0.0)) if de < math.inf else 0.0
                n_prop += 1
                if leaf == 0:
                    take, logw_sub = True, w
                else:
                    p = 0.0 if w == -math.inf else 1.0 / (1.0 + math.exp(min(logw_sub - w, 700.0)))
                    take = hash_uniform(self.seed, draw, K_LEAF, (depth << 32) | leaf) < p
                    logw_sub = _logaddexp(logw_sub, w)
                pending = take
                if take:
                    u_sub = u1
                turn = False
                for c in range(idx_max, idx_min - 1, -1):
                    sub = r_sum_sub - ck_rs[c] + ck_r[c]
                    cen = sub - f(0.5) * (ck_r[c] + rc)
                    turn = turn or dot(im * ck_r[c], cen) <= 0.0 or dot(im * rc, cen) <= 0.0
                diverging = diverging or div
                sub_turning = sub_turning or turn
                if div or turn:
                    stop = True
                    break
            depth += 1
            if stop:
                turning = turning or sub_turning
                break
            p = min(1.0, math.exp(min(logw_sub - logw_tree, 0.0)))
            if hash_uniform(self.seed, draw, K_MERGE, depth - 1) < p:
                z_prop, g_prop = (zc.copy(), gc.copy()) if pending else (z_sub.copy(), g_sub.copy())
                u_prop = u_sub
            logw_tree = _logaddexp(logw_tree, logw_sub)
            r_sum = r_sum + r_sum_sub
            rl, rr = end[-1][1], end[+1][1]
            cen = r_sum - f(0.5) * (rl + rr)
            if dot(im * rl, cen) <= 0.0 or dot(im * rr, cen) <= 0.0:
                turning = True
                break
        return dict(z=z_prop, g=g_prop, u=u_prop, accept=sum_acc / max(n_prop, 1), depth=depth, n_leapfrog=n_prop,
                    diverging=diverging, turning=turning, h0=h0)


def effective_sample_size(x):
    """Bulk ESS of draws x[chain, draw] (Geyer initial monotone sequence on split chains; the estimator of
    numpyro.diagnostics.effective_sample_size / Stan)."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[None]
    half = x.shape[1] // 2
    x = np.concatenate([x[:, :half], x[:, half:2 * half]], axis=0)      # split chains
    m, n = x.shape
    xc = x - x.mean(axis=1, keepdims=True)
    nfft = 1 << int(np.ceil(np.log2(2 * n)))
    f = np.fft.rfft(xc, nfft, axis=1)
    acov = np.fft.irfft(f * np.conj(f), nfft, axis=1)[:, :n] / n
    var_within = acov[:, 0].mean() * n / (n - 1.0)
    var_plus = acov[:, 0].mean() + (x.mean(axis=1).var(ddof=1) if m > 1 else 0.0)
    if var_plus <= 0:
        return float(m * n)
    rho = 1.0 - (var_within - acov.mean(axis=0) * n / (n - 1.0)) / var_plus
    rho[0] = 1.0
    pairs = rho[: 2 * (n // 2)].reshape(-1, 2).sum(axis=1)             # P_k = rho_2k + rho_2k+1
    pos = np.where(pairs < 0)[0]
    k = pos[0] if len(pos) else len(pairs)
    pk = np.minimum.accumulate(pairs[:k]) if k > 0 else np.array([1.0])
    tau = -1.0 + 2.0 * pk.sum()
    return float(m * n / max(tau, 1.0 / math.log10(max(m * n, 10))))


class DevicePotential:
    """``potential(z)`` on the B200 path: z is a float32 CUDA tensor in ``csi_layout`` order ([hyper | local stars]).
    One libcsi launch + (sharded) one 6-double allreduce; the summed hyper-gradient replaces the local partials."""

    def __init__(self, model, group=None):
        from .parallel import ShardedPotential
        self.sp = ShardedPotential(model, group)
        self.model, self.nh = model, len(model.hyper_sites)
        self._scratch = self._dots_dev = self._dots_pin = None

    def __call__(self, z):
        sp = self.sp
        out, grad = sp.value_and_grad_device(z.contiguous())
        g = grad.clone()
        if self.nh:
            g[: self.nh] = out[1: 1 + self.nh].to(g.dtype)
        return float(out[0].item()), g

    # ---- the leapfrog and the U-turn test as three libcsi vector kernels (csi_leapfrog_pre / _post, csi_uturn) ----
    def _sharded(self):
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(self.sp.group) > 1

    def leapfrog(self, z, r, g, eps, inv_mass):
        """One velocity-Verlet step: returns (z1, r1, g1, U(z1), kinetic(r1)) with ONE host synchronisation.
        pre-kernel -> potential kernel (-> allreduce of the 6 sums unless fused) -> post-kernel -> 8 doubles D2H."""
        sp, lib, h = self.sp, self.model._lib, self.model._h
        if self._scratch is None or self._scratch.shape != z.shape:
            self._scratch = torch.empty_like(z)
        z1, r1, g1 = torch.empty_like(z), torch.empty_like(z), torch.empty_like(z)
        stream = torch.cuda.current_stream(z.device)
        kin = sp.out_dev[6:8]
        _check(lib.csi_leapfrog_pre(h, z.data_ptr(), r.data_ptr(), g.data_ptr(), inv_mass.data_ptr(), float(eps),
                                    z1.data_ptr(), self._scratch.data_ptr(), kin.data_ptr(), stream.cuda_stream))
        out, grad = sp.value_and_grad_device(z1)
        _check(lib.csi_leapfrog_post(h, self._scratch.data_ptr(), grad.data_ptr(), out.data_ptr(),
                                     inv_mass.data_ptr(), float(eps), r1.data_ptr(), g1.data_ptr(), kin.data_ptr(),
                                     stream.cuda_stream))
        if self._sharded():
            import torch.distributed as dist
            dist.all_reduce(sp.out_dev[6:7], group=sp.group)
        sp.out_pin.copy_(sp.out_dev, non_blocking=True)
        stream.synchronize()
        o = sp.out_pin
        return z1, r1, g1, float(o[0]), 0.5 * (float(o[6]) + float(o[7]))

    def uturn(self, r_left, r_right, r_sum, inv_mass):
        """The two inner products of the generalised U-turn criterion, one kernel + one host synchronisation."""
        lib, h = self.model._lib, self.model._h
        stream = torch.cuda.current_stream(r_sum.device)
        if self._dots_dev is None:
            self._dots_dev = torch.zeros(4, dtype=torch.float64, device=r_sum.device)
            self._dots_pin = torch.zeros(4, dtype=torch.float64).pin_memory()
        _check(lib.csi_uturn(h, r_left.data_ptr(), r_right.data_ptr(), r_sum.data_ptr(), inv_mass.data_ptr(),
                             self._dots_dev.data_ptr(), stream.cuda_stream))
        if self._sharded():
            import torch.distributed as dist
            dist.all_reduce(self._dots_dev[0:2], group=self.sp.group)
        self._dots_pin.copy_(self._dots_dev, non_blocking=True)
        stream.synchronize()
        d = self._dots_pin
        return float(d[0]) + float(d[2]), float(d[1]) + float(d[3])


class DeviceNUTS:
    """One chain with the tree logic on the device (``csi_nuts_*``, csrc/csi_nuts.cuh): a transition costs one host
    synchronisation per doubling, a leapfrog three stream-ordered launches inside a CUDA graph.  Adaptation follows
    numpyro's warm-up as the reference configures it (scripts/model_star.py:18-22): dual-averaging step size to
    ``target_accept_prob`` 0.8, ``find_heuristic_step_size``, regularised diagonal mass matrix in Stan windows.

    ``model`` must be conditioned (and, for ``world > 1``, sharded with the peer exchange switched on); the state
    vectors are torch tensors only because PyTorch owns the device memory."""

    def __init__(self, model, z0_flat, seed=0, target_accept=0.8, max_tree_depth=10, max_delta_energy=1000.0,
                 adapt_mass=True, star_offset=0, n_total=None):
        import ctypes as C
        from . import _lib
        self.C, self._lib = C, _lib
        self.lib, self.model = model._lib, model
        self.dev = torch.device("cuda", model.device)
        self.nh = len(model.hyper_sites)
        self.target, self.adapt_mass, self.seed = target_accept, adapt_mass, int(seed)
        self.z = torch.as_tensor(np.asarray(z0_flat, np.float32), device=self.dev).contiguous()
        self.g = torch.zeros_like(self.z)
        self.inv_mass = torch.ones_like(self.z)
        h = C.c_void_p()
        n_total = model.n_local if n_total is None else n_total
        torch.cuda.synchronize(self.dev)
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.csi_nuts_create(model._h, self.z.data_ptr(), self.g.data_ptr(), self.inv_mass.data_ptr(),
                                                max_tree_depth, float(max_delta_energy), self.seed, int(star_offset),
                                                int(n_total), C.byref(h)))
        self._s = h
        self.draw = 0
        self.n_evals = self.n_evals_warmup = 0
        self.launches = self.host_syncs = 0
        u = C.c_double()
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.csi_nuts_init_state(self._s, C.byref(u)))
        self.u = u.value
        self.n_evals += 1

    def __del__(self):
        if getattr(self, "_s", None):
            self.lib.csi_nuts_destroy(self._s)
            self._s = None

    def transition(self, step_size):
        st = self._lib.NutsStats()
        with torch.cuda.device(self.dev):
            self._lib.check(self.lib.csi_nuts_transition(self._s, float(step_size), self.draw, self.C.byref(st)))
        self.draw += 1
        self.u = st.potential
        self.n_evals += st.num_leapfrog
        self.launches += st.launches
        self.host_syncs += st.host_syncs
        return st

    def probe(self, step_size):
        de = self.C.c_double()
        with torch.cuda.device(self.dev):
            self._lib.check(self.lib.csi_nuts_probe_step(self._s, float(step_size), self.draw, self.C.byref(de)))
        self.draw += 1
        self.n_evals += 1
        return de.value

    def find_heuristic_step_size(self, step_size=1.0):
        log_target = math.log(self.target)
        direction = 0
        for _ in range(60):
            de = self.probe(step_size)
            new_dir = 1 if (math.isfinite(de) and -de > log_target) else -1
            if direction and new_dir != direction:
                break
            direction = new_dir
            step_size *= 2.0 ** direction
            if not (1e-10 < step_size < 1e10):
                break
        return step_size

    def run(self, num_warmup, num_samples, collect=None, progress=None, step_size=None):
        nh = self.nh
        collect = collect or (lambda z: z[:max(nh, 1)].detach().double().cpu().numpy().copy())
        if not math.isfinite(self.u):
            raise ValueError("non-finite potential at the initial point")
        step = self.find_heuristic_step_size() if step_size is None else step_size
        t0, kappa, gamma = 10.0, 0.75, 0.05
        da = dict(mu=math.log(10.0 * step), x_bar=0.0, h_bar=0.0, t=0)
        windows = NUTS.adaptation_windows(num_warmup) if self.adapt_mass else []
        ends = {e: s for s, e in windows}
        in_window = lambda i: any(s <= i < e for s, e in windows)  # noqa: E731
        z = self.z
        w_n, w_mean, w_m2 = 0, torch.zeros_like(z, dtype=torch.float64), torch.zeros_like(z, dtype=torch.float64)
        out = dict(samples=[], accept=[], depth=[], diverging=[], potential=[], n_leapfrog=[])
        for it in range(num_warmup + num_samples):
            warm = it < num_warmup
            st = self.transition(step)
            acc = st.accept_prob
            if warm:
                da["t"] += 1
                t = da["t"]
                da["h_bar"] = (1 - 1 / (t + t0)) * da["h_bar"] + (self.target - acc) / (t + t0)
                x = da["mu"] - math.sqrt(t) / gamma * da["h_bar"]
                eta = t ** (-kappa)
                da["x_bar"] = eta * x + (1 - eta) * da["x_bar"]
                step = math.exp(x)
                if in_window(it):
                    w_n += 1
                    d = z.double() - w_mean
                    w_mean += d / w_n
                    w_m2 += d * (z.double() - w_mean)
                if (it + 1) in ends and w_n > 1:
                    var = w_m2 / (w_n - 1)
                    var = var * (w_n / (w_n + 5.0)) + 1e-3 * (5.0 / (w_n + 5.0))      # Stan's regularisation
                    self.inv_mass.copy_(var.to(z.dtype))
                    w_n = 0; w_mean.zero_(); w_m2.zero_()
                    torch.cuda.current_stream(self.dev).synchronize()
                    step = self.find_heuristic_step_size(math.exp(da["x_bar"]))
                    da = dict(mu=math.log(10.0 * step), x_bar=0.0, h_bar=0.0, t=0)
                if it + 1 == num_warmup:
                    step = math.exp(da["x_bar"]) if da["t"] > 0 else step
                    self.n_evals_warmup = self.n_evals
                # the adaptation arithmetic above ran on torch's stream, the sampler has its own: the next
                # transition must not overwrite z / read inv_mass before those kernels are done
                torch.cuda.current_stream(self.dev).synchronize()
            else:
                out["samples"].append(collect(z))
                out["accept"].append(acc); out["depth"].append(st.depth); out["diverging"].append(bool(st.diverging))
                out["potential"].append(st.potential); out["n_leapfrog"].append(st.num_leapfrog)
            if progress:
                progress(it, warm, step, st.depth, acc)
        out.update(step_size=step, n_evals=self.n_evals, n_evals_warmup=self.n_evals_warmup, z_last=z,
                   launches=self.launches, host_syncs=self.host_syncs)
        return out


def split_rhat(x):
    """Split-chain potential scale reduction of draws x[chain, draw] (Gelman et al.; numpyro.diagnostics.split_gelman_rubin)."""
    x = np.asarray(x, np.float64)
    if x.ndim == 1:
        x = x[None]
    half = x.shape[1] // 2
    x = np.concatenate([x[:, :half], x[:, half:2 * half]], axis=0)
    n = x.shape[1]
    w = x.var(axis=1, ddof=1).mean()
    b = n * x.mean(axis=1).var(ddof=1)
    return float(math.sqrt(((n - 1) / n * w + b / n) / w)) if w > 0 else float("nan")


def sample_model(model, z0_flat, num_warmup, num_samples, seed=0, group=None, rank=0, world=1, collect=None,
                 device_tree=True, star_offset=0, n_total=None, progress=None, **kw):
    """NUTS over a conditioned (and, for world > 1, sharded) model; returns the run dict with hyper-site draws.
    ``device_tree=True`` (default): the device-side sampler ``DeviceNUTS``; False: the host-recursive ``NUTS`` with
    one synchronisation per leapfrog (round 1's driver, kept as a cross-check)."""
    if device_tree:
        nuts = DeviceNUTS(model, z0_flat, seed=seed, star_offset=star_offset, n_total=n_total, **kw)
        return nuts.run(num_warmup, num_samples, collect=collect, progress=progress)
    dev = torch.device("cuda", model.device)
    pot = DevicePotential(model, group)
    space = VectorSpace(n_hyper=pot.nh, group=group, sharded=world > 1)
    nuts = NUTS(pot, space, seed=seed, rank=rank, **kw)
    z0 = torch.as_tensor(np.asarray(z0_flat, dtype=np.float32), device=dev)
    out = nuts.run(z0, num_warmup, num_samples, collect=collect, progress=progress)
    out["n_evals_warmup"] = getattr(nuts, "n_evals_warmup", 0)
    return out
