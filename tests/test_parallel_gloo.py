"""Star-axis sharding over 2 processes with the gloo backend (CPU): the host-side logic of
scalable_stellar_inference_b200.parallel -- block bounds, problem slicing, the single 6-double SUM-allreduce and the
"hyper-prior on rank 0 only" rule (csi_set_shard) -- reproduces the unsharded oracle potential and hyper-gradient.
The per-shard numbers come from the oracle here (the CUDA library needs a GPU; see test_gpu_parity.py::
test_shard_sum_invariance for the same identity through libcsi)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity_util as PU
from oracle import model as M
from scalable_stellar_inference_b200.parallel import allreduce_out, shard_bounds, shard_problem

NAME, N = "hier_diag_seis", 301


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        spec, (z0, obs, diag, ml, ms) = PU.make_problem(NAME, N)
        zl, ol, dl, mll, msl = shard_problem(z0, obs, diag, ml, ms, rank, world)
        lo, hi = shard_bounds(N, rank, world)
        sub = PU.oracle_spec(NAME, hi - lo)
        u, g, terms = M.potential_and_grad(sub, zl, ol, dl, mll, msl, return_terms=True)
        # what libcsi returns on this rank: hyper-prior terms only on rank 0 (csi_set_shard)
        hyper = spec.hyper_sites()
        z_h = {k: z0[k] for k in hyper}
        prior_u, prior_g = _hyper_prior_only(spec, z0)
        out = np.array([u] + [g[k] for k in hyper], np.float64)
        if rank != 0:
            out -= np.array([prior_u] + [prior_g[k] for k in hyper])
        t = torch.from_numpy(out.copy())
        allreduce_out(t)
        ret[rank] = (t.numpy().copy(), {k: np.asarray(g[k]) for k in spec.star_sites()}, (lo, hi))
    finally:
        dist.destroy_process_group()


def _hyper_prior_only(spec, z0):
    """U and dU/dhyper of the five hyper-prior terms alone: difference of two evaluations that share the stars."""
    n = 2
    z = {k: (v if np.ndim(v) == 0 else np.asarray(v)[:n]) for k, v in z0.items()}
    sub = PU.oracle_spec(NAME, n)
    obs = np.zeros((n, len(sub.obs_idx)))
    u1, g1, t1 = M.potential_and_grad(sub, z, obs, 1.0, return_terms=True)
    u_star = -float(np.sum(t1["per_star"]))
    prior_u = u1 - u_star
    # hyper-gradient of the star terms alone by central differences is avoided: the prior gradient is analytic
    c = sub.const
    from oracle import dists as D
    g = {}
    g["Y_0"] = (z["Y_0"] - c["Y_0"]["loc"]) / c["Y_0"]["scale"] ** 2
    _, _, _, dl = M._tn_site(np.float64(z["dY_dZ"]), c["dY_dZ"])
    g["dY_dZ"] = -dl
    for k, ck in (("precision_Y", c["precision_Y"]), ("precision_a", c["precision_a"])):
        p = np.exp(z[k])
        g[k] = -(((ck["concentration"] - 1) / p - ck["rate"]) * p + 1.0)
    _, _, _, dl = M._tn_site(np.float64(z["mu_a"]), c["mu_a"])
    g["mu_a"] = -dl
    return prior_u, g


def test_shard_bounds_cover_everything():
    for n in (1, 2, 7, 128, 1001):
        for w in (1, 2, 4, 8):
            b = [shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


@pytest.mark.timeout(300)
def test_two_rank_allreduce_reproduces_unsharded_potential():
    world, port = 2, _free_port()
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(NAME, N)
    u, g = M.potential_and_grad(spec, z0, obs, diag, ml, ms)
    ref = np.array([u] + [g[k] for k in spec.hyper_sites()])
    for r in range(world):
        np.testing.assert_allclose(ret[r][0], ref, rtol=1e-10, atol=1e-8)
    for k in spec.star_sites():
        full = np.concatenate([ret[r][1][k] for r in range(world)])
        np.testing.assert_allclose(full, g[k], rtol=1e-12, atol=1e-12)


def test_block_bounds_are_round_aligned_and_cover_the_range():
    from scalable_stellar_inference_b200.parallel import block_bounds
    unit = 2 * 128 * 148
    for lo, hi, nb in ((0, 1_000_000, 6), (500_000, 1_000_000, 4), (0, 700, 6), (3, 3 + unit, 5), (0, 10 * unit, 3)):
        b = block_bounds(lo, hi, nb, unit)
        assert b[0][0] == lo and b[-1][1] == hi and len(b) <= nb
        assert all(x[1] == y[0] for x, y in zip(b, b[1:]))
        assert all((e - s) % unit == 0 for s, e in b[:-1])          # only the last block may be partial


def test_ramped_block_bounds_cover_the_range_with_small_end_blocks():
    from scalable_stellar_inference_b200.parallel import block_bounds
    unit = 2 * 128 * 148
    for lo, hi, nb in ((0, 1_000_000, 6), (0, 1_000_000, 8), (500_000, 1_000_000, 4), (0, 700, 6), (3, 3 + unit, 5)):
        b = block_bounds(lo, hi, nb, unit, ramp=True)
        assert b[0][0] == lo and b[-1][1] == hi and len(b) <= nb
        assert all(x[1] == y[0] and x[1] > x[0] for x, y in zip(b, b[1:]))
        sizes = [e - s for s, e in b]
        assert sum(1 for k in sizes if k % unit) <= 1                 # one partly filled round at most
        if len(b) >= 4:
            assert sizes[0] < max(sizes) and sizes[-1] < max(sizes)   # the exposed copies belong to small blocks
