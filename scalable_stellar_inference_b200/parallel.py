"""Star-axis data parallelism: one process per GPU, each owning a contiguous block of stars.

Stars are conditionally independent given the 5 hyper-parameters (src/celestify/numpyro_models.py:341-346:
everything per-star sits inside ``plate("star")`` / ``vmap``), so an evaluation is: every rank evaluates its
block with libcsi, then ONE sum-allreduce of 6 doubles ``[U, dU/dhyper]`` over NCCL (gloo on CPU for the tests of
the host logic).  Per-star gradients stay on the owning rank.  The reference has no equivalent: it only runs
independent chains per device (scripts/model_star.py:82-86).

PyTorch is plumbing here: device buffers, pinned host buffers, streams and the process group.
"""
from __future__ import annotations

import numpy as np


def shard_bounds(n: int, rank: int, world: int):
    """Contiguous block [lo, hi) of rank `rank`; blocks differ by at most one star."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_problem(z: dict, obs, diag, mh_loc, mh_scale, rank: int, world: int):
    """Slice a full problem (site dict + data arrays) down to the rank's block; hyper sites are replicated."""
    n = np.asarray(obs).shape[0]
    lo, hi = shard_bounds(n, rank, world)
    sl = slice(lo, hi)
    zl = {k: (v if np.ndim(v) == 0 else np.asarray(v)[sl]) for k, v in z.items()}
    f = lambda a: None if a is None else (a if np.ndim(a) == 0 else np.asarray(a)[sl])  # noqa: E731
    return zl, np.asarray(obs)[sl], f(diag), f(mh_loc), f(mh_scale)


def allreduce_out(out, group=None):
    """SUM-allreduce of the 6-double vector [U, dU/dhyper]; a no-op without an initialised process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


class ShardedPotential:
    """The repo's public evaluation object for 1..8 GPUs.

    ``model`` must already be sharded (``set_shard``) and conditioned on the local block.
    ``value_and_grad_device`` is stream-ordered on torch's current stream and never synchronises the host;
    ``value_and_grad_host`` is the end-to-end call: pinned host z -> device, evaluate, allreduce, device -> host.
    """

    def __init__(self, model, group=None):
        import torch
        self.torch = torch
        self.model, self.group = model, group
        self.device = torch.device("cuda", model.device)
        n = model.flat_size
        self.n_hyper = len(model.hyper_sites)
        self.z_dev = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.grad_dev = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.out_dev = torch.zeros(8, dtype=torch.float64, device=self.device)
        self.z_pin = torch.zeros(n, dtype=torch.float32).pin_memory()
        self.grad_pin = torch.zeros(n, dtype=torch.float32).pin_memory()
        self.out_pin = torch.zeros(8, dtype=torch.float64).pin_memory()
        self.h2d_bytes = n * 4
        self.d2h_bytes = n * 4 + 6 * 8
        self.launches = 0

    def value_and_grad_device(self, z_dev=None):
        torch = self.torch
        z = self.z_dev if z_dev is None else z_dev
        stream = torch.cuda.current_stream(self.device).cuda_stream
        self.model.potential_device(z.data_ptr(), self.out_dev.data_ptr(), self.grad_dev.data_ptr(), stream)
        self.launches += 1
        allreduce_out(self.out_dev[:6], self.group)
        return self.out_dev, self.grad_dev

    def value_and_grad_host(self, z_host=None):
        torch = self.torch
        if z_host is not None:
            self.z_pin.copy_(torch.as_tensor(z_host))
        self.z_dev.copy_(self.z_pin, non_blocking=True)
        self.value_and_grad_device()
        self.grad_pin.copy_(self.grad_dev, non_blocking=True)
        self.out_pin.copy_(self.out_dev, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return float(self.out_pin[0]), self.out_pin[1:6].numpy(), self.grad_pin.numpy()
