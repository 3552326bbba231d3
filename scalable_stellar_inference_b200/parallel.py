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


def block_bounds(lo: int, hi: int, n_blocks: int, unit: int, ramp: bool = False):
    """Cut stars [lo, hi) into <= n_blocks contiguous blocks whose sizes are whole multiples of ``unit`` (one full
    round of the persistent kernel: SMs x 2 tiles x 128 stars), remainder last, so that no block pays for a
    partly filled round.  ``ramp``: triangular sizes (1, 2, 3, ... 3, 2, 1 in proportion) instead of equal ones --
    the first H2D and the last D2H of a pipelined evaluation are exposed, so the end blocks are kept small and each
    block is at most ~2x its neighbour, which keeps the next copy inside the current kernel."""
    n = hi - lo
    units = max(1, -(-n // unit))
    nb = max(1, min(n_blocks, units))
    if ramp:
        weight = [min(b + 1, nb - b) for b in range(nb)]
        sizes = [1] * nb
        for _ in range(units - nb):
            b = min(range(nb), key=lambda k: (sizes[k] / weight[k], abs(k - (nb - 1) / 2)))
            sizes[b] += 1
    else:
        base, rem = divmod(units, nb)
        sizes = [base + (1 if b < rem else 0) for b in range(nb)]
    stars = [k * unit for k in sizes]
    if ramp:
        stars[max(range(nb), key=lambda k: sizes[k])] -= units * unit - n  # the partly filled round: a middle block
    out, pos = [], lo
    for b in range(nb):
        end = min(hi, pos + stars[b])
        if end > pos:
            out.append((pos, end))
        pos = end
    return out


def _parse_cpulist(text: str):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def _gpu_pci_bus_id(device: int):
    """PCI bus id ("0000:1b:00.0") of CUDA device ``device`` of this process, through torch (honours
    CUDA_VISIBLE_DEVICES) or, failing that, NVML."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device)
        return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        pass
    import os
    import pynvml
    pynvml.nvmlInit()
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    index = int(vis.split(",")[device]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else device
    bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
    bus = bus.decode() if isinstance(bus, bytes) else bus
    dom, rest = bus.lower().split(":", 1)
    return f"{dom[-4:]}:{rest}"


def gpu_local_cpus(device: int):
    """(cpus, how): the host cores that share a NUMA node / PCIe root with GPU ``device`` -- sysfs ``local_cpulist``
    of the GPU's PCI function, else its ``numa_node`` -> ``/sys/devices/system/node/nodeN/cpulist``, else
    ``nvidia-smi topo -C``'s answer; (None, reason) when the topology is not visible."""
    import os
    import subprocess
    try:
        bus = _gpu_pci_bus_id(device)
    except Exception as e:  # noqa: BLE001
        return None, f"no PCI bus id ({type(e).__name__})"
    base = f"/sys/bus/pci/devices/{bus}"
    try:
        with open(f"{base}/local_cpulist") as f:
            cpus = _parse_cpulist(f.read())
        if cpus:
            return cpus, f"sysfs local_cpulist of {bus}"
    except OSError:
        pass
    try:
        with open(f"{base}/numa_node") as f:
            node = int(f.read().strip())
        if node >= 0:
            with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
                return _parse_cpulist(f.read()), f"sysfs numa_node {node} of {bus}"
    except (OSError, ValueError):
        pass
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-C", "-i", str(device)], capture_output=True, text=True, timeout=20).stdout
        for line in out.splitlines():
            if "NUMA" in line.upper() and ":" in line:
                ids = line.split(":")[-1].strip()
                if ids and ids[0].isdigit():
                    node = int(ids.split(",")[0].split("-")[0])
                    with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
                        return _parse_cpulist(f.read()), f"nvidia-smi topo -C: NUMA node {node}"
    except Exception:  # noqa: BLE001
        pass
    return None, f"no topology in sysfs for {bus}"


def bind_host_to_device(device: int):
    """Restrict this process to the CPU cores next to GPU ``device`` (see ``gpu_local_cpus``), so that the pinned host
    buffers allocated afterwards are first-touched on that node and the launching thread runs there: the host-buffer
    path then does not cross the socket interconnect.  Returns a dict: ``local`` = the process now runs on the
    GPU's node (it was bound, or its affinity mask was inside the node already), ``bound`` = the mask was changed,
    ``how`` = where the topology came from / why nothing was done, ``previous`` = the old mask when it was changed
    (undo with ``os.sched_setaffinity(0, previous)``)."""
    import os
    cpus, how = gpu_local_cpus(device)
    if cpus is None:
        return dict(local=False, bound=False, how=how, previous=None)
    try:
        previous = os.sched_getaffinity(0)
        target = cpus & previous
        if not target:
            return dict(local=False, bound=False, how=how + "; none of those cores is available to this process", previous=None)
        if target == previous:
            return dict(local=True, bound=False, how=how + "; the process was confined to that node already", previous=None)
        os.sched_setaffinity(0, target)
        return dict(local=True, bound=True, how=how, previous=previous, cores=len(target))
    except (AttributeError, OSError) as e:
        return dict(local=False, bound=False, how=how + f"; sched_setaffinity: {e}", previous=None)


def slice_problem(z: dict, obs, diag, mh_loc, mh_scale, lo: int, hi: int):
    """Stars [lo, hi) of a full problem; hyper sites are replicated."""
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


def setup_peer_exchange(model, group=None):
    """Map every rank's exchange buffer into this process (PyTorch symmetric memory over NVLink peer access) and
    hand the pointers to libcsi: the 6-double allreduce is then done by the finalising block of the potential
    kernel itself (csi_set_peer_exchange) -- no NCCL call per evaluation.  Returns an object that must be kept
    alive, or None when symmetric memory is unavailable (the caller keeps the NCCL allreduce)."""
    import sys
    import torch
    import torch.distributed as dist
    try:
        import torch.distributed._symmetric_memory as symm_mem
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if world < 2 or world > 8:
            return None
        dev = torch.device("cuda", model.device)
        from ._lib import PEER_SLOT_DOUBLES
        n = 2 * world * PEER_SLOT_DOUBLES          # CSI_PEER_BUF_DOUBLES(world), include/csi.h
        buf = symm_mem.empty(n, dtype=torch.float64, device=dev)
        buf.zero_()
        hdl = symm_mem.rendezvous(buf, group if group is not None else dist.group.WORLD)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        torch.cuda.synchronize(dev)
        dist.barrier(group)                      # every buffer is zeroed before anybody's kernel writes into it
        model.set_peer_exchange(ptrs)
        return dict(buf=buf, handle=hdl, ptrs=ptrs, world=world, rank=rank)
    except Exception as e:  # noqa: BLE001 -- any failure here just means "use NCCL"
        print(f"[parallel] fused peer exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
        try:
            model.set_peer_exchange(None)
        except Exception:  # noqa: BLE001
            pass
        return None


class ShardedPotential:
    """The repo's public evaluation object for 1..8 GPUs.

    ``model`` must already be sharded (``set_shard``) and conditioned on the local block.
    ``value_and_grad_device`` is stream-ordered on torch's current stream and never synchronises the host;
    ``value_and_grad_host`` is the end-to-end call: pinned host z -> device, evaluate, allreduce, device -> host.
    """

    def __init__(self, model, group=None, fused=False):
        import torch
        self.torch = torch
        self.model, self.group = model, group
        self.device = torch.device("cuda", model.device)
        # fused=True: the kernel itself exchanges the 6 sums over peer memory (setup_peer_exchange)
        self.exchange = setup_peer_exchange(model, group) if fused else None
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
        if self.exchange is None:
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


class PipelinedHostPotential:
    """End-to-end evaluation through HOST buffers with the copies hidden behind the kernel.

    The local stars are cut into ``len(models)`` contiguous blocks, each with its own libcsi handle, device
    buffers, pinned host buffers and CUDA stream.  One evaluation issues, per block, H2D of the block's z -> the
    block's kernel -> D2H of its gradient and 6 sums; the copy engines and the SMs serve different blocks at the same
    time, so a step costs ~max(PCIe time, kernel time) instead of their sum.  The host-side layout is
    block-major -- ``z_views[b][site]`` / ``grad_views[b][site]`` are numpy views into the pinned buffers that a
    host-side sampler reads and writes in place (the hyper sites are replicated into every block by
    ``set_hyper``).  ``models[b]`` must be conditioned on block b and sharded as ``set_shard(rank * B + b, world * B)``
    so that the hyper-prior terms are counted once.
    """

    def __init__(self, models, group=None, graph=True):
        import torch
        self.torch, self.models, self.group = torch, list(models), group
        # graph: the whole step (every block's copies and kernel, on the blocks' streams) is captured once into a
        # CUDA graph and replayed with ONE host call per evaluation -- 4 x blocks Python-issued operations per step
        # otherwise, whose jitter on a busy host shows up directly in the end-to-end rate.
        self.use_graph, self.graph, self.graph_error = bool(graph), None, None
        self.device = torch.device("cuda", models[0].device)
        self.n_hyper = len(models[0].hyper_sites)
        self.blocks = []
        self.h2d_bytes = self.d2h_bytes = 0
        for m in self.models:
            n = m.flat_size
            b = dict(model=m, stream=torch.cuda.Stream(self.device),
                     z_pin=torch.zeros(n, dtype=torch.float32).pin_memory(),
                     grad_pin=torch.zeros(n, dtype=torch.float32).pin_memory(),
                     out_pin=torch.zeros(8, dtype=torch.float64).pin_memory(),
                     z_dev=torch.zeros(n, dtype=torch.float32, device=self.device),
                     grad_dev=torch.zeros(n, dtype=torch.float32, device=self.device),
                     out_dev=torch.zeros(8, dtype=torch.float64, device=self.device))
            self.blocks.append(b)
            self.h2d_bytes += n * 4
            self.d2h_bytes += n * 4 + 6 * 8
        self.z_views = [m.unflatten(b["z_pin"].numpy()) for m, b in zip(self.models, self.blocks)]
        self.grad_views = [m.unflatten(b["grad_pin"].numpy()) for m, b in zip(self.models, self.blocks)]
        self.launches = 0
        self._sum = torch.zeros(6, dtype=torch.float64, device=self.device)

    def set_hyper(self, values):
        for b in self.blocks:
            b["z_pin"][: self.n_hyper] = self.torch.as_tensor(np.asarray(values, np.float32))

    def load_blocks(self, z_blocks):
        """z_blocks[b]: flat float32 vector of block b in the block model's ``flatten`` order."""
        for b, z in zip(self.blocks, z_blocks):
            b["z_pin"].copy_(self.torch.as_tensor(np.asarray(z, np.float32)))

    def _issue(self, fork=None):
        """One evaluation of every block on its own stream; ``fork``: the capturing stream the block streams branch
        from and join back into (graph capture), None = plain asynchronous issue."""
        torch = self.torch
        for b in self.blocks:
            if fork is not None:
                b["stream"].wait_stream(fork)
            with torch.cuda.stream(b["stream"]):
                b["z_dev"].copy_(b["z_pin"], non_blocking=True)
                b["model"].potential_device(b["z_dev"].data_ptr(), b["out_dev"].data_ptr(), b["grad_dev"].data_ptr(),
                                            b["stream"].cuda_stream)
                b["grad_pin"].copy_(b["grad_dev"], non_blocking=True)
                b["out_pin"].copy_(b["out_dev"], non_blocking=True)
        if fork is not None:  # join only after every block has branched off, or the blocks would form a chain
            for b in self.blocks:
                fork.wait_stream(b["stream"])

    def _capture(self):
        torch = self.torch
        try:
            self._issue()                       # warm: first-use initialisation must not happen under capture
            torch.cuda.synchronize(self.device)
            self._graph_stream = torch.cuda.Stream(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self._graph_stream):
                self._issue(fork=torch.cuda.current_stream(self.device))
            self.graph = g
        except Exception as e:  # capture is an optimisation of the issue path, not of the arithmetic
            self.graph, self.graph_error, self.use_graph = None, repr(e), False
            torch.cuda.synchronize(self.device)

    def value_and_grad_host(self):
        """Returns (U, dU/dhyper[5], grad_views): per-star gradients are left in the pinned block buffers."""
        torch = self.torch
        if self.use_graph and self.graph is None:
            self._capture()
        if self.graph is not None:
            with torch.cuda.stream(self._graph_stream):
                self.graph.replay()
            self._graph_stream.synchronize()
        else:
            self._issue()
            for b in self.blocks:
                b["stream"].synchronize()
        self.launches += len(self.blocks)
        tot = np.sum([b["out_pin"][:6].numpy() for b in self.blocks], axis=0)
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            self._sum.copy_(torch.as_tensor(tot))
            dist.all_reduce(self._sum, group=self.group)
            tot = self._sum.cpu().numpy()
        return float(tot[0]), tot[1:6].copy(), self.grad_views


class ZeroCopyHostPotential:
    """End-to-end evaluation through HOST buffers with no copy calls at all: the kernel reads the sites straight
    from pinned host memory and writes the gradient and the 6 sums straight back into it.

    Pinned allocations are device-accessible (unified virtual addressing); ``csi_potential_host`` -- the C ABI's
    host-buffer call -- recognises page-locked buffers and launches the kernel on them.  The tcgen05 kernel fetches a star's sites with asynchronous
    copies into shared memory half a tile (~15 us) before it needs them and its gradient stores are fire-and-
    forget, so the PCIe round trips hide under the MMA stages: one launch per evaluation, every input byte crosses
    the bus once as a coalesced 128-byte read and every gradient byte once as a posted write, and the step costs
    ~max(kernel, bytes / PCIe rate) with no pipeline fill or drain.  ``z_view[site]`` / ``grad_view[site]`` are
    numpy views of the pinned buffers for a host-side sampler.  Observations stay resident in HBM (csi_set_data).
    """

    def __init__(self, model, group=None):
        import torch
        self.torch, self.model, self.group = torch, model, group
        self.device = torch.device("cuda", model.device)
        n = model.flat_size
        self.z_pin = torch.zeros(n, dtype=torch.float32).pin_memory()
        self.grad_pin = torch.zeros(n, dtype=torch.float32).pin_memory()
        self.out = np.zeros(8, np.float64)
        self.z_view = model.unflatten(self.z_pin.numpy())
        self.grad_view = model.unflatten(self.grad_pin.numpy())
        self.h2d_bytes, self.d2h_bytes = n * 4, n * 4 + 6 * 8
        self.launches = 0
        self._sum = torch.zeros(6, dtype=torch.float64, device=self.device)
        self._out_pin = None

    def load(self, z_flat):
        self.z_pin.copy_(self.torch.as_tensor(np.asarray(z_flat, np.float32)))

    def enqueue(self, stream=None):
        """Queue one evaluation on ``stream`` (default: torch's current stream) and return at once: ``csi_potential``
        on the page-locked buffers (with unified addressing a cudaHostAlloc pointer is its own device alias).  The
        kernel reads the sites from host memory and writes the gradient and the 6 sums back to it; call ``wait()``
        before touching them.  Needs the fused peer exchange when sharded (an NCCL allreduce of host memory would
        need a staging copy)."""
        torch = self.torch
        if self._out_pin is None:
            self._out_pin = torch.zeros(8, dtype=torch.float64).pin_memory()
        st = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        self.model.potential_device(self.z_pin.data_ptr(), self._out_pin.data_ptr(), self.grad_pin.data_ptr(), st)
        self.launches += 1

    def wait(self, stream=None):
        (stream or self.torch.cuda.current_stream(self.device)).synchronize()
        tot = self._out_pin[:6].numpy().copy()
        return float(tot[0]), tot[1:6], self.grad_view

    def value_and_grad_host(self):
        """Returns (U, dU/dhyper[5], grad_view); the per-star gradients are in the pinned buffer when it returns."""
        self.model.potential_host(self.z_pin.data_ptr(), self.out.ctypes.data, self.grad_pin.data_ptr())
        self.launches += 1
        tot = self.out[:6].copy()
        import torch.distributed as dist
        fused = getattr(self.model, "_peer_exchange_on", False)
        if not fused and dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            self._sum.copy_(self.torch.as_tensor(tot))
            dist.all_reduce(self._sum, group=self.group)
            tot = self._sum.cpu().numpy()
        return float(tot[0]), tot[1:6].copy(), self.grad_view


class InterleavedHostPotential:
    """End-to-end evaluation through HOST buffers for SEVERAL independent chains: chain c's host -> device copy, its
    kernel and its device -> host copy run on three streams, so that while the SMs evaluate one chain the copy engines
    bring the next chain's sites in and take the previous chain's gradient out.  Every step still moves its own inputs
    from page-locked host memory and its own results back (h2d_bytes / d2h_bytes per step), but a step costs
    ~max(kernel, copies) instead of the kernel slowed down by its own PCIe reads (``ZeroCopyHostPotential``: one chain,
    the kernel reads host memory itself, 0.2 ms slower at 1 M stars).  Steps of ONE chain depend on each other, steps
    of different chains do not -- the reference itself runs 4-10 chains (scripts/model_star.py:78) -- so a host-side
    sampler with >= 2 chains per GPU keeps the device busy this way.  One libcsi handle serves all chains (same
    observations); ``z_pin[c]`` / ``grad_pin[c]`` / ``out_pin[c]`` are the chains' page-locked buffers.
    Needs the in-kernel peer exchange when sharded."""

    def __init__(self, model, chains=2):
        import torch
        self.torch, self.model, self.chains = torch, model, chains
        self.device = torch.device("cuda", model.device)
        n = model.flat_size
        mk = lambda shape, dt, pin: (torch.zeros(shape, dtype=dt).pin_memory() if pin  # noqa: E731
                                     else torch.zeros(shape, dtype=dt, device=self.device))
        self.z_pin = [mk(n, torch.float32, True) for _ in range(chains)]
        self.grad_pin = [mk(n, torch.float32, True) for _ in range(chains)]
        self.out_pin = [mk(8, torch.float64, True) for _ in range(chains)]
        self.z_dev = [mk(n, torch.float32, False) for _ in range(chains)]
        self.grad_dev = [mk(n, torch.float32, False) for _ in range(chains)]
        self.out_dev = [mk(8, torch.float64, False) for _ in range(chains)]
        self.s_in, self.s_k, self.s_out = (torch.cuda.Stream(self.device) for _ in range(3))
        self.ev_in = [torch.cuda.Event() for _ in range(chains)]
        self.ev_k = [torch.cuda.Event() for _ in range(chains)]
        self.ev_out = [torch.cuda.Event() for _ in range(chains)]
        self._used = [False] * chains
        self.h2d_bytes, self.d2h_bytes = n * 4, n * 4 + 6 * 8
        self.launches = 0

    def load(self, c, z_flat):
        self.z_pin[c].copy_(self.torch.as_tensor(np.asarray(z_flat, np.float32)))

    def submit(self, c):
        """Queue one evaluation of chain c (its current ``z_pin[c]``); returns at once."""
        torch = self.torch
        with torch.cuda.stream(self.s_in):
            if self._used[c]:
                self.s_in.wait_event(self.ev_k[c])       # the previous kernel of this chain has read z_dev[c]
            self.z_dev[c].copy_(self.z_pin[c], non_blocking=True)
            self.ev_in[c].record(self.s_in)
        with torch.cuda.stream(self.s_k):
            self.s_k.wait_event(self.ev_in[c])
            if self._used[c]:
                self.s_k.wait_event(self.ev_out[c])      # grad_dev[c] / out_dev[c] have been copied out
            self.model.potential_device(self.z_dev[c].data_ptr(), self.out_dev[c].data_ptr(), self.grad_dev[c].data_ptr(),
                                        self.s_k.cuda_stream)
            self.ev_k[c].record(self.s_k)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_k[c])
            self.grad_pin[c].copy_(self.grad_dev[c], non_blocking=True)
            self.out_pin[c].copy_(self.out_dev[c], non_blocking=True)
            self.ev_out[c].record(self.s_out)
        self._used[c] = True
        self.launches += 1

    def result(self, c):
        """Wait for chain c's last evaluation: (U, dU/dhyper[5], gradient in ``grad_pin[c]``)."""
        self.ev_out[c].synchronize()
        o = self.out_pin[c].numpy()
        return float(o[0]), o[1:6].copy(), self.grad_pin[c]
