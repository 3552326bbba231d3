#!/usr/bin/env python
"""bench.py -- logp+grad evaluations/s of celestify's star models on the B200 path, one BASELINE.json configuration
per run (default: config 4, the one the metric is quoted on: hierarchical spec+phot, 1 M synthetic stars, BC table,
star axis sharded over N GPUs with one 6-double exchange per evaluation).

    python bench.py --gpus 1 --steps 20 --warmup 3                 # config 4
    python bench.py --config c2|c3|c4|c5 ...                       # the other BASELINE configurations
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 20 --warmup 3
    python bench.py --impl reference ...      # CPU restatement of the reference on the host cores

One "step" = one evaluation of U(z) and dU/dz for ALL stars (the thing numpyro's NUTS calls per leapfrog).
value = evaluations/s with z resident in HBM; e2e = the same from / to page-locked HOST buffers through the C ABI
(csi_potential on the buffers' device aliases: the kernel reads the sites and writes the gradient over PCIe itself),
K steps queued back to back and one host synchronisation at the end, as a device-side sampler drives it;
e2e.sync_per_step is the same with a host synchronisation after every step (csi_potential_host).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BANDS = ["G", "BP", "RP"]
FLOP_PER_STAR = 332_288          # emulator fwd + input-VJP (SURVEY 8d / BASELINE.md section 3)
HIDDEN_FLOP_PER_STAR = 327_680   # the 10 tensor-core contractions
MMA_PASSES = 3

# BASELINE.json configs[1..4] (configs[0] is the reference's single-star CPU smoke case: a parity test, not a bench line)
CONFIGS = {
    "c2": dict(level="multi", obs=["mag", "plx"], phot=True, stars=1_000,
               text="BASELINE config 2: MultiStarModel photometric, observables [mag(G,BP,RP), plx], kind=diag, BC table "
                    "70x26x18x13 (synthetic MIST stand-in)"),
    "c3": dict(level="hierarchical", obs=["log_Teff", "log_g"], phot=False, stars=10_000,
               text="BASELINE config 3: HierarchicalStarModel spectroscopic, observables [log_Teff, log_g], kind=diag"),
    "c4": dict(level="hierarchical", obs=["log_Teff", "log_g", "mag", "plx"], phot=True, stars=1_000_000,
               text="BASELINE config 4: HierarchicalStarModel spec+phot, observables [log_Teff, log_g, mag(G,BP,RP), plx], "
                    "kind=diag, BC table 70x26x18x13 (synthetic MIST stand-in)"),
    "c5": dict(level="hierarchical", obs=["log_Teff", "log_L", "log_Dnu", "log_numax"], phot=False, stars=100_000,
               text="BASELINE config 5: HierarchicalStarModel with asteroseismic likelihood, observables [log_Teff, log_L, "
                    "log_Dnu, log_numax], kind=diag"),
    # not a BASELINE configuration: the reference's only recorded hierarchical run (notebook/hierarchical-model-analysis.ipynb
    # cell [5]: 1000 stars, 4 chains, 16000 warm-up + 2000 samples, sampling 1970 s on 4 x A100) -- tools/nuts_bench.py
    "r1k": dict(level="hierarchical", obs=["log_Teff", "log_L", "log_Dnu"], phot=False, stars=1_000,
                text="the reference's recorded run: HierarchicalStarModel, 1000 stars, observables [log_Teff, log_L, log_Dnu], "
                     "kind=diag (notebook/hierarchical-model-analysis.ipynb cell [5])"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--stars", type=int, default=0, help="override the configuration's star count")
    ap.add_argument("--engine", default="tcgen05", choices=["tcgen05", "ffma"])
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="seconds of CPU work for the in-line cpu_baseline")
    ap.add_argument("--ref-budget", type=float, default=240.0, help="seconds the --impl reference run may take")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--nccl-allreduce", action="store_true", help="keep the NCCL allreduce instead of the in-kernel exchange")
    ap.add_argument("--e2e-blocks", type=int, default=6, help="star blocks of the pipelined host-buffer path")
    ap.add_argument("--e2e-mode", default="zerocopy", choices=["zerocopy", "pipelined"],
                    help="host-buffer path: the kernel reads / writes pinned host memory itself (one launch), or "
                         "star blocks with explicit H2D / D2H copies on streams")
    ap.add_argument("--e2e-chains", type=int, default=2, help="independent chains interleaved in the end-to-end measurement (1 = one chain, zero copy)")
    ap.add_argument("--e2e-equal", action="store_true", help="equal star blocks instead of the triangular ramp")
    ap.add_argument("--e2e-no-graph", action="store_true", help="issue the host-buffer path per call, no CUDA graph")
    a = ap.parse_args()
    a.cfg = CONFIGS[a.config]
    if a.stars <= 0:
        a.stars = a.cfg["stars"]
    return a


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops", 1590.0), d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json, burst)"
    return 1590.0, 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of THIS command
    (profiles/potential_ncu.json, written by tools/ncu_to_json.py); None when no capture is committed."""
    p = os.path.join(ROOT, "profiles", "potential_ncu.json")
    if not os.path.exists(p):
        return None, None
    d = json.load(open(p))
    return d.get("dram_read_bytes", 0) + d.get("dram_write_bytes", 0), d


def git_sha(path):
    try:
        return subprocess.run(["git", "log", "-1", "--format=%h", "--", path], cwd=ROOT, capture_output=True,
                              text=True).stdout.strip() or None
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([t.strip() for t in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        busy = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# CPU legs (the ONLY places that touch oracle/)
# ---------------------------------------------------------------------------------------------------------------
def host_threads():
    """Every host core this process may use -- set explicitly (torch.distributed.run exports OMP_NUM_THREADS=1,
    which would otherwise single-thread BLAS and torch without anybody noticing)."""
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count() or 1
    import torch
    torch.set_num_threads(cores)
    blas = None
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=cores)
        from threadpoolctl import threadpool_info
        blas = max((d.get("num_threads", 1) for d in threadpool_info()), default=None)
    except Exception:
        pass
    return cores, torch.get_num_threads(), blas


class CpuReference:
    """The reference's CPU path for one configuration, as two restatements timed side by side (BASELINE.md section 2;
    jax / numpyro are not installed, so the reference's own jit(value_and_grad(potential_fn)) cannot run):
    (i) oracle/model.py: numpy, fp32 emulator, hand-written reverse mode, BLAS threads = all cores;
    (ii) oracle/torch_ref.py: torch-CPU fp32 autograd value_and_grad, torch threads = all cores.
    The stars are evaluated in chunks of <= 50 000 (memory); a step over S stars is the sum of its chunks."""
    CHUNK = 50_000

    def __init__(self, cfg, n_stars):
        import torch
        from oracle import model as M, synth
        from oracle.bc import GridInterpolator, synthetic_grid
        from oracle.nets import load_emulator
        from oracle.torch_ref import TorchPotential
        self.M, self.torch = M, torch
        self.chunks = []
        left = n_stars
        grid = GridInterpolator(*synthetic_grid()) if cfg["phot"] else None
        emu = load_emulator(np.float32)
        while left > 0:
            m = min(left, self.CHUNK)
            if m == 1 and n_stars > 1:
                m, left = 2, left + 1
            spec = M.ModelSpec(cfg["level"], m, list(cfg["obs"]), bands=3 if cfg["phot"] else 0, bc=grid,
                               bc_mode="grid" if cfg["phot"] else "none", emulator=emu)
            prob = synth.make_problem(spec, seed=len(self.chunks))
            tp = TorchPotential(spec, prob[1], prob[2], prob[3], prob[4], dtype=torch.float32)
            self.chunks.append((spec, prob, tp))
            left -= m
        self.n = n_stars

    def step(self, impl):
        u = 0.0
        for spec, (z0, obs, diag, ml, ms), tp in self.chunks:
            u += tp.value_and_grad(z0)[0] if impl == "torch" else self.M.potential_and_grad(spec, z0, obs, diag, ml, ms)[0]
        return u

    def time(self, impl, steps, warmup):
        for _ in range(warmup):
            self.step(impl)
        t = []
        for _ in range(steps):
            t0 = time.perf_counter()
            self.step(impl)
            t.append(time.perf_counter() - t0)
        return t


def cpu_reference(cfg, n_total, budget_s, steps=None, warmup=1):
    """Times both CPU restatements on the FULL workload when `steps` + `warmup` calls of it fit into budget_s (per
    implementation), else on the largest star sample that does (then the full-workload rate is the measured rate
    scaled by stars, cost being linear in stars, and both are reported).  Returns the dict for `cpu_baseline`."""
    cores, torch_threads, blas_threads = host_threads()
    probe_n = min(n_total, 20_000)
    probe = CpuReference(cfg, probe_n)
    per_star = {}
    for impl in ("torch", "numpy"):
        probe.step(impl)
        t0 = time.perf_counter()
        probe.step(impl)
        per_star[impl] = (time.perf_counter() - t0) / probe_n
    calls = (steps if steps is not None else 3) + warmup
    res = {}
    for impl in ("torch", "numpy"):
        fit = int(budget_s / 2 / calls / per_star[impl])
        sample = n_total if fit >= n_total else max(min(probe_n, n_total), fit // 1000 * 1000)
        ref = probe if sample == probe_n else CpuReference(cfg, sample)
        k = steps if steps is not None else max(2, min(20, int(budget_s / 2 / (per_star[impl] * sample)) - warmup))
        t = ref.time(impl, k, warmup)
        med = float(np.median(t))
        res[impl] = dict(sample_stars=sample, calls=len(t), ms_per_call=med * 1e3, measured_evals_per_s=1.0 / med,
                         full_workload_evals_per_s=1.0 / (med * n_total / sample), total_s=float(np.sum(t)))
    best = max(res, key=lambda k: res[k]["full_workload_evals_per_s"])
    r = res[best]
    full = r["sample_stars"] == n_total
    names = {"torch": "oracle/torch_ref.py (torch-CPU fp32 autograd value_and_grad)",
             "numpy": "oracle/model.py (numpy, fp32 emulator, manual reverse mode)"}
    desc = (f"{names[best]}: {'the full workload' if full else str(r['sample_stars']) + ' of ' + str(n_total) + ' stars'} "
            f"per timed call, median of {r['calls']} calls = {r['ms_per_call']:.1f} ms"
            + ("" if full else ", scaled linearly in stars to the full workload"))
    return {"value": r["full_workload_evals_per_s"], "unit": "evals/s", "cores": cores, "kind": "port", "sample": desc,
            "threads": {"torch": torch_threads, "blas": blas_threads, "OMP_NUM_THREADS_env": os.environ.get("OMP_NUM_THREADS")},
            "implementation": best, "full_workload_timed": full, "measured": res,
            "ms_per_step": 1e3 / r["full_workload_evals_per_s"]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_reference(args.cfg, args.stars, args.ref_budget, steps=args.steps, warmup=max(args.warmup, 1))
    rate = cb["value"]
    line = {
        "impl": "reference", "metric": "logp+grad evals/s", "value": rate, "unit": "evals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / rate, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, "cpu"),
        "cpu_baseline": cb,
        "e2e": {"value": rate, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference JAX-CPU unavailable on this box (jax, numpyro, h5py not installed, no network): the CPU "
                "restatements of the same path are timed instead, the faster one is reported",
    }
    emit(line)


def workload_config(args, engine):
    cfg = args.cfg
    hier = cfg["level"] == "hierarchical"
    return {"workload": cfg["text"], "config": args.config,
            "stars": args.stars, "sites_per_star": (6 if hier else 8) + (1 if cfg["phot"] else 0),
            "hyper_sites": 5 if hier else 0, "engine": engine,
            "parallelism": f"star-sharded x{args.gpus}, one 6-double allreduce per evaluation"
                           + (" (fused into the kernel over NVLink peer memory)" if getattr(args, "fused", False) else
                              (" (NCCL)" if args.gpus > 1 else "")),
            "l2": "no-flush" if args.no_flush else "L2 flushed (256 MiB write) between timed evaluations"}


# ---------------------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import scalable_stellar_inference_b200 as ssi
    from scalable_stellar_inference_b200 import synthetic
    from scalable_stellar_inference_b200.parallel import ShardedPotential, shard_problem

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    cfg = args.cfg
    grid = synthetic.synthetic_bc_grid(bands=3) if cfg["phot"] else None

    def make_model():
        kw = dict(engine=args.engine, device=local)
        if cfg["phot"]:
            kw.update(bands=BANDS, bc="grid", bc_grid=grid)
        cls = ssi.HierarchicalStarModel if cfg["level"] == "hierarchical" else ssi.MultiStarModel
        return cls(args.stars, list(cfg["obs"]), **kw)

    model = make_model()
    z0, obs, diag, ml, ms, _ = synthetic.make_problem(model, args.stars)
    model.set_shard(rank, world)
    zl, obs_l, diag_l, ml_l, ms_l = shard_problem(z0, obs, diag, ml, ms, rank, world)
    model.condition(obs_l, diag_l, M_H_loc=ml_l, M_H_scale=ms_l)
    pot = ShardedPotential(model)
    u_nccl = None
    if world > 1 and not args.nccl_allreduce:
        # reference value through NCCL first, then switch the handle to the in-kernel exchange over peer memory
        pot.z_pin.copy_(torch.from_numpy(model.flatten(zl))); pot.z_dev.copy_(pot.z_pin)
        pot.value_and_grad_device(); torch.cuda.synchronize(dev)
        u_nccl = pot.out_dev[:6].clone()
        from scalable_stellar_inference_b200.parallel import setup_peer_exchange
        pot.exchange = setup_peer_exchange(model)
    fused = pot.exchange is not None
    args.fused = fused
    pot.z_pin.copy_(torch.from_numpy(model.flatten(zl)))
    pot.z_dev.copy_(pot.z_pin)
    flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        pot.value_and_grad_device()
    torch.cuda.synchronize(dev)
    u_first = float(pot.out_dev[0].item())
    if fused and u_nccl is not None:
        if not torch.allclose(pot.out_dev[:6], u_nccl, rtol=1e-12, atol=0.0):
            raise SystemExit(f"fused exchange {pot.out_dev[:6].tolist()} differs from NCCL {u_nccl.tolist()}")
    if not np.isfinite(u_first) and not os.environ.get("CSI_TC_DEBUG"):
        raise SystemExit(f"non-finite potential {u_first}")

    # ---- device-resident timing: K evaluations, events on the launching stream --------------------------------
    K = args.steps
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    launches0 = pot.launches
    barrier()
    for k in range(K):
        if flush is not None:
            flush.fill_(k & 0xFF)
        ev[k][0].record()
        stream = torch.cuda.current_stream(dev).cuda_stream
        model.potential_device(pot.z_dev.data_ptr(), pot.out_dev.data_ptr(), pot.grad_dev.data_ptr(), stream)
        pot.launches += 1
        ev[k][1].record()
        if world > 1 and not fused:
            dist.all_reduce(pot.out_dev[:6])
        ev[k][2].record()
    barrier()
    gpu_launches = pot.launches - launches0
    kern_ms = np.array([ev[k][0].elapsed_time(ev[k][1]) for k in range(K)])
    step_ms = np.array([ev[k][0].elapsed_time(ev[k][2]) for k in range(K)])
    total = torch.tensor([float(step_ms.sum())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.MAX)
    total_ms = float(total.item())

    # ---- end to end: z in pinned host memory -> evaluate -> gradient and sums in host memory, every step --------
    # zerocopy (default): the kernel reads the sites from / writes the gradient to pinned host memory itself
    # (parallel.ZeroCopyHostPotential: one launch, no copy calls, PCIe traffic hidden by the kernel's input staging).
    # pipelined: the local stars are cut into blocks (own handle + stream each) so that the H2D of block b+1, the
    # kernel of block b and the D2H of block b-1 overlap (parallel.PipelinedHostPotential).
    from scalable_stellar_inference_b200.parallel import PipelinedHostPotential, ZeroCopyHostPotential
    from scalable_stellar_inference_b200.parallel import bind_host_to_device, block_bounds, shard_bounds, slice_problem
    lo_r, hi_r = shard_bounds(args.stars, rank, world)
    numa = bind_host_to_device(local)   # pinned buffers and the launching thread next to the GPU's PCIe root
    prev_affinity = numa.get("previous")
    numa = {k: v for k, v in numa.items() if k != "previous"}
    e2e_il_api = None
    if args.e2e_mode == "zerocopy":
        # the device-resident model of this rank (same handle, shard and exchange) fed from host memory
        pipe = ZeroCopyHostPotential(model)
        pipe.load(model.flatten(zl))
        B, bounds = 1, [(lo_r, hi_r)]
        e2e_api = ("csi_potential on page-locked host buffers, K steps queued, one synchronisation "
                   "(parallel.ZeroCopyHostPotential.enqueue / wait)")
        if args.e2e_chains > 1:
            e2e_il_api = (f"{args.e2e_chains} independent chains interleaved: per step H2D copy of the chain's sites, csi_potential, D2H copy "
                       "of gradient and sums on three streams; a chain's step waits for its previous result "
                       "(parallel.InterleavedHostPotential.submit / result)")
    else:
        bounds = block_bounds(lo_r, hi_r, max(1, args.e2e_blocks), unit=2 * 128 * model.info()["sm_count"],
                              ramp=not args.e2e_equal)
        B = len(bounds)
        blocks, zb = [], []
        for b, (lo_b, hi_b) in enumerate(bounds):
            mb = make_model()
            mb.set_shard(rank * B + b, world * B)   # hyper-prior terms on the first block of rank 0 only
            zl_b, obs_b, diag_b, ml_b, ms_b = slice_problem(z0, obs, diag, ml, ms, lo_b, hi_b)
            mb.condition(obs_b, diag_b, M_H_loc=ml_b, M_H_scale=ms_b)
            blocks.append(mb); zb.append(mb.flatten(zl_b))
        pipe = PipelinedHostPotential(blocks, graph=not args.e2e_no_graph)
        pipe.load_blocks(zb)
        e2e_api = "parallel.PipelinedHostPotential.value_and_grad_host"
    for _ in range(2):
        pipe.value_and_grad_host()
    barrier()
    t_e2e = time.perf_counter()
    for k in range(K):
        u_h, gh_h, g_h = pipe.value_and_grad_host()      # synchronises the host: wall clock brackets real work
    torch.cuda.synchronize(dev)
    e2e_sync_ms = (time.perf_counter() - t_e2e) * 1e3
    barrier()
    nh0 = len(model.hyper_sites)
    e2e_ms = e2e_sync_ms
    e2e_async = args.e2e_mode == "zerocopy" and (world == 1 or fused)
    if e2e_async:
        # the same K evaluations queued back to back (stream-ordered csi_potential on the pinned buffers, every step
        # reads its sites from and writes its gradient and sums to host memory) with ONE host synchronisation at the
        # end: how a device-side sampler drives the path -- no host-paced barrier between the ranks' steps
        pipe.enqueue(); pipe.wait()
        barrier()
        t_e2e = time.perf_counter()
        for k in range(K):
            pipe.enqueue()
        u_h = pipe.wait()[0]
        e2e_ms = (time.perf_counter() - t_e2e) * 1e3
        barrier()
    # Several independent chains per GPU (the reference runs 4-10): chain c's H2D copy, kernel and D2H copy on three
    # streams (parallel.InterleavedHostPotential) -- every step still brings its inputs from page-locked host memory
    # and takes its gradient and sums back, but the copy engines work under the other chain's kernel.
    e2e_il_ms = None
    if args.e2e_chains > 1 and (world == 1 or fused):
        from scalable_stellar_inference_b200.parallel import InterleavedHostPotential
        il = InterleavedHostPotential(model, chains=args.e2e_chains)
        zflat = model.flatten(zl)
        for c in range(args.e2e_chains):
            il.load(c, zflat)
        for c in range(args.e2e_chains):
            il.submit(c)
        for c in range(args.e2e_chains):
            u_il = il.result(c)[0]
        if not (abs(u_il - u_first) <= 1e-7 * abs(u_first)) or not torch.equal(pot.grad_dev.cpu()[nh0:], il.grad_pin[0][nh0:]):
            raise SystemExit(f"interleaved host-buffer result {u_il} differs from the device-resident result {u_first}")
        barrier()
        t_e2e = time.perf_counter()
        for k in range(K):
            c = k % args.e2e_chains
            if k >= args.e2e_chains:
                il.result(c)                 # a chain's next step needs its previous result (host-side sampler)
            il.submit(c)
        for c in range(args.e2e_chains):
            il.result(c)
        e2e_il_ms = (time.perf_counter() - t_e2e) * 1e3
        barrier()
    e2e_total = torch.tensor([e2e_ms, e2e_sync_ms, e2e_il_ms if e2e_il_ms is not None else 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_total, op=dist.ReduceOp.MAX)
    # both host-buffer modes move the same bytes per step inside their timed region; the headline is the faster one
    # (max over ranks of each): interleaving wins while the copies are long enough to hide (1 GPU: 28 MB per step),
    # the single zero-copy chain wins once a rank's shard is small and the per-step host waits dominate (8 GPUs)
    use_il = e2e_il_ms is not None and float(e2e_total[2].item()) < float(e2e_total[0].item())
    clocks = sampler.stop()
    if prev_affinity is not None:
        os.sched_setaffinity(0, prev_affinity)   # the CPU baseline below uses every host core again
    if not (np.isfinite(u_h) and abs(u_h - u_first) <= 1e-7 * abs(u_first)) and not os.environ.get("CSI_TC_DEBUG"):
        raise SystemExit(f"e2e result {u_h} differs from the device-resident result {u_first}")
    nh_ = len(model.hyper_sites)
    if args.e2e_mode == "zerocopy" and not torch.equal(pot.grad_dev.cpu()[nh_:], pipe.grad_pin[nh_:]):
        raise SystemExit("zero-copy gradient differs from the device-resident gradient")

    if rank == 0:
        bf16_peak, hbm_peak, which = measured_peaks()
        ms_per_step = total_ms / K
        kern = float(kern_ms.mean())
        stars_local = model.n_local
        achieved = FLOP_PER_STAR * stars_local / (kern * 1e-3) / 1e12
        traffic, ncu = ncu_traffic()
        if ncu is not None and (ncu.get("stars") != stars_local or world != 1 or args.config != "c4"):
            traffic = None  # the capture is of the 1-GPU, 1 M-star config-4 launch only
        n_sites_, k_obs = len(model.star_sites), len(model.observable_indices)
        alg_bytes = (2 * n_sites_ + 2 * k_obs + 3) * 4 * stars_local   # sites in, gradients out, obs + diag, M_H prior
        line = {
            "metric": "logp+grad evals/s", "value": 1e3 / ms_per_step, "unit": "evals/s", "n_gpus": world,
            "steps": K, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (fp16 hi/lo split x3 MMA, fp32 accumulate)"
            if args.engine == "tcgen05" else "f32",
            "data": "synthetic", "config": workload_config(args, model.info()["engine"]),
            "star_evals_per_s": args.stars * 1e3 / ms_per_step,
            "e2e": {"value": 1e3 / (float(e2e_total[2].item() if use_il else e2e_total[0].item()) / K), "unit": "evals/s",
                    "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes,
                    "sync_per_step": {"value": 1e3 / (float(e2e_total[1].item()) / K), "unit": "evals/s",
                                      "api": "csi_potential_host, one host synchronisation per step"},
                    "single_chain_queued": {"value": 1e3 / (float(e2e_total[0].item()) / K), "unit": "evals/s",
                                            "api": "csi_potential on page-locked buffers (zero copy), K steps queued, one synchronisation"},
                    "interleaved_chains": None if e2e_il_ms is None else
                    {"value": 1e3 / (float(e2e_total[2].item()) / K), "unit": "evals/s", "chains": args.e2e_chains, "api": e2e_il_api},
                    "selected": "interleaved_chains" if use_il else "single_chain_queued",
                    "chains": args.e2e_chains if use_il else 1,
                    "host_syncs_in_timed_region": K if use_il else (1 if e2e_async else K),
                    "blocks": B, "block_stars": [hi_b - lo_b for lo_b, hi_b in bounds],
                    "host_numa": numa,
                    "mode": args.e2e_mode, "cuda_graph": getattr(pipe, "graph", None) is not None,
                    "api": e2e_il_api if use_il else e2e_api},
            "gpu_launches": int(gpu_launches), "gpu_launches_e2e_per_step": B,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": bf16_peak, "unit": "TFLOP/s",
                         "frac": achieved / bf16_peak, "traffic": traffic,
                         "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full)",
                         "traffic_source": None if traffic is None else
                         "committed ncu capture profiles/potential_ncu.json"
                         + (f" @ {git_sha('profiles/potential_ncu.json')}" if git_sha('profiles/potential_ncu.json') else "")
                         + " (a constant of that capture, not a measurement of this run)",
                         "algorithmic_bytes": alg_bytes,
                         "kernel": "potential_tc_kernel" if args.engine == "tcgen05" else "potential_ffma_kernel",
                         "kernel_ms": kern, "stars_per_launch": stars_local, "flop_per_star": FLOP_PER_STAR,
                         "peak_source": which + "; fp16/bf16 dense tensor peak",
                         "executed_mma_tflops": MMA_PASSES * HIDDEN_FLOP_PER_STAR * stars_local / (kern * 1e-3) / 1e12,
                         "note": "3 MMA passes per contraction (error-compensated fp16 split): algorithmic frac is "
                                 "capped at 1/3 of the fp16 peak; executed_mma_tflops / peak is the tensor-pipe share"},
            "potential": u_first,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference(args.cfg, args.stars, args.cpu_budget)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def emit(line: dict):
    """The ONE JSON line on the real stdout (library banners such as NCCL's version line go to stderr)."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


_REAL_STDOUT = None

if __name__ == "__main__":
    # keep stdout clean for the single JSON line: everything else that writes to fd 1 (NCCL prints
    # "NCCL version ..." there) is redirected to stderr for the whole run
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
