#!/usr/bin/env python
"""bench.py -- logp+grad evaluations/s of the hierarchical spec+phot model (BASELINE.json config 4: 1 M synthetic
stars, observables log_Teff, log_g, G/BP/RP magnitudes, parallax; BC table interpolation), star axis sharded over
N GPUs with one 6-double allreduce per evaluation.

    python bench.py --gpus 1 --steps 20 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 20 --warmup 3
    python bench.py --impl reference ...      # CPU restatement of the reference on the host cores

One "step" = one evaluation of U(z) and dU/dz for ALL stars (the thing numpyro's NUTS calls per leapfrog).
value = evaluations/s with z resident in HBM; e2e = the same through ShardedPotential.value_and_grad_host
(pinned host z -> device, evaluate, allreduce, gradient + U back to the host, every step).
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

OBSERVABLES = ["log_Teff", "log_g", "mag", "plx"]
BANDS = ["G", "BP", "RP"]
FLOP_PER_STAR = 332_288          # emulator fwd + input-VJP (SURVEY 8d / BASELINE.md section 3)
HIDDEN_FLOP_PER_STAR = 327_680   # the 10 tensor-core contractions
MMA_PASSES = 3


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--stars", type=int, default=1_000_000)
    ap.add_argument("--engine", default="tcgen05", choices=["tcgen05", "ffma"])
    ap.add_argument("--cpu-sample", type=int, default=20_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--nccl-allreduce", action="store_true", help="keep the NCCL allreduce instead of the in-kernel exchange")
    ap.add_argument("--e2e-blocks", type=int, default=6, help="star blocks of the pipelined host-buffer path")
    ap.add_argument("--e2e-mode", default="zerocopy", choices=["zerocopy", "pipelined"],
                    help="host-buffer path: the kernel reads / writes pinned host memory itself (one launch), or "
                         "star blocks with explicit H2D / D2H copies on streams")
    ap.add_argument("--e2e-equal", action="store_true", help="equal star blocks instead of the triangular ramp")
    ap.add_argument("--e2e-no-graph", action="store_true", help="issue the host-buffer path per call, no CUDA graph")
    return ap.parse_args()


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
def cpu_reference_rate(n_total, sample, budget_s=20.0, steps=None, warmup=2):
    """Times the CPU restatement of the reference (oracle port: numpy, fp32 emulator as in the reference's default
    precision, BLAS on all host cores) on `sample` stars of the same workload; returns evals/s extrapolated to
    n_total stars (cost is linear in stars) and a description."""
    from oracle import model as M, synth
    from oracle.bc import GridInterpolator, synthetic_grid
    from oracle.nets import load_emulator
    spec = M.ModelSpec("hierarchical", sample, list(OBSERVABLES), bands=3, bc=GridInterpolator(*synthetic_grid()),
                       bc_mode="grid", emulator=load_emulator(np.float32))
    z0, obs, diag, ml, ms = synth.make_problem(spec)
    for _ in range(warmup):
        M.potential_and_grad(spec, z0, obs, diag, ml, ms)
    times = []
    t_start = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        M.potential_and_grad(spec, z0, obs, diag, ml, ms)
        times.append(time.perf_counter() - t0)
        if steps is not None and len(times) >= steps:
            break
        if steps is None and (time.perf_counter() - t_start > budget_s or len(times) >= 200):
            break
    t = float(np.median(times))
    cores = os.cpu_count() or 1
    rate = 1.0 / (t * n_total / sample)
    desc = (f"oracle port (numpy, fp32 emulator, manual reverse mode), {sample} of {n_total} stars per timed call, "
            f"median of {len(times)} calls = {t * 1e3:.1f} ms, extrapolated linearly in stars")
    return rate, cores, desc, t


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rate, cores, desc, t = cpu_reference_rate(args.stars, args.cpu_sample, steps=args.steps, warmup=max(args.warmup, 1))
    line = {
        "impl": "reference", "metric": "logp+grad evals/s", "value": rate, "unit": "evals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / rate, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, "cpu"),
        "cpu_baseline": {"value": rate, "unit": "evals/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": rate, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference JAX-CPU unavailable on this box (jax, numpyro, h5py not installed, no network): the CPU "
                "restatement of the same path is timed instead",
    }
    emit(line)


def workload_config(args, engine):
    return {"workload": "BASELINE config 4: HierarchicalStarModel spec+phot, observables "
                        "[log_Teff, log_g, mag(G,BP,RP), plx], kind=diag, BC table 70x26x18x13 (synthetic MIST stand-in)",
            "stars": args.stars, "sites_per_star": 7, "hyper_sites": 5, "engine": engine,
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

    grid = synthetic.synthetic_bc_grid(bands=3)
    model = ssi.HierarchicalStarModel(args.stars, list(OBSERVABLES), bands=BANDS, bc="grid", bc_grid=grid,
                                      engine=args.engine, device=local)
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
    prev_affinity = bind_host_to_device(local)   # pinned buffers and the launching thread next to the GPU's PCIe root
    if args.e2e_mode == "zerocopy":
        # the device-resident model of this rank (same handle, shard and exchange) fed from host memory
        pipe = ZeroCopyHostPotential(model)
        pipe.load(model.flatten(zl))
        B, bounds = 1, [(lo_r, hi_r)]
        e2e_api = "csi_potential_host on page-locked buffers (parallel.ZeroCopyHostPotential.value_and_grad_host)"
    else:
        bounds = block_bounds(lo_r, hi_r, max(1, args.e2e_blocks), unit=2 * 128 * model.info()["sm_count"],
                              ramp=not args.e2e_equal)
        B = len(bounds)
        blocks, zb = [], []
        for b, (lo_b, hi_b) in enumerate(bounds):
            mb = ssi.HierarchicalStarModel(args.stars, list(OBSERVABLES), bands=BANDS, bc="grid", bc_grid=grid,
                                           engine=args.engine, device=local)
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
    e2e_ms = (time.perf_counter() - t_e2e) * 1e3
    barrier()
    e2e_total = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_total, op=dist.ReduceOp.MAX)
    clocks = sampler.stop()
    if prev_affinity is not None:
        os.sched_setaffinity(0, prev_affinity)   # the CPU baseline below uses every host core again
    if not (np.isfinite(u_h) and abs(u_h - u_first) <= 1e-7 * abs(u_first)) and not os.environ.get("CSI_TC_DEBUG"):
        raise SystemExit(f"e2e result {u_h} differs from the device-resident result {u_first}")
    if args.e2e_mode == "zerocopy" and os.environ.get("CSI_HOST_GATE_EXP", "0") == "0" \
            and not torch.equal(pot.grad_dev.cpu()[5:], pipe.grad_pin[5:]):
        raise SystemExit("zero-copy gradient differs from the device-resident gradient")

    if rank == 0:
        bf16_peak, hbm_peak, which = measured_peaks()
        ms_per_step = total_ms / K
        kern = float(kern_ms.mean())
        stars_local = model.n_local
        achieved = FLOP_PER_STAR * stars_local / (kern * 1e-3) / 1e12
        traffic, ncu = ncu_traffic()
        if ncu is not None and (ncu.get("stars") != stars_local or world != 1):
            traffic = None  # the capture is of the 1-GPU, 1 M-star launch only
        line = {
            "metric": "logp+grad evals/s", "value": 1e3 / ms_per_step, "unit": "evals/s", "n_gpus": world,
            "steps": K, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (fp16 hi/lo split x3 MMA, fp32 accumulate)",
            "data": "synthetic", "config": workload_config(args, model.info()["engine"]),
            "star_evals_per_s": args.stars * 1e3 / ms_per_step,
            "e2e": {"value": 1e3 / (float(e2e_total.item()) / K), "unit": "evals/s",
                    "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes,
                    "blocks": B, "block_stars": [hi_b - lo_b for lo_b, hi_b in bounds],
                    "host_bound_to_gpu_numa_node": prev_affinity is not None,
                    "mode": args.e2e_mode, "cuda_graph": getattr(pipe, "graph", None) is not None,
                    "api": e2e_api},
            "gpu_launches": int(gpu_launches), "gpu_launches_e2e_per_step": B,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": bf16_peak, "unit": "TFLOP/s",
                         "frac": achieved / bf16_peak, "traffic": traffic,
                         "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full)",
                         "algorithmic_bytes": 29 * 4 * stars_local,
                         "kernel": "potential_tc_kernel" if args.engine == "tcgen05" else "potential_ffma_kernel",
                         "kernel_ms": kern, "stars_per_launch": stars_local, "flop_per_star": FLOP_PER_STAR,
                         "peak_source": which + "; fp16/bf16 dense tensor peak",
                         "executed_mma_tflops": MMA_PASSES * HIDDEN_FLOP_PER_STAR * stars_local / (kern * 1e-3) / 1e12,
                         "note": "3 MMA passes per contraction (error-compensated fp16 split): algorithmic frac is "
                                 "capped at 1/3 of the fp16 peak; executed_mma_tflops / peak is the tensor-pipe share"},
            "potential": u_first,
        }
        if world == 1 and not args.no_cpu_baseline:
            rate, cores, desc, _ = cpu_reference_rate(args.stars, args.cpu_sample, budget_s=15.0)
            line["cpu_baseline"] = {"value": rate, "unit": "evals/s", "cores": cores, "kind": "port", "sample": desc}
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
