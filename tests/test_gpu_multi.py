"""The star-sharded path through libcsi on TWO GPUs (skipped on a 1-GPU box): in-kernel exchange over NVLink peer
memory == NCCL allreduce == the unsharded handle, CUDA-graph replay of the exchange, the sharded device-side
sampler, and the missing-peer -> NaN path.  -m gpu; run with `gpurun --gpus 2`."""
import os
import socket

import numpy as np
import pytest
import torch

import parity_util as PU

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    import torch.distributed as dist
    from scalable_stellar_inference_b200.nuts import DeviceNUTS
    from scalable_stellar_inference_b200.parallel import ShardedPotential, setup_peer_exchange, shard_bounds, shard_problem
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), CSI_PEER_TIMEOUT_MS="300")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    res = {}
    try:
        for name, n in (("hier_phot_grid", 3001), ("hier_diag_seis", 2500)):   # config 4 / config 5 shapes
            spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
            model = PU.product_model(name, n, "tcgen05", device=rank).set_shard(rank, world)
            zl, ol, dl, mll, msl = shard_problem(z0, obs, diag, ml, ms, rank, world)
            model.condition(ol, dl, M_H_loc=mll, M_H_scale=msl)
            pot = ShardedPotential(model)                      # NCCL allreduce of the 6 sums
            pot.z_dev.copy_(torch.from_numpy(model.flatten(zl)))
            pot.value_and_grad_device()
            torch.cuda.synchronize()
            nccl = pot.out_dev[:6].cpu().numpy().copy()
            g_nccl = pot.grad_dev.cpu().numpy().copy()
            pot.exchange = setup_peer_exchange(model)          # from here on: exchanged by the finalising block
            assert pot.exchange is not None, "symmetric memory unavailable"
            fused = []
            for _ in range(3):
                pot.value_and_grad_device()
                torch.cuda.synchronize()
                fused.append(pot.out_dev[:6].cpu().numpy().copy())
            g_fused = pot.grad_dev.cpu().numpy().copy()
            # the exchange inside a CUDA graph, replayed: the epoch lives on the device and advances per replay
            graph = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                with torch.cuda.graph(graph, stream=side):
                    pot.value_and_grad_device()
            replay = []
            for _ in range(3):
                pot.out_dev.zero_()
                graph.replay()
                torch.cuda.synchronize()
                replay.append(pot.out_dev[:6].cpu().numpy().copy())
            # sharded sampler: every rank must take the same decisions and end with the same hyper-parameters
            lo, hi = shard_bounds(n, rank, world)
            nuts = DeviceNUTS(model, model.flatten(zl), seed=5, max_tree_depth=5, star_offset=lo, n_total=n)
            step = 0.5 * nuts.find_heuristic_step_size()
            stats, draw0 = [], nuts.draw
            for _ in range(4):
                st = nuts.transition(step)
                stats.append((st.depth, st.num_leapfrog, st.potential, st.accept_prob))
            res[name] = dict(nccl=nccl, fused=fused, replay=replay, g_nccl=g_nccl, g_fused=g_fused, stats=stats,
                             step=step, draw0=draw0, hyper=nuts.z[: len(model.hyper_sites)].cpu().numpy().copy(),
                             stars=nuts.z[len(model.hyper_sites):].cpu().numpy().copy(), bounds=(lo, hi))
            del nuts
            dist.barrier()
        # missing peer: rank 1 does not launch; rank 0's finalising block must give up and return NaN, not hang
        if rank == 0:
            pot.value_and_grad_device()
            torch.cuda.synchronize()
            res["timeout"] = pot.out_dev[:6].cpu().numpy().copy()
        dist.barrier()
        ret[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_two_rank_fused_exchange_nccl_and_single_handle_agree():
    import torch.multiprocessing as mp
    from scalable_stellar_inference_b200.nuts import DeviceNUTS
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    r0, r1 = ret[0], ret[1]
    for name, n in (("hier_phot_grid", 3001), ("hier_diag_seis", 2500)):
        spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
        full = PU.product_model(name, n, "tcgen05").condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
        o_full, g_full = full.potential_and_grad_flat(full.flatten(z0))
        a, b = r0[name], r1[name]
        np.testing.assert_allclose(a["nccl"], o_full, rtol=1e-9, atol=1e-9 * np.abs(o_full).max())
        for k in range(3):   # fused totals: identical on both ranks, identical across calls and graph replays, == NCCL
            assert np.array_equal(a["fused"][k], b["fused"][k]) and np.array_equal(a["fused"][k], a["fused"][0])
            assert np.array_equal(a["replay"][k], a["fused"][0]) and np.array_equal(b["replay"][k], a["fused"][0])
        np.testing.assert_allclose(a["fused"][0], a["nccl"], rtol=1e-12, atol=0)
        nh = len(full.hyper_sites)
        assert np.array_equal(a["g_fused"][nh:], a["g_nccl"][nh:])            # per-star gradients untouched
        assert np.array_equal(a["g_fused"][:nh], a["fused"][0][1:1 + nh].astype(np.float32))   # hyper slots = totals
        # sharded sampler: same trees on both ranks, same hyper state, and the same chain as ONE unsharded handle
        assert a["stats"] == b["stats"] and np.array_equal(a["hyper"], b["hyper"])
        nuts = DeviceNUTS(full, full.flatten(z0), seed=5, max_tree_depth=5)
        nuts.draw = a["draw0"]   # the sharded run spent its first draws on the step-size probes
        single = []
        for _ in range(4):
            st = nuts.transition(a["step"])
            single.append((st.depth, st.num_leapfrog))
        same_trees = [s[:2] for s in a["stats"]] == single
        print(name, "sharded", a["stats"], "single", single)
        if same_trees:   # momenta depend on the GLOBAL star index: sharding changes only the rounding of the sums
            ns = len(full.star_sites)
            stars = np.concatenate([np.concatenate([r[name]["stars"].reshape(ns, -1)[s] for r in (r0, r1)]) for s in range(ns)])
            zs = nuts.z.cpu().numpy()
            assert np.allclose(zs[nh:], stars, rtol=0, atol=1e-3 * (1 + np.abs(stars).max()))
            assert np.allclose(zs[:nh], a["hyper"], rtol=1e-3, atol=1e-4)
    assert np.all(np.isnan(r0["timeout"])), r0["timeout"]
