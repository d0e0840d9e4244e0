#!/usr/bin/env python3
"""BASELINE.json configs[3] and configs[4] on N GPUs (one process per GPU; launch with torchrun for N > 1):

  config 4  batched mapping: 2^20 samples per iteration drawn from a replay pool with one shared seed, every rank takes
            its contiguous slice (fused forward + numerical-gradient queries + backward), the gradients / certainty
            increments / ts_update are exchanged by ONE peer-memory kernel per rank over NVLink (csrc/peer.cu;
            PINQ_EXCHANGE=nccl: NCCL all-reduces), then the replicated Adam step.  15 iterations, samples/s.
            Self-check: rank 0 also computes the un-sharded gradient of iteration 0 and reports its distance.
  config 5  mesher dense SDF grid: 0.1 m voxels over the map's AABB, contiguous linear ranges of the z-fastest grid per
            rank, no collective; grid points/s and time to the full grid (marching cubes excluded).

Prints one JSON line per config on rank 0.   python scripts/bench_configs.py [--side 400] [--voxel 0.1]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--side", type=float, default=400.0)
    ap.add_argument("--surface-points", type=int, default=12_000_000)
    ap.add_argument("--voxel", type=float, default=0.1)
    ap.add_argument("--batch", type=int, default=1 << 20)
    ap.add_argument("--pool", type=int, default=8_000_000)
    ap.add_argument("--iters", type=int, default=15)
    ap.add_argument("--skip", default="")
    args = ap.parse_args()
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = "cuda:%d" % local_rank
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    from pin_slam_b200 import synth
    from pin_slam_b200.parallel import MappingPeer, PeerExchange, dp_mapping_iteration
    from pin_slam_b200.utils.mapper import Mapper
    from pin_slam_b200.utils.mesher import Mesher
    from pin_slam_b200.utils.tools import setup_optimizer

    cfg, npts, dec = bench.build_map(argparse.Namespace(surface_points=args.surface_points, side=args.side), dev)
    m = int(npts.count())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    # ---------------------------------------------------------------- config 5: mesher grid
    if "5" not in args.skip:
        mesher = Mesher(cfg, npts, dec, None, None)
        pts = npts.neural_points
        lo, hi = pts.min(0).values.cpu().numpy(), pts.max(0).values.cpu().numpy()
        res = {}
        for rep in range(2):  # first pass warms the allocator
            barrier()
            e0.record()
            sdf, mask, first, dims, origin = mesher.query_grid(lo, hi, args.voxel, mask_min_nn_count=cfg.mesh_min_nn, rank=rank, world=world)
            e1.record()
            barrier()
            res = {"ms": max_over_ranks(e0.elapsed_time(e1)), "dims": [int(v) for v in dims], "inside": float(mask.float().mean())}
            del sdf, mask
        total = int(np.prod(res["dims"]))
        if rank == 0:
            print(json.dumps({"config": "mesher dense SDF grid (BASELINE configs[4])", "n_gpus": world, "voxel_m": args.voxel,
                              "grid_dims": res["dims"], "grid_points": total, "neural_points": m, "seconds_to_full_grid": res["ms"] * 1e-3,
                              "value": total / (res["ms"] * 1e-3), "unit": "grid points/s", "mask_fraction_rank0": res["inside"],
                              "collective": "none (contiguous linear ranges per rank)"}))
        torch.cuda.empty_cache()

    # ---------------------------------------------------------------- config 4: data-parallel mapping
    if "4" not in args.skip:
        g = torch.Generator().manual_seed(3)  # the same pool and the same draws on every rank
        pool = torch.from_numpy(synth.microbench_queries(npts.neural_points.cpu().numpy(), args.pool, seed=5)).to(dev)
        label = (torch.randn(args.pool, generator=g) * 0.1).to(dev)
        weight = torch.ones(args.pool, device=dev)
        tsp = torch.zeros(args.pool, dtype=torch.int64, device=dev)
        mapper = Mapper(cfg, None, npts, dec, None, None)
        exchange = os.environ.get("PINQ_EXCHANGE", "peer") if world > 1 else "none"
        peer = None
        rows, fdim = npts.local_geo_features.shape
        if exchange == "peer":
            peer = PeerExchange(MappingPeer.capacity(rows, fdim, dec), dev)
            torch.cuda.synchronize()
            dist.barrier()
        opt = setup_optimizer(cfg, list(npts.parameters()), list(dec.parameters()), None, None)
        draws = [torch.randint(0, args.pool, (args.batch,), generator=g).to(dev) for _ in range(args.iters + 1)]

        # self-check on iteration 0: exchanged gradient of the sharded batch vs the un-sharded gradient (same state)
        def rel(a, b):
            return float((a - b).norm() / b.norm().clamp_min(1e-30))

        idx = draws[0]
        feats0 = npts.local_geo_features.detach().clone()
        dec0 = [p.detach().clone() for p in dec.parameters()]
        cert0, ts0 = npts.local_point_certainties.clone(), npts.local_point_ts_update.clone()
        got, want = [], []
        for keep, kw in ((got, dict(peer=peer)), (want, dict(shard=(0, 1)))):
            for grp in opt.param_groups:
                grp["lr"] = 0.0  # only the gradients are of interest here
            dp_mapping_iteration(mapper, pool[idx], label[idx], tsp[idx], weight[idx], opt, keep_grads=keep, **kw)
            with torch.no_grad():
                npts.local_geo_features.copy_(feats0)
                for p, v in zip(dec.parameters(), dec0):
                    p.copy_(v)
            npts.local_point_certainties, npts.local_point_ts_update = cert0.clone(), ts0.clone()
            opt = setup_optimizer(cfg, list(npts.parameters()), list(dec.parameters()), None, None)
        check = {"rel_err_vs_unsharded": {k: rel(a, b) for k, a, b in zip(("features", "W1", "b1", "W2", "b2"), got, want)}}

        for it in range(2):  # warm-up
            idx = draws[it]
            dp_mapping_iteration(mapper, pool[idx], label[idx], tsp[idx], weight[idx], opt, peer=peer)
        barrier()
        e0.record()
        losses = []
        for it in range(args.iters):
            idx = draws[1 + it]
            losses.append(dp_mapping_iteration(mapper, pool[idx], label[idx], tsp[idx], weight[idx], opt, peer=peer))
        e1.record()
        barrier()
        ms_autograd = max_over_ranks(e0.elapsed_time(e1))
        err = peer.error() if peer is not None else 0
        if peer is not None:
            dist.barrier()
            peer.close()
            peer = None
        del draws
        torch.cuda.empty_cache()

        # the product form: Mapper.mapping's device-side loop (pinq_mapping_run), data-parallel through Mapper.dp
        from pin_slam_b200.parallel import DataParallelMapping
        cfg.bs, cfg.bs_new_sample = args.batch, 0
        mapper.global_coord_pool, mapper.coord_pool = pool, pool
        mapper.sdf_label_pool, mapper.weight_pool, mapper.time_pool = label, weight, tsp
        mapper.pool_sample_count = args.pool
        mapper.generator = torch.Generator(device=dev).manual_seed(11)  # the same Philox stream on every rank
        if world > 1:
            peer = PeerExchange(MappingPeer.capacity(rows, fdim, dec, with_params=True), dev)
            torch.cuda.synchronize()
            dist.barrier()
            mapper.dp = DataParallelMapping(peer)
        assert mapper._fused_mapping_supported()
        for _ in range(2):
            mapper._mapping_fused(args.iters)
        barrier()
        e0.record()
        mapper._mapping_fused(args.iters)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        err = max(err, peer.error() if peer is not None else 0)
        last = mapper.last_loss.clone()
        if world > 1:
            dist.all_reduce(last)
        same = torch.tensor([float(npts.local_geo_features.data.double().sum())], device=dev, dtype=torch.float64)
        lo_hi = [same.clone(), same.clone()]
        if world > 1:
            dist.all_reduce(lo_hi[0], op=dist.ReduceOp.MIN)
            dist.all_reduce(lo_hi[1], op=dist.ReduceOp.MAX)
        if rank == 0:
            n_eik = 6 * ((args.batch + cfg.gradient_decimation - 1) // cfg.gradient_decimation)
            print(json.dumps({"config": "batched mapping (BASELINE configs[3])", "n_gpus": world, "samples_per_iter": args.batch,
                              "queries_per_iter": args.batch + n_eik, "iters": args.iters, "pool": args.pool, "neural_points": m,
                              "ms_per_iter": ms / args.iters, "value": args.batch * args.iters / (ms * 1e-3), "unit": "samples/s",
                              "form": "device-side loop (pinq_mapping_run)" + (", pinq_peer_exchange with fused Adam per iteration" if world > 1 else ""),
                              "exchange": "peer" if world > 1 else "none", "peer_exchange_error": err,
                              "loss_bce_eik_first_last": [last[0].tolist(), last[-1].tolist()],
                              "replicas_identical": bool(float(lo_hi[0]) == float(lo_hi[1])),
                              "autograd_form": {"ms_per_iter": ms_autograd / args.iters, "exchange": exchange,
                                                "adam": "replicated torch.optim.Adam",
                                                "loss_first_last_rank0": [float(losses[0]), float(losses[-1])],
                                                "self_check": check}}))
        if peer is not None:
            dist.barrier()
            peer.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
