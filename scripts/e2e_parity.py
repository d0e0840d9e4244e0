#!/usr/bin/env python3
"""End-to-end run on synthetic 64-beam scans (BASELINE.json configs[0]/[1]): the CUDA path and, optionally,
the CPU oracle loop on identical scans and random draws.  Prints one JSON line (ATE vs ground truth, ATE
between the two trajectories, frames/s from the reference's own time_table columns).
  python scripts/e2e_parity.py --frames 100 --oracle-frames 20 --out gpurun_out/e2e.json"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_cfg(device, args):
    from pin_slam_b200.utils.config import kitti_config
    cfg = kitti_config(device)
    cfg.deskew = False           # synthetic scans are instantaneous
    cfg.correction_deg = 0.0     # scans are geometrically exact; keeps the +0.8 m initial motion guess
    cfg.silence = True
    if args.init_iter_ratio is not None:
        cfg.init_iter_ratio = args.init_iter_ratio
    return cfg


def _scan(job):
    from pin_slam_b200 import synth
    seed, pose, beams, azimuth, noise_seed = job
    return synth.lidar_scan(synth.StreetScene(seed), pose, synth.lidar_directions(beams, azimuth), np.random.default_rng(noise_seed))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=1, help="> 1: generate the scans in parallel with per-frame noise seeds")
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--oracle-frames", type=int, default=0)
    ap.add_argument("--beams", type=int, default=64)
    ap.add_argument("--azimuth", type=int, default=1875)
    ap.add_argument("--init-iter-ratio", type=int, default=None)
    ap.add_argument("--out", default=None)
    ap.add_argument("--cuda-repeats", type=int, default=1, help="extra CUDA runs on the same scans and draws (atomics order differs "
                    "from run to run): the spread of the CUDA path against itself and against the oracle")
    args = ap.parse_args()
    from pin_slam_b200 import synth
    from pin_slam_b200.slam import absolute_trajectory_error, run_slam

    scene = synth.StreetScene(42)
    gt = synth.trajectory(args.frames)
    dirs = synth.lidar_directions(args.beams, args.azimuth)
    rng = np.random.default_rng(0)
    t0 = time.time()
    if args.workers > 1:  # per-frame noise streams, scans ray-cast on all host cores (4 min -> seconds on the GPU box)
        import multiprocessing as mp
        with mp.get_context("fork").Pool(args.workers) as pool:
            frames = pool.map(_scan, [(42, gt[i], args.beams, args.azimuth, 1000 + i) for i in range(args.frames)])
    else:
        frames = [synth.lidar_scan(scene, gt[i], dirs, rng) for i in range(args.frames)]
    gen_s = time.time() - t0
    res = run_slam(make_cfg("cuda", args), frames, None, generator=torch.Generator().manual_seed(7))
    repeats = [run_slam(make_cfg("cuda", args), frames, None, generator=torch.Generator().manual_seed(7))["poses"]
               for _ in range(max(args.cuda_repeats - 1, 0))]
    tt = res["time_table"]
    out = {"frames": args.frames, "rays_per_scan": int(dirs.shape[0]), "points_per_scan_mean": float(np.mean([f.shape[0] for f in frames])),
           "scan_generation_s": gen_s,
           "ate_vs_gt_m": absolute_trajectory_error(res["poses"], gt, align=False),
           "ate_vs_gt_aligned_m": absolute_trajectory_error(res["poses"], gt, align=True),
           "final_pos_err_m": float(np.linalg.norm(res["poses"][-1][:3, 3] - gt[-1][:3, 3])),
           "frames_per_s_excl_first": float((args.frames - 1) / tt[1:].sum()) if args.frames > 1 else None,
           "frames_per_s_all": float(args.frames / tt.sum()),
           "mean_ms": {"preprocess": float(tt[1:, 0].mean() * 1e3), "tracking": float(tt[1:, 1].mean() * 1e3),
                       "map_prep": float(tt[1:, 2].mean() * 1e3), "training": float(tt[1:, 3].mean() * 1e3)},
           "first_frame_s": float(tt[0].sum()), "tracking_iters_mean": float(np.mean(res["tracking_iters"])),
           "tracking_iters": [int(v) for v in res["tracking_iters"]],
           "neural_points": int(res["neural_points"].count()), "local_points": int(res["neural_points"].local_count()),
           "pool_samples": int(res["mapper"].pool_sample_count)}
    if args.oracle_frames > 0:
        from oracle.slam_oracle import OracleSLAM
        from pin_slam_b200.utils.data_sampler import DataSampler
        torch.set_num_threads(os.cpu_count() or 1)
        cpu_cfg = make_cfg("cpu", args)
        orc = OracleSLAM(cpu_cfg, DataSampler(cpu_cfg), torch.Generator().manual_seed(7), seed=cpu_cfg.seed)
        n = min(args.oracle_frames, args.frames)
        poses, times = [], []
        for i in range(n):
            p, t = orc.step(frames[i], i)
            poses.append(p)
            times.append(t)
        poses, times = np.stack(poses), np.array(times)
        d = np.linalg.norm(res["poses"][:n, :3, 3] - poses[:, :3, 3], axis=1)
        out["oracle"] = {"frames": n, "cores": torch.get_num_threads(),
                         "ate_cuda_vs_oracle_m": absolute_trajectory_error(res["poses"][:n], poses, align=False),
                         "max_pos_diff_m": float(d.max()), "per_frame_pos_diff_m": [float(v) for v in d],
                         "ate_oracle_vs_gt_m": absolute_trajectory_error(poses, gt[:n], align=False),
                         "ate_cuda_vs_gt_same_frames_m": absolute_trajectory_error(res["poses"][:n], gt[:n], align=False),
                         "frames_per_s_excl_first": float((n - 1) / times[1:].sum()) if n > 1 else None,
                         "first_frame_s": float(times[0].sum()),
                         "tracking_iters_mean": float(np.mean(orc.tracking_iters)) if orc.tracking_iters else None,
                         "tracking_iters": [int(v) for v in orc.tracking_iters],
                         "neural_points": int(orc.s.points.shape[0])}
        if repeats:
            out["oracle"]["ate_cuda_repeats_vs_oracle_m"] = [absolute_trajectory_error(r[:n], poses, align=False) for r in repeats]
    if repeats:
        out["ate_cuda_repeats_vs_first_cuda_m"] = [absolute_trajectory_error(r, res["poses"], align=False) for r in repeats]
    line = json.dumps(out)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
