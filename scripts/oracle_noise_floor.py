#!/usr/bin/env python3
"""Trajectory noise floor of the reference's own algorithm (CPU oracle loop, no GPU needed).

The control the trajectory-parity gate needs ("What's weak" 1): the CPU oracle loop against ITSELF on the same
100 synthetic 64-beam scans, with (a) every scan coordinate perturbed by at most one fp32 ulp (6e-8 relative -- a 1e-8
relative perturbation is below fp32 resolution and would be a no-op) and (b) identical scans but another intra-op thread
count (different fp32 reduction order inside ATen).  The ATE between such runs is what "the same algorithm on the same data"
scatters by; the CUDA-vs-oracle ATE of scripts/e2e_parity.py is judged against max(1 cm, 1.5 x this floor).

  python scripts/oracle_noise_floor.py --frames 100 --out profiles/r3_oracle_noise_floor.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _scan(job):
    from pin_slam_b200 import synth
    seed, pose, beams, azimuth, noise_seed = job
    return synth.lidar_scan(synth.StreetScene(seed), pose, synth.lidar_directions(beams, azimuth), np.random.default_rng(noise_seed))


def make_frames(n, beams, azimuth, workers):
    import multiprocessing as mp
    from pin_slam_b200 import synth
    gt = synth.trajectory(n)
    with mp.get_context("fork").Pool(workers) as pool:
        frames = pool.map(_scan, [(42, gt[i], beams, azimuth, 1000 + i) for i in range(n)])
    return frames, gt


def ulp_perturb(frames, seed=123):
    """+-1 ulp on a random half of the coordinates."""
    rng = np.random.default_rng(seed)
    out = []
    for f in frames:
        bits = f.view(np.int32).copy()
        bits += rng.integers(-1, 2, bits.shape, dtype=np.int32) * (rng.random(bits.shape) < 0.5)
        out.append(bits.view(np.float32))
    return out


def run_oracle(frames, args, threads):
    from oracle.slam_oracle import OracleSLAM
    from pin_slam_b200.utils.config import kitti_config
    from pin_slam_b200.utils.data_sampler import DataSampler
    torch.set_num_threads(threads)
    cfg = kitti_config("cpu")
    cfg.deskew, cfg.correction_deg, cfg.silence = False, 0.0, True
    if args.init_iter_ratio is not None:
        cfg.init_iter_ratio = args.init_iter_ratio
    orc = OracleSLAM(cfg, DataSampler(cfg), torch.Generator().manual_seed(7), seed=cfg.seed)
    poses = []
    t0 = time.time()
    for i, f in enumerate(frames):
        poses.append(orc.step(f, i)[0])
    return np.stack(poses), time.time() - t0, [int(v) for v in orc.tracking_iters]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--beams", type=int, default=64)
    ap.add_argument("--azimuth", type=int, default=1875)
    ap.add_argument("--init-iter-ratio", type=int, default=None)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--alt-threads", type=int, default=1, help="thread count of control (b); 0 = skip")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    from pin_slam_b200.slam import absolute_trajectory_error as ate

    frames, gt = make_frames(args.frames, args.beams, args.azimuth, min(os.cpu_count() or 1, 16))
    base, t_base, it_base = run_oracle(frames, args, args.threads)
    pert, t_pert, it_pert = run_oracle(ulp_perturb(frames), args, args.threads)
    d = np.linalg.norm(base[:, :3, 3] - pert[:, :3, 3], axis=1)
    out = {"frames": args.frames, "rays_per_scan": args.beams * args.azimuth, "threads": args.threads,
           "perturbation": "+-1 fp32 ulp on a random half of the scan coordinates",
           "ate_oracle_vs_perturbed_oracle_m": ate(base, pert, align=False), "max_pos_diff_m": float(d.max()),
           "per_frame_pos_diff_m": [round(float(v), 5) for v in d],
           "ate_oracle_vs_gt_m": ate(base, gt, align=False), "ate_perturbed_vs_gt_m": ate(pert, gt, align=False),
           "oracle_s": [t_base, t_pert], "tracking_iters_mean": [float(np.mean(it_base)), float(np.mean(it_pert))]}
    if args.alt_threads > 0 and args.alt_threads != args.threads:
        alt, t_alt, it_alt = run_oracle(frames, args, args.alt_threads)
        d2 = np.linalg.norm(base[:, :3, 3] - alt[:, :3, 3], axis=1)
        out["other_thread_count"] = {"threads": args.alt_threads, "ate_vs_base_m": ate(base, alt, align=False),
                                     "max_pos_diff_m": float(d2.max()), "oracle_s": t_alt,
                                     "per_frame_pos_diff_m": [round(float(v), 5) for v in d2]}
    line = json.dumps(out)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
