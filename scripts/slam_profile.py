#!/usr/bin/env python3
"""Where a SLAM frame's time goes on the CUDA path: runs a few synthetic frames, then profiles one mapping(15) call and one
tracking call with torch.profiler (host + device).  python scripts/slam_profile.py > gpurun_out/slam_profile.log"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pin_slam_b200 import synth  # noqa: E402
from pin_slam_b200.slam import run_slam  # noqa: E402
from scripts.e2e_parity import make_cfg  # noqa: E402


class A:
    init_iter_ratio = None


def main():
    frames_n = 12
    scene = synth.StreetScene(42)
    gt = synth.trajectory(frames_n)
    dirs = synth.lidar_directions(64, 1875)
    rng = np.random.default_rng(0)
    frames = [synth.lidar_scan(scene, gt[i], dirs, rng) for i in range(frames_n)]
    res = run_slam(make_cfg("cuda", A), frames, None, generator=torch.Generator().manual_seed(7))
    tt = res["time_table"]
    print("mean ms per frame (preprocess, tracking, map_prep, training):", (tt[1:].mean(0) * 1e3).round(2))
    mapper, tracker, ds = res["mapper"], res["tracker"], res["dataset"]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mapper.mapping(15)
    torch.cuda.synchronize()
    print("mapping(15): %.2f ms" % ((time.perf_counter() - t0) * 1e3))
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        mapper.mapping(15)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=35, max_name_column_width=60))
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        tracker.tracking(ds.cur_source_points, ds.cur_pose_guess_torch)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=25, max_name_column_width=60))


if __name__ == "__main__":
    main()
