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


def _wrap(cls, name):
    fn = getattr(cls, name)

    def wrapped(*a, **k):
        with torch.profiler.record_function("pin::%s.%s" % (cls.__name__, name)):
            return fn(*a, **k)
    setattr(cls, name, wrapped)


def main():
    from torch.profiler import ProfilerActivity, profile
    from pin_slam_b200.dataset.slam_dataset import SLAMDataset
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.data_sampler import DataSampler
    from pin_slam_b200.utils.mapper import Mapper
    from pin_slam_b200.utils.tracker import Tracker
    for cls, names in ((SLAMDataset, ("read_frame", "preprocess_frame", "update_odom_pose")), (Tracker, ("tracking", "registration_step")),
                       (Mapper, ("process_frame", "mapping", "_draw_batches")), (DataSampler, ("sample",)),
                       (NeuralPoints, ("update", "reset_local_map", "assign_local_to_global", "query_certainty", "_refresh_candidates"))):
        for nm in names:
            _wrap(cls, nm)
    do_profile = os.environ.get("PROFILE", "1") != "0"
    seeded = os.environ.get("SEEDED", "1") != "0"
    frames_n = 18 if do_profile else 40
    scene = synth.StreetScene(42)
    gt = synth.trajectory(frames_n)
    dirs = synth.lidar_directions(64, 1875)
    rng = np.random.default_rng(0)
    frames = [synth.lidar_scan(scene, gt[i], dirs, rng) for i in range(frames_n)]
    prof = profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA])
    state = {}

    def on_frame(fid, dataset, npts, mapper):
        state.setdefault("timing", []).append(dict(mapper.timing))
        if fid == 9 and do_profile:
            torch.cuda.synchronize()
            prof.__enter__()
        if fid == 15 and do_profile:
            torch.cuda.synchronize()
            prof.__exit__(None, None, None)

    res = run_slam(make_cfg("cuda", A), frames, None, generator=torch.Generator().manual_seed(7) if seeded else None, on_frame=on_frame)
    tt = res["time_table"]
    print("mean ms per frame (preprocess, tracking, map_prep, training):", (tt[4:].mean(0) * 1e3).round(2))
    keys = state["timing"][-1].keys()
    print("process_frame ms:", {k: round(float(np.mean([t[k] for t in state["timing"][4:]])), 2) for k in keys})
    print("tracking iterations:", res["tracking_iters"])
    if not do_profile:
        return
    ka = prof.key_averages()
    print("---- ranges (6 frames) ----")
    for e in sorted([e for e in ka if e.key.startswith("pin::")], key=lambda e: -e.cpu_time_total):
        print("%-44s calls %4d  cpu_total %9.1f us  per call %8.1f us  cuda_total %9.1f us" %
              (e.key, e.count, e.cpu_time_total, e.cpu_time_total / e.count, e.device_time_total))
    print(ka.table(sort_by="self_cpu_time_total", row_limit=40, max_name_column_width=60))
    print(ka.table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=60))


if __name__ == "__main__":
    main()
