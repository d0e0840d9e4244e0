"""A/B of the short SLAM run (tests/test_gpu_slam.py) with and without the brick index: prints the ATEs."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pin_slam_b200 import synth
from pin_slam_b200.slam import absolute_trajectory_error, run_slam
from pin_slam_b200.utils.config import kitti_config

n_frames = int(os.environ.get("FRAMES", "6"))
scene = synth.StreetScene(42)
gt = synth.trajectory(n_frames)
dirs = synth.lidar_directions(32, 450)
rng = np.random.default_rng(0)
frames = [synth.lidar_scan(scene, gt[i], dirs, rng) for i in range(n_frames)]
cfg = kitti_config("cuda")
cfg.deskew = False; cfg.correction_deg = 0.0; cfg.init_iter_ratio = 10; cfg.bs = 4096; cfg.silence = True
out = {}
for rep in range(int(os.environ.get("REPS", "2"))):
    res = run_slam(cfg, frames, None, generator=torch.Generator().manual_seed(7))
    out[rep] = res["poses"]
    print("rep", rep, "brick", os.environ.get("PINQ_BRICK", "1"), "ate_gt", absolute_trajectory_error(res["poses"], gt, align=False),
          "points", res["neural_points"].count())
print("rep0 vs rep1", absolute_trajectory_error(out[0], out[1], align=False) if len(out) > 1 else None)
np.save(os.path.join(ROOT, "gpurun_out", "poses_brick%s.npy" % os.environ.get("PINQ_BRICK", "1")), out[0])
