#!/usr/bin/env python3
"""Experiment: the bench step with a smaller search stencil (R = 1, 27 cells) -- how the search time scales with the candidates visited."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pin_slam_b200.utils.config as C
_orig = C.kitti_config
def patched(device):
    cfg = _orig(device)
    cfg.num_nei_cells = int(os.environ.get("EXP_NEI", "1"))
    cfg.search_alpha = float(os.environ.get("EXP_ALPHA", "1.0"))
    return cfg
C.kitti_config = patched
import bench
sys.argv = ["bench.py", "--steps", "10", "--warmup", "3", "--no-e2e", "--no-cpu-baseline", "--slam-frames", "0", "--no-extra-legs", "--driver-frames", "0"]
bench.main()
