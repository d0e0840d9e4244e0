#!/usr/bin/env python3
"""Runs the reference's UNCHANGED driver `pin_slam.py` (oracle/_ref or /root/reference, through oracle/ref_shim.py) on a
synthetic KITTI-layout sequence -- TEST / BENCH INFRASTRUCTURE, not product code.

  --mode dropin      pin_slam_b200.dropin.install() first: model.neural_points / model.decoder / utils.tracker /
                     utils.mapper / utils.mesher resolve to the B200 mirrors, everything else (config, dataset, pose graph,
                     loop detector, result writers) stays the reference's own code            (BASELINE configs[1])
  --mode reference   nothing replaced: the reference end to end, `--device cuda` = its stock PyTorch-eager path on the same
                     GPU (the comparator a user of the reference would run), `--device cpu` = BASELINE configs[0]

Sequence: `--frames` motion-distorted 64-beam sweeps of the seeded street scene (120 000 rays,
Velodyne timing, the inverse of the KITTI intrinsic correction applied), written as sequences/00/velodyne/*.bin + poses +
calib.  Config = the reference's config/lidar_slam/run_kitti.yaml with only the paths, `o3d_vis_on: False` and the `pgo:`
section removed (oracle/ref_shim.write_kitti_yaml): deskew ON, kitti_correct ON.

Prints one JSON line: frames/s from the driver's own time_table columns (pin_slam.py:343; first frame = 600 warm-up
iterations reported separately), ATE against ground truth (eval/eval_traj_utils.py:13-56 = the reference's own function),
the estimated poses (for ATE between runs)."""
import argparse
import importlib
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="dropin", choices=["dropin", "reference"])
    ap.add_argument("--device", default="cuda")
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--beams", type=int, default=64)
    ap.add_argument("--azimuth", type=int, default=1875)
    ap.add_argument("--workers", type=int, default=min(os.cpu_count() or 1, 16))
    ap.add_argument("--data-dir", default=None, help="reuse / keep the generated sequence here")
    ap.add_argument("--iters", type=int, default=None, help="TEST ONLY: override optimizer.iters (smoke runs on CPU)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    import torch
    from oracle import ref_shim
    from pin_slam_b200 import synth

    keep = args.data_dir is not None
    root = args.data_dir or tempfile.mkdtemp(prefix="pinq_kitti_")
    t0 = time.time()
    vel = os.path.join(root, "sequences", "00", "velodyne")
    if not (os.path.isdir(vel) and len(os.listdir(vel)) >= args.frames):
        frames, gt = synth.make_sweeps(args.frames, args.beams, args.azimuth, args.workers, correct_deg=0.195)
        paths = synth.write_kitti_sequence(root, frames, gt)
        del frames
    else:
        paths = {"pc_path": vel, "pose_path": os.path.join(root, "poses", "00.txt"),
                 "calib_path": os.path.join(root, "sequences", "00", "calib.txt")}
    gen_s = time.time() - t0
    extra = {"setting": {"end_frame": args.frames - 1}}
    if args.iters is not None:
        extra["optimizer"] = {"iters": args.iters}
    out_root = os.path.join(root, "experiments")
    yaml_path = ref_shim.write_kitti_yaml(os.path.join(root, "run_kitti_synth.yaml"), paths, out_root, extra=extra)
    with open(yaml_path) as f:
        text = f.read()
    if args.device != "cuda":  # the reference reads `device` from the setting section (utils/config.py)
        text = text.replace("setting:\n", "setting:\n  device: %s\n" % args.device, 1)
        with open(yaml_path, "w") as f:
            f.write(text)

    ref_shim.install()
    if args.mode == "dropin":
        import pin_slam_b200.dropin as dropin
        replaced = dropin.install()
    else:
        replaced = []
    # instrumentation only: remember the dataset object the unchanged driver creates (it returns nothing)
    ds_mod = importlib.import_module("dataset.slam_dataset")
    seen = {}
    inner_init = ds_mod.SLAMDataset.__init__

    def recording_init(self, *a, **k):
        inner_init(self, *a, **k)
        seen["dataset"] = self
    ds_mod.SLAMDataset.__init__ = recording_init
    pin_slam = importlib.import_module("pin_slam")
    ref_shim._patch_get_time()
    argv0 = sys.argv
    sys.argv = ["pin_slam.py", yaml_path]
    cwd = os.getcwd()
    # setup_experiment runs `git rev-parse HEAD` in the working directory (utils/tools.py:73): use this repository, or --
    # on the GPU box, whose snapshot has no .git -- a scratch repository with one empty commit
    import subprocess
    work = ROOT
    if subprocess.run(["git", "rev-parse", "HEAD"], cwd=ROOT, capture_output=True).returncode != 0:
        work = os.path.join(root, "gitwork")
        os.makedirs(work, exist_ok=True)
        subprocess.run(["git", "init", "-q"], cwd=work, check=True)
        subprocess.run(["git", "-c", "user.name=pinq", "-c", "user.email=pinq@localhost", "commit", "-q", "--allow-empty", "-m", "scratch"],
                       cwd=work, check=True)
    os.chdir(work)
    t1 = time.time()
    try:
        pin_slam.run_pin_slam()
    finally:
        sys.argv = argv0
        os.chdir(cwd)
    wall = time.time() - t1
    ds = seen["dataset"]
    tt = np.array(ds.time_table)                      # [frames, 5]: preprocess, tracking, map prep, training, pgo
    est = np.stack(ds.odom_poses)
    gt = np.stack(ds.gt_poses)
    ev = importlib.import_module("eval.eval_traj_utils")
    ate_rot, ate_trans, _ = ev.absolute_error(list(gt), list(est), False)
    n = est.shape[0]
    native = [ln.split()[-1] for ln in open("/proc/self/maps") if "libpinq" in ln]
    out = {"driver": "reference pin_slam.py, unchanged" + (" + pin_slam_b200.dropin" if args.mode == "dropin" else ""),
           "mode": args.mode, "device": args.device, "replaced_modules": replaced, "frames": int(n),
           "rays_per_scan": args.beams * args.azimuth, "deskew": bool(ds.config.deskew),
           "kitti_correct": bool(ds.config.kitti_correction_on), "config": "run_kitti.yaml (paths, o3d_vis_on False, pgo removed)",
           "frames_per_s": float((n - 1) / tt[1:, :4].sum()) if n > 1 else None,
           "first_frame_s": float(tt[0, :4].sum()),
           "ms_per_frame": {k: float(tt[1:, i].mean() * 1e3) for i, k in enumerate(("preprocess", "tracking", "map_prep", "training"))} if n > 1 else None,
           "median_ms_per_frame": float(np.median(tt[1:, :4].sum(1)) * 1e3) if n > 1 else None,
           "preprocess_ms_by_frame": [round(float(v) * 1e3, 2) for v in tt[:, 0]][:40],
           "stage_ms_by_frame": [[round(float(v) * 1e3, 2) for v in row[:4]] for row in tt],
           "wall_s_whole_run": wall, "scan_generation_s": gen_s,
           "ate_vs_ground_truth_m": float(ate_trans), "ate_rot_deg": float(ate_rot),
           "libpinq_loaded": sorted(set(os.path.basename(p) for p in native)),
           "reference_root": ref_shim.REFERENCE_ROOT, "torch_threads": torch.get_num_threads(),
           "poses": est[:, :3, :].reshape(n, 12).tolist()}
    line = json.dumps(out)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")
    if not keep:
        shutil.rmtree(root, ignore_errors=True)


if __name__ == "__main__":
    main()
