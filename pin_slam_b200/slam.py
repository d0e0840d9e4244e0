"""run_slam -- the per-frame loop of the reference driver (pin_slam.py:31-377) over the classes of this
package, without the visualiser, loop closure, PGO and file outputs (all out of scope)."""
from __future__ import annotations

import time

import numpy as np
import torch

from .dataset.slam_dataset import SLAMDataset
from .model.decoder import Decoder
from .model.neural_points import NeuralPoints
from .utils.mapper import Mapper
from .utils.mesher import Mesher
from .utils.tools import freeze_decoders, get_time
from .utils.tracker import Tracker


def run_slam(config, frames=None, gt_poses=None, max_frames=None, seed=None, generator=None, on_frame=None):
    """Returns dict(poses [n,4,4] f64 estimated odometry, time_table [n,4] s (preprocess, tracking,
    map-prep, training: the reference's own columns, pin_slam.py:343), objects...)."""
    torch.manual_seed(config.seed if seed is None else seed)
    geo_mlp = Decoder(config, config.geo_mlp_hidden_dim, config.geo_mlp_level, 1)
    neural_points = NeuralPoints(config)
    dataset = SLAMDataset(config, frames, gt_poses)
    tracker = Tracker(config, neural_points, geo_mlp, None, None)
    mapper = Mapper(config, dataset, neural_points, geo_mlp, None, None)
    mapper.generator = generator
    mesher = Mesher(config, neural_points, geo_mlp, None, None)
    n = dataset.total_pc_count if max_frames is None else min(max_frames, dataset.total_pc_count)
    poses, table, iters = [], [], []
    for frame_id in range(n):
        used_frame_id = dataset.processed_frame
        T0 = get_time()
        dataset.read_frame(frame_id)
        T1 = get_time()
        dataset.preprocess_frame(frame_id)
        T2 = get_time()
        if used_frame_id > 0:
            if config.track_on:
                cur_pose, _, _, valid = tracker.tracking(dataset.cur_source_points, dataset.cur_pose_guess_torch)
                dataset.lose_track = not valid
                mapper.lose_track = not valid
                dataset.update_odom_pose(cur_pose)
                iters.append(tracker.last_iterations)
            else:
                dataset.update_odom_pose(dataset.cur_pose_guess_torch)
        neural_points.travel_dist = torch.tensor(np.array(dataset.travel_dist), device=config.device, dtype=config.dtype)
        T3 = get_time()
        if not mapper.lose_track and not dataset.stop_status:
            mapper.process_frame(dataset.cur_point_cloud_torch, None, dataset.cur_pose_torch, used_frame_id,
                                 config.dynamic_filter_on and used_frame_id > 0)  # pin_slam.py:235-236
        else:
            neural_points.reset_local_map(dataset.cur_pose_torch[:3, 3], None, used_frame_id)
            mapper.static_mask = None
        T4 = get_time()
        cur_iter_num = config.iters * config.init_iter_ratio if used_frame_id == 0 else config.iters
        if used_frame_id == config.freeze_after_frame:
            freeze_decoders(geo_mlp, None, None, config)
        if used_frame_id % config.mapping_freq_frame == 0:
            mapper.mapping(cur_iter_num)
        T5 = get_time()
        table.append([T2 - T1, T3 - T2, T4 - T3, T5 - T4])
        poses.append(dataset.cur_pose_ref.copy() if used_frame_id > 0 or not config.track_on else dataset.cur_pose_ref.copy())
        dataset.processed_frame += 1
        if on_frame is not None:
            on_frame(frame_id, dataset, neural_points, mapper)
    return {"poses": np.stack(poses), "time_table": np.array(table), "tracking_iters": iters, "dataset": dataset,
            "neural_points": neural_points, "decoder": geo_mlp, "mapper": mapper, "tracker": tracker, "mesher": mesher}


def absolute_trajectory_error(est: np.ndarray, gt: np.ndarray, align: bool = True):
    """ATE RMSE in metres after Horn alignment (reference: eval/eval_traj_utils.py:13-56, numpy only)."""
    P_e = est[:, :3, 3].T
    P_g = gt[:, :3, 3].T
    if align:
        mu_e, mu_g = P_e.mean(1, keepdims=True), P_g.mean(1, keepdims=True)
        W = (P_g - mu_g) @ (P_e - mu_e).T
        U, _, Vt = np.linalg.svd(W)
        S = np.eye(3)
        if np.linalg.det(U) * np.linalg.det(Vt) < 0:
            S[2, 2] = -1
        R = U @ S @ Vt
        t = mu_g - R @ mu_e
        P_e = R @ P_e + t
    err = np.linalg.norm(P_e - P_g, axis=0)
    return float(np.sqrt(np.mean(err ** 2)))
