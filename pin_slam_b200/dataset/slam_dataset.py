"""SLAMDataset -- per-frame bookkeeping around the hot path (reference: dataset/slam_dataset.py:48-430).

Kept: frame reading (KITTI .bin folder or in-memory synthetic scans), voxel down-sampling through the
libpinq kernel, cropping, KITTI intrinsic correction, the constant-velocity pose guess, source-point
down-sampling, odometry / travel-distance bookkeeping and stop detection -- i.e. everything
`pin_slam.py` touches between two calls into the hot path.  Out of scope: Open3D / .ply / .pcd / .las
readers, semantic labels, plots.  Deskewing (utils/tools.py:540-567) runs as one kernel, with the Velodyne per-point
times of slam_dataset.py:247-249 when the scan carries none.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch

from ..utils.tools import deskewing, velodyne_point_ts, voxel_down_sample_torch


def crop_frame(points, ts, min_z_th=-3.0, max_z_th=100.0, min_range=2.75, max_range=100.0):
    """reference: dataset/slam_dataset.py:827-834."""
    dist = torch.norm(points[:, :3], dim=1)
    keep = (dist > min_range) & (dist < max_range) & (points[:, 2] > min_z_th) & (points[:, 2] < max_z_th)
    return points[keep], (ts[keep] if ts is not None else None)


def intrinsic_correct(points, correct_deg=0.0):
    """reference: dataset/slam_dataset.py:837-854 (KITTI vertical-angle calibration)."""
    if correct_deg == 0.0:
        return points
    dist = torch.norm(points[:, :3], dim=1)
    v_ang = torch.asin(points[:, 2] / dist)
    v_ang_c = v_ang + correct_deg / 180.0 * math.pi
    hor_scale = torch.cos(v_ang_c) / torch.cos(v_ang)
    points[:, 0] *= hor_scale
    points[:, 1] *= hor_scale
    points[:, 2] = dist * torch.sin(v_ang_c)
    return points


def tranmat_close_to_identity(mat, rot_thre, tran_thre):
    """reference: utils/tools.py:622-635."""
    return bool(np.all(np.abs(mat[:3, :3] - np.identity(3)) < rot_thre) and np.all(mat[:3, 3] < tran_thre))


class SLAMDataset:
    def __init__(self, config, frames=None, gt_poses=None):
        """`frames`: optional list of float32 [N,3+] numpy scans in the sensor frame (synthetic sequence);
        otherwise KITTI-layout .bin files are listed from config.pc_path (slam_dataset.py:631-646)."""
        self.config = config
        self.silence = config.silence
        self.dtype = config.dtype
        self.device = config.device
        self.frames = frames
        if frames is None:
            names = sorted(f for f in os.listdir(config.pc_path) if f.endswith(".bin"))
            self.pc_filenames = names
            self.total_pc_count = len(names)
        else:
            self.total_pc_count = len(frames)
        self.gt_pose_provided = gt_poses is not None
        self.poses_ref = None
        if gt_poses is not None:
            gt_poses = np.asarray(gt_poses, dtype=np.float64)
            begin_inv = np.linalg.inv(gt_poses[0]) if config.first_frame_ref else np.eye(4)
            self.poses_ref = np.stack([begin_inv @ p for p in gt_poses])
        self.odom_poses, self.gt_poses, self.pgo_poses, self.travel_dist = [], [], [], []
        self.time_table = []
        self.processed_frame = 0
        self.lose_track = False
        self.consecutive_lose_track_frame = 0
        self.last_pose_ref = np.eye(4)
        self.last_odom_tran = np.eye(4)
        self.cur_pose_ref = np.eye(4)
        if config.kitti_correction_on:
            self.last_odom_tran[0, 3] = config.max_range * 1e-2  # slam_dataset.py:126-127
        self.stop_count = 0
        self.stop_status = False
        self.cur_point_cloud_torch = None
        self.cur_point_ts_torch = None
        self.cur_sem_labels_torch = None
        self.cur_source_points = None
        self.cur_source_colors = None
        self.cur_source_normals = None
        self.cur_pose_torch = None
        self.cur_pose_guess_torch = None

    def read_frame(self, frame_id):
        """reference: dataset/slam_dataset.py:199-231."""
        if self.gt_pose_provided:
            self.cur_pose_ref = self.poses_ref[frame_id]
            self.gt_poses.append(self.cur_pose_ref)
        else:
            self.cur_pose_ref = np.eye(4)
        self.cur_pose_torch = torch.tensor(self.cur_pose_ref, device=self.device, dtype=self.dtype)
        if self.frames is not None:
            pc = self.frames[frame_id]
        else:
            pc = np.fromfile(os.path.join(self.config.pc_path, self.pc_filenames[frame_id]), dtype=np.float32).reshape(-1, 4)
        self.cur_point_cloud_torch = torch.tensor(pc[:, :3], device=self.device, dtype=self.dtype)
        self.cur_point_ts_torch = None
        if self.config.deskew:  # slam_dataset.py:228-257 (KITTI .bin carries no per-point time: Velodyne yaw convention)
            if self.config.lidar_type_guess != "velodyne":
                raise NotImplementedError("per-point times are derived for Velodyne-type sweeps only")
            self.cur_point_ts_torch = velodyne_point_ts(self.cur_point_cloud_torch)

    def preprocess_frame(self, frame_id=0):
        """reference: dataset/slam_dataset.py:260-374."""
        cfg = self.config
        crop_max_range = cfg.max_range
        train_voxel_m = (crop_max_range / cfg.max_range) * cfg.vox_down_m
        source_voxel_m = (crop_max_range / cfg.max_range) * cfg.source_vox_down_m
        n0 = self.cur_point_cloud_torch.shape[0]
        if cfg.rand_downsample:
            idx = torch.randint(0, n0, (int(n0 * cfg.rand_down_r),), device=self.device)
        else:
            idx = voxel_down_sample_torch(self.cur_point_cloud_torch[:, :3].contiguous(), train_voxel_m)
        self.cur_point_cloud_torch = self.cur_point_cloud_torch[idx]
        if self.cur_point_ts_torch is not None:
            self.cur_point_ts_torch = self.cur_point_ts_torch[idx]
        self.cur_point_cloud_torch, self.cur_point_ts_torch = crop_frame(self.cur_point_cloud_torch, self.cur_point_ts_torch,
                                                                         cfg.min_z, cfg.max_z, cfg.min_range, crop_max_range)
        if cfg.kitti_correction_on:
            self.cur_point_cloud_torch = intrinsic_correct(self.cur_point_cloud_torch, cfg.correction_deg)
        if self.processed_frame == 0:
            if cfg.track_on:
                self.odom_poses.append(self.cur_pose_ref)
            if self.gt_pose_provided and frame_id > 0:
                self.last_odom_tran = np.linalg.inv(self.poses_ref[frame_id - 1]) @ self.cur_pose_ref
            self.travel_dist.append(0.0)
            self.last_pose_ref = self.cur_pose_ref
        else:
            if cfg.uniform_motion_on and not self.lose_track:
                guess = self.last_pose_ref @ self.last_odom_tran
            else:
                guess = self.last_pose_ref
            if not cfg.track_on and self.gt_pose_provided:
                guess = self.poses_ref[frame_id]
            self.cur_pose_guess_torch = torch.tensor(guess, dtype=torch.float64, device=self.device)
            src = self.cur_point_cloud_torch.clone()
            idx = voxel_down_sample_torch(src[:, :3].contiguous(), source_voxel_m)
            self.cur_source_points = src[idx][:, :3].contiguous()
            # deskewing of the source points with the motion of the previous interval (slam_dataset.py:367-370)
            if cfg.deskew and not self.lose_track and self.cur_point_ts_torch is not None:
                self.cur_source_points = deskewing(self.cur_source_points, self.cur_point_ts_torch[idx],
                                                   torch.tensor(self.last_odom_tran, device=self.device, dtype=self.dtype))

    def update_odom_pose(self, cur_pose_torch):
        """reference: dataset/slam_dataset.py:376-430."""
        cfg = self.config
        self.cur_pose_torch = cur_pose_torch.detach()
        self.cur_pose_ref = self.cur_pose_torch.cpu().numpy().astype(np.float64)
        self.last_odom_tran = np.linalg.inv(self.last_pose_ref) @ self.cur_pose_ref
        if tranmat_close_to_identity(self.last_odom_tran, 1e-3, cfg.voxel_size_m * 0.1):
            self.stop_count += 1
        else:
            self.stop_count = 0
        self.stop_status = self.stop_count > cfg.stop_frame_thre
        if self.odom_poses is not None:
            self.odom_poses.append(self.odom_poses[-1] @ self.last_odom_tran)
        d = float(np.linalg.norm(self.last_odom_tran[:3, 3]))
        if d > cfg.surface_sample_range_m * 40.0:
            self.lose_track = True
        self.travel_dist.append(self.travel_dist[-1] + d)
        self.last_pose_ref = self.cur_pose_ref
        # deskewing of the cloud used for mapping with the motion just estimated (slam_dataset.py:419-422)
        if cfg.deskew and not self.lose_track and self.cur_point_ts_torch is not None:
            self.cur_point_cloud_torch = deskewing(self.cur_point_cloud_torch.contiguous(), self.cur_point_ts_torch,
                                                   torch.tensor(self.last_odom_tran, device=self.device, dtype=self.dtype))
        self.consecutive_lose_track_frame = self.consecutive_lose_track_frame + 1 if self.lose_track else 0
        if self.consecutive_lose_track_frame > 20:
            raise RuntimeError("Lose track for a long time, system failed")
