"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's per-frame SLAM loop
(pin_slam.py:96-260, utils/mapper.py:110-361 + 425-593, dataset/slam_dataset.py:260-430) on top of
oracle/pin_oracle.py.  Used by the end-to-end trajectory parity check (ATE between the CUDA path and this
loop on the same synthetic scans and the same random draws) and as the reference arm's frames/s leg.
"""
from __future__ import annotations

import math
import time

import numpy as np
import torch

from . import pin_oracle as O


def _crop(points, cfg):
    d = torch.norm(points[:, :3], dim=1)
    keep = (d > cfg.min_range) & (d < cfg.max_range) & (points[:, 2] > cfg.min_z) & (points[:, 2] < cfg.max_z)
    return points[keep]


def _close_to_identity(mat, rot_thre, tran_thre):
    return bool(np.all(np.abs(mat[:3, :3] - np.identity(3)) < rot_thre) and np.all(mat[:3, 3] < tran_thre))


class OracleSLAM:
    def __init__(self, cfg, sampler, generator, seed=42):
        """cfg: a Config with the reference's field names; sampler: DataSampler-like (CPU tensors)."""
        self.cfg = cfg
        self.gen = generator
        self.sampler = sampler
        p = O.MapParams(voxel_size_m=cfg.voxel_size_m, buffer_size=int(cfg.buffer_size), num_nei_cells=cfg.num_nei_cells,
                        search_alpha=cfg.search_alpha, query_nn_k=cfg.query_nn_k, feature_dim=cfg.feature_dim,
                        feature_std=cfg.feature_std, weighted_first=cfg.weighted_first,
                        local_map_radius=cfg.local_map_radius,
                        local_map_travel_dist_ratio=cfg.local_map_travel_dist_ratio, use_mid_ts=cfg.use_mid_ts,
                        sdf_scale=cfg.logistic_gaussian_ratio * cfg.sigma_sigmoid_m)
        self.s = O.new_map(p)
        torch.manual_seed(seed)
        lin1 = torch.nn.Linear(cfg.feature_dim + 3, cfg.geo_mlp_hidden_dim)
        lout = torch.nn.Linear(cfg.geo_mlp_hidden_dim, 1)
        self.d = O.DecoderParams(lin1.weight.detach().clone(), lin1.bias.detach().clone(), lout.weight.detach().clone(),
                                 lout.bias.detach().clone(), sdf_scale=p.sdf_scale)
        self.decoder_frozen = False
        self.coord_pool = torch.empty((0, 3))
        self.global_pool = torch.empty((0, 3))
        self.label_pool = torch.empty((0,))
        self.weight_pool = torch.empty((0,))
        self.time_pool = torch.empty((0,), dtype=torch.int64)
        self.new_idx = None
        self.cur_sample_count = 0
        self.pool_sample_count = 0
        self.lose_track = False
        self.stop_status = False
        self.stop_count = 0
        self.travel_dist = []
        self.odom_poses = []
        self.last_pose = np.eye(4)
        self.last_odom_tran = np.eye(4)
        if cfg.kitti_correction_on:
            self.last_odom_tran[0, 3] = cfg.max_range * 1e-2
        self.processed = 0
        self.tracking_iters = []

    def _randint(self, high, size):
        return torch.randint(0, high, size, generator=self.gen)

    # ---- dataset side ---------------------------------------------------------------------------
    def preprocess(self, scan):
        cfg = self.cfg
        pc = torch.tensor(scan[:, :3], dtype=torch.float32)
        pc = pc[O.voxel_down_sample(pc, cfg.vox_down_m)]
        pc = _crop(pc, cfg)
        src = None
        guess = None
        if self.processed == 0:
            self.odom_poses.append(np.eye(4))
            self.travel_dist.append(0.0)
            self.last_pose = np.eye(4)
        else:
            guess = self.last_pose @ self.last_odom_tran if (cfg.uniform_motion_on and not self.lose_track) else self.last_pose
            src = pc[O.voxel_down_sample(pc, cfg.source_vox_down_m)]
        return pc, src, guess

    def update_odom(self, pose):
        cfg = self.cfg
        self.last_odom_tran = np.linalg.inv(self.last_pose) @ pose
        self.stop_count = self.stop_count + 1 if _close_to_identity(self.last_odom_tran, 1e-3, cfg.voxel_size_m * 0.1) else 0
        self.stop_status = self.stop_count > cfg.stop_frame_thre
        self.odom_poses.append(self.odom_poses[-1] @ self.last_odom_tran)
        d = float(np.linalg.norm(self.last_odom_tran[:3, 3]))
        if d > cfg.surface_sample_range_m * 40.0:
            self.lose_track = True
        self.travel_dist.append(self.travel_dist[-1] + d)
        self.last_pose = pose

    # ---- mapper side ----------------------------------------------------------------------------
    def process_frame(self, pc, pose, frame_id):
        cfg, s = self.cfg, self.s
        pose_t = torch.tensor(pose, dtype=torch.float32)
        origin = pose_t[:3, 3]
        coord, label, _, _, _, weight = self.sampler.sample(pc, None, None, None, generator=self.gen)
        self.cur_sample_count = label.shape[0]
        upd = coord[torch.abs(label) < cfg.surface_sample_range_m * cfg.map_surface_ratio, :]
        upd = O.transform_points(upd, pose_t)
        O.map_update(s, upd, origin, frame_id)
        self.coord_pool = torch.cat((self.coord_pool, coord), 0)
        self.weight_pool = torch.cat((self.weight_pool, weight), 0)
        self.label_pool = torch.cat((self.label_pool, label), 0)
        self.time_pool = torch.cat((self.time_pool, torch.full((coord.shape[0],), frame_id, dtype=torch.int64)), 0)
        self.global_pool = torch.cat((self.global_pool, O.transform_points(coord, pose_t)), 0)
        if (frame_id + 1) % cfg.pool_filter_freq == 0:
            mask = torch.sum((self.global_pool - origin) ** 2, dim=-1) < cfg.window_radius ** 2
            true_idx = torch.nonzero(mask).squeeze()
            if true_idx.shape[0] > cfg.pool_capacity:
                disc = self._randint(true_idx.shape[0], (true_idx.shape[0] - cfg.pool_capacity,))
                mask[true_idx[disc]] = False
            self.coord_pool, self.global_pool = self.coord_pool[mask], self.global_pool[mask]
            self.label_pool, self.weight_pool, self.time_pool = self.label_pool[mask], self.weight_pool[mask], self.time_pool[mask]
            self.cur_sample_count = int(mask[-self.cur_sample_count:].sum())
            self.pool_sample_count = int(mask.sum())
        else:
            self.cur_sample_count = coord.shape[0]
            self.pool_sample_count = self.coord_pool.shape[0]
        if cfg.bs_new_sample > 0:
            cur = self.global_pool[-self.cur_sample_count:]
            cur_label = self.label_pool[-self.cur_sample_count:]
            O.set_search_neighborhood(s, 1, 0.0)
            cert = O.query_certainty(s, cur)
            O.set_search_neighborhood(s, cfg.num_nei_cells, cfg.search_alpha)
            self.new_idx = torch.where((cert < cfg.new_certainty_thre) & (torch.abs(cur_label) < cfg.surface_sample_range_m * 3.0))[0]
            self.new_idx = self.new_idx + (self.pool_sample_count - self.cur_sample_count)

    def get_batch(self):
        cfg = self.cfg
        if cfg.bs_new_sample > 0 and self.new_idx is not None and not self.lose_track and not self.stop_status:
            k = self.new_idx.shape[0]
            if k > 0:
                bs_new = min(k, cfg.bs_new_sample)
                hist = self._randint(self.pool_sample_count, (cfg.bs - bs_new,))
                new = self.new_idx[self._randint(k, (bs_new,))]
                index = torch.cat((hist, new), 0)
            else:
                index = self._randint(self.pool_sample_count, (cfg.bs,))
        else:
            index = self._randint(self.pool_sample_count, (cfg.bs,))
        return self.global_pool[index], self.label_pool[index], self.time_pool[index], self.weight_pool[index]

    def mapping(self, iters):
        cfg, s, d = self.cfg, self.s, self.d
        s.l_feats = s.l_feats.detach().requires_grad_(True)
        dec_params = [] if self.decoder_frozen else [t.requires_grad_(True) for t in d.tensors()]
        groups = []
        if dec_params:
            groups.append({"params": dec_params, "lr": cfg.lr})
        groups.append({"params": [s.l_feats], "lr": cfg.lr})
        opt = torch.optim.Adam(groups, betas=(0.9, 0.99), eps=cfg.adam_eps)
        for _ in range(iters):
            coord, label, ts, weight = self.get_batch()
            loss, _, _ = O.mapping_loss(s, d, coord, label, ts, weight, weight_e=cfg.weight_e,
                                        decimation=cfg.gradient_decimation, eps=cfg.voxel_size_m * cfg.num_grad_step_ratio)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
        s.l_feats = s.l_feats.detach()
        for t in d.tensors():
            t.requires_grad_(False)
        O.assign_local_to_global(s)

    # ---- the loop -------------------------------------------------------------------------------
    def step(self, scan, frame_id):
        cfg, s = self.cfg, self.s
        t0 = time.perf_counter()
        pc, src, guess = self.preprocess(scan)
        t1 = time.perf_counter()
        pose = np.eye(4)
        if self.processed > 0:
            T, valid, it, _ = O.tracking(s, self.d, src, torch.tensor(guess, dtype=torch.float64), iter_n=cfg.reg_iter_n,
                                         term_deg=cfg.reg_term_thre_deg, term_m=cfg.reg_term_thre_m,
                                         surface_sample_range_m=cfg.surface_sample_range_m,
                                         min_grad_norm=cfg.reg_min_grad_norm, max_grad_norm=cfg.reg_max_grad_norm,
                                         GM_dist=cfg.reg_GM_dist_m, GM_grad=cfg.reg_GM_grad, lm_lambda=cfg.reg_lm_lambda,
                                         max_sdf_std=cfg.surface_sample_range_m * cfg.max_sdf_std_ratio, min_nn=cfg.query_nn_k)
            self.tracking_iters.append(it)
            self.lose_track = not valid
            pose = T.numpy().astype(np.float64)
            self.update_odom(pose)
        s.travel_dist = torch.tensor(np.array(self.travel_dist), dtype=torch.float32)
        t2 = time.perf_counter()
        if not self.lose_track and not self.stop_status:
            self.process_frame(pc, pose, self.processed)
        else:
            O.reset_local_map(s, torch.tensor(pose[:3, 3], dtype=torch.float32), self.processed)
        t3 = time.perf_counter()
        iters = cfg.iters * cfg.init_iter_ratio if self.processed == 0 else cfg.iters
        if self.processed == cfg.freeze_after_frame:
            self.decoder_frozen = True
        self.mapping(iters)
        t4 = time.perf_counter()
        self.processed += 1
        return pose, [t1 - t0, t2 - t1, t3 - t2, t4 - t3]
