"""Config -- the parameters the neural-point SDF query path reads, with the reference's names,
defaults and YAML layout (reference: utils/config.py:10-325 defaults, :327-585 load()).

Only the keys that the hot path and its harness consume are kept; semantic / colour / ROS /
visualiser / PGO knobs are out of scope but their *switches* are present and
default to off so that code written against the reference's Config keeps working.
"""
from __future__ import annotations

import os

import torch
import yaml


class Config:
    def __init__(self):
        # settings (config.py:15-44)
        self.name = "dummy"
        self.run_path = ""
        self.output_root = ""
        self.pc_path = ""
        self.pose_path = ""
        self.calib_path = ""
        self.first_frame_ref = False
        self.begin_frame = 0
        self.end_frame = 0
        self.every_frame = 1
        self.seed = 42
        self.device = "cuda"
        self.gpu_id = "0"
        self.dtype = torch.float32
        self.tran_dtype = torch.float64
        self.adaptive_mode = False
        self.kitti_correction_on = False
        self.correction_deg = 0.0
        self.deskew = True
        self.valid_ts_in_points = True
        # process (config.py:56-69)
        self.min_range = 2.5
        self.max_range = 60.0
        self.adaptive_range_on = False
        self.min_z = -4.0
        self.max_z = 60.0
        self.rand_downsample = True
        self.vox_down_m = 0.05
        self.rand_down_r = 1.0
        self.estimate_normal = True
        # switches of out-of-scope subsystems (kept off)
        self.semantic_on = False
        self.sem_class_count = 20
        self.color_map_on = False
        self.color_on = False
        self.color_channel = 0
        self.lidar_type_guess = "velodyne"  # config.py:54
        self.dynamic_filter_on = False
        self.dynamic_certainty_thre = 4.0   # config.py:231
        self.dynamic_sdf_ratio_thre = 1.5   # config.py:232
        # neural points (config.py:88-127)
        self.weighted_first = True
        self.layer_norm_on = False
        self.voxel_size_m = 0.3
        self.num_nei_cells = 2
        self.query_nn_k = 6
        self.use_mid_ts = False
        self.search_alpha = 0.2
        self.buffer_size = int(5e7)
        self.feature_dim = 8
        self.feature_std = 0.0
        self.from_sample_points = True
        self.from_all_samples = False
        self.map_surface_ratio = 0.5
        self.diff_ts_local = 400.0
        self.local_map_travel_dist_ratio = 5.0
        self.local_map_radius = 50.0
        self.prune_map_on = False
        self.max_prune_certainty = 2.0
        self.use_gaussian_pe = False
        self.pos_encoding_freq = 200
        self.pos_encoding_band = 0
        self.pos_input_dim = 3
        self.pos_encoding_base = 2
        # sampler (config.py:136-143)
        self.surface_sample_range_m = 0.25
        self.surface_sample_n = 3
        self.free_sample_begin_ratio = 0.3
        self.free_sample_end_dist_m = 1.0
        self.free_front_n = 2
        self.free_behind_n = 1
        # replay pool (config.py:146-151)
        self.window_radius = 50.0
        self.pool_capacity = int(1e7)
        self.bs_new_sample = 2048
        self.new_certainty_thre = 1.0
        self.pool_filter_freq = 10
        self.new_sample_ratio_thre = 0.01
        # tracking (config.py:154-173)
        self.track_on = True
        self.photometric_loss_on = False
        self.photometric_loss_weight = 0.01
        self.source_vox_down_m = 0.8
        self.uniform_motion_on = True
        self.reg_min_grad_norm = 0.5
        self.reg_max_grad_norm = 2.0
        self.max_sdf_ratio = 5.0
        self.max_sdf_std_ratio = 1.0
        self.reg_dist_div_grad_norm = False
        self.reg_GM_dist_m = 0.5
        self.reg_GM_grad = 0.2
        self.reg_lm_lambda = 1e-4
        self.reg_iter_n = 50
        self.reg_term_thre_deg = 0.01
        self.reg_term_thre_m = 0.0005
        self.eigenvalue_check = True
        self.consist_wieght_on = True
        self.stop_frame_thre = 20
        # decoder (config.py:178-191)
        self.mlp_bias_on = True
        self.geo_mlp_level = 1
        self.geo_mlp_hidden_dim = 64
        self.sem_mlp_level = 1
        self.sem_mlp_hidden_dim = 64
        self.color_mlp_level = 1
        self.color_mlp_hidden_dim = 64
        self.freeze_after_frame = 40
        # loss (config.py:195-231)
        self.main_loss_type = "bce"
        self.sigma_sigmoid_m = 0.1
        self.logistic_gaussian_ratio = 0.55
        self.proj_correction_on = False
        self.loss_weight_on = False
        self.behind_dropoff_on = False
        self.dist_weight_on = True
        self.dist_weight_scale = 0.8
        self.numerical_grad = True
        self.gradient_decimation = 10
        self.num_grad_step_ratio = 0.2
        self.ekional_loss_on = True
        self.ekional_add_to = "all"
        self.weight_e = 0.5
        self.consistency_loss_on = False
        # optimizer (config.py:234-247)
        self.mapping_freq_frame = 1
        self.ba_freq_frame = 0
        self.iters = 15
        self.init_iter_ratio = 40
        self.opt_adam = True
        self.bs = 16384
        self.lr = 0.01
        self.lr_pose = 1e-3
        self.weight_decay = 0.0
        self.adam_eps = 1e-15
        # loop closure / pgo switches (out of scope, off)
        self.global_loop_on = False
        self.local_map_context = False
        self.use_gt_loop = False
        self.pgo_on = False
        # eval (config.py:281-318)
        self.wandb_vis_on = False
        self.silence = False
        self.o3d_vis_on = False
        self.mc_res_m = 0.1
        self.pad_voxel = 2
        self.skip_top_voxel = 2
        self.mc_mask_on = True
        self.mesh_min_nn = 8
        self.mesh_freq_frame = 10
        self.infer_bs = 4096
        self.save_map = False
        self.save_mesh = False

    def load(self, config_file):
        with open(os.path.abspath(config_file)) as f:
            args = yaml.safe_load(f)
        self.load_dict(args)

    def load_dict(self, args):
        """Same section/key names and derived values as the reference's load() (config.py:327-585)."""
        s = args.get("setting", None)
        if s is not None:
            self.name = s.get("name", "pin_slam")
            self.output_root = s.get("output_root", "./experiments")
            self.pc_path = s.get("pc_path", "")
            self.pose_path = s.get("pose_path", "")
            self.calib_path = s.get("calib_path", "")
            self.first_frame_ref = s.get("first_frame_ref", False)
            self.begin_frame = s.get("begin_frame", 0)
            self.end_frame = s.get("end_frame", 999999)
            self.every_frame = s.get("every_frame", 1)
            self.seed = s.get("random_seed", self.seed)
            self.device = s.get("device", "cuda")
            self.gpu_id = s.get("gpu_id", "0")
            self.kitti_correction_on = s.get("kitti_correct", False)
            if self.kitti_correction_on:
                self.correction_deg = s.get("correct_deg", self.correction_deg)
            self.deskew = s.get("deskew", True)
            self.valid_ts_in_points = s.get("valid_ts", True)
        p = args.get("process", None)
        if p is not None:
            self.min_range = p.get("min_range_m", self.min_range)
            self.max_range = p.get("max_range_m", self.max_range)
            self.min_z = p.get("min_z_m", self.min_z)
            self.max_z = p.get("max_z_m", self.max_z)
            self.rand_downsample = p.get("rand_downsample", False)
            if self.rand_downsample:
                self.rand_down_r = p.get("rand_down_r", self.rand_down_r)
            else:
                self.vox_down_m = p.get("vox_down_m", self.max_range * 1e-3)
            self.estimate_normal = p.get("estimate_normal", False)
            self.dynamic_filter_on = p.get("dynamic_filter_on", False)
            self.adaptive_range_on = p.get("adaptive_range_on", False)
        sm = args.get("sampler", None)
        if sm is not None:
            self.surface_sample_range_m = sm.get("surface_sample_range_m", self.vox_down_m * 3.0)
            self.free_sample_begin_ratio = sm.get("free_sample_begin_ratio", self.free_sample_begin_ratio)
            self.free_sample_end_dist_m = sm.get("free_sample_end_dist_m", self.surface_sample_range_m * 4.0)
            self.surface_sample_n = sm.get("surface_sample_n", self.surface_sample_n)
            self.free_front_n = sm.get("free_front_sample_n", self.free_front_n)
            self.free_behind_n = sm.get("free_behind_sample_n", self.free_behind_n)
        n = args.get("neuralpoints", None)
        if n is not None:
            self.voxel_size_m = n.get("voxel_size_m", self.vox_down_m * 5.0)
            self.query_nn_k = n.get("query_nn_k", self.query_nn_k)
            self.num_nei_cells = n.get("num_nei_cells", self.num_nei_cells)
            self.search_alpha = n.get("search_alpha", self.search_alpha)
            self.feature_dim = n.get("feature_dim", self.feature_dim)
            self.use_gaussian_pe = n.get("pos_encoding_gaussian", False)
            self.pos_encoding_band = n.get("pos_encoding_band", 0)
            self.weighted_first = n.get("weighted_first", self.weighted_first)
            self.from_sample_points = n.get("from_sample_points", self.from_sample_points)
            if self.from_sample_points:
                self.map_surface_ratio = n.get("map_surface_ratio", self.map_surface_ratio)
            self.max_prune_certainty = n.get("max_prune_certainty", self.max_prune_certainty)
            self.use_mid_ts = n.get("use_mid_ts", self.use_mid_ts)
            self.local_map_travel_dist_ratio = n.get("local_map_travel_dist_ratio", self.local_map_travel_dist_ratio)
        d = args.get("decoder", None)
        if d is not None:
            self.geo_mlp_level = d.get("mlp_level", self.geo_mlp_level)
            self.geo_mlp_hidden_dim = d.get("mlp_hidden_dim", self.geo_mlp_hidden_dim)
            self.freeze_after_frame = d.get("freeze_after_frame", self.freeze_after_frame)
        ls = args.get("loss", None)
        if ls is not None:
            self.main_loss_type = ls.get("main_loss_type", "bce")
            self.sigma_sigmoid_m = ls.get("sigma_sigmoid_m", self.vox_down_m)
            self.loss_weight_on = ls.get("loss_weight_on", self.loss_weight_on)
            if self.loss_weight_on:
                self.dist_weight_scale = ls.get("dist_weight_scale", self.dist_weight_scale)
                self.behind_dropoff_on = ls.get("behind_dropoff_on", self.behind_dropoff_on)
            self.ekional_loss_on = ls.get("ekional_loss_on", self.ekional_loss_on)
            self.weight_e = float(ls.get("weight_e", self.weight_e))
            self.numerical_grad = ls.get("numerical_grad_on", self.numerical_grad)
            if not self.numerical_grad:
                self.gradient_decimation = 1
            else:
                self.gradient_decimation = ls.get("grad_decimation", self.gradient_decimation)
                self.num_grad_step_ratio = ls.get("num_grad_step_ratio", self.num_grad_step_ratio)
            self.consistency_loss_on = ls.get("consistency_loss_on", self.consistency_loss_on)
        c = args.get("continual", None)
        if c is not None:
            self.pool_capacity = int(float(c.get("pool_capacity", self.pool_capacity)))
            self.bs_new_sample = int(c.get("batch_size_new_sample", self.bs_new_sample))
            self.new_certainty_thre = float(c.get("new_certainty_thre", self.new_certainty_thre))
            self.pool_filter_freq = c.get("pool_filter_freq", 1)
        # NOTE the reference stores the YAML sub-dict (truthy) here, not a bool (config.py:475, 496)
        self.track_on = args.get("tracker", False)
        if self.track_on:
            t = args["tracker"]
            self.uniform_motion_on = t.get("uniform_motion_on", self.uniform_motion_on)
            self.source_vox_down_m = t.get("source_vox_down_m", self.vox_down_m * 10)
            self.reg_iter_n = t.get("iter_n", self.reg_iter_n)
            self.reg_min_grad_norm = t.get("min_grad_norm", self.reg_min_grad_norm)
            self.reg_max_grad_norm = t.get("max_grad_norm", self.reg_max_grad_norm)
            self.reg_GM_grad = t.get("GM_grad", self.reg_GM_grad)
            self.reg_GM_dist_m = t.get("GM_dist", self.reg_GM_dist_m)
            self.reg_lm_lambda = float(t.get("lm_lambda", self.reg_lm_lambda))
            self.reg_term_thre_deg = float(t.get("term_deg", self.reg_term_thre_deg))
            self.reg_term_thre_m = float(t.get("term_m", self.reg_term_thre_m))
            self.eigenvalue_check = t.get("eigenvalue_check", True)
            if args.get("pgo", False):
                raise NotImplementedError("pose-graph optimisation (gtsam) is out of scope: remove the `pgo:` section")
        o = args.get("optimizer", None)
        if o is not None:
            self.mapping_freq_frame = o.get("mapping_freq_frame", 1)
            self.iters = o.get("iters", self.iters)
            self.init_iter_ratio = o.get("init_iter_ratio", self.init_iter_ratio)
            self.bs = o.get("batch_size", self.bs)
            self.lr = float(o.get("learning_rate", self.lr))
            self.ba_freq_frame = o.get("ba_freq_frame", 0)
        e = args.get("eval", None)
        if e is not None:
            self.wandb_vis_on = e.get("wandb_vis_on", False)
            self.silence = e.get("silence_log", self.silence)
            self.o3d_vis_on = False  # graphics API: unavailable on this box
            self.mesh_freq_frame = e.get("mesh_freq_frame", self.mesh_freq_frame)
            self.mesh_min_nn = e.get("mesh_min_nn", self.mesh_min_nn)
            self.skip_top_voxel = e.get("skip_top_voxel", self.skip_top_voxel)
            self.mc_res_m = e.get("mc_res_m", self.voxel_size_m)
            self.save_map = e.get("save_map", False)
            self.save_mesh = e.get("save_mesh", False)
        # associated parameters (config.py:568-576)
        self.infer_bs = self.bs * 64
        self.window_radius = max(self.max_range, 6.0)
        self.local_map_radius = self.max_range + 2.0


RUN_KITTI = {
    # config/lidar_slam/run_kitti.yaml with the `pgo:` section removed (gtsam unavailable) and the
    # visualiser off -- BASELINE.json configs[0]/[1]
    "setting": {"name": "test_kitti", "output_root": "./experiments", "kitti_correct": True, "correct_deg": 0.195},
    "process": {"min_range_m": 3.0, "max_range_m": 80.0, "vox_down_m": 0.08, "min_z_m": -3.5},
    "sampler": {"surface_sample_range_m": 0.25, "surface_sample_n": 4, "free_sample_begin_ratio": 0.3,
                "free_sample_end_dist_m": 1.0, "free_front_sample_n": 2},
    "neuralpoints": {"voxel_size_m": 0.4, "feature_dim": 8, "query_nn_k": 6, "search_alpha": 0.5,
                     "weighted_first": False},
    "decoder": {"freeze_after_frame": 30},
    "loss": {"main_loss_type": "bce", "sigma_sigmoid_m": 0.08, "loss_weight_on": True, "dist_weight_scale": 0.8,
             "ekional_loss_on": True, "weight_e": 0.5},
    "continual": {"batch_size_new_sample": 2000, "pool_capacity": 2e7},
    "tracker": {"source_vox_down_m": 0.8, "iter_n": 100, "GM_grad": 0.1, "GM_dist": 0.2},
    "optimizer": {"iters": 15, "batch_size": 16384},
    "eval": {"wandb_vis_on": False, "o3d_vis_on": False, "silence_log": True, "mesh_freq_frame": 50,
             "mesh_min_nn": 15, "save_map": False},
}


def kitti_config(device: str = "cuda") -> Config:
    cfg = Config()
    cfg.load_dict(RUN_KITTI)
    cfg.device = device
    return cfg
