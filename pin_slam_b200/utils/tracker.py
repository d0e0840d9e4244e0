"""Tracker -- point-to-implicit-model registration (reference: utils/tracker.py:19-520).

Same constructor, `tracking()` signature, return tuple and validity control flow as the reference.
Per Gauss-Newton iteration the host makes ONE call (pinq_tracking_step) that enqueues four launches --
transform by the current pose (passed by value), search + decode (pinq_query_fused_fwd: sdf, d sdf/dx, std,
nn_count in one pass, replacing query_feature -> Decoder.sdf -> get_gradient) and the fused
mask/weights/normal-equation reduction (pinq_gn_accumulate) -- and ONE 256-byte device->host read into
pinned memory; the 6x6 solve runs on the host in fp64 exactly as the reference does (tracker.py:491).  Colour / normal / semantic inputs are out of scope.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np
import torch

from .. import _pinq as P


class Tracker:
    def __init__(self, config, neural_points, geo_decoder, sem_decoder, color_decoder):
        self.config = config
        self.silence = config.silence
        self.neural_points = neural_points
        self.geo_decoder = geo_decoder
        self.sem_decoder = sem_decoder
        self.color_decoder = color_decoder
        self.device = config.device
        self.dtype = config.dtype
        self.sdf_scale = config.logistic_gaussian_ratio * config.sigma_sigmoid_m
        self._gn_out = torch.zeros(32, dtype=torch.float64, device=self.device)
        self._gn_host = torch.zeros(32, dtype=torch.float64).pin_memory()
        self._gn_host_np = self._gn_host.numpy()
        self._t12 = np.zeros((3, 4), dtype=np.float32)
        self._t12_ptr = self._t12.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        self._step_buf, self._label_keepalive = {}, None
        self.last_iterations = 0
        # device-side Gauss-Newton loop (pinq_tracking_run): pose + bookkeeping live in 32 doubles on the device
        self.device_loop = True      # False: one host round trip per iteration (A/B tests; always used with vis_result)
        self.loop_chunk = 8          # iterations enqueued per stream synchronise
        self._ts_dev = torch.zeros(P.TRACK_STATE_DOUBLES, dtype=torch.float64, device=self.device)
        self._ts_in = torch.zeros(P.TRACK_STATE_DOUBLES, dtype=torch.float64).pin_memory()
        self._ts_out = torch.zeros(P.TRACK_STATE_DOUBLES, dtype=torch.float64).pin_memory()
        self._ts_in_np, self._ts_out_np = self._ts_in.numpy(), self._ts_out.numpy()
        self._ts_in_flags, self._ts_out_flags = self._ts_in_np[24:].view(np.int32), self._ts_out_np[24:].view(np.int32)

    # ------------------------------------------------------------------------------------------
    def tracking(self, source_points, init_pose=None, source_colors=None, source_normals=None, source_semantics=None,
                 source_sdf=None, cur_ts=None, loop_reg: bool = False, vis_result: bool = False):
        """reference: utils/tracker.py:39-174 -> (T [4,4] f64, cov_mat | None, weight_pc | None, valid_flag)."""
        if source_colors is not None or source_normals is not None:
            raise NotImplementedError("photometric / normal-weighted registration is out of scope")
        cfg = self.config
        if init_pose is None:
            init_pose = torch.eye(4, dtype=torch.float64, device=self.device)
        T = init_pose.detach().to("cpu", torch.float64).numpy().copy()
        cov_mat = None
        min_grad_norm, max_grad_norm = cfg.reg_min_grad_norm, cfg.reg_max_grad_norm
        GM_dist = cfg.reg_GM_dist_m if cfg.reg_GM_dist_m > 0 else None
        GM_grad = cfg.reg_GM_grad if cfg.reg_GM_grad > 0 else None
        iter_n = cfg.reg_iter_n
        max_valid_final_sdf_residual_cm = cfg.surface_sample_range_m * 0.5 * 100.0
        min_valid_ratio = 0.2
        if loop_reg:
            max_valid_final_sdf_residual_cm = cfg.surface_sample_range_m * 0.6 * 100.0
            min_valid_ratio = 0.15
        max_increment_sdf_residual_ratio = 1.1
        eigenvalue_ratio_thre = 0.01
        min_valid_points = 30
        converged = False
        valid_flag = True
        last_sdf_residual_cm = 1e5
        src = source_points.detach().to(self.dtype).contiguous()
        source_point_count = src.shape[0]
        sdf_residual_cm = 0.0
        valid_point_count = 0
        eigenvalues = None
        cur_points = torch.empty_like(src)
        step_args = self._tracking_args(cur_points, src, source_sdf, min_grad_norm, max_grad_norm, GM_dist, GM_grad)
        it = 0
        if self.device_loop and not vis_result:
            # chunks of iterations run on the device without a host round trip (csrc/gn.cu: k_track_transform, fused query,
            # k_gn_accumulate, k_track_solve); the loop's break conditions are evaluated there
            st_in, fl_in = self._ts_in_np, self._ts_in_flags
            st_in[:] = 0.0
            st_in[:16] = T.reshape(-1)
            st_in[16] = last_sdf_residual_cm
            fl_in[1] = 1  # valid_flag
            loop = P.TrackingLoop()
            loop.state, loop.h_state = P.ptr(self._ts_dev), self._ts_out.data_ptr()
            loop.iter_n, loop.min_valid_points = int(iter_n), int(min_valid_points)
            loop.lm_lambda, loop.term_thre_deg, loop.term_thre_m = float(cfg.reg_lm_lambda), float(cfg.reg_term_thre_deg), float(cfg.reg_term_thre_m)
            loop.min_valid_ratio, loop.max_increment_ratio = float(min_valid_ratio), float(max_increment_sdf_residual_ratio)
            lib, stream = P.load(), torch.cuda.current_stream()
            first = 0
            while first < iter_n:
                loop.first_iter, loop.iters = first, min(int(self.loop_chunk), iter_n - first)
                loop.h_state_in = self._ts_in.data_ptr() if first == 0 else None
                P.check(lib.pinq_tracking_run(step_args, loop, stream.cuda_stream))
                stream.synchronize()
                first += loop.iters
                if self._ts_out_flags[0]:
                    break
            fl = self._ts_out_flags
            T = self._ts_out_np[:16].reshape(4, 4).copy()
            it, valid_point_count, sdf_residual_cm = int(fl[3]), int(fl[4]), float(self._ts_out_np[17])
            valid_flag = bool(fl[1])
            if not valid_flag and not self.silence:
                print("(Warning) registration failed: " + ("wrong optimization" if fl[5] == 1 else "not enough valid points"))
            iter_n = 0  # the host loop below is skipped
        for i in range(iter_n):
            it = i + 1
            # one call per iteration: transform by T, fused query (sdf, gradient, std), normal-equation reduction
            reg = self._gauss_newton_step(step_args, T, cfg.reg_lm_lambda, (vis_result and converged))
            delta_T, cov_mat, eigenvalues, _, valid_point_count, sdf_residual_cm, _ = reg
            T = delta_T @ T
            if (sdf_residual_cm - last_sdf_residual_cm) / last_sdf_residual_cm > max_increment_sdf_residual_ratio:
                if not self.silence:
                    print("(Warning) registration failed: wrong optimization")
                valid_flag = False
            else:
                last_sdf_residual_cm = sdf_residual_cm
            if (valid_point_count < min_valid_points) or (1.0 * valid_point_count / source_point_count < min_valid_ratio):
                if not self.silence:
                    print("(Warning) registration failed: not enough valid points")
                valid_flag = False
            if not valid_flag or converged:
                break
            with np.errstate(invalid="ignore"):
                rot_angle_deg = math.degrees(np.arccos((np.trace(delta_T[:3, :3]) - 1) / 2))
            tran_m = float(np.linalg.norm(delta_T[:3, 3]))
            # NOTE `and` binds tighter than `or` (tracker.py:135); NaN angles compare False as in torch
            if abs(rot_angle_deg) < cfg.reg_term_thre_deg and tran_m < cfg.reg_term_thre_m or i == iter_n - 2:
                converged = True
        self.last_iterations = it
        if sdf_residual_cm > max_valid_final_sdf_residual_cm:
            if not self.silence:
                print("(Warning) registration failed: too large final residual")
            valid_flag = False
        if eigenvalues is not None:
            min_eigenvalue = float(np.min(eigenvalues))
            if cfg.eigenvalue_check and min_eigenvalue < valid_point_count * eigenvalue_ratio_thre:
                valid_flag = False
        T_out = torch.tensor(T, dtype=torch.float64, device=self.device)
        if not valid_flag:  # NOTE: if not valid, just take the initial guess (tracker.py:170-172)
            T_out = init_pose
            cov_mat = None
        return T_out, cov_mat, None, valid_flag

    # ------------------------------------------------------------------------------------------
    def query_source_points(self, coord, ts, bs, query_sdf=True, query_sdf_grad=True, query_color=False,
                            query_color_grad=False, query_sem=False, query_mask=True, query_certainty=True,
                            query_locally=True, mask_min_nn_count: int = 4):
        """reference: utils/tracker.py:176-275 -> (sdf, sdf_grad, None, None, None, mc_mask, certainty, sdf_std)."""
        if query_color or query_sem or query_color_grad:
            raise NotImplementedError("colour / semantic queries are out of scope")
        n = coord.shape[0]
        dev = coord.device
        sdf_pred = torch.zeros(n, device=dev)
        sdf_std = torch.zeros(n, device=dev)
        sdf_grad = torch.zeros((n, 3), device=dev) if query_sdf_grad else None
        mc_mask = torch.zeros(n, device=dev, dtype=torch.bool) if query_mask else None
        certainty = torch.zeros(n, device=dev) if query_certainty else None
        with torch.no_grad():
            for head in range(0, n, bs):
                tail = min(head + bs, n)
                out = self.neural_points.query_sdf(coord[head:tail], self.geo_decoder, None, False, query_locally,
                                                   want_grad=query_sdf_grad, want_std=True,
                                                   want_certainty=query_certainty)
                sdf_pred[head:tail] = out["sdf"]
                sdf_std[head:tail] = out["std"]
                if query_sdf_grad:
                    sdf_grad[head:tail] = out["grad"]
                if query_mask:
                    mc_mask[head:tail] = out["nn_count"] >= mask_min_nn_count
                if query_certainty:
                    certainty[head:tail] = out["certainty"]
        return sdf_pred, sdf_grad, None, None, None, mc_mask, certainty, sdf_std

    def _tracking_args(self, xyz, source, sdf_labels, min_grad_norm, max_grad_norm, GM_dist, GM_grad):
        """pinq_tracking_args for `xyz` [n,3] (the buffer the fused query reads); built once per tracking() call --
        the map does not change between Gauss-Newton iterations."""
        cfg, npts, dev = self.config, self.neural_points, self.device
        n = xyz.shape[0]
        if self._step_buf.get("n") != n:
            self._step_buf = {"n": n, "sdf": torch.empty(n, dtype=torch.float32, device=dev),
                              "grad": torch.empty((n, 3), dtype=torch.float32, device=dev),
                              "std": torch.empty(n, dtype=torch.float32, device=dev),
                              "nn": torch.empty(n, dtype=torch.int32, device=dev),
                              "perm": torch.empty(n, dtype=torch.int32, device=dev) if n >= npts.bin_queries_min else None}
        b = self._step_buf
        a = npts._query_args(xyz, self.geo_decoder.view(), None, False, True)
        a.sdf, a.grad_x, a.sdf_std, a.nn_count, a.perm = (P.ptr(b["sdf"]), P.ptr(b["grad"]), P.ptr(b["std"]), P.ptr(b["nn"]),
                                                         P.ptr(b["perm"]))
        t = P.TrackingArgs()
        t.query = a
        t.source_xyz = P.ptr(source)
        self._label_keepalive = sdf_labels.detach().float().contiguous() if sdf_labels is not None else None
        t.sdf_label = P.ptr(self._label_keepalive)
        t.min_nn = int(cfg.query_nn_k)
        t.min_grad_norm, t.max_grad_norm = float(min_grad_norm), float(max_grad_norm)
        t.max_sdf_std = float(cfg.surface_sample_range_m * cfg.max_sdf_std_ratio)
        t.gm_dist, t.gm_grad = float(GM_dist or 0.0), float(GM_grad or 0.0)
        t.gn_out, t.h_gn_out = P.ptr(self._gn_out), self._gn_host.data_ptr()
        return t

    def _gauss_newton_step(self, step_args, T, lm_lambda, vis_weight_pc=False):
        """Enqueue pinq_tracking_step (pose T applied to the source points; None: points as given), wait for the
        32 doubles and solve on the host."""
        t12 = None
        if T is not None:
            self._t12[:] = T[:3, :]  # float64 -> float32 as the reference's tensor cast (tracker.py:94)
            t12 = self._t12_ptr
        P.check(P.load().pinq_tracking_step(step_args, t12, P.stream()))
        torch.cuda.current_stream().synchronize()
        g = self._gn_host_np
        valid_point_count = int(round(g[29]))
        if valid_point_count < 10:  # tracker.py:310-312
            return np.eye(4), None, None, None, valid_point_count, 0.0, 0.0
        residual_cm = g[28] / valid_point_count * 100.0
        Td, cov_mat, eig = solve_normal_equations(g, valid_point_count, lm_lambda, vis_weight_pc, vis_weight_pc)
        return Td, cov_mat, eig, None, valid_point_count, residual_cm, None

    def registration_step(self, points, normals, sdf_labels, colors, cur_ts, min_grad_norm, max_grad_norm, GM_dist=None,
                          GM_grad=None, lm_lambda=0.0, vis_weight_pc=False):
        """reference: utils/tracker.py:277-452 + implicit_reg :468-520.
        -> (delta_T ndarray[4,4] f64, cov_mat, eigenvalues, None, valid_point_count, residual_cm, None)"""
        pts = points.detach().to(self.dtype).contiguous()
        args = self._tracking_args(pts, None, sdf_labels, min_grad_norm, max_grad_norm, GM_dist, GM_grad)
        return self._gauss_newton_step(args, None, lm_lambda, vis_weight_pc)


def solve_normal_equations(g: np.ndarray, count: int, lm_lambda: float, require_cov=False, require_eigen=False):
    """Host side of implicit_reg (utils/tracker.py:468-520) from the fused reduction of pinq_gn_accumulate:
    g[0:21] = upper triangle of sum w J^T J, g[21:27] = sum w r J, g[27] = sum w  (w un-normalised)."""
    norm = 2.0 * g[27] / count  # w /= 2 mean(w), tracker.py:394
    H = np.zeros((6, 6))
    iu = np.triu_indices(6)
    H[iu] = g[:21]
    H = H + H.T - np.diag(np.diag(H))
    H /= norm
    b = -g[21:27] / norm
    H0 = H.copy()
    H = H + lm_lambda * np.diag(np.diag(H))
    t = np.linalg.inv(H) @ b
    T = np.eye(4)
    T[:3, :3] = expmap(t[:3])
    T[:3, 3] = t[3:]
    eig = np.linalg.eigvals(H0[3:, 3:]).real if require_eigen else None
    cov = None  # needs sum w r^2 (visualisation only, tracker.py:507-510): not accumulated
    return T, cov, eig


def skew(v):
    """reference: utils/tracker.py:573-578."""
    return np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])


def expmap(axis_angle: np.ndarray):
    """Rodrigues (reference: utils/tracker.py:580-589); a zero rotation yields NaNs there as here."""
    angle = np.linalg.norm(axis_angle)
    with np.errstate(invalid="ignore", divide="ignore"):
        S = skew(axis_angle / angle)
    return np.eye(3) + S * np.sin(angle) + (S @ S) * (1.0 - np.cos(angle))
