"""Math helpers on the neural-point query path, backed by libpinq kernels.

Mirrors the names of the reference's utils/tools.py for the functions that sit on the hot path
(voxel_down_sample_torch :409, voxel_down_sample_min_value_torch :444, transform_torch :386, deskewing :540,
get_time :269, freeze_model :187, setup_optimizer :89).  Everything here requires CUDA tensors.
"""
from __future__ import annotations

import time

import torch
import torch.nn as nn
from torch import optim

from .. import _pinq as P


def get_time():
    """Stage-level wall clock (reference: utils/tools.py:269-274).  Kept out of the kernels' way: the
    fused path never calls it between launches."""
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    return time.time()


def _next_pow2(n: int) -> int:
    return 1 << max(1, (int(n) - 1).bit_length())


def _voxel_down_sample(points: torch.Tensor, voxel_size: float, value, mode: int) -> torch.Tensor:
    lib = P.load()
    pts = points.detach()
    if pts.dtype != torch.float32:
        pts = pts.float()
    pts = pts.contiguous()
    n = pts.shape[0]
    dev = pts.device
    if n == 0:
        return torch.empty(0, dtype=torch.int64, device=dev)
    st = P.stream()
    scratch = torch.empty(8, dtype=torch.float32, device=dev)
    keys = torch.empty(n, dtype=torch.int64, device=dev)
    q = torch.empty(n, dtype=torch.int32, device=dev)
    val = None
    if mode == 1:
        val = value.detach().float().contiguous()
    P.check(lib.pinq_vds_prepare(P.ptr(pts), n, float(voxel_size), P.ptr(val), mode, P.ptr(scratch), P.ptr(keys),
                                 P.ptr(q), st))
    cap = _next_pow2(2 * n)
    tab_k = torch.empty(cap, dtype=torch.int64, device=dev)
    tab_b = torch.empty(cap, dtype=torch.int64, device=dev)
    P.check(lib.pinq_vds_select(P.ptr(keys), P.ptr(q), n, P.ptr(tab_k), P.ptr(tab_b), cap, st))
    out_idx = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.empty(1, dtype=torch.int32, device=dev)
    P.check(lib.pinq_vds_emit(P.ptr(tab_k), P.ptr(tab_b), cap, n, P.ptr(keys), P.ptr(out_idx), P.ptr(cnt), st))
    m = int(cnt.item())
    # the reference returns winners in ascending order of the unique voxel key (torch.unique sorts)
    order = torch.argsort(keys[:m])
    return out_idx[:m][order].to(torch.int64)


def voxel_down_sample_torch(points: torch.Tensor, voxel_size: float) -> torch.Tensor:
    """Index of the point closest to each voxel centre (reference: utils/tools.py:409-442)."""
    return _voxel_down_sample(points, voxel_size, None, 0)


def voxel_down_sample_min_value_torch(points: torch.Tensor, voxel_size: float, value: torch.Tensor) -> torch.Tensor:
    """Index of the point with the smallest quantised `value` per voxel (reference: utils/tools.py:444-477)."""
    return _voxel_down_sample(points, voxel_size, value, 1)


def transform_torch(points: torch.Tensor, transformation: torch.Tensor) -> torch.Tensor:
    """points [N,3] fp32, transformation [4,4] (any dtype) -> [N,3] (reference: utils/tools.py:386-399)."""
    lib = P.load()
    pts = points.detach().contiguous()
    T12 = transformation.detach().to(device=pts.device, dtype=torch.float32)[:3, :].contiguous()
    out = torch.empty_like(pts)
    P.check(lib.pinq_transform_points(P.ptr(pts), pts.shape[0], P.ptr(T12), P.ptr(out), P.stream()))
    return out


def deskewing(points: torch.Tensor, ts: torch.Tensor, pose: torch.Tensor, ts_mid_pose: float = 0.5) -> torch.Tensor:
    """Motion undistortion of a sweep, in place like the reference (utils/tools.py:540-567): ts normalised to [0, 1] over the
    cloud, tau = ts - ts_mid_pose, p <- slerp(I, R, tau) p + tau t with pose = T_last<-cur.  One kernel (pinq_deskew)."""
    if ts is None:
        return points
    lib = P.load()
    if points.dtype != torch.float32 or not points.is_contiguous():
        raise P.PinqError("deskewing works in place on a contiguous float32 cloud")
    t = ts.detach().reshape(-1).to(torch.float32).contiguous()
    n = points.shape[0]
    if n == 0:
        return points
    lo, hi = torch.aminmax(t)
    minmax = torch.stack((lo, hi))
    T = pose.detach().to(device=points.device, dtype=torch.float32).contiguous()
    P.check(lib.pinq_deskew(P.ptr(points), n, points.shape[1], P.ptr(t), P.ptr(minmax), P.ptr(T), float(ts_mid_pose), P.stream()))
    return points


def velodyne_point_ts(points: torch.Tensor) -> torch.Tensor:
    """Per-point time in [0, 1] of a scan without time stamps, Velodyne convention (dataset/slam_dataset.py:247-249)."""
    lib = P.load()
    pts = points.detach()
    if pts.dtype != torch.float32 or not pts.is_contiguous():
        pts = pts.float().contiguous()
    ts = torch.empty(pts.shape[0], dtype=torch.float32, device=pts.device)
    P.check(lib.pinq_velodyne_point_ts(P.ptr(pts), pts.shape[0], pts.shape[1], P.ptr(ts), P.stream()))
    return ts


def transform_batch_torch(points: torch.Tensor, transformation: torch.Tensor) -> torch.Tensor:
    """Per-point transforms [N,4,4] (reference: utils/tools.py:401-407); off the hot path -> torch ops."""
    p = torch.matmul(transformation[:, :3, :3].to(points), points.unsqueeze(-1)) + transformation[:, :3, 3:].to(points)
    return p.squeeze(-1)


def apply_quaternion_rotation(quat: torch.Tensor, points: torch.Tensor) -> torch.Tensor:
    """Passive rotation p' = q p q^-1 (reference: utils/tools.py:316-323)."""
    w = quat[..., 0].unsqueeze(-1)
    u = -quat[..., 1:]
    t = 2 * torch.linalg.cross(u, points)
    return points + w * t + torch.linalg.cross(u, t)


def freeze_model(model: nn.Module):
    """reference: utils/tools.py:187-190 (hazard H11: the backward checks requires_grad per call)."""
    for child in model.children():
        for param in child.parameters():
            param.requires_grad = False


def unfreeze_model(model: nn.Module):
    for child in model.children():
        for param in child.parameters():
            param.requires_grad = True


def freeze_decoders(geo_decoder, sem_decoder, color_decoder, config):
    freeze_model(geo_decoder)


def setup_optimizer(config, neural_point_feat, mlp_geo_param=None, mlp_sem_param=None, mlp_color_param=None,
                    poses=None, lr_ratio=1.0):
    """reference: utils/tools.py:89-116 -- fresh Adam(betas=(0.9, 0.99), eps=adam_eps) over decoder + features."""
    lr_cur = config.lr * lr_ratio
    groups = []
    if mlp_geo_param is not None:
        groups.append({"params": mlp_geo_param, "lr": lr_cur, "weight_decay": config.weight_decay})
    if poses is not None:
        groups.append({"params": poses, "lr": config.lr_pose, "weight_decay": config.weight_decay})
    groups.append({"params": neural_point_feat, "lr": lr_cur, "weight_decay": config.weight_decay})
    if config.opt_adam:
        return optim.Adam(groups, betas=(0.9, 0.99), eps=config.adam_eps)
    return optim.SGD(groups, momentum=0.9)
