"""DataSampler -- 1 + surface_n + front_n + behind_n samples along every LiDAR ray with projective SDF
labels and weights (reference: utils/data_sampler.py:20-192).  On CUDA tensors the ~40 element-wise
torch ops of the reference are one kernel (`pinq_ray_samples`, csrc/pool.cu) fed with the same three random tensors the
reference draws, in the same order; `sample_torch` is that formulation restated with torch ops (parity partner, and the
form the CPU oracle is fed from).  A torch.Generator may be passed so that the CUDA path and the CPU oracle can be fed
identical samples."""
from __future__ import annotations

import torch

from .. import _pinq as P


class DataSampler:
    def __init__(self, config):
        self.config = config
        self.dev = config.device
        self.fused = True  # False: the torch-op formulation on every device (A/B tests)

    def sample(self, points_torch, normal_torch=None, sem_label_torch=None, color_torch=None, generator=None):
        if not (self.fused and points_torch.is_cuda):
            return self.sample_torch(points_torch, normal_torch, sem_label_torch, color_torch, generator)
        cfg = self.config
        dev = points_torch.device
        rng_dev = generator.device if generator is not None else dev
        pts_in = points_torch.detach().float().contiguous()
        n = pts_in.shape[0]
        n_surf, n_front, n_behind = int(cfg.surface_sample_n), int(cfg.free_front_n), int(cfg.free_behind_n)
        n_all = 1 + n_surf + n_front + n_behind
        # the reference's draws, in its order: randn(n * n_surf), rand(n * n_front), rand(n * n_behind)
        r1 = torch.randn(n * n_surf, 1, device=rng_dev, generator=generator).to(dev)
        r2 = torch.rand(n * n_front, 1, device=rng_dev, generator=generator).to(dev)
        r3 = torch.rand(n * n_behind, 1, device=rng_dev, generator=generator).to(dev)
        pts = torch.empty((n * n_all, 3), dtype=torch.float32, device=dev)
        label = torch.empty(n * n_all, dtype=torch.float32, device=dev)
        weight = torch.empty(n * n_all, dtype=torch.float32, device=dev)
        P.check(P.load().pinq_ray_samples(P.ptr(pts_in), n, P.ptr(r1) if n_surf else None, P.ptr(r2) if n_front else None,
                                          P.ptr(r3) if n_behind else None, n_surf, n_front, n_behind,
                                          float(cfg.surface_sample_range_m), float(cfg.free_sample_begin_ratio),
                                          float(cfg.free_sample_end_dist_m), float(cfg.max_range), float(cfg.dist_weight_scale),
                                          int(bool(cfg.dist_weight_on)), int(bool(cfg.behind_dropoff_on)),
                                          P.ptr(pts), P.ptr(label), P.ptr(weight), P.stream()))
        return pts, label, None, None, None, weight

    def sample_torch(self, points_torch, normal_torch=None, sem_label_torch=None, color_torch=None, generator=None):
        cfg = self.config
        dev = points_torch.device
        rng_dev = generator.device if generator is not None else dev
        n = points_torch.shape[0]
        n_surf, n_front, n_behind = cfg.surface_sample_n, cfg.free_front_n, cfg.free_behind_n
        n_all = 1 + n_surf + n_front + n_behind
        sigma_ratio = 2.0
        dist = torch.linalg.norm(points_torch, dim=1, keepdim=True)

        def randn(k):
            return torch.randn(k, 1, device=rng_dev, generator=generator).to(dev)

        def rand(k):
            return torch.rand(k, 1, device=rng_dev, generator=generator).to(dev)

        # part 0: the measured point; part 1: gaussian around the surface (data_sampler.py:46-55)
        disp0 = torch.zeros_like(dist)
        ratio0 = torch.ones_like(dist)
        disp1 = randn(n * n_surf) * cfg.surface_sample_range_m
        ratio1 = disp1 / dist.repeat(n_surf, 1) + 1.0
        # part 2: free space in front (data_sampler.py:76-83)
        d2 = dist.repeat(n_front, 1)
        fmax = 1.0 - sigma_ratio * cfg.surface_sample_range_m / d2
        ratio2 = rand(n * n_front) * (fmax - cfg.free_sample_begin_ratio) + cfg.free_sample_begin_ratio
        disp2 = (ratio2 - 1.0) * d2
        # part 3: free space behind (data_sampler.py:90-97)
        d3 = dist.repeat(n_behind, 1)
        bmax = cfg.free_sample_end_dist_m / d3 + 1.0
        bmin = 1.0 + sigma_ratio * cfg.surface_sample_range_m / d3
        ratio3 = rand(n * n_behind) * (bmax - bmin) + bmin
        disp3 = (ratio3 - 1.0) * d3

        disp = torch.cat((disp0, disp1, disp2, disp3), 0)
        ratio = torch.cat((ratio0, ratio1, ratio2, ratio3), 0)
        rep_dist = dist.repeat(n_all, 1)
        pts = points_torch.repeat(n_all, 1) * ratio
        weight = torch.ones_like(rep_dist)
        n_surface_total = n * (n_surf + 1)
        if cfg.dist_weight_on:  # data_sampler.py:127-128
            weight[:n_surface_total] = 1 + cfg.dist_weight_scale * 0.5 - (rep_dist[:n_surface_total] / cfg.max_range) * cfg.dist_weight_scale
        if cfg.behind_dropoff_on:
            dmin, dmax = 0.2 * cfg.free_sample_end_dist_m, cfg.free_sample_end_dist_m
            weight = weight * (torch.clamp((dmax - disp) / (dmax - dmin), min=0.0, max=1.0) * 0.8 + 0.2)
        weight[n_surface_total:] *= -1.0  # sign flags free-space samples
        sdf_label = disp.squeeze(1)
        # ray-wise order (data_sampler.py:169-175)
        pts = pts.reshape(n_all, -1, 3).transpose(0, 1).reshape(-1, 3)
        sdf_label = sdf_label.reshape(n_all, -1).transpose(0, 1).reshape(-1) * (-1)
        weight = weight.reshape(n_all, -1).transpose(0, 1).reshape(-1)
        return pts.contiguous(), sdf_label.contiguous(), None, None, None, weight.contiguous()
