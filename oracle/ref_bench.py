"""TEST INFRASTRUCTURE ONLY -- runs the UNMODIFIED reference (through oracle/ref_shim.py; /root/reference here, the staged
byte-for-byte copy oracle/_ref/ on the GPU box) on the bench workload, for the timed baselines of bench.py:

  * `bench.py --impl reference`         the reference's own CPU path (kind = "reference"), map built by its own update()
  * `cpu_baseline` of the GPU arm       the same chain on a bounded sample
  * `reference_cuda_eager`              the same unmodified code with device = cuda on the same B200 (stock PyTorch eager)

Nothing under pin_slam_b200/ imports this module and nothing here calls libpinq.

The chain timed is the reference's mapping query (utils/mapper.py:461-468 + 572):
    query_feature(x, ts)  ->  Decoder.sdf  ->  sum_k w_k s_k  ->  backward()
"""
from __future__ import annotations

import os
import time

import torch

from . import ref_shim


def reference_config(device, local_map_radius=1.0e4):
    cfg = ref_shim.kitti_config(device)
    cfg.local_map_radius = local_map_radius  # microbench: the whole synthetic map is the local map (config 3)
    cfg.silence = True
    return cfg


def _assign_features(npts, device):
    """Same seeded features / certainties as bench.build_map."""
    g = torch.Generator(device="cpu").manual_seed(0)
    m = npts.neural_points.shape[0]
    feats = torch.randn((m + 1, npts.geo_feature_dim), generator=g) * 0.1
    feats[-1] = 0
    cert = torch.rand(m, generator=g) * 50
    npts.geo_features = feats.to(device)
    npts.point_certainties = cert.to(device)


def build_reference_map(surface_points_np, device="cpu"):
    """Map of the microbench built by the reference's own NeuralPoints.update (model/neural_points.py:205-270)."""
    ns = ref_shim.load()
    cfg = reference_config(device)
    npts = ns.NeuralPoints(cfg)
    npts.travel_dist = torch.zeros(2, device=device)
    raw = torch.from_numpy(surface_points_np).to(device)
    npts.update(raw, torch.zeros(3, device=device), torch.eye(3, device=device), 0)
    del raw
    _assign_features(npts, device)
    npts.reset_local_map(torch.zeros(3, device=device), torch.eye(3, device=device), 0)
    torch.manual_seed(42)
    dec = ns.Decoder(cfg, cfg.geo_mlp_hidden_dim, cfg.geo_mlp_level, 1)
    return cfg, npts, dec


def reference_map_from_state(state, device):
    """Reference NeuralPoints holding a map given as tensors (`state`: dict with the reference's attribute names) --
    used by the comparator legs of the GPU arm so that they time the reference's query on the very same map."""
    ns = ref_shim.load()
    cfg = reference_config(device)
    cfg.buffer_size = int(state["buffer_pt_index"].shape[0])
    npts = ns.NeuralPoints(cfg)
    npts.travel_dist = state["travel_dist"].to(device)
    npts.buffer_pt_index = state["buffer_pt_index"].to(device=device, dtype=torch.int64)
    for k in ("neural_points", "point_orientations", "geo_features", "point_ts_create", "point_ts_update", "point_certainties"):
        setattr(npts, k, state[k].to(device))
    npts.reset_local_map(torch.zeros(3, device=device), torch.eye(3, device=device), 0)
    torch.manual_seed(42)
    dec = ns.Decoder(cfg, cfg.geo_mlp_hidden_dim, cfg.geo_mlp_level, 1)
    with torch.no_grad():
        dec.layers[0].weight.copy_(state["W1"].to(device))
        dec.layers[0].bias.copy_(state["b1"].to(device))
        dec.lout.weight.copy_(state["W2"].to(device).reshape(dec.lout.weight.shape))
        dec.lout.bias.copy_(state["b2"].to(device).reshape(dec.lout.bias.shape))
    return cfg, npts, dec


def mean_nonempty_slots(npts, q, sample=100_000):
    """V-bar of SURVEY.md 8(d) from the reference's own table: mean number of non-empty probed slots per query."""
    x = q[:sample]
    cells = torch.floor(x / npts.resolution).to(torch.int64)[:, None, :] + npts.neighbor_dx[None]
    h = torch.fmod((cells * npts.primes).sum(-1), npts.buffer_size)
    h = torch.where(h < 0, h + npts.buffer_size, h)
    return float((npts.buffer_pt_index[h] >= 0).sum(1).float().mean())


def reference_step(npts, dec, x, ts):
    """fwd + bwd of the unmodified reference chain; returns the predicted sdf [n]."""
    npts.local_geo_features.grad = None
    for p in dec.parameters():
        p.grad = None
    geo_feature, _, weight_knn, _, _ = npts.query_feature(x, ts)          # utils/mapper.py:461
    sdf_pred = dec.sdf(geo_feature)                                       # :466
    sdf_pred = torch.sum(sdf_pred * weight_knn, dim=1).squeeze(1)         # :468
    sdf_pred.backward(torch.ones_like(sdf_pred))                          # :572 with upstream gradient = ones
    return sdf_pred.detach()


def time_reference(npts, dec, x, ts, steps, warmup, device="cpu"):
    """(queries/s, seconds per step) of `reference_step` on `x` -- wall clock on CPU, synchronised wall clock on cuda."""
    if device == "cpu":
        torch.set_num_threads(os.cpu_count() or 1)
    sync = torch.cuda.synchronize if str(device).startswith("cuda") else (lambda: None)
    for _ in range(warmup):
        reference_step(npts, dec, x, ts)
    sync()
    t0 = time.perf_counter()
    for _ in range(steps):
        reference_step(npts, dec, x, ts)
    sync()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return x.shape[0] / dt, dt
