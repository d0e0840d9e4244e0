"""TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of PIN-SLAM's neural-point SDF query path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module, and only as the checker or the
timed CPU baseline -- never as part of the product path (``pin_slam_b200``), which
must fail loudly when its CUDA library is missing.

The algorithm is restated from the reference (kelly7707/PIN_SLAM, all Python/PyTorch;
citations are ``path:line`` relative to the reference root).  It is written as plain
functions over a ``MapState`` record using torch **CPU** fp32/int64 tensor arithmetic,
because the reference's observable results are defined by exactly these ATen CPU
semantics (true fp32 division, no FMA contraction, sequential ``index_put_``).

Parity pin: the reference publishes no tests or golden vectors,
so the pin is the reference code itself.  ``oracle/make_golden.py`` imports the
unmodified reference through ``oracle/ref_shim.py`` in the build container and
writes ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks this file
against those vectors (bit-exact for slots / indices / counts, 1e-6 for floats), and
``tests/test_oracle_vs_reference.py`` re-checks live whenever ``/root/reference`` exists.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Optional

import torch

PRIMES = (73856093, 19349669, 83492791)  # model/neural_points.py:69
I64 = torch.int64


# --------------------------------------------------------------------------------------
# state
# --------------------------------------------------------------------------------------
@dataclass
class MapParams:
    """The subset of utils/config.py that the path reads (values = run_kitti.yaml resolved)."""

    voxel_size_m: float = 0.4
    buffer_size: int = 50_000_000
    num_nei_cells: int = 2
    search_alpha: float = 0.5
    query_nn_k: int = 6
    feature_dim: int = 8
    feature_std: float = 0.0
    weighted_first: bool = False
    local_map_radius: float = 82.0
    local_map_travel_dist_ratio: float = 5.0
    use_mid_ts: bool = False
    sdf_scale: float = 0.55 * 0.08  # model/decoder.py:51-52


@dataclass
class MapState:
    p: MapParams
    table: torch.Tensor = None            # buffer_pt_index [B] int64, -1 = empty          neural_points.py:73
    points: torch.Tensor = None           # neural_points [M,3] f32                          :76
    quats: torch.Tensor = None            # point_orientations [M,4] (w,x,y,z)               :77
    feats: torch.Tensor = None            # geo_features [M+1,F], last row = padding         :78
    ts_create: torch.Tensor = None        # [M] int64                                        :83
    ts_update: torch.Tensor = None        # [M] int64                                        :84
    certainty: torch.Tensor = None        # point_certainties [M] f32                        :85
    l_points: torch.Tensor = None         # local_* copies                                   :88-93
    l_quats: torch.Tensor = None
    l_feats: torch.Tensor = None          # [Ml+1,F]
    l_certainty: torch.Tensor = None
    l_ts_update: torch.Tensor = None
    local_mask: torch.Tensor = None       # [M+1] bool
    global2local: torch.Tensor = None     # [M+1] int64
    travel_dist: torch.Tensor = None      # [frames] f32, set by the driver (pin_slam.py:134)
    cur_ts: int = 0
    after_pgo: bool = False
    temporal_local_map_on: bool = True    # :53
    neighbor_dx: torch.Tensor = None      # [Kn,3] int64
    max_valid_dist2: float = 0.0
    extra: dict = field(default_factory=dict)

    @property
    def diff_travel_dist_local(self) -> float:   # neural_points.py:55
        return self.p.local_map_radius * self.p.local_map_travel_dist_ratio


def new_map(p: MapParams) -> MapState:
    """model/neural_points.py:69-98 -- empty map."""
    s = MapState(p=p)
    s.table = torch.full((int(p.buffer_size),), -1, dtype=I64)
    s.points = torch.empty((0, 3))
    s.quats = torch.empty((0, 4))
    s.feats = torch.zeros((1, p.feature_dim))
    s.ts_create = torch.empty((0,), dtype=I64)
    s.ts_update = torch.empty((0,), dtype=I64)
    s.certainty = torch.empty((0,))
    set_search_neighborhood(s, p.num_nei_cells, p.search_alpha)
    return s


def set_search_neighborhood(s: MapState, num_nei_cells: int, search_alpha: float) -> None:
    """model/neural_points.py:430-457 -- spherical stencil of probed cells (meshgrid 'ij' order)."""
    r = torch.arange(-num_nei_cells, num_nei_cells + 1, dtype=I64)
    gx, gy, gz = torch.meshgrid(r, r, r, indexing="ij")
    dx = torch.stack((gx, gy, gz), dim=-1).reshape(-1, 3)
    keep = (dx * dx).sum(-1) < (num_nei_cells + search_alpha) ** 2
    s.neighbor_dx = dx[keep]
    s.max_valid_dist2 = 3 * ((num_nei_cells + 1) * s.p.voxel_size_m) ** 2


# --------------------------------------------------------------------------------------
# voxel hashing
# --------------------------------------------------------------------------------------
def cells_of(points: torch.Tensor, res: float) -> torch.Tensor:
    """floor(x / res) as int64 -- true fp32 division (hazard H1).  neural_points.py:214, 465"""
    return (points / res).floor().to(I64)


def signed_hash(cells: torch.Tensor, buffer_size: int) -> torch.Tensor:
    """fmod keeps the dividend's sign (hazard H2).  neural_points.py:216, 472"""
    primes = torch.tensor(PRIMES, dtype=I64)
    return torch.fmod((cells * primes).sum(-1), int(buffer_size))


def slot_of(cells: torch.Tensor, buffer_size: int) -> torch.Tensor:
    """Effective table slot: a negative hash indexes from the end (Python/torch negative index)."""
    h = signed_hash(cells, buffer_size)
    return torch.where(h < 0, h + int(buffer_size), h)


def voxel_down_sample(points: torch.Tensor, voxel_size: float) -> torch.Tensor:
    """utils/tools.py:409-442 -- index of the point closest to each voxel centre, in ascending
    order of the (quirky, scalar-V) linear voxel key (hazard H5)."""
    lo = torch.floor(points.min(dim=0)[0] / voxel_size).long()
    g = torch.floor(points / voxel_size)
    ctr = (g + 0.5) * voxel_size
    d = ((points - ctr) ** 2).sum(dim=1) ** 0.5
    d = d / d.max() * 999
    gi = g.long() - lo
    V = gi.max()
    key = gi[:, 0] + gi[:, 1] * V + gi[:, 2] * V * V
    return _argmin_per_key(key, d.long())


def voxel_down_sample_min_value(points: torch.Tensor, voxel_size: float, value: torch.Tensor) -> torch.Tensor:
    """utils/tools.py:444-477 -- same grouping, winner = smallest quantised ``value``."""
    lo = torch.floor(points.min(dim=0)[0] / voxel_size).long()
    g = torch.floor(points / voxel_size)
    v = value / value.max() * 999
    gi = g.long() - lo
    V = gi.max()
    key = gi[:, 0] + gi[:, 1] * V + gi[:, 2] * V * V
    return _argmin_per_key(key, v.long())


def _argmin_per_key(key: torch.Tensor, q: torch.Tensor) -> torch.Tensor:
    """tools.py:433-441: decimal packing ``i + q*10^digits`` then scatter-amin then ``% 10^digits``."""
    n = key.shape[0]
    uniq, inv = torch.unique(key, return_inverse=True)
    pack_base = 10 ** len(str(n - 1))
    packed = torch.arange(n, dtype=I64) + q * pack_base
    best = torch.empty(uniq.shape, dtype=I64).scatter_reduce_(0, inv, packed, reduce="amin", include_self=False)
    return best % pack_base


def put_last_wins(table: torch.Tensor, index: torch.Tensor, values: torch.Tensor) -> None:
    """``table[index] = values`` with the semantics of a SEQUENTIAL index_put_: among duplicate indices the
    highest batch position wins (hazard H4).  The reference relies on ATen's CPU index_put_, which is only
    sequential on one thread -- with several threads (and on CUDA) duplicate slots race and the reference's
    own result varies from run to run (observed: 2-12 differing slots out of 17 k on an 8-thread run of the
    golden scenario).  The oracle, the golden vectors (generated with one thread) and the CUDA kernels all
    implement the deterministic single-thread outcome."""
    index = torch.where(index < 0, index + table.shape[0], index)
    uniq, inv = torch.unique(index, return_inverse=True)
    pos = torch.arange(index.shape[0], dtype=I64)
    last = torch.full(uniq.shape, -1, dtype=I64).scatter_reduce_(0, inv, pos, reduce="amax", include_self=True)
    table[uniq] = values[last]


# --------------------------------------------------------------------------------------
# map maintenance
# --------------------------------------------------------------------------------------
def map_update(s: MapState, points: torch.Tensor, sensor_position: torch.Tensor, cur_ts: int) -> None:
    """model/neural_points.py:205-270."""
    res = s.p.voxel_size_m
    keep = voxel_down_sample(points, res)
    sp = points[keep]
    h = signed_hash(cells_of(sp, res), s.p.buffer_size)
    old = s.table[h]
    if s.points.shape[0] > 0:
        d2 = ((s.points[old] - sp) ** 2).sum(-1)                       # old == -1 reads the last point (H8)
        dtravel = s.travel_dist[cur_ts] - s.travel_dist[s.ts_update[old]]
        upd = (old == -1) | (d2 > 3 * res ** 2) | (dtravel > s.diff_travel_dist_local)
    else:
        upd = torch.ones(old.shape, dtype=torch.bool)
    added = sp[upd]
    n_new = added.shape[0]
    m0 = s.points.shape[0]
    val = s.table[h]
    val[upd] = torch.arange(n_new, dtype=I64) + m0
    put_last_wins(s.table, h, val)                                       # sequential: last writer wins (H4)
    s.points = torch.cat((s.points, added), 0)
    ident = torch.zeros((n_new, 4))
    ident[:, 0] = 1.0
    s.quats = torch.cat((s.quats, ident), 0)
    ts_new = torch.full((n_new,), int(cur_ts), dtype=I64)
    s.ts_create = torch.cat((s.ts_create, ts_new), 0)
    s.ts_update = torch.cat((s.ts_update, ts_new), 0)
    new_f = s.p.feature_std * torch.randn(n_new + 1, s.p.feature_dim)
    s.feats = torch.cat((s.feats[:-1], new_f), 0)
    s.certainty = torch.cat((s.certainty, torch.zeros(n_new)), 0)
    reset_local_map(s, sensor_position, cur_ts)


def reset_local_map(s: MapState, sensor_position: torch.Tensor, cur_ts: int) -> None:
    """model/neural_points.py:272-313."""
    s.cur_ts = cur_ts
    d2s = ((s.points - sensor_position) ** 2).sum(-1)
    ts_used = ((s.ts_create + s.ts_update) / 2).long() if s.p.use_mid_ts else s.ts_create
    dtravel = torch.abs(s.travel_dist[cur_ts] - s.travel_dist[ts_used])
    m = (d2s < s.p.local_map_radius ** 2) & (dtravel < s.diff_travel_dist_local)
    s.l_points = s.points[m]
    s.l_quats = s.quats[m]
    s.l_certainty = s.certainty[m]
    s.l_ts_update = s.ts_update[m]
    m = torch.cat((m, torch.tensor([True])))
    s.local_mask = m
    # QUIRK (hazard H13, found by the golden vectors): the reference writes torch.full_like(local_mask, -1).long()
    # with a BOOL local_mask, i.e. True -> 1: points outside the local map remap to local index 1, not -1
    # (model/neural_points.py:301); only the padding entry is -1 (:305).
    g2l = torch.full_like(m, -1).long()
    nz = torch.nonzero(m).flatten()
    g2l[nz] = torch.arange(nz.shape[0], dtype=I64)
    g2l[-1] = -1
    s.global2local = g2l
    s.l_feats = s.feats[m].clone()


def assign_local_to_global(s: MapState) -> None:
    """model/neural_points.py:315-324."""
    m = s.local_mask
    s.points[m[:-1]] = s.l_points
    s.quats[m[:-1]] = s.l_quats
    s.feats[m] = s.l_feats.detach()
    s.certainty[m[:-1]] = s.l_certainty
    s.ts_update[m[:-1]] = s.l_ts_update


def recreate_hash(s: MapState, sensor_position, kept_points: bool = True, with_ts: bool = True, cur_ts: int = 0) -> None:
    """model/neural_points.py:372-428."""
    res = s.p.voxel_size_m
    s.table = torch.full((int(s.p.buffer_size),), -1, dtype=I64)
    if with_ts:
        ts_used = ((s.ts_create + s.ts_update) / 2).long() if s.p.use_mid_ts else s.ts_create
        keep = voxel_down_sample_min_value(s.points, res, torch.abs(ts_used - cur_ts).float())
    else:
        keep = voxel_down_sample_min_value(s.points, res, -s.certainty)
    if kept_points:
        h = signed_hash(cells_of(s.points[keep], res), s.p.buffer_size)
        put_last_wins(s.table, h, keep)
    else:
        s.points = s.points[keep]
        s.quats = s.quats[keep]
        s.ts_create = s.ts_create[keep]
        s.ts_update = s.ts_update[keep]
        s.certainty = s.certainty[keep]
        s.feats = s.feats[torch.cat((keep, torch.tensor([-1])))]
        h = signed_hash(cells_of(s.points, res), s.p.buffer_size)
        put_last_wins(s.table, h, torch.arange(s.points.shape[0], dtype=I64))
    if sensor_position is not None:
        reset_local_map(s, sensor_position, cur_ts)


def prune_map(s: MapState, prune_certainty_thre: float) -> bool:
    """model/neural_points.py:328-353 -- drop inactive (travel distance since the last update beyond the window) and
    uncertain neural points when there are more than 100 of them; the caller re-creates hash and local map."""
    diff = torch.abs(s.travel_dist[s.cur_ts] - s.travel_dist[s.ts_update])
    prune = (diff > s.diff_travel_dist_local) & (s.certainty < prune_certainty_thre)
    if int(prune.sum()) <= 100:
        return False
    s.points, s.quats = s.points[~prune], s.quats[~prune]
    s.ts_create, s.ts_update, s.certainty = s.ts_create[~prune], s.ts_update[~prune], s.certainty[~prune]
    s.feats = s.feats[~torch.cat((prune, torch.tensor([False])))]
    return True


def rotmat_to_quat(R: torch.Tensor) -> torch.Tensor:
    """utils/tools.py:326-334 -- (w, x, y, z); valid for rotations with 1 + trace > 0."""
    qw = torch.sqrt(1.0 + R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]) / 2.0
    qx = (R[:, 2, 1] - R[:, 1, 2]) / (4.0 * qw)
    qy = (R[:, 0, 2] - R[:, 2, 0]) / (4.0 * qw)
    qz = (R[:, 1, 0] - R[:, 0, 1]) / (4.0 * qw)
    return torch.stack((qw, qx, qy, qz), dim=1)


def quat_multiply(q1: torch.Tensor, q2: torch.Tensor) -> torch.Tensor:
    """utils/tools.py:356-369 -- q1 (x) q2."""
    w1, x1, y1, z1 = torch.unbind(q1, dim=1)
    w2, x2, y2, z2 = torch.unbind(q2, dim=1)
    return torch.stack((w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2), dim=1)


def adjust_map(s: MapState, pose_diff: torch.Tensor) -> None:
    """model/neural_points.py:355-370 -- after pose-graph optimisation every neural point is moved by the pose correction
    of the frame that created it (utils/tools.py:401-407 transform_batch_torch) and its orientation is rotated; from
    then on queries rotate x - p_k into the point's frame (after_pgo)."""
    s.after_pgo = True
    used = ((s.ts_create + s.ts_update) / 2).long() if s.p.use_mid_ts else s.ts_create
    T = pose_diff[used]
    s.points = (torch.matmul(T[:, :3, :3].to(s.points), s.points.unsqueeze(-1)) + T[:, :3, 3:].to(s.points)).squeeze(-1)
    s.quats = quat_multiply(rotmat_to_quat(pose_diff[:, :3, :3])[used], s.quats)


# --------------------------------------------------------------------------------------
# neighbour search + feature query
# --------------------------------------------------------------------------------------
def radius_search(s: MapState, x: torch.Tensor, time_filtering: bool = False):
    """model/neural_points.py:459-509 -> (dist2 [N,Kn] f32, idx [N,Kn] int64 global, -1 invalid)."""
    cells = cells_of(x, s.p.voxel_size_m)[..., None, :] + s.neighbor_dx
    idx = s.table[signed_hash(cells, s.p.buffer_size)]
    if time_filtering:
        dtravel = torch.abs(s.travel_dist[s.cur_ts] - s.travel_dist[s.ts_create[idx]])
        idx[~(dtravel < s.diff_travel_dist_local)] = -1
    d2 = ((s.points[idx] - x.view(-1, 1, 3)) ** 2).sum(-1)              # (dx^2+dy^2)+dz^2, no FMA (H3)
    d2[idx == -1] = s.max_valid_dist2
    idx[d2 > s.max_valid_dist2] = -1
    return d2, idx


def query_certainty(s: MapState, x: torch.Tensor) -> torch.Tensor:
    """model/neural_points.py:511-525 (global certainties, max over the stencil)."""
    _, idx = radius_search(s, x)
    c = s.certainty[idx]
    c[idx < 0] = 0.0
    return c.max(dim=-1)[0]


def rotate_passive(quat: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """utils/tools.py:316-323."""
    w = quat[..., 0:1]
    u = -quat[..., 1:]
    t = 2 * torch.linalg.cross(u, v)
    return v + w * t + torch.linalg.cross(u, t)


def query_feature(s: MapState, x: torch.Tensor, query_ts: Optional[torch.Tensor] = None,
                  training_mode: bool = True, query_locally: bool = True):
    """model/neural_points.py:528-674 (geo feature only).
    -> (feat [N,K,F+3] or [N,F+3], weight [N,K,1], nn_count [N] i64, certainty [N], knn_idx [N,K] i64)
    ``knn_idx`` (local or global index, -1 invalid) is extra, for index-parity checks."""
    K, F = s.p.query_nn_k, s.p.feature_dim
    d2, idx = radius_search(s, x, time_filtering=s.temporal_local_map_on and query_locally)
    if query_locally:
        idx = s.global2local[idx]
    nn = (idx >= 0).sum(-1)
    d2[idx == -1] = 9e3
    sd2, order = torch.sort(d2, dim=1)
    sidx = idx.gather(1, order)
    d2 = sd2[:, :K]
    idx = sidx[:, :K]
    knn_idx = idx.clone()
    valid = idx >= 0
    feats = s.l_feats if query_locally else s.feats
    pts = s.l_points if query_locally else s.points
    cert = s.l_certainty if query_locally else s.certainty
    quat = s.l_quats if query_locally else s.quats
    f = torch.zeros(x.shape[0], K, F)
    f[valid] = feats[idx[valid]]
    c = cert[idx]
    v = x.view(-1, 1, 3) - pts[idx]
    if s.after_pgo:
        v = rotate_passive(quat[idx], v)
    v[~valid] = torch.zeros(1, 3)
    fv = torch.cat((f, v), dim=2)
    eps = 1e-15
    w = 1.0 / (d2 + eps)
    w[~valid] = 0.0
    w[nn == 0] = eps
    w = torch.div(w, w.sum(1, keepdim=True))
    w[~valid] = 0.0
    with torch.no_grad():
        if training_mode:
            idx[~valid] = 0
            if query_locally:
                s.l_certainty.scatter_add_(0, idx.flatten(), w.detach().flatten())
                if query_ts is not None:
                    t = query_ts.view(-1, 1).repeat(1, K)
                    t[~valid] = 0
                    s.l_ts_update.scatter_reduce_(0, idx.flatten(), t.flatten(), reduce="amax", include_self=True)
            else:
                s.certainty.scatter_add_(0, idx.flatten(), w.detach().flatten())
        c[~valid] = 0.0
        qc = (c * w).sum(1)
    w = w.unsqueeze(-1)
    if s.p.weighted_first:
        fv = (fv * w).sum(1)
    return fv, w, nn, qc, knn_idx


# --------------------------------------------------------------------------------------
# decoder
# --------------------------------------------------------------------------------------
@dataclass
class DecoderParams:
    """layers.0.weight [H,F+3], layers.0.bias [H], lout.weight [1,H], lout.bias [1]  (model/decoder.py:43-49)."""
    W1: torch.Tensor
    b1: torch.Tensor
    W2: torch.Tensor
    b2: torch.Tensor
    sdf_scale: float = 0.55 * 0.08

    def tensors(self):
        return [self.W1, self.b1, self.W2, self.b2]


def new_decoder(in_dim: int = 11, hidden: int = 64, seed: int = 42, sdf_scale: float = 0.55 * 0.08) -> DecoderParams:
    """nn.Linear default initialisation under a fixed seed (same draw order as model/decoder.py:43-49)."""
    g = torch.Generator().manual_seed(seed)

    def lin(o, i):
        bound = 1.0 / math.sqrt(i)
        W = (torch.rand(o, i, generator=g) * 2 - 1) * bound
        b = (torch.rand(o, generator=g) * 2 - 1) * bound
        return W, b

    W1, b1 = lin(hidden, in_dim)
    W2, b2 = lin(1, hidden)
    return DecoderParams(W1, b1, W2, b2, sdf_scale)


def decoder_sdf(d: DecoderParams, feat: torch.Tensor) -> torch.Tensor:
    """model/decoder.py:66-88 (one hidden level, ReLU)."""
    h = torch.relu(torch.nn.functional.linear(feat, d.W1, d.b1))
    out = torch.nn.functional.linear(h, d.W2, d.b2).squeeze(1)
    return out * d.sdf_scale


# --------------------------------------------------------------------------------------
# fused-query semantics (what the CUDA kernels are compared with)
# --------------------------------------------------------------------------------------
def query_sdf(s: MapState, d: DecoderParams, x: torch.Tensor, query_ts=None, training_mode=False,
              query_locally=True, want_grad=False, want_std=False):
    """query_feature -> Decoder.sdf -> sum_k w_k s_k, as composed in utils/tracker.py:233-253,
    utils/mapper.py:461-468, 671-681.  Returns dict(sdf, std, grad, nn, certainty, knn_idx, w)."""
    if want_grad:
        x = x.detach().clone().requires_grad_(True)
    fv, w, nn, qc, knn = query_feature(s, x, query_ts, training_mode, query_locally)
    pred = decoder_sdf(d, fv)
    out = {"nn": nn, "certainty": qc, "knn_idx": knn, "w": w.detach().squeeze(-1)}
    if s.p.weighted_first:
        sdf = pred.squeeze(-1) if pred.dim() > 1 else pred
        std = None
    else:
        mean = (pred * w).sum(1)
        if want_std:
            var = (w * (pred - mean.unsqueeze(-1)) ** 2).sum(1)
            std = torch.sqrt(var).squeeze(1).detach()
        else:
            std = None
        sdf = mean.squeeze(1)
    out["std"] = std
    if want_grad:
        out["grad"] = torch.autograd.grad(sdf, x, torch.ones_like(sdf), retain_graph=False)[0].detach()
    out["sdf"] = sdf if (sdf.requires_grad and not want_grad) else sdf.detach()
    return out


def mesher_query(s: MapState, d: DecoderParams, x: torch.Tensor, min_nn: int, query_locally: bool = False):
    """utils/mesher.py:85-134 -- forward-only; MLP evaluated only where nn_count >= 1."""
    with torch.no_grad():
        fv, w, nn, _, _ = query_feature(s, x, None, False, query_locally)
        pm = nn >= 1
        pred = torch.zeros(x.shape[0], fv.shape[1], 1)
        pred[pm] = decoder_sdf(d, fv[pm])
        sdf = (pred * w).sum(1).squeeze(1)
        return sdf, nn >= min_nn


def grid_coords(min_bound, max_bound, voxel_size: float, pad_voxel: int = 0, skip_top_voxel: int = 0):
    """utils/mesher.py:138-180 -- z-fastest meshgrid of query points."""
    import numpy as np
    min_bound = np.asarray(min_bound, dtype=np.float64)
    max_bound = np.asarray(max_bound, dtype=np.float64)
    num = (np.ceil((max_bound - min_bound) / voxel_size) + pad_voxel * 2).astype(np.int_)
    origin = min_bound - pad_voxel * voxel_size
    origin[2] -= voxel_size
    num[2] += 1
    num[2] -= skip_top_voxel
    ax = [torch.arange(int(n), dtype=torch.int16) for n in num]
    gx, gy, gz = torch.meshgrid(*ax, indexing="ij")
    coord = torch.stack((gx.flatten(), gy.flatten(), gz.flatten())).transpose(0, 1).float()
    coord *= voxel_size
    coord += torch.tensor(origin, dtype=torch.float32)
    return coord, num, origin


# --------------------------------------------------------------------------------------
# mapping loss (utils/mapper.py:443-573, utils/loss.py:40-47)
# --------------------------------------------------------------------------------------
def bce_loss(pred, label, sigma, weight):
    """utils/loss.py:40-47 with weighted=True, mean reduction."""
    return torch.nn.functional.binary_cross_entropy_with_logits(pred / sigma, torch.sigmoid(label / sigma), weight=weight)


def numerical_gradient(s: MapState, d: DecoderParams, x: torch.Tensor, eps: float) -> torch.Tensor:
    """utils/mapper.py:683-711 (two-sided; the six shifted batches run in training mode, hazard H9)."""
    n = x.shape[0]
    e = torch.eye(3) * eps
    xs = torch.cat((x + e[0], x - e[0], x + e[1], x - e[1], x + e[2], x - e[2]), 0)
    fv, w, _, _, _ = query_feature(s, xs)
    v = (decoder_sdf(d, fv) * w).sum(1).squeeze(1).unsqueeze(-1)
    gx = (v[:n] - v[n:2 * n]) / (2 * eps)
    gy = (v[2 * n:3 * n] - v[3 * n:4 * n]) / (2 * eps)
    gz = (v[4 * n:5 * n] - v[5 * n:]) / (2 * eps)
    return torch.cat([gx, gy, gz], dim=1)


def mapping_loss(s: MapState, d: DecoderParams, coord, sdf_label, ts, weight, *, sigma=None, weight_e=0.5,
                 decimation=10, eps=0.4 * 0.2):
    """One iteration's loss of utils/mapper.py:443-547 for the run_kitti settings
    (bce + eikonal on numerical gradients, 'all').  ``s.l_feats`` / decoder tensors may require grad."""
    sigma = d.sdf_scale if sigma is None else sigma
    fv, w, _, _, _ = query_feature(s, coord, ts)
    pred = (decoder_sdf(d, fv) * w).sum(1).squeeze(1)
    g = numerical_gradient(s, d, coord[::decimation], eps)
    loss = bce_loss(pred, sdf_label, sigma, torch.abs(weight).detach())
    eik = ((g.norm(2, dim=-1) - 1.0) ** 2).mean()
    return loss + weight_e * eik, pred, g


# --------------------------------------------------------------------------------------
# tracker (utils/tracker.py)
# --------------------------------------------------------------------------------------
def skew(v):
    """utils/tracker.py:573-578."""
    S = torch.zeros(3, 3, dtype=v.dtype)
    S[0, 1], S[0, 2], S[1, 2] = -v[2], v[1], -v[0]
    return S - S.T


def expmap(a):
    """utils/tracker.py:580-589 (Rodrigues)."""
    ang = a.norm()
    S = skew(a / ang)
    return torch.eye(3, dtype=a.dtype) + S * torch.sin(ang) + (S @ S) * (1.0 - torch.cos(ang))


def gauss_newton_system(points, grad, residual, weight, lm_lambda):
    """utils/tracker.py:468-491 -> (H [6,6] f32 after LM damping, b [6] f32, H before damping)."""
    J = torch.cat([torch.cross(points, grad, dim=-1), grad], -1)
    H = J.T @ (weight * J)
    H0 = H.clone()
    H = H + lm_lambda * torch.diag(torch.diag(H))
    b = -(J * weight).T @ residual
    return H, b, H0


def registration_step(s: MapState, d: DecoderParams, points, *, min_grad_norm=0.5, max_grad_norm=2.0,
                      GM_dist=0.2, GM_grad=0.1, lm_lambda=1e-4, max_sdf_std=0.25, min_nn=6):
    """utils/tracker.py:277-452 (LiDAR path: no colours/normals, sdf labels = 0).
    -> (delta_T [4,4] f64, valid_count, residual_cm, aux dict)"""
    q = query_sdf(s, d, points, None, False, True, want_grad=True, want_std=True)
    sdf, g, std, nn = q["sdf"], q["grad"], q["std"], q["nn"]
    gn = g.norm(dim=-1)
    valid = (nn >= min_nn) & (gn < max_grad_norm) & (gn > min_grad_norm) & (std < max_sdf_std)
    vp = points[valid]
    n_valid = vp.shape[0]
    aux = {"valid": valid, "sdf": sdf, "grad": g, "std": std, "nn": nn}
    if n_valid < 10:
        return torch.eye(4, dtype=torch.float64), n_valid, 0.0, aux
    gn_v, sdf_v, g_v = gn[valid], sdf[valid], g[valid]
    res = sdf_v - 0.0
    res_cm = torch.mean(torch.abs(res)).item() * 100.0
    w_grad = ((GM_grad / (GM_grad ** 2 + (gn_v - 1.0) ** 2)) ** 2).unsqueeze(1)
    w_res = ((GM_dist / (GM_dist ** 2 + res ** 2)) ** 2).unsqueeze(1)
    w = w_res * w_grad
    w = w / (2.0 * torch.mean(w))
    H, b, H0 = gauss_newton_system(vp, g_v, res, w, lm_lambda)
    t = torch.linalg.inv(H.to(torch.float64)) @ b.to(torch.float64)
    T = torch.eye(4, dtype=torch.float64)
    T[:3, :3] = expmap(t[:3])
    T[:3, 3] = t[3:]
    aux.update({"H": H, "b": b, "H0": H0, "w": w})
    return T, n_valid, res_cm, aux


def transform_points(points, T):
    """utils/tools.py:386-399."""
    ph = torch.cat([points, torch.ones(points.shape[0], 1).to(points)], dim=1)
    return torch.matmul(ph, T.to(points).T)[:, :3]


def tracking(s: MapState, d: DecoderParams, source_points, init_pose, *, iter_n=100, term_deg=0.01, term_m=0.0005,
             surface_sample_range_m=0.25, **kw):
    """utils/tracker.py:39-174 (control flow incl. validity checks; no visualisation outputs).
    -> (T [4,4] f64, valid_flag, iterations, last residual cm)"""
    T = init_pose.clone()
    max_final_res_cm = surface_sample_range_m * 0.5 * 100.0
    last_res, valid_flag, converged = 1e5, True, False
    n_src = source_points.shape[0]
    it = 0
    res_cm = 0.0
    for i in range(iter_n):
        it = i + 1
        cur = transform_points(source_points, T)
        dT, n_valid, res_cm, _ = registration_step(s, d, cur, **kw)
        T = dT @ T
        if (res_cm - last_res) / last_res > 1.1:
            valid_flag = False
        else:
            last_res = res_cm
        if n_valid < 30 or 1.0 * n_valid / n_src < 0.2:
            valid_flag = False
        if not valid_flag or converged:
            break
        rot_deg = torch.acos((torch.trace(dT[:3, :3]) - 1) / 2) * 180.0 / math.pi
        tran = dT[:3, 3].norm()
        if abs(rot_deg) < term_deg and tran < term_m or i == iter_n - 2:
            converged = True
    if res_cm > max_final_res_cm:
        valid_flag = False
    if not valid_flag:
        T = init_pose
    return T, valid_flag, it, res_cm


# --------------------------------------------------------------------------------------
# frame data path in front of the hot path (dataset/slam_dataset.py, utils/tools.py:540-567)
# --------------------------------------------------------------------------------------
def velodyne_point_ts(points: torch.Tensor) -> torch.Tensor:
    """dataset/slam_dataset.py:247-249: per-point time of a sweep without time stamps (clockwise from the -x axis)."""
    yaw = -torch.atan2(points[:, 1], points[:, 0])
    return 0.5 * (yaw / math.pi + 1.0)


def rotmat_slerp_from_identity(R: torch.Tensor, steps: torch.Tensor) -> torch.Tensor:
    """roma.rotmat_slerp(I, R, steps) -- roma is an un-vendored dependency of the reference (utils/tools.py:24, 560;
    roma >= 1.4: rotation matrix -> unit quaternion -> rotation vector, scaled by every step, back to a matrix; shortest
    arc; steps outside [0, 1] extrapolate).  Restated as exp(step * log R) with Rodrigues' formula -> [N,3,3]."""
    cos = torch.clamp(((R[0, 0] + R[1, 1] + R[2, 2]) - 1.0) * 0.5, -1.0, 1.0)
    w = torch.stack([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    n = torch.linalg.norm(w)
    if float(n) <= 1e-12:
        return torch.eye(3, dtype=R.dtype).expand(steps.shape[0], 3, 3).clone()
    axis = w / n
    ang = steps * torch.atan2(0.5 * n, cos)
    Kx = torch.zeros(3, 3, dtype=R.dtype)
    Kx[0, 1], Kx[0, 2], Kx[1, 0], Kx[1, 2], Kx[2, 0], Kx[2, 1] = -axis[2], axis[1], axis[2], -axis[0], -axis[1], axis[0]
    eye = torch.eye(3, dtype=R.dtype)
    return eye[None] + torch.sin(ang)[:, None, None] * Kx[None] + (1.0 - torch.cos(ang))[:, None, None] * (Kx @ Kx)[None]


def deskew(points: torch.Tensor, ts: torch.Tensor, pose: torch.Tensor, ts_mid_pose: float = 0.5) -> torch.Tensor:
    """utils/tools.py:540-567 (returns a new tensor; the reference works in place): pose = T_last<-cur."""
    if ts is None:
        return points
    t = ts.reshape(-1).to(points.dtype)
    t = (t - torch.min(t)) / (torch.max(t) - torch.min(t)) - ts_mid_pose
    R = rotmat_slerp_from_identity(pose[:3, :3].to(points.dtype), t)
    return (R @ points[:, :3].unsqueeze(-1)).squeeze(-1) + t[:, None] * pose[:3, 3].to(points.dtype)


def crop_frame(points, ts, min_z, max_z, min_range, max_range):
    """dataset/slam_dataset.py:827-834."""
    dist = torch.norm(points[:, :3], dim=1)
    keep = (dist > min_range) & (dist < max_range) & (points[:, 2] > min_z) & (points[:, 2] < max_z)
    return points[keep], (ts[keep] if ts is not None else None)


def intrinsic_correct(points, correct_deg):
    """dataset/slam_dataset.py:837-854 (KITTI vertical-angle calibration); returns a new tensor."""
    if correct_deg == 0.0:
        return points
    dist = torch.norm(points[:, :3], dim=1)
    v = torch.asin(points[:, 2] / dist)
    vc = v + correct_deg / 180.0 * math.pi
    hs = torch.cos(vc) / torch.cos(v)
    return torch.stack([points[:, 0] * hs, points[:, 1] * hs, dist * torch.sin(vc)], 1)
