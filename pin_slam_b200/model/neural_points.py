"""NeuralPoints -- the neural-point map and its voxel-hash index, host side of libpinq.

Drop-in mirror of the reference class (model/neural_points.py:18-703): same constructor, method
names, argument meaning, return shapes/dtypes and externally touched attributes.  All arithmetic on
the path runs in the sm_100a kernels behind include/pinq.h; this file only owns the tensors and
sequences the launches.  There is no CPU path: a non-CUDA config.device raises.

Differences in representation (results identical):
  * the slot table is dense int32 (`_table`); `buffer_pt_index` is a lazily materialised int64 view;
  * the query kernels read an L2-resident view of that table (`_occ` occupancy words + rank-ordered
    16-byte candidate records `_cand_*` = xyz + global / filtered local index, csrc/index.cu), rebuilt
    whenever the table, the local map or `travel_dist` change;
  * `query_sdf()` is the fused entry used by Tracker / Mapper / Mesher (one autograd.Function).
"""
from __future__ import annotations

import os
import sys

import torch
import torch.nn as nn

from .. import _pinq as P
from ..utils.tools import voxel_down_sample_torch, voxel_down_sample_min_value_torch, transform_batch_torch


def _rotmat_to_quat(R):
    """reference: utils/tools.py:326-334."""
    qw = torch.sqrt(1.0 + R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]) / 2.0
    qx = (R[:, 2, 1] - R[:, 1, 2]) / (4.0 * qw)
    qy = (R[:, 0, 2] - R[:, 2, 0]) / (4.0 * qw)
    qz = (R[:, 1, 0] - R[:, 0, 1]) / (4.0 * qw)
    return torch.stack((qw, qx, qy, qz), dim=1)


def _quat_multiply(q1, q2):
    """reference: utils/tools.py:356-370."""
    w1, x1, y1, z1 = torch.unbind(q1, dim=1)
    w2, x2, y2, z2 = torch.unbind(q2, dim=1)
    return torch.stack((w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2), dim=1)


class NeuralPoints(nn.Module):
    def __init__(self, config) -> None:
        super().__init__()
        self.config = config
        self.silence = config.silence
        self.geo_feature_dim = config.feature_dim
        self.geo_feature_std = config.feature_std
        self.color_feature_dim = config.feature_dim
        self.color_feature_std = config.feature_std
        if config.pos_encoding_band > 0 or config.layer_norm_on or config.color_on:
            raise NotImplementedError("positional encoding / layer norm / colour features are out of scope")
        self.device = config.device
        if not str(self.device).startswith("cuda"):
            raise P.PinqError("pin_slam_b200.NeuralPoints runs on CUDA only (config.device=%r); no CPU fallback" % (self.device,))
        P.load()
        self.dtype = config.dtype
        self.idx_dtype = torch.int64
        self.resolution = config.voxel_size_m
        self.buffer_size = int(config.buffer_size)
        self.temporal_local_map_on = True
        self.local_map_radius = config.local_map_radius
        self.diff_travel_dist_local = config.local_map_radius * config.local_map_travel_dist_ratio
        self.diff_ts_local = config.diff_ts_local
        self.local_orientation = torch.eye(3, device=self.device)
        self.cur_ts = 0
        self.max_ts = 0
        self._travel_dist = None
        self.est_poses = None
        self.after_pgo = False
        self.primes = torch.tensor([73856093, 19349669, 83492791], dtype=self.idx_dtype, device=self.device)
        # global map
        self._table = torch.empty(self.buffer_size, dtype=torch.int32, device=self.device)
        P.check(P.load().pinq_table_clear(P.ptr(self._table), self.buffer_size, P.stream()))
        self.neural_points = torch.empty((0, 3), dtype=self.dtype, device=self.device)
        self.point_orientations = torch.empty((0, 4), dtype=self.dtype, device=self.device)
        self.geo_features = torch.zeros((1, self.geo_feature_dim), dtype=self.dtype, device=self.device)
        self.color_features = None
        self.point_ts_create = torch.empty((0), device=self.device, dtype=torch.long)
        self.point_ts_update = torch.empty((0), device=self.device, dtype=torch.long)
        self.point_certainties = torch.empty((0), dtype=self.dtype, device=self.device)
        # local map
        self.local_neural_points = torch.empty((0, 3), dtype=self.dtype, device=self.device)
        self.local_point_orientations = torch.empty((0, 4), dtype=self.dtype, device=self.device)
        self.local_geo_features = nn.Parameter()
        self.local_color_features = nn.Parameter()
        self.local_point_certainties = torch.empty((0), dtype=self.dtype, device=self.device)
        self.local_point_ts_update = torch.empty((0), device=self.device, dtype=torch.long)
        self.local_mask = None
        self.global2local = None
        # packed candidate records for the fused kernels (rebuilt lazily)
        self._cand_local = None
        self._cand_global = None
        self._cand_dirty = True
        self._index_dirty = True
        self._occ = None
        self._ws = None
        self.bin_queries_min = 32768  # fused queries of at least this many points are binned by brick first
        # fused search through the brick index when it is provably equivalent (PINQ_BRICK=0: slot probing only)
        self.use_brick_index = os.environ.get("PINQ_BRICK", "1") != "0"
        # the brick search packs a record rank into 24 key bits (csrc/query.cu kRankBits): maps of 2^24 = 16.7 M points or more
        # fall back to the slot-probing search (same results, tested by forcing the limit down)
        self.brick_max_points = 1 << 24
        self.exact_search_keys = False  # A/B tests: brick search with 64-bit keys for every query
        self.skip_in_range_test = True  # A/B tests: False keeps the per-candidate `d2 > max_valid_dist2` comparison
        self.zero_grad_in_search = True  # the fused forward clears the feature-gradient block of its backward (zero_fill)
        self.decode_wait_event = None  # torch.cuda.Event the fused query's decode waits for (see pinq_query_args)
        self.use_mma_sync = False  # A/B tests: the mma.sync decode / backward kernels instead of the tcgen05 ones
        self._brick_tab = self._brick_global = self._brick_local = self._brick_buf = self._brick_counters = None
        self._brick_host = self._brick_event = None
        self._brick_cap = 0
        self._brick_pending = False
        self._brick_safe_cache, self._stencil_radius = {}, 0
        self.set_search_neighborhood(num_nei_cells=config.num_nei_cells, search_alpha=config.search_alpha)
        self.memory_footprint = []
        self.to(self.device)

    # ---- attributes with side effects ------------------------------------------------------
    @property
    def travel_dist(self):
        return self._travel_dist

    @travel_dist.setter
    def travel_dist(self, value):
        # set by the driver every frame (reference: pin_slam.py:134); the travel-distance filter is
        # baked into the packed candidates, so they must be rebuilt.
        if value is not None:
            value = value.to(device=self.device, dtype=torch.float32).contiguous()
        self._travel_dist = value
        self._cand_dirty = True

    @property
    def buffer_pt_index(self):
        """Dense int64 view of the slot table (reference: model/neural_points.py:73); materialised on demand."""
        if self._table is None:
            return None
        return self._table.to(torch.int64)

    @buffer_pt_index.setter
    def buffer_pt_index(self, value):
        if value is None:
            self._table = None
        else:
            self._table = value.to(device=self.device, dtype=torch.int32).contiguous()
        self._cand_dirty = True
        self._index_dirty = True

    def is_empty(self):
        return self.neural_points.shape[0] == 0

    def count(self):
        return self.neural_points.shape[0]

    def local_count(self):
        return self.local_neural_points.shape[0]

    def print_memory(self):
        if not self.silence:
            print("# Global neural point: %d" % (self.count()))
            print("# Local  neural point: %d" % (self.local_count()))
        point_dim = self.config.feature_dim + 3 + 4
        cur_memory = self.count() * point_dim * 4 / 1024 / 1024
        if not self.silence:
            print("Memory consumption: %f (MB)" % cur_memory)
        self.memory_footprint.append(cur_memory)

    # ---- index helpers -----------------------------------------------------------------------
    def _slots(self, points):
        slots = torch.empty(points.shape[0], dtype=torch.int32, device=self.device)
        P.check(P.load().pinq_hash_points(P.ptr(points), points.shape[0], float(self.resolution), self.buffer_size, 0,
                                          P.ptr(slots), P.stream()))
        return slots

    def _compact(self, flags_u8):
        """indices (int32, ascending) of the set flags; one host sync for the count."""
        lib = P.load()
        n = flags_u8.shape[0]
        out = torch.empty(n, dtype=torch.int32, device=self.device)
        cnt = torch.empty(1, dtype=torch.int32, device=self.device)
        scratch = torch.empty(int(lib.pinq_compact_scratch_bytes(n)), dtype=torch.uint8, device=self.device)
        P.check(lib.pinq_compact_flags(P.ptr(flags_u8), n, P.ptr(out), P.ptr(cnt), P.ptr(scratch), P.stream()))
        return out[: int(cnt.item())]

    def _ensure_table(self):
        if self._table is None:
            self._table = torch.empty(self.buffer_size, dtype=torch.int32, device=self.device)
            P.check(P.load().pinq_table_clear(P.ptr(self._table), self.buffer_size, P.stream()))
            self._index_dirty = True
            self._cand_dirty = True

    # ---- map maintenance -----------------------------------------------------------------------
    def update(self, points: torch.Tensor, sensor_position: torch.Tensor, sensor_orientation: torch.Tensor, cur_ts):
        """reference: model/neural_points.py:205-270."""
        lib = P.load()
        st = P.stream()
        self._ensure_table()
        points = points.detach().to(self.dtype).contiguous()
        sample_idx = voxel_down_sample_torch(points, self.resolution)
        sample_points = points[sample_idx].contiguous()
        n = sample_points.shape[0]
        slots = self._slots(sample_points)
        hash_idx = torch.empty(n, dtype=torch.int32, device=self.device)
        P.check(lib.pinq_table_gather(P.ptr(self._table), P.ptr(slots), n, P.ptr(hash_idx), st))
        mask = torch.empty(n, dtype=torch.uint8, device=self.device)
        m0 = self.neural_points.shape[0]
        P.check(lib.pinq_update_mask(P.ptr(sample_points), P.ptr(hash_idx), n, P.ptr(self.neural_points),
                                     P.ptr(self.point_ts_update), m0, P.ptr(self._travel_dist), int(cur_ts),
                                     float(self.resolution), float(self.diff_travel_dist_local), P.ptr(mask), st))
        sel = self._compact(mask)
        new_point_count = sel.shape[0]
        sel64 = sel.to(torch.int64)
        added_pt = sample_points[sel64]
        values = hash_idx  # cur_pt_idx (neural_points.py:242-245)
        values[sel64] = torch.arange(new_point_count, dtype=torch.int32, device=self.device) + m0
        P.check(lib.pinq_table_scatter_last_wins(P.ptr(self._table), P.ptr(slots), P.ptr(values), n, st))
        self._index_dirty = True

        self.neural_points = torch.cat((self.neural_points, added_pt), 0)
        ident = torch.zeros((new_point_count, 4), dtype=self.dtype, device=self.device)
        ident[:, 0] = 1.0
        self.point_orientations = torch.cat((self.point_orientations, ident), 0)
        new_ts = torch.full((new_point_count,), int(cur_ts), device=self.device, dtype=torch.long)
        self.point_ts_create = torch.cat((self.point_ts_create, new_ts), 0)
        self.point_ts_update = torch.cat((self.point_ts_update, new_ts), 0)
        new_fts = self.geo_feature_std * torch.randn(new_point_count + 1, self.geo_feature_dim, device=self.device,
                                                     dtype=self.dtype)
        self.geo_features = torch.cat((self.geo_features[:-1], new_fts), 0)
        self.point_certainties = torch.cat((self.point_certainties,
                                            torch.zeros(new_point_count, device=self.device, dtype=self.dtype)), 0)
        self.reset_local_map(sensor_position, sensor_orientation, cur_ts)

    def reset_local_map(self, sensor_position, sensor_orientation, cur_ts: int, use_travel_dist: bool = True,
                        diff_ts_local: int = 50):
        """reference: model/neural_points.py:272-313."""
        if not use_travel_dist:
            raise NotImplementedError("time-difference local map (use_travel_dist=False) is not used by the reference")
        lib = P.load()
        self.cur_ts = cur_ts
        self.max_ts = max(self.max_ts, cur_ts)
        m = self.neural_points.shape[0]
        flags = torch.empty(m + 1, dtype=torch.uint8, device=self.device)
        flags[m] = 1  # padding element (neural_points.py:298)
        sp = sensor_position.detach().to(device=self.device, dtype=torch.float32).contiguous()
        P.check(lib.pinq_local_mask(P.ptr(self.neural_points), P.ptr(self.point_ts_create), P.ptr(self.point_ts_update),
                                    m, P.ptr(sp), P.ptr(self._travel_dist), int(cur_ts),
                                    float(self.local_map_radius ** 2), float(self.diff_travel_dist_local),
                                    int(bool(self.config.use_mid_ts)), P.ptr(flags), P.stream()))
        idx_pad = self._compact(flags).to(torch.int64)  # local point indices + the padding row
        idx = idx_pad[:-1]
        self.local_neural_points = self.neural_points[idx]
        self.local_point_orientations = self.point_orientations[idx]
        self.local_point_certainties = self.point_certainties[idx]
        self.local_point_ts_update = self.point_ts_update[idx]
        self.local_mask = flags.to(torch.bool)
        self._local_flags = flags
        # QUIRK kept for parity (hazard H13): the reference builds this with full_like(<bool mask>, -1).long(),
        # so points outside the local map remap to local index 1 (not -1); only the padding entry is -1
        # (reference: model/neural_points.py:301-305).
        g2l = torch.full((m + 1,), 1, dtype=torch.int64, device=self.device)
        g2l[idx] = torch.arange(idx.shape[0], dtype=torch.int64, device=self.device)
        g2l[m] = -1
        self.global2local = g2l
        self._local_idx = idx_pad
        self.local_geo_features = nn.Parameter(self.geo_features[idx_pad])
        self.local_orientation = sensor_orientation
        self._cand_dirty = True

    def assign_local_to_global(self):
        """reference: model/neural_points.py:315-324."""
        idx_pad = self._local_idx
        idx = idx_pad[:-1]
        self.neural_points[idx] = self.local_neural_points
        self.point_orientations[idx] = self.local_point_orientations
        self.geo_features[idx_pad] = self.local_geo_features.data
        self.point_certainties[idx] = self.local_point_certainties
        self.point_ts_update[idx] = self.local_point_ts_update

    def prune_map(self, prune_certainty_thre):
        """reference: model/neural_points.py:329-353."""
        diff = torch.abs(self._travel_dist[self.cur_ts] - self._travel_dist[self.point_ts_update])
        prune_mask = (diff > self.diff_travel_dist_local) & (self.point_certainties < prune_certainty_thre)
        prune_count = torch.sum(prune_mask).item()
        if prune_count > 100:
            keep = ~prune_mask
            self.neural_points = self.neural_points[keep]
            self.point_orientations = self.point_orientations[keep]
            self.point_ts_create = self.point_ts_create[keep]
            self.point_ts_update = self.point_ts_update[keep]
            self.point_certainties = self.point_certainties[keep]
            keep = torch.cat((keep, torch.tensor([True], device=self.device)), dim=0)
            self.geo_features = self.geo_features[keep]
            self._cand_dirty = True
            return True
        return False

    def adjust_map(self, pose_diff_torch):
        """reference: model/neural_points.py:355-370 (per-point SE(3) warp after pose-graph optimisation)."""
        self.after_pgo = True
        if self.config.use_mid_ts:
            used_ts = ((self.point_ts_create + self.point_ts_update) / 2).long()
        else:
            used_ts = self.point_ts_create
        self.neural_points = transform_batch_torch(self.neural_points, pose_diff_torch[used_ts]).contiguous()
        dq = _rotmat_to_quat(pose_diff_torch[:, :3, :3])
        self.point_orientations = _quat_multiply(dq[used_ts], self.point_orientations).to(self.dtype).contiguous()
        self._cand_dirty = True

    def recreate_hash(self, sensor_position, sensor_orientation, kept_points: bool = True, with_ts: bool = True,
                      cur_ts=0):
        """reference: model/neural_points.py:372-428."""
        lib = P.load()
        st = P.stream()
        self._table = None
        self._ensure_table()
        if with_ts:
            if self.config.use_mid_ts:
                ts_used = ((self.point_ts_create + self.point_ts_update) / 2).long()
            else:
                ts_used = self.point_ts_create
            ts_diff = torch.abs(ts_used - cur_ts).float()
            sample_idx = voxel_down_sample_min_value_torch(self.neural_points, self.resolution, ts_diff)
        else:
            sample_idx = voxel_down_sample_min_value_torch(self.neural_points, self.resolution, -self.point_certainties)
        if kept_points:
            sample_points = self.neural_points[sample_idx].contiguous()
            slots = self._slots(sample_points)
            values = sample_idx.to(torch.int32)
        else:
            self.neural_points = self.neural_points[sample_idx].contiguous()
            self.point_orientations = self.point_orientations[sample_idx].contiguous()
            self.point_ts_create = self.point_ts_create[sample_idx]
            self.point_ts_update = self.point_ts_update[sample_idx]
            self.point_certainties = self.point_certainties[sample_idx]
            pad = torch.cat((sample_idx, torch.tensor([-1], device=self.device, dtype=sample_idx.dtype)))
            self.geo_features = self.geo_features[pad]
            slots = self._slots(self.neural_points)
            values = torch.arange(self.neural_points.shape[0], dtype=torch.int32, device=self.device)
        P.check(lib.pinq_table_scatter_last_wins(P.ptr(self._table), P.ptr(slots), P.ptr(values), slots.shape[0], st))
        self._cand_dirty = True
        self._index_dirty = True
        if sensor_position is not None:
            self.reset_local_map(sensor_position, sensor_orientation, cur_ts)
        if not kept_points:
            self.print_memory()

    def set_search_neighborhood(self, num_nei_cells: int = 1, search_alpha: float = 1.0):
        """reference: model/neural_points.py:430-457."""
        dx = torch.arange(-num_nei_cells, num_nei_cells + 1, device=self.primes.device, dtype=self.primes.dtype)
        coords = torch.meshgrid(dx, dx, dx, indexing="ij")
        dx = torch.stack(coords, dim=-1).reshape(-1, 3)
        dx2 = torch.sum(dx ** 2, dim=-1)
        self.neighbor_dx = dx[dx2 < (num_nei_cells + search_alpha) ** 2]
        self.neighbor_K = self.neighbor_dx.shape[0]
        self._stencil_radius = int(num_nei_cells)
        self.max_valid_dist2 = 3 * ((num_nei_cells + 1) * self.resolution) ** 2
        self._stencil_i32 = self.neighbor_dx.to(torch.int32).contiguous()
        self._stencil_key = (int(num_nei_cells), float(search_alpha))
        # A query and a neural point of stencil cell dx are at most (|dx_i| + 1) cells apart per axis.  If even that reach stays
        # below max_valid_dist2 (run_kitti.yaml: 17 x 0.16 = 2.72 m^2 < 4.32 m^2), the reference's `dist2 > max_valid_dist2`
        # rejection (:499) can never fire for a candidate found through the brick index and the kernel skips the comparison
        # (PINQ_F_STENCIL_IN_RANGE).  0.1 % margin for the fp32 rounding of the cell assignment and of d^2.
        reach2 = float(((self.neighbor_dx.abs() + 1) ** 2).sum(-1).max()) * float(self.resolution) ** 2
        self._stencil_in_range = reach2 * 1.001 < float(self.max_valid_dist2)

    # ---- packed views for the kernels ---------------------------------------------------------
    def _refresh_candidates(self):
        """(Re)build the L2-resident read view of the slot table (csrc/index.cu): occupancy words +
        rank-ordered 16-byte candidate records (xyz + global index | local index after the
        travel-distance filter of model/neural_points.py:480-488 and the global2local remap :555)."""
        if not self._cand_dirty:
            return
        lib = P.load()
        st = P.stream()
        self._ensure_table()
        m = self.neural_points.shape[0]
        self.neural_points = self.neural_points.contiguous()
        if self._index_dirty or self._occ is None:
            nwords = int(lib.pinq_index_words(self.buffer_size))
            if self._occ is None:
                self._occ = torch.empty((nwords, 2), dtype=torch.int32, device=self.device)
                self._occ_scratch = torch.empty(int(lib.pinq_index_scratch_bytes(self.buffer_size)), dtype=torch.uint8,
                                                device=self.device)
                self._occ_count = torch.empty(1, dtype=torch.int32, device=self.device)
            P.check(lib.pinq_index_build(P.ptr(self._table), self.buffer_size, P.ptr(self._occ), P.ptr(self._occ_count),
                                         P.ptr(self._occ_scratch), st))
            self._index_dirty = False
        # occupied slots <= number of points ever inserted, so m rows always suffice (no host sync)
        self._cand_global = torch.empty((max(m, 1), 4), dtype=torch.float32, device=self.device)
        have_local = (self.global2local is not None and self.global2local.shape[0] == m + 1
                      and self._travel_dist is not None)
        self._cand_local = torch.empty((max(m, 1), 4), dtype=torch.float32, device=self.device) if have_local else None
        if m > 0:
            P.check(lib.pinq_index_pack(P.ptr(self._table), self.buffer_size, P.ptr(self._occ), P.ptr(self.neural_points),
                                        P.ptr(self.point_ts_create), P.ptr(self.global2local) if have_local else None,
                                        P.ptr(self._local_flags) if have_local else None,
                                        P.ptr(self._travel_dist) if have_local else None, int(self.cur_ts),
                                        float(self.diff_travel_dist_local), P.ptr(self._cand_global),
                                        P.ptr(self._cand_local) if have_local else None, st))
        self._build_bricks(m, have_local, st)
        self._cand_dirty = False

    # ---- brick index (csrc/brick.cuh) --------------------------------------------------------------
    def _brick_safe(self) -> bool:
        """The brick view equals the slot view iff every far-collision alias is rejected by the distance test,
        i.e. no two voxels closer than radius + ceil(sqrt(max_valid_dist2)/res) + 1 share a slot.  A property
        of (primes, buffer_size) only; checked once per (buffer_size, stencil) on the host, no device sync."""
        radius = int(self._stencil_radius)
        key = (int(self.buffer_size), radius, int(self.neighbor_K), float(self.max_valid_dist2), float(self.resolution))
        ok = self._brick_safe_cache.get(key)
        if ok is None:
            ok = radius <= 2 and self.neighbor_K <= 125
            if ok:
                import math
                rc = radius + int(math.ceil(math.sqrt(float(self.max_valid_dist2)) / float(self.resolution))) + 1
                d = torch.arange(-rc, rc + 1, dtype=torch.int64)
                off = torch.stack(torch.meshgrid(d, d, d, indexing="ij"), dim=-1).reshape(-1, 3)
                primes = torch.tensor([73856093, 19349669, 83492791], dtype=torch.int64)
                h = torch.remainder((off * primes).sum(-1), int(self.buffer_size))
                ok = int((h == 0).sum()) == 1  # only the zero offset maps to slot difference 0
            self._brick_safe_cache[key] = bool(ok)
        return bool(ok)

    def _build_bricks(self, m: int, have_local: bool, st):
        self._brick_tab = self._brick_global = self._brick_local = None
        if not self.use_brick_index or m == 0 or m >= self.brick_max_points or not self._brick_safe():
            return
        lib = P.load()
        self.check_index_invariants(block=False)
        cap = int(lib.pinq_brick_capacity(m))
        if self._brick_buf is None or self._brick_buf.numel() != cap * 32:
            self._brick_buf = torch.empty(cap * 32, dtype=torch.uint8, device=self.device)
            self._brick_counters = torch.zeros(2, dtype=torch.int32, device=self.device)
        self._brick_global = torch.empty((m, 4), dtype=torch.float32, device=self.device)
        self._brick_local = torch.empty((m, 4), dtype=torch.float32, device=self.device) if have_local else None
        P.check(lib.pinq_brick_build(P.ptr(self._cand_global), P.ptr(self._cand_local) if have_local else None,
                                     P.ptr(self._occ_count), m, float(self.resolution), 0, P.ptr(self._brick_buf), cap,
                                     P.ptr(self._brick_counters), P.ptr(self._brick_global),
                                     P.ptr(self._brick_local) if have_local else None, st))
        self._brick_tab, self._brick_cap = self._brick_buf, cap
        self._brick_pending = True

    def check_index_invariants(self, block: bool = True):
        """Raises if the last brick build saw two findable points in one voxel (the slot table and the point
        positions disagree, e.g. points were moved without recreate_hash).  Non-blocking callers check the
        counters of the PREVIOUS build, so the hot path never waits for the device."""
        if not getattr(self, "_brick_pending", False) or self._brick_counters is None:
            return
        if not block:
            if self._brick_host is None:
                self._brick_host = torch.zeros(2, dtype=torch.int32).pin_memory()
                self._brick_event = torch.cuda.Event()
            elif self._brick_event.query() and int(self._brick_host[1]) != 0:
                raise P.PinqError("brick index: %d voxels hold two findable points" % int(self._brick_host[1]))
            self._brick_host.copy_(self._brick_counters, non_blocking=True)
            self._brick_event.record()
            return
        bad = int(self._brick_counters[1].item())
        if bad:
            raise P.PinqError("brick index: %d voxels hold two findable points" % bad)

    def _map_view(self, local: bool) -> P.MapView:
        self._refresh_candidates()
        rec = self._cand_local if local else self._cand_global
        if local and rec is None:
            raise P.PinqError("local query before reset_local_map()/travel_dist were set")
        brec = self._brick_local if local else self._brick_global
        use_brick = self._brick_tab is not None and brec is not None and self._brick_safe()
        return P.MapView(P.ptr(self._occ), P.ptr(rec), self.buffer_size, float(self.resolution),
                         float(self.max_valid_dist2), int(self.neighbor_K), P.ptr(self._stencil_i32),
                         int(self.neural_points.shape[0]), P.ptr(self._brick_tab) if use_brick else None,
                         P.ptr(brec) if use_brick else None, int(self._brick_cap) if use_brick else 0,
                         int(self._stencil_radius) if use_brick else 0,
                         P.ptr(self._stencil_masks()) if use_brick else None)

    def _stencil_masks(self):
        """Brick-voxel membership masks of the current stencil (pinq_stencil_masks_build), built once per
        (num_nei_cells, search_alpha) -- process_frame switches between two stencils every frame."""
        cache = self.__dict__.setdefault("_stencil_mask_cache", {})
        key = self._stencil_key
        if key not in cache:
            out = torch.empty(512, dtype=torch.int64, device=self.device)
            P.check(P.load().pinq_stencil_masks_build(P.ptr(self._stencil_i32), int(self.neighbor_K),
                                                      int(self._stencil_radius), P.ptr(out), P.stream()))
            cache[key] = (out, self._stencil_i32)
        return cache[key][0]

    def _workspace(self, n: int):
        nbytes = int(P.load().pinq_query_workspace_bytes(n))
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        return self._ws

    @staticmethod
    def _as_query(points):
        x = points.detach()
        if x.dtype != torch.float32:
            x = x.float()
        return x.contiguous()

    # ---- neighbour search (API form) ------------------------------------------------------------
    def radius_neighborhood_search(self, points: torch.Tensor, time_filtering: bool = False):
        """reference: model/neural_points.py:459-509 -> (dist2 [N,Kn] f32, neighb_idx [N,Kn] i64 global)."""
        x = self._as_query(points)
        n = x.shape[0]
        d2 = torch.empty((n, self.neighbor_K), dtype=torch.float32, device=self.device)
        idx = torch.empty((n, self.neighbor_K), dtype=torch.int64, device=self.device)
        mv = self._map_view(False)
        P.check(P.load().pinq_neighbor_search(mv, P.ptr(x), n, 0, P.ptr(self.neural_points), P.ptr(self.point_ts_create),
                                              P.ptr(self._travel_dist), int(self.cur_ts),
                                              float(self.diff_travel_dist_local), int(bool(time_filtering)), P.ptr(d2),
                                              P.ptr(idx), P.stream()))
        return d2, idx

    def query_certainty(self, query_points: torch.Tensor):
        """reference: model/neural_points.py:511-525."""
        x = self._as_query(query_points)
        out = torch.empty(x.shape[0], dtype=torch.float32, device=self.device)
        mv = self._map_view(False)
        P.check(P.load().pinq_query_certainty(mv, P.ptr(x), x.shape[0], 0, P.ptr(self.neural_points),
                                              P.ptr(self.point_certainties), P.ptr(out), P.stream()))
        return out

    # ---- query ------------------------------------------------------------------------------------
    def _query_args(self, x, decoder_view, query_ts, training_mode, query_locally):
        flags = 0
        if query_locally:
            flags |= P.F_LOCAL
        if training_mode:
            flags |= P.F_TRAINING
        if self.after_pgo:
            flags |= P.F_ROTATE
        if self.use_mma_sync:
            flags |= P.F_MMA_SYNC
        if self.exact_search_keys:
            flags |= P.F_EXACT_KEYS
        if getattr(self, "_stencil_in_range", False) and getattr(self, "skip_in_range_test", True):
            flags |= P.F_STENCIL_IN_RANGE
        feats = self.local_geo_features if query_locally else self.geo_features
        cert = self.local_point_certainties if query_locally else self.point_certainties
        quats = self.local_point_orientations if query_locally else self.point_orientations
        ts_upd = self.local_point_ts_update if query_locally else None  # global queries never update ts (:647-648)
        a = P.QueryArgs()
        a.map = self._map_view(query_locally)
        if decoder_view is not None:
            a.dec = decoder_view
        a.xyz = P.ptr(x)
        a.n = x.shape[0]
        a.flags = flags
        a.k = int(self.config.query_nn_k)
        a.feature_dim = int(self.geo_feature_dim)
        a.features = P.ptr(feats.detach())
        a.certainties = P.ptr(cert)
        a.quats = P.ptr(quats.contiguous()) if self.after_pgo else None
        pts = self.local_neural_points if query_locally else self.neural_points
        a.point_xyz = P.ptr(pts.contiguous())
        if training_mode and query_ts is not None and ts_upd is not None:
            a.ts_update = P.ptr(ts_upd)
            self._ts_keepalive = query_ts.to(torch.int64).contiguous()
            a.query_ts = P.ptr(self._ts_keepalive)
        a.workspace = P.ptr(self._workspace(x.shape[0]))
        ev = getattr(self, "decode_wait_event", None)
        if ev is not None:  # data-parallel callers: the decode follows this (recorded) event, binning + search do not
            a.decode_wait_event = ev.cuda_event
        return a

    def query_feature(self, query_points: torch.Tensor, query_ts: torch.Tensor = None, training_mode: bool = True,
                      query_locally: bool = True, query_geo_feature: bool = True, query_color_feature: bool = False):
        """Un-fused API form (reference: model/neural_points.py:528-674).  Returns
        (geo_features_vector [N,K,F+3] | [N,F+3], None, weight_vector [N,K,1], nn_counts [N] i64,
        queried_certainty [N]).  Differentiable w.r.t. the feature table (and the query points) through
        `_QueryFeatureFn`; Tracker / Mapper / Mesher use the fused `query_sdf` instead."""
        if not query_geo_feature and not query_color_feature:
            sys.exit("you need to at least query one kind of feature")
        if query_color_feature and self.color_features is not None:
            raise NotImplementedError("colour features are out of scope")
        feats = self.local_geo_features if query_locally else self.geo_features
        fv, w, nn, qc = _QueryFeatureFn.apply(self, query_points, feats, query_ts, training_mode, query_locally)
        w = w.unsqueeze(-1)
        if self.config.weighted_first:
            fv = torch.sum(fv * w, dim=1)
        return fv, None, w, nn, qc

    def query_sdf(self, query_points: torch.Tensor, decoder, query_ts: torch.Tensor = None, training_mode: bool = True,
                  query_locally: bool = True, want_grad: bool = False, want_std: bool = False,
                  want_certainty: bool = False):
        """Fused query: neighbour search + gather + IDW + MLP + weighted combine in one kernel.
        Equivalent to query_feature -> Decoder.sdf -> sum_k w_k s_k (reference: utils/tracker.py:233-253,
        utils/mapper.py:461-468, 671-681).  Returns dict(sdf [N], nn_count [N] i32, and on request
        grad [N,3], std [N], certainty [N]).  `sdf` is differentiable w.r.t. the feature table and the
        decoder parameters (hand-derived backward kernel)."""
        if self.config.weighted_first:
            raise NotImplementedError("fused kernels implement weighted_first=False (run_kitti.yaml:25)")
        feats = self.local_geo_features if query_locally else self.geo_features
        W1, b1, W2, b2 = decoder.fused_params()
        sdf, nn, grad, std, cert = _FusedSDFFn.apply(self, decoder, query_points, feats, W1, b1, W2, b2, query_ts,
                                                    training_mode, query_locally, want_grad, want_std, want_certainty)
        return {"sdf": sdf, "nn_count": nn, "grad": grad if want_grad else None, "std": std if want_std else None,
                "certainty": cert if want_certainty else None}

    def clear_temp(self, clean_more: bool = False):
        """reference: model/neural_points.py:678-693 (called before torch.save of the whole module)."""
        self._table = None
        self._cand_local = None
        self._cand_global = None
        self._cand_dirty = True
        self._index_dirty = True
        self._occ = None
        self._occ_scratch = None
        self._occ_count = None
        self._brick_tab = self._brick_global = self._brick_local = self._brick_buf = self._brick_counters = None
        self._brick_host = self._brick_event = None
        self._brick_pending = False
        self._ws = None
        self._stencil_mask_cache = {}
        self._ts_keepalive = None
        self._local_idx = None
        self.local_neural_points = None
        self.local_point_orientations = None
        self.local_geo_features = nn.Parameter()
        self.local_color_features = nn.Parameter()
        self.local_point_certainties = None
        self.local_point_ts_update = None
        self.local_mask = None
        self.global2local = None
        if clean_more:
            self.point_ts_create = None
            self.point_ts_update = None
            self.point_certainties = None


class _FusedSDFFn(torch.autograd.Function):
    """pinq_query_fused_fwd / pinq_query_fused_bwd as one autograd node."""

    @staticmethod
    def forward(ctx, npts, decoder, points, feats, W1, b1, W2, b2, query_ts, training_mode, query_locally, want_grad,
                want_std, want_certainty):
        lib = P.load()
        dev = npts.device
        x = NeuralPoints._as_query(points)
        n = x.shape[0]
        K = int(npts.config.query_nn_k)
        need_bwd = any(ctx.needs_input_grad[i] for i in (2, 3, 4, 5, 6, 7))
        need_dx = want_grad or ctx.needs_input_grad[2]
        a = npts._query_args(x, decoder.view(), query_ts, training_mode, query_locally)
        sdf = torch.empty(n, dtype=torch.float32, device=dev)
        nn_count = torch.empty(n, dtype=torch.int32, device=dev)
        a.sdf = P.ptr(sdf)
        a.nn_count = P.ptr(nn_count)
        grad = std = cert = stage = relu_mask = perm = None
        if n >= npts.bin_queries_min:  # large batches are processed brick by brick (csrc/sort.cu)
            perm = torch.empty(n, dtype=torch.int32, device=dev)
            a.perm = P.ptr(perm)
        if need_dx:
            grad = torch.empty((n, 3), dtype=torch.float32, device=dev)
            a.grad_x = P.ptr(grad)
        if want_std:
            std = torch.empty(n, dtype=torch.float32, device=dev)
            a.sdf_std = P.ptr(std)
        if need_bwd:
            # what the backward re-reads: the stage blocks (K indices, IDW weights, x - p_k; 124 B/query) and the
            # relu'(pre) bits of the hidden layer (48 B/query) -- nothing of the decoder is re-evaluated
            stage = torch.empty(int(lib.pinq_stage_bytes(n)), dtype=torch.uint8, device=dev)
            a.stage = P.ptr(stage)
            relu_mask = torch.empty(int(lib.pinq_relu_mask_words(n)), dtype=torch.int32, device=dev)
            a.relu_mask = P.ptr(relu_mask)
        if want_certainty:
            cert = torch.empty(n, dtype=torch.float32, device=dev)
            a.certainty = P.ptr(cert)
        g_feats = None
        if (ctx.needs_input_grad[3] and npts.zero_grad_in_search and getattr(npts, "feature_grad_sink", None) is None
                and feats.dtype == torch.float32 and feats.is_contiguous() and feats.numel() % 4 == 0
                and 16 * n >= 4 * feats.numel()):
            # the feature-gradient block of THIS query's backward, cleared by the neighbour search on its way
            # (pinq_query_args.zero_fill) instead of a 53 MB memset between forward and backward
            g_feats = torch.empty_like(feats)
            a.zero_fill, a.zero_fill_bytes = P.ptr(g_feats), 4 * feats.numel()
        P.check(lib.pinq_query_fused_fwd(a, P.stream()))
        ctx.g_feats = g_feats
        ctx.npts, ctx.decoder, ctx.query_locally = npts, decoder, query_locally
        ctx.after_pgo = npts.after_pgo
        if need_bwd:
            ctx.n = n
            ctx.save_for_backward(feats, stage, relu_mask, perm, grad if ctx.needs_input_grad[2] else None)
        empty = torch.empty(0, device=dev)
        outs = (sdf, nn_count, grad if grad is not None else empty, std if std is not None else empty,
                cert if cert is not None else empty)
        ctx.mark_non_differentiable(*outs[1:])
        return outs

    @staticmethod
    def backward(ctx, g_sdf, *unused):
        lib = P.load()
        npts, decoder = ctx.npts, ctx.decoder
        feats, stage, relu_mask, perm, dx = ctx.saved_tensors
        g = g_sdf.contiguous().float()
        b = P.BwdArgs()
        b.dec = decoder.view()
        b.n = ctx.n
        b.flags = P.F_MMA_SYNC if npts.use_mma_sync else 0
        b.k = int(npts.config.query_nn_k)
        b.feature_dim = int(npts.geo_feature_dim)
        b.features = P.ptr(feats.detach())
        b.stage = P.ptr(stage)
        b.relu_mask = P.ptr(relu_mask)
        b.perm = P.ptr(perm)
        b.grad_sdf = P.ptr(g)
        g_feats = gW1 = gb1 = gW2 = gb2 = None
        # gradient sinks (data-parallel runs, parallel.PeerExchange): the kernel accumulates straight into caller-owned
        # buffers -- views of the rank's peer block -- and autograd sees no gradient for those inputs (p.grad stays None)
        f_sink = getattr(npts, "feature_grad_sink", None)
        d_sink = getattr(decoder, "grad_sink", None)
        if ctx.needs_input_grad[3]:
            if f_sink is not None:
                if f_sink.shape != feats.shape or f_sink.dtype != torch.float32:
                    raise P.PinqError("feature_grad_sink must be a float32 tensor shaped like local_geo_features")
                b.grad_features = P.ptr(f_sink)
            else:
                g_feats, ctx.g_feats = ctx.g_feats, None  # cleared by the forward's search; a second backward gets a fresh one
                if g_feats is None:
                    g_feats = torch.zeros_like(feats)
                b.grad_features = P.ptr(g_feats)
        if any(ctx.needs_input_grad[i] for i in (4, 5, 6, 7)):  # decoder not frozen (hazard H11)
            W1, b1, W2, b2 = decoder.fused_params()
            if d_sink is not None:
                b.grad_W1, b.grad_b1, b.grad_W2, b.grad_b2 = (P.ptr(t) for t in d_sink)
            else:
                # one zero-filled buffer (one memset launch instead of four), the four gradients are views of it
                sizes = [t.numel() for t in (W1, b1, W2, b2)]
                offs = [0]
                for sz in sizes:
                    offs.append(offs[-1] + ((sz + 3) // 4) * 4)  # 16-byte aligned starts
                flat = torch.zeros(offs[-1], dtype=W1.dtype, device=W1.device)
                gW1, gb1, gW2, gb2 = (flat[o:o + sz].view(t.shape) for o, sz, t in zip(offs, sizes, (W1, b1, W2, b2)))
                b.grad_W1, b.grad_b1, b.grad_W2, b.grad_b2 = P.ptr(gW1), P.ptr(gb1), P.ptr(gW2), P.ptr(gb2)
        P.check(lib.pinq_query_fused_bwd(b, P.stream()))
        g_x = g.unsqueeze(-1) * dx if ctx.needs_input_grad[2] else None
        return (None, None, g_x, g_feats, gW1, gb1, gW2, gb2, None, None, None, None, None, None)


class _QueryFeatureFn(torch.autograd.Function):
    """Un-fused query_feature with a (torch-op) backward into the feature table; compatibility path only."""

    @staticmethod
    def forward(ctx, npts, points, feats, query_ts, training_mode, query_locally):
        lib = P.load()
        dev = npts.device
        x = NeuralPoints._as_query(points)
        n = x.shape[0]
        K, F = int(npts.config.query_nn_k), int(npts.geo_feature_dim)
        a = npts._query_args(x, None, query_ts, training_mode, query_locally)
        fv = torch.empty((n, K, F + 3), dtype=torch.float32, device=dev)
        w = torch.empty((n, K), dtype=torch.float32, device=dev)
        nn32 = torch.empty(n, dtype=torch.int32, device=dev)
        cert = torch.empty(n, dtype=torch.float32, device=dev)
        knn_idx = torch.empty((n, K), dtype=torch.int32, device=dev)
        a.knn_w, a.nn_count, a.certainty, a.knn_idx = P.ptr(w), P.ptr(nn32), P.ptr(cert), P.ptr(knn_idx)
        P.check(lib.pinq_query_feature(a, P.ptr(fv), P.stream()))
        nn64 = nn32.to(torch.int64)
        ctx.save_for_backward(knn_idx)
        ctx.rows = feats.shape[0]
        ctx.F = F
        ctx.mark_non_differentiable(w, nn64, cert)
        return fv, w, nn64, cert

    @staticmethod
    def backward(ctx, g_fv, g_w, g_nn, g_c):
        (knn_idx,) = ctx.saved_tensors
        g_feats = None
        if ctx.needs_input_grad[2]:
            valid = knn_idx >= 0
            g_feats = torch.zeros((ctx.rows, ctx.F), dtype=g_fv.dtype, device=g_fv.device)
            g_feats.index_add_(0, knn_idx[valid].to(torch.int64), g_fv[..., : ctx.F][valid])
        if ctx.needs_input_grad[1]:
            raise NotImplementedError("d/dx through the un-fused query_feature: use query_sdf(want_grad=True)")
        return None, None, g_feats, None, None, None
