"""Mapper -- online training of the neural-point map (reference: utils/mapper.py:27-733).

Same constructor, attributes (`lose_track`, `static_mask`, `used_poses`, the `*_pool` tensors,
`train_less`, `new_idx`) and method names as the reference.  `mapping()` runs, per iteration, the fused
query kernels for the 16 384 pool samples and for the 6 x ceil(bs/10) numerical-gradient samples
(training mode, hazard H9), the BCE + eikonal loss on [N] tensors in PyTorch, the hand-derived backward
kernel (feature scatter-add + decoder gradients, skipped per flag when the decoder is frozen, H11) and
the Adam step -- all of it inside `pinq_mapping_run` (one device-side loop per frame; data-parallel through `Mapper.dp`).
`process_frame()` keeps the replay pool with the kernels of csrc/pool.cu (`_pool_update_torch` is the
reference's torch formulation, kept as parity partner) and uses the libpinq kernels for the ray sampler, the map update and
the new-sample certainty query.
Bundle adjustment (pypose), semantics, colour and the Open3D helpers are out of scope.
"""
from __future__ import annotations

import math

import ctypes as C

import torch

from .. import _pinq as P
from .data_sampler import DataSampler
from .loss import sdf_bce_loss
from .tools import get_time, setup_optimizer, transform_torch


class Mapper:
    def __init__(self, config, dataset, neural_points, geo_mlp, sem_mlp, color_mlp):
        self.config = config
        self.silence = config.silence
        self.dataset = dataset
        self.neural_points = neural_points
        self.geo_mlp = geo_mlp
        self.sem_mlp = sem_mlp
        self.color_mlp = color_mlp
        self.device = config.device
        self.dtype = config.dtype
        self.used_poses = None
        self.lose_track = False
        self.require_gradient = False
        if config.ekional_loss_on or config.proj_correction_on or config.consistency_loss_on:
            self.require_gradient = True
        if config.numerical_grad and not config.proj_correction_on and not config.consistency_loss_on:
            self.require_gradient = False
        if self.require_gradient or config.proj_correction_on or config.consistency_loss_on:
            raise NotImplementedError("analytic-gradient / projection-correction / consistency losses are unused by "
                                      "the reference configs and out of scope")
        self.total_iter = 0
        self.sdf_scale = config.logistic_gaussian_ratio * config.sigma_sigmoid_m
        self.sampler = DataSampler(config)
        self.ray_sample_count = 1 + config.surface_sample_n + config.free_behind_n + config.free_front_n
        self.new_idx = None
        self.ba_done_flag = False
        self.train_less = False
        self.static_mask = None
        self.generator = None  # optional torch.Generator for reproducible sampling / batches
        dev, dt = self.device, self.dtype
        self.coord_pool = torch.empty((0, 3), device=dev, dtype=dt)
        self.global_coord_pool = torch.empty((0, 3), device=dev, dtype=dt)
        self.sdf_label_pool = torch.empty((0), device=dev, dtype=dt)
        self.color_pool = None
        self.sem_label_pool = None
        self.normal_label_pool = None
        self.weight_pool = torch.empty((0), device=dev, dtype=dt)
        self.time_pool = torch.empty((0), device=dev, dtype=torch.long)
        self.cur_sample_count = 0
        self.pool_sample_count = 0
        self.timing = {}
        self.fused_mapping = True   # False: run the iterations as separate autograd steps (A/B tests)
        self.fused_pool = True      # False: replay-pool upkeep with torch ops (A/B tests)
        self._pool_store = [None, None]
        self._pool_ws = None
        self.dp = None              # parallel.DataParallelMapping: shard every batch over the ranks
        self.last_loss = None
        self._fused_buf, self._fused_last = {}, None
        self._draw_host, self._draw_event = {}, None

    # ------------------------------------------------------------------------------------------
    def dynamic_filter(self, points_torch, type_2_on: bool = False):
        """reference: utils/mapper.py:79-108."""
        with torch.no_grad():
            out = self.neural_points.query_sdf(points_torch, self.geo_mlp, None, False, True, want_grad=type_2_on,
                                               want_certainty=True)
        cfg = self.config
        static_mask = (out["certainty"] < cfg.dynamic_certainty_thre) | (out["sdf"] < cfg.dynamic_sdf_ratio_thre * cfg.voxel_size_m)
        if type_2_on:
            grad_norm = out["grad"].norm(dim=-1)
            static_mask = static_mask & ((grad_norm > 0.3) | (out["certainty"] < 0.5))
        return static_mask

    def _randint(self, high, size):
        if self.generator is not None:
            return torch.randint(0, high, size, device=self.generator.device, generator=self.generator).to(self.device)
        return torch.randint(0, high, size, device=self.device)

    def process_frame(self, point_cloud_torch, frame_label_torch, cur_pose_torch, frame_id: int,
                      filter_dynamic: bool = False, samples=None):
        """reference: utils/mapper.py:110-321.  `samples` = (coord, sdf_label, weight) overrides the sampler
        (used to feed the CUDA path and the oracle identical draws)."""
        cfg = self.config
        T0 = get_time()
        frame_origin_torch = cur_pose_torch[:3, 3]
        frame_orientation_torch = cur_pose_torch[:3, :3]
        frame_point_torch = point_cloud_torch[:, :3]
        self.static_mask = torch.ones(frame_point_torch.shape[0], dtype=torch.bool, device=self.device)
        if filter_dynamic:
            self.static_mask = self.dynamic_filter(transform_torch(frame_point_torch, cur_pose_torch))
            frame_point_torch = frame_point_torch[self.static_mask]
        T1 = get_time()
        if samples is None:
            coord, sdf_label, _, _, _, weight = self.sampler.sample(frame_point_torch, None, None, None,
                                                                    generator=self.generator)
        else:
            coord, sdf_label, weight = samples
        time_repeat = torch.full((coord.shape[0],), int(frame_id), dtype=torch.long, device=self.device)
        self.cur_sample_count = sdf_label.shape[0]
        self.pool_sample_count = self.sdf_label_pool.shape[0]
        T2 = get_time()
        if cfg.from_sample_points:
            if cfg.from_all_samples:
                update_points = coord
            else:
                update_points = coord[torch.abs(sdf_label) < cfg.surface_sample_range_m * cfg.map_surface_ratio, :]
                update_points = transform_torch(update_points.contiguous(), cur_pose_torch)
        else:
            update_points = transform_torch(frame_point_torch.contiguous(), cur_pose_torch)
        if cfg.prune_map_on and self.neural_points.prune_map(cfg.max_prune_certainty):
            self.neural_points.recreate_hash(None, None, True, True, frame_id)
        self.neural_points.update(update_points, frame_origin_torch, frame_orientation_torch, frame_id)
        T3 = get_time()
        if cfg.track_on and getattr(self.dataset, "odom_poses", None):
            import numpy as np
            self.used_poses = torch.tensor(np.array(self.dataset.odom_poses), device=self.device, dtype=torch.float64)
        elif getattr(self.dataset, "gt_pose_provided", False):
            import numpy as np
            self.used_poses = torch.tensor(np.array(self.dataset.gt_poses), device=self.device, dtype=torch.float64)
        global_coord = transform_torch(coord.contiguous(), cur_pose_torch)
        T3_1 = T3
        if self.fused_pool and (frame_id + 1) % cfg.pool_filter_freq == 0:
            self._pool_update_fused(coord, global_coord, sdf_label, weight, int(frame_id), frame_origin_torch)
        else:
            T3_1 = self._pool_update_torch(coord, global_coord, sdf_label, weight, time_repeat, frame_id, frame_origin_torch)
        T3_2 = get_time()
        if cfg.bs_new_sample > 0:  # mapper.py:274-307: learn more where the map is new
            cur = self.global_coord_pool[-self.cur_sample_count:]
            n_cur = cur.shape[0]
            cert = torch.zeros(n_cur, device=self.device)
            cur_label = self.sdf_label_pool[-self.cur_sample_count:]
            self.neural_points.set_search_neighborhood(num_nei_cells=1, search_alpha=0.0)
            bs = cfg.infer_bs
            for head in range(0, n_cur, bs):
                tail = min(head + bs, n_cur)
                cert[head:tail] = self.neural_points.query_certainty(cur[head:tail].contiguous())
            self.neural_points.set_search_neighborhood(num_nei_cells=cfg.num_nei_cells, search_alpha=cfg.search_alpha)
            self.new_idx = torch.where((cert < cfg.new_certainty_thre) &
                                       (torch.abs(cur_label) < cfg.surface_sample_range_m * 3.0))[0]
            self.new_idx += (self.pool_sample_count - self.cur_sample_count)
            new_sample_count = self.new_idx.shape[0]
            self.train_less = bool(cfg.adaptive_mode and new_sample_count / self.cur_sample_count < cfg.new_sample_ratio_thre)
        T4 = get_time()
        self.timing = {"dynamic_filter_ms": (T1 - T0) * 1e3, "sampling_ms": (T2 - T1) * 1e3,
                       "map_update_ms": (T3 - T2) * 1e3, "pool_ms": (T4 - T3) * 1e3,
                       "pool_transform_ms": (T3_1 - T3) * 1e3, "pool_filter_ms": (T3_2 - T3_1) * 1e3}

    def _pool_update_torch(self, coord, global_coord, sdf_label, weight, time_repeat, frame_id, frame_origin_torch):
        """Replay-pool upkeep with torch ops, the reference's formulation (utils/mapper.py:209-263): the parity partner
        of `_pool_update_fused` and the path for frames on which the pool is not filtered."""
        cfg = self.config
        self.coord_pool = torch.cat((self.coord_pool, coord), 0)
        self.weight_pool = torch.cat((self.weight_pool, weight), 0)
        self.sdf_label_pool = torch.cat((self.sdf_label_pool, sdf_label), 0)
        self.time_pool = torch.cat((self.time_pool, time_repeat), 0)
        self.global_coord_pool = torch.cat((self.global_coord_pool, global_coord), 0)
        T3_1 = get_time()
        if (frame_id + 1) % cfg.pool_filter_freq == 0:  # mapper.py:229-263
            rel = self.global_coord_pool - frame_origin_torch.to(self.global_coord_pool)
            filter_mask = torch.sum(rel ** 2, dim=-1) < cfg.window_radius ** 2
            true_indices = torch.nonzero(filter_mask).squeeze()
            pool_sample_count = true_indices.shape[0]
            if pool_sample_count > cfg.pool_capacity:
                discard_count = pool_sample_count - cfg.pool_capacity
                discarded = self._randint(pool_sample_count, (discard_count,))
                filter_mask[true_indices[discarded]] = False
            self.coord_pool = self.coord_pool[filter_mask]
            self.global_coord_pool = self.global_coord_pool[filter_mask]
            self.sdf_label_pool = self.sdf_label_pool[filter_mask]
            self.weight_pool = self.weight_pool[filter_mask]
            self.time_pool = self.time_pool[filter_mask]
            self.cur_sample_count = int(filter_mask[-self.cur_sample_count:].sum().item())
            self.pool_sample_count = int(filter_mask.sum().item())
        else:
            self.cur_sample_count = coord.shape[0]
            self.pool_sample_count = self.coord_pool.shape[0]
        return T3_1

    def _pool_rows_target(self, rows, n_new=0):
        """Rows to allocate for a pool buffer that must hold `rows` now.  The pool never exceeds pool_capacity after the
        discard (utils/mapper.py:237-246), so [old pool | new samples] is bounded by pool_capacity + one frame of samples:
        allocate that ONCE (run_kitti: 2e7 rows x 40 B = 0.9 GB per half of the ping-pong store) instead of doubling --
        every doubling (cudaMalloc + the free of the old half) stalled a frame by 6-33 ms (profiles/r4a_driver100.json: 8 such
        frames in 100).  Capacities too large to reserve up front (> 6e7 rows) keep the geometric growth."""
        cap_cfg = int(self.config.pool_capacity)
        self._pool_frame_rows = max(getattr(self, "_pool_frame_rows", 0), int(n_new))
        if cap_cfg <= 60_000_000:
            bound = cap_cfg + max(2 * self._pool_frame_rows, 1 << 21)
            if rows <= bound:
                return bound
        return max(2 * int(rows), 1 << 22 if rows > (1 << 16) else 1 << 16)

    def _pool_store_side(self, rows):
        """The half of the ping-pong store the next pool is written to: not the one the pool tensors live in now."""
        live = self.coord_pool.data_ptr() if self.coord_pool.numel() else 0
        side = 0
        if self._pool_store[0] is not None and live == self._pool_store[0]["coord"].data_ptr():
            side = 1
        st = self._pool_store[side]
        if st is None or st["ts"].shape[0] < rows:
            cap = self._pool_rows_target(rows)
            dev = self.device
            st = {"coord": torch.empty((cap, 3), dtype=torch.float32, device=dev),
                  "global_coord": torch.empty((cap, 3), dtype=torch.float32, device=dev),
                  "sdf_label": torch.empty(cap, dtype=torch.float32, device=dev),
                  "weight": torch.empty(cap, dtype=torch.float32, device=dev),
                  "ts": torch.empty(cap, dtype=torch.int64, device=dev)}
            self._pool_store[side] = st
        return st

    def _pool_update_fused(self, coord, global_coord, sdf_label, weight, frame_id, frame_origin_torch):
        """utils/mapper.py:209-263 on libpinq (csrc/pool.cu): window flags over [old pool | new samples] -> order-preserving
        compaction [-> random discard above pool_capacity, same torch.randint draw as the reference] -> one move of the
        survivors into the other half of a preallocated ping-pong store.  One host read-back (two counters) per frame."""
        import time as _time
        cfg, dev = self.config, self.device
        lib = P.load()
        t_begin = _time.perf_counter()
        n_old, n_new = int(self.sdf_label_pool.shape[0]), int(sdf_label.shape[0])
        n_tot = n_old + n_new
        if n_tot == 0:
            self.cur_sample_count = self.pool_sample_count = 0
            return
        f32 = lambda t: t.contiguous().float()  # noqa: E731
        old = P.PoolView(P.ptr(f32(self.coord_pool)), P.ptr(f32(self.global_coord_pool)), P.ptr(f32(self.sdf_label_pool)),
                         P.ptr(f32(self.weight_pool)), P.ptr(self.time_pool.contiguous())) if n_old else P.PoolView()
        c_new, g_new, l_new, w_new = f32(coord), f32(global_coord), f32(sdf_label), f32(weight)
        new = P.PoolView(P.ptr(c_new), P.ptr(g_new), P.ptr(l_new), P.ptr(w_new), None) if n_new else P.PoolView()
        origin = frame_origin_torch.to(device=dev, dtype=torch.float32).contiguous()
        ws = self._pool_ws
        if ws is None or ws["rows"] < n_tot:  # work buffers persist across frames: sized once for the steady state
            rows = self._pool_rows_target(n_tot, n_new)
            ws = self._pool_ws = {"rows": rows, "flags": torch.empty(rows, dtype=torch.uint8, device=dev),
                                  "kept": torch.empty(rows, dtype=torch.int32, device=dev),
                                  "scratch": torch.empty(int(lib.pinq_compact_scratch_bytes(rows)), dtype=torch.uint8, device=dev)}
        flags, kept, scratch = ws["flags"], ws["kept"], ws["scratch"]
        counts = torch.zeros(2, dtype=torch.int32, device=dev)  # [survivors, survivors among the new samples]
        st = P.stream()
        P.check(lib.pinq_pool_flags(old.global_coord, n_old, new.global_coord, n_new, P.ptr(origin),
                                    float(cfg.window_radius ** 2), P.ptr(flags), st))
        P.check(lib.pinq_compact_flags(P.ptr(flags), n_tot, P.ptr(kept), P.ptr(counts[0:1]), P.ptr(scratch), st))
        if n_tot > cfg.pool_capacity:  # only then can the survivors exceed the capacity (mapper.py:237-246)
            survivors = int(counts[0].item())
            if survivors > cfg.pool_capacity:
                discarded = self._randint(survivors, (survivors - int(cfg.pool_capacity),)).contiguous()
                P.check(lib.pinq_pool_discard(P.ptr(flags), P.ptr(kept), P.ptr(discarded), discarded.shape[0], st))
                P.check(lib.pinq_compact_flags(P.ptr(flags), n_tot, P.ptr(kept), P.ptr(counts[0:1]), P.ptr(scratch), st))
        t_enq = _time.perf_counter()
        dst = self._pool_store_side(n_tot)
        t_store = _time.perf_counter()
        out = P.PoolView(P.ptr(dst["coord"]), P.ptr(dst["global_coord"]), P.ptr(dst["sdf_label"]), P.ptr(dst["weight"]), P.ptr(dst["ts"]))
        P.check(lib.pinq_pool_move(old, n_old, new, n_new, frame_id, P.ptr(kept), P.ptr(counts[0:1]), out, P.ptr(counts[1:2]), st))
        t_move = _time.perf_counter()
        total, kept_new = (int(v) for v in counts.tolist())
        t_read = _time.perf_counter()
        self.pool_timing = {"flags_compact_enqueue_ms": (t_enq - t_begin) * 1e3, "store_ms": (t_store - t_enq) * 1e3,
                            "move_enqueue_ms": (t_move - t_store) * 1e3, "readback_ms": (t_read - t_move) * 1e3, "rows": n_tot}
        self.coord_pool, self.global_coord_pool = dst["coord"][:total], dst["global_coord"][:total]
        self.sdf_label_pool, self.weight_pool, self.time_pool = dst["sdf_label"][:total], dst["weight"][:total], dst["ts"][:total]
        self.cur_sample_count, self.pool_sample_count = kept_new, total

    def get_batch(self, global_coord=False):
        """reference: utils/mapper.py:323-361."""
        cfg = self.config
        stop_status = getattr(self.dataset, "stop_status", False)
        if cfg.bs_new_sample > 0 and self.new_idx is not None and not self.lose_track and not stop_status:
            new_idx_count = self.new_idx.shape[0]
            if new_idx_count > 0:
                bs_new = min(new_idx_count, cfg.bs_new_sample)
                index_history = self._randint(self.pool_sample_count, (cfg.bs - bs_new,))
                index_new = self.new_idx[self._randint(new_idx_count, (bs_new,))]
                index = torch.cat((index_history, index_new), dim=0)
            else:
                index = self._randint(self.pool_sample_count, (cfg.bs,))
        else:
            index = self._randint(self.pool_sample_count, (cfg.bs,))
        coord = self.global_coord_pool[index, :] if global_coord else self.coord_pool[index, :]
        return coord, self.sdf_label_pool[index], self.time_pool[index], None, None, None, self.weight_pool[index]

    def free_pool(self):
        self.coord_pool = self.weight_pool = self.sdf_label_pool = self.time_pool = None
        self.sem_label_pool = self.color_pool = self.normal_label_pool = None

    # ------------------------------------------------------------------------------------------
    def mapping(self, iter_count):
        """reference: utils/mapper.py:425-593.  The iterations run inside libpinq (`pinq_mapping_run`: batch gather,
        fused query, closed-form loss gradient, backward kernel, multi-tensor Adam -- 8 launches per iteration and no
        host work between them); configurations the fused loop does not cover take the autograd form of the same
        kernels (`_mapping_autograd`)."""
        if self.train_less:
            iter_count = max(1, iter_count - 5)
        if self.fused_mapping and self._fused_mapping_supported():
            self._mapping_fused(iter_count)
        else:
            self._mapping_autograd(iter_count)
        self.neural_points.assign_local_to_global()

    def _mapping_autograd(self, iter_count):
        """The reference's loop verbatim (get_batch -> loss -> backward -> Adam) over the fused autograd node."""
        cfg = self.config
        neural_point_feat = list(self.neural_points.parameters())
        geo_mlp_param = list(self.geo_mlp.parameters())
        opt = setup_optimizer(cfg, neural_point_feat, geo_mlp_param, None, None)
        for _ in range(iter_count):
            coord, sdf_label, ts, _, _, _, weight = self.get_batch(global_coord=not self.ba_done_flag)
            cur_loss, _ = self.mapping_loss(coord, sdf_label, ts, weight)
            opt.zero_grad(set_to_none=True)
            cur_loss.backward(retain_graph=False)
            opt.step()
            self.total_iter += 1

    def _trainable_decoder(self):
        """None if the decoder is frozen (H11), the four fused tensors if all train, False for a partial freeze."""
        params = list(self.geo_mlp.parameters())
        trainable = [p for p in params if p.requires_grad]
        if not trainable:
            return None
        fused = self.geo_mlp.fused_params() if self.geo_mlp.fused_supported() else ()
        if len(trainable) == len(params) == 4 and all(a is b for a, b in zip(params, fused)):
            return fused
        return False

    def _fused_mapping_supported(self) -> bool:
        cfg = self.config
        return bool(cfg.opt_adam and cfg.main_loss_type == "bce" and not cfg.weighted_first and self.geo_mlp.fused_supported()
                    and (not cfg.ekional_loss_on or cfg.weight_e <= 0 or cfg.numerical_grad)
                    and cfg.ekional_add_to in ("all", "surface", "freespace") and self._trainable_decoder() is not False
                    and self.neural_points.local_geo_features.requires_grad and self.pool_sample_count > 0
                    and self.sdf_label_pool.shape[0] == self.pool_sample_count)

    def _draw_batches(self, iters, bs_hist, bs_new, new_count):
        """The uniform indices of all iterations, drawn in get_batch's order (history, then new samples, per
        iteration) so a seeded generator yields the sequence the per-iteration form would."""
        gen, dev = self.generator, self.device
        if gen is None:
            hist = torch.randint(0, self.pool_sample_count, (iters, bs_hist), device=dev)
            new = torch.randint(0, new_count, (iters, bs_new), device=dev) if bs_new > 0 else None
            return hist, new
        on_cpu = gen.device.type == "cpu"
        if on_cpu and self._draw_event is not None:
            self._draw_event.synchronize()  # the previous frame's upload has left the pinned buffer
        key = (iters, bs_hist, bs_new)
        if on_cpu and self._draw_host.get("key") != key:
            self._draw_host = {"key": key, "hist": torch.empty((iters, bs_hist), dtype=torch.int64).pin_memory(),
                               "new": torch.empty((iters, max(bs_new, 1)), dtype=torch.int64).pin_memory()}
        hist_rows, new_rows = [], []
        for it in range(iters):
            h = torch.randint(0, self.pool_sample_count, (bs_hist,), device=gen.device, generator=gen)
            n = torch.randint(0, new_count, (bs_new,), device=gen.device, generator=gen) if bs_new > 0 else None
            if on_cpu:
                self._draw_host["hist"][it].copy_(h)
                if n is not None:
                    self._draw_host["new"][it].copy_(n)
            else:
                hist_rows.append(h)
                new_rows.append(n)
        if on_cpu:
            hist = self._draw_host["hist"].to(dev, non_blocking=True)
            new = self._draw_host["new"].to(dev, non_blocking=True) if bs_new > 0 else None
            self._draw_event = torch.cuda.Event()
            self._draw_event.record()
            return hist, new
        return torch.stack(hist_rows).to(dev), (torch.stack(new_rows).to(dev) if bs_new > 0 else None)

    def _mapping_fused(self, iters):
        cfg, npts, dev = self.config, self.neural_points, self.device
        lib = P.load()
        dp = self.dp
        bs = int(cfg.bs)
        bs_new = 0
        stop_status = getattr(self.dataset, "stop_status", False)
        if cfg.bs_new_sample > 0 and self.new_idx is not None and not self.lose_track and not stop_status:
            bs_new = min(int(self.new_idx.shape[0]), int(cfg.bs_new_sample))  # get_batch, mapper.py:336-343
        bs_hist = bs - bs_new
        eik_on = bool(cfg.ekional_loss_on and cfg.weight_e > 0)
        decim = int(cfg.gradient_decimation) if eik_on else 0
        hist, new = self._draw_batches(iters, bs_hist, bs_new, int(self.new_idx.shape[0]) if bs_new > 0 else 0)
        global_bs, first = bs, 0
        if dp is not None:
            # data-parallel form: the draws above are the GLOBAL batch (shared seed: identical on every rank); this rank
            # trains on the contiguous slice [first, last) of every row
            first, last = dp.shard(bs)
            hist = hist[:, min(first, bs_hist):min(last, bs_hist)].contiguous()
            if bs_new > 0:
                new = new[:, max(first, bs_hist) - bs_hist:max(last, bs_hist) - bs_hist].contiguous()
            bs, bs_hist = last - first, hist.shape[1]
            bs_new = bs - bs_hist
            if bs_new == 0:
                new = None
        eik_start = (decim - first % decim) % decim if decim > 0 else 0
        ng = (bs - eik_start + decim - 1) // decim if decim > 0 and eik_start < bs else 0
        n = bs + 6 * ng
        pool = (self.coord_pool if self.ba_done_flag else self.global_coord_pool).contiguous()
        buf = self._fused_buf
        if buf.get("n") != n or buf.get("bs") != bs:
            buf.clear()
            buf.update(n=n, bs=bs, xyz=torch.empty((n, 3), dtype=torch.float32, device=dev),
                       ts=torch.empty(n, dtype=torch.int64, device=dev),
                       label=torch.empty(bs, dtype=torch.float32, device=dev),
                       weight=torch.empty(bs, dtype=torch.float32, device=dev),
                       sdf=torch.empty(n, dtype=torch.float32, device=dev),
                       nn=torch.empty(n, dtype=torch.int32, device=dev),
                       grad_sdf=torch.empty(n, dtype=torch.float32, device=dev),
                       stage=torch.empty(int(lib.pinq_stage_bytes(n)), dtype=torch.uint8, device=dev),
                       relu=torch.empty(int(lib.pinq_relu_mask_words(n)), dtype=torch.int32, device=dev),
                       perm=torch.empty(n, dtype=torch.int32, device=dev) if n >= npts.bin_queries_min else None)
        decoder = self._trainable_decoder()
        dec_n = sum(int(t.numel()) for t in decoder) if decoder else 0
        nf = npts.local_geo_features.numel()
        if dp is not None:
            # parameters and gradients live in this rank's peer block for the duration of the call (csrc/peer.cu)
            moved = dp.begin(npts, self.geo_mlp if decoder else None)
        feats = npts.local_geo_features.data
        # one zero-filled block: [feature + decoder gradients,] loss log, eikonal counters; Adam leaves the grads zero
        n_own_grads = 0 if dp is not None else nf + dec_n
        zero = torch.zeros(n_own_grads + 2 * iters + iters, dtype=torch.float32, device=dev)
        moments = (torch.zeros if dp is not None else torch.empty)(2 * (nf + dec_n) + 64, dtype=torch.float32, device=dev)
        grad_block = moved["grads"] if dp is not None else zero
        g_feats, off = grad_block[:nf], nf
        a = npts._query_args(buf["xyz"], self.geo_mlp.view(), buf["ts"], True, True)
        a.sdf, a.nn_count, a.stage, a.relu_mask, a.perm = (P.ptr(buf["sdf"]), P.ptr(buf["nn"]), P.ptr(buf["stage"]),
                                                           P.ptr(buf["relu"]), P.ptr(buf["perm"]))
        m = P.MappingArgs()
        m.query = a
        b = m.bwd
        b.dec, b.n, b.flags, b.k, b.feature_dim = a.dec, n, (P.F_MMA_SYNC if npts.use_mma_sync else 0), a.k, a.feature_dim
        b.features, b.stage, b.relu_mask, b.perm = a.features, a.stage, a.relu_mask, a.perm
        b.grad_sdf, b.grad_features = P.ptr(buf["grad_sdf"]), P.ptr(g_feats)
        tensors = [(feats, g_feats)]
        if decoder:
            grads = []
            for t in decoder:
                grads.append(grad_block[off: off + t.numel()])
                off += t.numel()
                tensors.append((t.data, grads[-1]))
            b.grad_W1, b.grad_b1, b.grad_W2, b.grad_b2 = (P.ptr(g) for g in grads)
        loss = zero[n_own_grads: n_own_grads + 2 * iters]
        eik_count = zero[n_own_grads + 2 * iters:].view(torch.int32)
        mo = 0
        for i, (p, g) in enumerate(tensors):
            if not p.is_contiguous():
                raise P.PinqError("fused mapping needs contiguous parameters")
            cnt = p.numel()
            pad = (cnt + 3) // 4 * 4   # keeps every moment slice 16-byte aligned (vectorised Adam)
            m.tensors[i] = P.AdamTensor(P.ptr(p), P.ptr(g), P.ptr(moments[mo: mo + cnt]), P.ptr(moments[mo + pad: mo + pad + cnt]), cnt)
            mo += 2 * pad
        m.tensor_count = len(tensors)
        m.pool_xyz, m.pool_label, m.pool_weight, m.pool_ts = (P.ptr(pool), P.ptr(self.sdf_label_pool.contiguous()),
                                                              P.ptr(self.weight_pool.contiguous()), P.ptr(self.time_pool.contiguous()))
        m.draws_hist, m.draws_new = P.ptr(hist) if bs_hist > 0 else None, P.ptr(new) if bs_new > 0 else None
        m.new_idx = P.ptr(self.new_idx.contiguous()) if bs_new > 0 else None
        m.bs, m.bs_new, m.decimation = bs, bs_new, decim
        m.eik_mode = {"all": P.EIK_ALL, "surface": P.EIK_SURFACE, "freespace": P.EIK_FREESPACE}[cfg.ekional_add_to]
        m.weighted = int(bool(cfg.loss_weight_on))
        m.eps_step = float(cfg.voxel_size_m * cfg.num_grad_step_ratio)
        m.sigma, m.weight_e, m.surface_range = float(self.sdf_scale), float(cfg.weight_e), float(cfg.surface_sample_range_m)
        m.batch_xyz, m.batch_ts, m.batch_label, m.batch_weight = (P.ptr(buf["xyz"]), P.ptr(buf["ts"]), P.ptr(buf["label"]),
                                                                  P.ptr(buf["weight"]))
        m.grad_sdf, m.loss, m.eik_count = P.ptr(buf["grad_sdf"]), P.ptr(loss), P.ptr(eik_count)
        m.lr, m.beta1, m.beta2, m.adam_eps, m.weight_decay = float(cfg.lr), 0.9, 0.99, float(cfg.adam_eps), float(cfg.weight_decay)
        if dp is not None:
            m.global_bs, m.shard_first = global_bs, first
            exp_avg, exp_avg_sq = moments[:nf + dec_n], moments[(nf + dec_n + 3) // 4 * 4: (nf + dec_n + 3) // 4 * 4 + nf + dec_n]
            pa = dp.peer_args(iters, n_grad=nf + dec_n, exp_avg=exp_avg, exp_avg_sq=exp_avg_sq)
            m.peer = C.pointer(pa)
        try:
            P.check(lib.pinq_mapping_run(m, int(iters), P.stream()))
        except Exception:
            if dp is not None:
                dp.abort(moved)  # parameters back into torch-owned memory before the error travels up
            raise
        if dp is not None:
            dp.end(npts, self.geo_mlp if decoder else None, moved)  # certainty / ts_update exchange, parameters back home
        self.total_iter += iters
        self.last_loss = loss.view(iters, 2)   # (bce, eikonal) per iteration, on the device (data-parallel: this rank's part)
        self._fused_last = {"moments": moments, "tensors": tensors}  # read by the parity tests

    def mapping_loss(self, coord, sdf_label, ts, weight):
        """One iteration's loss (reference: utils/mapper.py:461-547): BCE on the fused SDF prediction plus the
        eikonal term on numerical gradients of every `gradient_decimation`-th sample."""
        cfg = self.config
        sdf_pred = self.neural_points.query_sdf(coord, self.geo_mlp, ts, True, True)["sdf"]
        weight = torch.abs(weight).detach()
        if cfg.main_loss_type != "bce":
            raise NotImplementedError("only the bce main loss is in scope (reference: mapper.py:515-524 '[used]')")
        cur_loss = sdf_bce_loss(sdf_pred, sdf_label, self.sdf_scale, weight, cfg.loss_weight_on)
        if cfg.ekional_loss_on and cfg.weight_e > 0:
            if not cfg.numerical_grad:
                raise NotImplementedError("analytic eikonal gradient in mapping is unused by the reference configs")
            g = self.get_numerical_gradient(coord[::cfg.gradient_decimation], sdf_pred[::cfg.gradient_decimation],
                                            cfg.voxel_size_m * cfg.num_grad_step_ratio)
            if cfg.ekional_add_to != "all":
                surface_mask = (torch.abs(sdf_label) < cfg.surface_sample_range_m)[::cfg.gradient_decimation]
                g = g[surface_mask] if cfg.ekional_add_to == "surface" else g[~surface_mask]
            eikonal_loss = ((g.norm(2, dim=-1) - 1.0) ** 2).mean()
            cur_loss = cur_loss + cfg.weight_e * eikonal_loss
        return cur_loss, sdf_pred

    def sdf(self, x, get_std=False):
        """reference: utils/mapper.py:671-681 (default training_mode=True, no timestamps: hazard H9)."""
        out = self.neural_points.query_sdf(x, self.geo_mlp, None, True, True, want_std=get_std)
        return out["sdf"], out["std"]

    def get_numerical_gradient(self, x, sdf_x=None, eps=0.02, two_side=True):
        """reference: utils/mapper.py:683-733."""
        N = x.shape[0]
        e = torch.eye(3, dtype=x.dtype, device=x.device) * eps
        if two_side:
            x_posneg = torch.cat((x + e[0], x - e[0], x + e[1], x - e[1], x + e[2], x - e[2]), dim=0)
            s = self.sdf(x_posneg)[0].unsqueeze(-1)
            gx = (s[:N] - s[N:2 * N]) / (2 * eps)
            gy = (s[2 * N:3 * N] - s[3 * N:4 * N]) / (2 * eps)
            gz = (s[4 * N:5 * N] - s[5 * N:]) / (2 * eps)
        else:
            x_all = torch.cat((x + e[0], x + e[1], x + e[2]), dim=0)
            s = self.sdf(x_all)[0].unsqueeze(-1)
            sx = sdf_x.unsqueeze(-1)
            gx, gy, gz = (s[:N] - sx) / eps, (s[N:2 * N] - sx) / eps, (s[2 * N:] - sx) / eps
        return torch.cat([gx, gy, gz], dim=1)

    def bundle_adjustment(self, iter_count, window_size: int = 50, use_lie_group: bool = False):
        raise NotImplementedError("bundle adjustment needs pypose and is off by default (reference: config.py:236)")
