#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference, imported through oracle/ref_shim.py) on CPU.  Run in the build container:

    python oracle/make_golden.py

The reference publishes no golden vectors; these files pin the oracle
(oracle/pin_oracle.py) -- and through it the CUDA path -- to the reference's own outputs on the GPU box,
where /root/reference does not exist.
"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from pin_slam_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def main():
    # One thread: ATen's CPU index_put_ is sequential only then, and `buffer_pt_index[hash] = idx` with duplicate
    # slots (model/neural_points.py:247) is a race otherwise -- the reference's result is run-to-run unstable
    # on a multi-threaded CPU (hazard H4).  The single-thread outcome (last writer wins) is the pinned one.
    torch.set_num_threads(1)
    ns = ref_shim.load()
    cfg = ref_shim.kitti_config("cpu")
    cfg.buffer_size = 2_000_003  # small table so the fixture can store it densely; same hash arithmetic
    torch.manual_seed(42)
    dec = ns.Decoder(cfg, 64, 1, 1)
    npts = ns.NeuralPoints(cfg)
    npts.travel_dist = torch.tensor([0.0, 3.0, 500.0])
    frames = []
    for ts in range(3):
        p = synth.surface_points(9000, 40.0, seed=100 + ts)
        p[:, 0] += 3.0 * ts
        frames.append(p)
        npts.update(torch.from_numpy(p), torch.tensor([3.0 * ts, 0.0, 0.0]), torch.eye(3), ts)
    # frame 2 is 500 m of travel away from frames 0/1 -> exercises the travel-distance window
    g = torch.Generator().manual_seed(1)
    feats = torch.randn(npts.geo_features.shape, generator=g) * 0.2
    feats[-1] = 0
    cert = torch.rand(npts.point_certainties.shape, generator=g) * 30
    npts.geo_features = feats.clone()
    npts.point_certainties = cert.clone()
    npts.reset_local_map(torch.tensor([6.0, 0.0, 0.0]), torch.eye(3), 2)
    q = torch.from_numpy(synth.microbench_queries(npts.neural_points.numpy(), 400, seed=3))

    out = {"frames0": frames[0], "frames1": frames[1], "frames2": frames[2], "queries": q.numpy(),
           "travel_dist": npts.travel_dist.numpy(), "buffer_size": np.int64(cfg.buffer_size),
           "features": feats.numpy(), "certainties": cert.numpy(),
           "W1": dec.layers[0].weight.detach().numpy(), "b1": dec.layers[0].bias.detach().numpy(),
           "W2": dec.lout.weight.detach().numpy(), "b2": dec.lout.bias.detach().numpy(),
           "sdf_scale": np.float64(dec.sdf_scale)}
    # ---- map state
    out["table"] = npts.buffer_pt_index.numpy().astype(np.int32)
    out["neural_points"] = npts.neural_points.numpy()
    out["ts_create"] = npts.point_ts_create.numpy()
    out["global2local"] = npts.global2local.numpy().astype(np.int32)
    out["local_mask"] = npts.local_mask.numpy()
    # ---- voxel down-sampling
    for name, vs in (("vds_04", 0.4), ("vds_008", 0.08)):
        out[name] = ns.tools.voxel_down_sample_torch(torch.from_numpy(frames[0]), vs).numpy().astype(np.int32)
    val = torch.rand(frames[1].shape[0], generator=g) * 20
    out["vds_min_value_in"] = val.numpy()
    out["vds_min_value"] = ns.tools.voxel_down_sample_min_value_torch(torch.from_numpy(frames[1]), 0.4, val).numpy().astype(np.int32)
    # ---- neighbour search
    for tf in (False, True):
        d2, idx = npts.radius_neighborhood_search(q, tf)
        out["rs_d2_%d" % tf] = d2.numpy()
        out["rs_idx_%d" % tf] = idx.numpy().astype(np.int32)
    npts.set_search_neighborhood(num_nei_cells=1, search_alpha=0.0)
    out["query_certainty"] = npts.query_certainty(q).numpy()
    npts.set_search_neighborhood(num_nei_cells=cfg.num_nei_cells, search_alpha=cfg.search_alpha)
    # ---- query_feature + decoder + gradient (tracking flavour, inference mode), local and global
    for local in (True, False):
        qq = q.clone().requires_grad_(True)
        fv, _, w, nn, qc = npts.query_feature(qq, training_mode=False, query_locally=local)
        pred = dec.sdf(fv)
        mean = torch.sum(pred * w, dim=1)
        std = torch.sqrt(torch.sum(w * (pred - mean.unsqueeze(-1)) ** 2, dim=1)).squeeze(1)
        sdf = mean.squeeze(1)
        grad = ns.tools.get_gradient(qq, sdf)
        tag = "L" if local else "G"
        out["fv_" + tag] = fv.detach().numpy()
        out["w_" + tag] = w.detach().numpy()
        out["nn_" + tag] = nn.numpy().astype(np.int32)
        out["qc_" + tag] = qc.numpy()
        out["sdf_" + tag] = sdf.detach().numpy()
        out["std_" + tag] = std.detach().numpy()
        out["grad_" + tag] = grad.detach().numpy()
    # ---- Gauss-Newton system (utils/tracker.py:468-520)
    pts = q[:300].clone()
    gg = torch.randn(300, 3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, 1.0])
    res = torch.randn(300, generator=g) * 0.05
    wt = torch.rand(300, 1, generator=g) + 0.1
    T, _, _ = ns.tracker_mod.implicit_reg(pts, gg, res, wt, lm_lambda=1e-4)
    out["gn_points"], out["gn_grad"], out["gn_res"], out["gn_w"], out["gn_T"] = pts.numpy(), gg.numpy(), res.numpy(), wt.numpy(), T.numpy()
    # ---- one registration step of the reference Tracker
    trk = ns.Tracker(cfg, npts, dec, None, None)
    src = (npts.local_neural_points[::3][:600] + 0.02 * torch.randn(600, 3, generator=g)).contiguous()
    Tr, _, _, _, valid_pts, res_cm, _ = trk.registration_step(src.clone(), None, torch.zeros(src.shape[0]), None, None,
                                                             0.0, 1e3, cfg.reg_GM_dist_m, cfg.reg_GM_grad, cfg.reg_lm_lambda)
    out["reg_src"], out["reg_T"] = src.numpy(), Tr.numpy()
    out["reg_valid"], out["reg_res_cm"] = np.int64(valid_pts.shape[0]), np.float64(res_cm)
    # ---- the whole Gauss-Newton loop, Tracker.tracking (utils/tracker.py:39-174): 3 iterations from a perturbed pose with
    # relaxed validity gates (random features: rough field), then once with run_kitti's own gates
    init = torch.eye(4, dtype=torch.float64)
    init[:3, 3] = torch.tensor([0.02, -0.01, 0.005], dtype=torch.float64)
    keep = (cfg.reg_iter_n, cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio)
    cfg.reg_iter_n, cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio = 3, 0.0, 1e3, 1e3
    T1, _, _, v1 = trk.tracking(src.clone(), init.clone())
    cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio = keep[1:]
    T2, _, _, v2 = trk.tracking(src.clone(), init.clone())
    cfg.reg_iter_n = keep[0]
    out["trk_init"] = init.numpy()
    out["trk_T_relaxed"], out["trk_valid_relaxed"] = T1.numpy(), np.int64(bool(v1))
    out["trk_T_default"], out["trk_valid_default"] = T2.numpy(), np.int64(bool(v2))
    print("tracking valid (relaxed, default):", bool(v1), bool(v2))
    # ---- mapping loss + gradients (utils/mapper.py:461-573), training-mode side effects included
    mp = ns.Mapper(cfg, None, npts, dec, None, None)
    n = 1024
    coord = torch.from_numpy(synth.microbench_queries(npts.neural_points.numpy(), n, seed=4))
    label = 0.2 * torch.randn(n, generator=g)
    weight = torch.rand(n, generator=g) + 0.5
    weight[::3] *= -1
    ts = torch.randint(0, 3, (n,), generator=g)
    cert_before = npts.local_point_certainties.clone()
    fv, _, w, _, _ = npts.query_feature(coord, ts)
    pred = torch.sum(dec.sdf(fv) * w, dim=1).squeeze(1)
    gnum = mp.get_numerical_gradient(coord[::cfg.gradient_decimation], pred[::cfg.gradient_decimation],
                                     cfg.voxel_size_m * cfg.num_grad_step_ratio)
    loss = ns.loss.sdf_bce_loss(pred, label, mp.sdf_scale, torch.abs(weight).detach(), cfg.loss_weight_on)
    loss = loss + cfg.weight_e * ((gnum.norm(2, dim=-1) - 1.0) ** 2).mean()
    for p in list(dec.parameters()) + [npts.local_geo_features]:
        p.grad = None
    loss.backward()
    out["map_coord"], out["map_label"], out["map_weight"], out["map_ts"] = coord.numpy(), label.numpy(), weight.numpy(), ts.numpy()
    out["map_loss"], out["map_pred"] = np.float64(loss.item()), pred.detach().numpy()
    out["map_gfeat"] = npts.local_geo_features.grad.numpy()
    out["map_gW1"], out["map_gb1"] = dec.layers[0].weight.grad.numpy(), dec.layers[0].bias.grad.numpy()
    out["map_gW2"], out["map_gb2"] = dec.lout.weight.grad.numpy(), dec.lout.bias.grad.numpy()
    out["map_cert_before"], out["map_cert_after"] = cert_before.numpy(), npts.local_point_certainties.numpy()
    out["map_ts_update_after"] = npts.local_point_ts_update.numpy()
    # ---- hazard H13: candidates inside the travel-distance window but OUTSIDE the local-map radius are
    # remapped to local index 1 by the reference (full_like on a bool mask).  Shrink the radius so that the
    # query set straddles the boundary.
    npts.local_map_radius = 9.0
    npts.reset_local_map(torch.tensor([6.0, 0.0, 0.0]), torch.eye(3), 2)
    qq = q.clone().requires_grad_(True)
    fv, _, w, nn, qc = npts.query_feature(qq, training_mode=False, query_locally=True)
    sdf = torch.sum(dec.sdf(fv) * w, dim=1).squeeze(1)
    grad = ns.tools.get_gradient(qq, sdf)
    out["h13_radius"] = np.float64(9.0)
    out["h13_fv"], out["h13_w"], out["h13_nn"] = fv.detach().numpy(), w.detach().numpy(), nn.numpy().astype(np.int32)
    out["h13_sdf"], out["h13_grad"], out["h13_qc"] = sdf.detach().numpy(), grad.detach().numpy(), qc.numpy()
    out["h13_local_count"] = np.int64(npts.local_count())
    # ---- Mesher.get_query_from_bbx + query_points (utils/mesher.py:138-180, 41-136): dense grid over a sub-volume,
    # global query (no time filter), mask = nn_count >= mesh_min_nn
    import types
    mesher = importlib.import_module("utils.mesher").Mesher(cfg, npts, dec, None, None)
    lo, hi = np.array([2.0, -4.0, -2.0]), np.array([9.1, 3.3, 1.05])
    bbx = types.SimpleNamespace(get_min_bound=lambda: lo.copy(), get_max_bound=lambda: hi.copy())
    coord, num_xyz, origin = mesher.get_query_from_bbx(bbx, 0.2, 1, 0)
    sdf_pred, _, _, mc_mask = mesher.query_points(coord, 4096, True, False, False, True, False, 8)
    out["mesh_lo"], out["mesh_hi"], out["mesh_voxel"], out["mesh_min_nn"] = lo, hi, np.float64(0.2), np.int64(8)
    out["mesh_num_xyz"], out["mesh_origin"] = np.asarray(num_xyz, dtype=np.int64), np.asarray(origin, dtype=np.float64)
    out["mesh_coord_first_last"] = np.stack((coord[0].numpy(), coord[-1].numpy()))
    out["mesh_sdf"], out["mesh_mask"] = np.asarray(sdf_pred, dtype=np.float32), np.asarray(mc_mask).astype(np.uint8)
    print("mesher grid", num_xyz, "inside fraction", float(np.mean(out["mesh_mask"])))
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, "ref_kitti_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB;", npts.count(), "neural points,", npts.local_count(), "local")


def pose_corrections(n_frames=3):
    """Per-frame SE(3) corrections as a pose-graph optimisation would hand to adjust_map (small rotations + shifts)."""
    T = torch.eye(4).repeat(n_frames, 1, 1)
    for i, (az, ax, t) in enumerate(((0.02, -0.01, (0.05, -0.02, 0.01)), (-0.03, 0.015, (-0.04, 0.03, 0.0)),
                                     (0.05, 0.02, (0.02, 0.06, -0.02)))[:n_frames]):
        cz, sz, cx, sx = np.cos(az), np.sin(az), np.cos(ax), np.sin(ax)
        Rz = torch.tensor([[cz, -sz, 0.0], [sz, cz, 0.0], [0.0, 0.0, 1.0]])
        Rx = torch.tensor([[1.0, 0.0, 0.0], [0.0, cx, -sx], [0.0, sx, cx]])
        T[i, :3, :3] = (Rz @ Rx).float()
        T[i, :3, 3] = torch.tensor(t)
    return T


def main_pgo():
    """Second fixture: the map-maintenance calls around a loop closure and the after_pgo branch of the
    query (model/neural_points.py:597-607) -- adjust_map, recreate_hash (both flavours), prune_map, query_feature + decoder
    + gradients with rotated neighbour offsets."""
    torch.set_num_threads(1)
    ns = ref_shim.load()
    cfg = ref_shim.kitti_config("cpu")
    cfg.buffer_size = 2_000_003
    torch.manual_seed(42)
    dec = ns.Decoder(cfg, 64, 1, 1)
    npts = ns.NeuralPoints(cfg)
    npts.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    frames = []
    for ts in range(3):
        p = synth.surface_points(9000, 40.0, seed=200 + ts)
        p[:, 0] += 3.0 * ts
        frames.append(p)
        npts.update(torch.from_numpy(p), torch.tensor([3.0 * ts, 0.0, 0.0]), torch.eye(3), ts)
    g = torch.Generator().manual_seed(2)
    feats = torch.randn(npts.geo_features.shape, generator=g) * 0.2
    feats[-1] = 0
    cert = torch.rand(npts.point_certainties.shape, generator=g) * 30
    npts.geo_features, npts.point_certainties = feats.clone(), cert.clone()
    out = {"frames0": frames[0], "frames1": frames[1], "frames2": frames[2], "buffer_size": np.int64(cfg.buffer_size),
           "features": feats.numpy(), "certainties": cert.numpy(),
           "W1": dec.layers[0].weight.detach().numpy(), "b1": dec.layers[0].bias.detach().numpy(),
           "W2": dec.lout.weight.detach().numpy(), "b2": dec.lout.bias.detach().numpy(), "sdf_scale": np.float64(dec.sdf_scale)}
    # ---- adjust_map (model/neural_points.py:355-370)
    pose_diff = pose_corrections(3)
    out["pose_diff"] = pose_diff.numpy()
    npts.adjust_map(pose_diff)
    out["adj_points"], out["adj_quats"] = npts.neural_points.numpy().copy(), npts.point_orientations.numpy().copy()
    # ---- recreate_hash, kept_points flavour, then the local map (model/neural_points.py:372-428)
    sensor = torch.tensor([6.0, 0.0, 0.0])
    npts.recreate_hash(sensor, torch.eye(3), True, True, 2)
    out["rh_table"] = npts.buffer_pt_index.numpy().astype(np.int32)
    out["rh_global2local"] = npts.global2local.numpy().astype(np.int32)
    # ---- after_pgo query: tracking flavour (inference) with gradient, then a training-flavour backward
    q = torch.from_numpy(synth.microbench_queries(npts.neural_points.numpy(), 400, seed=5))
    out["queries"] = q.numpy()
    qq = q.clone().requires_grad_(True)
    fv, _, w, nn, qc = npts.query_feature(qq, training_mode=False, query_locally=True)
    pred = dec.sdf(fv)
    sdf = torch.sum(pred * w, dim=1).squeeze(1)
    grad = ns.tools.get_gradient(qq, sdf)
    out["fv"], out["w"], out["nn"] = fv.detach().numpy(), w.detach().numpy(), nn.numpy().astype(np.int32)
    out["sdf"], out["grad"], out["qc"] = sdf.detach().numpy(), grad.detach().numpy(), qc.numpy()
    up = torch.randn(q.shape[0], generator=g)
    ts = torch.randint(0, 3, (q.shape[0],), generator=g)
    for p_ in list(dec.parameters()) + [npts.local_geo_features]:
        p_.grad = None
    fv, _, w, _, _ = npts.query_feature(q, ts)
    (torch.sum(dec.sdf(fv) * w, dim=1).squeeze(1) * up).sum().backward()
    out["bwd_up"], out["bwd_ts"] = up.numpy(), ts.numpy()
    out["bwd_gfeat"] = npts.local_geo_features.grad.numpy()
    out["bwd_gW1"], out["bwd_gb1"] = dec.layers[0].weight.grad.numpy(), dec.layers[0].bias.grad.numpy()
    out["bwd_gW2"], out["bwd_gb2"] = dec.lout.weight.grad.numpy(), dec.lout.bias.grad.numpy()
    npts.assign_local_to_global()
    out["after_query_certainties"] = npts.point_certainties.numpy().copy()
    out["after_query_ts_update"] = npts.point_ts_update.numpy().copy()
    # ---- prune_map (model/neural_points.py:328-353): frames 0/1 become inactive 500 m later
    npts.travel_dist = torch.tensor([0.0, 3.0, 500.0])
    npts.cur_ts = 2
    out["prune_travel_dist"] = npts.travel_dist.numpy()
    out["prune_thre"] = np.float64(15.0)
    pruned = npts.prune_map(15.0)
    out["pruned"] = np.int64(pruned)
    out["prune_points"], out["prune_features"] = npts.neural_points.numpy().copy(), npts.geo_features.numpy().copy()
    out["prune_ts_create"], out["prune_certainties"] = npts.point_ts_create.numpy().copy(), npts.point_certainties.numpy().copy()
    # ---- recreate_hash, compacting flavour keyed on certainty (kept_points=False, with_ts=False)
    npts.recreate_hash(None, None, False, False, 2)
    out["rh2_table"] = npts.buffer_pt_index.numpy().astype(np.int32)
    out["rh2_points"], out["rh2_features"] = npts.neural_points.numpy().copy(), npts.geo_features.numpy().copy()
    path = os.path.join(OUT, "ref_kitti_pgo.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB;", npts.count(), "neural points after prune + compaction")


def main_frames():
    """Third fixture: the per-frame data path of the mapper -- DataSampler.sample on a seeded RNG and three
    Mapper.process_frame calls (sampling, map update, replay-pool append + window filter, new-sample certainty query) of the
    unmodified reference.  The sampler's per-frame outputs are recorded by wrapping the bound method (instrumentation only)."""
    import types
    torch.set_num_threads(1)
    ns = ref_shim.load()
    cfg = ref_shim.kitti_config("cpu")
    cfg.buffer_size = 2_000_003
    cfg.window_radius = 14.0     # small enough that the moving sensor drops old samples from the pool
    cfg.silence = True
    cfg.pgo_on = False           # no pose graph in this repo's harness: poses = odometry
    out = {"buffer_size": np.int64(cfg.buffer_size), "window_radius": np.float64(cfg.window_radius)}
    # ---- DataSampler.sample (utils/data_sampler.py:20-192) on torch.manual_seed(5)
    sampler = importlib.import_module("utils.data_sampler").DataSampler(cfg)
    pts = torch.from_numpy(synth.surface_points(300, 40.0, seed=300))
    torch.manual_seed(5)
    coord, label, _, _, _, weight = sampler.sample(pts, None, None, None)
    out["s_points"], out["s_coord"], out["s_label"], out["s_weight"] = pts.numpy(), coord.numpy(), label.numpy(), weight.numpy()
    # ---- three frames through Mapper.process_frame (utils/mapper.py:110-321)
    torch.manual_seed(42)
    dec = ns.Decoder(cfg, 64, 1, 1)
    npts = ns.NeuralPoints(cfg)
    npts.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    dataset = types.SimpleNamespace(odom_poses=[np.eye(4)], gt_pose_provided=False, stop_status=False)
    mp = ns.Mapper(cfg, dataset, npts, dec, None, None)
    rec = []
    inner = mp.sampler.sample

    def recording_sample(*a, **k):
        r = inner(*a, **k)
        rec.append((r[0].clone(), r[1].clone(), r[5].clone()))
        return r
    mp.sampler.sample = recording_sample
    for f in range(3):
        p = synth.surface_points(2500, 40.0, seed=310 + f)
        pose = torch.eye(4, dtype=torch.float64)
        pose[0, 3] = 3.0 * f
        pose[1, 3] = 0.5 * f
        local = (torch.from_numpy(p).double() - pose[:3, 3]).float()   # sensor-frame scan (identity rotation)
        torch.manual_seed(100 + f)
        mp.process_frame(local, None, pose, f, False)
        dataset.odom_poses.append(pose.numpy())
        out["f%d_scan" % f], out["f%d_pose" % f] = local.numpy(), pose.numpy()
        out["f%d_coord" % f], out["f%d_label" % f], out["f%d_weight" % f] = (t.numpy() for t in rec[f])
        out["f%d_pool_coord" % f] = mp.coord_pool.numpy().copy()
        out["f%d_pool_global" % f] = mp.global_coord_pool.numpy().copy()
        out["f%d_pool_label" % f] = mp.sdf_label_pool.numpy().copy()
        out["f%d_pool_weight" % f] = mp.weight_pool.numpy().copy()
        out["f%d_pool_ts" % f] = mp.time_pool.numpy().copy()
        out["f%d_counts" % f] = np.array([mp.cur_sample_count, mp.pool_sample_count], dtype=np.int64)
        out["f%d_new_idx" % f] = mp.new_idx.numpy().copy()
        out["f%d_map_points" % f] = npts.neural_points.numpy().copy()
    # ---- Mapper.mapping (utils/mapper.py:425-593): four iterations incl. get_batch's draws and the Adam steps
    cfg.bs, cfg.bs_new_sample = 2048, 256
    mp.train_less = False
    out["map_bs"], out["map_bs_new"], out["map_iters"], out["map_seed"] = np.int64(2048), np.int64(256), np.int64(4), np.int64(77)
    out["map_features_before"] = npts.geo_features.detach().numpy().copy()
    torch.manual_seed(77)
    mp.mapping(4)
    out["map_features_after"] = npts.geo_features.detach().numpy().copy()
    out["map_certainties_after"] = npts.point_certainties.numpy().copy()
    out["map_ts_update_after"] = npts.point_ts_update.numpy().copy()
    for k, t in zip(("W1", "b1", "W2", "b2"), (dec.layers[0].weight, dec.layers[0].bias, dec.lout.weight, dec.lout.bias)):
        out["map_%s_after" % k] = t.detach().numpy().copy()
    path = os.path.join(OUT, "ref_kitti_frames.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB; pool", mp.pool_sample_count, "of", sum(r[0].shape[0] for r in rec), "samples kept")


def main_preprocess():
    """Fourth fixture: the frame data path in FRONT of the hot path of the unmodified reference --
    `deskewing` (utils/tools.py:540-567; roma.rotmat_slerp restated in oracle/ref_shim.py), three frames through
    SLAMDataset.read_frame / preprocess_frame / update_odom_pose (dataset/slam_dataset.py:199-430: Velodyne per-point
    times, voxel down-sampling, crop, KITTI intrinsic correction, constant-velocity guess, source down-sampling + deskew,
    cloud deskew after the pose update) on a KITTI-layout folder of motion-distorted synthetic sweeps, and
    Mapper.dynamic_filter (utils/mapper.py:79-108, both strategies) on the first fixture's map."""
    import shutil
    import tempfile
    import types
    torch.set_num_threads(1)
    ns = ref_shim.load()
    out = {}
    g = torch.Generator().manual_seed(11)
    # ---- deskewing alone
    pts = (torch.rand(4000, 3, generator=g) - 0.5) * torch.tensor([120.0, 120.0, 8.0])
    ts = torch.from_numpy(synth.velodyne_point_ts(pts.numpy().astype(np.float64)).astype(np.float32)).reshape(-1, 1)
    yaw, pitch = 0.021, -0.004
    cy, sy, cp, sp = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
    T = np.eye(4)
    T[:3, :3] = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]]) @ np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    T[:3, 3] = [0.91, 0.03, -0.01]
    Tt = torch.tensor(T, dtype=torch.float32)
    out["dk_points"], out["dk_ts"], out["dk_pose"] = pts.numpy().copy(), ts.numpy().copy(), Tt.numpy().copy()
    out["dk_out"] = ns.tools.deskewing(pts.clone(), ts.clone(), Tt).numpy().copy()
    out["dk_out_mid0"] = ns.tools.deskewing(pts.clone(), ts.clone(), Tt, ts_mid_pose=0.0).numpy().copy()
    # ---- three frames through the reference's SLAMDataset
    frames, gt = synth.make_sweeps(3, beams=24, azimuth=600, workers=1, correct_deg=0.195)
    root = tempfile.mkdtemp(prefix="pinq_kitti_")
    try:
        paths = synth.write_kitti_sequence(root, frames, gt)
        cfg = ns.Config()
        cfg.load(ref_shim.write_kitti_yaml(os.path.join(root, "run_kitti.yaml"), paths, os.path.join(root, "out")))
        cfg.device = "cpu"
        SLAMDataset = importlib.import_module("dataset.slam_dataset").SLAMDataset
        ds = SLAMDataset(cfg)
        out["pp_n"] = np.int64(3)
        out["pp_gt"] = np.stack(ds.poses_ref[:3]) if ds.gt_pose_provided else gt
        for i in range(3):
            out["pp%d_scan" % i] = frames[i]
            ds.read_frame(i)
            out["pp%d_ts_raw" % i] = ds.cur_point_ts_torch.numpy().copy()
            ds.preprocess_frame(i)
            out["pp%d_cloud" % i] = ds.cur_point_cloud_torch.numpy().copy()
            out["pp%d_ts" % i] = ds.cur_point_ts_torch.numpy().copy()
            if i > 0:
                out["pp%d_guess" % i] = ds.cur_pose_guess_torch.numpy().copy()
                out["pp%d_source" % i] = ds.cur_source_points.numpy().copy()
                # the tracker's answer is replaced by a fixed pose near the ground truth (the fixture pins the data path)
                est = ds.poses_ref[i].copy()
                est[:3, 3] += np.array([0.01, -0.005, 0.002]) * i
                out["pp%d_est" % i] = est
                ds.update_odom_pose(torch.tensor(est, dtype=torch.float64))
                out["pp%d_cloud_deskewed" % i] = ds.cur_point_cloud_torch.numpy().copy()
                out["pp%d_last_odom_tran" % i] = ds.last_odom_tran.copy()
            ds.processed_frame += 1
        out["pp_travel_dist"] = np.array(ds.travel_dist)
    finally:
        shutil.rmtree(root, ignore_errors=True)
    # ---- Mapper.dynamic_filter on the map of the first fixture
    G = {k: v for k, v in np.load(os.path.join(OUT, "ref_kitti_small.npz")).items()}
    cfg = ref_shim.kitti_config("cpu")
    cfg.buffer_size = int(G["buffer_size"])
    torch.manual_seed(42)
    dec = ns.Decoder(cfg, 64, 1, 1)
    with torch.no_grad():
        dec.layers[0].weight.copy_(torch.from_numpy(G["W1"]))
        dec.layers[0].bias.copy_(torch.from_numpy(G["b1"]))
        dec.lout.weight.copy_(torch.from_numpy(G["W2"]))
        dec.lout.bias.copy_(torch.from_numpy(G["b2"]))
    npts = ns.NeuralPoints(cfg)
    npts.travel_dist = torch.from_numpy(G["travel_dist"])
    for t in range(3):
        npts.update(torch.from_numpy(G["frames%d" % t]), torch.tensor([3.0 * t, 0.0, 0.0]), torch.eye(3), t)
    npts.geo_features = torch.from_numpy(G["features"]).clone()
    npts.point_certainties = torch.from_numpy(G["certainties"]).clone()
    npts.reset_local_map(torch.tensor([6.0, 0.0, 0.0]), torch.eye(3), 2)
    dataset = types.SimpleNamespace(odom_poses=[np.eye(4)], gt_pose_provided=False, stop_status=False)
    mp = ns.Mapper(cfg, dataset, npts, dec, None, None)
    q = torch.from_numpy(G["queries"]).clone()
    out["df_static_1"] = mp.dynamic_filter(q.clone(), False).numpy().copy()
    out["df_static_2"] = mp.dynamic_filter(q.clone(), True).detach().numpy().copy()
    out["df_thresholds"] = np.array([cfg.dynamic_certainty_thre, cfg.dynamic_sdf_ratio_thre], dtype=np.float64)
    path = os.path.join(OUT, "ref_kitti_preprocess.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB;", "static fraction", float(out["df_static_1"].mean()), float(out["df_static_2"].mean()))


if __name__ == "__main__":
    only = [a for a in sys.argv[1:] if a.startswith("--only=")]
    which = only[0].split("=")[1].split(",") if only else ["small", "pgo", "frames", "preprocess"]
    if "preprocess" in which:
        main_preprocess()
    if "small" in which:
        main()
    if "pgo" in which:
        main_pgo()
    if "frames" in which:
        main_frames()
