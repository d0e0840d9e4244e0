#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference, imported through oracle/ref_shim.py) on CPU.  Run in the build container:

    python oracle/make_golden.py

The reference publishes no golden vectors; these files pin the oracle
(oracle/pin_oracle.py) -- and through it the CUDA path -- to the reference's own outputs on the GPU box,
where /root/reference does not exist.
"""
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
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, "ref_kitti_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB;", npts.count(), "neural points,", npts.local_count(), "local")


if __name__ == "__main__":
    main()
