#!/usr/bin/env python3
"""Debug: frame-0 state of the CUDA loop vs the oracle loop (same scans, same draws)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from pin_slam_b200 import synth
from pin_slam_b200.slam import run_slam
from pin_slam_b200.utils.config import kitti_config
from pin_slam_b200.utils.data_sampler import DataSampler
from oracle.slam_oracle import OracleSLAM

ratio = int(sys.argv[1]) if len(sys.argv) > 1 else 1
nfr = int(sys.argv[2]) if len(sys.argv) > 2 else 1
def cfg_for(dev):
    c = kitti_config(dev); c.deskew = False; c.correction_deg = 0.0; c.silence = True; c.init_iter_ratio = ratio
    return c
scene = synth.StreetScene(42); gt = synth.trajectory(nfr + 1); dirs = synth.lidar_directions(64, 1875)
rng = np.random.default_rng(0)
frames = [synth.lidar_scan(scene, gt[i], dirs, rng) for i in range(nfr)]
res = run_slam(cfg_for("cuda"), frames, None, generator=torch.Generator().manual_seed(7))
cc = cfg_for("cpu")
orc = OracleSLAM(cc, DataSampler(cc), torch.Generator().manual_seed(7), seed=cc.seed)
for i in range(nfr):
    orc.step(frames[i], i)
m, mp, s = res["neural_points"], res["mapper"], orc.s
def rel(a, b):
    a = a.detach().double().cpu(); b = b.detach().double().cpu()
    return float((a - b).norm() / max(float(b.norm()), 1e-12))
print("points", m.count(), s.points.shape[0], "equal", m.count() == s.points.shape[0] and torch.equal(m.neural_points.cpu(), s.points))
print("pool", mp.pool_sample_count, orc.pool_sample_count, "coord rel", rel(mp.global_coord_pool, orc.global_pool) if mp.pool_sample_count == orc.pool_sample_count else None)
print("labels rel", rel(mp.sdf_label_pool, orc.label_pool), "weights rel", rel(mp.weight_pool, orc.weight_pool))
print("new_idx", mp.new_idx.shape[0], orc.new_idx.shape[0], torch.equal(mp.new_idx.cpu(), orc.new_idx))
print("feat rel", rel(m.geo_features, s.feats), "max abs", float((m.geo_features.cpu() - s.feats).abs().max()), "feat norm", float(s.feats.norm()))
dec = res["decoder"]
print("W1 rel", rel(dec.layers[0].weight, orc.d.W1), "W2 rel", rel(dec.lout.weight, orc.d.W2), "b1", rel(dec.layers[0].bias, orc.d.b1))
print("cert rel", rel(m.point_certainties, s.certainty), "ts_update eq", torch.equal(m.point_ts_update.cpu(), s.ts_update))
print("poses", [float(np.linalg.norm(res['poses'][i][:3,3]-np.array(orc.odom_poses[i])[:3,3])) for i in range(nfr)])
