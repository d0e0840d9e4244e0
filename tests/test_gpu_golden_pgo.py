"""CUDA path vs golden vectors of the UNMODIFIED reference for the calls around a loop closure:
NeuralPoints.adjust_map / recreate_hash / prune_map and the after_pgo branch of the fused query (x - p_k rotated into
each neural point's frame, model/neural_points.py:597-607) -- tests/golden/ref_kitti_pgo.npz, oracle/make_golden.py."""
import numpy as np
import pytest
import torch

from helpers import load_golden_pgo, rel_err

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


def test_after_pgo_sequence_matches_the_reference():
    from pin_slam_b200.model.decoder import Decoder
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config
    G = load_golden_pgo()
    cfg = kitti_config("cuda")
    cfg.buffer_size = int(G["buffer_size"])
    m = NeuralPoints(cfg)
    m.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    eye = torch.eye(3, device="cuda")
    for ts in range(3):
        m.update(T(G["frames%d" % ts]).cuda(), torch.tensor([3.0 * ts, 0.0, 0.0]).cuda(), eye, ts)
    m.geo_features, m.point_certainties = T(G["features"]).cuda(), T(G["certainties"]).cuda()
    dec = Decoder(cfg, 64, 1, 1)
    with torch.no_grad():
        for p, k in zip(dec.fused_params(), ("W1", "b1", "W2", "b2")):
            p.copy_(T(G[k]))
    # adjust_map: batched 3x3 products (cuBLAS on the device, ATen on the reference's CPU) -> tolerance, not bits
    m.adjust_map(T(G["pose_diff"]).cuda())
    assert m.after_pgo
    assert float((m.neural_points.cpu() - T(G["adj_points"])).abs().max()) < 1e-5
    assert float((m.point_orientations.cpu() - T(G["adj_quats"])).abs().max()) < 1e-6
    m.neural_points, m.point_orientations = T(G["adj_points"]).cuda(), T(G["adj_quats"]).cuda()  # identical inputs from here on
    # recreate_hash (kept points) + local map: bit-exact
    m.recreate_hash(torch.tensor([6.0, 0.0, 0.0]).cuda(), eye, True, True, 2)
    assert torch.equal(m.buffer_pt_index.cpu().int(), T(G["rh_table"]))
    assert torch.equal(m.global2local.cpu().int(), T(G["rh_global2local"]))
    # after_pgo query through the un-fused API and the fused kernels
    q = T(G["queries"]).cuda()
    fv, _, w, nn, qc = m.query_feature(q, training_mode=False, query_locally=True)
    assert torch.equal(nn.cpu().int(), T(G["nn"]))
    assert rel_err(fv, T(G["fv"])) < 1e-6 and rel_err(w, T(G["w"])) < 1e-6 and rel_err(qc, T(G["qc"])) < 1e-6
    out = m.query_sdf(q, dec, None, False, True, want_grad=True)
    assert torch.equal(out["nn_count"].cpu().int(), T(G["nn"]))
    assert rel_err(out["sdf"], T(G["sdf"])) < 1e-5
    assert rel_err(out["grad"], T(G["grad"])) < 1e-5
    # training-flavour backward with rotated offsets
    for p in list(dec.parameters()) + [m.local_geo_features]:
        p.grad = None
    o2 = m.query_sdf(q, dec, T(G["bwd_ts"]).cuda(), True, True)
    (o2["sdf"] * T(G["bwd_up"]).cuda()).sum().backward()
    assert rel_err(m.local_geo_features.grad, T(G["bwd_gfeat"])) < 1e-5
    for name, t in zip(("bwd_gW1", "bwd_gb1", "bwd_gW2", "bwd_gb2"), dec.fused_params()):
        assert rel_err(t.grad, T(G[name])) < 1e-5, name
    m.assign_local_to_global()
    # (the un-fused inference query above has no side effects; the training query has)
    assert rel_err(m.point_certainties, T(G["after_query_certainties"])) < 1e-5
    assert torch.equal(m.point_ts_update.cpu(), T(G["after_query_ts_update"]))
    # prune_map
    m.travel_dist, m.cur_ts = T(G["prune_travel_dist"]), 2
    m.point_certainties = T(G["after_query_certainties"]).cuda()
    assert m.prune_map(float(G["prune_thre"])) == bool(G["pruned"])
    assert torch.equal(m.neural_points.cpu(), T(G["prune_points"])) and torch.equal(m.point_ts_create.cpu(), T(G["prune_ts_create"]))
    assert torch.equal(m.geo_features.cpu(), T(G["prune_features"]))
    # recreate_hash, compacting flavour keyed on certainty: bit-exact table and arrays
    m.point_certainties = T(G["prune_certainties"]).cuda()
    m.recreate_hash(None, None, False, False, 2)
    assert torch.equal(m.buffer_pt_index.cpu().int(), T(G["rh2_table"]))
    assert torch.equal(m.neural_points.cpu(), T(G["rh2_points"])) and torch.equal(m.geo_features.cpu(), T(G["rh2_features"]))
