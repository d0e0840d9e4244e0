"""CPU: the oracle's map maintenance (adjust_map, recreate_hash, prune_map) and the after_pgo branch of the query vs golden
vectors produced by the UNMODIFIED reference (oracle/make_golden.py main_pgo -> tests/golden/ref_kitti_pgo.npz)."""
import numpy as np
import torch

from helpers import O, load_golden_pgo, rel_err


def T(a):
    return torch.from_numpy(np.asarray(a))


def test_after_pgo_sequence_matches_the_reference():
    G = load_golden_pgo()
    p = O.MapParams(buffer_size=int(G["buffer_size"]))
    s = O.new_map(p)
    s.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    for ts in range(3):
        O.map_update(s, T(G["frames%d" % ts]), torch.tensor([3.0 * ts, 0.0, 0.0]), ts)
    s.feats, s.certainty = T(G["features"]).clone(), T(G["certainties"]).clone()
    d = O.DecoderParams(*[T(G[k]).clone().requires_grad_(True) for k in ("W1", "b1", "W2", "b2")], sdf_scale=float(G["sdf_scale"]))
    # adjust_map
    O.adjust_map(s, T(G["pose_diff"]))
    assert s.after_pgo
    assert torch.equal(s.points, T(G["adj_points"]))
    assert torch.equal(s.quats, T(G["adj_quats"]))
    # recreate_hash (kept points, nearest-in-time winner) + local map
    O.recreate_hash(s, torch.tensor([6.0, 0.0, 0.0]), True, True, 2)
    assert torch.equal(s.table.int(), T(G["rh_table"]))
    assert torch.equal(s.global2local.int(), T(G["rh_global2local"]))
    # after_pgo query: feature tensor (rotated offsets), weights, sdf, gradient
    q = T(G["queries"]).clone().requires_grad_(True)
    fv, w, nn, qc, _ = O.query_feature(s, q, None, False, True)
    assert torch.equal(nn.int(), T(G["nn"]))
    assert rel_err(fv, T(G["fv"])) < 1e-6 and rel_err(w, T(G["w"])) < 1e-6 and rel_err(qc, T(G["qc"])) < 1e-6
    out = O.query_sdf(s, d, T(G["queries"]), None, False, True, want_grad=True)
    assert rel_err(out["sdf"], T(G["sdf"])) < 1e-6
    assert rel_err(out["grad"], T(G["grad"])) < 1e-5
    # rotation is not a no-op in this fixture
    ident = torch.zeros_like(s.quats)
    ident[:, 0] = 1
    assert float((s.quats - ident).abs().max()) > 1e-3
    # training-flavour backward
    s.l_feats = s.l_feats.detach().requires_grad_(True)
    o2 = O.query_sdf(s, d, T(G["queries"]), T(G["bwd_ts"]), True, True)
    (o2["sdf"] * T(G["bwd_up"])).sum().backward()
    assert rel_err(s.l_feats.grad, T(G["bwd_gfeat"])) < 1e-5
    for name, t in zip(("bwd_gW1", "bwd_gb1", "bwd_gW2", "bwd_gb2"), d.tensors()):
        assert rel_err(t.grad, T(G[name])) < 1e-5, name
    s.l_feats = s.l_feats.detach()
    O.assign_local_to_global(s)
    assert rel_err(s.certainty, T(G["after_query_certainties"])) < 1e-6
    assert torch.equal(s.ts_update, T(G["after_query_ts_update"]))
    # prune_map
    s.travel_dist, s.cur_ts = T(G["prune_travel_dist"]), 2
    assert O.prune_map(s, float(G["prune_thre"])) == bool(G["pruned"]) and bool(G["pruned"])
    assert torch.equal(s.points, T(G["prune_points"])) and torch.equal(s.ts_create, T(G["prune_ts_create"]))
    assert torch.equal(s.feats, T(G["prune_features"]))
    assert rel_err(s.certainty, T(G["prune_certainties"])) < 1e-6
    # recreate_hash, compacting flavour keyed on certainty
    s.certainty = T(G["prune_certainties"]).clone()
    O.recreate_hash(s, None, False, False, 2)
    assert torch.equal(s.table.int(), T(G["rh2_table"]))
    assert torch.equal(s.points, T(G["rh2_points"])) and torch.equal(s.feats, T(G["rh2_features"]))
