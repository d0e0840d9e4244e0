"""CUDA path vs golden vectors produced by the UNMODIFIED reference (tests/golden, oracle/make_golden.py).
This is the direct reference pin of the kernels: nothing of the oracle is involved."""
import numpy as np
import pytest
import torch

from helpers import golden_cuda_world, load_golden, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world():
    G = load_golden()
    m, dec, cfg = golden_cuda_world(G)
    return G, m, dec, cfg


def T(a):
    return torch.from_numpy(np.asarray(a))


def test_map_state_bit_exact(world):
    G, m, _, _ = world
    assert torch.equal(m.buffer_pt_index.cpu().int(), T(G["table"]))
    assert torch.equal(m.neural_points.cpu(), T(G["neural_points"]))
    assert torch.equal(m.point_ts_create.cpu(), T(G["ts_create"]))
    assert torch.equal(m.global2local.cpu().int(), T(G["global2local"]))
    assert torch.equal(m.local_mask.cpu(), T(G["local_mask"]))


def test_voxel_downsample_bit_exact(world):
    from pin_slam_b200.utils.tools import voxel_down_sample_min_value_torch, voxel_down_sample_torch
    G = world[0]
    assert torch.equal(voxel_down_sample_torch(T(G["frames0"]).cuda(), 0.4).cpu().int(), T(G["vds_04"]))
    assert torch.equal(voxel_down_sample_torch(T(G["frames0"]).cuda(), 0.08).cpu().int(), T(G["vds_008"]))
    got = voxel_down_sample_min_value_torch(T(G["frames1"]).cuda(), 0.4, T(G["vds_min_value_in"]).cuda())
    assert torch.equal(got.cpu().int(), T(G["vds_min_value"]))


@pytest.mark.parametrize("tf", [0, 1])
def test_radius_search_bit_exact(world, tf):
    G, m, _, _ = world
    d2, idx = m.radius_neighborhood_search(T(G["queries"]).cuda(), bool(tf))
    assert torch.equal(idx.cpu().int(), T(G["rs_idx_%d" % tf]))
    assert torch.equal(d2.cpu(), T(G["rs_d2_%d" % tf]))


def test_query_certainty(world):
    G, m, _, _ = world
    m.set_search_neighborhood(1, 0.0)
    try:
        got = m.query_certainty(T(G["queries"]).cuda()).cpu()
    finally:
        m.set_search_neighborhood(2, 0.5)
    assert torch.equal(got, T(G["query_certainty"]))


@pytest.mark.parametrize("tag,local", [("L", True), ("G", False)])
def test_query_feature_and_fused_sdf(world, tag, local):
    G, m, dec, _ = world
    q = T(G["queries"]).cuda()
    fv, _, w, nn, qc = m.query_feature(q, training_mode=False, query_locally=local)
    assert nn.dtype == torch.int64 and torch.equal(nn.cpu().int(), T(G["nn_" + tag]))
    assert torch.equal(fv.cpu(), T(G["fv_" + tag]))
    assert rel_err(w, T(G["w_" + tag])) < 1e-6
    assert rel_err(qc, T(G["qc_" + tag])) < 1e-6
    with torch.no_grad():
        out = m.query_sdf(q, dec, None, False, local, want_grad=True, want_std=True)
    assert rel_err(out["sdf"], T(G["sdf_" + tag])) < 1e-5
    assert rel_err(out["grad"], T(G["grad_" + tag])) < 1e-5
    assert rel_err(out["std"], T(G["std_" + tag])) < 1e-4
    # un-fused API chain == reference chain as well (Decoder.sdf on the [N,K,11] tensor)
    sdf_unfused = torch.sum(dec.sdf(fv) * w, dim=1).squeeze(1)
    assert rel_err(sdf_unfused, T(G["sdf_" + tag])) < 1e-5


def test_registration_step(world):
    from pin_slam_b200.utils.tracker import Tracker
    G, m, dec, cfg = world
    trk = Tracker(cfg, m, dec, None, None)
    Tm, _, _, _, n_valid, res_cm, _ = trk.registration_step(T(G["reg_src"]).cuda(), None, None, None, None, 0.0, 1e3,
                                                            cfg.reg_GM_dist_m, cfg.reg_GM_grad, cfg.reg_lm_lambda)
    assert n_valid == int(G["reg_valid"])
    assert abs(res_cm - float(G["reg_res_cm"])) < 1e-5 * max(1.0, float(G["reg_res_cm"]))
    assert np.allclose(Tm, G["reg_T"], rtol=1e-4, atol=1e-7)


def test_mapping_loss_and_gradients(world):
    from pin_slam_b200.utils.mapper import Mapper
    G, m, dec, cfg = world
    cert0, ts0 = m.local_point_certainties.clone(), m.local_point_ts_update.clone()
    m.local_point_certainties = T(G["map_cert_before"]).cuda()
    mp = Mapper(cfg, None, m, dec, None, None)
    for p in dec.parameters():
        p.grad = None
    m.local_geo_features.grad = None
    loss, pred = mp.mapping_loss(T(G["map_coord"]).cuda(), T(G["map_label"]).cuda(), T(G["map_ts"]).cuda(),
                                 T(G["map_weight"]).cuda())
    loss.backward()
    try:
        assert rel_err(pred, T(G["map_pred"])) < 1e-5
        assert abs(float(loss) - float(G["map_loss"])) < 1e-5 * abs(float(G["map_loss"]))
        assert rel_err(m.local_geo_features.grad, T(G["map_gfeat"])) < 1e-5
        assert rel_err(dec.layers[0].weight.grad, T(G["map_gW1"])) < 1e-5
        assert rel_err(dec.layers[0].bias.grad, T(G["map_gb1"])) < 1e-5
        assert rel_err(dec.lout.weight.grad, T(G["map_gW2"])) < 1e-5
        assert rel_err(dec.lout.bias.grad, T(G["map_gb2"])) < 1e-5
        assert rel_err(m.local_point_certainties, T(G["map_cert_after"])) < 1e-5
        assert torch.equal(m.local_point_ts_update.cpu(), T(G["map_ts_update_after"]))
    finally:
        m.local_point_certainties, m.local_point_ts_update = cert0, ts0


def test_h13_out_of_radius_candidates(world):
    """Reference quirk (neural_points.py:301): candidates outside the local-map radius gather local row 1."""
    G, m, dec, _ = world
    old = m.local_map_radius
    m.local_map_radius = float(G["h13_radius"])
    try:
        m.reset_local_map(torch.tensor([6.0, 0.0, 0.0]).cuda(), torch.eye(3).cuda(), 2)
        assert m.local_count() == int(G["h13_local_count"])
        q = T(G["queries"]).cuda()
        fv, _, w, nn, qc = m.query_feature(q, training_mode=False, query_locally=True)
        assert torch.equal(nn.cpu().int(), T(G["h13_nn"]))
        assert torch.equal(fv.cpu(), T(G["h13_fv"]))
        assert rel_err(w, T(G["h13_w"])) < 1e-6 and rel_err(qc, T(G["h13_qc"])) < 1e-6
        with torch.no_grad():
            out = m.query_sdf(q, dec, None, False, True, want_grad=True)
        assert rel_err(out["sdf"], T(G["h13_sdf"])) < 1e-5
        assert rel_err(out["grad"], T(G["h13_grad"])) < 1e-5
    finally:
        m.local_map_radius = old
        m.reset_local_map(torch.tensor([6.0, 0.0, 0.0]).cuda(), torch.eye(3).cuda(), 2)


@pytest.mark.parametrize("device_loop", [True, False])
def test_tracking_loop(world, device_loop):
    """Tracker.tracking against the reference's own loop output (utils/tracker.py:39-174): three Gauss-Newton iterations on
    the device-side loop (pinq_tracking_run) and on the per-iteration host loop; then run_kitti's validity gates on the same
    rough field -> invalid, the initial guess comes back."""
    from pin_slam_b200.utils.tracker import Tracker
    G, m, dec, cfg = world
    src, init = T(G["reg_src"]).cuda(), T(G["trk_init"]).cuda()
    old = (cfg.reg_iter_n, cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio)
    try:
        cfg.reg_iter_n, cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio = 3, 0.0, 1e3, 1e3
        trk = Tracker(cfg, m, dec, None, None)
        trk.device_loop = device_loop
        Tr, _, _, valid = trk.tracking(src, init)
        assert valid == bool(G["trk_valid_relaxed"]) and trk.last_iterations == 3
        assert np.allclose(Tr.cpu().numpy(), G["trk_T_relaxed"], rtol=1e-3, atol=2e-4)
        cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio = old[1:]
        Td, _, _, valid_d = trk.tracking(src, init)
        assert valid_d == bool(G["trk_valid_default"])
        assert np.array_equal(Td.cpu().numpy(), G["trk_T_default"])
    finally:
        cfg.reg_iter_n, cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, cfg.max_sdf_std_ratio = old


@pytest.mark.parametrize("world_size", [1, 3])
def test_mesher_grid(world, world_size):
    """Mesher against the reference's own grid query (utils/mesher.py:41-180): the API form (materialised coordinates) and
    the sharded in-kernel form (pinq_mesher_grid_query; concatenated rank ranges == the reference's full grid)."""
    from pin_slam_b200.utils.mesher import Mesher
    G, m, dec, cfg = world
    mesher = Mesher(cfg, m, dec, None, None)
    lo, hi, voxel, min_nn = G["mesh_lo"], G["mesh_hi"], float(G["mesh_voxel"]), int(G["mesh_min_nn"])
    coord, num, origin = mesher.get_query_from_bbx((lo, hi), voxel, 1, 0)
    assert [int(v) for v in num] == G["mesh_num_xyz"].tolist() and np.array_equal(origin, G["mesh_origin"])
    assert torch.equal(torch.stack((coord[0], coord[-1])).cpu(), T(G["mesh_coord_first_last"]))
    sdf, _, _, mask = mesher.query_points(coord, 4096, True, False, False, True, False, min_nn)
    assert np.array_equal(mask.astype(np.uint8), G["mesh_mask"])
    assert rel_err(torch.from_numpy(sdf), T(G["mesh_sdf"])) < 1e-5
    parts = [mesher.query_grid(lo, hi, voxel, pad_voxel=1, mask_min_nn_count=min_nn, rank=r, world=world_size) for r in range(world_size)]
    sdf_g = torch.cat([p[0] for p in parts]).cpu()
    mask_g = torch.cat([p[1] for p in parts]).cpu()
    assert torch.equal(mask_g, T(G["mesh_mask"]))
    assert rel_err(sdf_g, T(G["mesh_sdf"])) < 1e-5
