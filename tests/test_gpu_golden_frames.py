"""CUDA path vs golden vectors of the UNMODIFIED reference for the mapper's per-frame data path:
three Mapper.process_frame calls fed with the samples the reference drew -- map update from the samples, fused replay-pool
upkeep (csrc/pool.cu), new-sample certainty query -- tests/golden/ref_kitti_frames.npz (oracle/make_golden.py main_frames)."""
import numpy as np
import pytest
import torch

from helpers import load_golden_frames

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("fused_pool", [True, False])
def test_process_frame_matches_the_reference(fused_pool):
    from pin_slam_b200.model.decoder import Decoder
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config
    from pin_slam_b200.utils.mapper import Mapper
    G = load_golden_frames()
    cfg = kitti_config("cuda")
    cfg.buffer_size, cfg.window_radius, cfg.silence = int(G["buffer_size"]), float(G["window_radius"]), True
    torch.manual_seed(42)
    dec = Decoder(cfg, 64, 1, 1)
    m = NeuralPoints(cfg)
    m.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    mp = Mapper(cfg, None, m, dec, None, None)
    mp.fused_pool = fused_pool
    for f in range(3):
        samples = tuple(T(G["f%d_%s" % (f, k)]).cuda() for k in ("coord", "label", "weight"))
        mp.process_frame(T(G["f%d_scan" % f]).cuda(), None, T(G["f%d_pose" % f]).cuda(), f, False, samples=samples)
        assert torch.equal(m.neural_points.cpu(), T(G["f%d_map_points" % f])), f
        assert [mp.cur_sample_count, mp.pool_sample_count] == G["f%d_counts" % f].tolist(), f
        assert torch.equal(mp.coord_pool.cpu(), T(G["f%d_pool_coord" % f]))
        assert torch.equal(mp.sdf_label_pool.cpu(), T(G["f%d_pool_label" % f]))
        assert torch.equal(mp.weight_pool.cpu(), T(G["f%d_pool_weight" % f]))
        assert torch.equal(mp.time_pool.cpu(), T(G["f%d_pool_ts" % f]))
        assert float((mp.global_coord_pool.cpu() - T(G["f%d_pool_global" % f])).abs().max()) < 1e-5
        assert torch.equal(mp.new_idx.cpu(), T(G["f%d_new_idx" % f])), f


@pytest.mark.parametrize("fused_mapping", [True, False])
def test_mapping_iterations_match_the_reference(fused_mapping):
    """Mapper.mapping(4) after the three frames against the reference's trained map (get_batch's draws on seed 77, loss, Adam,
    write-back) -- the device-side loop (pinq_mapping_run) and the per-iteration autograd form.  Same statistical comparison
    as the CPU pin of the oracle (Adam with eps 1e-15 amplifies rounding noise of near-zero gradients to +-lr)."""
    from pin_slam_b200.model.decoder import Decoder
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config
    from pin_slam_b200.utils.mapper import Mapper
    G = load_golden_frames()
    cfg = kitti_config("cuda")
    cfg.buffer_size, cfg.window_radius, cfg.silence = int(G["buffer_size"]), float(G["window_radius"]), True
    torch.manual_seed(42)
    dec = Decoder(cfg, 64, 1, 1)
    m = NeuralPoints(cfg)
    m.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    mp = Mapper(cfg, None, m, dec, None, None)
    mp.used_poses = None
    for f in range(3):
        samples = tuple(T(G["f%d_%s" % (f, k)]).cuda() for k in ("coord", "label", "weight"))
        mp.process_frame(T(G["f%d_scan" % f]).cuda(), None, T(G["f%d_pose" % f]).cuda(), f, False, samples=samples)
    before = T(G["map_features_before"])
    assert torch.equal(m.geo_features.cpu(), before)
    cfg.bs, cfg.bs_new_sample = int(G["map_bs"]), int(G["map_bs_new"])
    mp.train_less, mp.fused_mapping = False, fused_mapping
    mp.generator = torch.Generator().manual_seed(int(G["map_seed"]))
    mp.mapping(int(G["map_iters"]))
    ref = T(G["map_features_after"])
    got = m.geo_features.detach().cpu()
    moved = (ref - before).abs() > 0
    assert float(moved.float().mean()) > 0.01
    assert float((((got - before).abs() > 0) != moved).float().mean()) < 1e-3      # the same rows were trained
    assert float(((got - ref).abs() > 1e-4)[moved].float().mean()) < 0.03
    for k, t in zip(("W1", "b1", "W2", "b2"), dec.fused_params()):
        assert float((t.detach().cpu() - T(G["map_%s_after" % k])).abs().max()) < 2e-3, k
    assert float((m.point_certainties.cpu() - T(G["map_certainties_after"])).abs().max()) < 1e-3
    assert torch.equal(m.point_ts_update.cpu(), T(G["map_ts_update_after"]))
