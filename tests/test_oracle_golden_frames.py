"""CPU: the per-frame data path of the mapper vs golden vectors of the UNMODIFIED reference
(oracle/make_golden.py main_frames -> tests/golden/ref_kitti_frames.npz): DataSampler.sample on a seeded RNG, and three
Mapper.process_frame calls (map update from the samples, replay-pool append + window filter, new-sample certainty query).
Pins (1) the torch formulation of the sampler that the CUDA kernel is tested against, (2) the oracle's SLAM loop
(oracle/slam_oracle.py), (3) Mapper._pool_update_torch, the partner of the fused pool kernels."""
import numpy as np
import torch

from helpers import O, load_golden_frames


def T(a):
    return torch.from_numpy(np.asarray(a))


def _cfg(G):
    from pin_slam_b200.utils.config import kitti_config
    cfg = kitti_config("cpu")
    cfg.buffer_size = int(G["buffer_size"])
    cfg.window_radius = float(G["window_radius"])
    cfg.silence = True
    return cfg


def test_sampler_formulation_matches_the_reference():
    from pin_slam_b200.utils.data_sampler import DataSampler
    G = load_golden_frames()
    torch.manual_seed(5)
    coord, label, _, _, _, weight = DataSampler(_cfg(G)).sample(T(G["s_points"]), None, None, None)
    assert torch.equal(coord, T(G["s_coord"]))
    assert torch.equal(label, T(G["s_label"]))
    assert torch.equal(weight, T(G["s_weight"]))


class _Replay:
    """Sampler stand-in handing out the samples the reference drew."""

    def __init__(self, G):
        self.G, self.f = G, 0

    def sample(self, *a, **k):
        f = self.f
        self.f += 1
        return T(self.G["f%d_coord" % f]), T(self.G["f%d_label" % f]), None, None, None, T(self.G["f%d_weight" % f])


def test_oracle_process_frame_matches_the_reference():
    from oracle.slam_oracle import OracleSLAM
    G = load_golden_frames()
    slam = OracleSLAM(_cfg(G), _Replay(G), torch.Generator().manual_seed(0))
    slam.s.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    for f in range(3):
        slam.process_frame(T(G["f%d_scan" % f]), G["f%d_pose" % f], f)
        assert torch.equal(slam.s.points, T(G["f%d_map_points" % f])), f
        assert [slam.cur_sample_count, slam.pool_sample_count] == G["f%d_counts" % f].tolist(), f
        assert torch.equal(slam.coord_pool, T(G["f%d_pool_coord" % f]))
        assert torch.equal(slam.label_pool, T(G["f%d_pool_label" % f])) and torch.equal(slam.weight_pool, T(G["f%d_pool_weight" % f]))
        assert torch.equal(slam.time_pool, T(G["f%d_pool_ts" % f]))
        assert float((slam.global_pool - T(G["f%d_pool_global" % f])).abs().max()) < 1e-5
        assert torch.equal(slam.new_idx, T(G["f%d_new_idx" % f])), f
    assert G["f2_counts"][1] < sum(G["f%d_coord" % f].shape[0] for f in range(3))   # the window filter dropped samples


def test_pool_update_torch_formulation_matches_the_reference():
    from pin_slam_b200.utils.mapper import Mapper
    G = load_golden_frames()
    cfg = _cfg(G)
    mp = Mapper(cfg, None, None, None, None, None)
    for f in range(3):
        coord, label, weight = T(G["f%d_coord" % f]), T(G["f%d_label" % f]), T(G["f%d_weight" % f])
        pose = T(G["f%d_pose" % f])
        gcoord = O.transform_points(coord, pose.float())
        ts = torch.full((coord.shape[0],), f, dtype=torch.long)
        mp.cur_sample_count = coord.shape[0]
        mp._pool_update_torch(coord, gcoord, label, weight, ts, f, pose[:3, 3])
        assert [mp.cur_sample_count, mp.pool_sample_count] == G["f%d_counts" % f].tolist(), f
        assert torch.equal(mp.coord_pool, T(G["f%d_pool_coord" % f])) and torch.equal(mp.time_pool, T(G["f%d_pool_ts" % f]))
        assert torch.equal(mp.sdf_label_pool, T(G["f%d_pool_label" % f])) and torch.equal(mp.weight_pool, T(G["f%d_pool_weight" % f]))
        assert float((mp.global_coord_pool - T(G["f%d_pool_global" % f])).abs().max()) < 1e-5


def test_oracle_mapping_iterations_match_the_reference():
    """Mapper.mapping(4) of the reference after the three frames (get_batch's draws on torch.manual_seed(77), BCE + eikonal
    loss, Adam, write-back).  Adam with eps 1e-15 turns rounding noise of a near-zero gradient into +-lr steps, so the
    comparison is: where the reference moved a feature, the oracle lands within 1e-4 for all but a few per cent."""
    from oracle.slam_oracle import OracleSLAM
    G = load_golden_frames()
    cfg = _cfg(G)
    slam = OracleSLAM(cfg, _Replay(G), torch.Generator().manual_seed(0))
    slam.s.travel_dist = torch.tensor([0.0, 3.0, 6.0])
    for f in range(3):
        slam.process_frame(T(G["f%d_scan" % f]), G["f%d_pose" % f], f)
    assert torch.equal(slam.s.feats, T(G["map_features_before"]))
    cfg.bs, cfg.bs_new_sample = int(G["map_bs"]), int(G["map_bs_new"])
    slam.gen = torch.Generator().manual_seed(int(G["map_seed"]))
    slam.mapping(int(G["map_iters"]))
    ref, before = T(G["map_features_after"]), T(G["map_features_before"])
    moved = (ref - before).abs() > 0
    assert float(moved.float().mean()) > 0.01
    assert torch.equal((slam.s.feats - before).abs() > 0, moved)                      # the same rows were trained
    assert float(((slam.s.feats - ref).abs() > 1e-4)[moved].float().mean()) < 0.03
    for k, t in zip(("W1", "b1", "W2", "b2"), slam.d.tensors()):
        assert float((t.detach() - T(G["map_%s_after" % k])).abs().max()) < 2e-3, k
    assert float((slam.s.certainty - T(G["map_certainties_after"])).abs().max()) < 1e-3
    assert torch.equal(slam.s.ts_update, T(G["map_ts_update_after"]))
