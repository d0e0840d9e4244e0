"""CUDA: the frame data path in front of the hot path against the fourth fixture generated from the unmodified
reference: the deskew kernel, the Velodyne per-point times, three frames through the SLAMDataset mirror (voxel down-sampling,
crop, KITTI correction, guess, source points + deskew, cloud deskew) and Mapper.dynamic_filter."""
import numpy as np
import pytest
import torch

from helpers import golden_cuda_world, load_golden, load_golden_preprocess, rel_err

pytestmark = pytest.mark.gpu
T = torch.from_numpy


def test_deskew_kernel_matches_the_reference():
    from pin_slam_b200.utils.tools import deskewing, velodyne_point_ts
    P = load_golden_preprocess()
    for key, mid in (("dk_out", 0.5), ("dk_out_mid0", 0.0)):
        pts = T(P["dk_points"]).cuda().contiguous()
        got = deskewing(pts, T(P["dk_ts"]).cuda(), T(P["dk_pose"]).cuda(), mid)
        assert got.data_ptr() == pts.data_ptr()                                  # in place, like the reference
        assert float((got.cpu() - T(P[key])).abs().max()) < 3e-5                  # coordinates up to 60 m in fp32
        assert rel_err(got, T(P[key])) < 1e-6
    ts = velodyne_point_ts(T(P["pp0_scan"]).cuda())
    assert float((ts.cpu() - T(P["pp0_ts_raw"]).reshape(-1)).abs().max()) < 1e-6


def test_dataset_mirror_preprocess_matches_the_reference():
    """dataset/slam_dataset.py:199-430 through pin_slam_b200.dataset.slam_dataset.SLAMDataset with deskew and the KITTI
    correction ON (run_kitti.yaml), the tracker's answers replaced by the fixture's poses."""
    from pin_slam_b200.dataset.slam_dataset import SLAMDataset
    from pin_slam_b200.utils.config import kitti_config
    P = load_golden_preprocess()
    cfg = kitti_config("cuda")
    assert cfg.deskew and cfg.kitti_correction_on and abs(cfg.correction_deg - 0.195) < 1e-12
    n = int(P["pp_n"])
    ds = SLAMDataset(cfg, [P["pp%d_scan" % i] for i in range(n)], P["pp_gt"])
    for i in range(n):
        ds.read_frame(i)
        assert float((ds.cur_point_ts_torch.cpu() - T(P["pp%d_ts_raw" % i]).reshape(-1)).abs().max()) < 1e-6
        ds.preprocess_frame(i)
        want = T(P["pp%d_cloud" % i])[:, :3]
        assert ds.cur_point_cloud_torch.shape[0] == want.shape[0]               # same voxel winners, same crop
        assert float((ds.cur_point_cloud_torch.cpu() - want).abs().max()) < 2e-5
        assert float((ds.cur_point_ts_torch.cpu() - T(P["pp%d_ts" % i]).reshape(-1)).abs().max()) < 1e-6
        if i > 0:
            assert np.allclose(ds.cur_pose_guess_torch.cpu().numpy(), P["pp%d_guess" % i], atol=1e-12)
            src = T(P["pp%d_source" % i])[:, :3]
            assert ds.cur_source_points.shape[0] == src.shape[0]
            assert float((ds.cur_source_points.cpu() - src).abs().max()) < 3e-5
            ds.update_odom_pose(torch.tensor(P["pp%d_est" % i], dtype=torch.float64, device="cuda"))
            assert np.allclose(ds.last_odom_tran, P["pp%d_last_odom_tran" % i], atol=1e-12)
            assert float((ds.cur_point_cloud_torch.cpu() - T(P["pp%d_cloud_deskewed" % i])[:, :3]).abs().max()) < 3e-5
        ds.processed_frame += 1
    assert np.allclose(np.array(ds.travel_dist), P["pp_travel_dist"], atol=1e-12)


def test_dynamic_filter_matches_the_reference():
    """Mapper.dynamic_filter (utils/mapper.py:79-108) through the fused query: both strategies."""
    from pin_slam_b200.utils.mapper import Mapper
    P = load_golden_preprocess()
    G = load_golden()
    m, dec, cfg = golden_cuda_world(G)
    assert (cfg.dynamic_certainty_thre, cfg.dynamic_sdf_ratio_thre) == tuple(P["df_thresholds"])
    mp = Mapper(cfg, None, m, dec, None, None)
    q = T(G["queries"]).cuda()
    s1 = mp.dynamic_filter(q.clone(), False).cpu()
    s2 = mp.dynamic_filter(q.clone(), True).cpu()
    assert int((s1 != T(P["df_static_1"])).sum()) == 0
    assert int((s2 != T(P["df_static_2"])).sum()) <= 1     # |grad| within rounding of the 0.3 threshold may flip one point
    assert 0.05 < float(s2.float().mean()) < 0.95
