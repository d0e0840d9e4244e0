"""CPU: the oracle's restatement of the frame data path in front of the hot path against the fourth
fixture generated from the unmodified reference (oracle/make_golden.py main_preprocess)."""
import numpy as np
import torch

from helpers import O, golden_oracle_world, load_golden, load_golden_preprocess, rel_err

T = torch.from_numpy


def test_deskewing_matches_the_reference():
    """utils/tools.py:540-567 with roma.rotmat_slerp restated: both ts_mid_pose settings."""
    P = load_golden_preprocess()
    for key, mid in (("dk_out", 0.5), ("dk_out_mid0", 0.0)):
        got = O.deskew(T(P["dk_points"]), T(P["dk_ts"]), T(P["dk_pose"]), mid)
        assert float((got - T(P[key])).abs().max()) < 2e-5      # coordinates up to 60 m, fp32: a few ulp
        assert rel_err(got, T(P[key])) < 1e-6
    moved = float((T(P["dk_out"]) - T(P["dk_points"])).norm(dim=1).max())
    assert 0.3 < moved < 2.5                                     # the pose is a real motion, not the identity


def test_preprocess_frames_match_the_reference():
    """dataset/slam_dataset.py:199-430 on three motion-distorted sweeps: per-point times, voxel down-sampling (0.08 m),
    crop, KITTI intrinsic correction, constant-velocity guess, source down-sampling (0.8 m) + deskew with the previous
    motion, cloud deskew with the estimated motion."""
    P = load_golden_preprocess()
    last_pose, last_odom = np.eye(4), np.eye(4)
    last_odom[0, 3] = 80.0 * 1e-2                                # kitti_correct: first motion guess (slam_dataset.py:126)
    for i in range(int(P["pp_n"])):
        scan = T(P["pp%d_scan" % i])                              # the .bin files carry a zero intensity column, dropped here
        ts = O.velodyne_point_ts(scan)
        assert float((ts - T(P["pp%d_ts_raw" % i]).reshape(-1)).abs().max()) < 1e-6
        idx = O.voxel_down_sample(scan, 0.08)
        pc, ts = scan[idx], ts[idx]
        pc, ts = O.crop_frame(pc, ts, -3.5, 60.0, 3.0, 80.0)
        pc = O.intrinsic_correct(pc, 0.195)
        assert pc.shape[0] == P["pp%d_cloud" % i].shape[0]
        assert torch.equal(ts, T(P["pp%d_ts" % i]).reshape(-1))
        # the cloud after preprocess_frame (of frames >= 1 it is deskewed in place later: the fixture holds both states)
        assert float((pc - T(P["pp%d_cloud" % i])[:, :3]).abs().max()) < 1e-5
        if i == 0:
            continue
        guess = last_pose @ last_odom
        assert np.allclose(guess, P["pp%d_guess" % i], atol=1e-12)
        sidx = O.voxel_down_sample(pc, 0.8)
        src = O.deskew(pc[sidx], ts[sidx], torch.tensor(last_odom, dtype=torch.float32))
        assert src.shape[0] == P["pp%d_source" % i].shape[0]
        assert float((src - T(P["pp%d_source" % i])[:, :3]).abs().max()) < 2e-5
        est = P["pp%d_est" % i]
        last_odom = np.linalg.inv(last_pose) @ est
        assert np.allclose(last_odom, P["pp%d_last_odom_tran" % i], atol=1e-12)
        last_pose = est
        cloud = O.deskew(pc, ts, torch.tensor(last_odom, dtype=torch.float32))
        assert float((cloud - T(P["pp%d_cloud_deskewed" % i])[:, :3]).abs().max()) < 2e-5
        if i == 1:
            last_pose_0 = np.eye(4)
            assert np.allclose(P["pp_travel_dist"][1], np.linalg.norm((np.linalg.inv(last_pose_0) @ est)[:3, 3]))
    # frame 0 sets last_pose = identity pose of the first frame; the loop above started from it


def test_dynamic_filter_matches_the_reference():
    """utils/mapper.py:79-108: strategy 1 (certain free space) and strategy 2 (+ gradient norm) on the first fixture's map."""
    P = load_golden_preprocess()
    G = load_golden()
    s, d = golden_oracle_world(G)
    q = T(G["queries"])
    out = O.query_sdf(s, d, q, None, False, True, want_grad=True)
    fv, w, nn, qc, _ = O.query_feature(s, q, None, False, True)
    cert_thre, ratio = P["df_thresholds"]
    static = (qc < cert_thre) | (out["sdf"] < ratio * 0.4)
    assert torch.equal(static, T(P["df_static_1"]))
    static2 = static & ((out["grad"].norm(dim=-1) > 0.3) | (qc < 0.5))
    assert torch.equal(static2, T(P["df_static_2"]))
    assert 0.05 < float(T(P["df_static_2"]).float().mean()) < 0.95
