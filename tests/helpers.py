"""Shared builders for the parity tests: the same seeded inputs go to the CUDA path and to the oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import pin_oracle as O  # noqa: E402  (tests may import the oracle)


def frames_of_points(n_frames=3, n_per_frame=40000, seed=0, side=60.0):
    """A few overlapping 'frames' of surface points (CPU float32 tensors) + sensor positions."""
    from pin_slam_b200 import synth

    out = []
    for f in range(n_frames):
        p = synth.surface_points(n_per_frame, side, seed=seed + f)
        p[:, 0] += 4.0 * f
        out.append((torch.from_numpy(p), torch.tensor([4.0 * f, 0.0, 0.0])))
    return out


def build_oracle_map(frames, params=None, travel=None):
    s = O.new_map(params or O.MapParams())
    n = len(frames)
    s.travel_dist = travel if travel is not None else torch.arange(n + 1, dtype=torch.float32) * 4.0
    for ts, (pts, sensor) in enumerate(frames):
        O.map_update(s, pts, sensor, ts)
    return s


def build_cuda_map(frames, travel=None, cfg=None):
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config

    cfg = cfg or kitti_config("cuda")
    m = NeuralPoints(cfg)
    n = len(frames)
    m.travel_dist = travel if travel is not None else torch.arange(n + 1, dtype=torch.float32) * 4.0
    eye = torch.eye(3, device="cuda")
    for ts, (pts, sensor) in enumerate(frames):
        m.update(pts.cuda(), sensor.cuda(), eye, ts)
    return m, cfg


def randomize_features(oracle_map, cuda_map=None, seed=3, std=0.1):
    """Features start exactly at 0 (feature_std = 0); give both sides the same non-trivial values."""
    g = torch.Generator().manual_seed(seed)
    f = torch.randn(oracle_map.feats.shape, generator=g) * std
    f[-1] = 0
    oracle_map.feats = f.clone()
    oracle_map.l_feats = f[oracle_map.local_mask].clone()
    c = torch.rand(oracle_map.certainty.shape, generator=g) * 50
    oracle_map.certainty = c.clone()
    oracle_map.l_certainty = c[oracle_map.local_mask[:-1]].clone()
    if cuda_map is not None:
        cuda_map.geo_features = f.cuda()
        cuda_map.point_certainties = c.cuda()
        with torch.no_grad():
            cuda_map.local_geo_features.data = f.cuda()[cuda_map._local_idx]
        cuda_map.local_point_certainties = c.cuda()[cuda_map._local_idx[:-1]]


def make_queries(points_cpu, n, seed=1):
    from pin_slam_b200 import synth

    return torch.from_numpy(synth.microbench_queries(points_cpu.numpy(), n, seed=seed))


def decoder_pair(cfg, seed=42):
    """(CUDA Decoder module, oracle DecoderParams) with identical weights."""
    from pin_slam_b200.model.decoder import Decoder

    torch.manual_seed(seed)
    dec = Decoder(cfg, cfg.geo_mlp_hidden_dim, cfg.geo_mlp_level, 1)
    d = O.DecoderParams(dec.layers[0].weight.detach().cpu().clone(), dec.layers[0].bias.detach().cpu().clone(),
                        dec.lout.weight.detach().cpu().clone(), dec.lout.bias.detach().cpu().clone(),
                        sdf_scale=dec.sdf_scale)
    return dec, d


def oracle_state_from(npts, dec):
    """CPU oracle state holding the same map (and decoder) as a CUDA NeuralPoints / Decoder pair."""
    p = O.MapParams(local_map_radius=float(npts.local_map_radius))
    s = O.MapState(p=p)
    s.table = npts.buffer_pt_index.cpu()
    s.points = npts.neural_points.cpu()
    s.quats = npts.point_orientations.cpu()
    s.feats = npts.geo_features.cpu()
    s.ts_create = npts.point_ts_create.cpu()
    s.ts_update = npts.point_ts_update.cpu()
    s.certainty = npts.point_certainties.cpu()
    s.travel_dist = npts.travel_dist.cpu()
    O.set_search_neighborhood(s, p.num_nei_cells, p.search_alpha)
    O.reset_local_map(s, torch.zeros(3), 0)
    d = O.DecoderParams(*[t.detach().cpu().clone() for t in dec.fused_params()], sdf_scale=dec.sdf_scale)
    return O, s, d


def to_double(obj):
    """Shallow copy of an oracle MapState / DecoderParams with every float32 tensor promoted to float64 (the exact-arithmetic
    arbiter of the noise-floor checks; run it under torch.set_default_dtype(torch.float64))."""
    import copy
    out = copy.copy(obj)
    for k, v in vars(obj).items():
        if torch.is_tensor(v) and v.dtype == torch.float32:
            setattr(out, k, v.double())
    return out


def row_outliers(got, ref, tol=1e-5):
    """Rows of `got` [R,F] further than tol (relative to the row's norm, floored by the median norm of the touched rows)
    from `ref` -> (bool mask [R], relative error of everything else)."""
    got, ref = got.detach().double().cpu(), ref.detach().double().cpu()
    rn = ref.norm(dim=1)
    floor = rn[rn > 0].median() if bool((rn > 0).any()) else torch.tensor(1.0, dtype=torch.float64)
    bad = (got - ref).norm(dim=1) > tol * torch.maximum(rn, floor)
    rest = float((got - ref)[~bad].norm()) / max(float(ref[~bad].norm()), 1e-30)
    return bad, rest


def rel_err(a, b, floor=1e-6):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm()) / max(float(b.norm()), floor)


# ---- golden fixture (generated from the unmodified reference by oracle/make_golden.py) -------------------
GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_kitti_small.npz")


def load_golden():
    return {k: v for k, v in np.load(GOLDEN).items()}


def load_golden_frames():
    """Third fixture (oracle/make_golden.py main_frames): DataSampler.sample and three Mapper.process_frame calls."""
    return {k: v for k, v in np.load(os.path.join(ROOT, "tests", "golden", "ref_kitti_frames.npz")).items()}


def load_golden_pgo():
    """Second fixture (oracle/make_golden.py main_pgo): map maintenance around a loop closure + the after_pgo query."""
    return {k: v for k, v in np.load(os.path.join(ROOT, "tests", "golden", "ref_kitti_pgo.npz")).items()}


def golden_oracle_world(G):
    """Replay the fixture's inputs through the oracle: (state, decoder)."""
    p = O.MapParams(buffer_size=int(G["buffer_size"]))
    s = O.new_map(p)
    s.travel_dist = torch.from_numpy(G["travel_dist"])
    for ts in range(3):
        O.map_update(s, torch.from_numpy(G["frames%d" % ts]), torch.tensor([3.0 * ts, 0.0, 0.0]), ts)
    s.feats = torch.from_numpy(G["features"]).clone()
    s.certainty = torch.from_numpy(G["certainties"]).clone()
    O.reset_local_map(s, torch.tensor([6.0, 0.0, 0.0]), 2)
    d = O.DecoderParams(*[torch.from_numpy(G[k]).clone() for k in ("W1", "b1", "W2", "b2")], sdf_scale=float(G["sdf_scale"]))
    return s, d


def golden_cuda_world(G):
    """Replay the fixture's inputs through the CUDA path: (NeuralPoints, Decoder, cfg)."""
    from pin_slam_b200.model.decoder import Decoder
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config

    cfg = kitti_config("cuda")
    cfg.buffer_size = int(G["buffer_size"])
    m = NeuralPoints(cfg)
    m.travel_dist = torch.from_numpy(G["travel_dist"])
    eye = torch.eye(3, device="cuda")
    for ts in range(3):
        m.update(torch.from_numpy(G["frames%d" % ts]).cuda(), torch.tensor([3.0 * ts, 0.0, 0.0]).cuda(), eye, ts)
    m.geo_features = torch.from_numpy(G["features"]).cuda()
    m.point_certainties = torch.from_numpy(G["certainties"]).cuda()
    m.reset_local_map(torch.tensor([6.0, 0.0, 0.0]).cuda(), eye, 2)
    dec = Decoder(cfg, 64, 1, 1)
    with torch.no_grad():
        dec.layers[0].weight.copy_(torch.from_numpy(G["W1"]))
        dec.layers[0].bias.copy_(torch.from_numpy(G["b1"]))
        dec.lout.weight.copy_(torch.from_numpy(G["W2"]))
        dec.lout.bias.copy_(torch.from_numpy(G["b2"]))
    return m, dec, cfg


def load_golden_preprocess():
    """Fourth fixture (oracle/make_golden.py main_preprocess): deskewing, three frames through the reference's SLAMDataset
    (Velodyne point times, voxel down-sampling, crop, KITTI correction, source points, deskew), Mapper.dynamic_filter."""
    return {k: v for k, v in np.load(os.path.join(ROOT, "tests", "golden", "ref_kitti_preprocess.npz")).items()}
