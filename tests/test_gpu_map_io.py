"""The pin_map.pth on-disk contract: the reference pickles the WHOLE NeuralPoints module after clear_temp()
(utils/tools.py:224-238, pin_slam.py end of run) and vis_pin_map.py:54-70 reloads it, calls
recreate_hash(neural_points[0], eye(3), False, False) and queries it globally through the Mesher.  The mirror keeps its
native views (table, occupancy index, bricks, candidate records) out of the pickle and rebuilds them from the SoA arrays."""
import io

import numpy as np
import pytest
import torch

from helpers import build_cuda_map, decoder_pair, frames_of_points, make_queries, randomize_features, build_oracle_map

pytestmark = pytest.mark.gpu


def test_pin_map_save_load_requery_round_trip(tmp_path):
    from pin_slam_b200.model.decoder import Decoder
    from pin_slam_b200.utils.mesher import Mesher

    frames = frames_of_points(3, 30000)
    s = build_oracle_map(frames)
    m, cfg = build_cuda_map(frames)
    randomize_features(s, m)
    dec, _ = decoder_pair(cfg)
    q = make_queries(s.points, 20000).cuda()
    eye = torch.eye(3, device="cuda")

    # what a loaded map is asked: global, inference-mode queries after recreate_hash(..., False, False)
    m.recreate_hash(m.neural_points[0], eye, False, False)
    with torch.no_grad():
        before = m.query_sdf(q, dec, None, False, False, want_grad=True)
        fv0, _, w0, nn0, _ = m.query_feature(q, training_mode=False, query_locally=False)
    table0 = m.buffer_pt_index.clone()
    mesher = Mesher(cfg, m, dec, None, None)
    lo, hi = s.points.min(0).values.numpy(), s.points.max(0).values.numpy()
    sdf_g0, mask_g0, _, dims0, _ = mesher.query_grid(lo, hi, 0.4, mask_min_nn_count=cfg.mesh_min_nn)

    m.clear_temp()                                                  # pin_slam.py: before save_implicit_map
    path = tmp_path / "pin_map.pth"
    torch.save({"neural_points": m, "geo_decoder": dec.state_dict()}, path)   # utils/tools.py:226-233
    assert path.stat().st_size < 40 * m.count() * 4 + (1 << 20)     # no 200 MB table, no index views in the file

    loaded = torch.load(path, weights_only=False)                  # vis_pin_map.py:54
    m2 = loaded["neural_points"]
    torch.manual_seed(0)
    dec2 = Decoder(cfg, cfg.geo_mlp_hidden_dim, cfg.geo_mlp_level, 1)
    dec2.load_state_dict(loaded["geo_decoder"])
    m2.recreate_hash(m2.neural_points[0], eye, False, False)        # vis_pin_map.py:70
    assert torch.equal(m2.buffer_pt_index, table0)
    with torch.no_grad():
        after = m2.query_sdf(q, dec2, None, False, False, want_grad=True)
        fv1, _, w1, nn1, _ = m2.query_feature(q, training_mode=False, query_locally=False)
    assert torch.equal(after["nn_count"], before["nn_count"])
    assert torch.equal(after["sdf"], before["sdf"]) and torch.equal(after["grad"], before["grad"])
    assert torch.equal(fv1, fv0) and torch.equal(w1, w0) and torch.equal(nn1, nn0)
    mesher2 = Mesher(cfg, m2, dec2, None, None)
    sdf_g1, mask_g1, _, dims1, _ = mesher2.query_grid(lo, hi, 0.4, mask_min_nn_count=cfg.mesh_min_nn)
    assert list(dims0) == list(dims1) and torch.equal(mask_g1, mask_g0) and torch.equal(sdf_g1, sdf_g0)
    assert float(mask_g1.float().mean()) > 0.01

    # the pickle survives an in-memory round trip too (torch.save to a buffer, as checkpoints do)
    buf = io.BytesIO()
    m2.clear_temp()
    torch.save(m2, buf)
    buf.seek(0)
    m3 = torch.load(buf, weights_only=False)
    m3.recreate_hash(m3.neural_points[0], eye, False, False)
    with torch.no_grad():
        again = m3.query_sdf(q, dec2, None, False, False, want_grad=True)
    assert torch.equal(again["sdf"], before["sdf"])
    assert np.isfinite(again["sdf"].cpu().numpy()).all()
