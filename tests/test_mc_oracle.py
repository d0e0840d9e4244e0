"""CPU: the marching-cubes restatement (oracle/mc_oracle.py, row f3) -- properties of the derived case table and of the
meshes it produces, and equality with the table libpinq derives in C++ (pinq_mc_tables is host-only: no GPU needed).
Parity with skimage.measure.marching_cubes is UNPINNED: scikit-image is not installed here nor vendored by the reference."""
from collections import Counter

import numpy as np

from oracle import mc_oracle as M


def _edges(f):
    return Counter(map(tuple, np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])))


def test_case_table_is_face_consistent_and_bounded():
    tri, ntri = M.tables()
    assert int(ntri.max()) == 5 and ntri[0] == 0 and ntri[255] == 0
    for case in range(256):
        used = tri[case][tri[case] >= 0]
        assert len(used) == 3 * ntri[case]
        cut = {e for e in range(12) if ((case >> M.edge_endpoints(e)[0]) & 1) != ((case >> M.edge_endpoints(e)[1]) & 1)}
        assert set(int(e) for e in used) == cut          # every intersected edge carries exactly the vertices used


def test_cpp_table_equals_python_table():
    from pin_slam_b200 import _pinq as P
    lib = P.load()
    tri = np.zeros((256, 16), np.int8)
    nt = np.zeros(256, np.uint8)
    assert lib.pinq_mc_tables(tri.ctypes.data, nt.ctypes.data) == 0
    t2, n2 = M.tables()
    assert np.array_equal(nt, n2) and np.array_equal(tri[:, :15], t2[:, :15]) and (tri[:, 15] == -1).all()


def test_sphere_is_closed_oriented_and_has_the_right_volume():
    n = 28
    g = np.stack(np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij"), -1).astype(np.float32)
    r = 9.3
    sdf = (np.linalg.norm(g - np.array([13.3, 14.1, 13.7], np.float32), axis=-1) - r).astype(np.float32)
    v, f = M.marching_cubes(sdf, np.ones_like(sdf, bool), 0.0, 0.5, (1.0, 2.0, 3.0))
    cnt = _edges(f)
    assert all(k == 1 and cnt[(b, a)] == 1 for (a, b), k in cnt.items())          # closed 2-manifold, consistently oriented
    a, b, c = (v[f[:, k]].astype(np.float64) for k in range(3))
    vol = np.einsum("ij,ij->i", a, np.cross(b, c)).sum() / 6
    assert vol < 0                                                                 # normals towards decreasing values (inside)
    assert abs(-vol / (4 / 3 * np.pi * (r * 0.5) ** 3) - 1) < 0.02
    assert len(v) - len(cnt) // 2 + len(f) == 2                                     # Euler characteristic of a sphere
    rad = np.linalg.norm(v - (np.array([13.3, 14.1, 13.7]) * 0.5 + np.array([1.0, 2.0, 3.0])), axis=1)
    assert np.abs(rad - r * 0.5).max() < 0.02


def test_random_field_surface_is_closed_except_at_the_grid_and_mask_boundary():
    rng = np.random.default_rng(0)
    s = rng.standard_normal((12, 13, 11)).astype(np.float32)           # every ambiguous configuration occurs
    v, f = M.marching_cubes(s, np.ones_like(s, bool))
    cnt = _edges(f)
    assert all(k == 1 for k in cnt.values())
    hi = np.array(s.shape) - 1
    for (a, b) in cnt:
        if (b, a) not in cnt:
            assert (((v[a] == 0) & (v[b] == 0)) | ((v[a] == hi) & (v[b] == hi))).any()   # open edges lie on the grid boundary
    m = np.ones_like(s, bool)
    m[4:8, 5:9, 3:7] = False
    v2, f2 = M.marching_cubes(s, m)
    assert 0 < len(f2) < len(f)
    # a masked-out corner removes all 8 cells around it: no vertex inside the removed block
    inside = ((v2 > np.array([3, 4, 2])) & (v2 < np.array([8, 9, 7]))).all(1)
    assert not inside.any()


def test_empty_and_degenerate_inputs():
    z = np.ones((4, 4, 4), np.float32)
    v, f = M.marching_cubes(z, np.ones_like(z, bool))
    assert v.shape == (0, 3) and f.shape == (0, 3)
    v, f = M.marching_cubes(-z, np.zeros_like(z, bool))
    assert len(f) == 0
    s = np.zeros((3, 3, 3), np.float32) + 1
    s[1, 1, 1] = -1
    v, f = M.marching_cubes(s, np.ones_like(s, bool))
    assert len(f) == 8 and len(v) == 6                                    # an octahedron around the single inside point
    s[1, 1, 1] = 0.0                                                      # exactly on the level: "outside" (v < level is false)
    assert len(M.marching_cubes(s, np.ones_like(s, bool))[1]) == 0
