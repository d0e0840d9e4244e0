"""CUDA marching cubes (csrc/mc.cu, row f3) against the CPU restatement (oracle/mc_oracle.py): identical vertices (bit for bit),
identical faces, on random fields with masks, on the reference-pinned mesher grid of the golden fixture, and sharded."""
from collections import Counter

import numpy as np
import pytest
import torch

from helpers import golden_cuda_world, load_golden
from oracle import mc_oracle as M

pytestmark = pytest.mark.gpu
T = torch.from_numpy


def _mesher():
    G = load_golden()
    m, dec, cfg = golden_cuda_world(G)
    from pin_slam_b200.utils.mesher import Mesher
    return G, Mesher(cfg, m, dec, None, None), cfg


@pytest.mark.parametrize("shape,seed", [((12, 13, 11), 0), ((33, 5, 64), 1), ((2, 2, 2), 2), ((70, 65, 3), 3)])
def test_cuda_mesh_equals_the_cpu_restatement_on_random_fields(shape, seed):
    _, mesher, _ = _mesher()
    rng = np.random.default_rng(seed)
    s = rng.standard_normal(shape).astype(np.float32)
    s[rng.random(shape) < 0.02] = 0.0                                   # values exactly on the level
    mask = rng.random(shape) < 0.9
    origin, voxel = (1.5, -2.25, 0.125), 0.1
    for mk in (None, mask):
        v_ref, f_ref = M.marching_cubes(s, np.ones(shape, bool) if mk is None else mk, 0.0, voxel, origin)
        v, f = mesher.extract_mesh(T(s).cuda(), None if mk is None else T(mk).cuda(), shape, origin, voxel)
        if len(f_ref):
            used = np.unique(f_ref)
            assert np.array_equal(v.cpu().numpy()[used], v_ref[used])                      # bit-identical coordinates
        assert np.array_equal(f.cpu().numpy(), f_ref)
        # allow_degenerate=True keeps the zero-area triangles
        v2, f2 = mesher.extract_mesh(T(s).cuda(), None if mk is None else T(mk).cuda(), shape, origin, voxel, allow_degenerate=True)
        assert f2.shape[0] >= f.shape[0] and v2.shape == v.shape
    vn, fn = mesher.mc_mesh(s, mask.astype(np.float64), voxel, np.array(origin))           # the reference's numpy signature
    assert vn.dtype == np.float64 and np.array_equal(fn, M.marching_cubes(s, mask, 0.0, voxel, origin)[1])


def test_mesh_of_the_reference_pinned_grid_and_its_shards():
    """The 38 x 39 x 19 grid whose SDF / mask the unmodified reference produced (golden fixture): the mesh of the CUDA query's
    own grid equals the CPU restatement's mesh of the same grid, is closed wherever the mask does not cut it, and the slabs of
    3 ranks (recon_grid_mesh) together hold exactly the same triangles with bit-identical coordinates."""
    G, mesher, cfg = _mesher()
    lo, hi, voxel, min_nn = G["mesh_lo"], G["mesh_hi"], float(G["mesh_voxel"]), int(G["mesh_min_nn"])
    sdf, mask, first, dims, origin = mesher.query_grid(lo, hi, voxel, pad_voxel=1, mask_min_nn_count=min_nn)
    dims = [int(v) for v in dims]
    # the fixture's map carries random features: its SDF is negative everywhere, so the iso level is the median of the
    # reference's own masked values (any level exercises the same code; 0.0 is what the mesher uses on a trained map)
    level = float(np.float32(np.median(G["mesh_sdf"][G["mesh_mask"] != 0])))
    v, f = mesher.extract_mesh(sdf, mask, dims, origin, voxel, level=level)
    v_ref, f_ref = M.marching_cubes(sdf.cpu().numpy().reshape(dims), mask.cpu().numpy().reshape(dims) != 0, level, voxel, origin)
    assert len(f_ref) > 500 and np.array_equal(f.cpu().numpy(), f_ref)
    assert np.abs(v.cpu().numpy() - v_ref).max() < 1e-6
    # the reference's own grid (fixture) gives the same triangle count up to cells whose SDF differs in the last bits
    v_g, f_g = M.marching_cubes(G["mesh_sdf"].reshape(dims).astype(np.float32), G["mesh_mask"].reshape(dims) != 0, level, voxel, origin)
    assert abs(len(f_g) - len(f_ref)) <= 0.01 * len(f_ref)
    cnt = Counter(map(tuple, np.concatenate([f_ref[:, [0, 1]], f_ref[:, [1, 2]], f_ref[:, [2, 0]]])))
    assert all(k == 1 for k in cnt.values())
    m3 = mask.cpu().numpy().reshape(dims) != 0
    cellok = np.ones([d - 1 for d in dims], bool)
    for c in range(8):
        cellok &= m3[(c & 1):dims[0] - 1 + (c & 1), ((c >> 1) & 1):dims[1] - 1 + ((c >> 1) & 1), ((c >> 2) & 1):dims[2] - 1 + ((c >> 2) & 1)]
    grid = (v_ref - np.asarray(origin, np.float32)) / np.float32(voxel)
    n_open = 0
    for (a, b) in cnt:
        if (b, a) in cnt:
            continue
        n_open += 1
        # an open edge lies on a cell face: the cell on its other side is masked out or outside the grid
        mid = (grid[a] + grid[b]) / 2
        base = np.floor(mid + 1e-4).astype(int)
        near = [tuple(base - np.array(d)) for d in np.ndindex(2, 2, 2)]
        assert any((min(q) < 0) or any(q[i] >= cellok.shape[i] for i in range(3)) or not cellok[q] for q in near)
    assert n_open > 0                                                        # the mask does cut this surface
    # ---- 3 ranks, no collective
    # (recon_grid_mesh pads like the reference's get_query_from_bbx default: no padding -> compare against its own whole grid)
    whole_v, whole_f, d0, o0 = mesher.recon_grid_mesh(lo, hi, voxel, min_nn, rank=0, world=1, level=level)
    tris = lambda vv, ff: Counter(map(lambda t: tuple(map(tuple, t)), vv.cpu().numpy()[ff.cpu().numpy()].round(7).tolist()))  # noqa: E731
    want = tris(whole_v, whole_f)
    got = Counter()
    for r in range(3):
        pv, pf, d1, o1 = mesher.recon_grid_mesh(lo, hi, voxel, min_nn, rank=r, world=3, level=level)
        assert list(d1) == list(d0)
        got += tris(pv, pf)
    assert got == want and sum(want.values()) == whole_f.shape[0] > 300
    # small chunks inside one rank walk the same slab
    cv, cf, _, _ = mesher.recon_grid_mesh(lo, hi, voxel, min_nn, max_points=int(d0[1]) * int(d0[2]) * 5, level=level)
    assert tris(cv, cf) == want


def test_large_sphere_volume_and_throughput():
    _, mesher, _ = _mesher()
    n = 256
    ax = torch.arange(n, device="cuda", dtype=torch.float32)
    x, y, z = torch.meshgrid(ax, ax, ax, indexing="ij")
    r = 100.3
    sdf = (torch.sqrt((x - 127.2) ** 2 + (y - 128.4) ** 2 + (z - 126.9) ** 2) - r).contiguous()
    del x, y, z
    v, f = mesher.extract_mesh(sdf, None, (n, n, n), (0.0, 0.0, 0.0), 1.0)
    a, b, c = (v[f[:, k]].double() for k in range(3))
    vol = float((a * torch.linalg.cross(b, c)).sum() / 6)
    assert abs(-vol / (4 / 3 * np.pi * r ** 3) - 1) < 2e-3
    area = float(torch.linalg.cross(b - a, c - a).norm(dim=1).sum() / 2)
    assert abs(area / (4 * np.pi * r ** 2) - 1) < 5e-3
    assert v.shape[0] - 3 * f.shape[0] // 2 + f.shape[0] == 2                  # closed surface of genus 0: V - E + F = 2
