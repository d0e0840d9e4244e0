"""Mesher -- dense SDF grid query (reference: utils/mesher.py:19-180).

Only the naturally sharded consumer of the hot path is in scope: `query_points` and
`get_query_from_bbx`.  `query_grid` is the B200 form of the two combined:
grid coordinates are generated inside the kernel from (origin, dims, voxel) -- no [N,3] tensor is
materialised -- and a linear range of the grid can be given so that ranks shard the grid without any
collective.  `extract_mesh` / `mc_mesh` / `recon_grid_mesh` are marching cubes on the device (row f3:
csrc/mc.cu replaces the host-side skimage call of utils/mesher.py:310-337); Open3D post-processing is out of scope.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from .. import _pinq as P


class Mesher:
    def __init__(self, config, neural_points, geo_decoder, sem_decoder, color_decoder):
        self.config = config
        self.silence = config.silence
        self.neural_points = neural_points
        self.geo_decoder = geo_decoder
        self.sem_decoder = sem_decoder
        self.color_decoder = color_decoder
        self.device = config.device
        self.cur_device = self.device
        self.dtype = config.dtype
        self.ts = 0
        self.global_transform = np.eye(4)

    def query_points(self, coord, bs, query_sdf=True, query_sem=False, query_color=False, query_mask=True,
                     query_locally=False, mask_min_nn_count: int = 4, out_torch: bool = False):
        """reference: utils/mesher.py:41-136 -> (sdf_pred, None, None, mc_mask); numpy float64 arrays unless
        out_torch (the reference allocates np.zeros(sample_count) for both outputs)."""
        if query_sem or query_color:
            raise NotImplementedError("semantic / colour meshing is out of scope")
        n = coord.shape[0]
        sdf_all = torch.zeros(n, device=self.device) if query_sdf else None
        mask_all = torch.zeros(n, device=self.device) if query_mask else None
        with torch.no_grad():
            for head in range(0, n, bs):
                tail = min(head + bs, n)
                batch = coord[head:tail]
                if not batch.is_cuda:
                    batch = batch.to(self.device)
                out = self.neural_points.query_sdf(batch, self.geo_decoder, None, False, query_locally)
                if query_sdf:
                    sdf_all[head:tail] = out["sdf"]
                if query_mask:
                    mask_all[head:tail] = (out["nn_count"] >= mask_min_nn_count).to(mask_all.dtype)
        if out_torch:
            return (sdf_all.cpu() if query_sdf else None), None, None, (mask_all.cpu() if query_mask else None)
        return (sdf_all.cpu().numpy().astype(np.float64) if query_sdf else None), None, None, \
               (mask_all.cpu().numpy().astype(np.float64) if query_mask else None)

    @staticmethod
    def grid_from_bounds(min_bound, max_bound, voxel_size, pad_voxel=0, skip_top_voxel=0):
        """Grid geometry of utils/mesher.py:152-160 -> (voxel_num_xyz int[3], voxel_origin float64[3])."""
        min_bound = np.asarray(min_bound, dtype=np.float64)
        max_bound = np.asarray(max_bound, dtype=np.float64)
        voxel_num_xyz = (np.ceil((max_bound - min_bound) / voxel_size) + pad_voxel * 2).astype(np.int_)
        voxel_origin = min_bound - pad_voxel * voxel_size
        voxel_origin[2] -= voxel_size
        voxel_num_xyz[2] += 1
        voxel_num_xyz[2] -= skip_top_voxel
        return voxel_num_xyz, voxel_origin

    def get_query_from_bbx(self, bbx, voxel_size, pad_voxel=0, skip_top_voxel=0):
        """reference: utils/mesher.py:138-180 (API form: materialises the [N,3] coordinates).  `bbx` is any
        object with get_min_bound()/get_max_bound() (an Open3D AABB in the reference) or a (min, max) pair."""
        if hasattr(bbx, "get_min_bound"):
            lo, hi = bbx.get_min_bound(), bbx.get_max_bound()
        else:
            lo, hi = bbx
        num, origin = self.grid_from_bounds(lo, hi, voxel_size, pad_voxel, skip_top_voxel)
        if int(num[0]) * int(num[1]) * int(num[2]) > 5e8:
            print("too many query points, use smaller chunks")
            return None, None, None
        ax = [torch.arange(int(k), dtype=torch.int16, device=self.cur_device) for k in num]
        x, y, z = torch.meshgrid(*ax, indexing="ij")
        coord = torch.stack((x.flatten(), y.flatten(), z.flatten())).transpose(0, 1).float()
        coord *= voxel_size
        coord += torch.tensor(origin, dtype=self.dtype, device=self.cur_device)
        return coord, num, origin

    def query_grid(self, min_bound, max_bound, voxel_size, pad_voxel=0, skip_top_voxel=0, query_locally=False,
                   mask_min_nn_count: int = 4, rank: int = 0, world: int = 1, chunk: int = 1 << 24):
        """Sharded dense-grid query: rank r evaluates the contiguous linear range [r*N/W, (r+1)*N/W) of the
        z-fastest grid.  Returns (sdf f32 [count] cuda, mask u8 [count] cuda, first, dims, origin)."""
        lib = P.load()
        npts = self.neural_points
        dims, origin = self.grid_from_bounds(min_bound, max_bound, voxel_size, pad_voxel, skip_top_voxel)
        total = int(dims[0]) * int(dims[1]) * int(dims[2])
        first = total * rank // world
        last = total * (rank + 1) // world
        count = last - first
        sdf = torch.empty(count, dtype=torch.float32, device=self.device)
        mask = torch.empty(count, dtype=torch.uint8, device=self.device)
        # the reference adds a float32 origin tensor (mesher.py:178)
        origin32 = (C.c_float * 3)(*[float(np.float32(v)) for v in origin])
        dims32 = (C.c_int32 * 3)(*[int(v) for v in dims])
        dummy = torch.zeros((1, 3), device=self.device)
        for off in range(0, count, chunk):
            c = min(chunk, count - off)
            a = npts._query_args(dummy, self.geo_decoder.view(), None, False, query_locally)
            a.n = c
            a.workspace = P.ptr(npts._workspace(c))
            P.check(lib.pinq_mesher_grid_query(a, origin32, dims32, float(voxel_size), first + off, c,
                                               int(mask_min_nn_count), sdf[off:].data_ptr(), mask[off:].data_ptr(),
                                               P.stream()))
        return sdf, mask, first, dims, origin

    # ---- iso-surface extraction (row f3 of SURVEY.md 8; reference: utils/mesher.py:310-337, 420-485) ----------------
    def extract_mesh(self, sdf: torch.Tensor, mask, dims, origin, voxel_size: float, level: float = 0.0,
                     allow_degenerate: bool = False, vertex_base: int = 0, x_first: int = 0):
        """Marching cubes on the device over a grid held in device memory (x-major, z fastest, as query_grid writes it).
        -> (verts float32 [V,3] world coordinates, faces int64 [T,3]) on the device.  Cells need all eight corners masked.
        `dims` describes the slab handed in; `x_first` = its first x layer in the whole grid whose origin is `origin`."""
        lib = P.load()
        dims32 = (C.c_int32 * 3)(*[int(v) for v in dims])
        n = int(dims[0]) * int(dims[1]) * int(dims[2])
        sdf = sdf.reshape(-1).contiguous()
        if sdf.numel() != n:
            raise P.PinqError("sdf grid does not match dims")
        m8 = None
        if mask is not None:
            m8 = mask.reshape(-1)
            m8 = (m8 if m8.dtype == torch.uint8 else (m8 != 0).to(torch.uint8)).contiguous()
        dev = sdf.device
        scratch = torch.empty(int(lib.pinq_mc_scratch_bytes(n)), dtype=torch.uint8, device=dev)
        counts = torch.zeros(2, dtype=torch.int32, device=dev)
        P.check(lib.pinq_mc_count(P.ptr(sdf), P.ptr(m8), dims32, float(level), P.ptr(scratch), P.ptr(counts), P.stream()))
        nv, nt = (int(v) for v in counts.tolist())          # one host read: the output sizes
        verts = torch.empty((nv, 3), dtype=torch.float32, device=dev)
        faces = torch.empty((nt, 3), dtype=torch.int32, device=dev)
        origin32 = (C.c_float * 3)(*[float(np.float32(v)) for v in origin])
        if nv or nt:
            P.check(lib.pinq_mc_emit(P.ptr(sdf), dims32, int(x_first), float(level), origin32, float(voxel_size), P.ptr(scratch),
                                     int(vertex_base), P.ptr(verts), P.ptr(faces), P.stream()))
        faces = faces.to(torch.int64)
        if not allow_degenerate and nt:  # skimage's allow_degenerate=False: zero-area triangles are dropped
            loc = faces - vertex_base
            a, b, c = (verts[loc[:, k]].double() for k in range(3))
            faces = faces[torch.linalg.cross(b - a, c - a).norm(dim=1) > 0]
        return verts, faces

    def mc_mesh(self, mc_sdf, mc_mask, voxel_size, mc_origin):
        """reference signature (utils/mesher.py:310-337): numpy grids [a,b,c] in, (verts float64 [V,3], faces [T,3]) out."""
        sdf = torch.as_tensor(np.ascontiguousarray(mc_sdf), dtype=torch.float32, device=self.device)
        mask = None if mc_mask is None else torch.as_tensor(np.ascontiguousarray(mc_mask) != 0, device=self.device)
        verts, faces = self.extract_mesh(sdf, mask, sdf.shape, np.asarray(mc_origin, dtype=np.float64).reshape(3), float(voxel_size))
        return verts.cpu().numpy().astype(np.float64), faces.cpu().numpy()

    def recon_grid_mesh(self, min_bound, max_bound, voxel_size, mask_min_nn_count: int = None, query_locally: bool = False,
                        rank: int = 0, world: int = 1, max_points: int = 1 << 28, level: float = 0.0):
        """Dense query + marching cubes of the slab of x-layers this rank owns (no collective: the slabs overlap by one layer of
        grid points, so every cell belongs to exactly one rank and boundary vertices are bit-identical on both sides).  Slabs
        larger than `max_points` grid points are walked in x-chunks.  -> (verts [V,3] f32, faces [T,3] i64, dims, origin)."""
        if mask_min_nn_count is None:
            mask_min_nn_count = self.config.mesh_min_nn
        lib = P.load()
        npts = self.neural_points
        dims, origin = self.grid_from_bounds(min_bound, max_bound, voxel_size)
        nx, ny, nz = (int(v) for v in dims)
        cells_x = nx - 1
        x0, x1 = cells_x * rank // world, cells_x * (rank + 1) // world      # this rank's cell layers [x0, x1)
        layer = ny * nz
        step = max(1, max_points // layer - 1)
        origin32 = (C.c_float * 3)(*[float(np.float32(v)) for v in origin])
        dims32 = (C.c_int32 * 3)(nx, ny, nz)
        dummy = torch.zeros((1, 3), device=self.device)
        all_v, all_f, vbase = [], [], 0
        for c0 in range(x0, x1, step):
            c1 = min(c0 + step, x1)
            count = (c1 - c0 + 1) * layer                                 # point layers c0 .. c1 inclusive
            sdf = torch.empty(count, dtype=torch.float32, device=self.device)
            mask = torch.empty(count, dtype=torch.uint8, device=self.device)
            a = npts._query_args(dummy, self.geo_decoder.view(), None, False, query_locally)
            for off in range(0, count, 1 << 24):
                c = min(1 << 24, count - off)
                a.n = c
                a.workspace = P.ptr(npts._workspace(c))
                P.check(lib.pinq_mesher_grid_query(a, origin32, dims32, float(voxel_size), c0 * layer + off, c, int(mask_min_nn_count),
                                                   sdf[off:].data_ptr(), mask[off:].data_ptr(), P.stream()))
            v, f = self.extract_mesh(sdf, mask, (c1 - c0 + 1, ny, nz), origin, voxel_size, level=level, vertex_base=vbase, x_first=c0)
            all_v.append(v)
            all_f.append(f)
            vbase += v.shape[0]
            del sdf, mask
        verts = torch.cat(all_v) if all_v else torch.empty((0, 3), device=self.device)
        faces = torch.cat(all_f) if all_f else torch.empty((0, 3), dtype=torch.int64, device=self.device)
        return verts, faces, dims, origin
