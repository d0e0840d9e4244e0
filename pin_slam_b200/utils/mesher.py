"""Mesher -- dense SDF grid query (reference: utils/mesher.py:19-180).

Only the naturally sharded consumer of the hot path is in scope: `query_points` and
`get_query_from_bbx`.  `query_grid` is the B200 form of the two combined:
grid coordinates are generated inside the kernel from (origin, dims, voxel) -- no [N,3] tensor is
materialised -- and a linear range of the grid can be given so that ranks shard the grid without any
collective.  Marching cubes / Open3D post-processing are out of scope (row f3).
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
