#!/usr/bin/env python3
"""Host-side cost of one public-API step (query_sdf + backward): cProfile over 200 steps of 65 536 queries (GPU time negligible)."""
import cProfile, pstats, os, sys, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from pin_slam_b200 import synth
class A: pass
args = A(); args.surface_points = 2_000_000; args.side = 150.0
cfg, npts, dec = bench.build_map(args, "cuda:0")
n = 65536
q = torch.from_numpy(synth.microbench_queries(npts.neural_points.cpu().numpy(), n, seed=1)).cuda()
ts = torch.zeros(n, dtype=torch.int64, device="cuda")
ones = torch.ones(n, device="cuda")
params = [npts.local_geo_features] + list(dec.parameters())
def step():
    for p in params: p.grad = None
    out = npts.query_sdf(q, dec, ts, True, True)
    out["sdf"].backward(ones)
for _ in range(20): step()
torch.cuda.synchronize()
import time
t = time.perf_counter()
for _ in range(200): step()
t_host = (time.perf_counter() - t) / 200 * 1e3
torch.cuda.synchronize()
print("host ms per step (enqueue only):", round(t_host, 4))
pr = cProfile.Profile(); pr.enable()
for _ in range(200): step()
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28); print(s.getvalue()[:6000])
