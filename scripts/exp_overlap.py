#!/usr/bin/env python3
"""Experiment: forward of batch i+1 (bin + search + decode) on one stream under the backward of batch i on another."""
import ctypes, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from pin_slam_b200 import _pinq as P, synth

class A: pass
args = A(); args.surface_points = 9_000_000; args.side = 350.0
dev = "cuda:0"; torch.cuda.set_device(0)
lib = P.load()
cfg, npts, dec = bench.build_map(args, dev)
n = 4 * 1024 * 1024
q = torch.from_numpy(synth.microbench_queries(npts.neural_points.cpu().numpy(), n, seed=1)).to(dev)
ts = torch.zeros(n, dtype=torch.int64, device=dev)
ones = torch.ones(n, device=dev)
sets = []
for s in range(2):
    a = npts._query_args(q, dec.view(), ts, True, True)
    d = dict(sdf=torch.empty(n, device=dev), nnc=torch.empty(n, dtype=torch.int32, device=dev),
             stage=torch.empty(int(lib.pinq_stage_bytes(n)), dtype=torch.uint8, device=dev),
             relu=torch.empty(int(lib.pinq_relu_mask_words(n)), dtype=torch.int32, device=dev),
             perm=torch.empty(n, dtype=torch.int32, device=dev))
    a.sdf, a.nn_count, a.stage, a.relu_mask, a.perm = P.ptr(d["sdf"]), P.ptr(d["nnc"]), P.ptr(d["stage"]), P.ptr(d["relu"]), P.ptr(d["perm"])
    b = P.BwdArgs(); b.dec = dec.view(); b.n, b.flags, b.k, b.feature_dim = n, 0, 6, 8
    b.features = P.ptr(npts.local_geo_features.detach())
    b.stage, b.relu_mask, b.perm, b.grad_sdf = P.ptr(d["stage"]), P.ptr(d["relu"]), P.ptr(d["perm"]), P.ptr(ones)
    d["gf"] = torch.zeros_like(npts.local_geo_features); d["gd"] = [torch.zeros_like(p) for p in dec.fused_params()]
    b.grad_features = P.ptr(d["gf"]); b.grad_W1, b.grad_b1, b.grad_W2, b.grad_b2 = (P.ptr(g) for g in d["gd"])
    sets.append((a, b, d))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def seq(steps):
    st = torch.cuda.current_stream().cuda_stream
    for i in range(steps):
        a, b, _ = sets[i % 2]
        P.check(lib.pinq_query_fused_fwd(a, st)); P.check(lib.pinq_query_fused_bwd(b, st))
def pipe(steps):
    fdone = [torch.cuda.Event() for _ in range(2)]; bdone = [torch.cuda.Event() for _ in range(2)]
    for e in bdone: e.record(s2)
    for i in range(steps):
        a, b, _ = sets[i % 2]
        s1.wait_event(bdone[i % 2])           # the set's previous backward has read its stage blocks
        P.check(lib.pinq_query_fused_fwd(a, s1.cuda_stream)); fdone[i % 2].record(s1)
        s2.wait_event(fdone[i % 2])
        P.check(lib.pinq_query_fused_bwd(b, s2.cuda_stream)); bdone[i % 2].record(s2)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
res = {}
for name, fn in (("sequential", seq), ("pipelined", pipe)):
    fn(4); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    e0.record(); s1.wait_event(e0); s2.wait_event(e0)
    fn(20)
    e1.record(); torch.cuda.synchronize()
    res[name] = e0.elapsed_time(e1) / 20
print(json.dumps(res))
