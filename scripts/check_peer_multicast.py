#!/usr/bin/env python3
"""torchrun check (N >= 2 GPUs): the NVLS form of pinq_peer_exchange (multimem.ld_reduce / multimem.st through the multicast
address of torch's symmetric memory) against the P2P form of the same kernel and against NCCL: plain sum over three epochs,
then the fused Adam (parameters identical on every rank, equal to the P2P form's within 1e-6, gradients zeroed)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pin_slam_b200.parallel import PeerExchange, PeerLayout  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
res = {"world": world}
n_sum, n_par = 3_000_000 + 12, 40_000
g = torch.Generator(device=dev).manual_seed(100 + rank)
mc = PeerExchange(PeerLayout(n_sum, 0, n_par), dev, grid=32, cooperative=True, multicast=True, timeout_ms=2000.0)
pp = PeerExchange(PeerLayout(n_sum, 0, n_par), dev, grid=32, cooperative=True, multicast=False, timeout_ms=2000.0)
res["multicast"] = bool(mc.mc_block) and not pp.mc_block
worst = 0.0
for rep in range(3):
    x = torch.randn(mc.layout.n_sum, device=dev, generator=g)
    mc.sum_region().copy_(x); pp.sum_region().copy_(x)
    want = x.clone(); dist.all_reduce(want)
    torch.cuda.synchronize(); dist.barrier()
    mc.exchange(); pp.exchange()
    torch.cuda.synchronize(); dist.barrier()
    worst = max(worst, float((mc.sum_region() - pp.sum_region()).abs().max().item()), float((mc.sum_region() - want).abs().max().item()))
    if world == 2:
        assert torch.equal(mc.sum_region(), pp.sum_region()), "two ranks: a + b has one order"
res["sum_max_abs_diff"] = worst
assert worst < 1e-5, worst
# fused Adam
state = lambda: dict(exp_avg=torch.zeros(n_par, device=dev), exp_avg_sq=torch.zeros(n_par, device=dev), lr=0.01, beta1=0.9, beta2=0.99, eps=1e-15)
sa, sb = state(), state()
p0 = torch.randn(n_par, device=dev, generator=torch.Generator(device=dev).manual_seed(7))
mc.param_region()[:n_par].copy_(p0); pp.param_region()[:n_par].copy_(p0)
for step in range(1, 4):
    x = torch.randn(mc.layout.n_sum, device=dev, generator=g) * 1e-3
    mc.sum_region().copy_(x); pp.sum_region().copy_(x)
    torch.cuda.synchronize(); dist.barrier()
    mc.exchange(adam=dict(sa, step=step, n=n_par)); pp.exchange(adam=dict(sb, step=step, n=n_par))
    torch.cuda.synchronize(); dist.barrier()
    a, b = mc.param_region()[:n_par], pp.param_region()[:n_par]
    rel = float(((a - b).abs().max() / b.abs().max()).item())
    assert rel < 1e-6, rel
    allp = [torch.empty_like(a) for _ in range(world)]
    dist.all_gather(allp, a.contiguous())
    assert all(torch.equal(allp[0], q) for q in allp), "replicas must stay bit-identical"
    assert torch.all(mc.sum_region()[:n_par] == 0), "fused zero_grad"
    assert torch.allclose(mc.sum_region()[n_par:], pp.sum_region()[n_par:], atol=1e-6)
res["adam_rel_diff"] = rel
res["errors"] = [mc.error(), pp.error()]
assert res["errors"] == [0, 0]
# timing of the 53 MB block of the bench
big_n = 13_250_000
for name, flag in (("nvls", True), ("p2p", False)):
    ex = PeerExchange(PeerLayout(big_n), dev, grid=int(os.environ.get("PINQ_PEER_GRID", "0")), cooperative=True, multicast=flag, timeout_ms=2000.0)
    for _ in range(3):
        ex.exchange()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ex.exchange()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res["exchange_53MB_ms_" + name] = float(t.item())
    assert ex.error() == 0
    ex.close()
if rank == 0:
    print(json.dumps(res))
dist.barrier()
dist.destroy_process_group()
