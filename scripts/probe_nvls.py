#!/usr/bin/env python3
"""Probe (torchrun, N GPUs): does torch's symmetric memory rendezvous work on this box, does it hand out an NVLS multicast
address, and what do torch's own multimem / two-shot all-reduce kernels take on a block of the bench's size (53 MB fp32)?"""
import json
import os
import sys

import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
out = {"world": world}
try:
    import torch.distributed._symmetric_memory as symm
    from torch._C._distributed_c10d import _SymmetricMemory as S
    out["has_multicast_support"] = bool(S.has_multicast_support(torch._C._autograd.DeviceType.CUDA, dev.index))
    n = 13_250_000
    t = symm.empty(n, dtype=torch.float32, device=dev)
    t.fill_(rank + 1.0)
    hdl = symm.rendezvous(t, dist.group.WORLD)
    out["multicast_ptr_nonzero"] = bool(hdl.multicast_ptr)
    out["buffer_ptrs"] = len(hdl.buffer_ptrs)
    out["signal_pad_size"] = int(hdl.signal_pad_size)
    gname = dist.group.WORLD.group_name

    def timed(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    for name, op in (("multimem_all_reduce_", lambda: torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname)),
                     ("two_shot_all_reduce_", lambda: torch.ops.symm_mem.two_shot_all_reduce_(t, "sum", gname)),
                     ("nccl_all_reduce", lambda: dist.all_reduce(t))):
        try:
            t.fill_(1e-3)
            out[name + "_ms"] = timed(op)
        except Exception as e:  # noqa: BLE001
            out[name + "_error"] = repr(e)[:300]
    t.fill_(rank + 1.0)
    torch.cuda.synchronize(); dist.barrier()
    try:
        torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname)
        torch.cuda.synchronize()
        out["multimem_sum_ok"] = bool(torch.all(t == world * (world + 1) / 2).item())
    except Exception as e:  # noqa: BLE001
        out["multimem_sum_error"] = repr(e)[:300]
except Exception as e:  # noqa: BLE001
    out["error"] = repr(e)[:600]
if rank == 0:
    print(json.dumps(out))
dist.barrier()
dist.destroy_process_group()
