"""Multi-GPU plumbing for the two paths that shard; one process per GPU,
torch.distributed (NCCL over NVLink on the box, gloo in the CPU tests).

* data-parallel mapping: every rank holds a replica of the map and the decoder, takes a contiguous slice
  of the global batch, runs the fused forward/backward kernels on it and exchanges
    - sum  : d loss / d local_geo_features, the packed decoder gradients, the certainty increments
    - max  : local_point_ts_update
  after which every rank applies the identical Adam step (replicated optimiser, no parameter broadcast).
* mesher grid: ranks evaluate disjoint linear ranges of the grid -- no collective (utils/mesher.py here).

Nothing below touches CUDA directly, so the exchange logic is testable with gloo on CPU tensors.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world_info(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_range(n: int, rank: int, world: int):
    """Contiguous slice [first, last) of n items for `rank`; slices tile [0, n) exactly."""
    return n * rank // world, n * (rank + 1) // world


def decimated_indices(first: int, last: int, step: int, device=None):
    """Global indices j in [first, last) with j % step == 0, relative to `first` -- so that the union over ranks of
    shard[rel] equals batch[::step] of the un-sharded batch (reference: utils/mapper.py:484 `coord[::decimation]`)."""
    start = (-first) % step
    if start >= last - first:
        return torch.empty(0, dtype=torch.int64, device=device)
    return torch.arange(start, last - first, step, device=device)


def exchange_gradients(params, sum_buffers=(), max_buffers=(), group=None):
    """All-reduce(sum) of every p.grad (small tensors packed into one flat buffer, large ones in place) and
    of `sum_buffers`; all-reduce(max) of `max_buffers`.  Returns the number of collectives issued."""
    rank, world = world_info(group)
    if world == 1:
        return 0
    calls = 0
    small, large = [], []
    for p in params:
        if p.grad is None:
            continue
        (small if p.grad.numel() <= 65536 else large).append(p.grad)
    for g in large:
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
        calls += 1
    if small:
        flat = torch.cat([g.reshape(-1) for g in small])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        calls += 1
        off = 0
        for g in small:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
    for b in sum_buffers:
        dist.all_reduce(b, op=dist.ReduceOp.SUM, group=group)
        calls += 1
    for b in max_buffers:
        dist.all_reduce(b, op=dist.ReduceOp.MAX, group=group)
        calls += 1
    return calls


def dp_mapping_iteration(mapper, coord, sdf_label, ts, weight, opt, group=None):
    """One data-parallel iteration of Mapper.mapping() (reference: utils/mapper.py:443-573) over a GLOBAL batch
    that every rank holds (indices are drawn with a shared seed): slice -> fused fwd/bwd -> exchange -> Adam.
    The per-rank loss is normalised by the GLOBAL counts so that the summed gradients equal the single-GPU
    gradient of the mean losses."""
    from .utils.loss import sdf_bce_loss

    cfg = mapper.config
    npts = mapper.neural_points
    rank, world = world_info(group)
    n = coord.shape[0]
    first, last = shard_range(n, rank, world)
    c, lab, t, w = coord[first:last], sdf_label[first:last], ts[first:last], torch.abs(weight[first:last]).detach()
    cert_before = npts.local_point_certainties.clone()
    pred = npts.query_sdf(c, mapper.geo_mlp, t, True, True)["sdf"]
    loss = sdf_bce_loss(pred, lab, mapper.sdf_scale, w, cfg.loss_weight_on, bce_reduction="sum") / n
    if cfg.ekional_loss_on and cfg.weight_e > 0:
        rel = decimated_indices(first, last, cfg.gradient_decimation, device=c.device)
        n_dec_global = (n + cfg.gradient_decimation - 1) // cfg.gradient_decimation
        if rel.numel() > 0:
            g = mapper.get_numerical_gradient(c[rel], pred[rel], cfg.voxel_size_m * cfg.num_grad_step_ratio)
            loss = loss + cfg.weight_e * ((g.norm(2, dim=-1) - 1.0) ** 2).sum() / n_dec_global
    params = [p for grp in opt.param_groups for p in grp["params"]]
    opt.zero_grad(set_to_none=True)
    loss.backward()
    for p in params:  # a rank whose slice never touched a parameter still has to join the collective
        if p.requires_grad and p.grad is None and p.numel() > 0:
            p.grad = torch.zeros_like(p)
    delta = npts.local_point_certainties - cert_before
    exchange_gradients(params, sum_buffers=[delta], max_buffers=[npts.local_point_ts_update], group=group)
    npts.local_point_certainties = cert_before + delta
    opt.step()
    return loss.detach()
