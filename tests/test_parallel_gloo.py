"""CPU, world_size 2, gloo: the exchange step of the data-parallel mapping path (pin_slam_b200/parallel.py)."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pin_slam_b200.parallel import decimated_indices, exchange_gradients, shard_range
        torch.manual_seed(0)  # same "global batch" and parameters on both ranks
        feats = torch.nn.Parameter(torch.randn(100_000, 8))  # "large": reduced in place
        W = torch.nn.Parameter(torch.randn(64, 11))           # "small": packed
        b = torch.nn.Parameter(torch.randn(64))
        frozen = torch.nn.Parameter(torch.randn(3), requires_grad=False)
        x = torch.randn(4096, 11)
        idx = torch.randint(0, 100_000, (4096,))
        n = x.shape[0]

        def loss_of(sl, norm):
            h = torch.relu((x[sl] + torch.nn.functional.pad(feats[idx[sl]], (0, 3))) @ W.T + b)
            return h.sum() / norm

        first, last = shard_range(n, rank, world)
        loss_of(slice(first, last), n).backward()
        cert = torch.zeros(50)
        cert[rank * 10:(rank + 1) * 10] = 1.0 + rank
        ts = torch.full((50,), rank, dtype=torch.int64)
        calls = exchange_gradients([feats, W, b, frozen], sum_buffers=[cert], max_buffers=[ts])
        got = [feats.grad.clone(), W.grad.clone(), b.grad.clone()]
        for p in (feats, W, b):
            p.grad = None
        loss_of(slice(0, n), n).backward()  # single-process reference on the whole batch
        ok = all(torch.allclose(a, p.grad, rtol=1e-4, atol=1e-6) for a, p in zip(got, (feats, W, b)))
        ok = ok and calls == 4 and frozen.grad is None
        ok = ok and torch.equal(cert[:20], torch.cat([torch.ones(10), 2 * torch.ones(10)])) and int(ts.min()) == world - 1
        rel = decimated_indices(first, last, 10)
        ok = ok and all((first + int(j)) % 10 == 0 for j in rel)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_gradient_exchange_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]
