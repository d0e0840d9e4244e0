"""Peer-memory exchange kernel (csrc/peer.cu, include/pinq.h pinq_peer_*): R "ranks" simulated on ONE GPU -- R peer blocks in
one process, R launches of the same kernel on R streams, the same flag barriers and P2P code path minus the link.  The sum is
checked bit-exactly against the same rank-ordered fp32 sum, the max bit-exactly, the fused Adam against torch.optim.Adam,
and a two-shard fused forward/backward + exchange against the un-sharded gradient (and, through it, against the oracle)."""
import pytest
import torch

from helpers import build_cuda_map, decoder_pair, frames_of_points, make_queries, randomize_features, build_oracle_map, rel_err

pytestmark = pytest.mark.gpu


def _peer(world, n_sum, n_max=0, n_param=0, grid=16, threads=0):
    from pin_slam_b200.parallel import PeerExchange, PeerLayout
    return PeerExchange(PeerLayout(n_sum, n_max, n_param), "cuda:0", simulate_world=world, grid=grid, timeout_ms=2000.0,
                        threads=threads)


@pytest.mark.parametrize("threads", [0, 128])  # 512-thread CTAs (stand-alone form) / 128-thread CTAs (overlapped form)
@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("n_sum,n_max", [(4, 2), (1000, 0), (262144 + 12, 4098), (3_000_000, 300_000)])
def test_allreduce_sum_and_max_bit_exact(world, n_sum, n_max, threads):
    peer = _peer(world, n_sum, n_max, grid=(8 if world == 8 else 16) * (4 if threads == 128 else 1), threads=threads)
    try:
        g = torch.Generator(device="cuda").manual_seed(world * 7 + n_sum)
        ins = [torch.randn(peer.layout.n_sum, device="cuda", generator=g) for _ in range(world)]
        mx = [torch.randint(-5, 1 << 40, (peer.layout.n_max,), device="cuda", generator=g) for _ in range(world)]
        for rep in range(3):  # epochs 1..3 over the same pads
            for r in range(world):
                peer.sum_region(r).copy_(ins[r] * (rep + 1))
                if n_max:
                    peer.max_region(r).copy_(mx[r] - rep)
            peer.exchange()
            want = ins[0] * (rep + 1)
            for r in range(1, world):
                want = want + ins[r] * (rep + 1)  # the kernel's order: ((b0 + b1) + b2) + ...
            for r in range(world):
                assert torch.equal(peer.sum_region(r), want), (world, r, rep)
                if n_max:
                    assert torch.equal(peer.max_region(r), torch.stack(mx).max(0).values - rep)
        assert peer.error() == 0
    finally:
        peer.close()


def test_prefix_exchange_leaves_the_rest_alone():
    peer = _peer(2, 4096, 512)
    try:
        for r in range(2):
            peer.sum_region(r).fill_(r + 1.0)
            peer.max_region(r).fill_(r + 10)
        peer.exchange(n_sum=1024, n_max=100)
        for r in range(2):
            assert torch.all(peer.sum_region(r)[:1024] == 3.0) and torch.all(peer.sum_region(r)[1024:] == r + 1.0)
            assert torch.all(peer.max_region(r)[:100] == 11) and torch.all(peer.max_region(r)[100:] == r + 10)
        assert peer.error() == 0
    finally:
        peer.close()


@pytest.mark.parametrize("threads", [0, 128])
@pytest.mark.parametrize("world", [2, 4])
def test_fused_adam_matches_torch_adam(world, threads):
    n_par, n_extra = 40000, 1000
    peer = _peer(world, n_par + n_extra, 0, n_par, threads=threads)
    try:
        g = torch.Generator(device="cuda").manual_seed(11)
        p0 = torch.randn(n_par, device="cuda", generator=g)
        ref = torch.nn.Parameter(p0.clone())
        opt = torch.optim.Adam([ref], lr=0.01, betas=(0.9, 0.99), eps=1e-15)  # utils/tools.py:112, config.py defaults
        state = {"exp_avg": [torch.zeros(n_par, device="cuda") for _ in range(world)],
                 "exp_avg_sq": [torch.zeros(n_par, device="cuda") for _ in range(world)],
                 "lr": 0.01, "beta1": 0.9, "beta2": 0.99, "eps": 1e-15}
        for r in range(world):
            peer.param_region(r).copy_(p0)
        for step in range(1, 4):
            grads = [torch.randn(peer.layout.n_sum, device="cuda", generator=g) * 1e-3 for _ in range(world)]
            for r in range(world):
                peer.sum_region(r).copy_(grads[r])
            peer.exchange(adam=dict(state, step=step))
            tot = grads[0].clone()
            for r in range(1, world):
                tot += grads[r]
            ref.grad = tot[:n_par].clone()
            opt.step()
            for r in range(world):
                assert rel_err(peer.param_region(r), ref.detach()) < 1e-6
                assert torch.equal(peer.param_region(r), peer.param_region(0))      # replicas stay bit-identical
                assert torch.all(peer.sum_region(r)[:n_par] == 0)                   # fused zero_grad
                assert torch.equal(peer.sum_region(r)[n_par:], tot[n_par:])         # plain tail of the sum region
        assert peer.error() == 0
    finally:
        peer.close()


def test_barrier_timeout_sets_the_error_word_instead_of_hanging():
    """Rank 1 never launches: rank 0's kernel must give up after timeout_ms and flag the error word."""
    from pin_slam_b200 import _pinq as P
    from pin_slam_b200.parallel import PeerExchange, PeerLayout
    peer = PeerExchange(PeerLayout(1024), "cuda:0", simulate_world=2, grid=4, timeout_ms=20.0)
    try:
        P.check(peer.lib.pinq_peer_exchange(peer.make_args(0, 1, 1024, 0), P.stream()))
        torch.cuda.synchronize()
        assert peer.error() != 0
    finally:
        peer.close()


def test_two_shard_backward_plus_exchange_equals_unsharded_gradient():
    from pin_slam_b200.parallel import MappingPeer, PeerExchange
    frames = frames_of_points(3, 40000)
    s = build_oracle_map(frames)
    m, cfg = build_cuda_map(frames)
    randomize_features(s, m)
    dec, _ = decoder_pair(cfg)
    q = make_queries(s.points, 8000).cuda()
    g = torch.Generator().manual_seed(5)
    up = torch.randn(q.shape[0], generator=g).cuda()
    ts = torch.randint(0, 3, (q.shape[0],), generator=g).cuda()
    cert0, ts0 = m.local_point_certainties.clone(), m.local_point_ts_update.clone()
    params = [m.local_geo_features] + list(dec.fused_params())
    try:
        for p in params:
            p.grad = None
        out = m.query_sdf(q, dec, ts, True, True)
        (out["sdf"] * up).sum().backward()
        want = [p.grad.clone() for p in params]
        for p in params:
            p.grad = None
        rows, F = m.local_geo_features.shape
        world = 2
        peer = PeerExchange(MappingPeer.capacity(rows, F, dec), "cuda:0", simulate_world=world, grid=16, timeout_ms=2000.0)
        try:
            half = q.shape[0] // 2
            carves = [MappingPeer(peer, rows, F, dec, r=r) for r in range(world)]
            for r, (a, b) in enumerate(((0, half), (half, q.shape[0]))):
                carves[r].attach(m, dec)
                o = m.query_sdf(q[a:b], dec, ts[a:b], True, True)
                (o["sdf"] * up[a:b]).sum().backward()
                MappingPeer.detach(m, dec)
            assert all(p.grad is None for p in params)  # the gradients went to the sinks
            peer.exchange(n_sum=carves[0].n_sum, n_max=0)
            for r in range(world):
                assert torch.equal(carves[r].grad_features, carves[0].grad_features)
                assert rel_err(carves[r].grad_features, want[0]) < 1e-5
                for got, w in zip(carves[r].grad_decoder, want[1:]):
                    assert rel_err(got, w) < 1e-5
            assert peer.error() == 0
        finally:
            peer.close()
    finally:
        m.local_point_certainties, m.local_point_ts_update = cert0, ts0
        m.feature_grad_sink = None
        dec.grad_sink = None


def _replica(frames, seed_pool=31, pool_n=12000):
    """One rank's replica: map, decoder and a mapper with a replay pool -- identical on every call."""
    from pin_slam_b200.utils.mapper import Mapper
    s = build_oracle_map(frames)
    m, cfg = build_cuda_map(frames)
    randomize_features(s, m, std=0.3)
    dec, _ = decoder_pair(cfg)
    cfg.bs, cfg.bs_new_sample = 2048, 256
    g = torch.Generator().manual_seed(seed_pool)
    mp = Mapper(cfg, None, m, dec, None, None)
    coord = make_queries(s.points, pool_n, seed=seed_pool).cuda()
    mp.global_coord_pool, mp.coord_pool = coord, coord.clone()
    mp.sdf_label_pool = (0.2 * torch.randn(pool_n, generator=g)).cuda()
    w = torch.rand(pool_n, generator=g) + 0.5
    w[::3] *= -1
    mp.weight_pool = w.cuda()
    mp.time_pool = torch.randint(0, 3, (pool_n,), generator=g).cuda()
    mp.pool_sample_count = pool_n
    mp.new_idx = torch.arange(pool_n - 900, pool_n, device="cuda")
    return s, m, cfg, dec, mp


@pytest.mark.parametrize("world,iters", [(2, 1), (2, 5), (4, 3)])
def test_data_parallel_fused_mapping_matches_single_device(world, iters):
    """pinq_mapping_run in its data-parallel form (sharded global batch, global loss normalisation, pinq_peer_exchange with
    the fused Adam between backward and the next iteration; certainty / ts exchange at the end) on `world` simulated ranks
    -- `world` replicas of the map on one GPU, one stream each -- against the single-device loop on the same draws."""
    from pin_slam_b200.parallel import DataParallelMapping, MappingPeer, PeerExchange
    frames = frames_of_points(3, 40000)
    s, m_ref, cfg, dec_ref, mp_ref = _replica(frames)
    probe = make_queries(s.points, 4000, seed=77).cuda()
    start = m_ref.local_geo_features.data.clone()
    mp_ref.generator = torch.Generator().manual_seed(5)
    mp_ref._mapping_fused(iters)
    torch.cuda.synchronize()
    with torch.no_grad():
        sdf_ref = m_ref.query_sdf(probe, dec_ref, None, False, True)["sdf"].clone()

    reps = [_replica(frames) for _ in range(world)]
    rows, fdim = reps[0][1].local_geo_features.shape
    peer = PeerExchange(MappingPeer.capacity(rows, fdim, reps[0][3], with_params=True), "cuda:0", simulate_world=world,
                        grid=8, timeout_ms=3000.0)
    try:
        torch.cuda.synchronize()
        streams = [torch.cuda.Stream() for _ in range(world)]
        for r, (_, m, _, dec, mp) in enumerate(reps):   # everything below is asynchronous: the "ranks" meet on the device
            mp.generator = torch.Generator().manual_seed(5)
            mp.dp = DataParallelMapping(peer, rank=r)
            with torch.cuda.stream(streams[r]):
                mp._mapping_fused(iters)
        torch.cuda.synchronize()
        assert peer.error() == 0
        f_ref = m_ref.local_geo_features.data
        moved = (f_ref - start).abs() > 0
        assert float(moved.float().mean()) > 0.01
        for r, (_, m, _, dec, mp) in enumerate(reps):
            f = m.local_geo_features.data
            assert torch.equal(f, reps[0][1].local_geo_features.data)          # replicas stay bit-identical
            for a, b in zip(dec.parameters(), reps[0][3].parameters()):
                assert torch.equal(a.data, b.data)
            # Adam (eps 1e-15) moves an element by ~lr * sign(g) when g is rounding noise: compare like the autograd-loop test
            assert float(((f - f_ref).abs() > 1e-4)[moved].float().mean()) < 0.02
            for a, b in zip(dec.parameters(), dec_ref.parameters()):
                assert float((a.data - b.data).abs().max()) < 1e-3
            with torch.no_grad():
                sdf = m.query_sdf(probe, dec, None, False, True)["sdf"]
            assert float((sdf - sdf_ref).abs().mean()) < 1e-4 * max(1.0, float(sdf_ref.abs().mean()))
            assert rel_err(m.local_point_certainties, m_ref.local_point_certainties) < 1e-5
            assert torch.equal(m.local_point_ts_update, m_ref.local_point_ts_update)
            assert f.data_ptr() != peer.param_region(r).data_ptr()             # parameters are back in torch-owned memory
        # the logged losses are per-rank partial sums of the global means
        loss_dp = sum(mp.last_loss for (_, _, _, _, mp) in reps)
        assert rel_err(loss_dp, mp_ref.last_loss) < 1e-4
    finally:
        peer.close()
