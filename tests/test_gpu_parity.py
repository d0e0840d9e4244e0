"""CUDA path vs the CPU oracle on identical seeded inputs (through the C-ABI).  Bit-exact for slots,
neighbour indices and counts; 1e-5 relative for SDF and gradients (BASELINE.json north_star)."""
import pytest
import torch

from helpers import (O, build_cuda_map, build_oracle_map, decoder_pair, frames_of_points, make_queries,
                     randomize_features, rel_err)

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def maps():
    frames = frames_of_points(3, 40000)
    s = build_oracle_map(frames)
    m, cfg = build_cuda_map(frames)
    randomize_features(s, m)
    dec, d = decoder_pair(cfg)
    q = make_queries(s.points, 6000)
    return s, m, cfg, dec, d, q


def test_voxel_downsample_bit_exact():
    from pin_slam_b200.utils.tools import voxel_down_sample_torch, voxel_down_sample_min_value_torch

    pts, _ = frames_of_points(1, 50000, seed=7)[0]
    for vs in (0.4, 0.08, 0.8):
        ref = O.voxel_down_sample(pts, vs)
        got = voxel_down_sample_torch(pts.cuda(), vs).cpu()
        assert torch.equal(ref, got), vs
    val = torch.rand(pts.shape[0]) * 30
    ref = O.voxel_down_sample_min_value(pts, 0.4, val)
    got = voxel_down_sample_min_value_torch(pts.cuda(), 0.4, val.cuda()).cpu()
    assert torch.equal(ref, got)


def test_map_update_slots_bit_exact(maps):
    s, m, *_ = maps
    assert m.count() == s.points.shape[0]
    assert torch.equal(m.neural_points.cpu(), s.points)
    assert torch.equal(m.buffer_pt_index.cpu(), s.table)
    assert torch.equal(m.point_ts_create.cpu(), s.ts_create)
    assert torch.equal(m.global2local.cpu(), s.global2local)
    assert torch.equal(m.local_mask.cpu(), s.local_mask)
    assert torch.equal(m.local_neural_points.cpu(), s.l_points)


@pytest.mark.parametrize("time_filtering", [False, True])
def test_radius_search_bit_exact(maps, time_filtering):
    s, m, _, _, _, q = maps
    d2_ref, idx_ref = O.radius_search(s, q, time_filtering)
    d2, idx = m.radius_neighborhood_search(q.cuda(), time_filtering)
    assert idx.dtype == torch.int64 and d2.dtype == torch.float32
    assert torch.equal(idx.cpu(), idx_ref)
    assert torch.equal(d2.cpu(), d2_ref)


def test_query_certainty(maps):
    s, m, _, _, _, q = maps
    O.set_search_neighborhood(s, 1, 0.0)
    m.set_search_neighborhood(1, 0.0)
    try:
        ref = O.query_certainty(s, q)
        got = m.query_certainty(q.cuda()).cpu()
    finally:
        O.set_search_neighborhood(s, 2, 0.5)
        m.set_search_neighborhood(2, 0.5)
    assert torch.equal(ref, got)


@pytest.mark.parametrize("locally", [True, False])
def test_query_feature_unfused(maps, locally):
    s, m, _, _, _, q = maps
    fv_r, w_r, nn_r, qc_r, knn_r = O.query_feature(s, q, None, False, locally)
    fv, _, w, nn, qc = m.query_feature(q.cuda(), training_mode=False, query_locally=locally)
    assert torch.equal(nn.cpu(), nn_r)
    assert fv.shape == fv_r.shape and w.shape == w_r.shape
    assert torch.equal(fv.cpu(), fv_r)          # gathers + one subtraction: exact
    assert rel_err(w, w_r) < 1e-6
    assert rel_err(qc, qc_r) < 1e-6


@pytest.mark.parametrize("locally", [True, False])
def test_fused_forward_tracking_flavour(maps, locally):
    s, m, _, dec, d, q = maps
    ref = O.query_sdf(s, d, q, None, False, locally, want_grad=True, want_std=True)
    with torch.no_grad():
        out = m.query_sdf(q.cuda(), dec, None, False, locally, want_grad=True, want_std=True, want_certainty=True)
    assert torch.equal(out["nn_count"].cpu().long(), ref["nn"])
    assert rel_err(out["sdf"], ref["sdf"]) < TOL
    assert rel_err(out["grad"], ref["grad"]) < TOL
    assert rel_err(out["std"], ref["std"]) < 1e-4   # sqrt of a cancelling sum: looser, not a north_star output
    assert rel_err(out["certainty"], ref["certainty"]) < TOL
    # element-wise with an absolute floor for near-zero SDF
    err = (out["sdf"].cpu() - ref["sdf"]).abs()
    assert float((err / (ref["sdf"].abs() + 1e-3)).max()) < 1e-4


def test_knn_indices_bit_exact(maps):
    s, m, _, dec, d, q = maps
    from pin_slam_b200 import _pinq as P
    _, _, _, _, knn_ref = O.query_feature(s, q, None, False, True)
    x = q.cuda()
    a = m._query_args(x, dec.view(), None, False, True)
    knn = torch.empty((x.shape[0], 6), dtype=torch.int32, device="cuda")
    sdf = torch.empty(x.shape[0], device="cuda")
    a.knn_idx, a.sdf = P.ptr(knn), P.ptr(sdf)
    P.check(P.load().pinq_query_fused_fwd(a, P.stream()))
    assert torch.equal(knn.cpu().long(), knn_ref)


def test_fused_backward_mapping_flavour(maps):
    s, m, _, dec, d, q = maps
    g = torch.Generator().manual_seed(5)
    up = torch.randn(q.shape[0], generator=g)
    ts = torch.randint(0, 3, (q.shape[0],), generator=g)
    # oracle: autograd through the restated reference path
    s2 = build_state_copy(s)
    s2.l_feats.requires_grad_(True)
    dp = O.DecoderParams(*[t.clone().requires_grad_(True) for t in d.tensors()], sdf_scale=d.sdf_scale)
    ref = O.query_sdf(s2, dp, q, ts, True, True)
    (ref["sdf"] * up).sum().backward()
    # CUDA
    cert0 = m.local_point_certainties.clone()
    ts0 = m.local_point_ts_update.clone()
    for p in dec.parameters():
        p.grad = None
    m.local_geo_features.grad = None
    out = m.query_sdf(q.cuda(), dec, ts.cuda(), True, True)
    (out["sdf"] * up.cuda()).sum().backward()
    try:
        assert rel_err(out["sdf"], ref["sdf"]) < TOL
        assert rel_err(m.local_geo_features.grad, s2.l_feats.grad) < TOL
        assert rel_err(dec.layers[0].weight.grad, dp.W1.grad) < TOL
        assert rel_err(dec.layers[0].bias.grad, dp.b1.grad) < TOL
        assert rel_err(dec.lout.weight.grad, dp.W2.grad) < TOL
        assert rel_err(dec.lout.bias.grad, dp.b2.grad) < TOL
        # training side effects (hazard H9): certainty += w (fp32 atomics -> tolerance), ts amax exact
        assert rel_err(m.local_point_certainties, s2.l_certainty) < TOL
        assert torch.equal(m.local_point_ts_update.cpu(), s2.l_ts_update)
    finally:
        m.local_point_certainties = cert0
        m.local_point_ts_update = ts0


def build_state_copy(s):
    import copy
    s2 = copy.copy(s)
    for name in ("l_feats", "l_certainty", "l_ts_update", "certainty", "feats"):
        setattr(s2, name, getattr(s, name).detach().clone())
    return s2


def test_empty_and_tiny_batches(maps):
    s, m, _, dec, d, q = maps
    with torch.no_grad():
        out = m.query_sdf(torch.empty((0, 3), device="cuda"), dec, None, False, True, want_grad=True)
    assert out["sdf"].shape == (0,)
    one = q[:1]
    ref = O.query_sdf(s, d, one, None, False, True)
    with torch.no_grad():
        out = m.query_sdf(one.cuda(), dec, None, False, True)
    assert rel_err(out["sdf"], ref["sdf"]) < TOL
    far = torch.full((5, 3), 1e4)
    with torch.no_grad():
        out = m.query_sdf(far.cuda(), dec, None, False, True, want_grad=True, want_std=True)
    assert int(out["nn_count"].sum()) == 0 and float(out["sdf"].abs().sum()) == 0.0


def test_cpu_tensor_is_refused(maps):
    _, m, _, dec, _, q = maps
    from pin_slam_b200 import _pinq as P
    with pytest.raises(P.PinqError):
        m.query_sdf(q, dec, None, False, True)


def test_binned_processing_order_changes_nothing(maps):
    """Spatial binning (csrc/sort.cu) only permutes the processing order: per-query outputs are bit-identical
    to the unbinned run, the training side effects and all gradients agree to atomics-order tolerance."""
    s, m, cfg, dec, d, _ = maps
    q = make_queries(s.points, 50000, seed=11).cuda()  # ragged: not a multiple of 32
    ts = torch.full((q.shape[0],), 2, dtype=torch.int64, device="cuda")
    res = {}
    cert0 = m.local_point_certainties.clone()
    tsu0 = m.local_point_ts_update.clone()
    for name, thr in (("plain", 1 << 60), ("binned", 0)):
        m.bin_queries_min = thr
        m.local_point_certainties = cert0.clone()
        m.local_point_ts_update = tsu0.clone()
        m.local_geo_features.grad = None
        for p_ in dec.parameters():
            p_.grad = None
        out = m.query_sdf(q, dec, ts, True, True, want_grad=True, want_std=True, want_certainty=True)
        (out["sdf"] * torch.linspace(0.5, 1.5, q.shape[0], device="cuda")).sum().backward()
        res[name] = dict(out=out, gf=m.local_geo_features.grad.clone(), gd=[p_.grad.clone() for p_ in dec.parameters()],
                         cert=m.local_point_certainties.clone(), tsu=m.local_point_ts_update.clone())
    m.bin_queries_min = 32768
    m.local_point_certainties, m.local_point_ts_update = cert0, tsu0
    a, b = res["plain"], res["binned"]
    for key in ("sdf", "nn_count", "grad", "std", "certainty"):
        assert torch.equal(a["out"][key], b["out"][key]), key
    assert torch.equal(a["tsu"], b["tsu"])
    assert rel_err(b["cert"], a["cert"]) < 1e-6
    assert rel_err(b["gf"], a["gf"]) < 1e-6
    for x, y in zip(a["gd"], b["gd"]):
        assert rel_err(y, x) < 1e-5


def test_brick_index_equals_slot_probing(maps):
    """The brick index (csrc/brick.cuh) is a read view of the same table: the fused query through it returns
    bit-identical per-query results to the 81-probe slot path, for local and global queries."""
    s, m, cfg, dec, d, _ = maps
    q = make_queries(s.points, 40000, seed=5).cuda()
    assert m._brick_safe(), "run_kitti buffer_size has no near collisions"
    res = {}
    for use in (False, True):
        m.use_brick_index = use
        m._cand_dirty = True
        for local in (True, False):
            with torch.no_grad():
                out = m.query_sdf(q, dec, None, False, local, want_grad=True, want_std=True, want_certainty=True)
            view = m._map_view(local)
            assert bool(view.brick_tab) == use
            res[(use, local)] = out
    m.check_index_invariants()
    for local in (True, False):
        a, b = res[(False, local)], res[(True, local)]
        for key in ("sdf", "nn_count", "grad", "std", "certainty"):
            assert torch.equal(a[key], b[key]), (local, key)
    # every local slot record is present exactly once in brick order
    n_occ = int(m._occ_count.item())
    assert int(m._brick_counters[0].item()) == n_occ
    assert torch.equal(torch.sort(m._cand_global[:n_occ, 3].view(torch.int32)).values,
                       torch.sort(m._brick_global[:n_occ, 3].view(torch.int32)).values)


def test_brick_rank_limit_falls_back_to_slot_probing(maps):
    """Maps with 2^24 points or more do not fit the brick search's 24-bit record rank: NeuralPoints then serves the same queries
    through the slot-probing search.  The limit is forced down to exercise that switch on a small map: no brick view is built,
    results are bit-identical to the brick path, and the view comes back when the limit is restored."""
    s, m, cfg, dec, d, _ = maps
    q = make_queries(s.points, 20000, seed=9).cuda()
    with torch.no_grad():
        ref = m.query_sdf(q, dec, None, False, True, want_grad=True, want_std=True)
    assert bool(m._map_view(True).brick_tab)
    old = m.brick_max_points
    m.brick_max_points = 1000            # < the map's point count
    m._cand_dirty = True
    try:
        with torch.no_grad():
            out = m.query_sdf(q, dec, None, False, True, want_grad=True, want_std=True)
        assert not bool(m._map_view(True).brick_tab)
        for key in ("sdf", "nn_count", "grad", "std"):
            assert torch.equal(out[key], ref[key]), key
    finally:
        m.brick_max_points = old
        m._cand_dirty = True
    with torch.no_grad():
        m.query_sdf(q[:100], dec, None, False, True)
    assert bool(m._map_view(True).brick_tab)


def test_skipped_distance_test_changes_nothing(maps):
    """PINQ_F_STENCIL_IN_RANGE: with run_kitti.yaml's stencil no candidate of a stencil cell can be farther than
    max_valid_dist2 (17 x 0.16 m^2 < 4.32 m^2), so the brick search compiles the `dist2 > max_valid_dist2` rejection
    (model/neural_points.py:499) out.  Same neighbours, counts and sdf with the comparison kept -- incl. far queries whose
    candidates sit in the outermost stencil cells -- and the oracle's neighbours."""
    s, m, _, dec, d, _ = maps
    from pin_slam_b200 import _pinq as P
    assert m._stencil_in_range
    pts = s.points
    lo, hi = pts.min(0).values, pts.max(0).values
    g = torch.Generator().manual_seed(31)
    far = lo + (hi - lo) * torch.rand((30000, 3), generator=g)
    q = torch.cat([make_queries(s.points, 60000, seed=37), far]).cuda()
    res = {}
    try:
        for skip in (True, False):
            m.skip_in_range_test = skip
            a = m._query_args(q, dec.view(), None, False, True)
            assert bool(a.flags & P.F_STENCIL_IN_RANGE) == skip
            knn = torch.empty((q.shape[0], 6), dtype=torch.int32, device="cuda")
            sdf = torch.empty(q.shape[0], device="cuda")
            nn = torch.empty(q.shape[0], dtype=torch.int32, device="cuda")
            a.knn_idx, a.sdf, a.nn_count = P.ptr(knn), P.ptr(sdf), P.ptr(nn)
            P.check(P.load().pinq_query_fused_fwd(a, P.stream()))
            torch.cuda.synchronize()
            res[skip] = (knn, sdf, nn)
    finally:
        m.skip_in_range_test = True
    for x, y in zip(res[True], res[False]):
        assert torch.equal(x, y)
    _, _, nn_ref, _, knn_ref = O.query_feature(s, q.cpu(), None, False, True)
    assert torch.equal(res[True][0].cpu().long(), knn_ref)
    assert torch.equal(res[True][2].cpu().long(), nn_ref.long())


def test_search_clears_the_gradient_block_of_its_backward(maps):
    """pinq_query_args.zero_fill: the fused forward clears the feature-gradient block its backward accumulates into (one store
    per thread of the neighbour search; a memset when the batch is smaller than the block or takes the slot-probing search).
    The block is handed out by the allocator right after a poisoned tensor of the same size was freed: a store the search
    skipped would show up as 1e3 in the gradient.  Same gradients as with the memset of `torch.zeros_like`, and the oracle's."""
    s, m, _, dec, d, _ = maps
    from pin_slam_b200 import _pinq as P
    rows = m.local_geo_features.shape[0]
    n = max(4 * rows, 50000)
    q = make_queries(s.points, n, seed=43)
    up = torch.randn(n, generator=torch.Generator().manual_seed(6))
    # the C-ABI contract, both branches: search stores (n >= bytes / 16) and memset (n < bytes / 16)
    for nq in (n, 1000):
        buf = torch.full((rows, m.local_geo_features.shape[1]), 7.0, device="cuda")
        a = m._query_args(q[:nq].cuda(), dec.view(), None, False, True)
        sdf = torch.empty(nq, device="cuda")
        nn = torch.empty(nq, dtype=torch.int32, device="cuda")
        a.sdf, a.nn_count = P.ptr(sdf), P.ptr(nn)
        a.zero_fill, a.zero_fill_bytes = P.ptr(buf), buf.numel() * 4
        P.check(P.load().pinq_query_fused_fwd(a, P.stream()))
        torch.cuda.synchronize()
        assert torch.all(buf == 0), nq
    a.zero_fill_bytes = 8
    assert P.load().pinq_query_fused_fwd(a, P.stream()) == -1  # PINQ_ERR_BAD_ARG: size not a multiple of 16
    # the autograd path
    grads = {}
    cert0, ts0 = m.local_point_certainties.clone(), m.local_point_ts_update.clone()
    try:
        for flag in (True, False):
            m.zero_grad_in_search = flag
            m.local_geo_features.grad = None
            for p_ in dec.parameters():
                p_.grad = None
            junk = torch.full_like(m.local_geo_features, 1e3)
            del junk
            out = m.query_sdf(q.cuda(), dec, None, True, True)
            (out["sdf"] * up.cuda()).sum().backward()
            grads[flag] = m.local_geo_features.grad.clone()
    finally:
        m.zero_grad_in_search = True
        m.local_point_certainties, m.local_point_ts_update = cert0, ts0
    assert float(grads[True].abs().max()) < 100.0
    assert rel_err(grads[True], grads[False]) < 1e-5
    s2 = build_state_copy(s)
    s2.l_feats.requires_grad_(True)
    ref = O.query_sdf(s2, d, q[:20000], None, False, True)
    m.local_geo_features.grad = None
    out = m.query_sdf(q[:20000].cuda(), dec, None, False, True)   # 20 000 < 2 x rows or not: either branch must match the oracle
    (out["sdf"] * up[:20000].cuda()).sum().backward()
    (ref["sdf"] * up[:20000]).sum().backward()
    assert rel_err(m.local_geo_features.grad, s2.l_feats.grad) < TOL


def test_truncated_key_search_equals_exact_keys(maps):
    """Brick search fast pass (32-bit keys: 23 bits of d^2 | walk identity, csrc/query.cu) against the 64-bit-key pass
    forced for every query (PINQ_F_EXACT_KEYS): same K neighbours in the same order, also for queries built to tie
    (midpoints of two neural points, copies of neural points, a query repeated many times)."""
    s, m, _, dec, d, _ = maps
    from pin_slam_b200 import _pinq as P
    pts = s.points
    g = torch.Generator().manual_seed(23)
    i = torch.randint(0, pts.shape[0], (20000,), generator=g)
    j = torch.randint(0, pts.shape[0], (20000,), generator=g)
    near = (pts[i] - pts[j]).norm(dim=1) < 0.8
    mid = ((pts[i] + pts[j]) * 0.5)[near]  # exact d^2 ties whenever the halving is exact
    q = torch.cat([make_queries(s.points, 60000, seed=29), mid, pts[i[:5000]], pts[i[:1]].repeat(4000, 1)]).cuda()
    res = {}
    try:
        for exact in (False, True):
            m.exact_search_keys = exact
            a = m._query_args(q, dec.view(), None, False, True)
            knn = torch.empty((q.shape[0], 6), dtype=torch.int32, device="cuda")
            sdf = torch.empty(q.shape[0], device="cuda")
            a.knn_idx, a.sdf = P.ptr(knn), P.ptr(sdf)
            P.check(P.load().pinq_query_fused_fwd(a, P.stream()))
            torch.cuda.synchronize()
            res[exact] = (knn, sdf)
    finally:
        m.exact_search_keys = False
    assert torch.equal(res[False][0], res[True][0])
    assert torch.equal(res[False][1], res[True][1])
    _, _, _, _, knn_ref = O.query_feature(s, q.cpu(), None, False, True)
    assert torch.equal(res[False][0].cpu().long(), knn_ref)


@pytest.mark.parametrize("n", [1, 33, 127, 129, 6000])
def test_tcgen05_kernels_match_mma_sync(maps, n):
    """The tcgen05 / TMEM decode and backward (csrc/decode_tc.cuh, bwd_tc.cuh) against the mma.sync kernels on ragged
    batch sizes (partial warps, partial 128-row tiles): same neighbours, sdf and every gradient to rounding.

    Two upstream gradients.  |up| (no cancellation): plain 1e-5 relative.  Signed up = linspace(-1, 1): the decoder gradients
    are then CANCELLING sums over all rows, whose fp32 error is bounded relative to the sum of the terms' magnitudes, not to
    the (small) result -- on this very input the reference's own arithmetic is 1.6e-5 (W2) / 7e-6 (b2) away from its fp64
    evaluation and either kernel 2e-6 .. 1.3e-5, varying with the order of the atomics (profiles/r3_tc_vs_mma_cancellation.txt).
    The signed case is therefore gated at 1e-5 of max(|result|, |result of the |up| pass|) -- the backward-error form."""
    s, m, cfg, dec, d, _ = maps
    q = make_queries(s.points, max(n, 8), seed=17)[:n].cuda()
    ts = torch.full((n,), 1, dtype=torch.int64, device="cuda")
    up = torch.linspace(-1.0, 1.0, n, device="cuda") if n > 1 else torch.ones(1, device="cuda")
    cert0 = m.local_point_certainties.clone()
    tsu0 = m.local_point_ts_update.clone()
    res = {}
    try:
        for sign, upstream in (("signed", up), ("abs", up.abs())):
            for name, flag in (("tc", False), ("mma", True)):
                m.use_mma_sync = flag
                m.local_point_certainties = cert0.clone()
                m.local_point_ts_update = tsu0.clone()
                m.local_geo_features.grad = None
                for p_ in dec.parameters():
                    p_.grad = None
                out = m.query_sdf(q, dec, ts, True, True, want_std=True)
                (out["sdf"] * upstream).sum().backward()
                res[(sign, name)] = dict(sdf=out["sdf"].detach().clone(), std=out["std"].clone(), nn=out["nn_count"].clone(),
                                         gf=m.local_geo_features.grad.clone(), gd=[p_.grad.clone() for p_ in dec.parameters()])
    finally:
        m.use_mma_sync = False
        m.local_point_certainties, m.local_point_ts_update = cert0, tsu0
    a, b = res[("abs", "mma")], res[("abs", "tc")]
    assert torch.equal(a["nn"], b["nn"])
    assert rel_err(b["sdf"], a["sdf"]) < 2e-6
    assert rel_err(b["std"], a["std"]) < 1e-4
    if float(a["gf"].abs().sum()) > 0:
        assert rel_err(b["gf"], a["gf"]) < 1e-5
    for x, y in zip(a["gd"], b["gd"]):
        if float(x.abs().sum()) > 0:
            assert rel_err(y, x) < 1e-5
    sa, sb = res[("signed", "mma")], res[("signed", "tc")]
    assert torch.equal(sa["nn"], sb["nn"]) and torch.equal(sb["sdf"], b["sdf"])
    if float(sa["gf"].abs().sum()) > 0:
        assert rel_err(sb["gf"], sa["gf"]) < 1e-5            # a feature row sums few terms: no cancellation to speak of
    for x, y, mag in zip(sa["gd"], sb["gd"], a["gd"]):
        scale = max(float(x.double().norm()), float(mag.double().norm()))
        if scale > 0:
            assert float((y.double() - x.double()).norm()) < 1e-5 * scale


def test_kernel_timing_marks(maps):
    """pinq_kernel_timing / pinq_kernel_times: CUDA-event time of each phase of the last fused forward + backward."""
    import ctypes
    from pin_slam_b200 import _pinq as P
    s, m, cfg, dec, d, _ = maps
    q = make_queries(s.points, 40000, seed=3).cuda()
    lib = P.load()
    cert0 = m.local_point_certainties.clone()
    tsu0 = m.local_point_ts_update.clone()
    lib.pinq_kernel_timing(1)
    try:
        out = m.query_sdf(q, dec, None, True, True)
        out["sdf"].sum().backward()
        torch.cuda.synchronize()
        buf = (ctypes.c_float * 8)()
        assert lib.pinq_kernel_times(buf, 8) == 7
    finally:
        lib.pinq_kernel_timing(0)
        m.local_point_certainties, m.local_point_ts_update = cert0, tsu0
    bin_, search, decode, bwd = buf[0], buf[1], buf[3], buf[6]
    assert 0 < bin_ < 50 and 0 < search < 50 and 0 < decode < 50 and 0 < bwd < 50
