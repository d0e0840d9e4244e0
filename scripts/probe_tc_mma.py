import sys, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from helpers import *
frames = frames_of_points(3, 40000)
s = build_oracle_map(frames); m, cfg = build_cuda_map(frames); randomize_features(s, m)
dec, d = decoder_pair(cfg)
n = 6000
q = make_queries(s.points, n, seed=17)[:n].cuda()
ts = torch.full((n,), 1, dtype=torch.int64, device="cuda")
up = torch.linspace(-1.0, 1.0, n, device="cuda")
cert0, tsu0 = m.local_point_certainties.clone(), m.local_point_ts_update.clone()
def run(flag):
    m.use_mma_sync = flag
    m.local_point_certainties, m.local_point_ts_update = cert0.clone(), tsu0.clone()
    m.local_geo_features.grad = None
    for p_ in dec.parameters(): p_.grad = None
    out = m.query_sdf(q, dec, ts, True, True)
    (out["sdf"] * up).sum().backward()
    return [p_.grad.clone() for p_ in dec.parameters()], m.local_geo_features.grad.clone()
# oracle (fp32 and fp64) as arbiter
def orc(ss, dd, dt):
    import copy
    s2 = copy.copy(ss); s2.l_feats = ss.l_feats.clone().requires_grad_(True)
    s2.l_certainty = ss.l_certainty.clone(); s2.l_ts_update = ss.l_ts_update.clone()
    dp = O.DecoderParams(*[t.clone().requires_grad_(True) for t in dd.tensors()], sdf_scale=dd.sdf_scale)
    r = O.query_sdf(s2, dp, q.cpu().to(dt), ts.cpu(), True, True)
    (r["sdf"] * up.cpu().to(dt)).sum().backward()
    return [t.grad for t in dp.tensors()]
g32 = orc(s, d, torch.float32)
torch.set_default_dtype(torch.float64); g64 = orc(to_double(s), to_double(d), torch.float64); torch.set_default_dtype(torch.float32)
names = ["W1", "b1", "W2", "b2"]
for rep in range(4):
    tc, _ = run(False); mma, _ = run(True)
    print("rep", rep, {k: "tc-mma %.1e tc-o64 %.1e mma-o64 %.1e" % (rel_err(a, b), rel_err(a, c.reshape(a.shape)), rel_err(b, c.reshape(b.shape)))
                       for k, a, b, c in zip(names, tc, mma, g64)})
print("oracle32 vs oracle64", {k: "%.1e" % rel_err(a, b) for k, a, b in zip(names, g32, g64)})
m.use_mma_sync = False
