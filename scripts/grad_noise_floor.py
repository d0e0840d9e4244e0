#!/usr/bin/env python3
"""fp32 rounding floor of the reference's own mapping gradients (CPU, no GPU needed).

The golden fixture holds the UNMODIFIED reference's loss / gradients of one mapping iteration in fp32 (1 thread).  The same
iteration is replayed through the oracle in float64: the distance fp32-reference <-> fp64 is how far the reference's own
numbers are from the exact value of the formula, i.e. the spread any other correct fp32 implementation may show.  A parity
tolerance below that spread would test rounding order, not the algorithm.
  python scripts/grad_noise_floor.py [--out profiles/r3_grad_noise_floor.json]"""
import argparse
import copy
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
from helpers import O, golden_oracle_world, load_golden, rel_err  # noqa: E402


def to_double(obj):
    out = copy.copy(obj)
    for k, v in vars(obj).items():
        if torch.is_tensor(v) and v.dtype == torch.float32:
            setattr(out, k, v.double())
    return out


def run(G, s, d, threads):
    torch.set_num_threads(threads)
    T = torch.from_numpy
    s2 = copy.copy(s)
    s2.l_feats = s.l_feats.clone().requires_grad_(True)
    s2.l_certainty = T(G["map_cert_before"]).clone().to(s.l_feats.dtype)
    s2.l_ts_update = s.l_ts_update.clone()
    dp = O.DecoderParams(*[t.clone().requires_grad_(True) for t in d.tensors()], sdf_scale=d.sdf_scale)
    dt = s.l_feats.dtype
    loss, pred, _ = O.mapping_loss(s2, dp, T(G["map_coord"]).to(dt), T(G["map_label"]).to(dt), T(G["map_ts"]), T(G["map_weight"]).to(dt))
    loss.backward()
    return {"pred": pred.detach(), "gfeat": s2.l_feats.grad, "gW1": dp.W1.grad, "gb1": dp.b1.grad, "gW2": dp.W2.grad, "gb2": dp.b2.grad}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    G = load_golden()
    s, d = golden_oracle_world(G)
    T = torch.from_numpy
    ref = {"pred": T(G["map_pred"]), "gfeat": T(G["map_gfeat"]), "gW1": T(G["map_gW1"]), "gb1": T(G["map_gb1"]),
           "gW2": T(G["map_gW2"]), "gb2": T(G["map_gb2"])}
    out = {}
    for name, (ss, dd, th) in {"oracle_fp32_1thread": (s, d, 1), "oracle_fp32_8threads": (s, d, 8),
                               "oracle_fp64": (to_double(s), to_double(d), 1)}.items():
        try:
            torch.set_default_dtype(torch.float64 if name.endswith("fp64") else torch.float32)
            r = run(G, ss, dd, th)
            out[name] = {k: rel_err(r[k], ref[k].reshape(r[k].shape)) for k in ref}
        except Exception as e:
            out[name] = {"error": repr(e)[:200]}
    torch.set_default_dtype(torch.float32)
    print(json.dumps({"rel_err_vs_unmodified_reference_fp32": out}, indent=1))
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"rel_err_vs_unmodified_reference_fp32": out}, f, indent=1)


if __name__ == "__main__":
    main()
