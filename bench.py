#!/usr/bin/env python3
"""bench.py -- SDF queries/s (fwd+bwd) of the neural-point query path on B200 (BASELINE.json metric).

Workload (BASELINE.json configs[2]): isolated query microbench,
M ~ 1 M neural points (0.4 m voxels over a ~350 m synthetic street scene), N = 4 Mi unsorted queries per
GPU (90 % near-surface, 10 % free space), run_kitti parameters (Kn = 81, K = 6, F = 8), local query with
travel-distance filtering, training mode (certainty scatter + ts amax), forward + hand-derived backward
(feature-table and decoder gradients), upstream gradient = ones.

One "step" = one fused forward launch + one backward launch over the N queries of this rank
(+ NCCL all-reduce of the feature / decoder gradients when N_gpus > 1).

  python bench.py --gpus 1 --steps 10 --warmup 3
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference      # the CPU restatement of the reference path on the host cores
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "sdf_queries_per_sec_fwd_bwd"
UNIT = "queries/s"


# kernels of libpinq launched by one step (fused forward + backward of a binned training-mode batch):
# k_bin_count, k_bin_tile_sums, k_bin_tile_scan, k_bin_tile_apply, k_bin_scatter, k_stencil_masks, k_search_brick,
# k_decode_tc, k_bwd_tc (the resolve step and the certainty scatter run in the tail of k_search_brick unless the
# caller asks for the queried certainty; through NeuralPoints k_stencil_masks runs once per stencil, not per step:
# 8 launches per step, as in profiles/r2n_launches.csv, the ncu list of the same command).
LAUNCHES_PER_STEP = 8
KERNEL_NAMES = {"bin": "k_bin_* (5 kernels)", "search": "k_search_brick", "resolve": "k_resolve",
                "decode": "k_decode_tc", "scatter": "k_certainty_scatter", "bwd": "k_bwd_tc"}


def ncu_traffic():
    """dram bytes (read + write) per launch of each kernel from the committed `ncu --set full` capture of this
    command (profiles/ncu_traffic.json, written by profiles/summarize.py); {} when no capture is committed."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="pinq", choices=["pinq", "reference"])
    ap.add_argument("--queries", type=int, default=4 * 1024 * 1024, help="queries per GPU per step")
    ap.add_argument("--surface-points", type=int, default=9_000_000, help="raw surface samples -> ~1 M neural points")
    ap.add_argument("--side", type=float, default=350.0)
    ap.add_argument("--cpu-sample", type=int, default=200_000, help="queries per CPU-baseline step")
    ap.add_argument("--eager-sample", type=int, default=1 << 20, help="queries per step of the reference CUDA-eager comparator")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra-legs", action="store_true", help="skip flavours / sweep / configs[3] mapping / configs[4] mesher")
    ap.add_argument("--driver-frames", type=int, default=100, help="N=1 only: frames of the run through the reference's unchanged "
                    "pin_slam.py (drop-in and reference CUDA-eager); 0 = skip")
    ap.add_argument("--slam-seeded", action="store_true", help="SLAM leg: a seeded CPU generator for the sampler and the batch draws "
                    "(reproducible, slower) instead of torch's device RNG (what the reference uses)")
    ap.add_argument("--slam-frames", type=int, default=30, help="N=1 only: frames of the synthetic 64-beam SLAM run reported "
                    "under 'slam' (BASELINE configs[1]; 0 = skip)")
    return ap.parse_args()


def bind_to_gpu_numa_node(local_rank):
    """Pin this rank's host threads (and with them the first-touch placement of its pinned buffers) to the NUMA node its GPU
    hangs off: without it cudaHostAlloc'ed pages land on whichever node the thread happened to run on and half the ranks of a
    two-socket box pull their batches across the socket interconnect.  Returns what was done for the JSON line."""
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
            node = int(f.read().strip())
        if node < 0:
            return {"numa_node": node, "bound": False}
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return {"numa_node": node, "bound": False}
        os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "bound": True, "cpus": len(cpus)}
    except Exception as e:  # no sysfs / no permission: leave the scheduler alone
        return {"bound": False, "why": str(e)[:80]}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------
def build_map(args, device):
    from pin_slam_b200 import synth
    from pin_slam_b200.model.decoder import Decoder
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config

    cfg = kitti_config(device)
    cfg.local_map_radius = 1.0e4  # microbench: the whole synthetic map is the local map (config 3)
    npts = NeuralPoints(cfg)
    npts.travel_dist = torch.zeros(2)
    raw = torch.from_numpy(synth.surface_points(args.surface_points, args.side, seed=0)).to(device)
    npts.update(raw, torch.zeros(3, device=device), torch.eye(3, device=device), 0)
    del raw
    g = torch.Generator(device="cpu").manual_seed(0)
    m = npts.count()
    feats = torch.randn((m + 1, cfg.feature_dim), generator=g) * 0.1
    feats[-1] = 0
    cert = torch.rand(m, generator=g) * 50
    npts.geo_features = feats.to(device)
    npts.point_certainties = cert.to(device)
    with torch.no_grad():
        npts.local_geo_features.data = npts.geo_features[npts._local_idx].clone()
    npts.local_point_certainties = npts.point_certainties[npts._local_idx[:-1]].clone()
    torch.manual_seed(42)
    dec = Decoder(cfg, cfg.geo_mlp_hidden_dim, cfg.geo_mlp_level, 1)
    return cfg, npts, dec


def mean_nonempty_slots(npts, q, sample=100_000):
    """V-bar of SURVEY.md section 8(d): mean number of non-empty probed slots per query."""
    x = q[:sample]
    cells = torch.floor(x / npts.resolution).to(torch.int64)[:, None, :] + npts.neighbor_dx[None]
    h = torch.fmod((cells * npts.primes).sum(-1), npts.buffer_size)
    h = torch.where(h < 0, h + npts.buffer_size, h)
    return float((npts._table[h] >= 0).sum(1).float().mean())


_CLOCK_CHILD = r"""
import sys, time
import pynvml as nv
nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(int(sys.argv[1]))
period = float(sys.argv[2])
mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
bits = (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
        ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap))
print("ready", mx, flush=True)
while True:
    t = time.time()
    sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
    print(t, sm, "|".join(n for n, b in bits if r & int(b)), flush=True)
    time.sleep(period)
"""


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region by a separate process polling NVML every few ms
    (time-stamped readings, filtered to the region afterwards).  A separate process because NVML calls made from the
    benchmark's own process were measured to stall its kernel launches (an in-process polling thread cost 5 % at 2 ms and
    whole milliseconds per step on a 2-GPU box); the timed region of the default run is tens of milliseconds, shorter than
    nvidia-smi's fastest loop, hence NVML rather than `nvidia-smi -lms`."""

    def __init__(self, gpu_index):
        self.proc, self.mx, self.t0, self.t1 = None, None, None, None
        period = float(os.environ.get("PINQ_CLOCK_POLL_MS", "4")) * 1e-3
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[gpu_index]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else gpu_index
        try:
            env = dict(os.environ)
            env.pop("CUDA_VISIBLE_DEVICES", None)
            self.proc = subprocess.Popen([sys.executable, "-c", _CLOCK_CHILD, str(phys), str(period)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True, env=env)
            first = self.proc.stdout.readline().split()  # blocks until NVML is up: start-up stays outside the timed region
            if len(first) != 2 or first[0] != "ready":
                raise RuntimeError("clock sampler did not start")
            self.mx = float(first[1])
        except Exception:
            if self.proc is not None:
                self.proc.kill()
            self.proc = None

    def begin(self):
        self.t0 = time.time()

    def end(self):
        self.t1 = time.time()

    def stop(self):
        if self.proc is None:
            return None
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            return None
        sm, reasons = [], set()
        for line in out.splitlines():
            f = line.split()
            try:
                t, clk = float(f[0]), float(f[1])
            except (ValueError, IndexError):
                continue
            if self.t0 is not None and self.t1 is not None and not (self.t0 <= t <= self.t1):
                continue
            sm.append(clk)
            if len(f) > 2:
                reasons.update(x for x in f[2].split("|") if x)
        if not sm:
            return None
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": self.mx, "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvml, separate process, readings inside the timed region"}


# ------------------------------------------------------------------------------------------------
def reference_legs(args, npts, dec, q_host, ts_host):
    """Comparator legs of the GPU arm (N = 1, rank 0), both running the UNMODIFIED reference (oracle/_ref through
    oracle/ref_shim.py) on the very map the CUDA path just used: (1) cpu_baseline -- its CPU path on the host cores on a
    bounded sample; (2) reference_cuda_eager -- the same code with device = cuda on this B200 (stock PyTorch eager), the
    comparator a user of the reference would run."""
    from oracle import ref_bench, ref_shim

    if not ref_shim.available():
        return None, {"unavailable": "oracle/_ref not staged"}
    state = {"buffer_pt_index": npts.buffer_pt_index, "travel_dist": npts.travel_dist}
    for k in ("neural_points", "point_orientations", "geo_features", "point_ts_create", "point_ts_update", "point_certainties"):
        state[k] = getattr(npts, k)
    state["W1"], state["b1"], state["W2"], state["b2"] = [t.detach() for t in dec.fused_params()]
    state = {k: v.detach().clone() for k, v in state.items()}
    sample = min(args.cpu_sample, q_host.shape[0])
    cpu = eager = None
    try:
        _, rn, rd = ref_bench.reference_map_from_state(state, "cpu")
        qps, dt = ref_bench.time_reference(rn, rd, q_host[:sample].contiguous(), ts_host[:sample].contiguous(), 3, 1, "cpu")
        cpu = {"value": qps, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "reference",
               "sample": "%d queries of the same set, fwd+bwd, 3 steps after 1 warm-up: unmodified reference "
                         "query_feature -> Decoder.sdf -> sum w s -> backward() on torch CPU fp32, same map" % sample}
        del rn, rd
    except Exception as e:
        cpu = {"error": repr(e)[:300]}
    try:
        dev = str(npts.neural_points.device)
        _, rn, rd = ref_bench.reference_map_from_state(state, dev)
        ne = min(args.eager_sample, q_host.shape[0])
        xe, te = q_host[:ne].to(dev), ts_host[:ne].to(dev)
        qps, dt = ref_bench.time_reference(rn, rd, xe, te, 5, 2, dev)
        eager = {"queries_per_s": qps, "ms_per_step": dt * 1e3, "queries_per_step": ne,
                 "what": "unmodified reference query_feature -> Decoder.sdf -> sum w s -> backward(), device=cuda (PyTorch eager) "
                         "on the same B200 and the same map, 5 steps after 2 warm-ups, host-synchronised wall clock"}
        del rn, rd, xe, te
        torch.cuda.empty_cache()
    except Exception as e:
        eager = {"error": repr(e)[:300]}
    return cpu, eager


def run_reference(args):
    """--impl reference: the UNMODIFIED reference (oracle/_ref = byte-for-byte copy of /root/reference staged by
    oracle/make_ref.py, imported through oracle/ref_shim.py) on the host cores with all threads: the map is built by the
    reference's own NeuralPoints.update, each step is query_feature -> Decoder.sdf -> sum_k w_k s_k -> backward() on a
    bounded sample of the GPU arm's query set.  This process never imports pin_slam_b200._pinq (no libpinq.so)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_bench, ref_shim
    from pin_slam_b200 import synth  # numpy-only generators shared with the GPU arm

    if not ref_shim.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not staged (run python oracle/make_ref.py in the build container)"}))
        return
    torch.set_num_threads(os.cpu_count() or 1)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    raw = synth.surface_points(args.surface_points, args.side, seed=0)
    cfg, npts, dec = ref_bench.build_reference_map(raw, "cpu")
    del raw
    m = int(npts.neural_points.shape[0])
    q = torch.from_numpy(synth.microbench_queries(npts.neural_points.numpy(), args.queries, seed=1))  # rank 0's query set
    vbar = ref_bench.mean_nonempty_slots(npts, q)
    sample = min(args.cpu_sample, args.queries)
    x, ts = q[:sample].contiguous(), torch.zeros(sample, dtype=torch.int64)
    qps, dt = ref_bench.time_reference(npts, dec, x, ts, args.steps, args.warmup, "cpu")
    cores = torch.get_num_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, m, vbar, world),
        "cpu_baseline": {"value": qps, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "%d queries/step (the first of the GPU arm's %d), unmodified reference "
                                   "NeuralPoints.query_feature -> Decoder.sdf -> sum w s -> backward(), torch CPU fp32, "
                                   "map built by the reference's own update()" % (sample, args.queries)},
        "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_root": os.path.relpath(ref_shim.REFERENCE_ROOT, ROOT) if ref_shim.REFERENCE_ROOT.startswith(ROOT) else ref_shim.REFERENCE_ROOT,
    }
    print(json.dumps(line))


def _scan_job(job):
    from pin_slam_b200 import synth
    pose, noise_seed = job
    return synth.lidar_scan(synth.StreetScene(42), pose, synth.lidar_directions(64, 1875), np.random.default_rng(noise_seed))


def slam_leg(n_frames, device_rng=True):
    """BASELINE.json's second metric, SLAM frames/s (configs[1]): the full per-frame loop -- preprocess, tracking, map update,
    replay pool, 15 mapping iterations (600 on the first frame) -- on synthetic 64-beam scans (120 000 rays each, SURVEY.md
    8(d)), run_kitti parameters.  Scans are ray-cast on the host cores beforehand (not timed); times are the reference's own
    per-stage columns (pin_slam.py:343) taken with a device synchronize on both sides of every stage."""
    import multiprocessing as mp

    from pin_slam_b200 import synth
    from pin_slam_b200.slam import absolute_trajectory_error, run_slam
    from pin_slam_b200.utils.config import kitti_config

    gt = synth.trajectory(n_frames)
    with mp.get_context("spawn").Pool(min(os.cpu_count() or 1, 16)) as pool:
        frames = pool.map(_scan_job, [(gt[i], 1000 + i) for i in range(n_frames)])
    cfg = kitti_config("cuda")
    cfg.deskew, cfg.correction_deg, cfg.silence = False, 0.0, True  # harness deviations
    stages = []
    res = run_slam(cfg, frames, None, generator=None if device_rng else torch.Generator().manual_seed(7),
                   on_frame=lambda fid, dataset, npts, mapper: stages.append(dict(mapper.timing)))
    tt = res["time_table"]
    return {"frames": n_frames, "rays_per_scan": 120000, "rng": "device" if device_rng else "seeded cpu generator", "frames_per_s": float((n_frames - 1) / tt[1:].sum()),
            "first_frame_s": float(tt[0].sum()),
            "ms_per_frame": {k: float(tt[1:, i].mean() * 1e3) for i, k in enumerate(("preprocess", "tracking", "map_update", "mapping"))},
            "map_update_stages_ms": {k: float(np.mean([t[k] for t in stages[1:] if k in t])) for k in (stages[-1] if stages else {})},
            "median_ms_per_frame": float(np.median(tt[1:].sum(1)) * 1e3),
            "per_frame_ms": [round(float(v) * 1e3, 1) for v in tt.sum(1)],
            "ate_vs_ground_truth_m": float(absolute_trajectory_error(res["poses"], gt, align=False)),
            "note": "frames/s excludes the first frame (600 warm-up iterations)"}


def workload_config(args, m, vbar, world):
    """The `config` object -- identical on both arms (the reference arm rebuilds the same map and query set on the host)."""
    return {"workload": "isolated SDF query microbench (BASELINE configs[2]): fwd+bwd, K=6, Kn=81, F=8, local+time filter, "
                        "training mode, feature+decoder grads",
            "neural_points": int(m), "queries_per_gpu": int(args.queries), "voxel_size_m": 0.4,
            "buffer_size": 50000000, "map_side_m": float(args.side), "surface_samples": int(args.surface_points),
            "l2_policy": "inputs larger than L2 (200 MB slot table + 48 MB queries + 32 MB features per step)",
            "mean_nonempty_slots": round(vbar, 2),
            "reference_arm_queries_per_step": int(min(args.cpu_sample, args.queries)),
            "parallelism": "replicated map, queries sharded x%d" % world}


# ------------------------------------------------------------------------------------------------
# Extra legs carried inside the same JSON line: other flavours of the path, the configuration
# sweep of BASELINE configs[2], configs[3] (batched mapping, strong scaling) and configs[4] (mesher grid + marching cubes).
def _max_over_ranks(ms, dev, world):
    import torch.distributed as dist
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def flavour_legs(npts, dec, q, vbar, hbm):
    """Tracking flavour (fwd + d sdf/dx + std, inference mode; SURVEY 8(d): 577 + 20 V-bar B/query) at the bench batch and in
    the latency regime of a Gauss-Newton iteration (15 000 source points); mesher flavour (forward only, global map, no time
    filter, coordinates generated in-kernel; 533 + 12 V-bar B/point with V-bar measured on the grid points)."""
    from pin_slam_b200.utils.mesher import Mesher
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = {}

    def timed(fn, reps, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    with torch.no_grad():
        n = q.shape[0]
        ms = timed(lambda: npts.query_sdf(q, dec, None, False, True, want_grad=True, want_std=True), 10)
        bq = 577.0 + 20.0 * round(vbar, 2)
        out["tracking"] = {"queries_per_s": n / (ms * 1e-3), "ms": ms, "queries": n, "bytes_per_query": bq,
                           "frac": n / (ms * 1e-3) * bq / (hbm * 1e9), "outputs": "sdf, d sdf/dx, std, nn_count"}
        q15 = q[:15000].contiguous()
        ms15 = timed(lambda: npts.query_sdf(q15, dec, None, False, True, want_grad=True, want_std=True), 200, 20)
        out["tracking"]["latency_regime"] = {"queries": 15000, "us_per_call": ms15 * 1e3, "queries_per_s": 15000 / (ms15 * 1e-3),
                                             "what": "one Gauss-Newton iteration's query (host launch path included)"}
        mesher = Mesher(npts.config, npts, dec, None, None)
        pts = npts.neural_points
        lo = pts.min(0).values.cpu().numpy()
        hi = lo + np.array([64.0, 64.0, pts.max(0).values.cpu().numpy()[2] - lo[2]])
        hi[:2] = np.minimum(hi[:2] + 100.0, pts.max(0).values.cpu().numpy()[:2])   # a 164 m x 164 m tile of the map at 0.1 m
        dims, origin = mesher.grid_from_bounds(lo, hi, 0.1)
        total = int(np.prod(dims))
        for _ in range(2):
            e0.record()
            sdf, mask, _, _, _ = mesher.query_grid(lo, hi, 0.1, mask_min_nn_count=npts.config.mesh_min_nn)
            e1.record()
            torch.cuda.synchronize()
            msg = e0.elapsed_time(e1)
        # V-bar of the grid points (global table, no time filter)
        g = torch.randint(0, total, (100_000,), device=q.device)
        iz, r = g % int(dims[2]), g // int(dims[2])
        iy, ix = r % int(dims[1]), r // int(dims[1])
        gp = torch.stack((ix, iy, iz), 1).float() * 0.1 + torch.tensor(origin, dtype=torch.float32, device=q.device)
        vg = mean_nonempty_slots(npts, gp)
        bm = 533.0 + 12.0 * vg
        out["mesher"] = {"points_per_s": total / (msg * 1e-3), "ms": msg, "grid_points": total, "grid_dims": [int(v) for v in dims],
                         "mean_nonempty_slots": vg, "bytes_per_point": bm, "frac": total / (msg * 1e-3) * bm / (hbm * 1e9),
                         "masked_in_fraction": float((mask != 0).float().mean())}
        del sdf, mask
    return out


def sweep_legs(args, hbm):
    """BASELINE configs[2] spans 1-10 M queries against 0.5-5 M neural points: the same step at the small and the large end."""
    from pin_slam_b200 import synth
    out = []
    for name, n, surf, side in (("1M x 0.5M", 1_000_000, 3_000_000, 200.0), ("10M x 5M", 10_000_000, 30_000_000, 640.0)):
        try:
            cfg, npts, dec = build_map(argparse.Namespace(surface_points=surf, side=side), "cuda")
            q = torch.from_numpy(synth.microbench_queries(npts.neural_points.cpu().numpy(), n, seed=1)).cuda()
            ts = torch.zeros(n, dtype=torch.int64, device="cuda")
            ones = torch.ones(n, device="cuda")
            vbar = mean_nonempty_slots(npts, q)
            params = [npts.local_geo_features] + list(dec.parameters())

            def step():
                for p in params:
                    p.grad = None
                o = npts.query_sdf(q, dec, ts, True, True)
                o["sdf"].backward(ones)
            for _ in range(3):
                step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(10):
                step()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            bq = 788.0 + 20.0 * round(vbar, 2)
            out.append({"config": name, "queries": n, "neural_points": int(npts.count()), "map_side_m": side, "mean_nonempty_slots": round(vbar, 2),
                        "ms_per_step": ms, "value": n / (ms * 1e-3), "unit": UNIT, "bytes_per_query": bq,
                        "frac": n / (ms * 1e-3) * bq / (hbm * 1e9), "brick_index": bool(npts._brick_safe() and npts.use_brick_index)})
            del npts, dec, q, ts, ones, params
            torch.cuda.empty_cache()
        except Exception as e:  # the headline must not depend on this leg
            out.append({"config": name, "error": repr(e)[:200]})
    return out


def mapping_leg(dev, world, rank, iters=15, batch=1 << 20, pool_n=8_000_000):
    """BASELINE configs[3]: batched mapping, 2^20 samples per iteration drawn with one shared seed, every rank trains on its
    slice, pinq_peer_exchange with the fused Adam per iteration (strong scaling: the global batch is fixed)."""
    import torch.distributed as dist
    from pin_slam_b200 import synth
    from pin_slam_b200.parallel import DataParallelMapping, MappingPeer, PeerExchange
    from pin_slam_b200.utils.mapper import Mapper
    cfg, npts, dec = build_map(argparse.Namespace(surface_points=12_000_000, side=400.0), dev)
    g = torch.Generator().manual_seed(3)
    pool = torch.from_numpy(synth.microbench_queries(npts.neural_points.cpu().numpy(), pool_n, seed=5)).to(dev)
    label = (torch.randn(pool_n, generator=g) * 0.1).to(dev)
    weight = torch.ones(pool_n, device=dev)
    tsp = torch.zeros(pool_n, dtype=torch.int64, device=dev)
    mapper = Mapper(cfg, None, npts, dec, None, None)
    cfg.bs, cfg.bs_new_sample = batch, 0
    mapper.global_coord_pool, mapper.coord_pool = pool, pool
    mapper.sdf_label_pool, mapper.weight_pool, mapper.time_pool = label, weight, tsp
    mapper.pool_sample_count = pool_n
    mapper.generator = torch.Generator(device=dev).manual_seed(11)  # the same Philox stream on every rank
    peer = None
    rows, fdim = npts.local_geo_features.shape
    if world > 1:
        peer = PeerExchange(MappingPeer.capacity(rows, fdim, dec, with_params=True), dev, grid=int(os.environ.get("PINQ_MAP_PEER_GRID", "48")),
                            threads=int(os.environ.get("PINQ_MAP_PEER_THREADS", "0")),
                            multicast=os.environ.get("PINQ_MAP_PEER_MULTICAST", "0") == "1")  # NVLS form: not measured on this leg
        torch.cuda.synchronize()
        dist.barrier()
        mapper.dp = DataParallelMapping(peer)
    for _ in range(2):
        mapper._mapping_fused(iters)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    mapper._mapping_fused(iters)
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = _max_over_ranks(e0.elapsed_time(e1), dev, world)
    n_eik = 6 * ((batch + cfg.gradient_decimation - 1) // cfg.gradient_decimation)
    res = {"samples_per_iter": batch, "queries_per_iter": batch + n_eik, "iters": iters, "neural_points": int(npts.count()),
           "ms_per_iter": ms / iters, "samples_per_s": batch * iters / (ms * 1e-3), "scaling": "strong",
           "form": "device-side loop (pinq_mapping_run)" + (", pinq_peer_exchange with fused Adam per iteration" if world > 1 else "")}
    if peer is not None:
        res["peer_exchange_error"] = peer.error()
        res["peer_grid"] = peer.grid
        dist.barrier()
        peer.close()
    del mapper, pool, label, weight, tsp
    torch.cuda.empty_cache()
    return res, (cfg, npts, dec)


def mesher_leg(world_map, dev, world, rank, voxel=0.1):
    """BASELINE configs[4]: 0.1 m grid over the 400 m map, x-slabs per rank, no collective: the dense query alone
    (grid points/s, time to the full grid) and query + marching cubes (triangles/s)."""
    import torch.distributed as dist
    from pin_slam_b200.utils.mesher import Mesher
    cfg, npts, dec = world_map
    mesher = Mesher(cfg, npts, dec, None, None)
    pts = npts.neural_points
    lo, hi = pts.min(0).values.cpu().numpy(), pts.max(0).values.cpu().numpy()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    res = {}
    for rep in range(2):  # first pass warms the allocator
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        sdf, mask, first, dims, origin = mesher.query_grid(lo, hi, voxel, mask_min_nn_count=cfg.mesh_min_nn, rank=rank, world=world)
        e1.record()
        torch.cuda.synchronize()
        ms_q = _max_over_ranks(e0.elapsed_time(e1), dev, world)
        del sdf, mask
    total = int(np.prod(dims))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    v, f, _, _ = mesher.recon_grid_mesh(lo, hi, voxel, cfg.mesh_min_nn, rank=rank, world=world)
    e1.record()
    torch.cuda.synchronize()
    ms_m = _max_over_ranks(e0.elapsed_time(e1), dev, world)
    tri = torch.tensor([float(f.shape[0]), float(v.shape[0])], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tri)
    res = {"voxel_m": voxel, "grid_dims": [int(x) for x in dims], "grid_points": total, "neural_points": int(npts.count()),
           "seconds_to_full_grid": ms_q * 1e-3, "grid_points_per_s": total / (ms_q * 1e-3),
           "seconds_query_plus_marching_cubes": ms_m * 1e-3, "triangles": int(tri[0].item()), "vertices": int(tri[1].item()),
           "triangles_per_s": float(tri[0].item()) / (ms_m * 1e-3), "collective": "none (x-slabs per rank, one shared layer)"}
    return res


def slam_driver_leg(frames):
    """BASELINE configs[1] through the reference's UNCHANGED pin_slam.py with pin_slam_b200.dropin (deskew and KITTI correction
    on, run_kitti.yaml), and the same driver without the drop-in on the same GPU (reference CUDA-eager) -- separate processes."""
    import tempfile
    out = {}
    data = tempfile.mkdtemp(prefix="pinq_kitti_bench_")
    try:
        for key, extra in (("dropin", ["--mode", "dropin"]), ("reference_cuda_eager", ["--mode", "reference", "--device", "cuda"])):
            dst = os.path.join(data, key + ".json")
            cmd = [sys.executable, os.path.join(ROOT, "scripts", "run_reference_driver.py"), "--frames", str(frames), "--data-dir", data,
                   "--out", dst] + extra
            p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
            if p.returncode == 0 and os.path.exists(dst):
                with open(dst) as f:
                    out[key] = json.load(f)
            else:
                out[key] = {"error": (p.stderr or p.stdout)[-400:]}
        if "poses" in out.get("dropin", {}) and "poses" in out.get("reference_cuda_eager", {}):
            a = np.array(out["dropin"]["poses"])[:, [3, 7, 11]]
            b = np.array(out["reference_cuda_eager"]["poses"])[:, [3, 7, 11]]
            out["ate_dropin_vs_reference_cuda_eager_m"] = float(np.sqrt(((a - b) ** 2).sum(1).mean()))
        for k in ("dropin", "reference_cuda_eager"):
            out.get(k, {}).pop("poses", None)
        if out.get("dropin", {}).get("frames_per_s") and out.get("reference_cuda_eager", {}).get("frames_per_s"):
            out["speedup_vs_reference_cuda_eager"] = out["dropin"]["frames_per_s"] / out["reference_cuda_eager"]["frames_per_s"]
    finally:
        import shutil
        shutil.rmtree(data, ignore_errors=True)
    return out


# ------------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import torch.distributed as dist
    from pin_slam_b200 import _pinq as P
    from pin_slam_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the neural-point query path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = "cuda:%d" % local_rank
    host_binding = bind_to_gpu_numa_node(local_rank) if os.environ.get("PINQ_NUMA_BIND", "1") != "0" else {"bound": False}
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout for the one JSON line
        dist.init_process_group("nccl", device_id=torch.device(dev))
    P.load()
    lib = P.load()

    cfg, npts, dec = build_map(args, dev)
    # every rank holds a replica of the map; its queries are its own shard (seeded by rank) -> weak scaling
    q_host = torch.from_numpy(synth.microbench_queries(npts.neural_points.cpu().numpy(), args.queries, seed=1 + rank))
    ts_host = torch.zeros(args.queries, dtype=torch.int64)
    q = q_host.to(dev)
    ts = ts_host.to(dev)
    vbar = mean_nonempty_slots(npts, q)
    n = args.queries
    ones = torch.ones(n, device=dev)
    params = [npts.local_geo_features] + list(dec.parameters())

    # data-parallel exchange step: the feature-table and decoder gradients are summed over the ranks.
    # Default: they accumulate straight into this rank's cudaIpc-shared peer block and ONE kernel per rank reduces them
    # over NVLink (pinq_peer_exchange, csrc/peer.cu); PINQ_EXCHANGE=nccl selects the NCCL all-reduce A/B partner.
    peer = carve = None
    exchange = "none"
    if world > 1:
        exchange = os.environ.get("PINQ_EXCHANGE", "peer")
        if exchange == "peer":
            from pin_slam_b200.parallel import MappingPeer, PeerExchange
            rows, fdim = npts.local_geo_features.shape
            # overlapped with the next step's search (below): a plain launch of FEW CTAs -- they take SM slots as the search's
            # CTAs retire (a cooperative grid would wait for an empty GPU); 32 CTAs x 512 threads keep ~4 MB of P2P loads in
            # flight, which is what the links need (measured at N = 2: 148 / 64 / 32 CTAs -> 2.131 / 2.088 / 2.078 ms per step)
            ov = os.environ.get("PINQ_EXCHANGE_OVERLAP", "1") != "0"
            peer = PeerExchange(MappingPeer.capacity(rows, fdim, dec), dev, grid=int(os.environ.get("PINQ_PEER_GRID", "32" if ov else "0")),
                                cooperative=not ov, threads=int(os.environ.get("PINQ_PEER_THREADS", "0")))
            carve = MappingPeer(peer, rows, fdim, dec)
            n_ex = carve.n_sum - (rows - 1)  # features + decoder (the certainty increments are not part of this step)
            npts.feature_grad_sink, dec.grad_sink = carve.grad_features, tuple(carve.grad_decoder)
            grad_block = peer.sum_region()[:n_ex]
            torch.cuda.synchronize()
            dist.barrier()

    def zero_grads():
        if peer is not None:
            grad_block.zero_()
        for p in params:
            p.grad = None

    skip_exchange = [False]
    # The peer exchange of step i runs on a side stream UNDER the binning + neighbour search of step i + 1 (they read query
    # points and map geometry only); the decode of step i + 1 -- which in training reads the parameters the exchange's
    # optimizer would have updated -- waits for it (pinq_query_args.decode_wait_event), and so does everything after it.
    overlap = peer is not None and os.environ.get("PINQ_EXCHANGE_OVERLAP", "1") != "0"
    # high priority: the exchange's few CTAs take the SM slots the search frees FIRST instead of queueing behind its 32 768 CTAs
    side = torch.cuda.Stream(priority=int(os.environ.get("PINQ_EXCHANGE_PRIORITY", "-1"))) if overlap else None
    ev_bwd, ev_x = (torch.cuda.Event(), torch.cuda.Event()) if overlap else (None, None)
    if overlap:
        ev_x.record(torch.cuda.current_stream())

    def exchange_grads():
        if skip_exchange[0]:
            return
        if peer is not None:
            peer.exchange(n_sum=n_ex, n_max=0)
        elif world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params[1:]])
            dist.all_reduce(npts.local_geo_features.grad)
            dist.all_reduce(flat)

    def step(x, t):
        if overlap and not skip_exchange[0]:
            main = torch.cuda.current_stream()
            npts.decode_wait_event = ev_x
            out = npts.query_sdf(x, dec, t, True, True)   # bin + search now, decode after the previous exchange + zero
            npts.decode_wait_event = None
            out["sdf"].backward(ones)                     # accumulates into the (zeroed) peer block
            ev_bwd.record(main)
            with torch.cuda.stream(side):
                side.wait_event(ev_bwd)
                peer.exchange(n_sum=n_ex, n_max=0)
                grad_block.zero_()
                ev_x.record(side)
            return out["sdf"]
        zero_grads()
        out = npts.query_sdf(x, dec, t, True, True)
        out["sdf"].backward(ones)
        exchange_grads()
        return out["sdf"]

    def drain():
        if overlap:
            torch.cuda.current_stream().wait_event(ev_x)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if overlap:
        zero_grads()
    for _ in range(max(args.warmup, 3)):
        step(q, ts)
    drain()
    sampler = ClockSampler(local_rank) if rank == 0 else None  # NVML start-up (tens of ms) stays outside the timed region
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if sampler:
        sampler.begin()
    ev0.record()
    for i in range(args.steps):
        step(q, ts)
    drain()  # the timed region ends when the last exchange has landed
    ev1.record()
    barrier()
    if sampler:
        sampler.end()
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = world * n / (ms_step * 1e-3)

    # ---- diagnostics (N > 1): every rank's own time for the timed region, and the step without its exchange
    rank_ms = no_exchange_ms = None
    if world > 1:
        allt = [torch.zeros(1, device=dev, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(allt, torch.tensor([ms_total / args.steps], device=dev, dtype=torch.float64))
        rank_ms = [float(x.item()) for x in allt]
        skip_exchange[0] = True
        barrier()
        ev0.record()
        for _ in range(args.steps):
            step(q, ts)
        ev1.record()
        barrier()
        tn = torch.tensor([ev0.elapsed_time(ev1) / args.steps], device=dev, dtype=torch.float64)
        dist.all_reduce(tn, op=dist.ReduceOp.MAX)
        no_exchange_ms = float(tn.item())
        skip_exchange[0] = False

    # ---- timeline of the overlapped step (PINQ_TRACE_OVERLAP=1): where the exchange sits under the next forward
    overlap_trace = None
    if overlap and os.environ.get("PINQ_TRACE_OVERLAP", "0") != "0":
        main = torch.cuda.current_stream()
        E = lambda: torch.cuda.Event(enable_timing=True)
        rec = []
        barrier()
        for i in range(12):
            e = {k: E() for k in ("q0", "q1", "b", "xs", "xe", "z")}
            e["q0"].record(main)
            npts.decode_wait_event = ev_x
            out = npts.query_sdf(q, dec, ts, True, True)
            npts.decode_wait_event = None
            e["q1"].record(main)
            out["sdf"].backward(ones)
            e["b"].record(main)
            ev_bwd.record(main)
            with torch.cuda.stream(side):
                side.wait_event(ev_bwd)
                e["xs"].record(side)
                peer.exchange(n_sum=n_ex, n_max=0)
                e["xe"].record(side)
                grad_block.zero_()
                e["z"].record(side)
                ev_x.record(side)
            rec.append(e)
        drain()
        barrier()
        med = lambda v: float(sorted(v)[len(v) // 2])
        r = rec[2:]
        overlap_trace = {
            "forward_ms": med([e["q0"].elapsed_time(e["q1"]) for e in r]),
            "backward_ms": med([e["q1"].elapsed_time(e["b"]) for e in r]),
            "exchange_under_search_ms": med([e["xs"].elapsed_time(e["xe"]) for e in r]),
            "zero_ms": med([e["xe"].elapsed_time(e["z"]) for e in r]),
            "backward_end_to_exchange_start_ms": med([e["b"].elapsed_time(e["xs"]) for e in r]),
            "exchange_end_to_next_forward_end_ms": med([rec[i]["z"].elapsed_time(rec[i + 1]["q1"]) for i in range(2, len(rec) - 1)]),
            "step_ms": med([rec[i]["q0"].elapsed_time(rec[i + 1]["q0"]) for i in range(2, len(rec) - 1)]),
        }

    # ---- the exchange step alone (N > 1): zeroed gradient block -> summed block, max over ranks
    exchange_ms = None
    if world > 1:
        if peer is None:
            step(q, ts)  # leaves p.grad populated for the NCCL form
        barrier()
        xe0, xe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        xe0.record()
        for _ in range(10):
            exchange_grads()
        xe1.record()
        barrier()
        tx = torch.tensor([xe0.elapsed_time(xe1) / 10], device=dev, dtype=torch.float64)
        dist.all_reduce(tx, op=dist.ReduceOp.MAX)
        exchange_ms = float(tx.item())

    # ---- per-kernel timing (CUDA events on the launching stream, recorded by the library between its kernels)
    K = 6
    a = npts._query_args(q, dec.view(), ts, True, True)
    sdf = torch.empty(n, device=dev)
    nnc = torch.empty(n, dtype=torch.int32, device=dev)
    stage = torch.empty(int(lib.pinq_stage_bytes(n)), dtype=torch.uint8, device=dev)
    relu_mask = torch.empty(int(lib.pinq_relu_mask_words(n)), dtype=torch.int32, device=dev)
    perm = torch.empty(n, dtype=torch.int32, device=dev)
    a.sdf, a.nn_count, a.stage, a.relu_mask, a.perm = P.ptr(sdf), P.ptr(nnc), P.ptr(stage), P.ptr(relu_mask), P.ptr(perm)
    b = P.BwdArgs()
    b.dec = dec.view()
    b.n, b.flags, b.k, b.feature_dim = n, 0, K, 8
    b.features = P.ptr(npts.local_geo_features.detach())
    b.stage, b.relu_mask, b.perm, b.grad_sdf = P.ptr(stage), P.ptr(relu_mask), P.ptr(perm), P.ptr(ones)
    gf = torch.zeros_like(npts.local_geo_features)
    gd = [torch.zeros_like(p) for p in dec.fused_params()]
    b.grad_features = P.ptr(gf)
    b.grad_W1, b.grad_b1, b.grad_W2, b.grad_b2 = (P.ptr(g) for g in gd)
    st = P.stream()
    reps = max(3, min(args.steps, 10))
    phases = ["bin", "search", "resolve", "decode", "scatter", None, "bwd"]
    acc = {k: [] for k in phases if k}
    buf = (ctypes.c_float * 8)()
    lib.pinq_kernel_timing(1)
    for i in range(reps + 2):
        P.check(lib.pinq_query_fused_fwd(a, st))
        P.check(lib.pinq_query_fused_bwd(b, st))
        torch.cuda.synchronize()
        lib.pinq_kernel_times(buf, 8)
        if i >= 2:
            for j, k in enumerate(phases):
                if k and buf[j] >= 0:
                    acc[k].append(float(buf[j]))
    lib.pinq_kernel_timing(0)
    kms = {k: float(np.mean(v)) for k, v in acc.items() if v and float(np.mean(v)) > 0.01}  # empty phases: event gaps
    ms_f = sum(v for k, v in kms.items() if k != "bwd")
    ms_b = kms.get("bwd", 0.0)
    hbm, peak_kind = peaks()
    # algorithmic bytes per query of the WHOLE path, each datum counted once with compact types:
    # B_q(fwd+bwd) = 788 + 20 V-bar.  roofline.frac is reproducible from `value`, config.mean_nonempty_slots and
    # MEASURED_PEAKS.json alone: frac = (value / n_gpus) * B_q / (hbm_gbs * 1e9).
    bq = 788.0 + 20.0 * round(vbar, 2)
    ach = (value / world) * bq / 1e9
    # per kernel: measured time (CUDA events recorded by the library between its kernels on the launching stream), the
    # DRAM bytes ncu measured for one launch of the same command (profiles/ncu_traffic.json, `ncu --set full`), and the
    # unit that limits it according to that capture (profiles/*_ncu_full_summary.csv) -- none of them is DRAM-bound.
    traffic = ncu_traffic()
    limiter = {"bin": "launch latency (small kernels) + histogram atomics",
               "search": "instruction issue at 36 of 64 warps (integer walk + top-K; ALU pipe half rate), dependent loads",
               "decode": "ALU pipe (relu + relu' bit of the CUDA-core epilogue of the tcgen05 product run on the half-rate pipe)",
               "bwd": "shared-memory data pipe (int8 tile stores + UMMA operand reads), 16 warps per SM",
               "resolve": "L2 gathers", "scatter": "L2 atomics"}
    kernels = {}
    for k, ms in kms.items():
        e = {"kernel": KERNEL_NAMES[k], "ms": ms, "share_of_step": ms / (ms_f + ms_b), "limiter": limiter.get(k)}
        if traffic.get(k):
            e["ncu_dram_bytes_per_launch"] = traffic[k]
            e["ncu_dram_bytes_queries"] = traffic.get("queries_per_launch", 4194304)
            e["dram_gbs_at_measured_ms"] = traffic[k] * (n / e["ncu_dram_bytes_queries"]) / (ms * 1e-3) / 1e9
            e["dram_frac_of_peak"] = e["dram_gbs_at_measured_ms"] / hbm
        kernels[k] = e
    tr = [traffic.get(k) for k in kms if traffic.get(k)]
    roofline = {"bound": "hbm", "achieved": ach, "peak": hbm, "peak_kind": peak_kind + " (burst copy)", "unit": "GB/s",
                "frac": ach / hbm, "traffic": float(sum(tr)) if tr else None,
                "bytes_per_query": bq, "formula": "SURVEY 8(d): (788 + 20 * mean_nonempty_slots) B/query * queries/s per GPU / hbm_gbs",
                "dominant_kernel": KERNEL_NAMES[max(kms, key=kms.get)], "kernels": kernels,
                "ms_fwd_kernels": ms_f, "ms_bwd_kernel": ms_b,
                "note": "`traffic` = dram bytes read+written per step summed over the kernels (ncu capture of the same command); "
                        "it is BELOW the algorithmic bytes because binned queries share candidate records and feature rows in L1/L2: "
                        "the step is issue / shared-memory bound, not DRAM bound"}

    # ---- end to end through the public API with HOST buffers ---------------------------------------------
    e2e = None
    if not args.no_e2e:
        # frame ids fit 32 bits: they cross PCIe as int32 and are widened to the API's int64 on the device
        qh, th = q_host.pin_memory(), ts_host.to(torch.int32).pin_memory()
        out_h = torch.empty(n, dtype=torch.float32).pin_memory()
        # The batch crosses PCIe in chunks on a copy stream, double-buffered against the fused forward + backward of
        # the previous chunk (gradients accumulate over the chunks as over one batch); results go back per chunk.
        nchunk = int(os.environ.get("PINQ_E2E_CHUNKS", "1"))  # per-chunk fixed costs (binning, launches) favour few chunks
        if n % nchunk or n < nchunk * 65536:
            nchunk = 1
        cn = n // nchunk
        copy_stream = torch.cuda.Stream()
        back_stream = torch.cuda.Stream()  # results return on their own stream: D2H overlaps the next chunk's kernels
        xb = [torch.empty((cn, 3), dtype=torch.float32, device=dev) for _ in range(2)]
        tb = [torch.empty(cn, dtype=torch.int64, device=dev) for _ in range(2)]
        tb32 = [torch.empty(cn, dtype=torch.int32, device=dev) for _ in range(2)]
        copied = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        ones_c = ones[:cn]
        chunk_no = [0]

        def e2e_step():
            zero_grads()
            main = torch.cuda.current_stream()
            for c in range(nchunk):
                bi = chunk_no[0] % 2  # buffers alternate across step boundaries too: the next step's first copy overlaps this step's kernels
                chunk_no[0] += 1
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(consumed[bi])  # the chunk that used this buffer is done
                    xb[bi].copy_(qh[c * cn:(c + 1) * cn], non_blocking=True)
                    tb32[bi].copy_(th[c * cn:(c + 1) * cn], non_blocking=True)
                    copied[bi].record(copy_stream)
                main.wait_event(copied[bi])
                tb[bi].copy_(tb32[bi])
                out = npts.query_sdf(xb[bi], dec, tb[bi], True, True)
                out["sdf"].backward(ones_c)
                consumed[bi].record(main)
                sdf_c = out["sdf"].detach()
                back_stream.wait_event(consumed[bi])
                with torch.cuda.stream(back_stream):
                    out_h[c * cn:(c + 1) * cn].copy_(sdf_c, non_blocking=True)
                sdf_c.record_stream(back_stream)
            exchange_grads()
            if not pipelined:
                main.wait_stream(back_stream)  # the step ends when its results are on the host

        # steps are pipelined like a prefetching data loader: the next step's kernels do not wait for this step's results to
        # reach the host (they return on their own stream); the timed region ends when the LAST result is on the host.
        # PINQ_E2E_PIPELINE=0: every step waits for its own results before the next one starts.
        pipelined = os.environ.get("PINQ_E2E_PIPELINE", "1") != "0"
        for _ in range(max(5, args.warmup)):  # PCIe link and pinned pages warm
            e2e_step()
        # the copies alone, all ranks at once: the host-side ceiling of this leg (bytes of one step over the slowest rank's time)
        torch.cuda.current_stream().wait_stream(back_stream)
        barrier()
        ce0, ce1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ce0.record()
        for _ in range(5):
            xb[0].copy_(qh[:cn], non_blocking=True)
            tb32[0].copy_(th[:cn], non_blocking=True)
            out_h[:cn].copy_(sdf[:cn], non_blocking=True)
        ce1.record()
        barrier()
        tc = torch.tensor([ce0.elapsed_time(ce1) / 5], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tc, op=dist.ReduceOp.MAX)
        copy_ms = float(tc.item()) * nchunk
        torch.cuda.current_stream().wait_stream(back_stream)
        barrier()
        ev0.record()
        t_host = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        t_host = (time.perf_counter() - t_host) / args.steps * 1e3  # host time to ENQUEUE a step (the GPU runs behind it)
        torch.cuda.current_stream().wait_stream(back_stream)
        ev1.record()
        barrier()
        te = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": world * n / (float(te.item()) / args.steps * 1e-3), "unit": UNIT, "chunks": nchunk, "steps_pipelined": pipelined,
               "h2d_bytes_per_step": n * 12 + n * 4, "d2h_bytes_per_step": n * 4,
               "host_enqueue_ms_per_step": t_host, "copies_alone_ms_per_step": copy_ms, "copies_alone_gbs_per_gpu": (n * 20) / (copy_ms * 1e-3) / 1e9,
               "host_binding": host_binding}

    cpu_baseline = cuda_eager = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_baseline, cuda_eager = reference_legs(args, npts, dec, q_host, ts_host)

    peer_err = peer.error() if peer is not None else 0
    had_peer = peer is not None
    peer_form = None if peer is None else ("NVLS multimem.ld_reduce / multimem.st" if peer.mc_block else "P2P loads / stores")
    config_line = workload_config(args, npts.count(), vbar, world)
    slam = flav = sweep = mapping = mesher_res = None
    if rank == 0 and world == 1 and not args.no_extra_legs:
        try:
            flav = flavour_legs(npts, dec, q, vbar, hbm)
        except Exception as e:  # the headline line must not depend on these legs
            flav = {"error": repr(e)[:300]}
    if world > 1 and peer is not None:  # the microbench's peer block is not needed any more: free it before the other legs
        npts.feature_grad_sink = dec.grad_sink = None
        dist.barrier()
        peer.close()
        peer = grad_block = carve = None
    del q, ts, ones, stage, relu_mask, perm, gf, sdf, nnc
    npts = dec = None
    torch.cuda.empty_cache()
    if not args.no_extra_legs:
        try:
            mapping, world_map = mapping_leg(dev, world, rank)
            mesher_res = mesher_leg(world_map, dev, world, rank)
            del world_map
            torch.cuda.empty_cache()
        except Exception as e:
            mapping = mapping or {"error": repr(e)[:300]}
            mesher_res = mesher_res or {"error": repr(e)[:300]}
    if rank == 0 and world == 1 and not args.no_extra_legs:
        sweep = sweep_legs(args, hbm)
    if rank == 0 and world == 1 and args.slam_frames > 1:
        try:
            slam = slam_leg(args.slam_frames, not args.slam_seeded)
        except Exception as e:
            slam = {"error": repr(e)[:200]}
        if args.driver_frames > 1:
            try:
                slam["reference_driver"] = slam_driver_leg(args.driver_frames)
            except Exception as e:
                slam["reference_driver"] = {"error": repr(e)[:300]}
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_line,
                "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e,
                "gpu_launches": LAUNCHES_PER_STEP * args.steps, "clocks": clocks}
        if slam is not None:
            line["slam"] = slam
        for key, val in (("flavours", flav), ("mapping", mapping), ("mesher", mesher_res), ("sweep", sweep)):
            if val is not None:
                line[key] = val
        if cuda_eager is not None:
            line["reference_cuda_eager"] = cuda_eager
        line["exchange"] = {"none": "none", "nccl": "NCCL all-reduce of feature+decoder grads",
                            "peer": "feature+decoder grads summed by one peer-memory kernel per rank over NVLink "
                                    "(pinq_peer_exchange)"}[exchange]
        if had_peer:
            line["gpu_launches"] += args.steps
            line["peer_exchange_error"] = peer_err
            line["exchange_overlapped_with_next_search"] = bool(overlap)
            line["exchange_form"] = peer_form
            if overlap_trace:
                line["overlap_trace"] = overlap_trace
        if exchange_ms is not None:
            line["exchange_ms"] = exchange_ms
            line["ms_per_step_by_rank"] = rank_ms
            line["ms_per_step_without_exchange"] = no_exchange_ms
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
