"""Seeded synthetic inputs: a street-like scene, 64-beam LiDAR scans, and the query-microbench
workload of SURVEY.md section 8(d) (BASELINE.json configs[0..2]).  numpy only; no network, no datasets.
"""
from __future__ import annotations

import numpy as np


# --------------------------------------------------------------------------------------------
# surface sampling for the isolated query microbench (config 3)
# --------------------------------------------------------------------------------------------
def surface_points(n: int, side: float, seed: int = 0) -> np.ndarray:
    """n points on a ground plane + facades + boxes inside a side x side square (float32 [n,3])."""
    rng = np.random.default_rng(seed)
    n_ground = int(n * 0.55)
    n_wall = int(n * 0.25)
    n_box = n - n_ground - n_wall
    half = side / 2
    g = np.empty((n_ground, 3), np.float32)
    g[:, 0] = rng.uniform(-half, half, n_ground)
    g[:, 1] = rng.uniform(-half, half, n_ground)
    g[:, 2] = -1.73 + 0.02 * rng.standard_normal(n_ground) + 0.15 * np.sin(g[:, 0] * 0.05) * np.cos(g[:, 1] * 0.07)
    # facades: vertical planes on a street grid every ~40 m
    w = np.empty((n_wall, 3), np.float32)
    lane = rng.integers(-int(half // 40), int(half // 40) + 1, n_wall) * 40.0
    along = rng.uniform(-half, half, n_wall)
    axis = rng.integers(0, 2, n_wall)
    w[:, 0] = np.where(axis == 0, lane + 9.0, along)
    w[:, 1] = np.where(axis == 0, along, lane + 9.0)
    w[:, 2] = rng.uniform(-1.7, 6.0, n_wall)
    w[:, :2] += 0.02 * rng.standard_normal((n_wall, 2)).astype(np.float32)
    # boxes ("cars")
    nb = max(1, n_box // 400)
    centres = np.stack([rng.uniform(-half, half, nb), rng.uniform(-half, half, nb), np.full(nb, -0.9)], 1)
    which = rng.integers(0, nb, n_box)
    face = rng.integers(0, 3, n_box)
    ext = np.array([2.2, 0.9, 0.8])
    b = centres[which] + rng.uniform(-1, 1, (n_box, 3)) * ext
    sign = rng.choice([-1.0, 1.0], n_box)
    for a in range(3):
        sel = face == a
        b[sel, a] = centres[which[sel], a] + sign[sel] * ext[a]
    return np.concatenate([g, w, b.astype(np.float32)], 0).astype(np.float32)


def microbench_queries(neural_points: np.ndarray, n: int, seed: int = 1) -> np.ndarray:
    """90 % = random neural point + N(0, 0.1 m)^3, 10 % = uniform in the map AABB; unsorted (config 3)."""
    rng = np.random.default_rng(seed)
    n_far = n // 10
    n_near = n - n_far
    pick = rng.integers(0, neural_points.shape[0], n_near)
    near = neural_points[pick] + (0.1 * rng.standard_normal((n_near, 3))).astype(np.float32)
    lo, hi = neural_points.min(0), neural_points.max(0)
    far = rng.uniform(lo, hi, (n_far, 3)).astype(np.float32)
    q = np.concatenate([near, far], 0).astype(np.float32)
    rng.shuffle(q, axis=0)
    return q


# --------------------------------------------------------------------------------------------
# synthetic 64-beam LiDAR sequence (configs 1/2)
# --------------------------------------------------------------------------------------------
class StreetScene:
    """Ground plane + two facades with relief + boxes + poles, ray-cast analytically."""

    def __init__(self, seed: int = 42, length: float = 260.0):
        rng = np.random.default_rng(seed)
        self.ground_z = -1.73
        self.wall_y = (rng.uniform(8.0, 15.0), -rng.uniform(8.0, 15.0))
        nb = 200
        c = np.stack([rng.uniform(-30, length, nb), rng.uniform(-7, 7, nb), np.zeros(nb)], 1)
        c[:, 1] = np.where(np.abs(c[:, 1]) < 2.5, np.sign(c[:, 1] + 1e-3) * (2.5 + np.abs(c[:, 1])), c[:, 1])
        half = np.stack([rng.uniform(1.5, 2.4, nb), rng.uniform(0.7, 1.0, nb), rng.uniform(0.6, 0.9, nb)], 1)
        c[:, 2] = self.ground_z + half[:, 2]
        # relief boxes on the facades
        nr = 120
        rc = np.stack([rng.uniform(-30, length, nr), np.where(rng.random(nr) < 0.5, self.wall_y[0], self.wall_y[1]),
                       rng.uniform(0.0, 5.0, nr)], 1)
        rh = np.stack([rng.uniform(1.0, 3.0, nr), rng.uniform(0.5, 1.5, nr), rng.uniform(1.0, 3.0, nr)], 1)
        self.box_c = np.concatenate([c, rc], 0)
        self.box_h = np.concatenate([half, rh], 0)
        npole = 300
        self.pole_xy = np.stack([rng.uniform(-30, length, npole), rng.choice([-1.0, 1.0], npole) * rng.uniform(5.0, 7.5, npole)], 1)
        self.pole_r = rng.uniform(0.08, 0.2, npole)
        self.pole_top = rng.uniform(3.0, 8.0, npole)

    def raycast(self, origin: np.ndarray, dirs: np.ndarray, max_range: float = 120.0) -> np.ndarray:
        """Ranges [n] along unit directions `dirs` [n,3] from `origin` [3]; inf = no return."""
        if origin.ndim == 2:
            return self.raycast_moving(origin, dirs, max_range)
        n = dirs.shape[0]
        t = np.full(n, np.inf)
        o = origin.astype(np.float64)
        d = dirs.astype(np.float64)
        with np.errstate(divide="ignore", invalid="ignore"):
            # ground
            tg = (self.ground_z - o[2]) / d[:, 2]
            t = np.where((tg > 0) & (tg < t), tg, t)
            # facades (infinite planes y = const, z < 12)
            for wy in self.wall_y:
                tw = (wy - o[1]) / d[:, 1]
                z = o[2] + tw * d[:, 2]
                ok = (tw > 0) & (tw < t) & (z < 12.0)
                t = np.where(ok, tw, t)
            # boxes: slab test, only boxes near the sensor
            near = np.abs(self.box_c[:, 0] - o[0]) < max_range
            inv = 1.0 / d
            for c, h in zip(self.box_c[near], self.box_h[near]):
                t1 = (c - h - o) * inv
                t2 = (c + h - o) * inv
                tmin = np.minimum(t1, t2).max(1)
                tmax = np.maximum(t1, t2).min(1)
                ok = (tmax >= np.maximum(tmin, 0.0)) & (tmin > 0) & (tmin < t)
                t = np.where(ok, tmin, t)
            # poles: vertical cylinders
            nearp = np.abs(self.pole_xy[:, 0] - o[0]) < 60.0
            a = d[:, 0] ** 2 + d[:, 1] ** 2
            for (px, py), r, top in zip(self.pole_xy[nearp], self.pole_r[nearp], self.pole_top[nearp]):
                ox, oy = o[0] - px, o[1] - py
                bq = ox * d[:, 0] + oy * d[:, 1]
                cq = ox * ox + oy * oy - r * r
                disc = bq * bq - a * cq
                tp = (-bq - np.sqrt(np.maximum(disc, 0.0))) / a
                z = o[2] + tp * d[:, 2]
                ok = (disc > 0) & (tp > 0) & (tp < t) & (z < top) & (z > self.ground_z)
                t = np.where(ok, tp, t)
        t[t > max_range] = np.inf
        return t


def _raycast_moving(self, origins: np.ndarray, dirs: np.ndarray, max_range: float = 120.0) -> np.ndarray:
    """The same scene with one origin PER RAY (`origins` [n,3]): a scan taken while the sensor moves."""
    n = dirs.shape[0]
    t = np.full(n, np.inf)
    o = origins.astype(np.float64)
    d = dirs.astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        tg = (self.ground_z - o[:, 2]) / d[:, 2]
        t = np.where((tg > 0) & (tg < t), tg, t)
        for wy in self.wall_y:
            tw = (wy - o[:, 1]) / d[:, 1]
            z = o[:, 2] + tw * d[:, 2]
            ok = (tw > 0) & (tw < t) & (z < 12.0)
            t = np.where(ok, tw, t)
        ox_mid = float(np.median(o[:, 0]))
        near = np.abs(self.box_c[:, 0] - ox_mid) < max_range
        inv = 1.0 / d
        for c, h in zip(self.box_c[near], self.box_h[near]):
            t1 = (c - h - o) * inv
            t2 = (c + h - o) * inv
            tmin = np.minimum(t1, t2).max(1)
            tmax = np.maximum(t1, t2).min(1)
            ok = (tmax >= np.maximum(tmin, 0.0)) & (tmin > 0) & (tmin < t)
            t = np.where(ok, tmin, t)
        nearp = np.abs(self.pole_xy[:, 0] - ox_mid) < 60.0
        a = d[:, 0] ** 2 + d[:, 1] ** 2
        for (px, py), r, top in zip(self.pole_xy[nearp], self.pole_r[nearp], self.pole_top[nearp]):
            ox, oy = o[:, 0] - px, o[:, 1] - py
            bq = ox * d[:, 0] + oy * d[:, 1]
            cq = ox * ox + oy * oy - r * r
            disc = bq * bq - a * cq
            tp = (-bq - np.sqrt(np.maximum(disc, 0.0))) / a
            z = o[:, 2] + tp * d[:, 2]
            ok = (disc > 0) & (tp > 0) & (tp < t) & (z < top) & (z > self.ground_z)
            t = np.where(ok, tp, t)
    t[t > max_range] = np.inf
    return t


StreetScene.raycast_moving = _raycast_moving


def lidar_directions(n_beams: int = 64, n_azimuth: int = 1875) -> np.ndarray:
    elev = np.deg2rad(np.linspace(2.0, -24.8, n_beams))
    azim = np.linspace(-np.pi, np.pi, n_azimuth, endpoint=False)
    e, a = np.meshgrid(elev, azim, indexing="ij")
    return np.stack([np.cos(e) * np.cos(a), np.cos(e) * np.sin(a), np.sin(e)], -1).reshape(-1, 3)


def trajectory(n_frames: int = 100, step: float = 0.9) -> np.ndarray:
    """Ground-truth sensor poses [n,4,4] (float64): +x at `step` m/frame, yaw 0.15 sin(2 pi i / 100)."""
    poses = np.tile(np.eye(4), (n_frames, 1, 1))
    x = y = 0.0
    for i in range(n_frames):
        yaw = 0.15 * np.sin(2 * np.pi * i / 100.0)
        c, s = np.cos(yaw), np.sin(yaw)
        poses[i, :3, :3] = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
        poses[i, :3, 3] = (x, y, 0.0)
        x += step * c
        y += step * s
    return poses


def lidar_scan(scene: StreetScene, pose: np.ndarray, dirs_sensor: np.ndarray, rng, noise: float = 0.02,
               max_range: float = 120.0) -> np.ndarray:
    """One scan in the SENSOR frame, float32 [n,3]; no-return rays dropped."""
    R, o = pose[:3, :3], pose[:3, 3]
    t = scene.raycast(o, dirs_sensor @ R.T, max_range)
    ok = np.isfinite(t)
    r = t[ok] + noise * rng.standard_normal(int(ok.sum()))
    return (dirs_sensor[ok] * r[:, None]).astype(np.float32)


# --------------------------------------------------------------------------------------------
# KITTI-like sequence on disk (configs 0/1 through the reference's own driver): motion-distorted sweeps with the
# Velodyne timestamp convention the reference assumes, the inverse of its intrinsic correction, KITTI file layout
# --------------------------------------------------------------------------------------------
def _se3_log(T):
    R, t = T[:3, :3], T[:3, 3]
    cos = np.clip((np.trace(R) - 1.0) * 0.5, -1.0, 1.0)
    th = np.arccos(cos)
    w = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    if th < 1e-9:
        return np.zeros(3), t.copy()
    return w / (2.0 * np.sin(th)) * th, t.copy()


def _rot_exp(w):
    """Rodrigues for a batch of rotation vectors [n,3] -> [n,3,3]."""
    th = np.linalg.norm(w, axis=1)
    k = w / np.maximum(th, 1e-30)[:, None]
    K = np.zeros((w.shape[0], 3, 3))
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0], K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -k[:, 2], k[:, 1], k[:, 2], -k[:, 0], -k[:, 1], k[:, 0]
    return np.eye(3)[None] + np.sin(th)[:, None, None] * K + (1.0 - np.cos(th))[:, None, None] * (K @ K)


def velodyne_point_ts(dirs_sensor: np.ndarray) -> np.ndarray:
    """Per-point time in [0, 1) of a sweep as the reference derives it for KITTI scans without timestamps
    (dataset/slam_dataset.py:247-249: clockwise from the -x axis)."""
    yaw = -np.arctan2(dirs_sensor[:, 1], dirs_sensor[:, 0])
    return 0.5 * (yaw / np.pi + 1.0)


def lidar_sweep(scene: StreetScene, pose_prev, pose, pose_next, dirs_sensor: np.ndarray, rng, noise: float = 0.02,
                max_range: float = 120.0, correct_deg: float = 0.0) -> np.ndarray:
    """One MOTION-DISTORTED sweep in the sensor frame (float32 [n,3]): ray j is cast at time tau_j = ts_j - 0.5 of the
    frame from the pose  pose . exp(tau_j log(M)),  M = the motion of the preceding (tau < 0) / following (tau > 0) frame
    interval (rotation by slerp, translation linear -- exactly the model utils/tools.py:540-567 undoes), and recorded in the
    sensor frame AT THAT TIME.  `correct_deg`: apply the inverse of the reference's KITTI intrinsic correction
    (dataset/slam_dataset.py:837-854) so that its corrected cloud is geometrically exact."""
    tau = velodyne_point_ts(dirs_sensor) - 0.5
    Mp = np.linalg.inv(pose_prev) @ pose if pose_prev is not None else np.linalg.inv(pose) @ pose_next
    Mn = np.linalg.inv(pose) @ pose_next if pose_next is not None else Mp
    wp, tp = _se3_log(Mp)
    wn, tn = _se3_log(Mn)
    neg = (tau < 0)[:, None]
    w = np.where(neg, wp[None], wn[None]) * tau[:, None]
    tr = np.where(neg, tp[None], tn[None]) * tau[:, None]
    Rt = _rot_exp(w)                                          # sensor(tau) -> sensor(mid)
    Rw = pose[:3, :3][None] @ Rt                              # sensor(tau) -> world
    ow = (pose[:3, :3] @ tr.T).T + pose[:3, 3][None]
    dw = np.einsum("nij,nj->ni", Rw, dirs_sensor)
    t = scene.raycast_moving(ow, dw, max_range)
    ok = np.isfinite(t)
    r = t[ok] + noise * rng.standard_normal(int(ok.sum()))
    pts = dirs_sensor[ok] * r[:, None]
    if correct_deg != 0.0:
        dist = np.linalg.norm(pts, axis=1)
        v = np.arcsin(pts[:, 2] / dist)
        v2 = v - np.deg2rad(correct_deg)
        sc = np.cos(v2) / np.cos(v)
        pts = np.stack([pts[:, 0] * sc, pts[:, 1] * sc, dist * np.sin(v2)], 1)
    return pts.astype(np.float32)


def _sweep_job(job):
    seed, pp, p, pn, beams, azimuth, noise_seed, correct_deg = job
    return lidar_sweep(StreetScene(seed), pp, p, pn, lidar_directions(beams, azimuth), np.random.default_rng(noise_seed),
                       correct_deg=correct_deg)


def make_sweeps(n_frames: int, beams: int = 64, azimuth: int = 1875, workers: int = 1, correct_deg: float = 0.0, seed: int = 42):
    """(list of motion-distorted sweeps, ground-truth poses [n,4,4]); per-frame noise seeds 1000 + i."""
    gt = trajectory(n_frames + 1)
    jobs = [(seed, gt[i - 1] if i > 0 else None, gt[i], gt[i + 1], beams, azimuth, 1000 + i, correct_deg) for i in range(n_frames)]
    if workers > 1:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(workers) as pool:
            frames = pool.map(_sweep_job, jobs)
    else:
        frames = [_sweep_job(j) for j in jobs]
    return frames, gt[:n_frames]


def write_kitti_sequence(root: str, frames, gt_poses) -> dict:
    """KITTI odometry layout the reference's readers expect (dataset/slam_dataset.py:631-646, 721-764):
    sequences/00/velodyne/%06d.bin (float32 [n,4]: x y z intensity), poses/00.txt (12 floats per row), calib.txt (Tr = I)."""
    import os
    vel = os.path.join(root, "sequences", "00", "velodyne")
    os.makedirs(vel, exist_ok=True)
    os.makedirs(os.path.join(root, "poses"), exist_ok=True)
    for i, f in enumerate(frames):
        np.concatenate([f[:, :3], np.zeros((f.shape[0], 1), np.float32)], 1).astype(np.float32).tofile(os.path.join(vel, "%06d.bin" % i))
    with open(os.path.join(root, "poses", "00.txt"), "w") as fh:
        for p in gt_poses:
            fh.write(" ".join("%.9e" % v for v in p[:3, :].reshape(-1)) + "\n")
    calib = os.path.join(root, "sequences", "00", "calib.txt")
    with open(calib, "w") as fh:
        ident = "1 0 0 0 0 1 0 0 0 0 1 0"
        for k in ("P0", "P1", "P2", "P3", "Tr"):
            fh.write("%s: %s\n" % (k, ident))
    return {"pc_path": vel, "pose_path": os.path.join(root, "poses", "00.txt"), "calib_path": calib}
