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
