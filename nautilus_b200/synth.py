"""Seeded synthetic 2D lidar scans.

There is no bag data in this environment, so every test, the oracle checks and
bench.py share this generator.  A "world" is a random rectilinear floor plan;
a scan is a 1080-beam, 270 degree sweep ray-cast from a pose inside it, with
Gaussian range noise, returned as float32 (x, y) points in the sensor frame --
the same thing the reference's LaserScanToPointCloud hands to the matcher
(/root/reference/src/input/pointcloud_helpers.cc:28-48).

numpy only: this module is host-side input synthesis, not part of the hot path.
"""
from __future__ import annotations

import numpy as np

N_BEAMS = 1080
FOV = np.deg2rad(270.0)
ANGLE_MIN = -0.75 * np.pi
MAX_RANGE = 30.0
RANGE_NOISE = 0.01


def make_world(seed: int) -> np.ndarray:
    """Wall segments of one floor plan as float64 [n, 4] = (x0, y0, x1, y1)."""
    rng = np.random.default_rng(1234 + int(seed))
    w = rng.uniform(8.0, 25.0)
    h = rng.uniform(6.0, 20.0)
    x0, x1, y0, y1 = -w / 2, w / 2, -h / 2, h / 2
    segs = [(x0, y0, x1, y0), (x1, y0, x1, y1), (x1, y1, x0, y1), (x0, y1, x0, y0)]
    for _ in range(int(rng.integers(2, 9))):
        if rng.random() < 0.5:  # horizontal interior wall
            y = rng.uniform(y0 + 1.0, y1 - 1.0)
            a = rng.uniform(x0, x1 - 1.0)
            b = min(x1, a + rng.uniform(1.0, 0.5 * w))
            segs.append((a, y, b, y))
        else:  # vertical interior wall
            x = rng.uniform(x0 + 1.0, x1 - 1.0)
            a = rng.uniform(y0, y1 - 1.0)
            b = min(y1, a + rng.uniform(1.0, 0.5 * h))
            segs.append((x, a, x, b))
    return np.asarray(segs, dtype=np.float64)


def world_extent(world: np.ndarray) -> tuple[float, float]:
    """Half width / half height of the outer box (first four segments)."""
    return float(np.abs(world[:4, [0, 2]]).max()), float(np.abs(world[:4, [1, 3]]).max())


def cast_scan(world: np.ndarray, pose, rng: np.random.Generator,
              n_beams: int = N_BEAMS) -> np.ndarray:
    """Ray-cast one scan from pose (x, y, theta); float32 [n, 2] sensor-frame points."""
    px, py, pth = pose
    ang = pth + ANGLE_MIN + np.arange(n_beams) * (FOV / n_beams)
    dx, dy = np.cos(ang)[:, None], np.sin(ang)[:, None]
    ax, ay = world[:, 0][None, :], world[:, 1][None, :]
    ex, ey = world[:, 2][None, :] - ax, world[:, 3][None, :] - ay
    denom = dx * ey - dy * ex
    with np.errstate(divide="ignore", invalid="ignore"):
        t = ((ax - px) * ey - (ay - py) * ex) / denom      # distance along the beam
        u = ((ax - px) * dy - (ay - py) * dx) / denom      # position along the wall
    hit = (np.abs(denom) > 1e-12) & (t > 1e-3) & (u >= 0.0) & (u <= 1.0)
    t = np.where(hit, t, np.inf)
    r = t.min(axis=1)
    r = r + rng.normal(0.0, RANGE_NOISE, size=r.shape)
    keep = np.isfinite(r) & (r < MAX_RANGE) & (r > 0.05)
    a = (ANGLE_MIN + np.arange(n_beams) * (FOV / n_beams))[keep]
    r = r[keep]
    pts = np.stack([r * np.cos(a), r * np.sin(a)], axis=1)
    return np.ascontiguousarray(pts, dtype=np.float32)


def free_pose(world: np.ndarray, rng: np.random.Generator, margin: float = 1.0):
    hw, hh = world_extent(world)
    return (rng.uniform(-hw + margin, hw - margin),
            rng.uniform(-hh + margin, hh - margin),
            rng.uniform(-np.pi, np.pi))


def compose(pose, dx, dy, dth):
    """pose o (dx, dy, dth): offset expressed in the pose's own frame."""
    x, y, th = pose
    c, s = np.cos(th), np.sin(th)
    return (x + c * dx - s * dy, y + s * dx + c * dy, th + dth)


def make_pair(seed: int, trans_range: float = 2.0, max_rot: float = 0.4,
              n_beams: int = N_BEAMS):
    """One (query a, target b) pair; returns (a, b, truth) with truth the
    transform that maps a's points onto b's frame, (x, y, theta)."""
    world = make_world(seed)
    rng = np.random.default_rng(99991 + int(seed))
    pose_b = free_pose(world, rng)
    dx, dy = rng.uniform(-0.8 * trans_range, 0.8 * trans_range, size=2)
    dth = rng.uniform(-max_rot, max_rot)
    hw, hh = world_extent(world)
    pose_a = compose(pose_b, dx, dy, dth)
    pose_a = (float(np.clip(pose_a[0], -hw + 0.3, hw - 0.3)),
              float(np.clip(pose_a[1], -hh + 0.3, hh - 0.3)), pose_a[2])
    b = cast_scan(world, pose_b, rng, n_beams)
    a = cast_scan(world, pose_a, rng, n_beams)
    # a's frame expressed in b's frame
    cb, sb = np.cos(pose_b[2]), np.sin(pose_b[2])
    ddx, ddy = pose_a[0] - pose_b[0], pose_a[1] - pose_b[1]
    truth = (cb * ddx + sb * ddy, -sb * ddx + cb * ddy, pose_a[2] - pose_b[2])
    return a, b, truth


def make_scan_set(n_scans: int, group: int = 5, seed: int = 0,
                  trans_range: float = 2.0, max_rot: float = 0.2,
                  n_beams: int = N_BEAMS, return_poses: bool = False):
    """n_scans scans in groups of `group` taken near one base pose of one world
    (the shape of a loop-closure candidate list: several revisits of a place).

    Returns (xy, offsets): concatenated float32 [sum n_i, 2] and int64 [n_scans+1].
    Scans inside a group are within 0.8*trans_range and 2*max_rot of each other.
    """
    clouds, poses = [], []
    for g in range((n_scans + group - 1) // group):
        world = make_world(seed * 1000003 + g)
        rng = np.random.default_rng(7777 + seed * 1000003 + g)
        base = free_pose(world, rng, margin=1.5)
        hw, hh = world_extent(world)
        for _ in range(group):
            if len(clouds) == n_scans:
                break
            dx, dy = rng.uniform(-0.4 * trans_range, 0.4 * trans_range, size=2)
            p = compose(base, dx, dy, rng.uniform(-max_rot, max_rot))
            p = (float(np.clip(p[0], -hw + 0.3, hw - 0.3)),
                 float(np.clip(p[1], -hh + 0.3, hh - 0.3)), p[2])
            clouds.append(cast_scan(world, p, rng, n_beams))
            poses.append(p)
    offsets = np.zeros(n_scans + 1, dtype=np.int64)
    offsets[1:] = np.cumsum([len(c) for c in clouds])
    xy = np.ascontiguousarray(np.concatenate(clouds, axis=0))
    if return_poses:
        return xy, offsets, np.asarray(poses, dtype=np.float64)
    return xy, offsets


def make_pair_list(n_scans: int, group: int = 5):
    """All ordered (a, b) pairs inside each group, self-pairs included:
    group**2 pairs per group => n_scans*group pairs; each target reused `group`
    times.  Sorted by target so pairs sharing a table are adjacent."""
    a, b = [], []
    for s in range(n_scans):
        g0 = (s // group) * group
        for q in range(g0, min(g0 + group, n_scans)):
            a.append(q)
            b.append(s)
    return np.asarray(a, dtype=np.int32), np.asarray(b, dtype=np.int32)
