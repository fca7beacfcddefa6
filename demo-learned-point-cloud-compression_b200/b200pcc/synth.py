"""Seeded synthetic frames shaped like the BASELINE.json configurations.

Frames have the capturer's layout (sender/capturer/capturer.py:79-134; sampled into a GOP at
sender/encoder/encoder.py:95-129): ``{"points": int16 [N,3] unique voxels, "colors": float64 [N,3] in
[0,1]}``.  numpy only: tests, bench.py and the CPU baseline all draw their inputs from here.
"""
import numpy as np


def _colors(points: np.ndarray, rng) -> np.ndarray:
    p = points.astype(np.float64)
    span = np.maximum(p.max(axis=0) - p.min(axis=0), 1.0)
    u = (p - p.min(axis=0)) / span
    r = 0.5 + 0.4 * np.sin(2.1 * np.pi * u[:, 0]) * np.cos(1.3 * np.pi * u[:, 2])
    g = 0.2 + 0.6 * u[:, 1] * (1 - 0.5 * u[:, 0])
    b = 0.5 + 0.45 * np.cos(1.7 * np.pi * (u[:, 0] + u[:, 1] + u[:, 2]))
    c = np.stack([r, g, b], axis=1) + rng.normal(0, 0.02, (p.shape[0], 3))
    return np.clip(c, 0.0, 1.0)


def _unique_voxels(p: np.ndarray) -> np.ndarray:
    v = np.floor(p).astype(np.int64)
    key = ((v[:, 0] + 32768) << 32) | ((v[:, 1] + 32768) << 16) | (v[:, 2] + 32768)
    _, idx = np.unique(key, return_index=True)
    return v[np.sort(idx)]


def _ellipsoid_shell(rng, centre, radii, n):
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.asarray(centre, float) + d * np.asarray(radii, float)


def _area(radii):
    a, b, c = radii
    p = 1.6075
    return 4 * np.pi * (((a * b) ** p + (a * c) ** p + (b * c) ** p) / 3) ** (1 / p)


def _humanoid(rng, scale, centre=(512, 512, 512), density=5.0):
    parts = [  # (offset, radii) in units of `scale`
        ((0, 0, 40), (95, 60, 150)),      # torso
        ((0, 0, 235), (48, 52, 58)),      # head
        ((-128, 0, 70), (28, 28, 135)),   # arms
        ((128, 0, 70), (28, 28, 135)),
        ((-45, 0, -230), (38, 40, 170)),  # legs
        ((45, 0, -230), (38, 40, 170)),
    ]
    pts = []
    for off, rad in parts:
        rad = tuple(r * scale for r in rad)
        n = int(_area(rad) * density) + 16
        pts.append(_ellipsoid_shell(rng, np.asarray(centre) + np.asarray(off) * scale, rad, n))
    return np.concatenate(pts)


def frame_8ivfb(seed: int = 1, n_target: int = 800_000, grid: int = 1024):
    """cfg-2: humanoid-like closed surface, ~n_target unique voxels in [0, grid)^3."""
    rng = np.random.default_rng(seed)
    scale = 1.0 * (n_target / 800_000) ** 0.5 * (grid / 1024)
    v = None
    for _ in range(4):
        rs = np.random.default_rng(seed)
        v = _unique_voxels(_humanoid(rs, scale, centre=(grid // 2,) * 3))
        v = v[np.all((v >= 0) & (v < grid), axis=1)]
        ratio = n_target / max(len(v), 1)
        if abs(ratio - 1) < 0.01:
            break
        scale *= ratio ** 0.5
    return {"points": v.astype(np.int16), "colors": _colors(v, rng)}


def frame_cpu100k(seed: int = 0, n: int = 100_100, grid: int = 1024):
    """cfg-1: 3 ellipsoid shells + 2 planes, truncated to n voxels (tests/test_pipeline.py:21-22 size)."""
    rng = np.random.default_rng(seed)
    pts = [
        _ellipsoid_shell(rng, (400, 500, 520), (150, 110, 90), 700_000),
        _ellipsoid_shell(rng, (640, 520, 480), (90, 120, 140), 600_000),
        _ellipsoid_shell(rng, (520, 380, 600), (70, 60, 110), 300_000),
    ]
    u = rng.uniform(0, 1, (200_000, 2))
    pts.append(np.stack([300 + 400 * u[:, 0], 300 + 400 * u[:, 1], np.full(len(u), 330.0)], axis=1))
    pts.append(np.stack([300 + 400 * u[:, 0], np.full(len(u), 700.0), 330 + 300 * u[:, 1]], axis=1))
    v = _unique_voxels(np.concatenate(pts))
    v = v[np.all((v >= 0) & (v < grid), axis=1)]
    sel = rng.permutation(len(v))[:n]
    v = v[np.sort(sel)]
    return {"points": v.astype(np.int16), "colors": _colors(v, rng)}


def frame_zed(seed: int = 2, n_target: int = 300_000, t: float = 0.0):
    """cfg-3: depth-derived single sheet with step edges; coordinates include negatives like the
    recorded ZED frames (evaluation/data/*/frame_*.pkl: x in [-149,85], z in [-354,-102])."""
    rng = np.random.default_rng(seed)
    w = int(round((n_target * 640 / 480) ** 0.5))
    h = int(round(n_target / w))
    xs, ys = np.meshgrid(np.arange(w) - w // 2, np.arange(h) - h // 2, indexing="xy")
    u, v = xs / w, ys / h
    depth = -230 + 60 * np.sin(2.2 * np.pi * (u + 0.05 * t)) * np.cos(1.4 * np.pi * v) + 25 * np.cos(5 * np.pi * v + t)
    depth = depth - 70 * (((u - 0.1 * np.sin(t)) ** 2 + (v + 0.05) ** 2) < 0.04)   # foreground blob (step edge)
    depth = depth - 40 * (np.abs(u + 0.3) < 0.05)                                   # a pole
    p = np.stack([xs.ravel(), ys.ravel(), np.floor(depth).ravel()], axis=1).astype(np.float64)
    vox = _unique_voxels(p)
    return {"points": vox.astype(np.int16), "colors": _colors(vox, rng)}


def small_frame(seed: int = 0, n_target: int = 20_000, offset=(0, 0, 0)):
    """Small surface patch for tests (two shells + a plane), coordinates straddle zero."""
    rng = np.random.default_rng(seed)
    s = (n_target / 20_000) ** 0.5
    pts = [
        _ellipsoid_shell(rng, (10, -5, 20), (28 * s, 22 * s, 34 * s), int(60_000 * s * s)),
        _ellipsoid_shell(rng, (-30, 25, -10), (16 * s, 18 * s, 14 * s), int(20_000 * s * s)),
    ]
    u = rng.uniform(-1, 1, (int(30_000 * s * s), 2))
    pts.append(np.stack([45 * s * u[:, 0], 45 * s * u[:, 1], np.full(len(u), -40.0 * s)], axis=1))
    v = _unique_voxels(np.concatenate(pts) + np.asarray(offset, float))
    return {"points": v.astype(np.int16), "colors": _colors(v, rng)}


def make_gop(frames, timestamps=None):
    """GOP dict as built by sender/encoder/encoder.py:95-129."""
    return {"frames": list(frames), "timestamps": dict(timestamps or {})}
