"""Oracle (test infrastructure): the capturer's frame preparation, restated in numpy.

Follows sender/capturer/capturer.py:79-128 (`Capturer.get_zed_frames` after `retrieve_measure`): the
XYZRGBA measure [h, w, 4] float32 is split into points and the three low bytes of the packed colour
float (:84-90), points that are not finite or farther than `depth_clip` are dropped (:93-96), the cloud
is voxel-down-sampled with Open3D in float64 (:99-106), the averaged points are rounded to the voxel grid
and cast to int16 (:107), rows that collide after rounding are dropped with `np.unique(..., axis=0,
return_index=True)` (:110-112) and, with `max_points`, only the rows with the largest z are kept (:117-120).

Open3D is not vendored under the reference tree (sender/capturer/requirements.txt pins `open3d`, no
version) and cannot be installed here, so `voxel_down_sample` is restated from its published algorithm
[UPSTREAM-RECALL, open3d/geometry/PointCloud.cpp `PointCloud::VoxelDownSample`]:

    voxel_min_bound = min over points (per axis) - voxel_size * 0.5
    voxel index     = floor((p - voxel_min_bound) / voxel_size)   per axis, float64
    each voxel accumulates point and colour with `+=` in input order and returns sum / count

Two things the reference leaves to chance are fixed here, and the CUDA path follows the same choice:
  * Open3D returns the voxels in `std::unordered_map` iteration order.  That order decides which colour
    survives when two voxels round to the same int16 cell (np.unique keeps the first).  Here the voxels
    come out in ascending (ix, iy, iz) order, so the voxel with the smallest index triple wins.
  * `np.argpartition` leaves the choice among rows tied at the z threshold, and the order of the kept
    rows, to the implementation.  Here ties go to the rows that come first in the (x, y, z) order of
    np.unique, and the kept rows stay in that order (the encoder sorts its input anyway,
    sender/encoder/codec_pipeline.py:262-266).

PARITY UNPINNED: the reference holds no recorded camera measure next to its recorded frames, so the
restatement is checked against a second, scalar statement of the same rules (`voxel_down_sample_scalar`)
and against the properties the recorded frames show (int16, unique rows, colours in [0, 1]).
"""
import numpy as np


def split_measure(data):
    """capturer.py:84-90: XYZRGBA float32 [h, w, 4] (or [n, 4]) -> points f32 [n,3], colours u8-valued [n,3]."""
    data = np.ascontiguousarray(data, dtype=np.float32).reshape(-1, 4)
    points = data[:, :3].reshape(-1, 3)
    float_colors = np.ascontiguousarray(data[:, 3]).reshape(-1, 1)
    int_colors = float_colors.view(np.uint32)
    colors = np.stack([((int_colors >> (8 * i)) & 0xFF) for i in range(3)], axis=-1).reshape(-1, 3)
    return points, colors


def valid_mask(points, depth_clip):
    """capturer.py:93-94.  The norm of a float32 array is evaluated in float32 and compared with the
    Python float as a float32 (NumPy 2 promotion)."""
    with np.errstate(invalid="ignore", over="ignore"):
        distances = np.linalg.norm(points, axis=1)
        return np.isfinite(points).all(axis=1) & (distances <= np.float32(depth_clip))


def voxel_indices(points64, voxel_size):
    voxel_min_bound = points64.min(axis=0) - np.float64(voxel_size) * 0.5
    return np.floor((points64 - voxel_min_bound) / np.float64(voxel_size)).astype(np.int64)


def voxel_down_sample(points64, colors64, voxel_size):
    """Open3D PointCloud::VoxelDownSample, voxels in ascending (ix, iy, iz) order.  The per-voxel sums
    are taken with `+=` in input order (element j of every voxel is added in step j), not with numpy's
    pairwise reduction."""
    n = points64.shape[0]
    if n == 0:
        return np.zeros((0, 3)), np.zeros((0, 3))
    idx = voxel_indices(points64, voxel_size)
    order = np.lexsort((idx[:, 2], idx[:, 1], idx[:, 0]))       # stable: input order inside a voxel
    sidx = idx[order]
    head = np.ones(n, bool)
    head[1:] = (sidx[1:] != sidx[:-1]).any(axis=1)
    start = np.flatnonzero(head)
    count = np.diff(np.append(start, n))
    vals = np.concatenate([points64, colors64], axis=1)[order]
    acc = np.zeros((len(start), 6))
    for j in range(int(count.max())):
        live = count > j
        acc[live] += vals[start[live] + j]
    acc /= count[:, None].astype(np.float64)
    return acc[:, :3], acc[:, 3:]


def voxel_down_sample_scalar(points64, colors64, voxel_size):
    """The same rules as a dictionary of accumulators, one point at a time (small inputs only)."""
    idx = voxel_indices(points64, voxel_size) if len(points64) else np.zeros((0, 3), np.int64)
    cells = {}
    for i in range(points64.shape[0]):
        key = (int(idx[i, 0]), int(idx[i, 1]), int(idx[i, 2]))
        a = cells.get(key)
        if a is None:
            cells[key] = [points64[i].copy(), colors64[i].copy(), 1]
        else:
            a[0] += points64[i]
            a[1] += colors64[i]
            a[2] += 1
    keys = sorted(cells)
    pts = np.array([cells[k][0] / float(cells[k][2]) for k in keys]).reshape(-1, 3)
    col = np.array([cells[k][1] / float(cells[k][2]) for k in keys]).reshape(-1, 3)
    return pts, col


def select_largest_z(points, max_points):
    """capturer.py:117-120 with the tie rule of the header: rows of the `max_points` largest z, ties to
    the earlier row, result in row order."""
    if max_points is None or points.shape[0] <= max_points:
        return np.arange(points.shape[0])
    order = np.argsort(-points[:, 2].astype(np.int64), kind="stable")[:max_points]
    return np.sort(order)


def capture_frame(data, voxel_size=0.01, depth_clip=-1.5, max_points=None, down_sample=voxel_down_sample):
    """XYZRGBA measure -> {"points": int16 [m,3], "colors": float64 [m,3]} (capturer.py:79-124)."""
    points, colors = split_measure(data)
    mask = valid_mask(points, depth_clip)
    points = points[mask]
    colors = colors[mask]
    p, c = down_sample(points.astype(np.float64), colors.astype(np.float64) / 255.0, voxel_size)
    p = np.round(p / voxel_size).astype(np.int16)
    if p.shape[0]:
        _, unique_indices = np.unique(p, axis=0, return_index=True)
        p = p[unique_indices]
        c = c[unique_indices]
    keep = select_largest_z(p, max_points)
    return {"points": p[keep].reshape(-1, 3), "colors": c[keep].reshape(-1, 3)}


def synth_measure(h=480, w=640, seed=0, holes=0.03, z_near=0.45, z_far=2.2):
    """A ZED-shaped XYZRGBA measure (sender/capturer/capturer.py:65-67, 81: 640x480, metres, right-handed
    y-up, camera looking down -z): smooth background surface + a nearer blob with step edges, NaN / inf
    holes as the SDK marks occlusions, packed RGBA colour floats from a procedural texture."""
    rng = np.random.default_rng(seed)
    v, u = np.mgrid[0:h, 0:w].astype(np.float64)
    fx = fy = 0.82 * w
    cx, cy = (w - 1) / 2.0, (h - 1) / 2.0
    depth = z_far - 0.5 * np.sin(u / w * 2.1 + 0.3) * np.cos(v / h * 1.7) - 0.25 * (v / h)
    blob = ((u - cx * 0.9) / (0.22 * w)) ** 2 + ((v - cy * 1.05) / (0.38 * h)) ** 2
    depth = np.where(blob < 1.0, z_near + 0.35 * np.sqrt(np.clip(1.0 - blob, 0, 1)) * -1.0 + 0.55, depth)
    depth = depth + rng.normal(0.0, 0.0008, size=depth.shape)
    x = (u - cx) / fx * depth
    y = -(v - cy) / fy * depth
    z = -depth
    r = (127.5 + 127.5 * np.sin(u / 23.0)).astype(np.uint32)
    g = (127.5 + 127.5 * np.cos(v / 17.0)).astype(np.uint32)
    b = ((u + v) % 256).astype(np.uint32)
    rgba = (r | (g << 8) | (b << 16) | (np.uint32(255) << 24)).astype(np.uint32)
    data = np.empty((h, w, 4), np.float32)
    data[..., 0] = x
    data[..., 1] = y
    data[..., 2] = z
    data[..., 3] = rgba.view(np.float32) if rgba.flags.c_contiguous else np.ascontiguousarray(rgba).view(np.float32)
    hole = rng.random((h, w)) < holes
    kind = rng.integers(0, 3, size=(h, w))
    data[hole & (kind == 0), :3] = np.nan
    data[hole & (kind == 1), 2] = -np.inf
    data[hole & (kind == 2), 0] = np.inf
    return data
