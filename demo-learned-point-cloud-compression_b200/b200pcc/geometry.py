"""Lossless coder for the stride-8 latent coordinates (the slot the reference fills with a G-PCC
``tmc3`` subprocess + temp PLY files: shared/utils.py:169-240, codec_pipeline.py:441-462,
codec_parallel.py:266-289).  G-PCC is out of scope; this is the minimal
stand-in that keeps the container drop-in: a breadth-first octree occupancy-byte stream over the
coordinates divided by 8, exactly the quantity the reference hands to tmc3 (shared/utils.py:173).
Ny is ~1.5 % of the input points, so this runs on the host in vectorised numpy.
"""
import struct

import numpy as np

_MAGIC = b"OCT1"


def _spread3(v: np.ndarray) -> np.ndarray:
    v = v.astype(np.uint64) & np.uint64(0xFFFF)
    v = (v | (v << np.uint64(16))) & np.uint64(0x001F0000FF0000FF)
    v = (v | (v << np.uint64(8))) & np.uint64(0x100F00F00F00F00F)
    v = (v | (v << np.uint64(4))) & np.uint64(0x10C30C30C30C30C3)
    v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
    return v


def _compact3(x: np.ndarray) -> np.ndarray:
    x = x & np.uint64(0x1249249249249249)
    x = (x | (x >> np.uint64(2))) & np.uint64(0x10C30C30C30C30C3)
    x = (x | (x >> np.uint64(4))) & np.uint64(0x100F00F00F00F00F)
    x = (x | (x >> np.uint64(8))) & np.uint64(0x001F0000FF0000FF)
    x = (x | (x >> np.uint64(16))) & np.uint64(0x001F00000000FFFF)
    return x & np.uint64(0xFFFF)


def _unique_sorted(a: np.ndarray) -> np.ndarray:
    if a.size == 0:
        return a
    keep = np.ones(a.shape[0], dtype=bool)
    keep[1:] = a[1:] != a[:-1]
    return a[keep]


def encode_points(points: np.ndarray, stride: int = 8) -> bytes:
    """points: int [n,3] multiples of `stride` (one frame).  Returns the octree stream."""
    p = np.asarray(points, dtype=np.int64)
    n = p.shape[0]
    if n == 0:
        return _MAGIC + struct.pack("<iBiii", 0, 0, 0, 0, 0)
    q = np.floor_divide(p, stride)
    mn = q.min(axis=0)
    q = q - mn
    depth = max(int(q.max()).bit_length(), 1)
    m = _spread3(q[:, 0]) | (_spread3(q[:, 1]) << np.uint64(1)) | (_spread3(q[:, 2]) << np.uint64(2))
    m = _unique_sorted(np.sort(m))
    out = [_MAGIC, struct.pack("<iBiii", int(m.shape[0]), depth, int(mn[0]), int(mn[1]), int(mn[2]))]
    for level in range(depth):
        shift = np.uint64(3 * (depth - level - 1))
        child = _unique_sorted(m >> shift)           # nodes of level+1, ascending
        parent = child >> np.uint64(3)
        bits = (np.uint64(1) << (child & np.uint64(7))).astype(np.uint8)
        start = np.ones(child.shape[0], dtype=bool)
        start[1:] = parent[1:] != parent[:-1]
        occ = np.bitwise_or.reduceat(bits, np.nonzero(start)[0])
        out.append(occ.astype(np.uint8).tobytes())
    return b"".join(out)


_BIT_TABLE = np.array([[(b >> j) & 1 for j in range(8)] for b in range(256)], dtype=bool)


def decode_points(stream: bytes, stride: int = 8) -> np.ndarray:
    """-> int32 [n,3] in Morton order (multiples of `stride`)."""
    if stream[:4] != _MAGIC:
        raise ValueError("bad geometry stream")
    n, depth, mx, my, mz = struct.unpack_from("<iBiii", stream, 4)
    pos = 4 + struct.calcsize("<iBiii")
    if n == 0:
        return np.zeros((0, 3), dtype=np.int32)
    nodes = np.zeros(1, dtype=np.uint64)
    for _ in range(depth):
        cnt = nodes.shape[0]
        if pos + cnt > len(stream):
            raise ValueError("truncated geometry stream")
        occ = np.frombuffer(stream, dtype=np.uint8, count=cnt, offset=pos)
        pos += cnt
        mask = _BIT_TABLE[occ]                       # [cnt, 8]
        idx_node, idx_child = np.nonzero(mask)        # row-major: node ascending, child ascending
        nodes = (nodes[idx_node] << np.uint64(3)) | idx_child.astype(np.uint64)
    if nodes.shape[0] != n:
        raise ValueError("geometry stream point count mismatch")
    x = _compact3(nodes).astype(np.int64) + mx
    y = _compact3(nodes >> np.uint64(1)).astype(np.int64) + my
    z = _compact3(nodes >> np.uint64(2)).astype(np.int64) + mz
    return (np.stack([x, y, z], axis=1) * stride).astype(np.int32)
