"""Oracle (test infrastructure): the "OCT2" lossless coordinate stream, restated in numpy.

The reference codes the stride-8 latent coordinates of each frame with a G-PCC `tmc3` subprocess over
temporary PLY files (shared/utils.py:169-240 `gpcc_encode` / `gpcc_decode`, called at
sender/encoder/codec_pipeline.py:441-462 and receiver/decoder/codec_parallel.py:266-289).  tmc3 is an
external binary that is not in the reference tree, so its bitstream cannot be reproduced; what the
reference fixes is the contract -- a per-frame byte string from which exactly the same coordinate set
comes back (shared/utils.py:173 divides by 8 before coding, :235 multiplies back).  The product fills
that slot with a breadth-first octree occupancy stream built on the device
(csrc/octree.cu); this file is the independent scalar/numpy statement of the same format:

    "OCT2" | u32 n | u32 depth | u64 root | u32 nodes[depth] | occupancy bytes of level 0 .. depth-1

with m = morton3(x + 32768, y + 32768, z + 32768) >> 3*log2_stride for every point,
depth = ceil(bits(min m XOR max m) / 3), root = m >> 3*depth, and bit c of a node's byte set iff its
child with octant c = dx + 2dy + 4dz exists.  Nodes of a level are in ascending Morton order.
"""
import struct

import numpy as np

from . import sparse_ref as S

MAGIC = b"OCT2"


def _morton(points, log2_stride):
    C = np.concatenate([np.zeros((len(points), 1), np.int64), np.asarray(points, np.int64)], axis=1)
    m = S.morton_key(C) & np.uint64(0x0000FFFFFFFFFFFF)
    return np.unique(m >> np.uint64(3 * log2_stride))


def encode_points(points, log2_stride=3) -> bytes:
    """points: int [n,3] (one frame; multiples of 2**log2_stride, any order, duplicates dropped)."""
    m = _morton(points, log2_stride)
    n = int(m.shape[0])
    if n == 0:
        return MAGIC + struct.pack("<IIQ", 0, 0, 0)
    x = int(m[0]) ^ int(m[-1])
    depth = (x.bit_length() + 2) // 3
    root = int(m[0]) >> (3 * depth)
    levels, counts = [], []
    for l in range(depth):
        child = np.unique(m >> np.uint64(3 * (depth - l - 1)))       # nodes of level l+1
        parent = child >> np.uint64(3)
        occ = np.zeros(0, np.uint8)
        uniq, start = np.unique(parent, return_index=True)
        bits = (np.uint64(1) << (child & np.uint64(7))).astype(np.uint8)
        occ = np.bitwise_or.reduceat(bits, start)
        levels.append(occ.tobytes())
        counts.append(len(occ))
    return MAGIC + struct.pack("<IIQ", n, depth, root) + struct.pack(f"<{depth}I", *counts) + b"".join(levels)


def decode_points(stream: bytes, log2_stride=3):
    """-> int32 [n,3] in ascending Morton order."""
    if len(stream) < 20 or stream[:4] != MAGIC:
        raise ValueError("bad OCT2 stream")
    n, depth, root = struct.unpack_from("<IIQ", stream, 4)
    if depth > 16 or len(stream) < 20 + 4 * depth:
        raise ValueError("bad OCT2 header")
    counts = struct.unpack_from(f"<{depth}I", stream, 20)
    pos = 20 + 4 * depth
    nodes = [root] if n else []
    for l in range(depth):
        if counts[l] != len(nodes) or pos + counts[l] > len(stream):
            raise ValueError("inconsistent OCT2 stream")
        nxt = []
        for node, occ in zip(nodes, stream[pos:pos + counts[l]]):     # scalar on purpose
            if occ == 0:
                raise ValueError("empty OCT2 node")
            for c in range(8):
                if occ >> c & 1:
                    nxt.append(node << 3 | c)
        pos += counts[l]
        nodes = nxt
    if len(nodes) != n:
        raise ValueError("OCT2 point count mismatch")
    m = np.asarray(nodes, dtype=np.uint64) << np.uint64(3 * log2_stride)

    def compact(v):
        v = v & np.uint64(0x1249249249249249)
        v = (v | (v >> np.uint64(2))) & np.uint64(0x10C30C30C30C30C3)
        v = (v | (v >> np.uint64(4))) & np.uint64(0x100F00F00F00F00F)
        v = (v | (v >> np.uint64(8))) & np.uint64(0x001F0000FF0000FF)
        v = (v | (v >> np.uint64(16))) & np.uint64(0x001F00000000FFFF)
        return (v & np.uint64(0xFFFF)).astype(np.int64) - 32768

    if n == 0:
        return np.zeros((0, 3), np.int32)
    return np.stack([compact(m), compact(m >> np.uint64(1)), compact(m >> np.uint64(2))], axis=1).astype(np.int32)
