"""Lossless coder for the stride-8 latent coordinates -- the slot the reference fills with a G-PCC
``tmc3`` subprocess + temporary PLY files (shared/utils.py:169-240, codec_pipeline.py:441-462,
codec_parallel.py:266-289).  G-PCC itself is an external binary outside the
path; the container slot (one byte string per frame, ``points_streams[i]``) is kept and filled with
"OCT2" breadth-first octree occupancy streams that are built and expanded on the device by
``pcc_octree_encode`` / ``pcc_octree_decode`` (csrc/octree.cu): the coordinates never visit the host.
Only the stream headers are touched here (the decoder needs the per-frame point counts to size its
buffers before anything runs).
"""
import struct

import numpy as np
import torch

from . import ops

_MAGIC = b"OCT2"
LOG2_STRIDE = 3          # the latent coordinates live at tensor stride 8 (shared/utils.py:173)


def encode_device(cmap):
    """CoordMap of the latents (with frame_off_dev / frame_counts from build_pyramid) ->
    [(device buffer, device size)] per frame; see model.fetch_strings()."""
    if cmap.frame_off_dev is None:
        raise ops._lib.PccError("coordinate map carries no per-frame offsets (build_pyramid with num_frames)")
    return ops.octree_encode(cmap.keys, cmap.frame_off_dev, cmap.frame_counts, cmap.log2_stride)


def stream_points(stream: bytes) -> int:
    """Point count stated by a stream's header."""
    if len(stream) < 20 or bytes(stream[:4]) != _MAGIC:
        raise ValueError("bad geometry stream")
    return struct.unpack_from("<I", stream, 4)[0]


def pack_streams(points_streams):
    """Host-side layout of the decoder's geometry input: (blob uint8 with every stream at a 4-byte
    aligned offset, int64 [nf] offsets, int32 [nf + nf+1] = lengths | row offsets, counts)."""
    nf = len(points_streams)
    counts = [stream_points(s) for s in points_streams]
    offs = np.zeros(nf, dtype=np.int64)
    pos = 0
    for b, s in enumerate(points_streams):
        offs[b] = pos
        pos += (len(s) + 3) // 4 * 4
    blob = np.zeros(max(pos, 4), dtype=np.uint8)
    for b, s in enumerate(points_streams):
        blob[offs[b]:offs[b] + len(s)] = np.frombuffer(s, dtype=np.uint8)
    meta32 = np.zeros(2 * nf + 1, dtype=np.int32)
    meta32[:nf] = [len(s) for s in points_streams]
    meta32[nf + 1:] = np.cumsum(counts)
    return blob, offs, meta32, counts


def decode_device(blob_d: torch.Tensor, offs_d: torch.Tensor, meta32_d: torch.Tensor, counts):
    """-> (keys int64 ascending, status) on the device, no host sync."""
    nf = len(counts)
    return ops.octree_decode(blob_d, offs_d, meta32_d[:nf], meta32_d[nf:], nf, int(sum(counts)), LOG2_STRIDE)
