"""Bitstream containers of the reference, written with ``struct`` instead of the
``bitstream`` package's bit strings (sender/encoder/codec_pipeline.py:464-517 <->
receiver/decoder/codec_parallel.py:173-216 for the batched layout; codec_single.py:269-337 <->
receiver/decoder/codec_single.py:90-136 for the per-frame layout).  Field order, widths and
endianness are the reference's (int32 / float64, big endian); the y/z strings are this repository's
PCL1 lane containers and ``points_i`` is the octree stream of geometry.py instead of a tmc3 stream.
"""
import struct
from typing import List, Sequence, Tuple


def make_bitstream_batched(y_string: bytes, z_string: bytes, y_shape: int, z_shape: int,
                           points_streams: Sequence[bytes], ks, q) -> bytes:
    """One blob per quality per GOP (codec_pipeline.py:464-517)."""
    num_frames = len(points_streams)
    parts = [struct.pack(">idd", num_frames, float(q[0]), float(q[1])),
             struct.pack(">iiii", int(y_shape), int(z_shape), len(y_string), len(z_string)),
             y_string, z_string]
    for i in range(num_frames):
        p = points_streams[i]
        parts.append(struct.pack(">iiii", len(p), int(ks[0][i]), int(ks[1][i]), int(ks[2][i])))
        parts.append(p)
    return b"".join(parts)


def read_bitstream_batched(data: bytes, copy: bool = True):
    """-> y_strings, z_strings, y_shape, z_shape, points_streams, ks, q (codec_parallel.py:173-216).
    copy=False returns memoryviews into `data` instead of bytes objects."""
    mv = memoryview(data)
    pos = 0

    def take(fmt):
        nonlocal pos
        sz = struct.calcsize(fmt)
        if pos + sz > len(mv):
            raise ValueError("truncated bitstream")
        v = struct.unpack_from(fmt, mv, pos)
        pos += sz
        return v

    def take_bytes(n):
        nonlocal pos
        if n < 0 or pos + n > len(mv):
            raise ValueError("truncated bitstream")
        b = mv[pos:pos + n]
        pos += n
        return bytes(b) if copy else b

    num_frames, q_g, q_a = take(">idd")
    y_shape, z_shape, y_len, z_len = take(">iiii")
    y_string = take_bytes(y_len)
    z_string = take_bytes(z_len)
    points_streams, ks = [], [[], [], []]
    if num_frames < 0 or num_frames > 65534:
        raise ValueError("bad frame count")
    for _ in range(num_frames):
        plen, k1, k2, k3 = take(">iiii")
        ks[0].append(k1)
        ks[1].append(k2)
        ks[2].append(k3)
        points_streams.append(take_bytes(plen))
    return [y_string], [z_string], y_shape, z_shape, points_streams, ks, (q_g, q_a)


def make_bitstream(y_strings: Sequence[bytes], z_strings: Sequence[bytes], y_shapes: Sequence[int],
                   z_shapes: Sequence[int], points_streams: Sequence[bytes], ks, q) -> bytes:
    """Per-frame layout (codec_single.py:269-337)."""
    num_frames = len(y_strings)
    parts = [struct.pack(">idd", num_frames, float(q[0]), float(q[1]))]
    for i in range(num_frames):
        p, ys, zs = points_streams[i], y_strings[i], z_strings[i]
        parts.append(struct.pack(">iiiiiiii", int(y_shapes[i]), int(z_shapes[i]), len(p), len(ys), len(zs),
                                 int(ks[0][i]), int(ks[1][i]), int(ks[2][i])))
        parts += [p, ys, zs]
    return b"".join(parts)


def read_bitstream(data: bytes):
    """Per-frame layout reader (receiver/decoder/codec_single.py:90-136)."""
    mv = memoryview(data)
    pos = 0
    num_frames, q_g, q_a = struct.unpack_from(">idd", mv, pos)
    pos += struct.calcsize(">idd")
    y_strings, z_strings, y_shapes, z_shapes, points_streams, ks = [], [], [], [], [], [[], [], []]
    for _ in range(num_frames):
        ysh, zsh, plen, ylen, zlen, k1, k2, k3 = struct.unpack_from(">iiiiiiii", mv, pos)
        pos += 32
        if min(plen, ylen, zlen) < 0 or pos + plen + ylen + zlen > len(mv):
            raise ValueError("truncated bitstream")
        points_streams.append(bytes(mv[pos:pos + plen])); pos += plen
        y_strings.append(bytes(mv[pos:pos + ylen])); pos += ylen
        z_strings.append(bytes(mv[pos:pos + zlen])); pos += zlen
        y_shapes.append(ysh); z_shapes.append(zsh)
        ks[0].append(k1); ks[1].append(k2); ks[2].append(k3)
    return y_strings, z_strings, y_shapes, z_shapes, points_streams, ks, (q_g, q_a)


# ---- the same containers through the C ABI (include/pcc_b200.h section 6) ------------------------------
# What a C/C++ host links against; the struct versions above stay the readable statement of the layout and
# the tests require both to produce identical bytes.

def _spans(items):
    import ctypes as C
    from ._lib import Span
    arr = (Span * max(len(items), 1))()
    keep = []
    for i, b in enumerate(items):
        buf = (C.c_uint8 * max(len(b), 1)).from_buffer_copy(bytes(b) or b"\0")
        keep.append(buf)
        arr[i].data = C.addressof(buf)
        arr[i].len = len(b)
    return arr, keep


def _ks_table(ks, n):
    import ctypes as C
    tab = (C.c_int32 * (3 * max(n, 1)))()
    for l in range(3):
        for i in range(n):
            tab[l * max(n, 1) + i] = int(ks[l][i])
    return tab


def make_bitstream_batched_native(y_string, z_string, y_shape, z_shape, points_streams, ks, q) -> bytes:
    import ctypes as C
    from ._lib import lib, check
    n = len(points_streams)
    pts, keep = _spans(points_streams)
    yz, keep2 = _spans([y_string, z_string])
    size = lib.pcc_container_batched_bytes(n, len(y_string), len(z_string), pts)
    out = (C.c_uint8 * max(size, 1))()
    written = C.c_int64(0)
    check(lib.pcc_container_write_batched(n, float(q[0]), float(q[1]), int(y_shape), int(z_shape), yz[0], yz[1], pts,
                                          _ks_table(ks, n), max(n, 1), out, size, C.byref(written)))
    return bytes(memoryview(out)[:written.value])


def read_bitstream_batched_native(data: bytes, max_frames: int = 65534):
    import ctypes as C
    from ._lib import lib, Span, PccError
    buf = (C.c_uint8 * max(len(data), 1)).from_buffer_copy(bytes(data) or b"\0")
    nf, qg, qa, ysh, zsh = C.c_int32(0), C.c_double(0), C.c_double(0), C.c_int32(0), C.c_int32(0)
    y, z = Span(), Span()
    pts = (Span * max_frames)()
    ks = (C.c_int32 * (3 * max_frames))()
    rc = lib.pcc_container_parse_batched(buf, len(data), max_frames, C.byref(nf), C.byref(qg), C.byref(qa), C.byref(ysh),
                                         C.byref(zsh), C.byref(y), C.byref(z), pts, ks, max_frames)
    if rc:
        raise ValueError(f"truncated bitstream ({lib.pcc_last_error().decode()})")
    base = C.addressof(buf)
    cut = lambda s: bytes(memoryview(buf)[s.data - base:s.data - base + s.len]) if s.len else b""
    n = nf.value
    return ([cut(y)], [cut(z)], ysh.value, zsh.value, [cut(pts[i]) for i in range(n)],
            [[ks[l * max_frames + i] for i in range(n)] for l in range(3)], (qg.value, qa.value))


def make_bitstream_native(y_strings, z_strings, y_shapes, z_shapes, points_streams, ks, q) -> bytes:
    import ctypes as C
    from ._lib import lib, check
    n = len(y_strings)
    ys, k1 = _spans(y_strings)
    zs, k2 = _spans(z_strings)
    ps, k3 = _spans(points_streams)
    size = lib.pcc_container_frames_bytes(n, ys, zs, ps)
    out = (C.c_uint8 * max(size, 1))()
    written = C.c_int64(0)
    ysh = (C.c_int32 * max(n, 1))(*[int(v) for v in y_shapes])
    zsh = (C.c_int32 * max(n, 1))(*[int(v) for v in z_shapes])
    check(lib.pcc_container_write_frames(n, float(q[0]), float(q[1]), ysh, zsh, ys, zs, ps, _ks_table(ks, n), max(n, 1), out,
                                         size, C.byref(written)))
    return bytes(memoryview(out)[:written.value])


def read_bitstream_native(data: bytes, max_frames: int = 65534):
    import ctypes as C
    from ._lib import lib, Span
    buf = (C.c_uint8 * max(len(data), 1)).from_buffer_copy(bytes(data) or b"\0")
    nf, qg, qa = C.c_int32(0), C.c_double(0), C.c_double(0)
    ysh, zsh = (C.c_int32 * max_frames)(), (C.c_int32 * max_frames)()
    y, z, pts = (Span * max_frames)(), (Span * max_frames)(), (Span * max_frames)()
    ks = (C.c_int32 * (3 * max_frames))()
    rc = lib.pcc_container_parse_frames(buf, len(data), max_frames, C.byref(nf), C.byref(qg), C.byref(qa), ysh, zsh, y, z,
                                        pts, ks, max_frames)
    if rc:
        raise ValueError(f"truncated bitstream ({lib.pcc_last_error().decode()})")
    base = C.addressof(buf)
    cut = lambda s: bytes(memoryview(buf)[s.data - base:s.data - base + s.len]) if s.len else b""
    n = nf.value
    return ([cut(y[i]) for i in range(n)], [cut(z[i]) for i in range(n)], [ysh[i] for i in range(n)],
            [zsh[i] for i in range(n)], [cut(pts[i]) for i in range(n)],
            [[ks[l * max_frames + i] for i in range(n)] for l in range(3)], (qg.value, qa.value))
