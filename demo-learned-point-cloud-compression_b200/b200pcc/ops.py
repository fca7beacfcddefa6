"""Tensor-level wrappers over the C ABI (include/pcc_b200.h).

torch is used for device memory and streams only: every function hands raw ``data_ptr()``s and the
current CUDA stream to libpcc_b200.so.  Nothing here computes on the host.
"""
import ctypes as C
import threading
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, CdfTables

# Per-thread state.  The C calls release the GIL and most of them launch several kernels that share a
# scratch buffer, so two callers (the containers use 3-worker pools: encoder.py:50,75, decoder.py:47)
# must never share a workspace: another thread's kernels could be enqueued between two kernels of one
# call.  Workspaces, the cached stream handle and the deferred-check list are therefore thread local.
_TLS = threading.local()


class KernelProfiler:
    """Optional per-family CUDA-event timing on the launching stream (bench.py's roofline leg).
    A record is (family, algorithmic bytes, flops, start event, stop event, dense-equivalent flops): `flops` is
    2 P Cin Cout with P the pairs counted from the kernel map, the dense equivalent counts every
    (offset, output row) whether a neighbour exists or not -- what the tensor pipe actually executes."""

    def __init__(self):
        self.records = []

    def summary(self):
        out = {}
        for fam, nbytes, flops, e0, e1, dense in self.records:
            d = out.setdefault(fam, {"launches": 0, "ms": 0.0, "bytes": 0, "flops": 0, "dense_flops": 0})
            d["launches"] += 1
            d["ms"] += e0.elapsed_time(e1)
            d["bytes"] += nbytes
            d["flops"] += flops
            d["dense_flops"] += dense
        return out


PROFILER = None


class _prof:
    def __init__(self, family, nbytes=0, flops=0, dense_flops=None):
        self.family, self.nbytes, self.flops = family, int(nbytes), int(flops)
        self.dense = int(flops if dense_flops is None else dense_flops)

    def __enter__(self):
        if PROFILER is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if PROFILER is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILER.records.append((self.family, self.nbytes, self.flops, self.e0, e1, self.dense))
        return False


def kmap_pairs(nbr_ptr: int, nbr_ld: int, n_out: int, kvol: int, device) -> int:
    """Pairs of a kernel map given by raw address (measurement only: synchronises)."""
    cnt = torch.empty(1, dtype=torch.int64, device=device)
    check(lib.pcc_kmap_count_pairs(nbr_ptr, nbr_ld, n_out, kvol, cnt.data_ptr(), _stream()))
    return int(cnt.item())


def _stream() -> int:
    st = getattr(_TLS, "stream", None)
    if st is not None:
        return st
    return torch.cuda.current_stream().cuda_stream


class stream_scope:
    """Resolve the current CUDA stream once for a whole compress()/decompress() call (the per-launch
    torch.cuda.current_stream() lookup is a measurable share of the host time at ~170 launches/frame).
    The cached handle is thread local."""

    def __enter__(self):
        self.prev = (getattr(_TLS, "stream", None), getattr(_TLS, "stream_obj", None))
        so = torch.cuda.current_stream()
        _TLS.stream_obj = so
        _TLS.stream = so.cuda_stream
        return self

    def __exit__(self, *exc):
        _TLS.stream, _TLS.stream_obj = self.prev
        return False


def stream_obj():
    """The torch Stream the current call runs on (cached by stream_scope; event.record(stream_obj()) skips torch's
    current-stream lookup, which costs more than the record itself)."""
    so = getattr(_TLS, "stream_obj", None)
    return so if so is not None else torch.cuda.current_stream()


def _ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    return t.data_ptr()


def workspace(nbytes: int, device, slot: str = "ws") -> torch.Tensor:
    """Growing per-thread, per-device scratch buffer (stream ordered reuse on the current stream)."""
    ws = getattr(_TLS, "ws", None)
    if ws is None:
        ws = _TLS.ws = {}
    key = (str(device), slot)
    buf = ws.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes * 1.25), 1 << 20), dtype=torch.uint8, device=device)
        ws[key] = buf
    return buf


def _need_cuda(t: torch.Tensor, dtype=None):
    if not t.is_cuda:
        raise _lib.PccError("b200pcc kernels need CUDA tensors; there is no CPU fallback")
    if dtype is not None and t.dtype != dtype:
        raise _lib.PccError(f"expected dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise _lib.PccError("expected a contiguous tensor")


# ------------------------------------------------------------------------------------------------
# coordinates
# ------------------------------------------------------------------------------------------------
def keys_from_coords(coords: torch.Tensor) -> torch.Tensor:
    """[N,4] int32 or float32 (b,x,y,z) -> uint64 keys stored as int64 tensor."""
    _need_cuda(coords)
    n = coords.shape[0]
    keys = torch.empty(n, dtype=torch.int64, device=coords.device)
    if coords.dtype == torch.int32:
        check(lib.pcc_keys_from_coords_i32(_ptr(coords), n, _ptr(keys), _stream()))
    elif coords.dtype == torch.float32:
        check(lib.pcc_keys_from_coords_f32(_ptr(coords), n, _ptr(keys), _stream()))
    else:
        raise _lib.PccError(f"coordinates must be int32 or float32, got {coords.dtype}")
    return keys


def ingest_frame(points_i16: torch.Tensor, colors_f64: torch.Tensor, batch: int, keys_out: torch.Tensor,
                 feats_out: torch.Tensor) -> None:
    _need_cuda(points_i16, torch.int16)
    _need_cuda(colors_f64, torch.float64)
    n = points_i16.shape[0]
    check(lib.pcc_ingest_frame(_ptr(points_i16), _ptr(colors_f64), n, batch, _ptr(keys_out), _ptr(feats_out),
                               _stream()))


def ingest_points(points_i16: torch.Tensor, batch: int, keys_out: torch.Tensor) -> None:
    _need_cuda(points_i16, torch.int16)
    check(lib.pcc_ingest_points(_ptr(points_i16), points_i16.shape[0], batch, _ptr(keys_out), _stream()))


def ingest_features(colors_f64: torch.Tensor, src_row: Optional[torch.Tensor], m: int, c_dst: int = 16) -> torch.Tensor:
    """bf16 feature rows (1,r,g,b,0...) [m, c_dst] of the colour rows src_row[i]."""
    _need_cuda(colors_f64, torch.float64)
    out = torch.empty((m, c_dst), dtype=torch.bfloat16, device=colors_f64.device)
    check(lib.pcc_ingest_features(_ptr(colors_f64), _ptr(src_row), m, _ptr(out), c_dst, _stream()))
    return out


def pack_visualizer(keys: torch.Tensor, feats3: torch.Tensor, offset: float = 100.0):
    """-> (points float32 [n,3] = xyz + offset, colours uint8 [n,3]): the client's playout buffer
    (receiver/client/client.py:139-146) built on the device."""
    _need_cuda(keys, torch.int64)
    _need_cuda(feats3, torch.float32)
    n = keys.shape[0]
    pts = torch.empty((n, 3), dtype=torch.float32, device=keys.device)
    col = torch.empty((n, 3), dtype=torch.uint8, device=keys.device)
    check(lib.pcc_pack_visualizer(_ptr(keys), _ptr(feats3), n, float(offset), _ptr(pts), _ptr(col), _stream()))
    return pts, col


def ingest_features_bf16(colors_bf16: torch.Tensor, src_row: Optional[torch.Tensor], m: int, c_dst: int = 16) -> torch.Tensor:
    """The same from colours already rounded to bf16 [n,3] (staged with PCC_STAGE_F64_TO_BF16)."""
    _need_cuda(colors_bf16, torch.bfloat16)
    out = torch.empty((m, c_dst), dtype=torch.bfloat16, device=colors_bf16.device)
    check(lib.pcc_ingest_features_bf16(_ptr(colors_bf16), _ptr(src_row), m, _ptr(out), c_dst, _stream()))
    return out


def unpack_output(keys: torch.Tensor, feats3: Optional[torch.Tensor], want_points: bool = True):
    """-> (points int32 [n,3] or None, colours float32 [n,3] clipped like the reference's pack_batches, or
    None when feats3 is None)."""
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    if feats3 is not None:
        _need_cuda(feats3, torch.float32)
        if feats3.shape != (n, 3):
            raise _lib.PccError(f"unpack_output: expected [{n}, 3] colours, got {tuple(feats3.shape)}")
    pts = torch.empty((n, 3), dtype=torch.int32, device=keys.device) if want_points else None
    col = torch.empty((n, 3), dtype=torch.float32, device=keys.device) if feats3 is not None else None
    check(lib.pcc_unpack_output(_ptr(keys), _ptr(feats3), n, _ptr(pts), _ptr(col), _stream()))
    return pts, col


def coords_from_keys(keys: torch.Tensor) -> torch.Tensor:
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    out = torch.empty((n, 4), dtype=torch.int32, device=keys.device)
    check(lib.pcc_coords_from_keys(_ptr(keys), n, _ptr(out), _stream()))
    return out


# Deferred size checks.  A caller that already knows a data-dependent count (the decoder reads them
# from the bitstream) passes it as `expect`: the op then returns without a host sync and files
# (device count, expected value, message) here; the pipeline verifies all of them with one copy at the
# end of the call (verify_deferred) and raises if any differs.
def _deferred_list():
    d = getattr(_TLS, "deferred", None)
    if d is None:
        d = _TLS.deferred = []
    return d


def _defer(cnt: torch.Tensor, expect: int, what: str):
    _deferred_list().append((cnt, int(expect), what))


def fault_fetch(device) -> torch.Tensor:
    """Device int32 [1] holding the device fault word (cleared by the fetch, stream ordered): non-zero when a
    kernel since the last fetch gave up instead of trapping (pcc_b200.h: PCC_FAULT_*)."""
    dst = torch.empty(1, dtype=torch.int32, device=device)
    check(lib.pcc_fault_fetch(_ptr(dst), _stream()))
    return dst


def fault_message(code: int) -> str:
    return f"device fault {int(code)}: {lib.pcc_fault_string(int(code)).decode()}"


def verify_deferred(extra_status=(), device=None):
    """One launch and one device->host copy for every deferred count and status word -- and, when `device` is given,
    the device fault word; raises PccError on mismatch.  (pcc_fetch_words gathers the words; the copy lands in a pinned
    buffer of this thread.  Counters must be int32 device words.)"""
    items = list(_deferred_list())
    _TLS.deferred = []
    words = [c for c, _, _ in items] + [t for t, _ in extra_status]
    if not words and device is None:
        return
    if len(words) >= 47 or any(w.dtype != torch.int32 for w in words):      # (not on the codec's path) the general way
        tensors = [w.reshape(1).to(torch.int32) for w in words]
        if device is not None:
            tensors.append(fault_fetch(device))
        host = torch.cat(tensors).cpu().numpy()
    else:
        dev = device if device is not None else words[0].device
        n = len(words)
        srcs = (C.c_void_p * max(n, 1))(*[w.data_ptr() for w in words])
        dst = torch.empty(n + 1, dtype=torch.int32, device=dev)
        check(lib.pcc_fetch_words(srcs, n, 1 if device is not None else 0, _ptr(dst), _stream()))
        pin = getattr(_TLS, "verify_pin", None)
        if pin is None or pin.numel() < n + 1:
            pin = _TLS.verify_pin = torch.empty(64, dtype=torch.int32, pin_memory=True)
        pin[:n + 1].copy_(dst, non_blocking=True)
        stream_obj().synchronize()
        host = pin[:n + 1].numpy().copy() if device is not None else pin[:n].numpy().copy()
    if device is not None and int(host[-1]) != 0:
        raise _lib.PccError(fault_message(int(host[-1])))
    for (c, e, what), v in zip(items, host[:len(items)]):
        if int(v) != e:
            raise _lib.PccError(f"{what}: device count {int(v)} != expected {e}")
    for (t, what), v in zip(extra_status, host[len(items):]):
        if int(v) != 0:
            raise _lib.PccError(f"{what} (status {int(v)})")


def clear_deferred():
    _TLS.deferred = []


def capture_voxelize(xyzrgba: torch.Tensor, voxel_size: float, depth_clip: float, max_points: Optional[int] = None):
    """XYZRGBA measure float32 [n_px, 4] -> (points int16 [cap,3], colours float64 [cap,3], count int32 [1]), all on
    the device, nothing waited for (sender/capturer/capturer.py:84-120; see pcc_capture_voxelize)."""
    _need_cuda(xyzrgba, torch.float32)
    if xyzrgba.dim() != 2 or xyzrgba.shape[1] != 4:
        raise _lib.PccError(f"capture_voxelize: expected [n, 4] float32, got {tuple(xyzrgba.shape)}")
    n = xyzrgba.shape[0]
    dev = xyzrgba.device
    mp = int(max_points) if max_points is not None else 0
    if mp < 0:
        raise _lib.PccError("capture_voxelize: max_points must be positive or None")
    cap = min(n, mp) if mp > 0 else n
    pts = torch.empty((cap, 3), dtype=torch.int16, device=dev)
    col = torch.empty((cap, 3), dtype=torch.float64, device=dev)
    cnt = torch.empty(1, dtype=torch.int32, device=dev)
    wsb = lib.pcc_capture_voxelize_workspace_bytes(n)
    ws = workspace(wsb, dev)
    with _prof("capture_voxelize", n * 16 + cap * 30):
        check(lib.pcc_capture_voxelize(_ptr(xyzrgba), n, float(voxel_size), float(depth_clip), mp, _ptr(pts), _ptr(col),
                                       cap, _ptr(cnt), _ptr(ws), ws.numel(), _stream()))
    return pts, col, cnt


def sort_unique(keys: torch.Tensor, want_src: bool = True, expect: Optional[int] = None, sync: bool = True,
                key_bits: int = 64):
    """-> (unique sorted keys, src_row or None).  One host sync for the count, none with `expect`.
    key_bits: the caller guarantees that bits [key_bits, 64) of every key are zero (frame index below
    2^(key_bits - 48)); the radix sort then skips those digit passes."""
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    dev = keys.device
    out = torch.empty(n, dtype=torch.int64, device=dev)
    # sync=False callers may read all n rows before they know the count: rows past it must be benign (-1 = no row)
    src = (torch.empty(n, dtype=torch.int32, device=dev) if sync or expect is not None else
           torch.full((n,), -1, dtype=torch.int32, device=dev)) if want_src else None
    cnt = torch.empty(1, dtype=torch.int32, device=dev)
    wsb = lib.pcc_sort_unique_workspace_bytes(n)
    ws = workspace(wsb, dev)
    with _prof("sort_unique", n * (16 + 12 + (8 if want_src else 0))):
        check(lib.pcc_sort_unique_bits(_ptr(keys), n, int(key_bits), _ptr(out), _ptr(src), _ptr(cnt), _ptr(ws), ws.numel(),
                                       _stream()))
    if expect is not None:
        m = int(expect)
        _defer(cnt, m, "sort_unique")
    elif not sync:
        return out, src, cnt             # the caller reads cnt when it needs the size
    else:
        m = int(cnt.item())
    return out[:m], (src[:m] if want_src else None)


def coords_downsample(keys: torch.Tensor, log2_stride: int):
    """-> (parent keys, parent_row[int32, n])."""
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    dev = keys.device
    out = torch.empty(n, dtype=torch.int64, device=dev)
    parent = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.empty(1, dtype=torch.int32, device=dev)
    wsb = lib.pcc_coords_downsample_workspace_bytes(n)
    ws = workspace(wsb, dev)
    check(lib.pcc_coords_downsample(_ptr(keys), n, log2_stride, _ptr(out), _ptr(parent), _ptr(cnt), _ptr(ws),
                                    ws.numel(), _stream()))
    m = int(cnt.item())
    return out[:m], parent


def coords_pyramid_expect(keys: torch.Tensor, log2_stride: int, levels: int, expect_last: int):
    """Decoder variant: only the last level is used and its size is known from the bitstream, so the
    pyramid runs without any host sync; the device count is filed for verify_deferred()."""
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    dev = keys.device
    kout = torch.empty((levels, max(n, 1)), dtype=torch.int64, device=dev)
    par = torch.empty((levels, max(n, 1)), dtype=torch.int32, device=dev)
    meta = torch.empty(levels, dtype=torch.int32, device=dev)
    ws = workspace(lib.pcc_coords_pyramid_workspace_bytes(n), dev)
    check(lib.pcc_coords_pyramid(_ptr(keys), n, log2_stride, levels, _ptr(kout), _ptr(par), _ptr(meta), 0, None,
                                 _ptr(ws), ws.numel(), _stream()))
    _defer(meta[levels - 1], expect_last, "coords_pyramid (last level size vs bitstream)")
    return kout[levels - 1, :int(expect_last)]


_SIDE = {}


def _side_stream(device, name: str = "readback") -> torch.cuda.Stream:
    """Side streams per device: "readback" (small device->host copies the host waits for) and "h2d"
    (bulk host->device copies) are separate so a 20 MB copy never sits in front of a 100-byte one."""
    st = _SIDE.get((str(device), name))
    if st is None:
        st = _SIDE[(str(device), name)] = torch.cuda.Stream(device=device)
    return st


def coords_pyramid(keys: torch.Tensor, log2_stride: int, levels: int, num_frames: int, overlap=None, extra=None):
    """All `levels` stride-2 coarsenings of a sorted key set with ONE host sync.
    -> (list of level keys, list of parent_row [rows of the finer level], host counts [levels],
        host frame offsets [(levels+1), num_frames+1] (row 0 = the input set), the same on the device).
    overlap: optional callable that enqueues independent work on the current stream; it runs after the
    pyramid kernels are launched, and the counts are read back on a side stream, so the host wait (and
    the host-side bookkeeping that follows) is covered by that work instead of leaving the GPU idle.
    extra: optional device int32 counter fetched with the same wait (appended to the returned tuple)."""
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    dev = keys.device
    kout = torch.empty((levels, max(n, 1)), dtype=torch.int64, device=dev)
    par = torch.empty((levels, max(n, 1)), dtype=torch.int32, device=dev)
    nf1 = num_frames + 1
    meta = torch.empty(levels + (levels + 1) * nf1, dtype=torch.int32, device=dev)
    ws = workspace(lib.pcc_coords_pyramid_workspace_bytes(n), dev, "pyramid")
    with _prof("coords_pyramid", levels * n * (8 + 4 + 8)):
        check(lib.pcc_coords_pyramid(_ptr(keys), n, log2_stride, levels, _ptr(kout), _ptr(par), _ptr(meta),
                                     num_frames, meta.data_ptr() + 4 * levels if num_frames else None, _ptr(ws),
                                     ws.numel(), _stream()))
    extra_host = None
    if overlap is None:
        if extra is not None:
            both = torch.cat([meta, extra.view(-1)[:1]]).cpu().numpy()
            host, extra_host = both[:-1], int(both[-1])
        else:
            host = meta.cpu().numpy()              # the one sync
    else:
        ready = torch.cuda.Event()
        ready.record()
        overlap()
        side = _side_stream(dev)
        host_t = torch.empty(meta.numel() + 1, dtype=meta.dtype, pin_memory=True)
        with torch.cuda.stream(side):
            side.wait_event(ready)
            host_t[:-1].copy_(meta, non_blocking=True)
            if extra is not None:
                host_t[-1:].copy_(extra.view(-1)[:1], non_blocking=True)
            done = torch.cuda.Event()
            done.record()
        done.synchronize()                     # the one sync (pyramid kernels + 100-byte copy only)
        host = host_t.numpy()[:-1]
        if extra is not None:
            extra_host = int(host_t[-1])
    counts = [int(c) for c in host[:levels]]
    offs = host[levels:].reshape(levels + 1, nf1) if num_frames else None
    keys_l = [kout[l, :counts[l]] for l in range(levels)]
    n_fine = [n] + counts[:-1]
    par_l = [par[l, :n_fine[l]] for l in range(levels)]
    offs_dev = meta[levels:].view(levels + 1, nf1) if num_frames else None
    if extra is not None:
        return keys_l, par_l, counts, offs, offs_dev, extra_host
    return keys_l, par_l, counts, offs, offs_dev


def coords_children(keys: torch.Tensor, log2_stride_out: int) -> torch.Tensor:
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    out = torch.empty(8 * n, dtype=torch.int64, device=keys.device)
    check(lib.pcc_coords_children(_ptr(keys), n, log2_stride_out, _ptr(out), _stream()))
    return out


def hash_capacity(n: int) -> int:
    cap = 8
    while cap < 2 * n:
        cap *= 2
    return cap


def hash_build(keys: torch.Tensor):
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    cap = hash_capacity(n)
    tk = torch.empty(cap, dtype=torch.int64, device=keys.device)
    tv = torch.empty(cap, dtype=torch.int32, device=keys.device)
    with _prof("hash_build", n * 28):      # 16 B/coord read-equivalent + 12 B/slot written
        check(lib.pcc_hash_build(_ptr(keys), n, _ptr(tk), _ptr(tv), cap, _stream()))
    return tk, tv


def hash_query(queries: torch.Tensor, table) -> torch.Tensor:
    _need_cuda(queries, torch.int64)
    tk, tv = table
    nq = queries.shape[0]
    rows = torch.empty(nq, dtype=torch.int32, device=queries.device)
    check(lib.pcc_hash_query(_ptr(queries), nq, _ptr(tk), _ptr(tv), tk.shape[0], _ptr(rows), _stream()))
    return rows


def kmap_k3(keys: torch.Tensor, log2_stride: int, table) -> torch.Tensor:
    tk, tv = table
    n = keys.shape[0]
    nbr = torch.empty((27, n), dtype=torch.int32, device=keys.device)
    with _prof("kmap_k3_voxelhash", n * (8 + 27 * 4)):
        check(lib.pcc_kmap_k3(_ptr(keys), n, log2_stride, _ptr(tk), _ptr(tv), tk.shape[0], _ptr(nbr), _stream()))
    return nbr


def blockmap_build(keys: torch.Tensor, log2_stride: int, block_bound: Optional[int] = None):
    """4x4x4-block table of a sorted key set -> (table keys, table values [cap, 2] int64).
    block_bound: host-known upper bound on the number of occupied blocks (= coordinates two levels
    coarser); it sizes the table, which otherwise has to assume one block per voxel.  A compact table
    (a few MB instead of ~100 MB at 1.8 M rows) keeps the probes of the map kernel in L2."""
    _need_cuda(keys, torch.int64)
    n = keys.shape[0]
    cap = hash_capacity(n if block_bound is None else max(min(int(block_bound), n), 1))
    tk = torch.empty(cap, dtype=torch.int64, device=keys.device)
    tv = torch.empty((cap, 2), dtype=torch.int64, device=keys.device)
    with _prof("blockmap_build", n * 8 + cap * 8 + (n // 8) * 24):   # keys read, table cleared, ~n/8 entries
        check(lib.pcc_blockmap_build(_ptr(keys), n, log2_stride, _ptr(tk), _ptr(tv), cap, _stream()))
    return tk, tv


def _padded_map(kvol: int, n: int, device) -> torch.Tensor:
    """[kvol, n] int32 view of a buffer whose row pitch is padded to the 128-row tile of the conv
    kernel (its kernel-map loader then fetches a tile's rows with one bulk copy per offset)."""
    ld = max((n + 127) // 128 * 128, 128)
    return torch.empty((kvol, ld), dtype=torch.int32, device=device)[:, :n]


def kmap_k3_blocks(keys: torch.Tensor, log2_stride: int, btable) -> torch.Tensor:
    tk, tv = btable
    n = keys.shape[0]
    nbr = _padded_map(27, n, keys.device)
    with _prof("kmap_k3", n * (8 + 27 * 4)):   # key read once + 27 map entries written per output row
        check(lib.pcc_kmap_k3_blocks(_ptr(keys), n, log2_stride, _ptr(tk), _ptr(tv), tk.shape[0], _ptr(nbr),
                                     nbr.stride(0), _stream()))
    return nbr


def kmap_down2(child_keys: torch.Tensor, parent_row: torch.Tensor, log2_child_stride: int, n_parent: int):
    n = child_keys.shape[0]
    nbr = _padded_map(8, n_parent, child_keys.device)
    check(lib.pcc_kmap_down2(_ptr(child_keys), _ptr(parent_row), n, log2_child_stride, _ptr(nbr), n_parent,
                             nbr.stride(0), _stream()))
    return nbr


def frame_offsets(keys: torch.Tensor, num_frames: int) -> torch.Tensor:
    off = torch.empty(num_frames + 1, dtype=torch.int32, device=keys.device)
    check(lib.pcc_frame_offsets(_ptr(keys), keys.shape[0], num_frames, _ptr(off), _stream()))
    return off


def lex_order(keys: torch.Tensor, log2_stride: int = 0, frame_bits: int = 16) -> torch.Tensor:
    """perm[r] = row holding canonical (b,x,y,z) rank r.  log2_stride / frame_bits: what the caller knows about the
    keys (coordinates are multiples of 2^log2_stride, frame index < 2^frame_bits) -- fewer radix passes; the defaults
    assume nothing."""
    n = keys.shape[0]
    perm = torch.empty(n, dtype=torch.int32, device=keys.device)
    wsb = lib.pcc_lex_order_workspace_bytes(n)
    ws = workspace(wsb, keys.device)
    check(lib.pcc_lex_order_bits(_ptr(keys), n, int(log2_stride), int(frame_bits), _ptr(perm), _ptr(ws), ws.numel(),
                                 _stream()))
    return perm


# ------------------------------------------------------------------------------------------------
# coordinate coder (OCT2 streams)
# ------------------------------------------------------------------------------------------------
def octree_encode(keys: torch.Tensor, frame_off_dev: torch.Tensor, frame_counts, log2_stride: int):
    """Per-frame OCT2 streams of a sorted key set, left on the device.
    frame_off_dev: int32 [num_frames+1] device row offsets; frame_counts: the same row counts on the
    host (they size the output).  -> [(device uint8 buffer, device int32 size)] per frame."""
    _need_cuda(keys, torch.int64)
    _need_cuda(frame_off_dev, torch.int32)
    nf = len(frame_counts)
    dev = keys.device
    bounds = [int(lib.pcc_octree_max_bytes(int(c))) for c in frame_counts]
    offs = np.zeros(nf + 1, dtype=np.int64)
    offs[1:] = np.cumsum(bounds)
    out = torch.empty(max(int(offs[-1]), 4), dtype=torch.uint8, device=dev)
    out_off = torch.from_numpy(offs[:nf].copy()).pin_memory().to(dev, non_blocking=True)      # (pinned: see verify_deferred)
    sizes = torch.empty(max(nf, 1), dtype=torch.int32, device=dev)
    n = keys.shape[0]
    ws = workspace(lib.pcc_octree_encode_workspace_bytes(n, nf), dev, "octree")
    with _prof("octree_encode", n * 8 * 2 + n * 2):    # keys read by the level and the pack pass; ~1.3 n bytes out
        check(lib.pcc_octree_encode(_ptr(keys), n, _ptr(frame_off_dev), nf, log2_stride, _ptr(out), _ptr(out_off),
                                    _ptr(sizes), _ptr(ws), ws.numel(), _stream()))
    return [(out[int(offs[b]):int(offs[b + 1])], sizes[b:b + 1]) for b in range(nf)]


def octree_decode(streams: torch.Tensor, stream_off: torch.Tensor, stream_len: torch.Tensor,
                  row_off: torch.Tensor, num_frames: int, n_total: int, log2_stride: int):
    """streams: device uint8 blob holding every frame's stream at a 4-byte aligned offset.
    -> (keys int64 [n_total] ascending, status int32 [1]); no host sync."""
    _need_cuda(streams, torch.uint8)
    _need_cuda(stream_off, torch.int64)
    _need_cuda(stream_len, torch.int32)
    _need_cuda(row_off, torch.int32)
    dev = streams.device
    keys = torch.empty(n_total, dtype=torch.int64, device=dev)
    status = torch.empty(1, dtype=torch.int32, device=dev)
    ws = workspace(lib.pcc_octree_decode_workspace_bytes(n_total), dev, "octree")
    with _prof("octree_decode", n_total * 8 * 3):
        check(lib.pcc_octree_decode(_ptr(streams), _ptr(stream_off), _ptr(stream_len), _ptr(row_off), num_frames,
                                    n_total, log2_stride, _ptr(keys), _ptr(status), _ptr(ws), ws.numel(), _stream()))
    return keys, status


# ------------------------------------------------------------------------------------------------
# rows
# ------------------------------------------------------------------------------------------------
def rows_f32_to_bf16(src: torch.Tensor, c_dst: int, src_row: Optional[torch.Tensor] = None,
                     n_out: Optional[int] = None) -> torch.Tensor:
    _need_cuda(src, torch.float32)
    n = n_out if n_out is not None else (src_row.shape[0] if src_row is not None else src.shape[0])
    out = torch.empty((n, c_dst), dtype=torch.bfloat16, device=src.device)
    check(lib.pcc_rows_f32_to_bf16(_ptr(src), src.shape[1], _ptr(src_row), n, _ptr(out), c_dst, _stream()))
    return out


def rows_bf16_to_f32(src: torch.Tensor, c_dst: int) -> torch.Tensor:
    _need_cuda(src, torch.bfloat16)
    n = src.shape[0]
    out = torch.empty((n, c_dst), dtype=torch.float32, device=src.device)
    check(lib.pcc_rows_bf16_to_f32(_ptr(src), src.shape[1], n, _ptr(out), c_dst, _stream()))
    return out


def gather_rows(src: torch.Tensor, src_row: torch.Tensor) -> torch.Tensor:
    n = src_row.shape[0]
    out = torch.empty((n, src.shape[1]), dtype=src.dtype, device=src.device)
    if src.dtype == torch.bfloat16:
        check(lib.pcc_gather_rows_bf16(_ptr(src), src.shape[1], _ptr(src_row), n, _ptr(out), _stream()))
    elif src.dtype == torch.float32:
        check(lib.pcc_gather_rows_f32(_ptr(src), src.shape[1], _ptr(src_row), n, _ptr(out), _stream()))
    else:
        raise _lib.PccError(f"gather_rows: unsupported dtype {src.dtype}")
    return out


def zero_unkept_rows(feats: torch.Tensor, keep: torch.Tensor) -> None:
    """In place: rows of feats [n, c] (bf16) that are not listed in keep (ascending int32 row numbers) become zero."""
    _need_cuda(feats, torch.bfloat16)
    _need_cuda(keep, torch.int32)
    n, c = feats.shape
    with _prof("zero_unkept_rows", (n - keep.shape[0]) * c * 2 + keep.shape[0] * 4):
        check(lib.pcc_zero_unkept_rows(_ptr(keep), keep.shape[0], n, _ptr(feats), c, _stream()))


def gather_rows_i32(src: torch.Tensor, src_row: torch.Tensor) -> torch.Tensor:
    """Rows of an int32 matrix (coordinates) through the float32 row gather: a bit-for-bit copy."""
    _need_cuda(src, torch.int32)
    return gather_rows(src.view(torch.float32), src_row).view(torch.int32)


def gather_keys(src: torch.Tensor, src_row: torch.Tensor) -> torch.Tensor:
    n = src_row.shape[0]
    out = torch.empty(n, dtype=torch.int64, device=src.device)
    check(lib.pcc_gather_keys(_ptr(src), _ptr(src_row), n, _ptr(out), _stream()))
    return out


def topk_select(logits: torch.Tensor, keys: torch.Tensor, frame_off: torch.Tensor, k: torch.Tensor,
                num_frames: int, expect: Optional[int] = None) -> torch.Tensor:
    """-> kept rows (ascending).  One host sync for the count, none with `expect`."""
    _need_cuda(logits, torch.float32)
    n = keys.shape[0]
    dev = keys.device
    keep = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.empty(1, dtype=torch.int32, device=dev)
    wsb = lib.pcc_topk_workspace_bytes(n)
    ws = workspace(wsb, dev)
    with _prof("topk_select", n * (4 + 8 + 4)):
        check(lib.pcc_topk_select(_ptr(logits), _ptr(keys), n, _ptr(frame_off), _ptr(k), num_frames, _ptr(keep),
                                  _ptr(cnt), _ptr(ws), ws.numel(), _stream()))
    if expect is not None:
        m = int(expect)
        _defer(cnt, m, "topk_select")
    else:
        m = int(cnt.item())
    return keep[:m]


# ------------------------------------------------------------------------------------------------
# sparse convolution
# ------------------------------------------------------------------------------------------------
def pack_weights_umma(w: np.ndarray, cin_pad: int) -> torch.Tensor:
    """w float32 [kvol, cin, cout] -> packed bf16 panels (uint8 CPU tensor)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    kvol, cin, cout = w.shape
    nb = lib.pcc_spconv_packed_weight_bytes(kvol, cin_pad, cout)
    out = np.empty(nb, dtype=np.uint8)
    check(lib.pcc_spconv_pack_weights(w.ctypes.data, kvol, cin, cout, cin_pad, out.ctypes.data))
    return torch.from_numpy(out)


def spconv_umma(x: torch.Tensor, nbr: Optional[torch.Tensor], n_out: int, kvol: int, w_packed: torch.Tensor,
                bias: Optional[torch.Tensor], cout: int, residual: Optional[torch.Tensor] = None,
                relu: bool = False, out_bf16: Optional[torch.Tensor] = None,
                out_f32: Optional[torch.Tensor] = None, col0: int = 0, cstore: int = 0) -> None:
    """Writes columns [col0, col0+cout) of the given output tensors (cstore in 1..15: narrow head, only
    the first cstore columns exist in the output).  This wrapper runs ~50 times per frame: it is kept
    free of anything that is not needed to make the call (the profiling leg is a separate branch)."""
    if not x.is_cuda or x.dtype is not torch.bfloat16 or not x.is_contiguous():
        raise _lib.PccError("spconv_umma: expected a contiguous CUDA bfloat16 feature matrix (there is no CPU fallback)")
    n_in, cin = x.shape
    ob = None if out_bf16 is None else out_bf16.data_ptr() + col0 * 2
    of = None if out_f32 is None else out_f32.data_ptr() + col0 * 4
    ldo = (out_bf16 if out_bf16 is not None else out_f32).shape[1]
    rp, ldr = None, 0
    if residual is not None:
        rp = residual.data_ptr() + col0 * 2
        ldr = residual.shape[1]
    nbr_p, nbr_ld = None, 0
    if nbr is not None:
        nbr_ld, cs = nbr.stride()
        if cs != 1 or nbr.dtype is not torch.int32:
            raise _lib.PccError("kernel map must be int32 [kvol, n] with unit column stride")
        nbr_p = nbr.data_ptr()
    bias_p = None if bias is None else bias.data_ptr()
    if PROFILER is None:
        rc = lib.pcc_spconv_umma(x.data_ptr(), n_in, cin, nbr_p, nbr_ld, n_out, kvol, w_packed.data_ptr(), bias_p, cout,
                                 rp, ldr, int(relu), ob, of, ldo, cstore, _stream())
        if rc:
            check(rc)
        return
    cw = cstore if 0 < cstore < cout else cout
    nbytes = 2 * n_in * cin + n_out * (cw * ((2 if ob else 0) + (4 if of else 0)) + cout * (2 if rp else 0)) + \
        (4 * kvol * n_out if nbr is not None else 0) + 2 * kvol * cin * cout
    pairs = kmap_pairs(nbr_p, nbr_ld, n_out, kvol, x.device) if nbr is not None else n_out
    with _prof(f"spconv_umma_k{kvol}_c{cin}x{cout}", nbytes, 2 * pairs * cin * cout, 2 * kvol * n_out * cin * cout):
        check(lib.pcc_spconv_umma(x.data_ptr(), n_in, cin, nbr_p, nbr_ld, n_out, kvol, w_packed.data_ptr(), bias_p, cout,
                                  rp, ldr, int(relu), ob, of, ldo, cstore, _stream()))


_OCTET_BLOCKS = None


def _octet_pairs(nbr_ptr: int, nbr_ld: int, n_par: int, device) -> int:
    """Child-level pairs of an octet convolution from its PARENT map (measurement only): a present parent offset D
    stands for 4^(number of zero axes of D) (source child, output child) blocks."""
    global _OCTET_BLOCKS
    if _OCTET_BLOCKS is None:
        _OCTET_BLOCKS = [4 ** [dx, dy, dz].count(0) for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]
    return sum(_OCTET_BLOCKS[j] * kmap_pairs(nbr_ptr + 4 * j * nbr_ld, nbr_ld, n_par, 1, device) for j in range(27))


def pack_weights_octet(w: np.ndarray, cout_pad: int) -> torch.Tensor:
    """w float32 [27, 32, cout] -> panels for pcc_spconv_octet (uint8 CPU tensor), columns padded to cout_pad."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    kvol, cin, cout = w.shape
    if kvol != 27:
        raise _lib.PccError("octet weights need a k = 3 kernel")
    out = np.empty(lib.pcc_spconv_octet_packed_weight_bytes(cin, cout_pad), dtype=np.uint8)
    check(lib.pcc_spconv_octet_pack_weights(w.ctypes.data, cin, cout, cout_pad, out.ctypes.data))
    return torch.from_numpy(out)


def spconv_octet(x: torch.Tensor, nbr_par: torch.Tensor, n_par: int, w_packed: torch.Tensor, bias: Optional[torch.Tensor],
                 cout: int, residual: Optional[torch.Tensor] = None, relu: bool = False,
                 out_bf16: Optional[torch.Tensor] = None, out_f32: Optional[torch.Tensor] = None, cstore: int = 0) -> None:
    """k3 convolution on complete octets (x has 8 n_par rows; nbr_par is the PARENTS' [27, n_par] kernel map)."""
    if not x.is_cuda or x.dtype is not torch.bfloat16 or not x.is_contiguous() or x.shape[0] != 8 * n_par:
        raise _lib.PccError("spconv_octet: expected a contiguous CUDA bfloat16 matrix with 8 rows per parent")
    nbr_ld, cs = nbr_par.stride()
    if cs != 1 or nbr_par.dtype is not torch.int32:
        raise _lib.PccError("kernel map must be int32 [27, n] with unit column stride")
    ob = None if out_bf16 is None else out_bf16.data_ptr()
    of = None if out_f32 is None else out_f32.data_ptr()
    ldo = (out_bf16 if out_bf16 is not None else out_f32).shape[1]
    rp, ldr = (None, 0) if residual is None else (residual.data_ptr(), residual.shape[1])
    bias_p = None if bias is None else bias.data_ptr()
    args = (x.data_ptr(), n_par, x.shape[1], nbr_par.data_ptr(), nbr_ld, w_packed.data_ptr(), bias_p, cout, rp, ldr,
            int(relu), ob, of, ldo, cstore, _stream())
    if PROFILER is None:
        rc = lib.pcc_spconv_octet(*args)
        if rc:
            check(rc)
        return
    fam, nbytes, flops, dense = octet_family(x.shape[1], cout, cstore, n_par, bool(ob), bool(of), bool(rp),
                                             nbr_par.data_ptr(), nbr_ld, x.device)
    with _prof(fam, nbytes, flops, dense):
        check(lib.pcc_spconv_octet(*args))


def octet_family(cin, cout, cstore, n_par, has_bf16, has_f32, has_res, nbr_ptr, nbr_ld, device):
    """(family, algorithmic bytes, flops from counted pairs, dense-equivalent flops) of one pcc_spconv_octet launch."""
    n = 8 * n_par
    cw = cstore if 0 < cstore < cout else cout
    nbytes = 2 * n * cin + n * (cw * ((2 if has_bf16 else 0) + (4 if has_f32 else 0)) + cout * (2 if has_res else 0)) + \
        4 * 27 * n_par + 2 * 27 * cin * cout
    pairs = _octet_pairs(nbr_ptr, nbr_ld, n_par, device)
    return f"spconv_octet_k27_c{cin}x{cout}", nbytes, 2 * pairs * cin * cout, 2 * 27 * n * cin * cout


def spconv_simt(x: torch.Tensor, nbr: Optional[torch.Tensor], n_out: int, kvol: int, w_bf16: torch.Tensor,
                bias: Optional[torch.Tensor], cout: int, residual: Optional[torch.Tensor] = None,
                relu: bool = False, out_bf16: Optional[torch.Tensor] = None,
                out_f32: Optional[torch.Tensor] = None) -> None:
    _need_cuda(x, torch.bfloat16)
    if nbr is not None and not nbr.is_contiguous():
        nbr = nbr.contiguous()          # this comparator kernel takes an unpadded [kvol, n_out] map
    ldo = (out_bf16 if out_bf16 is not None else out_f32).shape[1]
    ldr = residual.shape[1] if residual is not None else 0
    cin = x.shape[1]
    nbytes = 2 * x.shape[0] * cin + n_out * cout * 4 + (4 * kvol * n_out if nbr is not None else 0)
    with _prof(f"spconv_simt_k{kvol}_c{cin}x{cout}", nbytes, 2 * kvol * n_out * cin * cout):
        check(lib.pcc_spconv_simt(_ptr(x), x.shape[0], x.shape[1], _ptr(nbr), n_out, kvol, _ptr(w_bf16),
                                  _ptr(bias), cout, _ptr(residual), ldr, int(relu), _ptr(out_bf16), _ptr(out_f32),
                                  ldo, _stream()))


# ------------------------------------------------------------------------------------------------
# entropy model
# ------------------------------------------------------------------------------------------------
def factorized_quant(z: torch.Tensor, perm: Optional[torch.Tensor], medians: torch.Tensor, want_f32=False):
    _need_cuda(z, torch.float32)
    n, c = z.shape
    sym = torch.empty((c, n), dtype=torch.int32, device=z.device)
    zb = torch.empty((n, c), dtype=torch.bfloat16, device=z.device)
    zf = torch.empty((n, c), dtype=torch.float32, device=z.device) if want_f32 else None
    check(lib.pcc_factorized_quant(_ptr(z), _ptr(perm), _ptr(medians), n, c, _ptr(sym), _ptr(zb), _ptr(zf),
                                   _stream()))
    return sym, zb, zf


def factorized_dequant(sym: torch.Tensor, perm: Optional[torch.Tensor], medians: torch.Tensor, want_f32=False):
    _need_cuda(sym, torch.int32)
    c, n = sym.shape
    zb = torch.empty((n, c), dtype=torch.bfloat16, device=sym.device)
    zf = torch.empty((n, c), dtype=torch.float32, device=sym.device) if want_f32 else None
    check(lib.pcc_factorized_dequant(_ptr(sym), _ptr(perm), _ptr(medians), n, c, _ptr(zb), _ptr(zf), _stream()))
    return zb, zf


def gauss_encode(y: torch.Tensor, params: torch.Tensor, prow: Optional[torch.Tensor],
                 perm: Optional[torch.Tensor], scale: torch.Tensor, table: torch.Tensor, bound: float):
    _need_cuda(y, torch.float32)
    _need_cuda(params, torch.float32)
    n, c = y.shape
    q = scale.shape[0]
    sym = torch.empty((q, c, n), dtype=torch.int32, device=y.device)
    idx = torch.empty((q, c, n), dtype=torch.int16, device=y.device)
    with _prof("gauss_encode", n * c * (12 + 6 * q)):   # y, sigma, mu read once; int32 sym + u16 idx per quality
        check(lib.pcc_gauss_encode(_ptr(y), _ptr(params), _ptr(prow), _ptr(perm), _ptr(scale), _ptr(table),
                                   table.shape[0], bound, n, c, q, _ptr(sym), _ptr(idx), _stream()))
    return sym, idx


def gauss_indexes(params: torch.Tensor, prow, perm, scale: torch.Tensor, table: torch.Tensor, bound: float,
                  n: int, c: int):
    idx = torch.empty((c, n), dtype=torch.int16, device=params.device)
    check(lib.pcc_gauss_indexes(_ptr(params), _ptr(prow), _ptr(perm), _ptr(scale), _ptr(table), table.shape[0],
                                bound, n, c, _ptr(idx), _stream()))
    return idx


def gauss_dequant(sym: torch.Tensor, params: torch.Tensor, prow, perm, scale: torch.Tensor,
                  gain: Optional[torch.Tensor], bound: float, want_f32=False):
    c, n = sym.shape
    yb = torch.empty((n, c), dtype=torch.bfloat16, device=sym.device)
    yf = torch.empty((n, c), dtype=torch.float32, device=sym.device) if want_f32 else None
    check(lib.pcc_gauss_dequant(_ptr(sym), _ptr(params), _ptr(prow), _ptr(perm), _ptr(scale), _ptr(gain), bound,
                                n, c, _ptr(yb), _ptr(yf), _stream()))
    return yb, yf


def pmf_to_quantized_cdf(pmf: np.ndarray, precision: int = 16) -> np.ndarray:
    pmf = np.ascontiguousarray(pmf, dtype=np.float32)
    out = np.empty(pmf.shape[0] + 1, dtype=np.int32)
    check(lib.pcc_pmf_to_quantized_cdf(pmf.ctypes.data, pmf.shape[0], precision, out.ctypes.data))
    return out


class DeviceCdfTables:
    """Quantised CDF tables resident on the device (flat cdf + per-table offset/size/symbol offset)."""

    def __init__(self, cdf_rows, offsets, device):
        sizes = np.array([len(r) for r in cdf_rows], dtype=np.int32)
        offs = np.zeros(len(cdf_rows), dtype=np.int32)
        offs[1:] = np.cumsum(sizes)[:-1]
        flat = np.concatenate([np.asarray(r, dtype=np.uint32) for r in cdf_rows]).astype(np.uint32)
        self.num_tables = len(cdf_rows)
        self.cdf = torch.from_numpy(flat.view(np.int32)).to(device)
        self.cdf_off = torch.from_numpy(offs).to(device)
        self.cdf_size = torch.from_numpy(sizes).to(device)
        self.offset = torch.from_numpy(np.asarray(offsets, dtype=np.int32)).to(device)
        self.struct = CdfTables(self.cdf.data_ptr(), self.cdf_off.data_ptr(), self.cdf_size.data_ptr(),
                                self.offset.data_ptr(), None, self.num_tables, int(flat.shape[0]))
        # decode search windows (built on the device once per table set)
        self.lut = torch.empty(lib.pcc_rans_lut_entries(self.num_tables), dtype=torch.int16, device=device)
        with torch.cuda.device(device):
            check(lib.pcc_rans_build_lut(C.byref(self.struct), self.lut.data_ptr(), _stream()))
        self.struct = CdfTables(self.cdf.data_ptr(), self.cdf_off.data_ptr(), self.cdf_size.data_ptr(),
                                self.offset.data_ptr(), self.lut.data_ptr(), self.num_tables, int(flat.shape[0]))


DEFAULT_SEG = 1 << 15


def auto_seg_size(n: int) -> int:
    """Symbols per segment (= per warp).  The coder's run time is its critical path, seg_size / 32
    dependent rounds, so streams are cut into >= ~64 segments; a segment costs 136 bytes of header
    (32 lane states + counts), which bounds it below.  The value travels in the container header."""
    seg = 4096
    while seg < DEFAULT_SEG and seg * 64 < n:
        seg *= 2
    return seg


def rans_encode(sym: torch.Tensor, idx: Optional[torch.Tensor], chan_len: int, tables: DeviceCdfTables,
                seg_size: Optional[int] = None):
    """sym: flat int32 device tensor.  Returns (device uint8 buffer, device int32 byte count)."""
    _need_cuda(sym, torch.int32)
    n = sym.numel()
    if seg_size is None:
        seg_size = auto_seg_size(n)
    dev = sym.device
    out = torch.empty(lib.pcc_rans_max_bytes(n, seg_size), dtype=torch.uint8, device=dev)
    nbytes = torch.empty(1, dtype=torch.int32, device=dev)
    wsb = lib.pcc_rans_workspace_bytes(n, seg_size)
    ws = workspace(wsb, dev, "rans")
    with _prof("rans_encode", n * 6):      # int32 symbol + u16 index per symbol (+ emitted bytes)
        check(lib.pcc_rans_encode(_ptr(sym), _ptr(idx), chan_len, n, seg_size, C.byref(tables.struct), _ptr(out),
                                  _ptr(nbytes), _ptr(ws), ws.numel(), _stream()))
    return out, nbytes


def rans_encode_batch(sym: torch.Tensor, idx: Optional[torch.Tensor], chan_len: int, tables: DeviceCdfTables,
                      seg_size: Optional[int] = None):
    """sym (and idx): [Q, ...] int32 (uint16) device tensors, one row of n symbols per quality level.  All Q
    streams are coded by one call (three launches, grid.y = level; codec_pipeline.py:414-432 is the batched
    call this matches).  Returns [(device uint8 buffer, device int32 byte count)] * Q, each an independent
    container identical to rans_encode of that row."""
    _need_cuda(sym, torch.int32)
    Q = sym.shape[0]
    n = sym.numel() // Q
    if seg_size is None:
        seg_size = auto_seg_size(n)
    dev = sym.device
    stride = (int(lib.pcc_rans_max_bytes(n, seg_size)) + 15) // 16 * 16
    out = torch.empty((Q, stride), dtype=torch.uint8, device=dev)
    nbytes = torch.empty(Q, dtype=torch.int32, device=dev)
    wsb = lib.pcc_rans_workspace_bytes(n, seg_size) * Q
    ws = workspace(wsb, dev, "rans")
    with _prof("rans_encode", Q * n * 6):
        check(lib.pcc_rans_encode_batch(_ptr(sym), _ptr(idx), chan_len, n, Q, seg_size, C.byref(tables.struct), _ptr(out),
                                        stride, _ptr(nbytes), _ptr(ws), ws.numel(), _stream()))
    return [(out[q], nbytes[q:q + 1]) for q in range(Q)]


def rans_decode(container: torch.Tensor, seg_size: int, idx: Optional[torch.Tensor], chan_len: int, n: int,
                tables: DeviceCdfTables) -> torch.Tensor:
    _need_cuda(container, torch.uint8)
    dev = container.device
    sym = torch.empty(n, dtype=torch.int32, device=dev)
    status = torch.empty(1, dtype=torch.int32, device=dev)
    with _prof("rans_decode", n * 6):
        check(lib.pcc_rans_decode(_ptr(container), container.numel(), seg_size, _ptr(idx), chan_len, n,
                                  C.byref(tables.struct), _ptr(sym), _ptr(status), _stream()))
    return sym, status
