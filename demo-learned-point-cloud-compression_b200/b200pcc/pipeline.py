"""Drop-in codec pipelines: ``CompressionPipeline(settings).compress(gop)`` and
``DecompressionPipeline().decompress(bytes)`` with the reference's signatures, output keys and
container (sender/encoder/codec_pipeline.py:21-236, receiver/decoder/codec_parallel.py:18-171;
contract in SURVEY.md 8b).  The reference runs one Python thread per stage joined by queues and
shares one result dict between them; here a call is one stream-ordered sequence of kernel launches,
serialised per pipeline object by a lock (the callers use a 3-worker pool: encoder.py:50,75,
decoder.py:47), so concurrent callers cannot receive each other's results.
"""
import os
import threading
import time
from typing import List

import numpy as np
import torch

from . import ops, container, geometry
from .model import ColorModel
from .sparse import CoordMap, SparseTensor

_MODEL_CACHE = {}
_MODEL_LOCK = threading.Lock()


def load_model(device, config="synth_small", state_dict=None) -> ColorModel:
    """Mirror of CompressionPipeline.load_model (codec_pipeline.py:56-72): build, load weights,
    move to the device, update() the entropy tables, eval()."""
    m = ColorModel(config)
    if state_dict is not None:
        m.load_state_dict(state_dict)
    m.to(device)
    m.update()
    m.eval()
    return m


def _shared_model(device, config):
    key = (str(device), config if isinstance(config, str) else id(config))
    with _MODEL_LOCK:
        if key not in _MODEL_CACHE:
            _MODEL_CACHE[key] = load_model(device, config)
        return _MODEL_CACHE[key]


_NVTX = os.environ.get("PCC_NVTX", "0") not in ("", "0")
ENC_STAGES = ("unpack", "analysis", "geometry_compression", "hyper_analysis", "factorized_model", "hyper_synthesis",
              "gaussian_model")
DEC_STAGES = ("bitstream_reading", "geometry_decompression", "factorized_model", "hyper_synthesis", "guassian_model",
              "synthesis_transform")


class _StageTimer:
    """CUDA-event stage timing resolved after the call's final sync (the reference brackets each
    stage with torch.cuda.synchronize(): codec_pipeline.py:270-291).  mark(name) closes the stage `name`.
    With PCC_NVTX=1 every stage is also an NVTX range named after the reference's stage key
    (`pcc.enc.analysis`, `pcc.dec.guassian_model`, ... -- codec_pipeline.py:218-225, codec_parallel.py:156-163) for
    Nsight timelines; `order` is the sequence of stage keys the caller is going to close."""

    def __init__(self, order=None, prefix="pcc"):
        self.marks = []
        self.order = order if _NVTX else None
        self.prefix = prefix
        self.open = False
        self.dev = None

    _POOL = threading.local()      # timing events are recycled per thread (creating one costs more than recording it)

    def mark(self, name):
        pools = getattr(self._POOL, "free", None)
        if pools is None:
            pools = self._POOL.free = {}
        if self.dev is None:
            self.dev = torch.cuda.current_device()       # an event belongs to the device it was first recorded on
        free = pools.setdefault(self.dev, [])
        ev = free.pop() if free else torch.cuda.Event(enable_timing=True)
        ev.record(ops.stream_obj())
        self.marks.append((name, ev, time.time()))
        if self.order is not None:
            if self.open:
                torch.cuda.nvtx.range_pop()
                self.open = False
            i = len(self.marks) - 1          # mark 0 is "start": stage order[i] begins now
            if i < len(self.order):
                torch.cuda.nvtx.range_push(f"{self.prefix}.{self.order[i]}")
                self.open = True

    def resolve(self):
        if self.open:
            torch.cuda.nvtx.range_pop()
            self.open = False
        out = {}
        for (n0, e0, _), (n1, e1, _) in zip(self.marks[:-1], self.marks[1:]):
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1) * 1e-3
        pools = getattr(self._POOL, "free", None)
        if pools is not None and self.dev is not None and len(pools.setdefault(self.dev, [])) < 64:
            pools[self.dev].extend(ev for _, ev, _ in self.marks)      # (resolve() runs after the call's final sync)
        self.marks = []
        return out


def _require_cuda(device):
    if not torch.cuda.is_available():
        raise RuntimeError("b200pcc needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")


_COPY_POOL = None
_COPY_THREADS = 4


def _parallel_copy(dst: np.ndarray, src: np.ndarray) -> None:
    """dst[:] = src for flat uint8 arrays headed for a DMA: anything sizeable goes through
    pcc_host_copy_stream (non-temporal stores, so the DMA engine does not read cache-dirty lines)."""
    n = src.size
    if n < (64 << 10):
        dst[:n] = src
        return
    from ._lib import lib
    lib.pcc_host_copy_stream(src.ctypes.data, n, dst.ctypes.data)


class _Stager:
    """Pinned host staging so the per-step host->device copies are asynchronous DMA."""

    def __init__(self):
        self.buf = None
        self.busy = None        # event recorded after the last DMA that reads self.buf

    def wait_free(self):
        """The pinned buffer is written by the host and read by an asynchronous DMA: before the host writes it
        again, the previous copy out of it must have finished (stage_frames() may be called back to back)."""
        if self.busy is not None:
            self.busy.synchronize()
            self.busy = None

    def mark_busy(self, stream=None):
        self.busy = torch.cuda.Event()
        self.busy.record(stream if stream is not None else torch.cuda.current_stream())

    def to_device(self, arrays: List[np.ndarray], device, pad4: bool = False) -> List[torch.Tensor]:
        """pad4: uint8 arrays come back with their length rounded up to 4 (zero filled)."""
        total = sum((a.nbytes + 255) // 256 * 256 for a in arrays)
        self.wait_free()
        if self.buf is None or self.buf.numel() < total:
            self.buf = torch.empty(max(total, 1 << 20), dtype=torch.uint8, pin_memory=True)
        dev = torch.empty(max(total, 1), dtype=torch.uint8, device=device)
        hv = self.buf.numpy()
        offs, pos = [], 0
        for a in arrays:
            b = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
            _parallel_copy(hv[pos:pos + b.size], b)
            hv[pos + b.size:pos + (b.size + 15) // 16 * 16] = 0      # readers may round a length up
            offs.append(pos)
            pos += (b.size + 255) // 256 * 256
        dev[:pos].copy_(self.buf[:pos], non_blocking=True)
        self.mark_busy()
        outs = []
        for a, o in zip(arrays, offs):
            if pad4 and a.dtype == np.uint8 and a.ndim == 1:
                outs.append(dev[o:o + (a.nbytes + 3) // 4 * 4])
            else:
                outs.append(dev[o:o + a.nbytes].view(_TORCH_DT[a.dtype.name]).view(a.shape))
        return outs


import os as _os
_STAGE_PIECE = int(float(_os.environ.get("PCC_STAGE_PIECE_MB", "1")) * (1 << 20))   # host copy and DMA are pipelined in pieces of this size
def _default_stage_threads() -> int:
    # six threads saturate the host copy on one GPU; ranks of one node share the cores (torchrun exports
    # LOCAL_WORLD_SIZE), so each takes its share but never fewer than two
    ranks = max(int(_os.environ.get("LOCAL_WORLD_SIZE", _os.environ.get("WORLD_SIZE", "1")) or 1), 1)
    try:
        cores = len(_os.sched_getaffinity(0))       # what this process may run on (a launcher may have pinned it)
    except (AttributeError, OSError):
        cores = _os.cpu_count() or 8
    return max(2, min(6, cores // ranks))


_STAGE_THREADS = int(_os.environ.get("PCC_STAGE_THREADS", "0")) or _default_stage_threads()
_STAGE_MODE = _os.environ.get("PCC_STAGE_MODE", "split")


def _stage_concat(stager: "_Stager", arrays: List[np.ndarray], torch_dtype, width: int, dev_buf: torch.Tensor, stream=None,
                  after=None, background=False, to_bf16=False):
    """Rows of all arrays back to back through pinned memory into dev_buf (viewed [rows, width]), on
    `stream` (after event `after`) when given, else on the current stream.  The host copy into pinned
    memory and the DMA run at similar speeds, so the data moves in pieces (`pcc_stage_h2d`: native
    threads copy a piece and enqueue its DMA themselves -- Python threads doing the same serialised on
    the interpreter lock, one acquisition per piece).  to_bf16: float64 sources are converted on the way
    (double -> float -> bf16, the device path's roundings) and arrive as 2 bytes per value; torch_dtype
    then names the device view.  background=True: the call runs on a helper thread
    (it releases the GIL) while the caller keeps enqueuing GPU work; the returned wait() joins it.
    -> (device view, completion event / wait function / None)."""
    global _COPY_POOL
    import ctypes as C
    from ._lib import lib, check
    total = sum(a.nbytes for a in arrays) // (4 if to_bf16 else 1)
    stager.wait_free()
    if stager.buf is None or stager.buf.numel() < total:
        stager.buf = torch.empty(max(total, 1 << 20), dtype=torch.uint8, pin_memory=True)
    st = stream if stream is not None else torch.cuda.current_stream()
    if stream is not None and after is not None:
        stream.wait_event(after)
    n = len(arrays)
    srcs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrays])
    sizes = (C.c_int64 * max(n, 1))(*[a.nbytes for a in arrays])
    pinned_ptr, dev_ptr, st_ptr, device = stager.buf.data_ptr(), dev_buf.data_ptr(), st.cuda_stream, dev_buf.device

    def job():
        with torch.cuda.device(device):      # a helper thread starts on device 0
            check(lib.pcc_stage_h2d(srcs, sizes, n, pinned_ptr, dev_ptr, _STAGE_PIECE, _STAGE_THREADS, int(to_bf16), st_ptr))
            stager.mark_busy(st)
        return arrays       # the sources stay alive until their copies are enqueued
    view = dev_buf[:total].view(torch_dtype).view(-1, width)
    if background and total > _STAGE_PIECE:
        if _COPY_POOL is None:
            from concurrent.futures import ThreadPoolExecutor
            _COPY_POOL = ThreadPoolExecutor(max_workers=_COPY_THREADS, thread_name_prefix="b200pcc-stage")
        future = _COPY_POOL.submit(job)

        def wait():
            future.result()
            ev = torch.cuda.Event()
            ev.record(st)
            return ev
        return view, wait
    if total:
        job()
    if background:       # small frame: copied inline (a helper would only get the GIL when this thread blocks)
        def wait():
            ev = torch.cuda.Event()
            ev.record(st)
            return ev
        return view, wait
    done = None
    if stream is not None:
        done = torch.cuda.Event()
        done.record(stream)
    return view, done


_TORCH_DT = {"int16": torch.int16, "float64": torch.float64, "int32": torch.int32, "float32": torch.float32,
             "uint8": torch.uint8, "int64": torch.int64}


def _gop_key_bits(num_frames: int) -> int:
    """Significant bits of the coordinate keys of a GOP: 48 Morton bits + the frame index (keys are b << 48 | morton)."""
    return 48 + max(0, (int(num_frames) - 1).bit_length())


class CompressionPipeline:
    def __init__(self, settings, device=None, config="synth_small", model: ColorModel = None):
        self.device = _require_cuda(device)
        self.settings = [list(map(float, s)) for s in settings]
        self.compression_model = model if model is not None else _shared_model(self.device, config)
        self._lock = threading.Lock()
        self._stager = _Stager()
        self._stager_c = _Stager()      # colours have their own pinned buffer (copied while the points are in flight)
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.keep_debug = False     # tests: keep stage tensors of the last call in self.debug
        self.debug = {}

    # ---- reference-shaped steps -------------------------------------------------------------------------
    def stage_frames(self, frames):
        """Host->device copy of the raw frames (pinned staging, async DMA).  Returns device tensors."""
        pts, cols = [], []
        for item in frames:
            if "points" not in item:      # "Sometimes we miss a frame?" (codec_pipeline.py:247-249)
                continue
            pts.append(np.ascontiguousarray(item["points"], dtype=np.int16))
            cols.append(np.ascontiguousarray(item["colors"], dtype=np.float64))
        staged = self._stager.to_device(pts + cols, self.device)
        self.h2d_bytes += sum(a.nbytes for a in pts) + sum(a.nbytes for a in cols)
        nf = len(pts)
        return {"points": staged[:nf], "colors": staged[nf:]}

    def _unpack_host_frames(self, frames, defer_features, speculate=False):
        """Host frames -> sparse tensor with the copies pipelined against the coordinate work: the points
        go first and their keys are sorted while native threads round the float64 colours to bf16 (the
        precision the stem convolution reads them in: double -> float -> bf16, the device path's roundings)
        into pinned memory, piece by piece, each piece's DMA following on a side stream; the colours are
        only waited for when the feature rows are built (after the coordinate pyramid when defer_features
        is set).  A frame of n points moves 12 n bytes instead of 30 n."""
        pts, cols = [], []
        for item in frames:
            if "points" not in item:      # "Sometimes we miss a frame?" (codec_pipeline.py:247-249)
                continue
            pts.append(np.ascontiguousarray(item["points"], dtype=np.int16))
            cols.append(np.ascontiguousarray(item["colors"], dtype=np.float64))
        nf = len(pts)
        counts = [int(p.shape[0]) for p in pts]
        total = sum(counts)
        dev = self.device
        pts_buf = torch.empty(max(total * 6, 8), dtype=torch.uint8, device=dev)
        cols_buf = torch.empty(max(total * 6, 8), dtype=torch.uint8, device=dev)     # bf16 [n,3]: converted while staged
        keys = torch.empty(total, dtype=torch.int64, device=dev)
        allocated = torch.cuda.Event()
        allocated.record()
        if total * 24 > _STAGE_PIECE:
            # one helper thread drives both copies back to back (the native calls release the GIL): points
            # first, then the colours, which therefore start converting at once instead of after this
            # thread has enqueued the sort -- they are on the device before the pyramid is
            global _COPY_POOL
            if _COPY_POOL is None:
                from concurrent.futures import ThreadPoolExecutor
                _COPY_POOL = ThreadPoolExecutor(max_workers=_COPY_THREADS, thread_name_prefix="b200pcc-stage")
            side = ops._side_stream(dev, "h2d")
            side.wait_event(allocated)
            pts_ready = threading.Event()
            state = {}

            def job():
                try:
                    with torch.cuda.device(dev):
                        state["pts_view"], _ = _stage_concat(self._stager, pts, torch.int16, 3, pts_buf, stream=side)
                        state["pts"] = torch.cuda.Event()
                        state["pts"].record(side)
                        pts_ready.set()
                        state["cols_view"], state["cols"] = _stage_concat(self._stager_c, cols, torch.bfloat16, 3, cols_buf,
                                                                          stream=side, to_bf16=True)
                finally:
                    pts_ready.set()
            future = _COPY_POOL.submit(job)
            pts_ready.wait()
            if "pts" not in state:
                future.result()          # raises what the helper raised
            torch.cuda.current_stream().wait_event(state["pts"])
            pts_d = state["pts_view"]

            def arrived():
                future.result()
                return state["cols"]
            cols_view = lambda: state["cols_view"]
        else:
            pts_d, _ = _stage_concat(self._stager, pts, torch.int16, 3, pts_buf)
            cols_d0, arrived = _stage_concat(self._stager_c, cols, torch.bfloat16, 3, cols_buf,
                                             stream=ops._side_stream(dev, "h2d"), after=allocated, background=True, to_bf16=True)
            cols_view = lambda: cols_d0
        pos = 0
        for b in range(nf):
            ops.ingest_points(pts_d[pos:pos + counts[b]], b, keys[pos:pos + counts[b]])
            pos += counts[b]
        ukeys_all, src_all, cnt = ops.sort_unique(keys, want_src=True, sync=False, key_bits=_gop_key_bits(nf))
        self.h2d_bytes += total * 12
        # speculate: capturer frames hold no duplicate points, so carry on as if all rows were unique and
        # let compress() check the device count with the pyramid's read-back (one host wait fewer)
        m = total if speculate else int(cnt.item())
        x = SparseTensor(_cmap=CoordMap(ukeys_all[:m], 0), _feats=None, _channels=4)
        x.num_frames = nf
        x._unique = (cnt, ukeys_all, src_all) if speculate else None

        def finish_features():
            torch.cuda.current_stream().wait_event(arrived())      # joins the copy workers
            rows = x.cmap.n
            x._feats = ops.ingest_features_bf16(cols_view(), src_all[:rows], rows, 16)
        if defer_features:
            x._finish_features = finish_features
        else:
            finish_features()
        return x

    def unpack_batch(self, gop, staged=None, defer_features=False, speculate=False):
        """frames list -> batched sparse tensor (codec_pipeline.py:239-267, shared/utils.py:10-42).
        `staged` (from stage_frames) skips the host->device copy: inputs already resident in HBM.
        defer_features: the caller runs x._finish_features() once it needs the feature rows."""
        frames = gop.pop("frames")
        if staged is None and _STAGE_MODE != "simple":
            return self._unpack_host_frames(frames, defer_features, speculate), gop
        if staged is None:
            staged = self.stage_frames(frames)
        pts, cols = staged["points"], staged["colors"]
        dev = self.device
        total = sum(int(p.shape[0]) for p in pts)
        keys = torch.empty(total, dtype=torch.int64, device=dev)
        feats = torch.empty((total, 4), dtype=torch.float32, device=dev)
        pos = 0
        nf = len(pts)
        for b in range(nf):
            n = int(pts[b].shape[0])
            ops.ingest_frame(pts[b], cols[b], b, keys[pos:pos + n], feats[pos:pos + n])
            pos += n
        if speculate:      # see _unpack_host_frames
            ukeys_all, src_all, cnt = ops.sort_unique(keys, want_src=True, sync=False, key_bits=_gop_key_bits(nf))
            x = SparseTensor(_cmap=CoordMap(ukeys_all, 0), _feats=ops.rows_f32_to_bf16(feats, 16, src_all), _channels=4)
            x._unique = (cnt, ukeys_all, src_all)
        else:
            ukeys, src = ops.sort_unique(keys, want_src=True, key_bits=_gop_key_bits(nf))
            x = SparseTensor(_cmap=CoordMap(ukeys, 0), _feats=ops.rows_f32_to_bf16(feats, 16, src), _channels=4)
        x.num_frames = nf
        return x, gop

    def _coordinate_pyramid(self, x, nf):
        """Every coarser coordinate map (g_a: 3, h_a: 2) with one host wait, which also settles the number
        of unique input points when unpack_batch() speculated on it.  The stride-1 kernel map and the stem
        convolution do not depend on the pyramid: they run while the host waits and does its bookkeeping."""
        g_a = self.compression_model.g_a
        unique = getattr(x, "_unique", None)
        x.cmap.build_pyramid(5, nf, overlap=lambda: g_a.prelaunch_stem(x), extra=unique[0] if unique else None)
        if unique is None:
            return x
        x._unique = None
        m = int(x.cmap.extra_host)
        if m == x.cmap.n:
            return x
        # duplicate input points (first wins): everything enqueued so far ran on rows that do not exist --
        # redo the coordinate work for the m unique rows
        _, ukeys_all, src_all = unique
        y = SparseTensor(_cmap=CoordMap(ukeys_all[:m], 0), _feats=None if x._feats is None else x._feats[:m],
                         _channels=x._channels)
        y.num_frames = nf
        y._finish_features = getattr(x, "_finish_features", None)
        if y._finish_features is not None:
            finish, y._finish_features = y._finish_features, None

            def finish_features():
                x.cmap = y.cmap
                finish()
                y._feats = x._feats
            y._finish_features = finish_features
        y.cmap.build_pyramid(5, nf, overlap=lambda: g_a.prelaunch_stem(y))
        return y

    def analysis_step(self, x):
        return self.compression_model.g_a(x)

    def hyper_analysis_step(self, y):
        return self.compression_model.entropy_model.h_a(y)

    def factorized_model_step_batched(self, z):
        """codec_pipeline.py:294-317.  The z string stays on the device until compress() collects all
        strings at once; z_hat = round(z - median) + median is what re-decoding the string would give."""
        em = self.compression_model.entropy_model
        perm = z.cmap.lex_perm()
        z_pending, zhat = em.entropy_bottleneck.compress_rows_device(z._feats, perm)
        z_hat = SparseTensor(_cmap=z.cmap, _feats=zhat, _channels=z._channels)
        if self.keep_debug:
            self.debug.update(z=z, z_hat=z_hat)
        return z_hat, [z_pending], [z.cmap.n]

    def hyper_synthesis_step(self, z_hat):
        return self.compression_model.entropy_model.h_s(z_hat)

    def geometry_compression_step(self, ycm):
        """Per-frame coordinate streams for the stride-8 latent coordinates (replaces gpcc_encode,
        codec_pipeline.py:441-462): built on the device right after the coordinate pyramid; the
        streams stay there until compress() fetches every string of the GOP at once."""
        return geometry.encode_device(ycm)

    def gaussian_model_step_batched(self, y, settings, gaussian_params):
        """All qualities in one pass (codec_pipeline.py:397-437): gather h_s rows at the y
        coordinates, gain, build_indexes, quantise, then one rANS container per quality (left on the
        device; compress() fetches every string of the GOP with one size sync)."""
        em = self.compression_model.entropy_model
        gc = em.gaussian_conditional
        prow = ops.hash_query(y.cmap.keys, gaussian_params.cmap.table())
        perm = y.cmap.lex_perm()
        scale = em.scale_for(settings, self.device)
        sym, idx = ops.gauss_encode(y._feats, gaussian_params._feats, prow, perm, scale, gc.table_d, gc.scale_bound)
        if self.keep_debug:
            self.debug.update(y=y, gaussian_params=gaussian_params, y_sym=sym, y_idx=idx, scale=scale)
        nq = len(settings)
        pending = gc._encode_device_batch(sym.reshape(nq, -1), idx.reshape(nq, -1), 1)
        return pending, [y.cmap.n]

    # ---- entry point -----------------------------------------------------------------------------------------
    def compress(self, data, staged=None):
        """gop dict -> (compressed_data {0: raw frames, i+1: bytes for settings[i]}, sideinfo)."""
        from .model import fetch_strings
        with self._lock, torch.cuda.device(self.device), ops.stream_scope():
            t_start = time.time()
            compressed_data = {0: data["frames"]}
            tm = _StageTimer(ENC_STAGES, "pcc.enc")
            tm.mark("start")
            x, sideinfo = self.unpack_batch(data, staged, defer_features=True, speculate=True)
            nf = x.num_frames
            x = self._coordinate_pyramid(x, nf)
            num_points = x.cmap.n
            ycm = x.cmap.coarse.coarse.coarse
            if getattr(x, "_finish_features", None) is not None:
                x._finish_features()          # waits for the colours, which travelled during the coordinate work
                x._finish_features = None
            tm.mark("unpack")
            y, k = self.analysis_step(x)
            tm.mark("analysis")
            # (after the analysis transform is enqueued: its host-side preparation then overlaps GPU work)
            geo_pending = self.geometry_compression_step(ycm)
            tm.mark("geometry_compression")
            z = self.hyper_analysis_step(y)
            tm.mark("hyper_analysis")
            z_hat, z_pending, z_shapes = self.factorized_model_step_batched(z)
            tm.mark("factorized_model")
            gaussian_params = self.hyper_synthesis_step(z_hat)
            tm.mark("hyper_synthesis")
            y_pending, y_shapes = self.gaussian_model_step_batched(y, self.settings, gaussian_params)
            tm.mark("gaussian_model")
            strings = fetch_strings(z_pending + y_pending + geo_pending, as_bytes=False)   # the GOP's two remaining syncs
            z_strings, y_strings = strings[:1], strings[1:1 + len(y_pending)]
            points_streams = strings[1 + len(y_pending):]
            self.d2h_bytes += sum(len(b) for b in strings)
            t7 = []
            for i, q in enumerate(self.settings):
                t0 = time.time()
                compressed_data[i + 1] = container.make_bitstream_batched(
                    y_strings[i], z_strings[0], y_shapes[0], z_shapes[0], points_streams, k, q)
                t7.append(time.time() - t0)
            st = tm.resolve()
            sideinfo["enc_time_measurements"] = {
                "analysis": st.get("analysis", 0.0) + st.get("unpack", 0.0),
                "hyper_analysis": st.get("hyper_analysis", 0.0),
                "factorized_model": st.get("factorized_model", 0.0),
                "hyper_synthesis": st.get("hyper_synthesis", 0.0),
                "geometry_compression": st.get("geometry_compression", 0.0),
                "gaussian_model": st.get("gaussian_model", 0.0),
                "bitstream_writing": t7,
            }
            bw = [8 * 6 * num_points if idx == 0 else len(d) * 8 for idx, (key, d) in enumerate(compressed_data.items())]
            sideinfo["gop_info"] = {"num_points": num_points, "bandwidth": bw,
                                    "bpp": [b / max(num_points, 1) for b in bw]}
            sideinfo.setdefault("timestamps", {})
            sideinfo["timestamps"]["codec_start"] = t_start
            sideinfo["timestamps"]["codec_end"] = time.time()
            return compressed_data, sideinfo


class DecompressionPipeline:
    def __init__(self, device=None, config="synth_small", model: ColorModel = None):
        self.device = _require_cuda(device)
        self.decompression_model = model if model is not None else _shared_model(self.device, config)
        self._lock = threading.Lock()
        self._stager = _Stager()
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.keep_debug = False
        self.debug = {}

    def read_bitstream_batched(self, compressed_data):
        return container.read_bitstream_batched(compressed_data, copy=False)

    def stage_bitstream(self, points_streams, y_strings, z_strings):
        """One pinned staging copy for everything the decoder needs on the device: the coordinate
        streams (+ their offsets / lengths / row offsets) and the two entropy-coded strings (padded to
        4 bytes for the lane decoder).  -> (device tensors, per-frame latent point counts)."""
        blob, offs, meta32, counts = geometry.pack_streams(points_streams)
        arrays = [blob, offs, meta32, np.frombuffer(y_strings[0], dtype=np.uint8),
                  np.frombuffer(z_strings[0], dtype=np.uint8)]
        dev = self._stager.to_device(arrays, self.device, pad4=True)
        self.h2d_bytes += sum(a.nbytes for a in arrays)
        return dev, counts

    def geometry_decompression_step(self, blob_d, offs_d, meta32_d, counts):
        """Per-frame coordinate streams -> the sorted stride-8 coordinate map of the latents, on the
        device (replaces gpcc_decode + stack_tensors, codec_parallel.py:266-289).  The octree expands
        in ascending key order, so no sort / de-duplication pass follows."""
        ykeys, status = geometry.decode_device(blob_d, offs_d, meta32_d, counts)
        ycm = CoordMap(ykeys, geometry.LOG2_STRIDE)
        ycm.num_frames = len(counts)
        return ycm, status

    def factorized_model_step_batched(self, z_string_dev, z_hdr, z_shape, ycm):
        """Rebuild the stride-32 coordinates from the y coordinates and decode z_hat
        (codec_parallel.py:291-318; g_s.down_conv is used for coordinates only).  No host sync: the
        sizes come from the bitstream and are verified at the end of decompress()."""
        m = self.decompression_model
        zkeys = ops.coords_pyramid_expect(ycm.keys, 3, 2, z_shape)
        zcm = CoordMap(zkeys, 5)
        zcm.num_frames = ycm.num_frames
        eb = m.entropy_model.entropy_bottleneck
        sym, status = eb._decode_device(z_string_dev, z_hdr, None, zcm.n, zcm.n * eb.channels)
        zhat, _ = ops.factorized_dequant(sym.view(eb.channels, zcm.n), zcm.lex_perm(), eb.medians_d)
        return SparseTensor(_cmap=zcm, _feats=zhat, _channels=eb.channels), status

    def hyper_synthesis_step(self, z_hat):
        return self.decompression_model.entropy_model.h_s(z_hat)

    def gaussian_model_step_batched(self, y_string_dev, y_hdr, ycm, q, gaussian_params):
        """codec_parallel.py:382-419."""
        em = self.decompression_model.entropy_model
        gc = em.gaussian_conditional
        cy = self.decompression_model.config["c_y"]
        prow = ops.hash_query(ycm.keys, gaussian_params.cmap.table())
        perm = ycm.lex_perm()
        scale = em.scale_for([list(q)], self.device)
        idx = ops.gauss_indexes(gaussian_params._feats, prow, perm, scale, gc.table_d, gc.scale_bound, ycm.n, cy)
        sym, status = gc._decode_device(y_string_dev, y_hdr, idx.view(-1), 1, ycm.n * cy)
        yb, _ = ops.gauss_dequant(sym.view(cy, ycm.n), gaussian_params._feats, prow, perm, scale[0].contiguous(),
                                  em.offset_gain_d, gc.scale_bound)
        y_hat = SparseTensor(_cmap=ycm, _feats=yb, _channels=cy)
        if self.keep_debug:
            self.debug.update(y_sym=sym.view(cy, ycm.n), y_idx=idx, y_hat=y_hat, gaussian_params=gaussian_params)
        return y_hat, status

    def hyper_synthesis(self, y_hat, ks):
        """Step 6 (the reference names the synthesis step `hyper_synthesis`, codec_parallel.py:465-472)."""
        return self.decompression_model.g_s(y_hat, k=ks)

    def _ship_points_early(self, keys):
        """Called by g_s once the output coordinates are final (before the colour head): points leave for
        the host on a side stream while the colour head runs."""
        n = keys.shape[0]
        pts_d, _ = ops.unpack_output(keys, None)
        ready = torch.cuda.Event()
        ready.record()
        pts_h = torch.empty((n, 3), dtype=torch.int32, pin_memory=True)
        side = ops._side_stream(self.device, "d2h")
        with torch.cuda.stream(side):
            side.wait_event(ready)
            pts_h.copy_(pts_d, non_blocking=True)
            done = torch.cuda.Event()
            done.record()
        pts_d.record_stream(side)
        self._early = (keys, pts_h, done)

    def pack_batches(self, pointcloud, num_frames):
        """codec_parallel.py:474-502: split per frame, NaN -> 0, clip colours to [0,1].  The clip runs
        on the device; points (int32 [N,3]) and colours (float32 [N,3]) land in pinned host memory that
        the returned arrays own."""
        n = pointcloud.cmap.n
        early = getattr(self, "_early", None)
        self._early = None
        if early is not None and early[0] is pointcloud.cmap.keys:
            _, pts_h, self._pts_done = early
            _, col_d = ops.unpack_output(pointcloud.cmap.keys, pointcloud._feats, want_points=False)
        else:
            pts_d, col_d = ops.unpack_output(pointcloud.cmap.keys, pointcloud._feats)
            pts_h = torch.empty((n, 3), dtype=torch.int32, pin_memory=True)
            pts_h.copy_(pts_d, non_blocking=True)
            self._pts_done = None
        col_h = torch.empty((n, 3), dtype=torch.float32, pin_memory=True)
        col_h.copy_(col_d, non_blocking=True)
        counts = getattr(pointcloud, "frame_counts", None)
        if counts is None:
            off_d = ops.frame_offsets(pointcloud.cmap.keys, num_frames)
            off = off_d.cpu().numpy()
        else:
            off = np.concatenate([[0], np.cumsum(counts)])
        self.d2h_bytes += 24 * n
        return pts_h, col_h, off

    @staticmethod
    def _finish_batches(pts_h, col_h, off, num_frames):
        P, Cc = pts_h.numpy(), col_h.numpy()
        return [{"points": P[off[i]:off[i + 1]], "colors": Cc[off[i]:off[i + 1]]} for i in range(num_frames)]

    def playout_buffers(self, x_hat, num_frames: int = None, offset: float = 100.0):
        """Device reconstruction (decompress(..., to_host=False)) -> one playout buffer per frame, as the
        client builds them for the visualizer (receiver/client/client.py:139-146): float32 (points + offset)
        bytes followed by uint8 (255 * colours) bytes.  Packed on the device: 15 bytes per point cross
        the bus instead of 24, and the host neither scales nor converts."""
        with self._lock, torch.cuda.device(self.device), ops.stream_scope():
            keys = x_hat.cmap.keys
            n = x_hat.cmap.n
            nf = num_frames if num_frames is not None else (int((keys[-1:] >> 48).item()) + 1 if n else 0)
            pts_d, col_d = ops.pack_visualizer(keys, x_hat._feats, offset)
            pts_h = torch.empty((n, 3), dtype=torch.float32, pin_memory=True)
            col_h = torch.empty((n, 3), dtype=torch.uint8, pin_memory=True)
            pts_h.copy_(pts_d, non_blocking=True)
            col_h.copy_(col_d, non_blocking=True)
            off = ops.frame_offsets(keys, nf).cpu().numpy()          # (synchronises: the copies above are done)
            torch.cuda.current_stream().synchronize()
            self.d2h_bytes += 15 * n
            P, Cc = pts_h.numpy(), col_h.numpy()
            return [P[off[i]:off[i + 1]].tobytes() + Cc[off[i]:off[i + 1]].tobytes() for i in range(nf)]

    def decompress(self, compressed_data, to_host=True):
        """bytes -> (frames, sideinfo).  to_host=False leaves the reconstruction on the device (returns
        the SparseTensor instead of the per-frame numpy dicts)."""
        with self._lock, torch.cuda.device(self.device), ops.stream_scope():
            t_start = time.time()
            ops.clear_deferred()
            tm = _StageTimer(DEC_STAGES, "pcc.dec")
            tm.mark("start")
            y_strings, z_strings, y_shape, z_shape, points_streams, ks, q = self.read_bitstream_batched(compressed_data)
            nf = len(points_streams)
            tm.mark("bitstream_reading")
            (blob_d, offs_d, meta32_d, y_dev, z_dev), frame_counts = self.stage_bitstream(points_streams, y_strings,
                                                                                          z_strings)
            if sum(frame_counts) != y_shape:
                raise ValueError(f"bitstream y shape {y_shape} != coordinate streams {sum(frame_counts)}")
            ycm, st_geo = self.geometry_decompression_step(blob_d, offs_d, meta32_d, frame_counts)
            tm.mark("geometry_decompression")
            z_hat, st_z = self.factorized_model_step_batched(z_dev, z_strings[0][:16], z_shape, ycm)
            tm.mark("factorized_model")
            gaussian_params = self.hyper_synthesis_step(z_hat)
            tm.mark("hyper_synthesis")
            y_hat, st_y = self.gaussian_model_step_batched(y_dev, y_strings[0][:16], ycm, q, gaussian_params)
            y_hat.frame_counts = frame_counts
            self._early = None
            if to_host:
                y_hat.on_geometry = self._ship_points_early
            tm.mark("guassian_model")
            x_hat = self.hyper_synthesis(y_hat, ks)
            tm.mark("synthesis_transform")
            packed = self.pack_batches(x_hat, nf) if to_host else None
            # the call's single blocking point: every deferred size check and both decoder status words
            ops.verify_deferred([(st_geo, "coordinate stream decode failed"), (st_z, "entropy decode of z failed"),
                                 (st_y, "entropy decode of y failed")], device=self.device)
            if to_host and getattr(self, "_pts_done", None) is not None:
                self._pts_done.synchronize()          # the side-stream copy of the points
                self._pts_done = None
            final_data = self._finish_batches(*packed, nf) if to_host else x_hat
            st = tm.resolve()
            sideinfo = {"time_measurements": {}, "timestamps": {}}
            for key in ("bitstream_reading", "geometry_decompression", "factorized_model", "hyper_synthesis",
                        "guassian_model", "synthesis_transform"):
                sideinfo["time_measurements"][key] = st.get(key, 0.0)
            sideinfo["timestamps"]["codec_start"] = t_start
            sideinfo["timestamps"]["codec_end"] = time.time()
            return final_data, sideinfo
