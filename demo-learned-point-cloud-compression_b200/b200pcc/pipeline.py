"""Drop-in codec pipelines: ``CompressionPipeline(settings).compress(gop)`` and
``DecompressionPipeline().decompress(bytes)`` with the reference's signatures, output keys and
container (sender/encoder/codec_pipeline.py:21-236, receiver/decoder/codec_parallel.py:18-171;
contract in SURVEY.md 8b).  The reference runs one Python thread per stage joined by queues and
shares one result dict between them; here a call is one stream-ordered sequence of kernel launches,
serialised per pipeline object by a lock (the callers use a 3-worker pool: encoder.py:50,75,
decoder.py:47), so concurrent callers cannot receive each other's results.
"""
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


class _StageTimer:
    """CUDA-event stage timing resolved after the call's final sync (the reference brackets each
    stage with torch.cuda.synchronize(): codec_pipeline.py:270-291)."""

    def __init__(self):
        self.marks = []

    def mark(self, name):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self.marks.append((name, ev, time.time()))

    def resolve(self):
        out = {}
        for (n0, e0, _), (n1, e1, _) in zip(self.marks[:-1], self.marks[1:]):
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1) * 1e-3
        return out


def _require_cuda(device):
    if not torch.cuda.is_available():
        raise RuntimeError("b200pcc needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")


class _Stager:
    """Pinned host staging so the per-step host->device copies are asynchronous DMA."""

    def __init__(self):
        self.buf = None

    def to_device(self, arrays: List[np.ndarray], device) -> List[torch.Tensor]:
        total = sum((a.nbytes + 255) // 256 * 256 for a in arrays)
        if self.buf is None or self.buf.numel() < total:
            self.buf = torch.empty(max(total, 1 << 20), dtype=torch.uint8, pin_memory=True)
        dev = torch.empty(max(total, 1), dtype=torch.uint8, device=device)
        hv = self.buf.numpy()
        offs, pos = [], 0
        for a in arrays:
            b = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
            hv[pos:pos + b.size] = b
            offs.append(pos)
            pos += (b.size + 255) // 256 * 256
        dev[:pos].copy_(self.buf[:pos], non_blocking=True)
        outs = []
        for a, o in zip(arrays, offs):
            t = dev[o:o + a.nbytes].view(_TORCH_DT[a.dtype.name]).view(a.shape)
            outs.append(t)
        return outs


_TORCH_DT = {"int16": torch.int16, "float64": torch.float64, "int32": torch.int32, "float32": torch.float32,
             "uint8": torch.uint8, "int64": torch.int64}


class CompressionPipeline:
    def __init__(self, settings, device=None, config="synth_small", model: ColorModel = None):
        self.device = _require_cuda(device)
        self.settings = [list(map(float, s)) for s in settings]
        self.compression_model = model if model is not None else _shared_model(self.device, config)
        self._lock = threading.Lock()
        self._stager = _Stager()
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

    def unpack_batch(self, gop, staged=None):
        """frames list -> batched sparse tensor (codec_pipeline.py:239-267, shared/utils.py:10-42).
        `staged` (from stage_frames) skips the host->device copy: inputs already resident in HBM."""
        frames = gop.pop("frames")
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
        ukeys, src = ops.sort_unique(keys, want_src=True)
        x = SparseTensor(_cmap=CoordMap(ukeys, 0), _feats=ops.rows_f32_to_bf16(feats, 16, src), _channels=4)
        x.num_frames = nf
        return x, gop

    def analysis_step(self, x):
        return self.compression_model.g_a(x)

    def hyper_analysis_step(self, y):
        return self.compression_model.entropy_model.h_a(y)

    def factorized_model_step_batched(self, z):
        em = self.compression_model.entropy_model
        perm = z.cmap.lex_perm()
        z_strings, zhat = em.entropy_bottleneck.compress_rows(z._feats, perm)
        z_hat = SparseTensor(_cmap=z.cmap, _feats=zhat, _channels=z._channels)
        if self.keep_debug:
            self.debug.update(z=z, z_hat=z_hat)
        return z_hat, z_strings, [z.cmap.n]

    def hyper_synthesis_step(self, z_hat):
        return self.compression_model.entropy_model.h_s(z_hat)

    def geometry_compression_step(self, y, num_frames):
        """Per-frame coordinate streams for the stride-8 latent coordinates (replaces gpcc_encode,
        codec_pipeline.py:441-462)."""
        C = y.C.cpu().numpy()
        self.d2h_bytes += C.nbytes
        streams = []
        for b in range(num_frames):
            streams.append(geometry.encode_points(C[C[:, 0] == b][:, 1:], 8))
        return streams

    def gaussian_model_step_batched(self, y, settings, gaussian_params):
        """All qualities in one pass (codec_pipeline.py:397-437): gather h_s rows at the y
        coordinates, gain, build_indexes, quantise, then one rANS container per quality."""
        em = self.compression_model.entropy_model
        gc = em.gaussian_conditional
        prow = ops.hash_query(y.cmap.keys, gaussian_params.cmap.table())
        perm = y.cmap.lex_perm()
        scale = em.scale_for(settings, self.device)
        sym, idx = ops.gauss_encode(y._feats, gaussian_params._feats, prow, perm, scale, gc.table_d, gc.scale_bound)
        if self.keep_debug:
            self.debug.update(y=y, gaussian_params=gaussian_params, y_sym=sym, y_idx=idx, scale=scale)
        strings = []
        pending = []
        for qi in range(len(settings)):
            pending.append(ops.rans_encode(sym[qi].reshape(-1), idx[qi].reshape(-1), 1, gc.tables, gc.seg_size))
        sizes = torch.cat([nb for _, nb in pending]).cpu().numpy()   # one sync for all qualities
        for (buf, _), nb in zip(pending, sizes):
            strings.append(buf[: int(nb)].cpu().numpy().tobytes())
            self.d2h_bytes += int(nb)
        return strings, [y.cmap.n]

    # ---- entry point -----------------------------------------------------------------------------------------
    def compress(self, data, staged=None):
        """gop dict -> (compressed_data {0: raw frames, i+1: bytes for settings[i]}, sideinfo)."""
        with self._lock, torch.cuda.device(self.device):
            t_start = time.time()
            compressed_data = {0: data["frames"]}
            tm = _StageTimer()
            tm.mark("start")
            x, sideinfo = self.unpack_batch(data, staged)
            num_points = x.cmap.n
            nf = x.num_frames
            tm.mark("unpack")
            y, k = self.analysis_step(x)
            tm.mark("analysis")
            z = self.hyper_analysis_step(y)
            tm.mark("hyper_analysis")
            z_hat, z_strings, z_shapes = self.factorized_model_step_batched(z)
            tm.mark("factorized_model")
            gaussian_params = self.hyper_synthesis_step(z_hat)
            tm.mark("hyper_synthesis")
            points_streams = self.geometry_compression_step(y, nf)
            tm.mark("geometry_compression")
            y_strings, y_shapes = self.gaussian_model_step_batched(y, self.settings, gaussian_params)
            tm.mark("gaussian_model")
            t7 = []
            for i, q in enumerate(self.settings):
                t0 = time.time()
                compressed_data[i + 1] = container.make_bitstream_batched(
                    y_strings[i], z_strings[0], y_shapes[0], z_shapes[0], points_streams, k, q)
                t7.append(time.time() - t0)
            torch.cuda.synchronize()
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
        return container.read_bitstream_batched(compressed_data)

    def geometry_decompression_step(self, points_streams):
        """Per-frame coordinate streams -> stacked (b,x,y,z) int32 (replaces gpcc_decode + stack_tensors,
        codec_parallel.py:266-289)."""
        out = []
        for b, s in enumerate(points_streams):
            p = geometry.decode_points(s, 8)
            out.append(np.concatenate([np.full((p.shape[0], 1), b, dtype=np.int32), p], axis=1))
        return np.concatenate(out, axis=0) if out else np.zeros((0, 4), np.int32)

    def factorized_model_step_batched(self, z_strings, z_shape, y_points_dev, num_frames):
        """Rebuild the stride-32 coordinates from the y coordinates and decode z_hat
        (codec_parallel.py:291-318; g_s.down_conv is used for coordinates only)."""
        m = self.decompression_model
        ykeys, _ = ops.sort_unique(ops.keys_from_coords(y_points_dev), want_src=False)
        ycm = CoordMap(ykeys, 3)
        zcm = ycm.downsample().downsample()
        if zcm.n != z_shape:
            raise ValueError(f"bitstream z shape {z_shape} != coordinates {zcm.n}")
        eb = m.entropy_model.entropy_bottleneck
        zhat, status = eb.decompress_rows(z_strings, zcm.n, zcm.lex_perm())
        return ycm, SparseTensor(_cmap=zcm, _feats=zhat, _channels=eb.channels), status

    def hyper_synthesis_step(self, z_hat):
        return self.decompression_model.entropy_model.h_s(z_hat)

    def gaussian_model_step_batched(self, y_strings, ycm, q, gaussian_params):
        """codec_parallel.py:382-419."""
        em = self.decompression_model.entropy_model
        gc = em.gaussian_conditional
        cy = self.decompression_model.config["c_y"]
        prow = ops.hash_query(ycm.keys, gaussian_params.cmap.table())
        perm = ycm.lex_perm()
        scale = em.scale_for([list(q)], self.device)
        idx = ops.gauss_indexes(gaussian_params._feats, prow, perm, scale, gc.table_d, gc.scale_bound, ycm.n, cy)
        sym, status = gc._decode(y_strings[0], idx.view(-1), 1, ycm.n * cy, self.device)
        yb, _ = ops.gauss_dequant(sym.view(cy, ycm.n), gaussian_params._feats, prow, perm, scale[0].contiguous(),
                                  em.offset_gain_d, gc.scale_bound)
        y_hat = SparseTensor(_cmap=ycm, _feats=yb, _channels=cy)
        if self.keep_debug:
            self.debug.update(y_sym=sym.view(cy, ycm.n), y_idx=idx, y_hat=y_hat, gaussian_params=gaussian_params)
        return y_hat, status

    def hyper_synthesis(self, y_hat, ks):
        """Step 6 (the reference names the synthesis step `hyper_synthesis`, codec_parallel.py:465-472)."""
        return self.decompression_model.g_s(y_hat, k=ks)

    def pack_batches(self, pointcloud, num_frames):
        """codec_parallel.py:474-502: split per frame, NaN -> 0, clip colours to [0,1]."""
        C = pointcloud.C.cpu().numpy()
        F = pointcloud.F.cpu().numpy()
        self.d2h_bytes += C.nbytes + F.nbytes
        off = np.searchsorted(C[:, 0], np.arange(num_frames + 1))
        batch = []
        for i in range(num_frames):
            pts = C[off[i]:off[i + 1], 1:]
            col = np.nan_to_num(F[off[i]:off[i + 1]], nan=0.0)
            col = np.clip(col * 255.0, 0, 255) / 255
            batch.append({"points": pts, "colors": col})
        return batch

    def decompress(self, compressed_data, to_host=True):
        """bytes -> (frames, sideinfo).  to_host=False leaves the reconstruction on the device (returns
        the SparseTensor instead of the per-frame numpy dicts)."""
        with self._lock, torch.cuda.device(self.device):
            t_start = time.time()
            tm = _StageTimer()
            tm.mark("start")
            y_strings, z_strings, y_shape, z_shape, points_streams, ks, q = self.read_bitstream_batched(compressed_data)
            nf = len(points_streams)
            tm.mark("bitstream_reading")
            y_points = self.geometry_decompression_step(points_streams)
            if y_points.shape[0] != y_shape:
                raise ValueError(f"bitstream y shape {y_shape} != decoded coordinates {y_points.shape[0]}")
            (y_points_dev,) = self._stager.to_device([y_points], self.device)
            self.h2d_bytes += y_points.nbytes + len(compressed_data)
            tm.mark("geometry_decompression")
            ycm, z_hat, st_z = self.factorized_model_step_batched(z_strings, z_shape, y_points_dev, nf)
            tm.mark("factorized_model")
            gaussian_params = self.hyper_synthesis_step(z_hat)
            tm.mark("hyper_synthesis")
            y_hat, st_y = self.gaussian_model_step_batched(y_strings, ycm, q, gaussian_params)
            tm.mark("guassian_model")
            x_hat = self.hyper_synthesis(y_hat, ks)
            tm.mark("synthesis_transform")
            final_data = self.pack_batches(x_hat, nf) if to_host else x_hat
            status = int(st_z.item()) | int(st_y.item())
            if status != 0:
                raise ops._lib.PccError(f"entropy decode failed (status {status})")
            st = tm.resolve()
            sideinfo = {"time_measurements": {}, "timestamps": {}}
            for key in ("bitstream_reading", "geometry_decompression", "factorized_model", "hyper_synthesis",
                        "guassian_model", "synthesis_transform"):
                sideinfo["time_measurements"][key] = st.get(key, 0.0)
            sideinfo["timestamps"]["codec_start"] = t_start
            sideinfo["timestamps"]["codec_end"] = time.time()
            return final_data, sideinfo
