"""The sequential, per-frame codec variant ("Single"): one y string and one z string PER FRAME per quality in the
per-frame container of SURVEY App. B.2, instead of one string pair per GOP.

Mirrors ``sender/encoder/codec_single.py:47-111, 114-243, 269-337`` (CompressionPipeline of that file) and
``receiver/decoder/codec_single.py:44-136, 224-265`` (selected by ``decoder_type == "Single"``,
receiver/decoder/decoder.py:41-44).  The transforms still run once over the whole batch -- frames are independent,
the batch index is only a coordinate column -- but the entropy models are driven frame by frame the way the
reference does (``for z_feat, shape in zip(z_feats, z_shapes)``, ``for y_feat, gaussian_param in zip(...)``): frame b's
symbols are the rows of frame b in canonical order, coded as their own PCL1 string.  The symbols, and therefore the
reconstruction, are identical to the batched pipeline's (tests/test_gpu_single.py).

Throughput is not the point of this mode (the reference uses it for single small frames); it reuses the batched
pipeline's kernels with one launch group per frame.
"""
import pickle
import time

import numpy as np
import torch

from . import ops, container, geometry
from .model import fetch_strings
from .pipeline import CompressionPipeline, DecompressionPipeline, _StageTimer
from .sparse import CoordMap, SparseTensor


def _frame_perms(cm: CoordMap, offsets):
    """Canonical (b,x,y,z) order restricted to each frame: frame b owns the Morton rows AND the canonical ranks
    [offsets[b], offsets[b+1]) (the batch index is the most significant part of both orders)."""
    perm = cm.lex_perm()
    return [(perm[o0:o1] - o0).contiguous() for o0, o1 in zip(offsets[:-1], offsets[1:])]


class CompressionPipelineSingle(CompressionPipeline):
    def compress(self, data):
        """data: the GOP dict ({"frames": [...]}), or a list of frame dicts like codec_single.py:114-131 takes.
        -> (compressed_data {0: pickle of the input, i+1: bytes for settings[i]}, sideinfo)."""
        with self._lock, torch.cuda.device(self.device), ops.stream_scope():
            t_start = time.time()
            frames = data["frames"] if isinstance(data, dict) else list(data)
            compressed_data = {0: pickle.dumps(data)}                              # codec_single.py:56
            sideinfo = {} if not isinstance(data, dict) else {k: v for k, v in data.items() if k != "frames"}
            sideinfo["frames"] = [{k: v for k, v in f.items() if k not in ("points", "colors")} for f in frames]
            model = self.compression_model
            em = model.entropy_model
            tm = _StageTimer()
            tm.mark("start")
            x, _ = self.unpack_batch({"frames": list(frames)}, None, defer_features=False, speculate=False)
            nf = x.num_frames
            offs = x.cmap.build_pyramid(5, nf)                                    # host row offsets of all six levels
            num_points = x.cmap.n
            y, k = self.analysis_step(x)                                           # step 1
            ycm = y.cmap
            y_off = [int(v) for v in offs[3]]
            tm.mark("analysis")
            geo_pending = self.geometry_compression_step(ycm)                      # step 6 (per-frame coordinate streams)
            tm.mark("geometry_compression")
            z = self.hyper_analysis_step(y)                                        # step 2
            tm.mark("hyper_analysis")
            z_off = [int(v) for v in offs[5]]
            z_perms = _frame_perms(z.cmap, z_off)
            eb = em.entropy_bottleneck
            z_pending, zhat_rows = [], []
            for b in range(nf):                                                    # step 3, frame by frame (:172-180)
                rows = z._feats[z_off[b]:z_off[b + 1]]
                if rows.shape[0] == 0:
                    z_pending.append(None)
                    continue
                sym, zb, _ = ops.factorized_quant(rows.contiguous(), z_perms[b], eb.medians_d)
                z_pending.append(eb._encode_device(sym.view(-1), None, rows.shape[0]))
                zhat_rows.append(zb)
            zhat = torch.cat(zhat_rows, dim=0) if zhat_rows else torch.empty((0, eb.channels), dtype=torch.bfloat16, device=self.device)
            z_hat = SparseTensor(_cmap=z.cmap, _feats=zhat, _channels=z._channels)
            tm.mark("factorized_model")
            gaussian_params = self.hyper_synthesis_step(z_hat)                     # step 4
            tm.mark("hyper_synthesis")
            gc = em.gaussian_conditional
            prow = ops.hash_query(ycm.keys, gaussian_params.cmap.table())          # features_at_coordinates (:213)
            y_perms = _frame_perms(ycm, y_off)
            scale = em.scale_for(self.settings, self.device)
            nq = len(self.settings)
            y_pending = [[None] * nf for _ in range(nq)]
            dbg_sym = []
            for b in range(nf):                                                    # step 5, frame by frame (:218-236)
                o0, o1 = y_off[b], y_off[b + 1]
                if o1 == o0:
                    continue
                sym, idx = ops.gauss_encode(y._feats[o0:o1].contiguous(), gaussian_params._feats, prow[o0:o1].contiguous(),
                                            y_perms[b], scale, gc.table_d, gc.scale_bound)
                dbg_sym.append(sym)
                for qi in range(nq):
                    y_pending[qi][b] = gc._encode_device(sym[qi].reshape(-1), idx[qi].reshape(-1), 1)
            if self.keep_debug:
                self.debug.update(y=y, z=z, z_hat=z_hat, gaussian_params=gaussian_params, y_sym_frames=dbg_sym, scale=scale)
            tm.mark("gaussian_model")
            flat = [p for p in z_pending if p is not None] + [p for row in y_pending for p in row if p is not None] + geo_pending
            strings = [bytes(s) for s in fetch_strings(flat)]
            it = iter(strings)
            z_strings = [next(it) if p is not None else b"" for p in z_pending]
            y_strings = [[next(it) if p is not None else b"" for p in row] for row in y_pending]
            points_streams = [next(it) for _ in geo_pending]
            self.d2h_bytes += sum(len(s) for s in strings)
            y_shapes = [y_off[b + 1] - y_off[b] for b in range(nf)]
            z_shapes = [z_off[b + 1] - z_off[b] for b in range(nf)]
            t7 = []
            for qi, q in enumerate(self.settings):                                 # step 7 (:269-337)
                t0 = time.time()
                compressed_data[qi + 1] = container.make_bitstream(y_strings[qi], z_strings, y_shapes, z_shapes,
                                                                   points_streams, k, q)
                t7.append(time.time() - t0)
            st = tm.resolve()
            sideinfo["enc_time_measurements"] = {key: st.get(key, 0.0) for key in
                                                 ("analysis", "hyper_analysis", "factorized_model", "hyper_synthesis",
                                                  "geometry_compression", "gaussian_model")}
            sideinfo["enc_time_measurements"]["bitstream_writing"] = t7
            bw = [8 * 6 * num_points] + [len(compressed_data[i + 1]) * 8 for i in range(nq)]
            sideinfo["gop_info"] = {"num_points": num_points, "bandwidth": bw, "bpp": [v / max(num_points, 1) for v in bw]}
            sideinfo.setdefault("timestamps", {})
            sideinfo["timestamps"]["codec_start"] = t_start
            sideinfo["timestamps"]["codec_end"] = time.time()
            return compressed_data, sideinfo


class DecompressionPipelineSingle(DecompressionPipeline):
    def read_bitstream(self, compressed_data):
        return container.read_bitstream(compressed_data)

    def decompress(self, compressed_data, to_host=True):
        """bytes in the per-frame container -> (frames, sideinfo) (receiver/decoder/codec_single.py:44-88 returns the
        frames only; decoder.py:58-62 unpacks two values, so the time measurements come along as in the parallel variant)."""
        with self._lock, torch.cuda.device(self.device), ops.stream_scope():
            t_start = time.time()
            ops.clear_deferred()
            tm = _StageTimer()
            tm.mark("start")
            y_strings, z_strings, y_shapes, z_shapes, points_streams, ks, q = self.read_bitstream(compressed_data)
            nf = len(points_streams)
            tm.mark("bitstream_reading")
            model = self.decompression_model
            em = model.entropy_model
            eb, gc = em.entropy_bottleneck, em.gaussian_conditional
            cy = model.config["c_y"]
            blob, offs, meta32, frame_counts = geometry.pack_streams(points_streams)
            if list(frame_counts) != [int(v) for v in y_shapes]:
                raise ValueError(f"bitstream y shapes {list(y_shapes)} != coordinate streams {list(frame_counts)}")
            arrays = [blob, offs, meta32] + [np.frombuffer(s, dtype=np.uint8) for s in list(y_strings) + list(z_strings)]
            dev = self._stager.to_device(arrays, self.device, pad4=True)
            self.h2d_bytes += sum(a.nbytes for a in arrays)
            y_devs, z_devs = dev[3:3 + nf], dev[3 + nf:3 + 2 * nf]
            ycm, st_geo = self.geometry_decompression_step(dev[0], dev[1], dev[2], frame_counts)        # step 2
            tm.mark("geometry_decompression")
            y_off = np.concatenate([[0], np.cumsum(frame_counts)]).astype(np.int64).tolist()
            z_off = np.concatenate([[0], np.cumsum(z_shapes)]).astype(np.int64).tolist()
            zkeys = ops.coords_pyramid_expect(ycm.keys, 3, 2, z_off[-1])                                 # g_s.down_conv x 2
            zcm = CoordMap(zkeys, 5)
            z_perms = _frame_perms(zcm, z_off)
            status, zhat_rows = [(st_geo, "coordinate stream decode failed")], []
            for b in range(nf):                                                                          # step 3
                n = z_off[b + 1] - z_off[b]
                if n == 0:
                    continue
                sym, st = eb._decode_device(z_devs[b], z_strings[b][:16], None, n, n * eb.channels)
                zb, _ = ops.factorized_dequant(sym.view(eb.channels, n), z_perms[b], eb.medians_d)
                zhat_rows.append(zb)
                status.append((st, f"entropy decode of z failed (frame {b})"))
            zhat = torch.cat(zhat_rows, dim=0) if zhat_rows else torch.empty((0, eb.channels), dtype=torch.bfloat16, device=self.device)
            z_hat = SparseTensor(_cmap=zcm, _feats=zhat, _channels=eb.channels)
            tm.mark("factorized_model")
            gaussian_params = self.hyper_synthesis_step(z_hat)                                           # step 4
            tm.mark("hyper_synthesis")
            prow = ops.hash_query(ycm.keys, gaussian_params.cmap.table())
            y_perms = _frame_perms(ycm, y_off)
            scale = em.scale_for([list(q)], self.device)
            y_rows, dbg_sym = [], []
            for b in range(nf):                                                                          # step 5
                o0, o1 = y_off[b], y_off[b + 1]
                n = o1 - o0
                if n == 0:
                    continue
                pr = prow[o0:o1].contiguous()
                idx = ops.gauss_indexes(gaussian_params._feats, pr, y_perms[b], scale, gc.table_d, gc.scale_bound, n, cy)
                sym, st = gc._decode_device(y_devs[b], y_strings[b][:16], idx.view(-1), 1, n * cy)
                yb, _ = ops.gauss_dequant(sym.view(cy, n), gaussian_params._feats, pr, y_perms[b], scale[0].contiguous(),
                                          em.offset_gain_d, gc.scale_bound)
                y_rows.append(yb)
                dbg_sym.append(sym.view(cy, n))
                status.append((st, f"entropy decode of y failed (frame {b})"))
            yb_all = torch.cat(y_rows, dim=0) if y_rows else torch.empty((0, cy), dtype=torch.bfloat16, device=self.device)
            y_hat = SparseTensor(_cmap=ycm, _feats=yb_all, _channels=cy)
            y_hat.frame_counts = list(frame_counts)
            self._early = None
            if self.keep_debug:
                self.debug.update(y_sym_frames=dbg_sym, y_hat=y_hat, gaussian_params=gaussian_params)
            tm.mark("guassian_model")
            x_hat = self.hyper_synthesis(y_hat, ks)                                                      # step 6
            tm.mark("synthesis_transform")
            packed = self.pack_batches(x_hat, nf) if to_host else None                                   # step 7
            ops.verify_deferred(status, device=self.device)
            final_data = self._finish_batches(*packed, nf) if to_host else x_hat
            st = tm.resolve()
            sideinfo = {"time_measurements": {key: st.get(key, 0.0) for key in
                                              ("bitstream_reading", "geometry_decompression", "factorized_model",
                                               "hyper_synthesis", "guassian_model", "synthesis_transform")},
                        "timestamps": {"codec_start": t_start, "codec_end": time.time()}}
            return final_data, sideinfo
