"""Oracle (test infrastructure): the whole codec path on the CPU — the stand-in for "MinkowskiEngine CPU
backend + CompressAI range coder" that BASELINE.json names as the CPU baseline (neither can be installed
here).  Stage order and data flow follow sender/encoder/codec_pipeline.py:196-437 and
receiver/decoder/codec_parallel.py:141-502; the entropy coder is the CompressAI-style single-stream
rans64 of rans_ref.c.  G-PCC (tmc3) is outside the path: latent coordinates are handed over as arrays.
"""
import time

import numpy as np

from . import sparse_ref as S
from . import entropy_ref as E
from . import rans_ref as R
from .model_ref import RefModel


class RefCodec:
    def __init__(self, cfg, sd, bf16=True):
        self.cfg, self.sd = cfg, sd
        self.model = RefModel(cfg, sd, bf16=bf16)
        self.table = E.scale_table()
        rows, _, offs = E.gaussian_tables(self.table)
        self.g_tables = R.Tables(rows, offs)
        n = len(cfg["eb_filters"]) + 1
        p = "entropy_model.entropy_bottleneck."
        mats = [sd[f"{p}matrices.{i}"] for i in range(n)]
        bias = [sd[f"{p}biases.{i}"] for i in range(n)]
        facs = [sd[f"{p}factors.{i}"] for i in range(n - 1)]
        rz, _, oz, self.medians = E.eb_tables(mats, bias, facs, sd[p + "quantiles"])
        self.z_tables = R.Tables(rz, oz)

    def scale(self, q):
        sd = self.sd
        p = "entropy_model.scale_nn."
        return E.scale_nn(q, sd[p + "w1"], sd[p + "b1"], sd[p + "w2"], sd[p + "b2"]) + np.float32(self.cfg["eps"])

    # ---- encoder (codec_pipeline.py:196-437) -----------------------------------------------------------
    def compress(self, frames, settings):
        t = {}
        t0 = time.time()
        cs, fs = [], []
        for b, f in enumerate(frames):
            p = np.asarray(f["points"]).astype(np.float64)
            cs.append(np.concatenate([np.full((len(p), 1), float(b)), p], axis=1))
            fs.append(np.concatenate([np.ones((len(p), 1)), np.asarray(f["colors"], np.float64)], axis=1))
        C, F = S.sparse_tensor(np.concatenate(cs), np.concatenate(fs).astype(np.float32))
        nf = len(frames)
        Cy, y, k = self.model.g_a(C, F, nf)
        t["analysis"] = time.time() - t0; t0 = time.time()
        Cz, z = self.model.h_a(Cy, y)
        t["hyper_analysis"] = time.time() - t0; t0 = time.time()
        z_sym, z_hat = E.factorized_quantize(z.T, self.medians)
        cz = z.shape[1]
        z_idx = np.repeat(np.arange(cz, dtype=np.int32), z.shape[0])
        z_string = R.ref64_encode(z_sym.reshape(-1), z_idx, self.z_tables)
        t["factorized_model"] = time.time() - t0; t0 = time.time()
        Cp, params = self.model.h_s(Cz, z_hat.T)
        t["hyper_synthesis"] = time.time() - t0; t0 = time.time()
        gp = S.features_at_coordinates(Cp, params, Cy)
        cy = y.shape[1]
        scales_hat, means_hat = gp[:, :cy].T, gp[:, cy:].T
        y_strings, syms = [], []
        for q in settings:
            s = self.scale(q)
            sym, idx = E.gaussian_encode_symbols(y.T, scales_hat, means_hat, s, self.table)
            y_strings.append(R.ref64_encode(sym.reshape(-1), idx.reshape(-1), self.g_tables))
            syms.append(sym)
        t["gaussian_model"] = time.time() - t0
        return {"y_strings": y_strings, "z_string": z_string, "y_coords": Cy, "k": k, "settings": list(settings),
                "num_frames": nf, "times": t, "debug": {"y": y, "z": z, "y_sym": syms, "params": gp, "C": C}}

    # ---- decoder (codec_parallel.py:141-502) ---------------------------------------------------------------
    def decompress(self, blob, qi=0):
        t = {}
        t0 = time.time()
        Cy = blob["y_coords"]
        q = blob["settings"][qi]
        Cz = self.model.down_conv_coords(self.model.down_conv_coords(Cy, 8), 16)
        cz, cy = self.cfg["c_z"], self.cfg["c_y"]
        z_idx = np.repeat(np.arange(cz, dtype=np.int32), len(Cz))
        z_sym = R.ref64_decode(blob["z_string"], z_idx, self.z_tables).reshape(cz, len(Cz))
        z_hat = (z_sym.astype(np.float32) + self.medians.reshape(-1, 1)).T
        t["factorized_model"] = time.time() - t0; t0 = time.time()
        Cp, params = self.model.h_s(Cz, z_hat)
        t["hyper_synthesis"] = time.time() - t0; t0 = time.time()
        gp = S.features_at_coordinates(Cp, params, Cy)
        scales_hat, means_hat = gp[:, :cy].T, gp[:, cy:].T
        s = self.scale(q)
        idx = E.build_indexes(scales_hat * s.reshape(-1, 1), self.table)
        sym = R.ref64_decode(blob["y_strings"][qi], idx.reshape(-1), self.g_tables).reshape(cy, len(Cy))
        y_hat = E.gaussian_dequant(sym, scales_hat, means_hat, s, self.sd["entropy_model.offset_gain"]).T
        t["guassian_model"] = time.time() - t0; t0 = time.time()
        C, colors = self.model.g_s(Cy, y_hat, blob["k"])
        t["synthesis_transform"] = time.time() - t0
        out = []
        for b in range(blob["num_frames"]):
            m = C[:, 0] == b
            col = np.clip(np.nan_to_num(colors[m], nan=0.0) * 255.0, 0, 255) / 255
            out.append({"points": C[m][:, 1:], "colors": col})
        return out, {"times": t, "y_sym": sym, "y_hat": y_hat}
