"""ColorModel-shaped object graph on the B200 kernels.

Mirrors the surface of ``unified.model.model.ColorModel`` that the reference pipelines touch
(sender/encoder/codec_pipeline.py:56-72, 273, 287, 305-306, 354, 418-430;
receiver/decoder/codec_parallel.py:302-307, 376, 394-409, 469):

    ColorModel(config) .load_state_dict .to .update() .eval()
      .g_a(x) -> (y, k)         .g_s(y_hat, k=ks)        .g_s.down_conv(t)
      .entropy_model.h_a(y)  .h_s(z_hat)  .scale_nn(q)  .eps  .get_offsets(stdev, scale)
      .entropy_model.entropy_bottleneck.compress(x) / .decompress(strings, size)
      .entropy_model.gaussian_conditional.build_indexes(s) / .compress(x, idx, means=) /
          .decompress(strings, idx) / .lower_bound_scale(s)

The layer graph is the canonical stand-in of config.py.  All sparse convolutions run through
libpcc_b200.so: the persistent tcgen05 gather-GEMM kernel (narrow heads zero-padded to N = 16).
"""
import math
import threading
from typing import List, Optional

import numpy as np
import torch

from . import ops
from . import config as cfgmod
from . import program as prg
from .sparse import CoordMap, SparseTensor
from ._lib import lib, check, PccError

SCALE_MIN, SCALE_MAX, SCALE_LEVELS = 0.11, 256.0, 64
SCALE_BOUND = 0.11
TAIL_MASS = 1e-9

# Run whole transforms as launch programs (program.py) instead of one Python call per layer.  The
# per-layer path stays: it is the readable statement of the graph, serves return_intermediates and
# degenerate (empty) inputs, and the tests require both paths to agree bit for bit.
USE_PROGRAMS = True
# k3 convolutions on the un-pruned candidates of a g_s level (complete octets) through pcc_spconv_octet with the
# parents' kernel map, where the layer shape allows it (C_in = 32); False = the general kernel on the children's map.
USE_OCTET = True


def _pad16(c: int) -> int:
    return (c + 15) // 16 * 16


# ------------------------------------------------------------------------------------------------------
# convolution layers
# ------------------------------------------------------------------------------------------------------
class Conv:
    """One sparse convolution's parameters, prepared for the kernels on ``.to(device)``."""

    def __init__(self, name: str, kvol: int, cin: int, cout: int, simt: bool = False):
        self.name, self.kvol, self.cin, self.cout = name, kvol, cin, cout
        self.simt = simt            # CUDA-core path (only the 1->1 coordinate-only down_conv)
        self.kernel = np.zeros((kvol, cin, cout), dtype=np.float32)
        self.bias = np.zeros(cout, dtype=np.float32)
        self.device = None
        self.narrow = cout <= 8
        self.cin_pad = _pad16(cin)

    def load(self, sd):
        k = np.asarray(sd[self.name + ".kernel"], dtype=np.float32)
        b = np.asarray(sd[self.name + ".bias"], dtype=np.float32).reshape(-1)
        if k.shape != self.kernel.shape or b.shape != self.bias.shape:
            raise ValueError(f"{self.name}: state_dict shape {k.shape}/{b.shape} != {self.kernel.shape}/{self.bias.shape}")
        self.kernel, self.bias = k.copy(), b.copy()

    def to(self, device):
        self.device = device
        self.bias_d = torch.from_numpy(self.bias).to(device)
        if self.simt:
            w = np.zeros((self.kvol, self.cin_pad, self.cout), dtype=np.float32)
            w[:, : self.cin] = self.kernel
            self.w_d = torch.from_numpy(w).to(device).to(torch.bfloat16).contiguous()
        elif self.narrow:
            # narrow heads (occupancy C->1, colour C->3) ride the tensor-core kernel with the output
            # channels zero-padded to one UMMA N=16 tile; only `cout` columns are stored.
            w = np.zeros((self.kvol, self.cin, 16), dtype=np.float32)
            w[:, :, : self.cout] = self.kernel
            b = np.zeros(16, dtype=np.float32)
            b[: self.cout] = self.bias
            self.w_d = ops.pack_weights_umma(w, self.cin_pad).to(device)
            self.bias_d = torch.from_numpy(b).to(device)
        else:
            self.w_d = ops.pack_weights_umma(self.kernel, self.cin_pad).to(device)
        self.w_oct = None
        if not self.simt and self.kvol == 27 and self.cin == 32 and self.cin_pad == 32 and (self.narrow or self.cout == 32):
            self.w_oct = ops.pack_weights_octet(self.kernel, 16 if self.narrow else 32).to(device)
        return self

    def octet(self, x: torch.Tensor, nbr_par: torch.Tensor, n_par: int, relu=False, residual: Optional[torch.Tensor] = None):
        """The same layer on complete octets (x: bf16 [8 n_par, 32]) with the parents' kernel map."""
        dev = x.device
        if self.narrow:
            out = torch.empty((8 * n_par, self.cout), dtype=torch.float32, device=dev)
            ops.spconv_octet(x, nbr_par, n_par, self.w_oct, self.bias_d, 16, None, relu, None, out, cstore=self.cout)
            return out
        out = torch.empty((8 * n_par, self.cout), dtype=torch.bfloat16, device=dev)
        ops.spconv_octet(x, nbr_par, n_par, self.w_oct, self.bias_d, self.cout, residual, relu, out, None)
        return out

    def __call__(self, x: torch.Tensor, nbr: Optional[torch.Tensor], n_out: int, relu=False,
                 residual: Optional[torch.Tensor] = None, out_f32=False):
        """x: bf16 [n_in, cin_pad].  Returns bf16 [n_out, cout] (or fp32 when out_f32)."""
        dev = x.device
        if self.simt:
            out = torch.empty((n_out, self.cout), dtype=torch.float32, device=dev)
            ops.spconv_simt(x, nbr, n_out, self.kvol, self.w_d, self.bias_d, self.cout, residual, relu, None, out)
            return out
        if self.narrow:
            assert residual is None
            out = torch.empty((n_out, self.cout), dtype=torch.float32, device=dev)
            ops.spconv_umma(x, nbr, n_out, self.kvol, self.w_d, self.bias_d, 16, None, relu, None, out,
                            cstore=self.cout)
            return out
        ob = of = None
        if out_f32:
            of = torch.empty((n_out, self.cout), dtype=torch.float32, device=dev)
        else:
            ob = torch.empty((n_out, self.cout), dtype=torch.bfloat16, device=dev)
        ops.spconv_umma(x, nbr, n_out, self.kvol, self.w_d, self.bias_d, self.cout, residual, relu, ob, of)
        return of if out_f32 else ob


class GenUp:
    """k=2, s=2 generative transposed convolution in its dense form: out viewed as
    [n_parent, 8*cout] = in [n_parent, cin] @ concat_j W[j]; child row 8p+j holds columns
    [j*cout, (j+1)*cout).  Column chunks of <= 256 go through the tcgen05 kernel."""

    def __init__(self, name: str, cin: int, cout: int):
        self.name, self.cin, self.cout = name, cin, cout
        self.kernel = np.zeros((8, cin, cout), dtype=np.float32)
        self.bias = np.zeros(cout, dtype=np.float32)
        self.cin_pad = _pad16(cin)

    def load(self, sd):
        k = np.asarray(sd[self.name + ".kernel"], dtype=np.float32)
        b = np.asarray(sd[self.name + ".bias"], dtype=np.float32).reshape(-1)
        if k.shape != self.kernel.shape or b.shape != self.bias.shape:
            raise ValueError(f"{self.name}: bad state_dict shapes")
        self.kernel, self.bias = k.copy(), b.copy()

    def to(self, device):
        wcat = np.concatenate([self.kernel[j] for j in range(8)], axis=1)  # [cin, 8*cout]
        bcat = np.tile(self.bias, 8)
        total = 8 * self.cout
        self.chunks = []
        c0 = 0
        while c0 < total:
            cw = min(256, total - c0)
            wp = ops.pack_weights_umma(wcat[None, :, c0:c0 + cw], self.cin_pad).to(device)
            bd = torch.from_numpy(bcat[c0:c0 + cw].copy()).to(device)
            self.chunks.append((c0, cw, wp, bd))
            c0 += cw
        return self

    def __call__(self, x: torch.Tensor, relu=True) -> torch.Tensor:
        n = x.shape[0]
        out = torch.empty((n, 8 * self.cout), dtype=torch.bfloat16, device=x.device)
        for c0, cw, wp, bd in self.chunks:
            ops.spconv_umma(x, None, n, 1, wp, bd, cw, None, relu, out, None, col0=c0)
        return out.view(8 * n, self.cout)


class _Module:
    def convs(self) -> List:
        out = []
        for v in self.__dict__.values():
            if isinstance(v, (Conv, GenUp)):
                out.append(v)
            elif isinstance(v, (list, tuple)):
                for e in v:
                    if isinstance(e, (Conv, GenUp)):
                        out.append(e)
                    elif isinstance(e, (list, tuple)):
                        out.extend(x for x in e if isinstance(x, (Conv, GenUp)))
        return out


def _resblock(c0: Conv, c1: Conv, cmap: CoordMap, x: torch.Tensor) -> torch.Tensor:
    nbr = cmap.nbr3()
    h = c0(x, nbr, cmap.n, relu=True)
    return c1(h, nbr, cmap.n, relu=True, residual=x)


class AnalysisTransform(_Module):
    def __init__(self, cfg):
        self.cfg = cfg
        self.stem = Conv("g_a.stem", 27, cfg["c_in"], cfg["stem"])
        self.down, self.blocks = [], []
        cprev = cfg["stem"]
        for i, c in enumerate(cfg["widths"]):
            self.down.append(Conv(f"g_a.down{i}", 8, cprev, c))
            self.blocks.append([Conv(f"g_a.block{i}_{b}.conv{j}", 27, c, c)
                                for b in range(cfg["blocks"]) for j in range(2)])
            cprev = c
        self.head = Conv("g_a.head", 1, cprev, cfg["c_y"])

    def __call__(self, x: SparseTensor):
        """g_a(x) -> (y, k): y at tensor stride 8 (fp32 features), k[level][frame] point counts
        (codec_pipeline.py:273, 505-507)."""
        cm = x.cmap
        num_frames = int(x.num_frames) if hasattr(x, "num_frames") else None
        f = x.feats_bf16(self.stem.cin_pad)
        if num_frames is None:
            num_frames = int((cm.keys[-1:] >> 48).item()) + 1 if cm.n else 0
        # all coarser coordinate maps (three for g_a, two more for h_a) and the per-frame counts of every
        # level in one go: coordinates do not depend on features, so this costs one host sync in total
        nlev = len(self.down)
        offs_h = cm.build_pyramid(nlev + 2, num_frames)
        if USE_PROGRAMS and offs_h is not None and cm.n > 0:
            yf, cm = self._program(f, cm, nlev)
            counts = [list(map(int, np.diff(offs_h[i]))) for i in range(nlev)]
            y = SparseTensor(_cmap=cm, _feats=yf, _channels=self.cfg["c_y"])
            y.num_frames = num_frames
            return y, counts[::-1]
        f = self.stem(f, cm.nbr3(), cm.n, relu=True)
        for i in range(nlev):
            coarse = cm.downsample()
            f = self.down[i](f, cm.nbr_down(), coarse.n, relu=True)
            cm.release_maps()
            cm = coarse
            blk = self.blocks[i]
            for b in range(0, len(blk), 2):
                f = _resblock(blk[b], blk[b + 1], cm, f)
        yf = self.head(f, None, cm.n, out_f32=True)
        if offs_h is None:
            counts = [[] for _ in range(nlev)]
        else:
            counts = [list(map(int, np.diff(offs_h[i]))) for i in range(nlev)]   # rows per frame at strides 1, 2, 4
        k = counts[::-1]
        y = SparseTensor(_cmap=cm, _feats=yf, _channels=self.cfg["c_y"])
        y.num_frames = num_frames
        return y, k


def _block_bound(cm):
    if cm.coarse is not None and cm.coarse.coarse is not None:
        return cm.coarse.coarse.n
    return cm.up2 if cm.up2 is not None else cm.up1


def _analysis_program(self, f, cm, nlev):
    """g_a as one launch program: stem, then per level kmap_down2 + down conv + block table + k3 map +
    the residual blocks, then the 1x1 head.  -> (y features fp32 tensor, stride-8 CoordMap)."""
    pr = prg.Program(f.device)
    nbr0 = cm.nbr3()
    stem_out = getattr(cm, "_stem_out", None)
    if stem_out is not None:      # launched ahead of the pyramid read-back (prelaunch_stem)
        cm._stem_out = None
        cur = stem_out.data_ptr()
        pr.keep.append(stem_out)
    else:
        cur = pr.alloc(cm.n * self.stem.cout * 2)
        pr.conv(self.stem, f.data_ptr(), cm.n, nbr0.data_ptr(), nbr0.stride(0), cm.n, out_bf16=cur, relu=True)
        pr.flush()          # the stem runs while the rest of the program is recorded
    for i in range(nlev):
        coarse = cm.coarse
        nd, ld = pr.kmap_down2(cm)
        c = self.down[i].cout
        o = pr.alloc(coarse.n * c * 2)
        pr.conv(self.down[i], cur, cm.n, nd, ld, coarse.n, out_bf16=o, relu=True)
        cm, cur = coarse, o
        nb, ld3 = pr.kmap_k3(cm.keys.data_ptr(), cm.n, cm.log2_stride, _block_bound(cm))
        if i == 0:
            pr.flush()          # the GPU has been waiting for the level sizes: hand it the first kernels at once
        cur = pr.resblocks(self.blocks[i], cur, cm.n, c, nb, ld3)
        if i == 0:
            pr.flush()
    yf = torch.empty((cm.n, self.cfg["c_y"]), dtype=torch.float32, device=f.device)
    pr.conv(self.head, cur, cm.n, 0, 0, cm.n, out_f32=yf.data_ptr())
    pr.keep += [f, nbr0, yf]
    pr.run()
    return yf, cm


def _prelaunch_stem(self, x):
    """Kernel map of the input coordinates + the stem convolution, enqueued before the host waits for the
    pyramid's level sizes (neither depends on them).  g_a picks the result up from the coordinate map."""
    cm = x.cmap
    nbr0 = cm.nbr3()
    if USE_PROGRAMS and x._feats is not None and cm.n > 0:
        cm._stem_out = self.stem(x.feats_bf16(self.stem.cin_pad), nbr0, cm.n, relu=True)


AnalysisTransform._program = _analysis_program
AnalysisTransform.prelaunch_stem = _prelaunch_stem


class SynthesisTransform(_Module):
    def __init__(self, cfg):
        self.cfg = cfg
        w = cfg["widths"]
        self.pre = Conv("g_s.pre", 1, cfg["c_y"], w[-1])
        ups = [w[1], w[0], cfg["stem"]]
        self.up, self.blocks, self.occ = [], [], []
        cprev = w[-1]
        for i, c in enumerate(ups):
            self.up.append(GenUp(f"g_s.up{i}", cprev, c))
            self.blocks.append([Conv(f"g_s.block{i}_{b}.conv{j}", 27, c, c)
                                for b in range(cfg["blocks"]) for j in range(2)])
            self.occ.append(Conv(f"g_s.occ{i}", 27, c, 1))
            cprev = c
        self.color = Conv("g_s.color", 27, cprev, 3)
        self.down_conv_w = Conv("g_s.down_conv", 8, 1, 1, simt=True)

    def down_conv(self, t: SparseTensor) -> SparseTensor:
        """g_s.down_conv(t): MinkowskiConvolution(1,1,k=2,s=2) used by the decoder only for its output
        coordinates (codec_parallel.py:302-303)."""
        cm = t.cmap
        coarse = cm.downsample()
        f = t.feats_bf16(self.down_conv_w.cin_pad)
        out = self.down_conv_w(f, cm.nbr_down(), coarse.n)
        r = SparseTensor(_cmap=coarse, _feats=out, _channels=1)
        if hasattr(t, "num_frames"):
            r.num_frames = t.num_frames
        return r

    def __call__(self, y_hat: SparseTensor, k=None, return_intermediates=False):
        """g_s(y_hat, k=ks): three generative up-samplings, each followed by occupancy logits and the
        per-frame top-k prune; colour head (codec_parallel.py:469).
        When y_hat carries `frame_counts` (rows per frame, host list -- the decoder knows them from the
        coordinate streams) every kept-row count follows from k on the host and no level syncs."""
        cm = y_hat.cmap
        num_frames = len(k[0])
        dev = cm.device
        f = y_hat.feats_bf16(self.pre.cin_pad)
        if USE_PROGRAMS and not return_intermediates and getattr(y_hat, "frame_counts", None) is not None and cm.n > 0:
            out = self._program(f, cm, k, num_frames, list(y_hat.frame_counts), getattr(y_hat, "on_geometry", None))
            if out is not None:
                return out
        f = self.pre(f, None, cm.n, relu=True)
        inter = []
        prev = getattr(y_hat, "frame_counts", None)
        kd_all = torch.tensor([list(map(int, k[i])) for i in range(3)], dtype=torch.int32, pin_memory=True).to(dev, non_blocking=True)   # (pinned: a pageable copy waits behind other callers' transfers)
        for i in range(3):
            f = self.up[i](f, relu=True)
            parent = cm
            cm = cm.children()
            blk = self.blocks[i]
            octet = USE_OCTET and self.occ[i].w_oct is not None and all(c.w_oct is not None for c in blk)
            if octet:
                # candidates = complete octets: the parents' kernel map drives every k3 layer of the level
                nbp = parent.nbr3()
                for b in range(0, len(blk), 2):
                    h = blk[b].octet(f, nbp, parent.n, relu=True)
                    f = blk[b + 1].octet(h, nbp, parent.n, relu=True, residual=f)
                logits = self.occ[i].octet(f, nbp, parent.n)
            else:
                for b in range(0, len(blk), 2):
                    f = _resblock(blk[b], blk[b + 1], cm, f)
                logits = self.occ[i](f, cm.nbr3(), cm.n)
            off = ops.frame_offsets(cm.keys, num_frames)
            expect = None
            if prev is not None:
                kept = [max(0, min(int(k[i][b]), 8 * int(prev[b]))) for b in range(num_frames)]
                expect, prev = sum(kept), kept
            keep = ops.topk_select(logits.view(-1), cm.keys, off, kd_all[i], num_frames, expect=expect)
            if return_intermediates:
                inter.append({"keys": cm.keys, "F": f, "logits": logits, "keep": keep})
            pruned = cm.subset(keep)
            if i == 2 and octet and self.color.w_oct is not None:
                # colour head without a kernel map of the kept set: the pruned candidates are zeroed in place and the head
                # runs over all candidates (complete octets, the parents' map is there already); on the kept rows that
                # is the convolution over the kept set.  The kept rows' colours are then picked out.
                fz = f.clone() if return_intermediates else f
                ops.zero_unkept_rows(fz, keep)
                colors = ops.gather_rows(self.color.octet(fz, nbp, parent.n), keep)
                f = None
            else:
                f = ops.gather_rows(f, keep)
            cm.release_maps()
            cm = pruned
        if f is not None:
            colors = self.color(f, cm.nbr3(), cm.n)
        out = SparseTensor(_cmap=cm, _feats=colors, _channels=3)
        out.num_frames = num_frames
        out.frame_counts = prev
        if return_intermediates:
            return out, inter
        return out


def _synthesis_program(self, f, cm, k, num_frames, prev, on_geometry=None):
    """g_s as one launch program (decoder: every kept-row count follows from k and the per-frame latent
    counts, so all sizes are known on the host).  Per level: generative up-conv, children coordinates,
    block table + k3 map, residual blocks, occupancy head, per-frame top-k, prune; then the colour head.
    Returns None when a level would be empty (the per-layer path handles degenerate inputs)."""
    dev = f.device
    kept_l, n_l = [], []
    n = cm.n
    for i in range(3):
        kept = [max(0, min(int(k[i][b]), 8 * int(prev[b]))) for b in range(num_frames)]
        if sum(kept) == 0:
            return None
        kept_l.append(kept)
        n_l.append(8 * n)
        n, prev = sum(kept), kept
    pr = prg.Program(dev)
    kd_all = torch.tensor([list(map(int, k[i])) for i in range(3)], dtype=torch.int32, pin_memory=True).to(dev, non_blocking=True)   # (pinned: a pageable copy waits behind other callers' transfers)
    counts = torch.empty(3, dtype=torch.int32, device=dev)
    n = cm.n
    cur = pr.alloc(n * self.pre.cout * 2)
    pr.conv(self.pre, f.data_ptr(), n, 0, 0, n, out_bf16=cur, relu=True)
    keys_ptr, l2, up1, up2 = cm.keys.data_ptr(), cm.log2_stride, cm.up1, cm.up2
    out_keys = None
    for i in range(3):
        up = self.up[i]
        c = up.cout
        cur = pr.gen_up(up, cur, n, relu=True)
        nc = 8 * n
        ck = pr.alloc(nc * 8)
        l2 -= 1
        pr.add(prg.OP_COORDS_CHILDREN, keys_ptr, n, l2, ck)
        octet = USE_OCTET and self.occ[i].w_oct is not None and all(cv.w_oct is not None for cv in self.blocks[i])
        logits = pr.alloc(nc * 4)
        if octet:
            # the level's k3 layers run on complete octets: kernel map of the n parents, not of the 8 n children
            bound_par = up2 if up2 is not None else (up1 if up1 is not None else n)
            nbp, ldp = pr.kmap_k3(keys_ptr, n, l2 + 1, bound_par)
            cur = pr.resblocks_octet(self.blocks[i], cur, n, c, nbp, ldp)
            pr.conv_octet(self.occ[i], cur, n, nbp, ldp, out_f32=logits)
            up1, up2 = n, up1
        else:
            up1, up2 = n, up1
            nb, ld = pr.kmap_k3(ck, nc, l2, up2 if up2 is not None else up1)
            cur = pr.resblocks(self.blocks[i], cur, nc, c, nb, ld)
            pr.conv(self.occ[i], cur, nc, nb, ld, nc, out_f32=logits)
        off = pr.alloc((num_frames + 1) * 4)
        pr.add(prg.OP_FRAME_OFFSETS, ck, nc, num_frames, off)
        m = sum(kept_l[i])
        keep = pr.alloc(nc * 4)
        wsb = int(prg.lib.pcc_topk_workspace_bytes(nc))
        ws = pr.alloc(wsb)
        pr.add(prg.OP_TOPK_SELECT, logits, ck, nc, off, kd_all[i].data_ptr(), num_frames, keep,
               counts.data_ptr() + 4 * i, ws, wsb)
        ops._defer(counts[i], m, "topk_select")
        if i == 2:
            out_keys = torch.empty(m, dtype=torch.int64, device=dev)
            pk = out_keys.data_ptr()
        else:
            pk = pr.alloc(m * 8)
        pr.add(prg.OP_GATHER_KEYS, ck, keep, m, pk)
        color_on_candidates = i == 2 and octet and self.color.w_oct is not None
        if i == 2 and on_geometry is not None:
            pr.flush()                 # the geometry is final here: the caller may start shipping it while the colours are computed
            on_geometry(out_keys)
        if color_on_candidates:
            # (see __call__) pruned candidates zeroed in place, colour head over the complete octets, kept rows picked out
            pr.add(prg.OP_ZERO_UNKEPT_ROWS, keep, m, nc, cur, c)
            col_all = pr.alloc(nc * 3 * 4)
            pr.conv_octet(self.color, cur, n, nbp, ldp, out_f32=col_all)
            colors = torch.empty((m, 3), dtype=torch.float32, device=dev)
            col_job = (col_all, keep, m)
        else:
            pf = pr.alloc(m * c * 2)
            pr.add(prg.OP_GATHER_ROWS_BF16, cur, c, keep, m, pf)
            cur = pf
        keys_ptr, n = pk, m
        if i < 2:
            pr.flush()          # this level runs while the next one is recorded
    if color_on_candidates:
        pr.add(prg.OP_GATHER_ROWS_F32, col_job[0], 3, col_job[1], col_job[2], colors.data_ptr())
    else:
        nb, ld = pr.kmap_k3(keys_ptr, n, l2, up2 if up2 is not None else up1)
        colors = torch.empty((n, 3), dtype=torch.float32, device=dev)
        pr.conv(self.color, cur, n, nb, ld, n, out_f32=colors.data_ptr())
    pr.keep += [f, kd_all, counts, out_keys, colors]
    pr.run()
    ocm = CoordMap(out_keys, l2)
    ocm.up1, ocm.up2 = up1, up2
    out = SparseTensor(_cmap=ocm, _feats=colors, _channels=3)
    out.num_frames = num_frames
    out.frame_counts = kept_l[2]
    return out


SynthesisTransform._program = _synthesis_program


class HyperAnalysis(_Module):
    def __init__(self, cfg):
        self.down0 = Conv("entropy_model.h_a.down0", 8, cfg["c_y"], cfg["c_h"])
        self.down1 = Conv("entropy_model.h_a.down1", 8, cfg["c_h"], cfg["c_z"])
        self.cfg = cfg

    def __call__(self, y: SparseTensor) -> SparseTensor:
        cm = y.cmap
        if (USE_PROGRAMS and y._feats.dtype == torch.float32 and y._feats.is_contiguous() and cm.n > 0
                and cm.coarse is not None and cm.coarse.coarse is not None and cm.coarse.coarse.n > 0):
            c1, c2 = cm.coarse, cm.coarse.coarse
            pr = prg.Program(y._feats.device)
            cp = self.down0.cin_pad
            f = pr.alloc(cm.n * cp * 2)
            pr.add(prg.OP_ROWS_F32_TO_BF16, y._feats.data_ptr(), y._feats.shape[1], 0, cm.n, f, cp)
            nd, ld = pr.kmap_down2(cm)
            h = pr.alloc(c1.n * self.down0.cout * 2)
            pr.conv(self.down0, f, cm.n, nd, ld, c1.n, out_bf16=h, relu=True)
            nd, ld = pr.kmap_down2(c1)
            z = torch.empty((c2.n, self.cfg["c_z"]), dtype=torch.float32, device=y._feats.device)
            pr.conv(self.down1, h, c1.n, nd, ld, c2.n, out_f32=z.data_ptr())
            pr.keep += [y._feats, z]
            pr.run()
            out = SparseTensor(_cmap=c2, _feats=z, _channels=self.cfg["c_z"])
            if hasattr(y, "num_frames"):
                out.num_frames = y.num_frames
            return out
        f = y.feats_bf16(self.down0.cin_pad)
        c1 = cm.downsample()
        f = self.down0(f, cm.nbr_down(), c1.n, relu=True)
        c2 = c1.downsample()
        z = self.down1(f, c1.nbr_down(), c2.n, out_f32=True)
        out = SparseTensor(_cmap=c2, _feats=z, _channels=self.cfg["c_z"])
        if hasattr(y, "num_frames"):
            out.num_frames = y.num_frames
        return out


class HyperSynthesis(_Module):
    def __init__(self, cfg):
        self.up0 = GenUp("entropy_model.h_s.up0", cfg["c_z"], cfg["c_h"])
        self.up1 = GenUp("entropy_model.h_s.up1", cfg["c_h"], cfg["c_h"])
        self.head = Conv("entropy_model.h_s.head", 1, cfg["c_h"], 2 * cfg["c_y"])
        self.cfg = cfg

    def __call__(self, z_hat: SparseTensor) -> SparseTensor:
        cm = z_hat.cmap
        f = z_hat.feats_bf16(self.up0.cin_pad)
        if USE_PROGRAMS and cm.n > 0:
            dev = f.device
            pr = prg.Program(dev)
            n = cm.n
            h = pr.gen_up(self.up0, f.data_ptr(), n, relu=True)
            k1 = pr.alloc(8 * n * 8)
            pr.add(prg.OP_COORDS_CHILDREN, cm.keys.data_ptr(), n, cm.log2_stride - 1, k1)
            h = pr.gen_up(self.up1, h, 8 * n, relu=True)
            k2 = torch.empty(64 * n, dtype=torch.int64, device=dev)
            pr.add(prg.OP_COORDS_CHILDREN, k1, 8 * n, cm.log2_stride - 2, k2.data_ptr())
            p = torch.empty((64 * n, 2 * self.cfg["c_y"]), dtype=torch.float32, device=dev)
            pr.conv(self.head, h, 64 * n, 0, 0, 64 * n, out_f32=p.data_ptr())
            pr.keep += [f, k2, p]
            pr.run()
            ocm = CoordMap(k2, cm.log2_stride - 2)
            ocm.up1, ocm.up2 = 8 * n, n
            return SparseTensor(_cmap=ocm, _feats=p, _channels=2 * self.cfg["c_y"])
        f = self.up0(f, relu=True)
        cm = cm.children()
        f = self.up1(f, relu=True)
        cm = cm.children()
        p = self.head(f, None, cm.n, out_f32=True)
        return SparseTensor(_cmap=cm, _feats=p, _channels=2 * self.cfg["c_y"])


# ------------------------------------------------------------------------------------------------------
# entropy models
# ------------------------------------------------------------------------------------------------------
def _std_round_clip(a):
    return a


class _CodedModel:
    """Shared by both entropy models: device CDF tables + PCL1 string coding."""

    tables: ops.DeviceCdfTables = None
    seg_size = None          # None: ops.auto_seg_size(n) per stream

    def _encode_device(self, sym_flat: torch.Tensor, idx_flat: Optional[torch.Tensor], chan_len: int):
        """-> (device byte buffer, device int32 size): no host sync; see fetch_strings()."""
        return ops.rans_encode(sym_flat, idx_flat, chan_len, self.tables, self.seg_size)

    def _encode_device_batch(self, sym: torch.Tensor, idx: Optional[torch.Tensor], chan_len: int):
        """sym / idx: [Q, n].  -> [(device byte buffer, device int32 size)] * Q from one batched coder call."""
        if sym.shape[0] == 1:
            return [self._encode_device(sym.reshape(-1), None if idx is None else idx.reshape(-1), chan_len)]
        return ops.rans_encode_batch(sym, idx, chan_len, self.tables, self.seg_size)

    def _encode(self, sym_flat: torch.Tensor, idx_flat: Optional[torch.Tensor], chan_len: int) -> bytes:
        return fetch_strings([self._encode_device(sym_flat, idx_flat, chan_len)])[0]

    def _decode_device(self, cont: torch.Tensor, hdr, idx_flat: Optional[torch.Tensor], chan_len: int, n: int):
        """cont: the container already on the device (uint8, 4-byte aligned, padded to 4); hdr: its
        first 16 bytes on the host.  -> (symbols, status) without a host sync."""
        if hdr is None or len(hdr) < 16:
            raise ops._lib.PccError("rANS container too short")
        h = np.frombuffer(bytes(hdr[:16]), dtype=np.uint32)
        if h[0] != 0x314C4350 or int(h[1]) != n:
            raise ops._lib.PccError("rANS container header mismatch")
        return ops.rans_decode(cont, int(h[2]), idx_flat, chan_len, n, self.tables)

    def _decode(self, string: bytes, idx_flat: Optional[torch.Tensor], chan_len: int, n: int, device):
        if len(string) < 16:
            raise ops._lib.PccError("rANS container too short")
        hdr = np.frombuffer(string[:16], dtype=np.uint32)
        if hdr[0] != 0x314C4350 or int(hdr[1]) != n:
            raise ops._lib.PccError("rANS container header mismatch")
        seg_size = int(hdr[2])
        arr = np.frombuffer(string, dtype=np.uint8)
        pad = (-len(arr)) % 4
        if pad:
            arr = np.concatenate([arr, np.zeros(pad, np.uint8)])
        cont = torch.from_numpy(arr.copy()).to(device)
        sym, status = ops.rans_decode(cont, seg_size, idx_flat, chan_len, n, self.tables)
        return sym, status


_FETCH = threading.local()


def fetch_strings(pending, as_bytes: bool = True):
    """[(device buffer, device int32 size)] -> [bytes] with ONE host wait in the steady state: the strings
    are packed back to back on the device (`pcc_pack_strings`: lengths + payloads in one buffer) and the
    host copies as many bytes as the previous call of this thread needed, plus a quarter; only when the
    header then says there is more does a second copy follow.  as_bytes=False returns memoryviews over the
    pinned host buffer, valid until this thread's next call (the container writer joins them at once)."""
    if not pending:
        return []
    import ctypes as C
    n = len(pending)
    dev = pending[0][0].device
    hdr = int(lib.pcc_pack_strings_header_bytes(n))
    cap = hdr + sum((int(buf.numel()) + 15) // 16 * 16 for buf, _ in pending)
    packed = torch.empty(cap, dtype=torch.uint8, device=dev)
    bufs = (C.c_void_p * n)(*[buf.data_ptr() for buf, _ in pending])
    lens = (C.c_void_p * n)(*[nb.data_ptr() for _, nb in pending])
    check(lib.pcc_pack_strings(bufs, lens, n, packed.data_ptr(), cap, ops._stream()))
    guess = min(cap, hdr + getattr(_FETCH, "guess", 1 << 18))
    host = getattr(_FETCH, "host", None)
    if host is None or host.numel() < guess:
        host = _FETCH.host = torch.empty(max(guess * 2, 1 << 20), dtype=torch.uint8, pin_memory=True)
    host[:guess].copy_(packed[:guess], non_blocking=True)
    torch.cuda.current_stream().synchronize()
    h32 = host[:hdr].numpy().view(np.uint32)
    sizes = h32[:n].tolist()
    nb = (n + 63) // 64
    end = int(h32[n + nb - 1])
    fault = int(h32[n + nb])           # device fault word, delivered (and cleared) with the strings
    if fault:
        raise PccError(ops.fault_message(fault))
    if end > cap:
        raise PccError(f"fetch_strings: the strings need {end} bytes, their buffers hold {cap}")
    if end > guess:            # first call, or a GOP much larger than the last one
        if host.numel() < end:
            bigger = torch.empty(end * 2, dtype=torch.uint8, pin_memory=True)
            bigger[:guess].copy_(host[:guess])
            host = _FETCH.host = bigger
        host[guess:end].copy_(packed[guess:end], non_blocking=True)
        torch.cuda.current_stream().synchronize()
    _FETCH.guess = (end - hdr) + (end - hdr) // 4 + 4096
    mv = memoryview(host.numpy())
    out, pos = [], hdr
    for nb in sizes:
        v = mv[pos:pos + nb]
        out.append(v.tobytes() if as_bytes else v)
        pos += (nb + 15) // 16 * 16
    return out


class EntropyBottleneck(_CodedModel):
    """CompressAI EntropyBottleneck surface used by the reference (codec_pipeline.py:305-306,
    codec_parallel.py:307): per-channel factorized prior, symbols = round(z - median_c)."""

    def __init__(self, channels: int, filters=(3, 3, 3, 3)):
        self.channels = channels
        self.filters = tuple(filters)
        self.matrices, self.biases, self.factors = [], [], []
        self.quantiles = None
        self.device = None

    def load(self, sd, prefix="entropy_model.entropy_bottleneck."):
        n = len(self.filters) + 1
        self.matrices = [np.asarray(sd[f"{prefix}matrices.{i}"], np.float32) for i in range(n)]
        self.biases = [np.asarray(sd[f"{prefix}biases.{i}"], np.float32) for i in range(n)]
        self.factors = [np.asarray(sd[f"{prefix}factors.{i}"], np.float32) for i in range(n - 1)]
        self.quantiles = np.asarray(sd[f"{prefix}quantiles"], np.float32)

    def _logits_cumulative(self, x: torch.Tensor) -> torch.Tensor:
        """CompressAI 1.2.4 EntropyBottleneck._logits_cumulative on float32 CPU tensors, op for op
        (softplus(matrix) @ logits, += bias, += tanh(factor) * tanh(logits)); SURVEY App. A.3."""
        logits = x
        for i in range(len(self.matrices)):
            m = torch.from_numpy(self.matrices[i])
            logits = torch.matmul(torch.nn.functional.softplus(m), logits)
            logits = logits + torch.from_numpy(self.biases[i])
            if i < len(self.factors):
                logits = logits + torch.tanh(torch.from_numpy(self.factors[i])) * torch.tanh(logits)
        return logits

    def factorized_pmf_f32(self):
        """The pmf rows EntropyBottleneck.update(force=True) hands to pmf_to_quantized_cdf: float32 torch
        arithmetic in CompressAI's order -> (list of float32 rows incl. the tail entry, minima, medians)."""
        q = torch.from_numpy(self.quantiles)
        medians = q[:, 0, 1]
        minima = torch.clamp(torch.ceil(medians - q[:, 0, 0]).int(), min=0)
        maxima = torch.clamp(torch.ceil(q[:, 0, 2] - medians).int(), min=0)
        pmf_start = medians - minima
        pmf_length = maxima + minima + 1
        max_length = int(pmf_length.max())
        samples = torch.arange(max_length)[None, :] + pmf_start[:, None, None]
        lower = self._logits_cumulative(samples - 0.5)
        upper = self._logits_cumulative(samples + 0.5)
        sign = -torch.sign(lower + upper)
        pmf = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))[:, 0, :]
        tail = torch.sigmoid(lower[:, 0, :1]) + torch.sigmoid(-upper[:, 0, -1:])
        rows = [torch.cat((pmf[c, :int(pmf_length[c])], tail[c])).numpy().astype(np.float32)
                for c in range(self.channels)]
        return rows, minima.numpy().astype(np.int64), pmf_length.numpy().astype(np.int64), medians.numpy()

    def update(self, device, force=True):
        prows, minima, pmf_length, medians = self.factorized_pmf_f32()
        rows = [ops.pmf_to_quantized_cdf(r) for r in prows]
        self._quantized_cdf = rows
        self._cdf_length = (pmf_length + 2).astype(np.int32)
        self._offset = (-minima).astype(np.int32)
        self.medians = np.asarray(medians, np.float32)
        self.device = device
        self.tables = ops.DeviceCdfTables(rows, self._offset, device)
        self.medians_d = torch.from_numpy(self.medians).to(device)
        return True

    # symbols for z [n, C] rows in key order, serialised channel-major in canonical column order
    def quantize_rows(self, z_rows: torch.Tensor, perm: torch.Tensor):
        return ops.factorized_quant(z_rows, perm, self.medians_d)

    def compress_rows(self, z_rows: torch.Tensor, perm: torch.Tensor):
        sym, zb, _ = self.quantize_rows(z_rows, perm)
        n = z_rows.shape[0]
        return [self._encode(sym.view(-1), None, n)], zb

    def compress_rows_device(self, z_rows: torch.Tensor, perm: torch.Tensor):
        """Like compress_rows but the string stays on the device: -> ((buffer, size), z_hat bf16)."""
        sym, zb, _ = self.quantize_rows(z_rows, perm)
        return self._encode_device(sym.view(-1), None, z_rows.shape[0]), zb

    def decompress_rows(self, strings, n: int, perm: torch.Tensor):
        sym, status = self._decode(strings[0], None, n, n * self.channels, self.device)
        zb, _ = ops.factorized_dequant(sym.view(self.channels, n), perm, self.medians_d)
        return zb, status

    # CompressAI-shaped entry points (x: [1, C, N] in canonical column order)
    def compress(self, x: torch.Tensor):
        assert x.dim() == 3 and x.shape[0] == 1
        rows = x[0].t().contiguous().float()
        return self.compress_rows(rows, None)[0]

    def decompress(self, strings, size):
        n = int(size[0]) if isinstance(size, (list, tuple)) else int(size)
        zb, status = self.decompress_rows(strings, n, None)
        if int(status.item()) != 0:
            raise ops._lib.PccError(f"rANS decode failed (status {int(status.item())})")
        return zb.float().t().unsqueeze(0)


class GaussianConditional(_CodedModel):
    """CompressAI GaussianConditional surface used by the reference (codec_pipeline.py:425-430,
    codec_parallel.py:398-403)."""

    def __init__(self):
        self.scale_bound = SCALE_BOUND
        self.scale_table = np.exp(np.linspace(math.log(SCALE_MIN), math.log(SCALE_MAX), SCALE_LEVELS)).astype(
            np.float32)
        self.device = None

    def gaussian_pmf_f32(self):
        """The pmf rows GaussianConditional.update() hands to pmf_to_quantized_cdf, computed as CompressAI 1.2.4
        does: float32 torch tensors throughout (scale table, |arange - centre|, (0.5 - s) / sigma, erfc), only the
        tail-mass quantile is a Python double (scipy.stats.norm.ppf there, ndtri here: the same function).
        -> (list of float32 rows incl. the tail entry, pmf_center int64)."""
        from scipy.special import ndtri
        t = torch.from_numpy(self.scale_table)
        multiplier = -float(ndtri(TAIL_MASS / 2))
        pmf_center = torch.ceil(t * multiplier).int()
        pmf_length = 2 * pmf_center + 1
        max_length = int(pmf_length.max())
        samples = torch.abs(torch.arange(max_length).int() - pmf_center[:, None]).float()
        scale = t.unsqueeze(1).float()

        def cumulative(x):      # _standardized_cumulative: 0.5 * erfc(-(2 ** -0.5) * x)
            return 0.5 * torch.erfc(float(-(2 ** -0.5)) * x)
        upper = cumulative((0.5 - samples) / scale)
        lower = cumulative((-0.5 - samples) / scale)
        pmf = upper - lower
        tail = 2 * lower[:, :1]
        rows = [torch.cat((pmf[i, :int(pmf_length[i])], tail[i])).numpy().astype(np.float32)
                for i in range(t.shape[0])]
        return rows, pmf_center.numpy().astype(np.int64)

    def update(self, device):
        prows, pmf_center = self.gaussian_pmf_f32()
        pmf_length = 2 * pmf_center + 1
        rows = [ops.pmf_to_quantized_cdf(r) for r in prows]
        self._quantized_cdf = rows
        self._cdf_length = (pmf_length + 2).astype(np.int32)
        self._offset = (-pmf_center).astype(np.int32)
        self.device = device
        self.tables = ops.DeviceCdfTables(rows, self._offset, device)
        self.table_d = torch.from_numpy(self.scale_table).to(device)

    def lower_bound_scale(self, s: torch.Tensor) -> torch.Tensor:
        return torch.clamp_min(s, self.scale_bound)

    def build_indexes(self, scales: torch.Tensor) -> torch.Tensor:
        """scales [B, C, N] (already multiplied by the gain) -> int32 indexes."""
        B, Cc, N = scales.shape
        flat = scales.reshape(B * Cc, N).t().contiguous().float()       # rows = columns, "channels" = B*C
        zeros = torch.zeros_like(flat)
        params = torch.cat([flat, zeros], dim=1).contiguous()
        ones = torch.ones(B * Cc, dtype=torch.float32, device=scales.device)
        idx = ops.gauss_indexes(params, None, None, ones, self.table_d, self.scale_bound, N, B * Cc)
        return idx.view(B, Cc, N).int()

    def compress(self, x: torch.Tensor, indexes: torch.Tensor, means: Optional[torch.Tensor] = None):
        """x, indexes, means [B, C, N] -> list of B strings (symbols = round(x - means))."""
        d = x if means is None else x - means
        sym = torch.round(d).clamp_(-2.0 ** 30, 2.0 ** 30).int().contiguous()
        out = []
        for b in range(x.shape[0]):
            out.append(self._encode(sym[b].reshape(-1), indexes[b].reshape(-1).to(torch.int16).contiguous(), 1))
        return out

    def decompress(self, strings, indexes: torch.Tensor):
        outs = []
        for b, s in enumerate(strings):
            idx = indexes[b].reshape(-1).to(torch.int16).contiguous()
            sym, status = self._decode(s, idx, 1, idx.numel(), indexes.device)
            if int(status.item()) != 0:
                raise ops._lib.PccError(f"rANS decode failed (status {int(status.item())})")
            outs.append(sym.view(indexes.shape[1:]).float())
        return torch.stack(outs)


class ScaleNN:
    """Rate-control gain q=(q_geo,q_att) -> per-channel gain (entropy_model.scale_nn,
    codec_pipeline.py:418, codec_parallel.py:394).  Evaluated on the host as sequential fp32
    accumulation: encoder and decoder (different machines in the demo) must agree bit for bit."""

    def __init__(self, hidden, cy):
        self.w1 = np.zeros((2, hidden), np.float32)
        self.b1 = np.zeros(hidden, np.float32)
        self.w2 = np.zeros((hidden, cy), np.float32)
        self.b2 = np.zeros(cy, np.float32)

    def load(self, sd, prefix="entropy_model.scale_nn."):
        for k in ("w1", "b1", "w2", "b2"):
            a = np.asarray(sd[prefix + k], np.float32)
            if a.shape != getattr(self, k).shape:
                raise ValueError(f"scale_nn.{k}: bad shape {a.shape}")
            setattr(self, k, a.copy())

    def numpy(self, q) -> np.ndarray:
        q = np.asarray(q, dtype=np.float32).reshape(-1, 2)
        outs = []
        for row in q:
            h = self.b1.copy()
            for k in range(2):
                h = h + self.w1[k] * row[k]
            h = np.maximum(h, np.float32(0))
            o = self.b2.copy()
            for k in range(h.shape[0]):
                o = o + self.w2[k] * h[k]
            outs.append((np.maximum(o, 0) + np.log1p(np.exp(-np.abs(o)))).astype(np.float32))
        return np.stack(outs)

    def __call__(self, q: torch.Tensor) -> torch.Tensor:
        out = self.numpy(q.detach().cpu().numpy())
        return torch.from_numpy(out).to(q.device)


class EntropyModel:
    def __init__(self, cfg):
        self.cfg = cfg
        self.h_a = HyperAnalysis(cfg)
        self.h_s = HyperSynthesis(cfg)
        self.entropy_bottleneck = EntropyBottleneck(cfg["c_z"], cfg["eb_filters"])
        self.gaussian_conditional = GaussianConditional()
        self.scale_nn = ScaleNN(cfg["scale_nn_hidden"], cfg["c_y"])
        self.eps = float(cfg["eps"])
        self.offset_gain = np.ones(cfg["c_y"], np.float32)
        self.offset_gain_d = None
        self._scale_cache = {}

    def get_offsets(self, stdev: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
        """Centroid shift (toward zero) of a unit bin for a Laplacian of std `stdev`, times the
        per-channel gain; broadcastable to [1, Cy, Ny] (codec_parallel.py:405)."""
        b = stdev * 0.70710678118654752
        ib = 1.0 / b
        far = ib * (1.0 / 12.0) - ib * ib * ib * (1.0 / 720.0)
        near = 0.5 / torch.tanh(0.5 / b) - b
        d = torch.where(b > 4.0, far, near).clamp(0.0, 0.5)
        g = torch.from_numpy(self.offset_gain).to(stdev.device).view(1, -1, 1)
        return d * g

    def scale_for(self, settings, device) -> torch.Tensor:
        """[Q, Cy] gains = scale_nn(q) + eps for a list of [q_geo, q_att]."""
        key = (tuple(tuple(float(v) for v in q) for q in settings), str(device))
        hit = self._scale_cache.get(key)
        if hit is None:            # the gains depend on the weights and q only: evaluate once per setting list
            s = self.scale_nn.numpy(np.asarray(settings, dtype=np.float32)) + np.float32(self.eps)
            hit = torch.from_numpy(s.astype(np.float32)).to(device)
            if len(self._scale_cache) > 256:
                self._scale_cache.clear()
            self._scale_cache[key] = hit
        return hit


class ColorModel:
    def __init__(self, config=None):
        if config is None:
            config = cfgmod.SYNTH_SMALL
        if isinstance(config, str):
            config = cfgmod.CONFIGS[config]
        self.config = dict(cfgmod.SYNTH_SMALL, **config)
        self.g_a = AnalysisTransform(self.config)
        self.g_s = SynthesisTransform(self.config)
        self.entropy_model = EntropyModel(self.config)
        self.device = None
        self._sd = None
        self.load_state_dict(cfgmod.make_state_dict(self.config))

    def _all_convs(self):
        em = self.entropy_model
        return self.g_a.convs() + self.g_s.convs() + em.h_a.convs() + em.h_s.convs()

    def state_dict(self):
        return dict(self._sd)

    def load_state_dict(self, sd, strict=True):
        sd = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in sd.items()}
        for c in self._all_convs():
            c.load(sd)
        em = self.entropy_model
        em.entropy_bottleneck.load(sd)
        em.scale_nn.load(sd)
        em.offset_gain = np.asarray(sd["entropy_model.offset_gain"], np.float32).copy()
        em._scale_cache = {}
        self._sd = sd
        if self.device is not None:
            self.to(self.device)
        return self

    def to(self, device):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("b200pcc.ColorModel runs on CUDA devices only (no CPU fallback)")
        self.device = device
        with torch.cuda.device(device):
            for c in self._all_convs():
                c.to(device)
            self.entropy_model.offset_gain_d = torch.from_numpy(self.entropy_model.offset_gain).to(device)
        return self

    def update(self, force=True):
        if self.device is None:
            raise RuntimeError("call .to(device) before .update()")
        with torch.cuda.device(self.device):
            self.entropy_model.gaussian_conditional.update(self.device)
            self.entropy_model.entropy_bottleneck.update(self.device, force=force)
        return True

    def eval(self):
        return self
