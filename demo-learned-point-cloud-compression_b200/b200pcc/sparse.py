"""SparseTensor-shaped host objects over device-resident Morton-key coordinate maps.

Mirrors the slice of ``ME.SparseTensor`` the reference touches: constructor kwargs
``coordinates, features, tensor_stride, device``; attributes ``.C .F .tensor_stride .device .dtype``;
method ``.features_at_coordinates(float coords)``.  Row order is ascending Morton key (ME leaves row
order implementation defined; the reference canonicalises with shared/utils.py:116-165, mirrored by
``sort_tensor`` / ``sort_points`` below).
"""
from typing import Optional

import torch

from . import ops


class CoordMap:
    """A sorted, duplicate-free coordinate set at one tensor stride plus lazily built companions
    (hash table, k=3 kernel map, links to the coarser map).  Plays the role of ME's coordinate
    manager entry; shared by every tensor living on these coordinates."""

    def __init__(self, keys: torch.Tensor, log2_stride: int):
        self.keys = keys
        self.n = int(keys.shape[0])
        self.log2_stride = int(log2_stride)
        self._table = None
        self._nbr3 = None
        self._coords = None
        self._lex = None
        self.coarse = None          # CoordMap at 2x stride
        self.parent_row = None      # int32 [n]: row in `coarse`
        self._nbr_down = None
        self.frame_off_host = None  # set by build_pyramid
        self.frame_off_dev = None   # int32 [num_frames+1] device row offsets (levels made by build_pyramid)
        self.frame_counts = None    # the same as host row counts
        self.num_frames = None      # frames in the set when the owner knows it (bounds the frame bits of sort keys)
        self.up1 = None             # host-known upper bounds on the number of coordinates one / two levels
        self.up2 = None             # coarser (exact where the coarser maps exist); None = unknown

    @property
    def stride(self) -> int:
        return 1 << self.log2_stride

    @property
    def device(self):
        return self.keys.device

    def table(self):
        if self._table is None:
            self._table = ops.hash_build(self.keys)
        return self._table

    def nbr3(self) -> torch.Tensor:
        """[27, n] kernel map of a k=3 stride-1 convolution on this map (block-table probing)."""
        if self._nbr3 is None:
            bound = self.up2 if self.up2 is not None else (self.up1 if self.up1 is not None else None)
            if self.coarse is not None and self.coarse.coarse is not None:
                bound = self.coarse.coarse.n          # occupied 4x4x4 blocks = coordinates two levels up
            bt = ops.blockmap_build(self.keys, self.log2_stride, bound)
            self._nbr3 = ops.kmap_k3_blocks(self.keys, self.log2_stride, bt)
        return self._nbr3

    def coords(self) -> torch.Tensor:
        if self._coords is None:
            self._coords = ops.coords_from_keys(self.keys)
        return self._coords

    def lex_perm(self) -> torch.Tensor:
        """perm[r] = row holding canonical (b,x,y,z) rank r."""
        if self._lex is None:
            # the set's coordinates are multiples of its stride; a set with per-frame counts knows its frame count
            nf = self.num_frames if self.num_frames is not None else (len(self.frame_counts) if self.frame_counts is not None else None)
            bits = 16 if nf is None else max(0, (int(nf) - 1).bit_length())
            self._lex = ops.lex_order(self.keys, self.log2_stride, bits)
        return self._lex

    def downsample(self) -> "CoordMap":
        """Output coordinates of a stride-2 convolution (c' = floor(c / 2t) * 2t)."""
        if self.coarse is None:
            keys, parent = ops.coords_downsample(self.keys, self.log2_stride + 1)
            self.coarse = CoordMap(keys, self.log2_stride + 1)
            self.parent_row = parent
        return self.coarse

    def build_pyramid(self, levels: int, num_frames: int = 0, overlap=None, extra=None):
        """Precompute the next `levels` coarser maps (and per-frame row counts of every level) with a
        single host sync, instead of one sync per downsample().  Returns the host frame offsets
        [(levels+1), num_frames+1] (row 0 = this map) or None.  extra: a device int32 counter fetched with
        the same wait; its value is left in self.extra_host."""
        have, cm = 0, self
        while cm.coarse is not None and have < levels:
            cm, have = cm.coarse, have + 1
        if have >= levels and (num_frames == 0 or self.frame_off_host is not None):
            return self.frame_off_host
        res = ops.coords_pyramid(self.keys, self.log2_stride, levels, num_frames, overlap, extra)
        keys_l, par_l, counts, offs, offs_dev = res[:5]
        self.extra_host = res[5] if extra is not None else None
        cm = self
        for l in range(levels):
            cm.coarse = CoordMap(keys_l[l], cm.log2_stride + 1)
            cm.parent_row = par_l[l]
            cm = cm.coarse
            if offs_dev is not None:      # per-frame row ranges of this level, on both sides
                cm.frame_off_dev = offs_dev[l + 1]
                cm.frame_counts = [int(c) for c in (offs[l + 1][1:] - offs[l + 1][:-1])]
        self.frame_off_host = offs
        return offs

    def nbr_down(self) -> torch.Tensor:
        """[8, n_coarse] kernel map of the k=2 stride-2 convolution onto ``downsample()``."""
        if self._nbr_down is None:
            coarse = self.downsample()
            self._nbr_down = ops.kmap_down2(self.keys, self.parent_row, self.log2_stride, coarse.n)
        return self._nbr_down

    def children(self) -> "CoordMap":
        """Generative k=2 s=2 transposed-convolution coordinates: child row 8p + octant."""
        assert self.log2_stride >= 1
        keys = ops.coords_children(self.keys, self.log2_stride - 1)
        cm = CoordMap(keys, self.log2_stride - 1)
        cm.up1, cm.up2 = self.n, self.up1
        return cm

    def subset(self, rows: torch.Tensor) -> "CoordMap":
        """Pruned map (rows ascending => still sorted)."""
        cm = CoordMap(ops.gather_keys(self.keys, rows), self.log2_stride)
        cm.up1, cm.up2 = self.up1, self.up2     # still upper bounds after pruning
        return cm

    def release_maps(self):
        self._table = None
        self._nbr3 = None
        self._nbr_down = None


def _log2(stride) -> int:
    if isinstance(stride, (list, tuple)):
        stride = stride[0]
    stride = int(stride)
    l = stride.bit_length() - 1
    if (1 << l) != stride:
        raise ValueError(f"tensor_stride must be a power of two, got {stride}")
    return l


class SparseTensor:
    def __init__(self, features=None, coordinates=None, tensor_stride=1, device=None, *, _cmap: CoordMap = None,
                 _feats: torch.Tensor = None, _channels: Optional[int] = None):
        self._order = None          # sort_tensor(): permutation that .C / .F present the rows in (canonical order)
        if _cmap is not None:
            self.cmap = _cmap
            self._feats = _feats
            self._channels = _channels if _channels is not None else (_feats.shape[1] if _feats is not None else 0)
            return
        if device is None:
            device = coordinates.device if coordinates.is_cuda else torch.device("cuda", torch.cuda.current_device())
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("b200pcc.SparseTensor lives on a CUDA device (no CPU backend)")
        coords = coordinates.to(device)
        if coords.dtype in (torch.float64, torch.float16, torch.bfloat16):
            coords = coords.float()
        if coords.dtype in (torch.int64, torch.int16, torch.uint8):
            coords = coords.int()
        coords = coords.contiguous()
        keys = ops.keys_from_coords(coords)
        ukeys, src = ops.sort_unique(keys, want_src=True)
        self.cmap = CoordMap(ukeys, _log2(tensor_stride))
        feats = features.to(device=device, dtype=torch.float32).contiguous()
        self._channels = feats.shape[1]
        self._feats = ops.gather_rows(feats, src)

    # ---- reference-visible surface -------------------------------------------------------------------
    @property
    def C(self) -> torch.Tensor:
        c = self.cmap.coords()
        return c if self._order is None else ops.gather_rows_i32(c, self._order)

    @property
    def F(self) -> torch.Tensor:
        f = self._feats
        if f.dtype == torch.bfloat16:
            f = ops.rows_bf16_to_f32(f, self._channels)
        elif f.shape[1] != self._channels:
            f = f[:, : self._channels].contiguous()
        return f if self._order is None else ops.gather_rows(f.contiguous(), self._order)

    @property
    def tensor_stride(self):
        s = self.cmap.stride
        return [s, s, s]

    @property
    def device(self):
        return self.cmap.device

    @property
    def dtype(self):
        return torch.float32

    @property
    def shape(self):
        return (self.cmap.n, self._channels)

    def features_at_coordinates(self, query_coordinates: torch.Tensor) -> torch.Tensor:
        """Rows at (on-grid) float query coordinates, zeros where absent (ME interpolation at grid
        nodes; codec_pipeline.py:401, codec_parallel.py:387)."""
        q = query_coordinates.to(self.device)
        if q.dtype != torch.float32 and q.dtype != torch.int32:
            q = q.float()
        rows = ops.hash_query(ops.keys_from_coords(q.contiguous()), self.cmap.table())
        f = self._feats if self._feats.dtype == torch.float32 else self.F
        return ops.gather_rows(f.contiguous(), rows)

    # ---- internal -----------------------------------------------------------------------------------
    def feats_bf16(self, c_pad: int) -> torch.Tensor:
        f = self._feats
        if f.dtype == torch.bfloat16 and f.shape[1] == c_pad:
            return f
        if f.dtype == torch.bfloat16:
            f = ops.rows_bf16_to_f32(f, self._channels)
        return ops.rows_f32_to_bf16(f.contiguous(), c_pad)


def sort_points(points: torch.Tensor) -> torch.Tensor:
    """shared/utils.py:145-165: rows of an [N,4] (b,x,y,z) tensor in canonical order."""
    keys = ops.keys_from_coords(points.contiguous() if points.dtype in (torch.int32, torch.float32)
                                else points.int().contiguous())
    skeys, src = ops.sort_unique(keys, want_src=True)
    cm = CoordMap(skeys, 0)
    perm = cm.lex_perm().long()
    return points[src.long()[perm]]


def sort_tensor(t: SparseTensor) -> SparseTensor:
    """shared/utils.py:116-141: the same tensor with its rows in canonical (b,x,y,z) order.  The reference builds a new
    ME.SparseTensor from the sorted coordinates and features; here the result shares the coordinate map and the feature
    storage (which stay in Morton order for the kernels) and presents `.C` / `.F` through the canonical permutation."""
    out = SparseTensor(_cmap=t.cmap, _feats=t._feats, _channels=t._channels)
    out._order = t.cmap.lex_perm()
    for name in ("num_frames", "frame_counts"):
        if hasattr(t, name):
            setattr(out, name, getattr(t, name))
    return out
