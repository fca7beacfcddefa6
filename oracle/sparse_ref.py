"""Oracle (test infrastructure): MinkowskiEngine semantics needed by the codec, in numpy.

MinkowskiEngine is unvendored (sender/encoder/Dockerfile:19-25); the conventions restated here are
SURVEY.md Appendix A.2, each isolated in one named function so it can be re-pointed.  Row order of
every coordinate set is the reference's canonical order (shared/utils.py:116-141 `sort_tensor`):
ascending (b, x, y, z).
"""
import ctypes as _C
import os as _os

import numpy as np

# ---- optional C core (sparse_ref.c): same definitions, OpenMP; selected with use_fast(True) ---------------
_FAST = False
_fastlib = None


def _fast():
    global _fastlib
    if _fastlib is None:
        here = _os.path.dirname(_os.path.abspath(__file__))
        so = _os.path.join(here, "_build", "libsparse_ref.so")
        if not _os.path.exists(so):
            import subprocess
            subprocess.check_call(["make", "-C", here], stdout=subprocess.DEVNULL)
        _fastlib = _C.CDLL(so)
        _fastlib.ref_count_pairs.restype = _C.c_int64
    return _fastlib


def use_fast(on=True):
    """Route kernel_map / conv_forward through the C core (full-size frames); returns the previous setting."""
    global _FAST
    prev, _FAST = _FAST, bool(on)
    return prev


def fast_threads():
    import multiprocessing
    return int(_os.environ.get("OMP_NUM_THREADS", 0) or multiprocessing.cpu_count())


def _p(a):
    return a.ctypes.data_as(_C.c_void_p)


# ---- canonical order ----------------------------------------------------------------------------
def lex_key(C):
    """Packed key whose ascending order is the signed lexicographic (b,x,y,z) order."""
    C = np.asarray(C, dtype=np.int64)
    return (C[:, 0] << 48) | ((C[:, 1] + 32768) << 32) | ((C[:, 2] + 32768) << 16) | (C[:, 3] + 32768)


def sort_order_reference(C):
    """argsort used by shared/utils.py:131-133 and :160-162 (weights 1e15, 1e10, 1e5, 1)."""
    w = np.array([1e15, 1e10, 1e5, 1], dtype=np.int64)
    return np.argsort((np.asarray(C, dtype=np.int64) * w).sum(axis=1), kind="stable")


def morton_key(C):
    """b << 48 | interleave(x+32768, y+32768, z+32768), x on the lowest bit of each triple.
    Only used for the top-k tie rule (lower Morton key wins a tie)."""
    C = np.asarray(C, dtype=np.int64)

    def spread(v):
        v = v.astype(np.uint64) & np.uint64(0xFFFF)
        v = (v | (v << np.uint64(16))) & np.uint64(0x001F0000FF0000FF)
        v = (v | (v << np.uint64(8))) & np.uint64(0x100F00F00F00F00F)
        v = (v | (v << np.uint64(4))) & np.uint64(0x10C30C30C30C30C3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
        return v

    m = spread(C[:, 1] + 32768) | (spread(C[:, 2] + 32768) << np.uint64(1)) | (spread(C[:, 3] + 32768) << np.uint64(2))
    return (C[:, 0].astype(np.uint64) << np.uint64(48)) | m


# ---- A.2.1 SparseTensor construction --------------------------------------------------------------
def sparse_tensor(coords, feats=None):
    """ME.SparseTensor(coordinates, features) as used at codec_pipeline.py:262-266: floor float
    coordinates to int32, drop duplicate coordinates (the first row in input order is kept), return
    rows in canonical order."""
    C = np.floor(np.asarray(coords, dtype=np.float64)).astype(np.int64)
    key = lex_key(C)
    order = np.argsort(key, kind="stable")  # stable: first occurrence first inside a run
    ks = key[order]
    first = np.ones(len(ks), dtype=bool)
    first[1:] = ks[1:] != ks[:-1]
    sel = order[first]
    Cn = C[sel].astype(np.int32)
    Fn = None if feats is None else np.asarray(feats)[sel]
    return Cn, Fn


# ---- A.2.2 strided convolution output coordinates --------------------------------------------------
def stride_coords(C, new_stride):
    """c' = floor(c / t') * t' per spatial axis, batch unchanged, de-duplicated, canonical order."""
    C = np.asarray(C, dtype=np.int64).copy()
    C[:, 1:] = np.floor_divide(C[:, 1:], new_stride) * new_stride
    key = lex_key(C)
    _, idx = np.unique(key, return_index=True)
    return C[idx].astype(np.int32)


# ---- A.2.3 kernel region ---------------------------------------------------------------------------
def region_offsets(kernel_size, stride):
    """Hyper-cube offsets; odd k centred, even k anchored at the output coordinate; axis 0 (x)
    enumerated fastest."""
    if kernel_size % 2 == 1:
        r = [(i - kernel_size // 2) * stride for i in range(kernel_size)]
    else:
        r = [i * stride for i in range(kernel_size)]
    return np.array([(dx, dy, dz) for dz in r for dy in r for dx in r], dtype=np.int64)


def lookup_rows(C_src, C_query):
    """Row of each query coordinate in C_src (canonical order) or -1."""
    ks = lex_key(C_src)
    order = np.argsort(ks, kind="stable")
    ks_sorted = ks[order]
    kq = lex_key(C_query)
    pos = np.searchsorted(ks_sorted, kq)
    pos_c = np.minimum(pos, len(ks_sorted) - 1) if len(ks_sorted) else pos
    hit = (pos < len(ks_sorted)) & (ks_sorted[pos_c] == kq) if len(ks_sorted) else np.zeros(len(kq), bool)
    return np.where(hit, order[pos_c] if len(ks_sorted) else 0, -1).astype(np.int64)


# ---- A.2.4 / A.2.5 kernel maps ----------------------------------------------------------------------
def kernel_map(C_in, C_out, offsets, transposed=False):
    """nbr[j, o] = row i of C_in with  C_in[i] = C_out[o] + d_j  (forward conv)  or
    C_in[i] = C_out[o] - d_j  (transposed conv: out[f] += in[c] W[j] where f = c + d_j)."""
    if _FAST and len(C_out):
        ks = lex_key(C_in)
        if len(ks) < 2 or bool(np.all(ks[1:] > ks[:-1])):      # canonical order (every caller's case)
            co = np.ascontiguousarray(C_out, dtype=np.int32)
            offs = np.ascontiguousarray(offsets, dtype=np.int64)
            nbr = np.empty((len(offs), len(co)), dtype=np.int64)
            _fast().ref_kernel_map(_p(ks), _C.c_int64(len(ks)), _p(co), _C.c_int64(len(co)), _p(offs),
                                   _C.c_int32(len(offs)), _C.c_int32(-1 if transposed else 1), _p(nbr))
            return nbr
    C_out = np.asarray(C_out, dtype=np.int64)
    nbr = np.full((len(offsets), len(C_out)), -1, dtype=np.int64)
    for j, d in enumerate(offsets):
        q = C_out.copy()
        if transposed:
            q[:, 1:] -= d
        else:
            q[:, 1:] += d
        ok = np.all((q[:, 1:] >= -32768) & (q[:, 1:] < 32768), axis=1)
        rows = np.full(len(q), -1, dtype=np.int64)
        if ok.any():
            rows[ok] = lookup_rows(C_in, q[ok])
        nbr[j] = rows
    return nbr


def generative_coords(C, out_stride):
    """Generative transposed convolution, k=2, s=2: 8 children c + {0,t'}^3 per input voxel."""
    C = np.asarray(C, dtype=np.int64)
    offs = region_offsets(2, out_stride)
    ch = np.repeat(C[:, None, :], 8, axis=1)
    ch[:, :, 1:] += offs[None, :, :]
    ch = ch.reshape(-1, 4)
    key = lex_key(ch)
    _, idx = np.unique(key, return_index=True)
    return ch[idx].astype(np.int32)


# ---- convolution ------------------------------------------------------------------------------------
def bf16_round(a):
    """Round float32 values to the nearest bfloat16 (ties to even); result stays float32."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    if _FAST and a.size:
        out = np.empty_like(a)
        _fast().ref_bf16_round(_p(a), _C.c_int64(a.size), _p(out))
        return out
    u = a.view(np.uint32)
    r = (u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) & np.uint32(0xFFFF0000)   # wraps never: finite max -> inf
    finite = (u & np.uint32(0x7F800000)) != np.uint32(0x7F800000)
    return np.where(finite, r, u).view(np.float32).reshape(a.shape)


def conv_forward(F_in, nbr, W, bias=None, residual=None, relu=False, bf16=False):
    """out[o] = act(bias + sum_j F_in[nbr[j,o]] @ W[j] (+ residual[o])), fp32 accumulation.
    nbr None = identity rows (kernel volume 1).  bf16=True rounds F_in and W to bfloat16 first
    (what the tensor-core path multiplies); the caller rounds outputs where the product stores bf16."""
    F_in = np.asarray(F_in, dtype=np.float32)
    W = np.asarray(W, dtype=np.float32)
    if bf16:
        F_in = bf16_round(F_in)
        W = bf16_round(W)
    n_out = F_in.shape[0] if nbr is None else nbr.shape[1]
    if _FAST and n_out and W.shape[2] <= 1024:
        F_c, W_c = np.ascontiguousarray(F_in), np.ascontiguousarray(W)
        out = np.empty((n_out, W.shape[2]), dtype=np.float32)
        nb = None if nbr is None else np.ascontiguousarray(nbr, dtype=np.int64)
        b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32).reshape(-1)
        r = None if residual is None else np.ascontiguousarray(residual, dtype=np.float32)
        rc = _fast().ref_conv_forward(_p(F_c), _C.c_int64(F_c.shape[0]), _C.c_int32(F_c.shape[1]),
                                      None if nb is None else _p(nb), _C.c_int64(n_out), _C.c_int32(W_c.shape[0]),
                                      _p(W_c), _C.c_int32(W_c.shape[2]), None if b is None else _p(b),
                                      None if r is None else _p(r), _C.c_int32(int(relu)), _p(out))
        if rc != 0:
            raise ValueError("ref_conv_forward: unsupported shape")
        return out
    out = np.zeros((n_out, W.shape[2]), dtype=np.float32)
    if nbr is None:
        out += F_in @ W[0]
    else:
        for j in range(W.shape[0]):
            rows = nbr[j]
            m = rows >= 0
            if m.any():
                out[m] += F_in[rows[m]] @ W[j]
    if bias is not None:
        out += np.asarray(bias, dtype=np.float32).reshape(1, -1)
    if residual is not None:
        out += np.asarray(residual, dtype=np.float32)
    if relu:
        out = np.maximum(out, 0.0)
    return out


def count_pairs(nbr):
    """P of SURVEY 8d: (offset, output row) pairs that have a neighbour."""
    return int((np.asarray(nbr) >= 0).sum())


# ---- pruning ------------------------------------------------------------------------------------------
def topk_mask(logits, C, ks):
    """Per-frame top-k (torch.topk + MinkowskiPruning inside g_s, codec_parallel.py:469).
    Frame b keeps its ks[b] largest logits; ties -> lower Morton key.  NaN ranks last, -0 == +0."""
    logits = np.asarray(logits, dtype=np.float32).reshape(-1).copy()
    logits[np.isnan(logits)] = -np.inf
    logits = logits + np.float32(0.0)
    C = np.asarray(C)
    mk = morton_key(C)
    keep = np.zeros(len(logits), dtype=bool)
    for b, k in enumerate(ks):
        rows = np.nonzero(C[:, 0] == b)[0]
        if len(rows) == 0 or k <= 0:
            continue
        order = np.lexsort((mk[rows], -logits[rows].astype(np.float64)))
        keep[rows[order[: int(k)]]] = True
    return keep


def features_at_coordinates(C_src, F_src, C_query):
    """SparseTensor.features_at_coordinates for on-grid queries (codec_pipeline.py:401): exact row
    lookup, zeros where the coordinate is absent."""
    rows = lookup_rows(C_src, C_query)
    out = np.zeros((len(rows), F_src.shape[1]), dtype=F_src.dtype)
    m = rows >= 0
    out[m] = F_src[rows[m]]
    return out
