"""Oracle (test infrastructure): ctypes front-end of rans_ref.c (build with `make -C oracle`)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "librans_ref.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(_SO)
        _lib.ref64_encode.restype = C.c_int64
        _lib.pcl1_encode.restype = C.c_int64
        _lib.pcl1_max_bytes.restype = C.c_int64
        _lib.pcl1_max_bytes.argtypes = [C.c_int64, C.c_int32]
    return _lib


class Tables:
    """Flat CDF tables: rows (ragged list of int arrays), symbol offsets."""

    def __init__(self, rows, offsets):
        self.sizes = np.array([len(r) for r in rows], dtype=np.int32)
        self.off = np.zeros(len(rows), dtype=np.int32)
        self.off[1:] = np.cumsum(self.sizes)[:-1]
        self.cdf = np.concatenate([np.asarray(r, dtype=np.int32) for r in rows]).astype(np.int32)
        self.offsets = np.asarray(offsets, dtype=np.int32)
        self.n = len(rows)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def ref64_encode(symbols, indexes, t: Tables) -> bytes:
    """CompressAI RansEncoder.encode_with_indexes restatement (single 64-bit stream)."""
    s = np.ascontiguousarray(symbols, dtype=np.int32).reshape(-1)
    i = np.ascontiguousarray(indexes, dtype=np.int32).reshape(-1)
    cap = 16 * s.size + 64
    out = np.empty(cap, dtype=np.uint8)
    n = lib().ref64_encode(_p(s), _p(i), C.c_int64(s.size), _p(t.cdf), _p(t.off), _p(t.sizes), _p(t.offsets),
                           _p(out), C.c_int64(cap))
    if n < 0:
        raise RuntimeError("ref64_encode failed")
    return out[:n].tobytes()


def ref64_decode(data: bytes, indexes, t: Tables) -> np.ndarray:
    i = np.ascontiguousarray(indexes, dtype=np.int32).reshape(-1)
    buf = np.frombuffer(data, dtype=np.uint8).copy()
    out = np.empty(i.size, dtype=np.int32)
    rc = lib().ref64_decode(_p(buf), C.c_int64(buf.size), _p(i), C.c_int64(i.size), _p(t.cdf), _p(t.off),
                            _p(t.sizes), _p(t.offsets), _p(out))
    if rc != 0:
        raise RuntimeError(f"ref64_decode failed ({rc})")
    return out


def pcl1_encode(symbols, indexes, chan_len, t: Tables, seg_size=1 << 15) -> bytes:
    s = np.ascontiguousarray(symbols, dtype=np.int32).reshape(-1)
    idx = None if indexes is None else np.ascontiguousarray(indexes, dtype=np.uint16).reshape(-1)
    cap = int(lib().pcl1_max_bytes(s.size, seg_size))
    out = np.zeros(cap, dtype=np.uint8)
    n = lib().pcl1_encode(_p(s), None if idx is None else _p(idx), C.c_int64(chan_len), C.c_int64(s.size),
                          C.c_int32(seg_size), _p(t.cdf.astype(np.uint32)), _p(t.off), _p(t.sizes), _p(t.offsets),
                          C.c_int32(t.n), _p(out), C.c_int64(cap))
    if n < 0:
        raise RuntimeError("pcl1_encode failed")
    return out[:n].tobytes()


def pcl1_decode(data: bytes, indexes, chan_len, n, t: Tables) -> np.ndarray:
    idx = None if indexes is None else np.ascontiguousarray(indexes, dtype=np.uint16).reshape(-1)
    buf = np.frombuffer(data, dtype=np.uint8).copy()
    out = np.empty(n, dtype=np.int32)
    rc = lib().pcl1_decode(_p(buf), C.c_int64(buf.size), None if idx is None else _p(idx), C.c_int64(chan_len),
                           C.c_int64(n), _p(t.cdf.astype(np.uint32)), _p(t.off), _p(t.sizes), _p(t.offsets),
                           C.c_int32(t.n), _p(out))
    if rc != 0:
        raise RuntimeError(f"pcl1_decode failed ({rc})")
    return out
