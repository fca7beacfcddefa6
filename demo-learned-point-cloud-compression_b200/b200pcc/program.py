"""Launch programs: a whole transform as one call into libpcc_b200.so (``pcc_run_program``).

The reference dispatches every sparse convolution from Python, one MinkowskiEngine call per layer and
one thread per stage (sender/encoder/codec_pipeline.py:31-54, 273-354; receiver/decoder/
codec_parallel.py:302-469).  Doing the same here costs ~12 us of host time per launch (tensor
allocation + argument marshalling), which made the encoder host bound (2.3 ms to enqueue 1.8 ms of
kernels).  A ``Program`` records the launches of a transform as rows of an int64 table -- intermediate
buffers are raw addresses inside a per-thread arena -- and hands the table to C in one call.  Sizes
change from frame to frame, so programs are rebuilt per call (a few us); nothing is cached or captured.
With ``ops.PROFILER`` set the entries run one by one under CUDA events (bench.py's roofline leg).
"""
import array
import ctypes as C
import threading

import numpy as np
import torch

from . import ops
from ._lib import lib, check, PccError

OP_SPCONV_UMMA = 1
OP_KMAP_DOWN2 = 2
OP_BLOCKMAP_BUILD = 3
OP_KMAP_K3_BLOCKS = 4
OP_ROWS_F32_TO_BF16 = 5
OP_COORDS_CHILDREN = 6
OP_GATHER_ROWS_BF16 = 7
OP_GATHER_KEYS = 8
OP_FRAME_OFFSETS = 9
OP_TOPK_SELECT = 10
OP_SPCONV_OCTET = 11
OP_ZERO_UNKEPT_ROWS = 12
OP_GATHER_ROWS_F32 = 13

_ROW = 20          # int64 words per entry: op + 19 arguments (pcc_op_t)
_TLS = threading.local()


def pad128(n: int) -> int:
    return max((n + 127) // 128 * 128, 128)


class Arena:
    """Per-thread bump allocator over torch uint8 chunks.  Addresses stay valid until the next
    ``reset()`` of the same thread; reuse is ordered by the stream every program of the thread runs on."""

    def __init__(self, device):
        self.device = device
        self.chunks = []
        self.pos = 0
        self.used = 0

    def reset(self):
        if len(self.chunks) > 1:       # grown during the last program: coalesce for the next one
            total = sum(c.numel() for c in self.chunks)
            self.chunks = [torch.empty(int(total * 1.25), dtype=torch.uint8, device=self.device)]
        self.pos = 0
        self.used = 0

    def alloc(self, nbytes: int) -> int:
        nbytes = (max(int(nbytes), 1) + 255) // 256 * 256
        if not self.chunks or self.pos + nbytes > self.chunks[-1].numel():
            # geometric growth: a cold arena reaches its working size in a few allocations
            total = sum(c.numel() for c in self.chunks)
            self.chunks.append(torch.empty(max(nbytes, 64 << 20, total), dtype=torch.uint8, device=self.device))
            self.pos = 0
        addr = self.chunks[-1].data_ptr() + self.pos      # torch allocations are 512-byte aligned
        self.pos += nbytes
        self.used += nbytes
        return addr


def arena(device) -> Arena:
    d = getattr(_TLS, "arenas", None)
    if d is None:
        d = _TLS.arenas = {}
    a = d.get(str(device))
    if a is None:
        a = d[str(device)] = Arena(device)
    return a


_PAD = [(0,) * k for k in range(_ROW + 1)]


class Program:
    def __init__(self, device):
        self.device = device
        self.words = array.array("q")     # entries back to back, _ROW int64 words each
        self.n = 0
        self.arena = arena(device)
        self.arena.reset()
        self.keep = []          # tensors that must outlive the launches (outputs, parameters)

    def alloc(self, nbytes: int) -> int:
        return self.arena.alloc(nbytes)

    def add(self, op: int, *args) -> None:
        w = self.words
        w.append(op)
        w.extend(args)
        w.extend(_PAD[_ROW - 1 - len(args)])
        self.n += 1

    # ---- typed helpers -------------------------------------------------------------------------------
    def conv(self, conv, in_ptr: int, n_in: int, nbr_ptr: int, nbr_ld: int, n_out: int, out_bf16: int = 0,
             out_f32: int = 0, ldo: int = 0, relu: bool = False, residual: int = 0, ldr: int = 0) -> None:
        """One model.Conv through pcc_spconv_umma (narrow heads: N = 16 tile, `cout` columns stored)."""
        cout = 16 if conv.narrow else conv.cout
        cstore = conv.cout if conv.narrow else 0
        self.add(OP_SPCONV_UMMA, in_ptr, n_in, conv.cin_pad, nbr_ptr, nbr_ld, n_out, conv.kvol, conv.w_d.data_ptr(),
                 conv.bias_d.data_ptr(), cout, residual, ldr, int(relu), out_bf16, out_f32, ldo or conv.cout, cstore)

    def conv_octet(self, conv, in_ptr: int, n_par: int, nbr_ptr: int, nbr_ld: int, out_bf16: int = 0, out_f32: int = 0,
                   relu: bool = False, residual: int = 0, ldr: int = 0) -> None:
        """One k3 model.Conv on complete octets through pcc_spconv_octet (parents' kernel map)."""
        cout = 16 if conv.narrow else conv.cout
        cstore = conv.cout if conv.narrow else 0
        self.add(OP_SPCONV_OCTET, in_ptr, n_par, conv.cin_pad, nbr_ptr, nbr_ld, conv.w_oct.data_ptr(), conv.bias_d.data_ptr(),
                 cout, residual, ldr, int(relu), out_bf16, out_f32, conv.cout, cstore)

    def resblocks_octet(self, blk, cur: int, n_par: int, c: int, nb: int, ld: int) -> int:
        n = 8 * n_par
        for b in range(0, len(blk), 2):
            h = self.alloc(n * c * 2)
            self.conv_octet(blk[b], cur, n_par, nb, ld, out_bf16=h, relu=True)
            o = self.alloc(n * c * 2)
            self.conv_octet(blk[b + 1], h, n_par, nb, ld, out_bf16=o, relu=True, residual=cur, ldr=c)
            cur = o
        return cur

    def gen_up(self, up, in_ptr: int, n: int, relu: bool = True) -> int:
        """model.GenUp (dense form of the k2/s2 generative transpose) -> address of [8n, cout] bf16."""
        out = self.alloc(n * 8 * up.cout * 2)
        for c0, cw, wp, bd in up.chunks:
            self.add(OP_SPCONV_UMMA, in_ptr, n, up.cin_pad, 0, 0, n, 1, wp.data_ptr(), bd.data_ptr(), cw, 0, 0, int(relu),
                     out + 2 * c0, 0, 8 * up.cout, 0)
        return out

    def kmap_down2(self, cm) -> tuple:
        """Kernel map of the k2/s2 convolution from CoordMap cm onto cm.coarse -> (address, pitch)."""
        ld = pad128(cm.coarse.n)
        nd = self.alloc(8 * ld * 4)
        self.add(OP_KMAP_DOWN2, cm.keys.data_ptr(), cm.parent_row.data_ptr(), cm.n, cm.log2_stride, nd, cm.coarse.n, ld)
        return nd, ld

    def kmap_k3(self, keys_ptr: int, n: int, log2_stride: int, block_bound) -> tuple:
        """Block table + [27, n] kernel map of a k3 convolution -> (address, pitch)."""
        cap = ops.hash_capacity(n if block_bound is None else max(min(int(block_bound), n), 1))
        tk = self.alloc(cap * 8)
        tv = self.alloc(cap * 16)
        self.add(OP_BLOCKMAP_BUILD, keys_ptr, n, log2_stride, tk, tv, cap)
        ld = pad128(n)
        nb = self.alloc(27 * ld * 4)
        self.add(OP_KMAP_K3_BLOCKS, keys_ptr, n, log2_stride, tk, tv, cap, nb, ld)
        return nb, ld

    def resblocks(self, blk, cur: int, n: int, c: int, nb: int, ld: int) -> int:
        for b in range(0, len(blk), 2):
            h = self.alloc(n * c * 2)
            self.conv(blk[b], cur, n, nb, ld, n, out_bf16=h, relu=True)
            o = self.alloc(n * c * 2)
            self.conv(blk[b + 1], h, n, nb, ld, n, out_bf16=o, relu=True, residual=cur, ldr=c)
            cur = o
        return cur

    # ---- execution -----------------------------------------------------------------------------------
    def _family(self, row):
        op = int(row[0])
        a = row[1:]
        if op == OP_SPCONV_UMMA:
            n_in, cin, nbr, n_out, kvol, cout, res, ob, of, cstore = (int(a[1]), int(a[2]), int(a[3]), int(a[5]),
                                                                     int(a[6]), int(a[9]), int(a[10]), int(a[13]),
                                                                     int(a[14]), int(a[16]))
            cw = cstore if 0 < cstore < cout else cout
            nbytes = 2 * n_in * cin + n_out * (cw * ((2 if ob else 0) + (4 if of else 0)) + cout * (2 if res else 0)) + \
                (4 * kvol * n_out if nbr else 0) + 2 * kvol * cin * cout
            pairs = ops.kmap_pairs(nbr, int(a[4]), n_out, kvol, self.device) if nbr else n_out
            return f"spconv_umma_k{kvol}_c{cin}x{cout}", nbytes, 2 * pairs * cin * cout, 2 * kvol * n_out * cin * cout
        if op == OP_SPCONV_OCTET:
            return ops.octet_family(int(a[2]), int(a[7]), int(a[14]), int(a[1]), bool(a[11]), bool(a[12]), bool(a[8]),
                                    int(a[3]), int(a[4]), self.device)
        if op == OP_KMAP_K3_BLOCKS:
            return "kmap_k3", int(a[1]) * (8 + 27 * 4), 0, 0
        if op == OP_BLOCKMAP_BUILD:
            return "blockmap_build", int(a[1]) * 8 + int(a[5]) * 8 + (int(a[1]) // 8) * 24, 0, 0
        if op == OP_TOPK_SELECT:
            return "topk_select", int(a[2]) * 16, 0, 0
        if op == OP_ZERO_UNKEPT_ROWS:
            return "zero_unkept_rows", (int(a[2]) - int(a[1])) * int(a[4]) * 2 + int(a[1]) * 4, 0, 0
        if op == OP_GATHER_ROWS_F32:
            return "gather_rows", int(a[3]) * (int(a[1]) * 8 + 4), 0, 0
        return {OP_KMAP_DOWN2: "kmap_down2", OP_ROWS_F32_TO_BF16: "rows_f32_to_bf16", OP_COORDS_CHILDREN: "coords_children",
                OP_GATHER_ROWS_BF16: "gather_rows", OP_GATHER_KEYS: "gather_keys", OP_FRAME_OFFSETS: "frame_offsets"}[op], 0, 0, 0

    def flush(self) -> None:
        """Run the entries recorded so far and start an empty table over the SAME arena: the GPU gets the
        first layers while the host is still recording the rest, and callers can act between two parts of a
        transform (e.g. start a copy on another stream)."""
        self.run()
        self.words = array.array("q")
        self.n = 0

    def run(self) -> None:
        if self.n == 0:
            return
        failed = C.c_int32(-1)
        st = ops._stream()
        addr, _ = self.words.buffer_info()
        if ops.PROFILER is None:
            rc = lib.pcc_run_program(addr, self.n, st, C.byref(failed))
            if rc:
                op = self.words[max(failed.value, 0) * _ROW]
                raise PccError(f"program entry {failed.value} (op {op}): {ops._lib.last_error()}")
            return
        table = np.frombuffer(self.words, dtype=np.int64).reshape(self.n, _ROW)
        for i in range(self.n):
            fam, nbytes, flops, dense = self._family(table[i])
            with ops._prof(fam, nbytes, flops, dense):
                check(lib.pcc_run_program(addr + i * _ROW * 8, 1, st, C.byref(failed)))
