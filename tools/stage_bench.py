"""Host -> device staging micro-benchmark (pcc_stage_h2d): time until every piece is enqueued and until
the data is on the device, for cold (freshly written) and warm sources, piece sizes and thread counts."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200"))
import torch
from b200pcc._lib import lib, check

dev = torch.device("cuda:0")
n = 800_000
pinned = torch.empty(n * 24, dtype=torch.uint8, pin_memory=True)
dbuf = torch.empty(n * 24, dtype=torch.uint8, device=dev)
st = torch.cuda.Stream()
base = np.random.rand(n, 3)


def run(src, piece, threads, convert=0):
    srcs = (C.c_void_p * 1)(src.ctypes.data)
    sizes = (C.c_int64 * 1)(src.nbytes)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    check(lib.pcc_stage_h2d(srcs, sizes, 1, pinned.data_ptr(), dbuf.data_ptr(), piece, threads, convert, st.cuda_stream))
    t1 = time.perf_counter()
    st.synchronize()
    t2 = time.perf_counter()
    return (t1 - t0) * 1e3, (t2 - t0) * 1e3


for convert in (0, 1):
  for piece_mb in (1, 2, 4):
    for threads in (1, 2, 4, 6, 8):
        warm, cold = [], []
        for rep in range(6):
            warm.append(run(base, int(piece_mb * (1 << 20)), threads, convert))
        for rep in range(6):
            src = base.copy()
            cold.append(run(src, int(piece_mb * (1 << 20)), threads, convert))
        w = np.median(np.asarray(warm[1:]), axis=0)
        c = np.median(np.asarray(cold[1:]), axis=0)
        print(f"{'f64->bf16' if convert else 'copy     '} piece {piece_mb:4} MB threads {threads}: warm enqueue {w[0]:.3f} done {w[1]:.3f} ms | fresh source enqueue {c[0]:.3f} done {c[1]:.3f} ms")
# plain pinned DMA of the same bytes, no host copy
torch.cuda.synchronize()
for rep in range(3):
    t0 = time.perf_counter()
    with torch.cuda.stream(st):
        dbuf.copy_(pinned, non_blocking=True)
    st.synchronize()
    t1 = time.perf_counter()
print(f"pinned -> device DMA alone, {n * 24 / 1e6:.1f} MB: {(t1 - t0) * 1e3:.3f} ms")
