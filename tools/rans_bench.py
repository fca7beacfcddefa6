#!/usr/bin/env python
"""Time the lane-rANS kernels on Gaussian-conditional symbols shaped like cfg-2's latents (GPU only)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)

import numpy as np
import torch

from b200pcc import ops
from b200pcc.model import GaussianConditional


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=940_000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--segs", type=str, default="32768,16384,8192,4096")
    ap.add_argument("--escape", type=float, default=0.006)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    gc = GaussianConditional()
    gc.update(dev)
    rng = np.random.default_rng(0)
    n = args.n
    idx = np.clip(rng.normal(22, 6, n), 0, 63).astype(np.int32)
    sym = np.rint(rng.normal(0, 1, n) * gc.scale_table[idx]).astype(np.int32)
    esc = rng.random(n) < args.escape
    sym[esc] += rng.integers(50, 4000, esc.sum()).astype(np.int32)
    ts, ti = torch.from_numpy(sym).to(dev), torch.from_numpy(idx.astype(np.int16)).to(dev)
    for seg in [int(s) for s in args.segs.split(",")]:
        te, td = [], []
        for r in range(args.reps + 1):
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            buf, nb = ops.rans_encode(ts, ti, 1, gc.tables, seg)
            e1.record()
            nbytes = int(nb.item())
            cont = buf[: (nbytes + 3) // 4 * 4].clone()
            torch.cuda.synchronize()
            e1b = torch.cuda.Event(enable_timing=True)
            e1b.record()
            dec, status = ops.rans_decode(cont, seg, ti, 1, n, gc.tables)
            e2.record()
            e2.synchronize()
            te.append(e0.elapsed_time(e1))
            td.append(e1b.elapsed_time(e2))
        assert int(status.item()) == 0 and torch.equal(dec, ts)
        print(f"seg={seg:6d}: encode {np.median(te[1:]) * 1e3:8.1f} us  decode {np.median(td[1:]) * 1e3:8.1f} us  "
              f"{8 * nbytes / n:.4f} bits/sym ({nbytes} B)")


if __name__ == "__main__":
    main()
