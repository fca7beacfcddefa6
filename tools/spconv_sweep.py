#!/usr/bin/env python
"""Sweep the tuning knobs of pcc_spconv_umma on cfg-2-shaped coordinate sets (GPU only).
Times one k=3 convolution per setting with CUDA events on the launching stream, L2 flushed between
repetitions.  Usage: python tools/spconv_sweep.py [--cin 32 --cout 32]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)

import numpy as np
import torch

from b200pcc import ops, synth, _lib
from b200pcc.sparse import CoordMap


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cin", type=int, default=32)
    ap.add_argument("--cout", type=int, default=32)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--one", action="store_true", help="default tuning only, candidates map only (for ncu)")
    ap.add_argument("--n-target", type=int, default=800_000, help="points of the synthetic frame (smaller = the coarser levels' sizes)")
    ap.add_argument("--kvol", type=int, default=27)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    f = synth.frame_8ivfb(seed=1, n_target=args.n_target)
    pts = torch.from_numpy(np.concatenate([np.zeros((len(f["points"]), 1), np.int32), f["points"].astype(np.int32)], 1)).to(dev)
    keys, _ = ops.sort_unique(ops.keys_from_coords(pts))
    fine = CoordMap(keys, 0)
    cand = fine.downsample().children()          # the g_s candidate set at stride 1 (8 children per stride-2 voxel)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rng = np.random.default_rng(0)
    for name, cm in ((("candidates", cand),) if args.one else (("fine", fine), ("candidates", cand))):
        n = cm.n
        nbr = cm.nbr3()
        pairs = float((nbr >= 0).sum().item()) / n
        x = torch.randn(n, args.cin, device=dev).to(torch.bfloat16)
        W = (rng.normal(0, 1, (27, args.cin, args.cout)) / np.sqrt(args.cin * 9)).astype(np.float32)
        wp = ops.pack_weights_umma(W, args.cin).to(dev)
        bias = torch.zeros(args.cout, device=dev)
        out = torch.empty(n, args.cout, dtype=torch.bfloat16, device=dev)
        alg = 2 * n * args.cin + 2 * n * args.cout + 4 * 27 * n
        print(f"{name}: n={n} pairs/row={pairs:.2f} algorithmic MB={alg / 1e6:.1f}")
        for groups, mt, cg in (((0, 4, 0),) if args.one else
                               ((0, 4, 0), (5, 4, 0), (4, 4, 0), (4, 2, 0), (4, 1, 0), (2, 4, 0), (2, 2, 0), (2, 1, 0))):
            if True:
                _lib.check(_lib.lib.pcc_spconv_set_tuning(groups, mt, cg))
                ts = []
                for r in range(args.reps + 2):
                    flush.fill_(r)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    ops.spconv_umma(x, nbr, n, 27, wp, bias, args.cout, None, True, out, None)
                    e1.record()
                    e1.synchronize()
                    ts.append(e0.elapsed_time(e1))
                t = float(np.median(ts[2:]))
                print(f"  groups={groups} T<={mt} cg={cg}: {t * 1e3:8.1f} us  {alg / t / 1e6:8.1f} GB/s "
                      f"{2 * pairs * n * args.cin * args.cout / t / 1e9:7.1f} TF/s(real pairs)")
        # kernel-map timings
        for r in range(0 if args.one else 3):
            flush.fill_(r)
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            bt = ops.blockmap_build(cm.keys, cm.log2_stride)
            e1.record()
            ops.kmap_k3_blocks(cm.keys, cm.log2_stride, bt)
            e2.record()
            e2.synchronize()
        if not args.one:
            print(f"  blockmap_build {e0.elapsed_time(e1) * 1e3:.1f} us, kmap_k3_blocks {e1.elapsed_time(e2) * 1e3:.1f} us")


if __name__ == "__main__":
    main()
