#!/usr/bin/env python
"""Pipeline timeline of pcc_spconv_octet's CTA 0 (pcc_debug_trace): per stage, when the gather group found its slot free,
had issued its copies, saw its rows landed, and when issuer 0 saw the stage full and had issued its MMAs (ns, relative).
Usage (GPU box): python tools/octet_trace.py [n_target]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200"))
from b200pcc import ops, synth  # noqa: E402
from b200pcc._lib import lib, check  # noqa: E402
from b200pcc.sparse import CoordMap  # noqa: E402

dev = torch.device("cuda:0")
n_target = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 800_000
f = synth.frame_8ivfb(seed=1, n_target=n_target)
C = np.concatenate([np.zeros((len(f["points"]), 1), np.int32), f["points"].astype(np.int32) // 2 * 2], axis=1)
pk, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(np.ascontiguousarray(C)).to(dev)))
parent = CoordMap(pk, 1)
n_par = parent.n
x = torch.randn((8 * n_par, 32), device=dev).to(torch.bfloat16)
r = torch.randn((8 * n_par, 32), device=dev).to(torch.bfloat16)
w = ops.pack_weights_octet(np.random.default_rng(0).normal(0, 0.05, (27, 32, 32)).astype(np.float32), 32).to(dev)
b = torch.zeros(32, device=dev)
out = torch.empty_like(x)
nbp = parent.nbr3()
for _ in range(3):
    ops.spconv_octet(x, nbp, n_par, w, b, 32, r, True, out, None)
def timed_traced(exp, reps=7):
    """Median time of the TRACE instantiation with the experiment bits of PCC_OCTET_EXP (see OctetParams::exp)."""
    os.environ["PCC_OCTET_EXP"] = str(exp)
    tb = torch.zeros(int(lib.pcc_debug_trace_words()), dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for rep in range(reps):
        flush.fill_(rep)
        check(lib.pcc_debug_trace(tb.data_ptr()))
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.spconv_octet(x, nbp, n_par, w, b, 32, r, True, out, None)
        b_.record()
        b_.synchronize()
        ts.append(a.elapsed_time(b_) * 1e3)
    os.environ["PCC_OCTET_EXP"] = "0"
    return float(np.median(ts[1:])), float(np.min(ts[1:]))


if "--exp" in sys.argv:
    names = {0: "baseline (trace instantiation)", 1: "no zero fill", 2: "gather waits with suspend hint", 4: "issuer waits with suspend hint",
             6: "gather + issuer waits with hint", 8: "epilogue without staging / stores", 16: "no row copies (MMAs on stale slots)",
             24: "no row copies, no epilogue stores", 17: "no copies, no zero fill", 9: "no zero fill, no epilogue stores"}
    ref = None
    for order in (0, 1):
        os.environ["PCC_OCTET_ORDER"] = str(order)
        for e in (0, 8, 16, 24, 0):
            med, mn = timed_traced(e)
            print(f"order {order} exp {e:3d} {names.get(e, ''):45s} median {med:7.1f} us  min {mn:7.1f} us")
            if e == 0:
                torch.cuda.synchronize()
                if ref is None:
                    ref = out.clone()
                elif order == 1:
                    d = (out.float() - ref.float()).abs()
                    print(f"   order 1 vs order 0 output: max abs diff {d.max().item():.4g}, differing elements {int((d > 0).sum())} of {d.numel()}")
    sys.exit(0)

words = int(lib.pcc_debug_trace_words())
buf = torch.zeros(words, dtype=torch.int64, device=dev)
check(lib.pcc_debug_trace(buf.data_ptr()))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ops.spconv_octet(x, nbp, n_par, w, b, 32, r, True, out, None)
e1.record()
torch.cuda.synchronize()
print(f"n_par {n_par} rows {8 * n_par} tiles {(n_par + 127) // 128} traced launch {e0.elapsed_time(e1) * 1e3:.1f} us")
t = buf.cpu().numpy().reshape(6, 128, 3)
t0 = t[t > 0].min()
print(" k  S slot | free   issued landed | mma:full issued | fill  wait_mma")
for k in range(80):
    g = k % 4
    a = t[g, k] - t0
    m = t[4, k] - t0
    print(f"{k:3d} {k % 20:2d} {g:4d} | {a[0]:6d} {a[1]:6d} {a[2]:6d} | {m[0]:8d} {m[1]:6d} | {a[2] - a[0]:5d} {m[0] - a[2]:6d}")

print("epilogue warp 0 (tile iteration, step): step start, after tmem load; last column = tile done")
for k in range(48):
    e = t[5, k] - t0
    print(f"{k // 16:2d} {k % 16:2d} | {e[0]:7d} {e[1]:7d} {e[2] if t[5, k, 2] else 0:7d}")
