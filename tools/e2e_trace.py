#!/usr/bin/env python
"""Host wall-clock trace of the end-to-end path (host buffers in and out): where do the 3 ms beyond the
device-resident step go?  (GPU only.)"""
import functools
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)
import torch
from b200pcc import CompressionPipeline, DecompressionPipeline, synth, ops, pipeline as P, model as M
from b200pcc.pipeline import load_model

dev = torch.device("cuda:0")
model = load_model(dev)
which = next((a.split("=")[1] for a in sys.argv if a.startswith("--frame=")), "8ivfb")
frame = {"8ivfb": lambda: synth.frame_8ivfb(seed=1), "zed": lambda: synth.frame_zed(seed=2),
         "100k": lambda: synth.frame_cpu100k(seed=0)}[which]()
enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
dec = DecompressionPipeline(device=dev, model=model)
for _ in range(4):
    c, _ = enc.compress({"frames": [frame]})
    dec.decompress(c[1])
torch.cuda.synchronize()
log = []


def wrap(obj, name, sync=False):
    f = getattr(obj, name)

    @functools.wraps(f)
    def g(*a, **k):
        t0 = time.perf_counter()
        r = f(*a, **k)
        if sync:
            torch.cuda.synchronize()
        log.append((name, t0, time.perf_counter()))
        return r
    setattr(obj, name, g)


SYNC = "--sync" in sys.argv
for n in ("stage_frames", "unpack_batch", "analysis_step", "hyper_analysis_step", "factorized_model_step_batched",
          "hyper_synthesis_step", "gaussian_model_step_batched", "geometry_compression_step"):
    wrap(enc, n, SYNC)
for n in ("read_bitstream_batched", "stage_bitstream", "geometry_decompression_step", "factorized_model_step_batched",
          "hyper_synthesis_step", "gaussian_model_step_batched", "hyper_synthesis", "pack_batches"):
    wrap(dec, n, SYNC)
wrap(M, "fetch_strings")
wrap(ops, "verify_deferred")
wrap(P.container, "make_bitstream_batched")
t0 = time.perf_counter()
c, _ = enc.compress({"frames": [frame]})
t1 = time.perf_counter()
out, _ = dec.decompress(c[1])
t2 = time.perf_counter()
print(f"compress {1e3 * (t1 - t0):.3f} ms, decompress {1e3 * (t2 - t1):.3f} ms (sync after each stage: {SYNC})")
for name, a, b in log:
    base = t0 if a < t1 else t1
    print(f"{name:34s} +{1e3 * (a - base):7.3f} ms  dur {1e3 * (b - a):7.3f} ms")
