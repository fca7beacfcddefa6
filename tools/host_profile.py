#!/usr/bin/env python
"""Where does the host time go?  Stage times (CUDA events) + cProfile of compress/decompress (GPU only)."""
import cProfile
import copy
import io
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)

import numpy as np
import torch

from b200pcc import CompressionPipeline, DecompressionPipeline, synth
from b200pcc.pipeline import load_model


def main():
    dev = torch.device("cuda:0")
    model = load_model(dev)
    frame = synth.frame_8ivfb(seed=1)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    for _ in range(3):
        c, s = enc.compress({"frames": [frame], "timestamps": {}})
        out, ds = dec.decompress(c[1])
    for name, fn in (("e2e", lambda: None),):
        t0 = time.time()
        n = 5
        for _ in range(n):
            c, s = enc.compress({"frames": [frame], "timestamps": {}})
            t1 = time.time()
            out, ds = dec.decompress(c[1])
        torch.cuda.synchronize()
        print(f"wall per step {(time.time() - t0) / n * 1e3:.2f} ms; bytes {len(c[1])}")
        print("enc stages (ms):", {k: (round(v * 1e3, 2) if not isinstance(v, list) else [round(x * 1e3, 2) for x in v])
                                   for k, v in s["enc_time_measurements"].items()})
        print("dec stages (ms):", {k: round(v * 1e3, 2) for k, v in ds["time_measurements"].items()})
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(5):
        c, s = enc.compress({"frames": [frame], "timestamps": {}})
        out, ds = dec.decompress(c[1])
    pr.disable()
    st = io.StringIO()
    pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(45)
    print(st.getvalue()[:9000])


if __name__ == "__main__":
    main()
