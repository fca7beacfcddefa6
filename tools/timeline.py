#!/usr/bin/env python
"""Kernel timeline of one resident-frame step (CUPTI via torch.profiler): where is the GPU idle?
Prints per phase the busy time, the idle time and the largest gaps with the kernels around them."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)

import torch
from torch.profiler import profile, ProfilerActivity

from b200pcc import CompressionPipeline, DecompressionPipeline, synth
from b200pcc.pipeline import load_model


def main():
    dev = torch.device("cuda:0")
    model = load_model(dev)
    which = next((a.split("=")[1] for a in sys.argv if a.startswith("--frame=")), "8ivfb")
    frame = {"8ivfb": lambda: synth.frame_8ivfb(seed=1), "zed": lambda: synth.frame_zed(seed=2),
             "100k": lambda: synth.frame_cpu100k(seed=0)}[which]()
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    staged = enc.stage_frames([frame])
    for _ in range(4):
        c, _ = enc.compress({"frames": [frame]}, staged=staged)
        dec.decompress(c[1], to_host=False)
    torch.cuda.synchronize()
    e2e = "--e2e" in sys.argv
    if e2e:
        for _ in range(3):
            c, _ = enc.compress({"frames": [frame]})
            dec.decompress(c[1])
        torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA] + ([] if e2e else [ProfilerActivity.CPU])) as prof:
        c, _ = (enc.compress({"frames": [frame]}) if e2e else enc.compress({"frames": [frame]}, staged=staged))
        torch.cuda.synchronize()
        dec.decompress(c[1], to_host=e2e)
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ks = sorted(((e.time_range.start, e.time_range.end, e.name) for e in evs), key=lambda t: t[0])
    if not ks:
        print("no CUDA events captured")
        return
    t0 = ks[0][0]
    busy = sum(b - a for a, b, _ in ks)
    span = ks[-1][1] - t0
    print(f"{len(ks)} device activities, span {span / 1e3:.3f} ms, busy {busy / 1e3:.3f} ms, idle {(span - busy) / 1e3:.3f} ms")
    # idle = the complement of the UNION of the activities (copies on side streams and programmatic dependent launches
    # overlap): a gap starts at the latest end seen so far, not at the end of the previously started activity
    gaps = []
    end, last = ks[0][1], ks[0][2]
    for a1, b1, n1 in ks[1:]:
        if a1 > end:
            gaps.append((a1 - end, (end - t0) / 1e3, last[:50], n1[:50]))
        if b1 > end:
            end, last = b1, n1
    union_idle = sum(g for g, *_ in gaps)
    print(f"union of activities: busy {(span - union_idle) / 1e3:.3f} ms, idle {union_idle / 1e3:.3f} ms in {len(gaps)} gaps")
    small = sum(g for g, *_ in gaps if 0 < g <= 10)
    print(f"sum of gaps <= 10 us: {small / 1e3:.3f} ms over {sum(1 for g, *_ in gaps if 0 < g <= 10)} gaps")
    if e2e:
        first_dec = next((i for i, k in enumerate(ks) if "octree_decode" in k[2]), len(ks))
        for i, (a, b, n) in enumerate(ks):
            if i < 12 or "Memcpy" in n or "ingest" in n or abs(i - first_dec) < 3 or i >= len(ks) - 4:
                print(f"  +{(a - t0) / 1e3:7.3f} ms  {(b - a):8.1f} us  {n[:80]}")
    for g, at, n0, n1 in sorted(gaps, key=lambda x: -x[0])[:40]:
        print(f"gap {g:8.1f} us at {at:7.3f} ms  after {n0:50s} before {n1}")


if __name__ == "__main__":
    main()
