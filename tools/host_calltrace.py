#!/usr/bin/env python
"""Host call timeline of one resident step: every call into b200pcc (and the C library through ctypes) with its start
and duration, as an indented tree of the calls that took >= --min-us (default 8).  Shows what the host does while the
GPU idles (after the pyramid read-back, between compress() and decompress()).  GPU only.
Usage: python tools/host_calltrace.py [--min-us 8] [--e2e]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from b200pcc import CompressionPipeline, DecompressionPipeline, synth  # noqa: E402
from b200pcc.pipeline import load_model  # noqa: E402

min_us = float(sys.argv[sys.argv.index("--min-us") + 1]) if "--min-us" in sys.argv else 8.0
e2e = "--e2e" in sys.argv
dev = torch.device("cuda:0")
model = load_model(dev)
frame = synth.frame_8ivfb(seed=1)
enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
dec = DecompressionPipeline(device=dev, model=model)
staged = None if e2e else enc.stage_frames([frame])


def step():
    c, _ = enc.compress({"frames": [frame]}, staged=staged) if not e2e else enc.compress({"frames": [frame]})
    return dec.decompress(c[1], to_host=e2e)


for _ in range(5):
    step()
torch.cuda.synchronize()

events = []      # (depth, name, t_start, t_end)
stack = []
PKG = os.path.join("demo-learned-point-cloud-compression_b200", "b200pcc")


def prof(frame_, event, arg):
    if event in ("call", "c_call"):
        if event == "call":
            fn = frame_.f_code.co_filename
            if PKG not in fn:
                stack.append(None)
                return
            name = f"{os.path.basename(fn)}:{frame_.f_code.co_name}"
        else:
            name = getattr(arg, "__qualname__", None) or getattr(arg, "__name__", repr(arg))
            mod = getattr(arg, "__module__", "") or ""
            if not (name.startswith("pcc_") or "synchronize" in name or name in ("cpu", "item", "copy_", "empty", "tolist")
                    or mod.startswith("torch")):
                stack.append(None)
                return
            name = "C:" + name
        stack.append((len([s for s in stack if s is not None]), name, time.perf_counter()))
    elif event in ("return", "c_return", "c_exception"):
        if stack:
            s = stack.pop()
            if s is not None:
                events.append((s[0], s[1], s[2], time.perf_counter()))


t0 = time.perf_counter()
sys.setprofile(prof)
step()
sys.setprofile(None)
torch.cuda.synchronize()
events.sort(key=lambda e: e[2])
print(f"one {'end-to-end' if e2e else 'resident'} step under the call tracer: {(time.perf_counter() - t0) * 1e3:.2f} ms "
      f"(tracing slows the host several times: read the proportions)")
for d, name, a, b in events:
    if (b - a) * 1e6 >= min_us:
        print(f"{(a - t0) * 1e6:9.1f} us {'  ' * d}{name}  {(b - a) * 1e6:.1f} us")
