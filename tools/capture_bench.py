"""Capturer frame preparation (pcc_capture_voxelize): time per ZED-sized measure with the
measure resident in HBM (CUDA events, L2 flushed between iterations), end to end from a host numpy measure
to the host frame dict, the per-kernel split (KernelProfiler) and the CPU oracle beside it.  One JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200"))
import torch  # noqa: E402
from b200pcc.capture import FrameVoxelizer  # noqa: E402
from b200pcc._lib import launch_count  # noqa: E402
from oracle import capture_ref as R  # noqa: E402

dev = torch.device("cuda:0")
out = {}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, (h, w, clip, mp) in {"zed_640x480_clip1.4": (480, 640, 1.4, 30000), "zed_640x480_clip3.0": (480, 640, 3.0, 30000),
                               "hd720_clip3.0": (720, 1280, 3.0, None)}.items():
    m = R.synth_measure(h, w, seed=0)
    v = FrameVoxelizer({"voxel_size": 0.005, "depth_clip": clip, "max_points": mp}, device=dev)
    md = torch.from_numpy(m).to(dev).reshape(-1, 4)
    for _ in range(5):
        pts, col, cnt = v.voxelize(md)
    torch.cuda.synchronize()
    rows = int(cnt.item())
    times = []
    l0 = launch_count()
    for _ in range(30):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        v.voxelize(md)
        b.record()
        b.synchronize()
        times.append(a.elapsed_time(b))
    launches = (launch_count() - l0) / 30
    e2e = []
    for _ in range(15):
        t0 = time.perf_counter()
        f = v.capture(m)
        e2e.append((time.perf_counter() - t0) * 1e3)
    t0 = time.perf_counter()
    ref = R.capture_frame(m, 0.005, clip, mp)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(ref["points"], f["points"]) and np.array_equal(ref["colors"], f["colors"])
    ms = float(np.median(times))
    alg = h * w * 16 + rows * 30
    out[name] = {"pixels": h * w, "rows": rows, "resident_ms_p50": ms, "resident_ms_min": float(np.min(times)),
                 "host_to_host_ms_p50": float(np.median(e2e)), "launches": launches, "algorithmic_bytes": alg,
                 "algorithmic_GBps": alg / ms / 1e6, "Mpixels_per_s": h * w / ms / 1e3, "cpu_oracle_ms": cpu_ms}
print(json.dumps(out))
