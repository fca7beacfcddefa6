"""Measurement of the BASELINE.json configurations that bench.py does not time:

  cfg-1  cpu-100k   one 100 100-point frame, end to end on the GPU (the CPU port's number is bench.py's cpu_baseline)
  cfg-3  zed-300k   ZED-shaped frames (negative coordinates) arriving at 30 fps, one frame per call through
                    compress() -> bytes -> decompress() with HOST buffers; per-frame latency p50/p95/p99 and the
                    worst lag behind the arrival clock
  cfg-3c capture    the demo's own operating point (shared/config.yaml: 640x480 ZED measure, depth_clip 1.4 m, 5 mm
                    voxels, max_points 30000) from the camera's XYZRGBA measure in host memory to the client's
                    playout buffer in host memory: FrameVoxelizer.capture_staged -> compress(staged=) ->
                    decompress(to_host=False) -> playout_buffers, frames arriving at the given rate
  cfg-5  ladder     one cfg-2 frame, Gaussian stage batched over Q settings (Q = 1, 3 reference ladder, 25 grid):
                    encode time vs Q, bits per point per level, decode of the best level

Prints one JSON line per configuration (stdout) -- copied to profiles/ by hand.  cfg-2 and cfg-4 are bench.py
(--gpus N).  Wall-clock latencies are taken around the public calls after a synchronize, like a caller sees them.
"""
import argparse
import copy
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200"))

import torch  # noqa: E402


def pct(a, q):
    return float(np.percentile(np.asarray(a, dtype=np.float64), q))


def roundtrip(enc, dec, frame, key=1):
    t0 = time.perf_counter()
    comp, side = enc.compress({"frames": [frame]})
    t1 = time.perf_counter()
    out, _ = dec.decompress(comp[key])
    t2 = time.perf_counter()
    return comp, side, out, (t1 - t0) * 1e3, (t2 - t1) * 1e3


def fresh(frame):
    # compress() pops "frames" and may keep views: give every call its own arrays like the capturer does
    return {"points": frame["points"].copy(), "colors": frame["colors"].copy()}


def cfg1(model, dev, reps):
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frame = synth.frame_cpu100k(seed=0)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    for _ in range(5):
        roundtrip(enc, dec, fresh(frame))
    e, d = [], []
    for _ in range(reps):
        comp, side, out, te, td = roundtrip(enc, dec, fresh(frame))
        e.append(te)
        d.append(td)
    n = len(frame["points"])
    assert out[0]["points"].shape == (n, 3)
    tot = np.asarray(e) + np.asarray(d)
    return {"config": "cfg-1 cpu-100k", "points": n, "frames": reps, "encode_ms_p50": pct(e, 50), "decode_ms_p50": pct(d, 50),
            "e2e_ms_p50": pct(tot, 50), "e2e_mpoints_s": n / pct(tot, 50) * 1e-3, "bytes": len(comp[1]),
            "bpp": 8.0 * len(comp[1]) / n, "buffers": "host numpy in, host numpy out"}


def cfg3(model, dev, nframes, fps):
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    distinct = [synth.frame_zed(seed=2, t=0.1 * i) for i in range(10)]
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    for i in range(6):
        roundtrip(enc, dec, fresh(distinct[i % 10]))
    period = 1.0 / fps
    lat, lag, enc_ms, dec_ms, sizes = [], [], [], [], []
    start = time.perf_counter()
    for i in range(nframes):
        arrival = start + i * period
        now = time.perf_counter()
        if now < arrival:
            time.sleep(arrival - now)
        f = fresh(distinct[i % 10])
        t0 = time.perf_counter()
        comp, side, out, te, td = roundtrip(enc, dec, f)
        t1 = time.perf_counter()
        lat.append((t1 - max(arrival, 0)) * 1e3)     # from arrival to reconstruction in host memory
        lag.append((t0 - arrival) * 1e3)
        enc_ms.append(te)
        dec_ms.append(td)
        sizes.append(len(comp[1]))
    n = float(np.mean([len(f["points"]) for f in distinct]))
    return {"config": "cfg-3 zed-300k", "points_per_frame": n, "frames": nframes, "arrival_fps": fps,
            "latency_ms": {"p50": pct(lat, 50), "p95": pct(lat, 95), "p99": pct(lat, 99), "max": max(lat)},
            "encode_ms_p50": pct(enc_ms, 50), "decode_ms_p50": pct(dec_ms, 50), "start_lag_ms_max": max(lag),
            "keeps_up": bool(max(lag) < period * 1e3), "bpp_mean": 8.0 * float(np.mean(sizes)) / n,
            "mbit_s_at_fps": 8.0 * float(np.mean(sizes)) * fps * 1e-6, "min_coordinate": int(min(f["points"].min() for f in distinct)),
            "buffers": "host numpy in, host numpy out; one frame per call"}


def cfg3c(model, dev, nframes, fps):
    sys.path.insert(0, ROOT)
    from b200pcc import CompressionPipeline, DecompressionPipeline
    from b200pcc.capture import FrameVoxelizer
    from oracle.capture_ref import synth_measure          # input generator only (the checker is not on the timed path)
    measures = [synth_measure(480, 640, seed=100 + i) for i in range(10)]
    vox = FrameVoxelizer({"voxel_size": 0.005, "depth_clip": 1.4, "max_points": 30000}, device=dev)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)

    def once(m):
        t0 = time.perf_counter()
        frame, staged = vox.capture_staged(m)
        t1 = time.perf_counter()
        comp, _ = enc.compress({"frames": [frame]}, staged=staged)
        t2 = time.perf_counter()
        x_hat, _ = dec.decompress(comp[1], to_host=False)
        bufs = dec.playout_buffers(x_hat, num_frames=1)
        t3 = time.perf_counter()
        return len(frame["points"]), len(comp[1]), len(bufs[0]), (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3

    for i in range(6):
        once(measures[i % 10])
    period = 1.0 / fps
    lat, lag, cap_ms, enc_ms, dec_ms, sizes, pts = [], [], [], [], [], [], []
    start = time.perf_counter()
    for i in range(nframes):
        arrival = start + i * period
        now = time.perf_counter()
        if now < arrival:
            time.sleep(arrival - now)
        t0 = time.perf_counter()
        n, nbytes, nplay, tc, te, td = once(measures[i % 10])
        t1 = time.perf_counter()
        assert nplay == 15 * n
        lat.append((t1 - arrival) * 1e3)
        lag.append((t0 - arrival) * 1e3)
        cap_ms.append(tc)
        enc_ms.append(te)
        dec_ms.append(td)
        sizes.append(nbytes)
        pts.append(n)
    n = float(np.mean(pts))
    return {"config": "cfg-3c capture-to-playout (shared/config.yaml operating point)", "measure": "640x480 XYZRGBA float32, host",
            "points_per_frame": n, "frames": nframes, "arrival_fps": fps,
            "latency_ms": {"p50": pct(lat, 50), "p95": pct(lat, 95), "p99": pct(lat, 99), "max": max(lat)},
            "capture_ms_p50": pct(cap_ms, 50), "encode_ms_p50": pct(enc_ms, 50), "decode_playout_ms_p50": pct(dec_ms, 50),
            "start_lag_ms_max": max(lag), "keeps_up": bool(max(lag) < period * 1e3), "bpp_mean": 8.0 * float(np.mean(sizes)) / n,
            "h2d_bytes_per_frame": 480 * 640 * 16, "d2h_bytes_per_frame": 15 * n + float(np.mean(sizes)),
            "buffers": "host measure in, host playout buffer out; the voxelised frame never leaves HBM"}


def cfg5(model, dev, reps):
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frame = synth.frame_8ivfb(seed=1)
    n = len(frame["points"])
    ladders = {"Q1": [[1.0, 1.0]], "Q3_reference_ladder": [[1.0, 0.0], [0.0, 1.0], [1.0, 1.0]],
               "Q25_grid": [[a, b] for a in (0.0, 0.25, 0.5, 0.75, 1.0) for b in (0.0, 0.25, 0.5, 0.75, 1.0)]}
    dec = DecompressionPipeline(device=dev, model=model)
    res = {"config": "cfg-5 ladder", "points": n, "levels": {}}
    for name, settings in ladders.items():
        enc = CompressionPipeline(settings, device=dev, model=model)
        for _ in range(4):
            enc.compress({"frames": [fresh(frame)]})
        ts = []
        for _ in range(reps):
            gop = {"frames": [fresh(frame)]}
            t0 = time.perf_counter()
            comp, side = enc.compress(gop)
            ts.append((time.perf_counter() - t0) * 1e3)
        bpp = [8.0 * len(comp[i + 1]) / n for i in range(len(settings))]
        tds = []
        for i in range(4 + max(reps // 2, 3)):
            t0 = time.perf_counter()
            out, _ = dec.decompress(comp[len(settings)])
            if i >= 4:
                tds.append((time.perf_counter() - t0) * 1e3)
        td = pct(tds, 50)
        assert out[0]["points"].shape == (n, 3)
        res["levels"][name] = {"Q": len(settings), "encode_ms_p50": pct(ts, 50), "encode_ms_per_level": pct(ts, 50) / len(settings),
                               "bpp_min": min(bpp), "bpp_max": max(bpp), "bpp": [round(b, 4) for b in bpp],
                               "decode_last_level_ms_p50": td}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--fps", type=float, default=30.0)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    from b200pcc.pipeline import load_model
    model = load_model(dev)
    todo = {"cfg1": lambda: cfg1(model, dev, args.reps), "cfg3": lambda: cfg3(model, dev, args.frames, args.fps),
            "cfg3c": lambda: cfg3c(model, dev, args.frames, args.fps),
            "cfg5": lambda: cfg5(model, dev, args.reps)}
    for k, fn in todo.items():
        if args.only and k not in args.only.split(","):
            continue
        r = fn()
        r["device"] = torch.cuda.get_device_name(0)
        r["model"] = "synth_small (seeded weights)"
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
