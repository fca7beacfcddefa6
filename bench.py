#!/usr/bin/env python
"""bench.py — codec Mpoints/s (encode + decode) of the learned point-cloud codec hot path.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference --steps K --warmup W     (CPU arm: the oracle port on host cores)

A "step" = compress() + decompress() of one GOP of one synthetic 8iVFB-shaped frame (cfg-2,
~800k coloured points in a 1024^3 grid), quality [1,1].  `value` is timed with the raw
frame already resident in HBM and the reconstruction left on the device; `e2e` goes through the public
CompressionPipeline.compress(gop) / DecompressionPipeline.decompress(bytes) calls with host buffers
(host->device copy of the frame and device->host copy of the reconstruction inside the timed region).
Frames are independent, so with N GPUs every rank codes its own frames (weak scaling, no collective on
the data path); the timed region is bracketed by barrier + synchronize and the max over ranks is used.
"""
import argparse
import copy
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")
for _p in (ROOT, PKG_DIR):
    if _p not in sys.path:
        sys.path.insert(0, _p)

if "reference" in sys.argv:
    # the CPU arm uses every host core: torchrun exports OMP_NUM_THREADS=1, which would pin the BLAS
    # pools of the port to one thread (must be undone before numpy is imported)
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count())

import numpy as np  # noqa: E402

METRIC = "codec Mpoints/s (enc+dec)"
UNIT = "Mpoints/s"
WORKLOAD = "cfg-2 8ivfb-800k: one synthetic 8iVFB-shaped frame (~800k coloured points, 1024^3 grid), Q=1 [1,1]"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.stop_flag = False
        self.max_mhz = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.2)

    def result(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def _crop_sample(frame, n_target):
    """Bounded CPU sample of the same workload: a horizontal slab of the frame with ~n_target points."""
    p = frame["points"]
    order = np.argsort(p[:, 2], kind="stable")
    mid = len(order) // 2
    sel = np.sort(order[max(mid - n_target // 2, 0): mid + n_target // 2])
    return {"points": p[sel], "colors": frame["colors"][sel]}


def _blas_threads():
    """Threads the numpy BLAS pool of this process actually uses (what the CPU port runs on)."""
    try:
        from threadpoolctl import threadpool_info
        n = [int(p.get("num_threads", 1)) for p in threadpool_info() if p.get("user_api") == "blas"]
        return max(n) if n else 1
    except Exception:
        return 1


def cpu_port_run(frame, sample_points, reps=1):
    """Oracle port (numpy + C rANS) timed on the host cores: encode + decode of the sample."""
    from b200pcc import config as cfgmod
    from oracle.codec_ref import RefCodec
    sample = _crop_sample(frame, sample_points)
    codec = RefCodec(cfgmod.SYNTH_SMALL, cfgmod.make_state_dict(), bf16=True)
    n = len(sample["points"])
    times = []
    for _ in range(reps):
        t0 = time.time()
        blob = codec.compress([sample], [[1.0, 1.0]])
        codec.decompress(blob, 0)
        times.append(time.time() - t0)
    return n, times


def run_reference(args):
    """--impl reference: the reference's own CPU path cannot be installed here (MinkowskiEngine,
    CompressAI and the model submodule are unvendored), so this arm times the oracle port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from b200pcc import synth
    frame = synth.frame_8ivfb(seed=1)
    sample_points = args.cpu_sample
    steps, warmup = args.steps, args.warmup
    n, times = cpu_port_run(frame, sample_points, reps=warmup + steps)
    t = times[warmup:]
    total = sum(t)
    value = n * len(t) / total / 1e6
    cores = _blas_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * total / len(t), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "model": "synth_small (seeded weights)",
                   "note": "CPU port of the path (numpy sparse conv + C rans64); reference deps not installable"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{n}-point slab of the cfg-2 frame, encode+decode, {len(t)} reps"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=150_000, help="points in the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--points", type=int, default=800_000)
    ap.add_argument("--breakdown", action="store_true", help="print per-kernel-family table to stderr")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer arm (short runs under ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner under
    # torchrun) are sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth, ops, _lib
    from b200pcc.pipeline import load_model

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    steps, warmup = args.steps, max(args.warmup, 3)
    # per-rank frames (weak scaling): a handful of distinct frames cycled through the steps
    n_frames = 2
    frames = [synth.frame_8ivfb(seed=1 + rank * 100 + i, n_target=args.points) for i in range(n_frames)]
    n_points = [len(f["points"]) for f in frames]
    model = load_model(dev)
    settings = [[1.0, 1.0]]
    enc = CompressionPipeline(settings, device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def gop(i):
        return {"frames": [frames[i % n_frames]], "timestamps": {}}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident arm (value) ----------------
    staged = [enc.stage_frames(g["frames"]) for g in (gop(0), gop(1))]
    torch.cuda.synchronize()
    for i in range(warmup):
        c, _ = enc.compress(gop(i), staged=staged[i % n_frames])
        dec.decompress(c[1], to_host=False)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    launches0 = _lib.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(steps)]
    pts_done = 0
    t_wall0 = time.time()
    for i in range(steps):
        flush.fill_(i & 0xFF)                  # L2 flush between timed iterations (outside the events)
        e0, e1, e2 = ev[i]
        e0.record()
        c, _ = enc.compress(gop(i), staged=staged[i % n_frames])
        e1.record()
        dec.decompress(c[1], to_host=False)
        e2.record()
        pts_done += n_points[i % n_frames]
    barrier()
    t_wall = time.time() - t_wall0
    launches = _lib.launch_count() - launches0
    clocks = sampler.result()
    enc_ms = [a.elapsed_time(b) for a, b, _ in ev]
    dec_ms = [b.elapsed_time(c_) for _, b, c_ in ev]
    step_ms = [a + b for a, b in zip(enc_ms, dec_ms)]
    total_s = sum(step_ms) * 1e-3

    # ---------------- per-kernel pass (roofline leg): the same steps once more with a CUDA-event pair
    # around every launch of this library (programs then run entry by entry); not part of `value` ----
    prof = ops.KernelProfiler()
    ops.PROFILER = prof
    prof_steps = min(steps, 5)
    for i in range(prof_steps):
        flush.fill_(i & 0xFF)
        c, _ = enc.compress(gop(i), staged=staged[i % n_frames])
        dec.decompress(c[1], to_host=False)
    torch.cuda.synchronize()
    ops.PROFILER = None

    # ---------------- end-to-end arm (public API, host buffers) ----------------
    enc2 = CompressionPipeline(settings, device=dev, model=model)
    dec2 = DecompressionPipeline(device=dev, model=model)
    out = None
    for i in range(0 if args.no_e2e else max(warmup, 4)):
        # warm up exactly like the timed loop: `out` keeps the previous reconstruction alive while the next
        # one is produced, so the pinned-buffer cache must hold two sets before timing starts
        c, _ = enc2.compress(gop(i))
        out, _ = dec2.decompress(c[1])
    enc2.h2d_bytes = enc2.d2h_bytes = dec2.h2d_bytes = dec2.d2h_bytes = 0
    barrier()
    e2e_ms = []
    e2e_steps = 0 if args.no_e2e else max(steps // 2, 3)
    for i in range(e2e_steps):
        flush.fill_(i & 0xFF)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        c, _ = enc2.compress(gop(i))
        out, _ = dec2.decompress(c[1])
        b.record()
        b.synchronize()
        e2e_ms.append(a.elapsed_time(b))
    barrier()
    if args.breakdown and rank == 0:
        print("e2e per-step ms: " + " ".join(f"{t:.2f}" for t in e2e_ms), file=sys.stderr)
    e2e_total = sum(e2e_ms) * 1e-3
    e2e_pts = sum(n_points[i % n_frames] for i in range(e2e_steps))

    # ---------------- reduce over ranks ----------------
    if world > 1:
        t = torch.tensor([total_s, e2e_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_s, e2e_total = float(t[0]), float(t[1])
        p = torch.tensor([pts_done, e2e_pts, launches], dtype=torch.float64, device=dev)
        dist.all_reduce(p, op=dist.ReduceOp.SUM)
        pts_done, e2e_pts, launches = float(p[0]), float(p[1]), int(p[2])

    if rank == 0:
        hbm_peak, tf_peak, which = _peaks()
        summ = prof.summary()
        fams = sorted(summ.items(), key=lambda kv: -kv[1]["ms"])
        kernel_ms = sum(v["ms"] for v in summ.values())
        top_name, _ = fams[0]
        # the dominant kernel = the largest launch shape of the family with the largest time share
        shapes = {}
        for fam, nbytes, flops, e0, e1 in prof.records:
            if fam == top_name:
                d = shapes.setdefault(nbytes, {"launches": 0, "ms": 0.0, "bytes": 0, "flops": 0})
                d["launches"] += 1
                d["ms"] += e0.elapsed_time(e1)
                d["bytes"] += nbytes
                d["flops"] += flops
        top_bytes = max(shapes)
        top = shapes[top_bytes]
        gbs = top["bytes"] / (top["ms"] * 1e-3) / 1e9
        tfs = top["flops"] / (top["ms"] * 1e-3) / 1e12
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tpath):
            best = None
            for ent in json.load(open(tpath)).get("kernels", []):
                err = abs(ent["algorithmic_bytes"] - top_bytes) / top_bytes
                if ent["family"] == top_name and err <= 0.05 and (best is None or err < best[0]):
                    best = (err, ent["dram_bytes_per_launch"])
            traffic = best[1] if best else None
        roof = {"bound": "hbm", "kernel": top_name, "achieved": gbs, "peak": hbm_peak, "unit": "GB/s",
                "frac": gbs / hbm_peak, "traffic": traffic, "peak_source": which,
                "algorithmic_bytes_per_launch": int(top_bytes),
                "launches_per_step": top["launches"] / prof_steps, "avg_launch_ms": top["ms"] / top["launches"],
                "share_of_kernel_time": top["ms"] / kernel_ms,
                "family_share_of_kernel_time": fams[0][1]["ms"] / kernel_ms, "tensor_tflops": tfs,
                "tensor_frac": tfs / tf_peak,
                "note": "largest launch shape of the top family; traffic = dram read+write bytes of the same "
                        "shape from the ncu --set full capture summarised under profiles/"}
        if args.breakdown:
            print(f"{'family':34s} {'n/step':>7s} {'ms/step':>9s} {'GB/s':>9s} {'TF/s':>8s}", file=sys.stderr)
            for name, v in fams:
                print(f"{name:34s} {v['launches'] / prof_steps:7.1f} {v['ms'] / prof_steps:9.3f} "
                      f"{v['bytes'] / max(v['ms'], 1e-9) / 1e6:9.1f} {v['flops'] / max(v['ms'], 1e-9) / 1e9:8.2f}",
                      file=sys.stderr)
            print(f"profiled kernel time/step {kernel_ms / prof_steps:.3f} ms (separate pass); step {np.mean(step_ms):.3f} ms "
                  f"(enc {np.mean(enc_ms):.3f}, dec {np.mean(dec_ms):.3f}); wall/step {1e3 * t_wall / steps:.3f} ms",
                  file=sys.stderr)
        line = {
            "metric": METRIC, "value": pts_done / total_s / 1e6, "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": 1e3 * total_s / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "points_per_frame": int(n_points[0]), "frames_per_step": 1,
                       "model": "synth_small (seeded weights; the demo_small weights are not in the reference tree)",
                       "l2": "256 MiB flush write between timed iterations", "parallelism": f"frame-parallel x{world}"},
            "latency_ms": {"encode_p50": float(np.median(enc_ms)), "decode_p50": float(np.median(dec_ms)),
                           "step_p50": float(np.median(step_ms))},
            "fps": steps / total_s,
            "e2e": None if not e2e_steps else {
                "value": e2e_pts / e2e_total / 1e6, "unit": UNIT,
                "h2d_bytes_per_step": int((enc2.h2d_bytes + dec2.h2d_bytes) / e2e_steps),
                "d2h_bytes_per_step": int((enc2.d2h_bytes + dec2.d2h_bytes) / e2e_steps),
                "ms_per_step": 1e3 * e2e_total / e2e_steps},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
        }
        if not args.no_cpu_baseline and world == 1:
            n, times = cpu_port_run(frames[0], args.cpu_sample, reps=1)
            line["cpu_baseline"] = {"value": n / times[0] / 1e6, "unit": UNIT, "cores": _blas_threads(), "kind": "port",
                                    "sample": f"{n}-point slab of the cfg-2 frame, encode+decode once "
                                              f"({times[0]:.1f} s; numpy BLAS threads + scalar C rANS)"}
        else:
            line["cpu_baseline"] = None
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
