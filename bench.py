#!/usr/bin/env python
"""bench.py — codec Mpoints/s (encode + decode) of the learned point-cloud codec hot path.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference --steps K --warmup W     (CPU arm: the oracle port on host cores)

A "step" = compress() + decompress() of one GOP of one synthetic 8iVFB-shaped frame (cfg-2,
~800k coloured points in a 1024^3 grid), quality [1,1].  `value` is timed with the raw
frame already resident in HBM and the reconstruction left on the device; `e2e` goes through the public
CompressionPipeline.compress(gop) / DecompressionPipeline.decompress(bytes) calls with host buffers
(host->device copy of the frame and device->host copy of the reconstruction inside the timed region).
Both are throughputs with --inflight (default 2) GOPs per GPU: caller threads, each with its own pipeline
pair and CUDA stream, as the reference's encoder / decoder services run 3-worker pools around the two calls;
the K steps are bracketed by one start and one end event that all caller streams are tied to.  The latencies
(`latency_ms`, `rank_step_ms`) and the one-caller throughput (`sequential`) come from a separate pass.
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
        # NVML in-process when the module is there (a query costs microseconds); nvidia-smi otherwise -- starting that
        # program every 0.2 s attaches to the driver each time, which showed up as stalls of the two caller threads
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self.stop_flag:
                try:
                    self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                    r = int(get_reasons(h))
                    for n in names:
                        if r & bits[n]:
                            self.reasons.add(n)
                except Exception:
                    pass
                time.sleep(0.05)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
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


def _load_plain(name):
    """b200pcc/<name>.py as a stand-alone module (synth, config: numpy only).  The CPU arm must not import the
    b200pcc package -- its __init__ loads libpcc_b200.so, and the reference process is meant to hold no GPU code."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(f"_plain_{name}", os.path.join(PKG_DIR, "b200pcc", f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _config(points_per_frame, world):
    """The `config` object of the JSON line: the same for the GPU arm and the reference arm."""
    return {"workload": WORKLOAD, "points_per_frame": int(points_per_frame), "frames_per_step": 1,
            "model": "synth_small (seeded weights; the demo_small weights are not in the reference tree)",
            "l2": "256 MiB flush write between timed iterations", "parallelism": f"frame-parallel x{world}"}


def _crop_sample(frame, n_target):
    """Bounded CPU sample of the same workload: a horizontal slab of the frame with ~n_target points."""
    p = frame["points"]
    order = np.argsort(p[:, 2], kind="stable")
    mid = len(order) // 2
    sel = np.sort(order[max(mid - n_target // 2, 0): mid + n_target // 2])
    return {"points": p[sel], "colors": frame["colors"][sel]}


def cpu_port_run(frame, sample_points, reps=1, settings=((1.0, 1.0),)):
    """Oracle port timed on the host cores: encode + decode.  numpy for the bookkeeping, the C/OpenMP core of
    oracle/sparse_ref.c for kernel maps, convolutions and bf16 rounding, scalar C rans64 for the entropy coder.
    sample_points = 0: the whole frame.  -> (points, [seconds per rep], [stage seconds per rep])"""
    cfgmod = _load_plain("config")
    from oracle.codec_ref import RefCodec
    from oracle import sparse_ref as S
    S.use_fast(True)
    sample = _crop_sample(frame, sample_points) if sample_points else frame
    codec = RefCodec(cfgmod.SYNTH_SMALL, cfgmod.make_state_dict(), bf16=True)
    n = len(sample["points"])
    times, stages = [], []
    q = [list(x) for x in settings]
    for _ in range(reps):
        t0 = time.time()
        blob = codec.compress([sample], q)
        _, dside = codec.decompress(blob, len(q) - 1)
        times.append(time.time() - t0)
        st = {"enc_" + k: v for k, v in blob["times"].items()}
        st.update({"dec_" + k: v for k, v in dside["times"].items()})
        stages.append(st)
    return n, times, stages


def _cpu_threads():
    """Threads the port runs on: the OpenMP team of the C core (= host cores unless OMP_NUM_THREADS says otherwise)."""
    from oracle import sparse_ref as S
    return S.fast_threads()


def cpu_cfg1_baseline(warmup=3, reps=10):
    """cfg-1 (one synthetic 100 100-point frame, seed 0), encode + decode, Q = 1 and Q = 3,
    3 warm-ups / 10 repetitions, p50, per-stage split with the reference's stage names
    (sender/encoder/codec_pipeline.py:218-225, receiver/decoder/codec_parallel.py:160-167)."""
    synth = _load_plain("synth")
    frame = synth.frame_cpu100k(seed=0)
    out = {"config": "cfg-1 cpu-100k", "points": int(len(frame["points"])), "warmup": warmup, "reps": reps,
           "cores": _cpu_threads(), "kind": "port"}
    for name, settings in (("Q1", ((1.0, 1.0),)), ("Q3", ((1.0, 0.0), (0.0, 1.0), (1.0, 1.0)))):
        n, times, stages = cpu_port_run(frame, 0, reps=warmup + reps, settings=settings)
        t = np.asarray(times[warmup:])
        p50 = float(np.median(t))
        out[name] = {"seconds_p50": p50, "mpoints_s": n / p50 / 1e6,
                     "stage_seconds_p50": {k: float(np.median([s[k] for s in stages[warmup:]])) for k in stages[0]}}
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU path (MinkowskiEngine CPU backend + CompressAI) cannot be installed
    here (unvendored dependencies, no network), so this arm times the oracle port -- on the SAME cfg-2 frame as the
    GPU arm (--cpu-sample N > 0 restricts it to an N-point slab and says so).  Nothing of b200pcc is imported."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    synth = _load_plain("synth")
    frame = synth.frame_8ivfb(seed=1, n_target=args.points)
    steps, warmup = args.steps, args.warmup
    n, times, stages = cpu_port_run(frame, args.cpu_sample, reps=warmup + steps)
    t = times[warmup:]
    total = sum(t)
    value = n * len(t) / total / 1e6
    cores = _cpu_threads()
    sample = (f"the whole cfg-2 frame ({n} points)" if not args.cpu_sample else f"{n}-point slab of the cfg-2 frame") + \
        f", encode+decode, {len(t)} timed reps after {warmup} warm-ups"
    cfg = _config(len(frame["points"]), args.gpus)
    if args.cpu_sample:
        cfg["workload"] += f" -- REFERENCE ARM SAMPLE: {n}-point slab"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * total / len(t), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "what": "oracle port: numpy + C/OpenMP sparse ops (oracle/sparse_ref.c) + scalar C rans64; "
                                 "MinkowskiEngine / CompressAI / the model submodule are not installable here",
                         "stage_seconds_p50": {k: float(np.median([s[k] for s in stages[warmup:]])) for k in stages[0]}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_cfg4(enc, dec, dev, rank, world, nframes, barrier, pool=4):
    """cfg-4 as BASELINE.json states it: a sequence of `nframes` cfg-2-shaped frames, frame i -> rank i mod G
    (b200pcc.shard), every rank coding its frames through the public compress() / decompress() calls with host
    buffers; then the sequence's bitstreams are gathered on every rank (NCCL all_gather of lengths and of padded
    payloads) OUTSIDE the timed region, its time reported.  Frame i of the sequence is pool frame i mod `pool`
    (seeds 1000 + i mod pool): generating 300 distinct 800 k-point frames would take the host several minutes.
    The digest of the gathered sequence must be the same for every G."""
    import hashlib
    import torch
    import torch.distributed as dist
    from b200pcc import shard
    synth = _load_plain("synth")
    mine = shard.frames_for_rank(nframes, rank, world)
    frames = {j: synth.frame_8ivfb(seed=1000 + j) for j in sorted({i % pool for i in mine})}
    for i in mine[:3]:
        c, _ = enc.compress({"frames": [frames[i % pool]], "timestamps": {}})
        dec.decompress(c[1])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    local, pts = {}, 0
    t0 = time.time()
    e0.record()
    for i in mine:
        c, _ = enc.compress({"frames": [frames[i % pool]], "timestamps": {}})
        dec.decompress(c[1])
        local[i] = c[1]
        pts += len(frames[i % pool]["points"])
    e1.record()
    e1.synchronize()
    own_s = e0.elapsed_time(e1) * 1e-3
    barrier()
    wall_s = time.time() - t0
    tg = time.time()
    seq = shard.gather_bitstreams(local, nframes, None, dev)
    torch.cuda.synchronize()
    gather_s = time.time() - tg
    stats = torch.tensor([own_s, wall_s, gather_s, float(pts)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        own_s, wall_s, gather_s, pts = float(mx[0]), float(mx[1]), float(mx[2]), float(sm[3])
    return {"config": "cfg-4 seq-300", "frames": nframes, "ranks": world, "sharding": "frame i -> rank i mod G (b200pcc.shard)",
            "distinct_frames": pool, "points_total": int(pts), "seconds_max_over_ranks": own_s,
            "mpoints_s": pts / own_s / 1e6, "frames_per_s": nframes / own_s, "wall_seconds_incl_barrier": wall_s,
            "gather_seconds": gather_s, "gather": "all_gather (NCCL) of lengths + padded payloads, outside the timed region"
            if world > 1 else "single rank: no collective", "sequence_bytes": sum(len(b) for b in seq),
            "sequence_sha256": hashlib.sha256(b"".join(seq)).hexdigest(), "buffers": "host numpy in, host numpy out"}


def run_configs(model, dev):
    """cfg-1 / cfg-3 / cfg-5 of BASELINE.json on one GPU (tools/bench_configs.py holds the procedures): cfg-1 end to
    end, cfg-3 frames arriving at 30 fps (p50 / p95 / p99 latency), cfg-5 encode time vs Q (1, 3, 25)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_configs as bc
    return {"cfg1": bc.cfg1(model, dev, 10), "cfg3": bc.cfg3(model, dev, 120, 30.0), "cfg5": bc.cfg5(model, dev, 6)}


def run_inflight(torch, pairs, streams, gop_of, staged_of, steps, flush, host_buffers, barrier):
    # Resident frames: the callers draw their steps from one counter and start together (183-187 Mpoints/s over five
    # runs; out of phase 175-187).  Host buffers: caller w takes the steps w, w + W, ... and starts w / W of 2.5 ms late,
    # so that one caller's copies fall on the other's kernels (4.55-4.82 ms/step against 4.75-5.09 in phase,
    # tools/inflight_bench.py --stagger-ms, profiles/r2u_*).
    """K steps (compress + decompress of one GOP each) issued by len(pairs) caller threads, worker w taking the steps
    w, w + W, ...; every worker has its own pipeline pair and CUDA stream, so the kernels of one GOP fill the gaps and the
    thin phases (lane coder, host turn-arounds) of the other.  Device time: a start event every worker stream waits
    for, an end event that waits for every worker.  The L2 flush of a step is inside the timed region.
    -> (seconds, points coded)."""
    import threading
    import itertools
    W = len(pairs)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ends = [torch.cuda.Event() for _ in range(W)]
    errors, pts = [], [0] * W
    ticket, ticket_lock = itertools.count(), threading.Lock()

    mine = [iter(range(w, steps, W)) for w in range(W)]

    def next_step(w):
        if host_buffers:      # fixed share: the callers keep their phase offset
            return next(mine[w], None)
        with ticket_lock:     # one counter: the callers finish together
            i = next(ticket)
        return i if i < steps else None

    def work(w):
        try:
            enc, dec = pairs[w]
            with torch.cuda.stream(streams[w]):
                streams[w].wait_event(start)
                if w and host_buffers:       # callers out of phase: one copies while the other computes
                    time.sleep(2.5e-3 * w / W)
                out = None
                while True:
                    i = next_step(w)
                    if i is None:
                        break
                    flush.fill_(i & 0xFF)
                    g, n = gop_of(i)
                    c, _ = enc.compress(g) if host_buffers else enc.compress(g, staged=staged_of(w, i))
                    out = dec.decompress(c[1], to_host=host_buffers)
                    pts[w] += n
                ends[w].record(streams[w])
                del out
        except BaseException as exc:      # noqa: BLE001  (reported by the caller; a worker must not die silently)
            errors.append(f"worker {w}: {exc!r}")

    # the interpreter hands the lock over every 5 ms by default: a caller coming back from a C call could wait that long
    # behind the other one's Python; 0.2 ms keeps the two launch sequences interleaved
    switch = sys.getswitchinterval()
    sys.setswitchinterval(2e-4)
    barrier()
    start.record()
    threads = [threading.Thread(target=work, args=(w,), name=f"bench-caller-{w}") for w in range(W)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    sys.setswitchinterval(switch)
    if errors:
        raise SystemExit("bench: " + "; ".join(errors))
    for w in range(W):
        torch.cuda.current_stream().wait_event(ends[w])
    end.record()
    barrier()
    return start.elapsed_time(end) * 1e-3, sum(pts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="CPU arm: points of a slab sample (0 = the whole frame)")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg-1 / cfg-3 / cfg-4 / cfg-5 legs")
    ap.add_argument("--seq-frames", type=int, default=300, help="frames of the cfg-4 sequence")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--points", type=int, default=800_000)
    ap.add_argument("--breakdown", action="store_true", help="print per-kernel-family table to stderr")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer arm (short runs under ncu)")
    ap.add_argument("--inflight", type=int, default=2,
                    help="GOPs in flight per GPU in the throughput arms: caller threads, each with its own pipeline pair and "
                         "CUDA stream (the reference runs 3-worker pools around compress() / decompress()); 1 = one caller")
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
        # the ranks of a node share its host cores (every GPU of the bench box reports the same CPU affinity): give each
        # rank its own slice, so that its Python thread and its staging threads do not migrate onto another rank's cores
        try:
            cores = sorted(os.sched_getaffinity(0))
            lw = int(os.environ.get("LOCAL_WORLD_SIZE", world))
            per = len(cores) // lw
            if per >= 2 and os.environ.get("PCC_BENCH_PIN", "1") != "0":
                os.sched_setaffinity(0, set(cores[local * per:(local + 1) * per]))
        except (AttributeError, OSError):
            pass
        dist.init_process_group("nccl", device_id=dev)

    steps, warmup = args.steps, max(args.warmup, 4)      # (every distinct frame is seen twice before timing starts)
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
    enc_ms = [a.elapsed_time(b) for a, b, _ in ev]
    dec_ms = [b.elapsed_time(c_) for _, b, c_ in ev]
    step_ms = [a + b for a, b in zip(enc_ms, dec_ms)]
    total_s = sum(step_ms) * 1e-3
    seq_value = pts_done / total_s / 1e6          # one caller: the sum of the steps' own device times (flush excluded)
    seq_ms_per_step = 1e3 * total_s / steps

    # ---------------- throughput: `inflight` GOPs per GPU, each caller thread with its own pipelines and stream ----------------
    W = max(1, args.inflight)
    pairs = [(enc, dec)] + [(CompressionPipeline(settings, device=dev, model=model), DecompressionPipeline(device=dev, model=model))
                            for _ in range(W - 1)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(W)]
    if W > 1:
        staged_w = [staged] + [[pairs[w][0].stage_frames(g["frames"]) for g in (gop(0), gop(1))] for w in range(1, W)]
        torch.cuda.synchronize()
        for w in range(1, W):                      # warm the new pipelines (workspaces, pinned buffers) one by one ...
            for i in range(max(warmup, 3)):
                c, _ = pairs[w][0].compress(gop(i), staged=staged_w[w][i % n_frames])
                pairs[w][1].decompress(c[1], to_host=False)
        run_inflight(torch, pairs, streams, lambda i: (gop(i), n_points[i % n_frames]),
                     lambda w, i: staged_w[w][i % n_frames], max(4 * W, 8), flush, False, barrier)     # ... and together
        launches0 = _lib.launch_count()
        total_s, pts_done = run_inflight(torch, pairs, streams, lambda i: (gop(i), n_points[i % n_frames]),
                                         lambda w, i: staged_w[w][i % n_frames], steps, flush, False, barrier)
        launches = _lib.launch_count() - launches0
    clocks = sampler.result()

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
    e2e_h2d, e2e_d2h = enc2.h2d_bytes + dec2.h2d_bytes, enc2.d2h_bytes + dec2.d2h_bytes
    e2e_seq = None if not e2e_steps else {"value": e2e_pts / e2e_total / 1e6, "ms_per_step": 1e3 * e2e_total / e2e_steps}
    if W > 1 and e2e_steps:
        pairs2 = [(enc2, dec2)] + [(CompressionPipeline(settings, device=dev, model=model), DecompressionPipeline(device=dev, model=model))
                                   for _ in range(W - 1)]
        for w in range(1, W):
            for i in range(max(warmup, 4)):
                c, _ = pairs2[w][0].compress(gop(i))
                pairs2[w][1].decompress(c[1])
        run_inflight(torch, pairs2, streams, lambda i: (gop(i), n_points[i % n_frames]), None, max(6 * W, 12), flush, True, barrier)
        for e_, d_ in pairs2:
            e_.h2d_bytes = e_.d2h_bytes = d_.h2d_bytes = d_.d2h_bytes = 0
        e2e_steps = steps
        e2e_total, e2e_pts = run_inflight(torch, pairs2, streams, lambda i: (gop(i), n_points[i % n_frames]), None, e2e_steps,
                                          flush, True, barrier)
        e2e_h2d = sum(e_.h2d_bytes + d_.h2d_bytes for e_, d_ in pairs2)
        e2e_d2h = sum(e_.d2h_bytes + d_.d2h_bytes for e_, d_ in pairs2)

    # ---------------- the other BASELINE.json configurations ----------------
    configs = None
    if not args.no_configs:
        configs = {"cfg4": run_cfg4(enc2, dec2, dev, rank, world, args.seq_frames, barrier)}
        if world == 1:
            configs.update(run_configs(model, dev))

    # ---------------- reduce over ranks ----------------
    mine = torch.tensor([float(np.median(step_ms)), float(np.median(e2e_ms)) if e2e_ms else 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        allr = torch.stack(allr).cpu().numpy()
    else:
        allr = mine.cpu().numpy()[None]
    rank_stats = {"resident_p50": {"min": float(allr[:, 0].min()), "median": float(np.median(allr[:, 0])), "max": float(allr[:, 0].max())},
                  "e2e_p50": {"min": float(allr[:, 1].min()), "median": float(np.median(allr[:, 1])), "max": float(allr[:, 1].max())}}
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
        for fam, nbytes, flops, e0, e1, dense in prof.records:
            if fam == top_name:
                d = shapes.setdefault(nbytes, {"launches": 0, "ms": 0.0, "bytes": 0, "flops": 0, "dense": 0})
                d["launches"] += 1
                d["ms"] += e0.elapsed_time(e1)
                d["bytes"] += nbytes
                d["flops"] += flops
                d["dense"] += dense
        top_bytes = max(shapes)
        top = shapes[top_bytes]
        gbs = top["bytes"] / (top["ms"] * 1e-3) / 1e9
        tfs = top["flops"] / (top["ms"] * 1e-3) / 1e12            # 2 P Cin Cout, P counted from the kernel map
        tfs_dense = top["dense"] / (top["ms"] * 1e-3) / 1e12      # every (offset, row) incl. absent neighbours
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
                "tensor_frac": tfs / tf_peak, "tensor_tflops_dense_equivalent": tfs_dense,
                "pairs_per_output_row": top["flops"] / max(top["dense"], 1) * 27.0 if "_k27_" in top_name else None,
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
            "config": _config(n_points[0], world),
            "latency_ms": {"encode_p50": float(np.median(enc_ms)), "decode_p50": float(np.median(dec_ms)),
                           "step_p50": float(np.median(step_ms)), "step_max": float(np.max(step_ms)),
                           "step_ms_rank0": [round(float(v), 3) for v in step_ms]},
            "fps": steps / total_s,
            "e2e": None if not e2e_steps else {
                "value": e2e_pts / e2e_total / 1e6, "unit": UNIT,
                "h2d_bytes_per_step": int(e2e_h2d / e2e_steps),
                "d2h_bytes_per_step": int(e2e_d2h / e2e_steps),
                "ms_per_step": 1e3 * e2e_total / e2e_steps},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
            # value / e2e are throughputs with `inflight` GOPs per GPU (caller threads with their own pipeline pair and
            # CUDA stream, like the reference's 3-worker pools around compress() / decompress()); the one-caller numbers
            # (sum of the steps' own device times) and the latencies above are measured in a separate pass
            "inflight": W,
            "sequential": {"value": seq_value, "ms_per_step": seq_ms_per_step, "e2e": e2e_seq,
                           "note": "one caller, rank 0: compress() then decompress() of a GOP, next GOP after that"},
        }
        line["rank_step_ms"] = rank_stats
        if configs is not None:
            line["configs"] = configs
        if not args.no_cpu_baseline and world == 1:
            # the same workload on the host cores: ONE encode + decode of the whole cfg-2 frame (about 10 s of CPU
            # work), then BASELINE.md section 3's procedure on cfg-1 (3 warm-ups, 10 repetitions, p50, stage split)
            n, times, stages = cpu_port_run(frames[0], args.cpu_sample, reps=1)
            line["cpu_baseline"] = {"value": n / times[0] / 1e6, "unit": UNIT, "cores": _cpu_threads(), "kind": "port",
                                    "sample": (f"the whole cfg-2 frame ({n} points)" if not args.cpu_sample else
                                               f"{n}-point slab of the cfg-2 frame") +
                                              f", encode+decode once ({times[0]:.1f} s; numpy + C/OpenMP sparse ops + scalar C rans64)",
                                    "stage_seconds": stages[0],
                                    "cfg1": cpu_cfg1_baseline()}
        else:
            line["cpu_baseline"] = None
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
