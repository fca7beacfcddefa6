#!/usr/bin/env python
"""Throughput with several GOPs in flight on one GPU: W worker threads, each with its own CompressionPipeline /
DecompressionPipeline pair and CUDA stream, code frames concurrently (the reference runs 3-worker pools around
compress() / decompress(): sender/encoder/encoder.py:50,75).  Device time from a start event all worker streams wait
for to an end event that waits for all of them.  Usage: python tools/inflight_bench.py [--workers 1,2,3] [--steps 24] [--e2e]"""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from b200pcc import CompressionPipeline, DecompressionPipeline, synth  # noqa: E402
from b200pcc.pipeline import load_model  # noqa: E402


def arg(name, default):
    return sys.argv[sys.argv.index(name) + 1] if name in sys.argv else default


def main():
    dev = torch.device("cuda:0")
    workers_list = [int(w) for w in arg("--workers", "1,2,3").split(",")]
    steps = int(arg("--steps", "24"))
    e2e = "--e2e" in sys.argv
    stagger_ms = float(arg("--stagger-ms", "0"))
    reps = int(arg("--reps", "1"))
    model = load_model(dev)
    frames = [synth.frame_8ivfb(seed=1 + i) for i in range(2)]
    npts = [len(f["points"]) for f in frames]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ref_bytes = {}
    for W in workers_list:
        encs = [CompressionPipeline([[1.0, 1.0]], device=dev, model=model) for _ in range(W)]
        decs = [DecompressionPipeline(device=dev, model=model) for _ in range(W)]
        streams = [torch.cuda.Stream(device=dev) for _ in range(W)]
        staged = [[None if e2e else encs[w].stage_frames([f]) for f in frames] for w in range(W)]
        torch.cuda.synchronize()
        start, ends = torch.cuda.Event(enable_timing=True), [torch.cuda.Event(enable_timing=True) for _ in range(W)]
        end = torch.cuda.Event(enable_timing=True)
        errors, digests = [], [None] * W

        def work(w, n, timed):
            try:
                with torch.cuda.stream(streams[w]):
                    if timed:
                        streams[w].wait_event(start)
                        if stagger_ms and w:
                            time.sleep(1e-3 * stagger_ms * w / W)      # callers out of phase: one encodes while the other decodes
                    for i in range(n):
                        j = (w + i) % 2
                        flush.fill_(i & 0xFF)          # L2 flush between this worker's iterations (inside the timed region)
                        c, _ = encs[w].compress({"frames": [frames[j]], "timestamps": {}}, staged=staged[w][j]) if not e2e else \
                            encs[w].compress({"frames": [frames[j]], "timestamps": {}})
                        out = decs[w].decompress(c[1], to_host=e2e)
                        if j == 0:
                            digests[w] = hash(c[1])
                    if timed:
                        ends[w].record(streams[w])
            except Exception as exc:      # noqa: BLE001
                errors.append(repr(exc))

        results = []
        for timed, n in [(False, 4)] + [(True, steps // W)] * reps:
            torch.cuda.synchronize()
            if timed:
                start.record()
            th = [threading.Thread(target=work, args=(w, n, timed)) for w in range(W)]
            for t in th:
                t.start()
            for t in th:
                t.join()
            if timed:
                for w in range(W):
                    torch.cuda.current_stream().wait_event(ends[w])
                end.record()
            torch.cuda.synchronize()
            if timed:
                results.append(start.elapsed_time(end))
        if errors:
            print(f"workers {W}: ERROR {errors[:2]}")
            continue
        ms = float(np.median(results))
        done = (steps // W) * W
        pts = sum(npts[(w + i) % 2] for w in range(W) for i in range(steps // W))
        ref_bytes.setdefault(0, digests[0])
        same = all(d == ref_bytes[0] for d in digests if d is not None)
        print(f"workers {W} stagger {stagger_ms} ms, {reps} reps, ms/step each: " + " ".join(f"{r / done:.3f}" for r in results))
        print(f"workers {W}: {done} steps in {ms:.2f} ms (median) -> {ms / done:.3f} ms/step, {pts / ms / 1e3:.1f} Mpoints/s "
              f"({'host buffers' if e2e else 'resident'}; flush inside; frame-0 bitstreams identical across workers: {same})")


if __name__ == "__main__":
    main()
