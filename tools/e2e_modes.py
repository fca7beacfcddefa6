import os, sys, time
ROOT = "/root/repo"
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from b200pcc import CompressionPipeline, DecompressionPipeline, synth
from b200pcc.pipeline import load_model
dev = torch.device("cuda:0")
model = load_model(dev)
frames = [synth.frame_8ivfb(seed=1 + i) for i in range(2)]
enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
dec = DecompressionPipeline(device=dev, model=model)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(4):
    c, _ = enc.compress({"frames": [frames[i % 2]]}); dec.decompress(c[1])
te, td = [], []
for i in range(12):
    flush.fill_(i); torch.cuda.synchronize()
    t0 = time.perf_counter(); c, _ = enc.compress({"frames": [frames[i % 2]]}); t1 = time.perf_counter()
    out, _ = dec.decompress(c[1]); t2 = time.perf_counter()
    te.append(t1 - t0); td.append(t2 - t1)
print(os.environ.get("PCC_STAGE_MODE"), os.environ.get("PCC_STAGE_PIECE_MB"), f"enc {1e3*np.median(te):.2f} ms (min {1e3*min(te):.2f}) dec {1e3*np.median(td):.2f} ms (min {1e3*min(td):.2f})")
