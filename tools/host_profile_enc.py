import os, sys, time, cProfile, pstats
ROOT = "/root/repo"
for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
    sys.path.insert(0, p)
import torch
from b200pcc import CompressionPipeline, DecompressionPipeline, synth, ops
from b200pcc.pipeline import load_model
dev = torch.device("cuda:0")
model = load_model(dev)
frame = synth.frame_8ivfb(seed=1)
enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
dec = DecompressionPipeline(device=dev, model=model)
staged = enc.stage_frames([frame])
for _ in range(5):
    c, _ = enc.compress({"frames": [frame]}, staged=staged)
    dec.decompress(c[1], to_host=False)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    c, _ = enc.compress({"frames": [frame]}, staged=staged)
pr.disable()
st = pstats.Stats(pr); st.sort_stats("cumulative").print_stats(45)
