import os, sys, time, cProfile, pstats, io
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
for _ in range(4):
    c, _ = enc.compress({"frames": [frame]}, staged=staged)
    dec.decompress(c[1], to_host=False)
torch.cuda.synchronize()
# wall-clock trace of ops-level calls
import functools
log = []
def wrap(mod, name):
    f = getattr(mod, name)
    @functools.wraps(f)
    def g(*a, **k):
        t0 = time.perf_counter(); r = f(*a, **k); log.append((name, t0, time.perf_counter())); return r
    setattr(mod, name, g)
for n in ("sort_unique", "coords_pyramid", "blockmap_build", "kmap_k3_blocks", "spconv_umma", "rows_f32_to_bf16", "octree_encode", "kmap_down2", "rans_encode", "gauss_encode", "hash_query", "lex_order", "factorized_quant", "hash_build", "ingest_frame", "octree_decode", "coords_pyramid_expect", "rans_decode", "topk_select", "gather_rows", "verify_deferred"):
    wrap(ops, n)
t00 = time.perf_counter()
c, _ = enc.compress({"frames": [frame]}, staged=staged)
t01 = time.perf_counter()
out = dec.decompress(c[1], to_host=False)
t02 = time.perf_counter()
print(f"compress {1e3*(t01-t00):.3f} ms decompress {1e3*(t02-t01):.3f} ms")
prev = t00
for name, a, b in log[:40]:
    print(f"{name:22s} start +{1e6*(a-t00):8.1f} us  dur {1e6*(b-a):7.1f} us  since-prev-end {1e6*(a-prev):7.1f}")
    prev = b
print("...decoder")
prev = t01
for name, a, b in [x for x in log if x[1] > t01][:30]:
    print(f"{name:22s} start +{1e6*(a-t01):8.1f} us  dur {1e6*(b-a):7.1f} us  since-prev-end {1e6*(a-prev):7.1f}")
    prev = b
