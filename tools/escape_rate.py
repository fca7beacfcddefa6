"""How many latent symbols of the cfg-2 frame leave their table's support (escape-coded by the rANS lanes)?
The coder's run time is its chain of rounds, and a round in which any of the 32 lanes escapes takes several
extra steps -- with the seeded synthetic weights that is the common case, with a trained model it is rare."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from b200pcc import CompressionPipeline, synth  # noqa: E402
from b200pcc.pipeline import load_model  # noqa: E402

dev = torch.device("cuda:0")
model = load_model(dev)
enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
enc.keep_debug = True
frame = synth.frame_8ivfb(seed=1)
comp, side = enc.compress({"frames": [frame]})
gc = model.entropy_model.gaussian_conditional
sym = enc.debug["y_sym"][0].cpu().numpy().astype(np.int64).ravel()
idx = enc.debug["y_idx"][0].cpu().numpy().astype(np.int64).ravel()
# table symbol v = sym - offset; support 0 .. size - 3 (size - 2 is the escape entry)
v = sym - gc._offset[idx]
maxv = gc._cdf_length[idx] - 2
esc = (v < 0) | (v >= maxv)
n = sym.size
rounds = esc[: n // 32 * 32].reshape(-1, 32).any(axis=1)
print(json.dumps({"symbols": int(n), "escape_fraction": float(esc.mean()), "rounds_with_an_escape": float(rounds.mean()),
                  "bytes": len(comp[1]), "bpp": 8.0 * len(comp[1]) / len(frame["points"]),
                  "index_histogram_top": np.bincount(idx, minlength=64)[:8].tolist()}))
