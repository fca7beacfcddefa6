"""Generates tests/golden/oracle_small.npz from the CPU oracle (the reference ships no golden vectors
and none of its dependencies can run here, SURVEY.md 8c: these fixtures pin the ORACLE, "parity
unpinned" with respect to the reference).  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200"))

from b200pcc import config as cfgmod, synth  # noqa: E402
from oracle.codec_ref import RefCodec  # noqa: E402

SEED, N = 31, 3000
frames = [synth.small_frame(seed=SEED, n_target=N)]
codec = RefCodec(cfgmod.SYNTH_SMALL, cfgmod.make_state_dict(), bf16=True)
blob = codec.compress(frames, [[1.0, 1.0]])
out, _ = codec.decompress(blob, 0)
np.savez_compressed(
    os.path.join(ROOT, "tests", "golden", "oracle_small.npz"),
    seed=SEED, n_target=N, points=frames[0]["points"], y_coords=blob["y_coords"], k=np.array(blob["k"]),
    y_sym=blob["debug"]["y_sym"][0].astype(np.int16), y_string=np.frombuffer(blob["y_strings"][0], np.uint8),
    z_string=np.frombuffer(blob["z_string"], np.uint8), dec_points=out[0]["points"].astype(np.int16),
    dec_colors=out[0]["colors"].astype(np.float32))
print("points", len(frames[0]["points"]), "Ny", len(blob["y_coords"]), "bytes", len(blob["y_strings"][0]))
