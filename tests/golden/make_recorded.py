"""Converts a few of the reference's RECORDED capturer frames (evaluation/data/test_sequence/frame_0000*.pkl,
written by sender/capturer/capturer.py:136-173) into tests/golden/recorded_frames.npz -- real codec inputs
(19 679 points in frame 0, negative int16 coordinates, already de-duplicated) for the coordinate / kernel-map /
round-trip tests.  The pickles are untrusted content: they are read with an unpickler that only
resolves numpy's array reconstruction.  Category (b) fixture: reference DATA, not code.  Run in the build
container (the reference tree does not exist on the GPU box):  python tests/golden/make_recorded.py"""
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SRC = "/root/reference/evaluation/data/test_sequence"
FRAMES = [0, 1, 2]


class ArraysOnly(pickle.Unpickler):
    _ALLOWED = {("numpy.core.multiarray", "_reconstruct"), ("numpy._core.multiarray", "_reconstruct"),
                ("numpy", "ndarray"), ("numpy", "dtype")}

    def find_class(self, module, name):
        if (module, name) not in self._ALLOWED:
            raise pickle.UnpicklingError(f"{module}.{name} is not allowed")
        return super().find_class(module, name)


out = {}
for i in FRAMES:
    with open(os.path.join(SRC, f"frame_{i:05d}.pkl"), "rb") as fh:
        d = ArraysOnly(fh).load()
    p, c = np.asarray(d["points"]), np.asarray(d["colors"])
    assert p.dtype == np.int16 and c.dtype == np.float64 and p.shape == c.shape
    out[f"points_{i}"], out[f"colors_{i}"] = p, c
    print(i, p.shape, p.min(axis=0), p.max(axis=0))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "recorded_frames.npz"), frames=np.array(FRAMES), **out)
