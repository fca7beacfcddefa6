"""Generates tests/golden/capture_small.npz from the CPU oracle (oracle/capture_ref.py): a 96x128 XYZRGBA
measure and the frame the capturer's preparation makes of it (sender/capturer/capturer.py:84-120) for two
settings.  The reference holds no camera measure next to its recorded frames, so this pins the ORACLE
("parity unpinned" with respect to the reference).  Run from the repo root:
python tests/golden/make_golden_capture.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import capture_ref as R  # noqa: E402

measure = R.synth_measure(96, 128, seed=5)
a = R.capture_frame(measure, voxel_size=0.005, depth_clip=1.4, max_points=None)
b = R.capture_frame(measure, voxel_size=0.01, depth_clip=3.0, max_points=700)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "capture_small.npz"), measure=measure,
                    a_points=a["points"], a_colors=a["colors"], b_points=b["points"], b_colors=b["colors"])
print("a", a["points"].shape, "b", b["points"].shape)
