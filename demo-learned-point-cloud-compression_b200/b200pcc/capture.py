"""Capturer frame preparation on the device (capture half).

Mirrors what `Capturer.get_zed_frames` does between `retrieve_measure` and the frame dict it returns
(sender/capturer/capturer.py:79-124): split the XYZRGBA measure, drop invalid / far points, voxel-down-sample
(Open3D, float64), round to the int16 voxel grid, drop duplicate rows, keep the `max_points` rows of largest
z.  The settings carry the reference's names and defaults (capturer.py:23-25, shared/config.yaml:5-7).  All of
it runs in libpcc_b200.so (`pcc_capture_voxelize`, csrc/capture.cu); there is no CPU fallback.

Two ways out:
  * `capture(measure)` -> the reference's frame dict with host numpy arrays (what the capturer serialises
    and pushes to the encoder, capturer.py:47-50, 124);
  * `capture_staged(measure)` -> the same frame resident in HBM, in the form
    `CompressionPipeline.compress(gop, staged=...)` takes, so a capturer that lives in the encoder's process
    hands its frames over without the 30 bytes per point crossing the bus twice.
"""
import time

import numpy as np
import torch

from . import ops
from ._lib import PccError


class FrameVoxelizer:
    def __init__(self, config=None, device="cuda:0"):
        config = config or {}
        self.depth_clip = config.get("depth_clip", -1.5)      # capturer.py:23 (the default keeps no point)
        self.voxel_size = config.get("voxel_size", 0.01)      # capturer.py:24
        self.max_points = config.get("max_points", None)      # capturer.py:25
        if not torch.cuda.is_available():
            raise PccError("b200pcc.capture needs a CUDA device; there is no CPU fallback")
        self.device = torch.device(device)
        self._pinned = None

    def _measure_to_device(self, measure):
        """[h, w, 4] / [n, 4] float32, host numpy or device tensor -> device tensor [n, 4]."""
        if isinstance(measure, torch.Tensor):
            t = measure.reshape(-1, 4)
            if t.dtype != torch.float32:
                raise PccError(f"XYZRGBA measure must be float32, got {t.dtype}")
            return t.to(self.device, non_blocking=True).contiguous()
        a = np.ascontiguousarray(measure, dtype=np.float32).reshape(-1, 4)
        n = a.shape[0]
        if self._pinned is None or self._pinned.shape[0] < n:
            self._pinned = torch.empty((max(n, 1), 4), dtype=torch.float32).pin_memory()
        self._pinned[:n].numpy()[...] = a
        return self._pinned[:n].to(self.device, non_blocking=True)

    def voxelize(self, measure):
        """-> (points int16 [cap,3], colours float64 [cap,3], count int32 [1]) on the device, not waited for."""
        with torch.cuda.device(self.device):
            return ops.capture_voxelize(self._measure_to_device(measure), self.voxel_size, self.depth_clip,
                                        self.max_points)

    @staticmethod
    def _count(cnt, cap):
        m = int(cnt.item())
        if m == -1:
            raise PccError("capture: voxel index beyond 21 bits (scene extent / voxel_size too large)")
        if m == -2:
            raise PccError("capture: a voxel coordinate does not fit int16")
        if m < 0 or m > cap:
            raise PccError(f"capture: inconsistent row count {m} (capacity {cap})")
        return m

    def capture_staged(self, measure):
        """-> (frame, staged): `frame` = {"points": int16 cuda [m,3], "colors": float64 cuda [m,3]} and
        `staged` = {"points": [..], "colors": [..]} for CompressionPipeline.compress(gop, staged=staged)."""
        pts, col, cnt = self.voxelize(measure)
        m = self._count(cnt, pts.shape[0])
        frame = {"points": pts[:m], "colors": col[:m]}
        return frame, {"points": [frame["points"]], "colors": [frame["colors"]]}

    def capture(self, measure, timestamp=None):
        """-> {"points": int16 ndarray [m,3], "colors": float64 ndarray [m,3], "timestamp": float}
        (capturer.py:124)."""
        timestamp = time.time() if timestamp is None else timestamp
        frame, _ = self.capture_staged(measure)
        return {"points": frame["points"].cpu().numpy(), "colors": frame["colors"].cpu().numpy(), "timestamp": timestamp}
