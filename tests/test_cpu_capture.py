"""CPU: the capture oracle (oracle/capture_ref.py, restating sender/capturer/capturer.py:84-120) against its
scalar twin, its golden fixture, and the invariants the reference's recorded frames show."""
import glob
import os
import pickle

import numpy as np
import pytest

from oracle import capture_ref as R

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "capture_small.npz")
RECORDED = "/root/reference/evaluation/data"


class _ArraysOnly(pickle.Unpickler):
    """The recorded frames are untrusted content: only numpy's array reconstruction may be resolved."""
    _ALLOWED = {("numpy.core.multiarray", "_reconstruct"), ("numpy._core.multiarray", "_reconstruct"),
                ("numpy", "ndarray"), ("numpy", "dtype")}

    def find_class(self, module, name):
        if (module, name) not in self._ALLOWED:
            raise pickle.UnpicklingError(f"{module}.{name} is not allowed")
        return super().find_class(module, name)


def _invariants(frame, max_points=None):
    p, c = frame["points"], frame["colors"]
    assert p.dtype == np.int16 and c.dtype == np.float64 and p.shape == c.shape and p.shape[1] == 3
    assert np.array_equal(np.lexsort((p[:, 2], p[:, 1], p[:, 0])), np.arange(len(p)))      # np.unique order
    assert len(np.unique(p, axis=0)) == len(p)
    assert len(p) == 0 or (c.min() >= 0.0 and c.max() <= 1.0)
    if max_points is not None:
        assert len(p) <= max_points


def test_split_measure_takes_the_low_three_bytes():
    data = np.zeros((2, 4), np.float32)
    data[:, :3] = [[1, 2, 3], [4, 5, 6]]
    data[:, 3] = np.array([0xFF332211, 0x00FFEEDD], np.uint32).view(np.float32)
    pts, col = R.split_measure(data)
    assert pts.tolist() == [[1, 2, 3], [4, 5, 6]]
    assert col.tolist() == [[0x11, 0x22, 0x33], [0xDD, 0xEE, 0xFF]]


def test_valid_mask_is_a_float32_comparison():
    pts = np.array([[0.6, 0.8, 0.0], [np.nan, 0, 0], [0, np.inf, 0], [0, 0, -1.0000001], [0, 0, -1.0000002]], np.float32)
    # float32(1.0000001) == float32(1 + 2**-23): kept only up to the float32 value of the clip
    assert R.valid_mask(pts, 1.0000001).tolist() == [True, False, False, True, False]
    assert R.valid_mask(pts, -1.5).tolist() == [False] * 5          # the reference's default keeps nothing


@pytest.mark.parametrize("vs,clip,mp", [(0.005, 1.4, None), (0.01, 3.0, 500), (0.02, 2.0, 64)])
def test_vectorised_statement_equals_scalar_statement(vs, clip, mp):
    m = R.synth_measure(90, 120, seed=11)
    a = R.capture_frame(m, vs, clip, mp)
    b = R.capture_frame(m, vs, clip, mp, down_sample=R.voxel_down_sample_scalar)
    assert np.array_equal(a["points"], b["points"]) and np.array_equal(a["colors"], b["colors"])
    _invariants(a, mp)
    assert len(a["points"]) > 50


def test_running_sum_is_not_a_pairwise_sum():
    """The per-voxel sums follow Open3D's `+=` in input order; numpy's reduction order would differ."""
    rng = np.random.default_rng(0)
    pts = rng.random((4000, 3)) * 0.009            # every point in one voxel of size 1 around the minimum
    col = rng.random((4000, 3))
    p, c = R.voxel_down_sample(pts, col, 1.0)
    seq = np.zeros(3)
    for row in col:
        seq += row
    assert p.shape == (1, 3) and np.array_equal(c[0], seq / 4000.0)


def test_max_points_keeps_the_largest_z():
    m = R.synth_measure(120, 160, seed=2)
    full = R.capture_frame(m, 0.01, 3.0, None)
    cut = R.capture_frame(m, 0.01, 3.0, 1000)
    assert len(full["points"]) > 1000 and len(cut["points"]) == 1000
    z = full["points"][:, 2]
    ref = np.argpartition(z, -1000)[-1000:]                     # capturer.py:118
    assert sorted(z[ref].tolist()) == sorted(cut["points"][:, 2].tolist())
    # every kept row is a row of the full frame with its colour
    lut = {tuple(p): tuple(c) for p, c in zip(full["points"].tolist(), full["colors"].tolist())}
    assert all(lut[tuple(p)] == tuple(c) for p, c in zip(cut["points"].tolist(), cut["colors"].tolist()))


def test_empty_and_all_invalid_measures():
    assert R.capture_frame(np.zeros((0, 4), np.float32), 0.01, 1.0)["points"].shape == (0, 3)
    bad = np.full((10, 4), np.nan, np.float32)
    out = R.capture_frame(bad, 0.01, 1.0, 5)
    assert out["points"].shape == (0, 3) and out["colors"].shape == (0, 3)


def test_golden_fixture():
    g = np.load(GOLDEN)
    a = R.capture_frame(g["measure"], voxel_size=0.005, depth_clip=1.4, max_points=None)
    b = R.capture_frame(g["measure"], voxel_size=0.01, depth_clip=3.0, max_points=700)
    assert np.array_equal(a["points"], g["a_points"]) and np.array_equal(a["colors"], g["a_colors"])
    assert np.array_equal(b["points"], g["b_points"]) and np.array_equal(b["colors"], g["b_colors"])


@pytest.mark.skipif(not os.path.isdir(RECORDED), reason="the reference tree is only present in the build container")
def test_recorded_frames_show_the_same_invariants():
    """The reference's recorded capturer output (evaluation/data/*/frame_*.pkl, written by
    capturer.py:130-139) is int16 / float64, unique, in np.unique's (x,y,z) order, colours in [0,1] and within
    shared/config.yaml's max_points -- what the oracle's frames are."""
    files = sorted(glob.glob(os.path.join(RECORDED, "*", "frame_*.pkl")))[::97]
    assert files
    for f in files:
        with open(f, "rb") as fh:
            d = _ArraysOnly(fh).load()
        _invariants({"points": d["points"], "colors": d["colors"]}, 30000)
        # colours are means of k/255 values over a voxel's points: 255 * colour * count is whole for a small count
        c = d["colors"][:200] * 255.0
        counts = np.arange(1, 129)[None, None, :]
        frac = np.abs(c[:, :, None] * counts - np.round(c[:, :, None] * counts))
        assert (frac.min(axis=2) < 1e-6).all()
