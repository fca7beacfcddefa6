"""GPU parity: pcc_capture_voxelize / FrameVoxelizer (csrc/capture.cu) against the capture oracle
(oracle/capture_ref.py, restating sender/capturer/capturer.py:84-120).  Bit-exact: int16 points, float64
colours compared with array_equal."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import capture_ref as R  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "capture_small.npz")


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _vox(dev, vs, clip, mp):
    from b200pcc.capture import FrameVoxelizer
    return FrameVoxelizer({"voxel_size": vs, "depth_clip": clip, "max_points": mp}, device=dev)


def _same(got, ref):
    assert got["points"].dtype == np.int16 and got["colors"].dtype == np.float64
    assert got["points"].shape == ref["points"].shape, (got["points"].shape, ref["points"].shape)
    assert np.array_equal(got["points"], ref["points"])
    assert np.array_equal(got["colors"], ref["colors"])


def test_golden_fixture(dev):
    g = np.load(GOLDEN)
    a = _vox(dev, 0.005, 1.4, None).capture(g["measure"])
    b = _vox(dev, 0.01, 3.0, 700).capture(g["measure"])
    _same(a, {"points": g["a_points"], "colors": g["a_colors"]})
    _same(b, {"points": g["b_points"], "colors": g["b_colors"]})


@pytest.mark.parametrize("h,w,seed,vs,clip,mp", [
    (480, 640, 0, 0.005, 1.4, 30000),          # shared/config.yaml:5-7 on a ZED-sized measure
    (480, 640, 1, 0.005, 3.0, 30000),          # more rows than max_points: the z cut is exercised
    (480, 640, 2, 0.01, 3.0, None),
    (120, 160, 3, 0.02, 2.0, 64),
    (97, 131, 4, 0.004, 1.0, 1),               # ragged size, a single row kept
])
def test_matches_oracle(dev, h, w, seed, vs, clip, mp):
    m = R.synth_measure(h, w, seed=seed)
    ref = R.capture_frame(m, vs, clip, mp)
    assert len(ref["points"]) > 0
    _same(_vox(dev, vs, clip, mp).capture(m), ref)


def test_long_runs_and_z_ties(dev):
    """Near-field content: hundreds of pixels per voxel (the running sums are long) and few distinct z
    values, so the max_points threshold falls inside a large group of tied rows."""
    rng = np.random.default_rng(7)
    n = 200000
    m = np.empty((n, 4), np.float32)
    m[:, 0] = rng.uniform(-0.2, 0.2, n)
    m[:, 1] = rng.uniform(-0.15, 0.15, n)
    m[:, 2] = -0.30 - 0.02 * rng.integers(0, 3, n) + rng.normal(0, 0.0004, n)
    m[:, 3] = rng.integers(0, 1 << 24, n).astype(np.uint32).view(np.float32)
    for mp in (None, 1500, 4000):
        ref = R.capture_frame(m, 0.01, 1.4, mp)
        _same(_vox(dev, 0.01, 1.4, mp).capture(m), ref)


def test_edge_cases(dev):
    v = _vox(dev, 0.01, 1.4, 100)
    out = v.capture(np.zeros((0, 4), np.float32))
    assert out["points"].shape == (0, 3) and out["colors"].shape == (0, 3)
    out = v.capture(np.full((1000, 4), np.nan, np.float32))
    assert out["points"].shape == (0, 3)
    # the reference's default depth_clip (-1.5) keeps no point (capturer.py:23, 94)
    from b200pcc.capture import FrameVoxelizer
    assert FrameVoxelizer({}, device=dev).capture(R.synth_measure(60, 80))["points"].shape == (0, 3)
    # one valid pixel among invalid ones
    one = np.full((33, 4), np.inf, np.float32)
    one[17] = [0.1, -0.2, -0.5, np.array([0x00804020], np.uint32).view(np.float32)[0]]
    _same(v.capture(one), R.capture_frame(one, 0.01, 1.4, 100))
    # identical pixels collapse into one row whose colour is the common colour
    same = np.tile(one[17], (500, 1))
    _same(v.capture(same), R.capture_frame(same, 0.01, 1.4, 100))


def test_out_of_range_is_reported_not_wrapped(dev):
    from b200pcc._lib import PccError
    far = np.zeros((4, 4), np.float32)
    far[:, 2] = [-1.0, -2.0, -3.0, -400.0]
    with pytest.raises(PccError):                 # 400 m / 1 mm = 400 000 > int16
        _vox(dev, 0.001, 1000.0, None).capture(far)
    with pytest.raises(PccError):                 # voxel index beyond 21 bits
        _vox(dev, 0.0001, 1000.0, None).capture(far)


def test_device_measure_and_tensor_input(dev):
    m = R.synth_measure(240, 320, seed=9)
    ref = R.capture_frame(m, 0.005, 1.4, 30000)
    v = _vox(dev, 0.005, 1.4, 30000)
    _same(v.capture(torch.from_numpy(m).to(dev)), ref)          # [h, w, 4] tensor already in HBM
    with pytest.raises(Exception):
        v.capture(torch.from_numpy(m.astype(np.float64)).to(dev))


def test_captured_frame_enters_the_codec_resident(dev):
    """capture_staged -> compress(gop, staged=...) codes exactly what the host route (capture -> numpy frame
    -> compress) codes: same bitstream, and the decoder returns a frame of as many points."""
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline, load_model
    model = load_model(dev)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    m = R.synth_measure(480, 640, seed=12)
    v = _vox(dev, 0.005, 1.4, 30000)
    host_frame = v.capture(m)
    host_frame.pop("timestamp")
    via_host, _ = enc.compress({"frames": [dict(host_frame)]}, staged=enc.stage_frames([host_frame]))
    frame, staged = v.capture_staged(m)
    resident, side = enc.compress({"frames": [frame]}, staged=staged)
    assert resident[1] == via_host[1]
    assert side["gop_info"]["num_points"] == len(host_frame["points"])
    frames, _ = dec.decompress(resident[1])
    got = np.asarray(frames[0]["points"])
    assert got.shape == host_frame["points"].shape      # top-k keeps exactly the coded number of points
    assert len(np.unique(got, axis=0)) == len(got)
