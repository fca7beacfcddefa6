"""GPU checks at BASELINE.json's full configurations through size-independent properties (the oracle
is too slow there): cfg-2 (8iVFB-shaped 800k frame), cfg-3 (ZED-shaped 300k frames, negative
coordinates, streamed), cfg-5 (quality ladder batched per frame), and the 3-caller thread contract of
the encoder / decoder containers (sender/encoder/encoder.py:50,75; receiver/decoder/decoder.py:47)."""
import copy
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def model(dev):
    from b200pcc.pipeline import load_model
    return load_model(dev)


def _roundtrip_properties(enc, dec, frames):
    """encode -> decode: exact point counts per frame (top-k keeps k), unique in-range points, colours
    in [0,1], encoder/decoder symbol identity, geometry streams reproduce the stride-8 coordinates."""
    from b200pcc import container
    enc.keep_debug = dec.keep_debug = True
    comp, side = enc.compress({"frames": copy.deepcopy(frames)})
    out, _ = dec.decompress(comp[1])
    assert len(out) == len(frames)
    for f_in, f_out in zip(frames, out):
        assert f_out["points"].shape == (len(f_in["points"]), 3)
        assert f_out["colors"].shape == (len(f_in["points"]), 3)
        assert len(np.unique(f_out["points"], axis=0)) == len(f_out["points"])
        assert 0.0 <= f_out["colors"].min() and f_out["colors"].max() <= 1.0
        # every decoded point lies inside an occupied stride-8 cell of the input (the latent coordinates
        # are coded losslessly and the synthesis only refines inside them)
        cells_in = set(map(tuple, np.unique(np.floor_divide(f_in["points"].astype(np.int64), 8), axis=0)))
        cells_out = set(map(tuple, np.unique(np.floor_divide(f_out["points"].astype(np.int64), 8), axis=0)))
        assert cells_out <= cells_in
    assert np.array_equal(dec.debug["y_sym"].cpu().numpy(), enc.debug["y_sym"].cpu().numpy()[0])
    assert np.array_equal(dec.debug["y_idx"].cpu().numpy(), enc.debug["y_idx"].cpu().numpy()[0])
    y_shape = container.read_bitstream_batched(comp[1])[2]
    assert y_shape == sum(len(np.unique(np.floor_divide(f["points"].astype(np.int64), 8), axis=0)) for f in frames)
    assert side["gop_info"]["num_points"] == sum(len(f["points"]) for f in frames)
    return comp, out


def test_cfg2_full_frame_properties(model, dev):
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frame = synth.frame_8ivfb(seed=1)
    assert abs(len(frame["points"]) - 800_000) < 8_000
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    comp, out = _roundtrip_properties(enc, dec, [frame])
    # idempotence of the coded representation: decoding twice, and re-encoding the same frame, are bit-stable
    out2, _ = dec.decompress(comp[1])
    assert np.array_equal(out[0]["points"], out2[0]["points"]) and np.array_equal(out[0]["colors"], out2[0]["colors"])
    comp2, _ = enc.compress({"frames": [copy.deepcopy(frame)]})
    assert comp2[1] == comp[1]


def test_cfg3_zed_stream_negative_coordinates(model, dev):
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frames = [synth.frame_zed(seed=2, t=0.1 * i) for i in range(3)]
    assert min(f["points"].min() for f in frames) < 0
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    # streamed one frame per call (the 30 fps latency path) and as one 3-frame GOP: same per-frame output
    singles = []
    for f in frames:
        _, out = _roundtrip_properties(enc, dec, [f])
        singles.append(out[0])
    _, gop_out = _roundtrip_properties(enc, dec, frames)
    for a, b in zip(singles, gop_out):
        assert np.array_equal(a["points"], b["points"])
        np.testing.assert_array_equal(a["colors"], b["colors"])


def test_cfg5_quality_ladder_batched_equals_single(model, dev):
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frame = synth.small_frame(seed=31, n_target=60_000)
    grid = [[a, b] for a in (0.0, 0.25, 0.5, 0.75, 1.0) for b in (0.0, 0.25, 0.5, 0.75, 1.0)]    # Q = 25
    enc = CompressionPipeline(grid, device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    comp, side = enc.compress({"frames": [copy.deepcopy(frame)]})
    assert sorted(comp.keys()) == list(range(26)) and len(side["gop_info"]["bpp"]) == 26
    sizes = []
    for qi, q in enumerate(grid):
        single = CompressionPipeline([q], device=dev, model=model)
        c1, _ = single.compress({"frames": [copy.deepcopy(frame)]})
        assert c1[1] == comp[qi + 1], q                 # the batched Gaussian stage changes nothing
        sizes.append(len(comp[qi + 1]))
    out, _ = dec.decompress(comp[25])
    assert out[0]["points"].shape == (len(frame["points"]), 3)
    assert len(set(sizes)) > 1                          # the gain network does move the rate


def test_three_concurrent_callers(model, dev):
    """Encoder.compress_batch / Decoder.decode_and_send run on 3-worker pools over ONE pipeline object."""
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frames = [synth.small_frame(seed=40 + i, n_target=12_000 + 3000 * i) for i in range(6)]
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    serial = []
    for f in frames:
        c, _ = enc.compress({"frames": [copy.deepcopy(f)]})
        o, _ = dec.decompress(c[1])
        serial.append((c[1], o[0]["points"].copy(), o[0]["colors"].copy()))
    results = [None] * len(frames)
    errors = []

    def work(i):
        try:
            c, _ = enc.compress({"frames": [copy.deepcopy(frames[i])]})
            o, _ = dec.decompress(c[1])
            results[i] = (c[1], o[0]["points"].copy(), o[0]["colors"].copy())
        except Exception as e:   # noqa: BLE001
            errors.append(e)
    for base in (0, 3):
        ts = [threading.Thread(target=work, args=(base + j,)) for j in range(3)]
        [t.start() for t in ts]
        [t.join() for t in ts]
    assert not errors, errors
    for a, b in zip(serial, results):
        assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_two_gops_in_flight_on_their_own_streams(model, dev):
    """bench.py's throughput arms: two caller threads, each with its OWN pipeline pair and CUDA stream, code different
    frames at the same time on one GPU (the kernels of the two GOPs interleave).  Every result must equal the serial
    one bit for bit: bitstream, points, colours."""
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    frames = [synth.small_frame(seed=60 + i, n_target=30_000 + 9000 * (i % 3)) for i in range(8)]
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    serial = []
    for f in frames:
        c, _ = enc.compress({"frames": [copy.deepcopy(f)]})
        o, _ = dec.decompress(c[1])
        serial.append((c[1], o[0]["points"].copy(), o[0]["colors"].copy()))
    W = 2
    pairs = [(CompressionPipeline([[1.0, 1.0]], device=dev, model=model), DecompressionPipeline(device=dev, model=model))
             for _ in range(W)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(W)]
    results = [None] * len(frames)
    errors = []

    def work(w):
        try:
            with torch.cuda.stream(streams[w]):
                for rep in range(2):
                    for i in range(w, len(frames), W):
                        c, _ = pairs[w][0].compress({"frames": [copy.deepcopy(frames[i])]})
                        o, _ = pairs[w][1].decompress(c[1])
                        results[i] = (c[1], o[0]["points"].copy(), o[0]["colors"].copy())
        except Exception as e:   # noqa: BLE001
            errors.append(e)
    ts = [threading.Thread(target=work, args=(w,)) for w in range(W)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    torch.cuda.synchronize()
    assert not errors, errors
    for a, b in zip(serial, results):
        assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
