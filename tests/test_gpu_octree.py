"""GPU parity: the OCT2 coordinate coder (pcc_octree_encode / pcc_octree_decode) vs the numpy
restatement in oracle/octree_ref.py -- byte-exact streams, bit-exact keys (integer work)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import octree_ref as O  # noqa: E402
from oracle import sparse_ref as S  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _frames_to_keys(frames, dev, log2_stride):
    """frames: list of int [n,3] arrays -> (sorted keys, device frame offsets, host counts)."""
    from b200pcc import ops
    cs = [np.concatenate([np.full((len(p), 1), b, np.int32), np.asarray(p, np.int32).reshape(-1, 3)], axis=1)
          for b, p in enumerate(frames)]
    C = np.concatenate(cs) if cs else np.zeros((0, 4), np.int32)
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C).to(dev)), want_src=False)
    off = ops.frame_offsets(keys, len(frames))
    counts = np.diff(off.cpu().numpy()).tolist()
    return keys, off, counts


def _encode(frames, dev, log2_stride=3):
    from b200pcc import ops
    from b200pcc.model import fetch_strings
    keys, off, counts = _frames_to_keys(frames, dev, log2_stride)
    pending = ops.octree_encode(keys, off, counts, log2_stride)
    return fetch_strings(pending), keys, counts


def _decode(streams, dev, log2_stride=3):
    from b200pcc import geometry, ops
    blob, offs, meta32, counts = geometry.pack_streams(streams)
    nf = len(streams)
    t = lambda a: torch.from_numpy(a).to(dev)
    m = t(meta32)
    keys, status = ops.octree_decode(t(blob), t(offs), m[:nf], m[nf:], nf, int(sum(counts)), log2_stride)
    return keys, int(status.item())


def _surface(seed, n, stride=8, offset=(0, 0, 0)):
    from b200pcc import synth
    p = synth.small_frame(seed=seed, n_target=n, offset=offset)["points"].astype(np.int64)
    return np.unique(np.floor_divide(p, stride) * stride, axis=0)


def test_streams_byte_exact_and_roundtrip(dev):
    frames = [_surface(3, 60000), _surface(4, 9000, offset=(5, -3, 2)), np.zeros((0, 3), np.int64),
              np.array([[-16, 8, 0]]), np.array([[-8, -8, -8], [0, 0, 0]])]
    streams, keys, counts = _encode(frames, dev)
    assert counts == [len(f) for f in frames]
    for f, s in zip(frames, streams):
        assert s == O.encode_points(f, 3)
    dk, status = _decode(streams, dev)
    assert status == 0 and torch.equal(dk, keys)
    # the decoder also expands streams written by the oracle
    dk2, status = _decode([O.encode_points(f, 3) for f in frames], dev)
    assert status == 0 and torch.equal(dk2, keys)


@pytest.mark.parametrize("log2_stride", [0, 1, 3, 5])
def test_strides_and_negative_coordinates(dev, log2_stride):
    rng = np.random.default_rng(log2_stride)
    st = 1 << log2_stride
    p = np.unique(rng.integers(-300, 300, (5000, 3)) * st, axis=0)
    p = p[(np.abs(p) < 32768 - st).all(axis=1)]
    streams, keys, _ = _encode([p], dev, log2_stride)
    assert streams[0] == O.encode_points(p, log2_stride)
    dk, status = _decode(streams, dev, log2_stride)
    assert status == 0 and torch.equal(dk, keys)
    from b200pcc import ops
    C = ops.coords_from_keys(dk).cpu().numpy()
    assert sorted(map(tuple, C[:, 1:])) == sorted(map(tuple, p))


def test_large_frame_is_sorted_and_lossless(dev):
    """cfg-2 scale property check (no oracle): ascending keys out, identical to the keys in."""
    from b200pcc import synth
    f = synth.frame_8ivfb(seed=1)
    p = f["points"].astype(np.int64)
    streams, keys, counts = _encode([p, np.floor_divide(p, 8) * 8], dev, 0)
    dk, status = _decode(streams, dev, 0)
    assert status == 0 and torch.equal(dk, keys)
    k = dk.cpu().numpy().view(np.uint64)
    assert np.all(k[1:] > k[:-1])
    assert len(streams[0]) < 2 * counts[0]


def test_corrupt_streams_are_reported(dev):
    from b200pcc import geometry
    good = O.encode_points(_surface(5, 20000), 3)
    n = int(np.frombuffer(good[4:8], np.uint32)[0])
    depth = int(np.frombuffer(good[8:12], np.uint32)[0])
    hdr = 20 + 4 * depth
    b = good[hdr + 3]
    other = 0xFF if bin(b).count("1") < 8 else 0x01          # a different child count
    zero_byte = good[:hdr + 3] + b"\x00" + good[hdr + 4:]
    changed = good[:hdr + 3] + bytes([other]) + good[hdr + 4:]
    wrong_n = good[:4] + np.uint32(n).tobytes() + np.uint32(depth + 1).tobytes() + good[12:]
    for bad, bit in ((good[: len(good) - 7], 1), (wrong_n, 1), (zero_byte, 2), (changed, 2)):
        keys, status = _decode([bad], dev)
        assert status & bit, (status, bit)
        assert keys.shape[0] == n
    with pytest.raises(ValueError):
        geometry.pack_streams([b"OCTX" + good[4:]])          # the host checks the magic before staging


def test_pipeline_geometry_streams_match_oracle(dev):
    """The streams inside a real bitstream are the oracle's streams of the latent coordinates."""
    from b200pcc import CompressionPipeline, container, synth
    frames = [synth.small_frame(seed=11, n_target=30000), synth.small_frame(seed=12, n_target=12000, offset=(-40, 7, 3))]
    enc = CompressionPipeline([[1.0, 1.0]], device=dev)
    c, _ = enc.compress({"frames": [dict(f) for f in frames]})
    _, _, y_shape, _, pts, _, _ = container.read_bitstream_batched(c[1])
    tot = 0
    for f, s in zip(frames, pts):
        p8 = np.unique(np.floor_divide(f["points"].astype(np.int64), 8) * 8, axis=0)
        assert s == O.encode_points(p8, 3)
        tot += len(p8)
    assert tot == y_shape
