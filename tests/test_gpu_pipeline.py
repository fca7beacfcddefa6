"""GPU parity: model transforms vs the oracle, and the drop-in compress()/decompress() round trip."""
import copy

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import sparse_ref as S  # noqa: E402
from oracle import entropy_ref as E  # noqa: E402
from oracle.model_ref import RefModel  # noqa: E402
from helpers import to_canonical, stack_frames  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def model(dev):
    from b200pcc.pipeline import load_model
    return load_model(dev)


@pytest.fixture(scope="module")
def frames():
    from b200pcc import synth
    return [synth.small_frame(seed=21, n_target=16000), synth.small_frame(seed=22, n_target=9000, offset=(3, 1, -2))]


def _gop(frames):
    return {"frames": copy.deepcopy(frames), "timestamps": {"capture": 0.0}}


def test_analysis_and_hyper_transforms_vs_oracle(model, dev, frames):
    from b200pcc.pipeline import CompressionPipeline
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    x, _ = enc.unpack_batch(_gop(frames))
    y, k = model.g_a(x)
    z = model.entropy_model.h_a(y)
    C, F = stack_frames(frames)
    Co, Fo = S.sparse_tensor(C.astype(np.float64), F)
    ref = RefModel(model.config, model.state_dict(), bf16=True)
    Cy, y_ref, k_ref = ref.g_a(Co, Fo, 2)
    assert k == k_ref
    Cg, yg = to_canonical(y.C.cpu().numpy(), y.F.cpu().numpy())
    assert np.array_equal(Cg, Cy)
    scale = np.abs(y_ref).mean()
    err = np.abs(yg - y_ref)
    # bf16 activations on both sides; differences come from fp32 summation order flipping a bf16 rounding
    assert err.mean() < 5e-3 * scale and err.max() < 0.15 * max(scale, 1.0), (err.mean(), err.max(), scale)
    Cz, z_ref = ref.h_a(Cy, yg)            # stage-wise: feed the product's y
    Czg, zg = to_canonical(z.C.cpu().numpy(), z.F.cpu().numpy())
    assert np.array_equal(Czg, Cz)
    np.testing.assert_allclose(zg, z_ref, rtol=2e-2, atol=2e-2 * np.abs(z_ref).mean())


def test_compress_decompress_roundtrip(model, dev, frames):
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    settings = [[1.0, 0.0], [0.0, 1.0], [1, 1]]          # shared/config.yaml:12-15
    enc = CompressionPipeline(settings, device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    enc.keep_debug = dec.keep_debug = True
    gop = _gop(frames)
    compressed, side = enc.compress(gop)
    assert sorted(compressed.keys()) == [0, 1, 2, 3]
    assert compressed[0] is not None and len(compressed[0]) == 2
    for key in ("analysis", "hyper_analysis", "factorized_model", "hyper_synthesis", "geometry_compression",
                "gaussian_model", "bitstream_writing"):
        assert key in side["enc_time_measurements"]
    assert len(side["enc_time_measurements"]["bitstream_writing"]) == 3
    n_in = sum(len(f["points"]) for f in frames)
    assert side["gop_info"]["num_points"] == n_in
    assert side["gop_info"]["bandwidth"][0] == 48 * n_in and len(side["gop_info"]["bpp"]) == 4
    assert side["timestamps"]["capture"] == 0.0 and side["timestamps"]["codec_end"] >= side["timestamps"]["codec_start"]
    enc_sym = enc.debug["y_sym"].cpu().numpy()
    enc_idx = enc.debug["y_idx"].cpu().numpy()
    for qi in range(3):
        blob = compressed[qi + 1]
        assert isinstance(blob, bytes)
        out, dside = dec.decompress(blob)
        assert list(dside["time_measurements"].keys()) == ["bitstream_reading", "geometry_decompression",
                                                           "factorized_model", "hyper_synthesis", "guassian_model",
                                                           "synthesis_transform"]
        # latents are coded losslessly and the decoder rebuilds identical indexes (SURVEY 3.2 inv. 3)
        assert np.array_equal(dec.debug["y_sym"].cpu().numpy(), enc_sym[qi])
        assert np.array_equal(dec.debug["y_idx"].cpu().numpy(), enc_idx[qi])
        assert torch.equal(dec.debug["gaussian_params"]._feats, enc.debug["gaussian_params"]._feats)
        assert len(out) == 2
        for f_in, f_out in zip(frames, out):
            assert f_out["points"].shape == (len(f_in["points"]), 3)     # top-k keeps exactly k points
            assert f_out["colors"].shape == (len(f_in["points"]), 3)
            assert f_out["colors"].min() >= 0.0 and f_out["colors"].max() <= 1.0
            pts = f_out["points"]
            assert len(np.unique(pts, axis=0)) == len(pts)
    # decoding twice gives identical output (determinism)
    out2, _ = dec.decompress(compressed[3])
    out3, _ = dec.decompress(compressed[3])
    for a, b in zip(out2, out3):
        assert np.array_equal(a["points"], b["points"]) and np.array_equal(a["colors"], b["colors"])


def test_synthesis_stagewise_vs_oracle(model, dev, frames):
    """Feed the decoder's y_hat to the oracle g_s; compare each up-stage on identical inputs."""
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    enc = CompressionPipeline([[1, 1]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    dec.keep_debug = True
    compressed, _ = enc.compress(_gop(frames))
    dec.decompress(compressed[1])
    y_hat = dec.debug["y_hat"]
    from b200pcc import container
    ks = container.read_bitstream_batched(compressed[1])[5]
    x_hat, inter = model.g_s(y_hat, k=ks, return_intermediates=True)
    from b200pcc import ops
    ref = RefModel(model.config, model.state_dict(), bf16=True)
    Cy, yh = to_canonical(y_hat.C.cpu().numpy(), y_hat.F.cpu().numpy())
    _, _, inter_ref = ref.g_s(Cy, yh, ks, return_intermediates=True)
    # level 0 runs on identical inputs: logits within tolerance, top-k identical when fed the same logits
    lv = inter[0]
    Cg = ops.coords_from_keys(lv["keys"]).cpu().numpy()
    Cg_c, lg = to_canonical(Cg, lv["logits"].cpu().numpy())
    assert np.array_equal(Cg_c, inter_ref[0]["C"])
    lr = inter_ref[0]["logits"]
    assert np.abs(lg - lr).max() < 0.05 * max(np.abs(lr).max(), 1.0)
    keep_ref = S.topk_mask(lv["logits"].cpu().numpy()[:, 0], Cg, ks[0])
    assert np.array_equal(np.nonzero(keep_ref)[0], lv["keep"].cpu().numpy())
    # geometry agreement with the fp32-ordered oracle decisions: report, require > 97 %
    got = set(map(tuple, x_hat.C.cpu().numpy()))
    C_ref, col_ref = ref.g_s(Cy, yh, ks)
    exp = set(map(tuple, C_ref))
    agree = len(got & exp) / len(exp)
    assert agree > 0.97, agree


def test_launch_programs_equal_per_layer_path(model, dev, frames):
    """pcc_run_program (whole transforms as one call) vs one Python call per layer: identical
    bitstreams and reconstructions, bit for bit."""
    from b200pcc import model as M
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    enc = CompressionPipeline([[1.0, 1.0], [0.25, 0.5]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    res = {}
    try:
        for flag in (True, False):
            M.USE_PROGRAMS = flag
            comp, _ = enc.compress(_gop(frames))
            out, _ = dec.decompress(comp[1])
            res[flag] = (comp, out)
    finally:
        M.USE_PROGRAMS = True
    assert res[True][0][1] == res[False][0][1] and res[True][0][2] == res[False][0][2]
    for a, b in zip(res[True][1], res[False][1]):
        assert np.array_equal(a["points"], b["points"]) and np.array_equal(a["colors"], b["colors"])


def test_host_staging_equals_resident_frames(model, dev):
    """compress(gop) stages host frames through pcc_stage_h2d (colours rounded double -> float -> bf16 by
    native threads on the way) while compress(gop, staged=...) reads float64 colours already in HBM
    (pcc_ingest_frame + pcc_rows_f32_to_bf16): same bitstream, for a frame large enough to be staged in
    several pieces, with duplicate points (first wins) and awkward colour values."""
    from b200pcc import synth
    from b200pcc.pipeline import CompressionPipeline
    big = synth.small_frame(seed=23, n_target=150_000)
    rng = np.random.default_rng(3)
    dup = rng.integers(0, len(big["points"]), 2000)
    big["points"] = np.concatenate([big["points"], big["points"][dup]])
    big["colors"] = np.concatenate([big["colors"], rng.random((2000, 3))])
    big["colors"][:6, 0] = [0.0, 1.0, 0.00390625 * 1.001, 1e-45, 0.3 + 2.0 ** -10, 1.0 - 2.0 ** -9]   # ties and denormals
    small = synth.small_frame(seed=24, n_target=5000)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    for frames in ([big], [small, big]):
        host, _ = enc.compress({"frames": copy.deepcopy(frames)})
        staged = enc.stage_frames(frames)
        resident, _ = enc.compress({"frames": copy.deepcopy(frames)}, staged=staged)
        assert host[1] == resident[1]


def test_stage_h2d_copy_and_convert(dev):
    """pcc_stage_h2d: several sources back to back, both modes, piece boundaries inside a source."""
    import ctypes as C
    from b200pcc._lib import lib, check
    rng = np.random.default_rng(11)
    srcs = [rng.random(n) * 2 - 0.5 for n in (1, 70_001, 0, 300_000)]
    srcs[1][:3] = [np.nan, np.inf, -0.0]
    total = sum(a.nbytes for a in srcs)
    pinned = torch.empty(total, dtype=torch.uint8, pin_memory=True)
    for convert in (0, 1):
        out = torch.zeros(total, dtype=torch.uint8, device=dev)
        ptrs = (C.c_void_p * len(srcs))(*[a.ctypes.data for a in srcs])
        sizes = (C.c_int64 * len(srcs))(*[a.nbytes for a in srcs])
        st = torch.cuda.current_stream()
        check(lib.pcc_stage_h2d(ptrs, sizes, len(srcs), pinned.data_ptr(), out.data_ptr(), 64 << 10, 3, convert, st.cuda_stream))
        st.synchronize()
        flat = np.concatenate(srcs)
        if convert:
            got = out[: total // 4].view(torch.bfloat16).cpu()
            want = torch.from_numpy(flat).to(torch.float32).to(torch.bfloat16)
            ok = torch.isnan(want) | (got.view(torch.int16) == want.view(torch.int16))
            assert bool(ok.all()) and bool(torch.isnan(got[1]))
        else:
            assert np.array_equal(out.cpu().numpy().view(np.float64), flat, equal_nan=True)


def test_duplicate_input_points_first_wins(model, dev):
    """compress() speculates that the input holds no duplicate points and checks the device count with the
    pyramid's read-back; a frame WITH duplicates must fall back and code exactly what its de-duplicated
    (first occurrence wins, ME.SparseTensor semantics) version codes -- host buffers and resident frames."""
    from b200pcc import synth
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    base = synth.small_frame(seed=25, n_target=30_000)
    rng = np.random.default_rng(4)
    dup = rng.integers(0, len(base["points"]), 500)
    frame = {"points": np.concatenate([base["points"], base["points"][dup]]),
             "colors": np.concatenate([base["colors"], rng.random((500, 3))])}
    order = rng.permutation(len(frame["points"]))
    frame = {"points": frame["points"][order], "colors": frame["colors"][order]}
    _, first = np.unique(frame["points"], axis=0, return_index=True)
    first.sort()
    dedup = {"points": frame["points"][first], "colors": frame["colors"][first]}
    assert len(dedup["points"]) == len(base["points"]) < len(frame["points"])
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    want, side_w = enc.compress({"frames": [copy.deepcopy(dedup)]})
    got, side_g = enc.compress({"frames": [copy.deepcopy(frame)]})
    assert got[1] == want[1] and side_g["gop_info"]["num_points"] == len(dedup["points"])
    resident, _ = enc.compress({"frames": [copy.deepcopy(frame)]}, staged=enc.stage_frames([frame]))
    assert resident[1] == want[1]
    out, _ = dec.decompress(got[1])
    assert out[0]["points"].shape == (len(dedup["points"]), 3)


def test_malformed_bitstreams_raise_and_leave_the_device_usable(model, dev):
    """decompress() on damaged input: truncated blobs, flipped bytes in the y / z strings and in the
    coordinate stream, inconsistent counts.  Every case either raises (ValueError / PccError) or returns
    frames of the advertised shape -- never a CUDA fault -- and a clean blob decodes identically afterwards."""
    import struct
    from b200pcc import synth, container
    from b200pcc._lib import PccError
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    frame = synth.small_frame(seed=26, n_target=12_000)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    comp, _ = enc.compress({"frames": [copy.deepcopy(frame)]})
    blob = comp[1]
    want, _ = dec.decompress(blob)
    y, z, ysh, zsh, pts, ks, q = container.read_bitstream_batched(blob)
    y0 = 36                      # App. B.1: 20-byte head + 16 bytes of shapes / lengths
    z0 = y0 + len(y[0])
    p0 = z0 + len(z[0]) + 16
    rng = np.random.default_rng(8)

    def damaged(lo, hi, count):
        b = bytearray(blob)
        for pos in rng.integers(lo, hi, count):
            b[pos] ^= 0xFF
        return bytes(b)

    cases = [blob[:10], blob[:y0 + 50], blob[:-1], blob[:p0 + 3],
             damaged(y0 + 400, z0, 1), damaged(y0 + 400, z0, 40), damaged(y0, y0 + 16, 2),      # y payload / header
             damaged(z0 + 150, p0 - 16, 3), damaged(z0, z0 + 16, 1),                             # z payload / header
             damaged(p0 + 40, len(blob), 2), damaged(p0, p0 + 12, 1),                            # octree bytes / header
             blob[:28] + struct.pack(">i", ysh + 1) + blob[32:],                                  # y shape off by one
             container.make_bitstream_batched(y[0], z[0], ysh, zsh, pts, [[2 ** 31 - 1], [0], [7]], q),   # absurd k
             container.make_bitstream_batched(y[0], z[0], ysh, zsh + 5, pts, ks, q)]             # wrong z shape
    outcomes = []
    for i, bad in enumerate(cases):
        try:
            out, _ = dec.decompress(bad)
            assert len(out) == 1 and out[0]["points"].ndim == 2 and out[0]["points"].shape[1] == 3, i
            assert out[0]["colors"].shape == out[0]["points"].shape, i
            outcomes.append("decoded")
        except (ValueError, PccError) as e:
            outcomes.append(type(e).__name__)
        torch.cuda.synchronize()          # a device fault would surface here
    assert outcomes[:4] == ["ValueError"] * 4 and "decoded" not in outcomes[11:12] + outcomes[13:], outcomes
    again, _ = dec.decompress(blob)
    assert np.array_equal(again[0]["points"], want[0]["points"]) and np.array_equal(again[0]["colors"], want[0]["colors"])


def test_playout_buffers_match_client_packing(model, dev, frames):
    """DecompressionPipeline.playout_buffers == what receiver/client/client.py:139-146 builds from the
    decoded frames: float32 (points + 100) bytes + uint8 (255 * colours) bytes, per frame."""
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    comp, _ = enc.compress(_gop(frames))
    host, _ = dec.decompress(comp[1])
    x_hat, _ = dec.decompress(comp[1], to_host=False)
    got = dec.playout_buffers(x_hat)
    assert len(got) == len(host) == len(frames)
    for buf, frame in zip(got, host):
        points = frame["points"] + 100
        colors = 255 * frame["colors"]
        want = np.array(points, dtype=np.float32).tobytes() + np.array(colors, dtype=np.uint8).tobytes()
        assert buf == want
    assert dec.playout_buffers(x_hat, num_frames=len(frames)) == got


def test_nvtx_stage_ranges_do_not_change_results(model, dev, frames, monkeypatch):
    """PCC_NVTX=1: every stage of compress() / decompress() is an NVTX range named after the reference's stage key;
    the ranges are balanced and the results are the ones without them."""
    from b200pcc import pipeline as P
    enc = P.CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = P.DecompressionPipeline(device=dev, model=model)
    c0, _ = enc.compress(_gop(frames))
    o0, _ = dec.decompress(c0[1])
    monkeypatch.setattr(P, "_NVTX", True)
    pushed, popped = [], []
    monkeypatch.setattr(torch.cuda.nvtx, "range_push", lambda name: pushed.append(name))
    monkeypatch.setattr(torch.cuda.nvtx, "range_pop", lambda: popped.append(1))
    c1, s1 = enc.compress(_gop(frames))
    o1, d1 = dec.decompress(c1[1])
    assert c1[1] == c0[1]
    for a, b in zip(o0, o1):
        assert np.array_equal(a["points"], b["points"]) and np.array_equal(a["colors"], b["colors"])
    assert pushed == [f"pcc.enc.{k}" for k in P.ENC_STAGES] + [f"pcc.dec.{k}" for k in P.DEC_STAGES]
    assert len(popped) == len(pushed)
    assert set(d1["time_measurements"]) >= set(P.DEC_STAGES)
