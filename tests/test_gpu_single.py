"""GPU: (1) the per-frame "Single" codec variant against the batched one, frame by frame; (2) INTEGRATION route B -- the
reference's own call sequence replayed against b200pcc.ColorModel / b200pcc.SparseTensor through nothing but the
model-level surface SURVEY 8b lists -- against route A (the drop-in pipelines): same bytes, same reconstruction.

Reference call sequences: sender/encoder/codec_single.py:47-111, 114-243, 269-337; receiver/decoder/codec_single.py:44-136;
sender/encoder/codec_pipeline.py:239-267, 294-317, 397-437; receiver/decoder/codec_parallel.py:291-318, 382-419, 469."""
import copy
import pickle

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


@pytest.fixture(scope="module")
def frames():
    from b200pcc import synth
    return [synth.small_frame(seed=71, n_target=14000), synth.small_frame(seed=72, n_target=5000, offset=(4, -6, 3)),
            synth.small_frame(seed=73, n_target=9000, offset=(-20, 2, 11))]


def _gop(frames):
    return {"frames": copy.deepcopy(frames), "timestamps": {"capture": 1.5}}


def test_single_variant_equals_batched_frame_by_frame(model, dev, frames):
    from b200pcc import (CompressionPipeline, DecompressionPipeline, CompressionPipelineSingle, DecompressionPipelineSingle,
                         container)
    settings = [[1.0, 0.0], [1.0, 1.0]]
    enc_b, dec_b = CompressionPipeline(settings, device=dev, model=model), DecompressionPipeline(device=dev, model=model)
    enc_s, dec_s = CompressionPipelineSingle(settings, device=dev, model=model), DecompressionPipelineSingle(device=dev, model=model)
    enc_b.keep_debug = enc_s.keep_debug = dec_s.keep_debug = True
    comp_b, side_b = enc_b.compress(_gop(frames))
    comp_s, side_s = enc_s.compress(_gop(frames))
    # key 0: the pickled input (codec_single.py:56); keys 1..Q: per-frame containers
    raw = pickle.loads(comp_s[0])
    assert np.array_equal(raw["frames"][1]["points"], frames[1]["points"])
    assert sorted(comp_s.keys()) == [0, 1, 2] and side_s["timestamps"]["capture"] == 1.5
    assert side_s["gop_info"]["num_points"] == side_b["gop_info"]["num_points"]
    assert len(side_s["enc_time_measurements"]["bitstream_writing"]) == 2
    # the symbols of frame b are the batched symbols of frame b's columns
    y_off = np.concatenate([[0], np.cumsum([s.shape[2] for s in enc_s.debug["y_sym_frames"]])])
    sym_b = enc_b.debug["y_sym"].cpu().numpy()
    for b, s in enumerate(enc_s.debug["y_sym_frames"]):
        assert np.array_equal(s.cpu().numpy(), sym_b[:, :, y_off[b]:y_off[b + 1]])
    for qi in range(2):
        ys, zs, ysh, zsh, pts, ks, q = container.read_bitstream(comp_s[qi + 1])
        yb, zb, ysh_b, zsh_b, pts_b, ks_b, q_b = container.read_bitstream_batched(comp_b[qi + 1])
        assert len(ys) == len(zs) == len(pts) == 3 and sum(ysh) == ysh_b and sum(zsh) == zsh_b
        assert pts == [bytes(p) for p in pts_b] and ks == ks_b and tuple(q) == tuple(q_b)
        assert all(s[:4] == b"PCL1" for s in ys + zs)
        out_s, info = dec_s.decompress(comp_s[qi + 1])
        out_b, _ = dec_b.decompress(comp_b[qi + 1])
        assert list(info["time_measurements"].keys()) == ["bitstream_reading", "geometry_decompression", "factorized_model",
                                                          "hyper_synthesis", "guassian_model", "synthesis_transform"]
        assert len(out_s) == len(out_b) == 3
        for a, b_ in zip(out_s, out_b):
            assert np.array_equal(a["points"], b_["points"]) and np.array_equal(a["colors"], b_["colors"])
    # a list of frame dicts is accepted like a GOP dict (codec_single.py:114-131 iterates the batch itself)
    comp_l, _ = enc_s.compress(copy.deepcopy(frames))
    assert comp_l[2] == comp_s[2]
    # single-frame GOP and a damaged stream
    one, _ = enc_s.compress({"frames": [copy.deepcopy(frames[2])]})
    out1, _ = dec_s.decompress(one[1])
    assert out1[0]["points"].shape == (len(frames[2]["points"]), 3)
    from b200pcc._lib import PccError
    with pytest.raises((ValueError, PccError)):
        dec_s.decompress(one[1][:-7])


def test_route_b_model_surface_equals_route_a(model, dev, frames):
    """The reference's call sequence against the model objects only: SparseTensor(coordinates=float, features=,
    tensor_stride=, device=), sort_tensor, features_at_coordinates(C.float()), scale_nn(q) + eps, build_indexes,
    gaussian_conditional.compress(x * scale, idx, means=) / .decompress, lower_bound_scale, get_offsets,
    entropy_bottleneck.compress / .decompress, g_s.down_conv x 2, g_s(y_hat, k=ks)."""
    from b200pcc import CompressionPipeline, DecompressionPipeline, SparseTensor, sort_tensor, sort_points, container
    em = model.entropy_model
    qset = [1.0, 1.0]
    # ---------------- route A
    enc, dec = CompressionPipeline([qset], device=dev, model=model), DecompressionPipeline(device=dev, model=model)
    dec.keep_debug = True
    comp, _ = enc.compress(_gop(frames))
    out_a, _ = dec.decompress(comp[1])
    y_a, z_a, ysh, zsh, pts_streams, ks_a, _ = container.read_bitstream_batched(comp[1])
    # ---------------- route B, encoder (codec_pipeline.py:239-267, 273, 287, 294-317, 354, 397-437)
    points = np.concatenate([np.concatenate([np.full((len(f["points"]), 1), b), f["points"]], axis=1)
                             for b, f in enumerate(frames)]).astype(np.float32)
    colors = np.concatenate([np.concatenate([np.ones((len(f["colors"]), 1)), f["colors"]], axis=1) for f in frames])
    x = SparseTensor(coordinates=torch.from_numpy(points).to(dev), features=torch.from_numpy(colors).float().to(dev), device=dev)
    y, k = model.g_a(x)
    assert k == ks_a
    y = sort_tensor(y)
    assert y.tensor_stride == [8, 8, 8] and y.F.shape == (ysh, model.config["c_y"])
    z = em.h_a(y)
    z = sort_tensor(z)
    z_feats = z.F.t().unsqueeze(0)
    z_strings = em.entropy_bottleneck.compress(z_feats)
    assert bytes(z_strings[0]) == bytes(z_a[0])                                        # same z string as route A
    z_hat_feats = em.entropy_bottleneck.decompress(z_strings, [z.F.shape[0]])
    z_hat = SparseTensor(coordinates=z.C.clone(), features=z_hat_feats[0].t(), tensor_stride=32, device=dev)
    gaussian_params = em.h_s(z_hat)
    gp = gaussian_params.features_at_coordinates(y.C.float())
    scales_hat, means_hat = gp.chunk(2, dim=1)
    scales_hat, means_hat = scales_hat.t().unsqueeze(0), means_hat.t().unsqueeze(0)
    q = torch.tensor([qset], dtype=torch.float, device=dev)
    scale = em.scale_nn(q) + em.eps
    scale = scale.unsqueeze(2)
    gc = em.gaussian_conditional
    indexes = gc.build_indexes(scales_hat * scale)
    y_strings = gc.compress(y.F.t().unsqueeze(0) * scale, indexes, means=means_hat * scale)
    assert bytes(y_strings[0]) == bytes(y_a[0])                                        # same y string as route A
    # ---------------- route B, decoder (codec_parallel.py:291-318, 376, 382-419, 469)
    y_points = y.C.clone()                                                             # (the decoder gets them from the coordinate streams)
    lat = SparseTensor(coordinates=y_points.clone(), features=torch.ones((y_points.shape[0], 1), device=dev), tensor_stride=8,
                       device=dev)
    lat = model.g_s.down_conv(model.g_s.down_conv(lat))
    z_points = sort_points(lat.C)
    assert torch.equal(z_points, z.C)
    zf = em.entropy_bottleneck.decompress(z_strings, [zsh])
    z_hat_d = SparseTensor(coordinates=z_points, features=zf[0].t(), tensor_stride=32, device=dev)
    gpd = em.h_s(z_hat_d).features_at_coordinates(sort_points(y_points).float())
    assert torch.equal(gpd, gp)                                                        # SURVEY 3 invariant 3: encoder == decoder
    s_hat, m_hat = gpd.chunk(2, dim=1)
    s_hat, m_hat = s_hat.t().unsqueeze(0), m_hat.t().unsqueeze(0)
    sc = (em.scale_nn(q) + em.eps).unsqueeze(2).repeat(1, 1, s_hat.shape[-1])
    rescale = torch.tensor(1.0) / sc
    idx_d = gc.build_indexes(s_hat * sc)
    assert torch.equal(idx_d, indexes)
    q_val = gc.decompress(y_strings, idx_d)
    sym_a = dec.debug["y_sym"]
    assert torch.equal(q_val[0].int(), sym_a)                                          # the symbols route A decoded
    q_abs, signs = q_val.abs(), torch.sign(q_val)
    stdev = gc.lower_bound_scale(s_hat * sc)
    q_off = (-1) * em.get_offsets(stdev, sc)
    q_off[q_abs < 0.0001] = 0
    y_hat_feat = signs * (q_abs + q_off)
    y_hat_feat = y_hat_feat * rescale + m_hat
    # route A's fused dequantiser computes the same expression: compare before anything is rounded to bf16
    ya = dec.debug["y_hat"]
    ya_sorted = sort_tensor(ya).F
    err = (ya_sorted - y_hat_feat[0].t()).abs()
    ref = y_hat_feat.abs().max().item()
    assert err.max().item() <= 2.0 ** -8 * max(ref, 1.0), (err.max().item(), ref)          # one bf16 step (route A stores bf16)
    y_hat = SparseTensor(coordinates=y_points, features=y_hat_feat[0].t(), tensor_stride=8, device=dev)
    x_hat = model.g_s(y_hat, k=ks_a)
    # reconstruction: the same points; colours equal wherever y_hat rounded to the same bf16 values (reported)
    C = x_hat.C.cpu().numpy()
    F = x_hat.F.cpu().numpy()
    same_y = bool(torch.equal(y_hat.feats_bf16(64), ya.feats_bf16(64)))
    for b, f_a in enumerate(out_a):
        m = C[:, 0] == b
        pb = C[m][:, 1:]
        assert pb.shape == f_a["points"].shape
        if same_y:
            col = np.clip(np.nan_to_num(F[m], nan=0.0) * 255.0, 0, 255) / 255
            assert np.array_equal(pb, f_a["points"]) and np.array_equal(col, f_a["colors"])
    from helpers import report
    report("route_b", y_hat_bf16_identical=same_y, y_hat_max_abs_diff=float(err.max().item()))
