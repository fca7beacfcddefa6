"""GPU parity, transform level: h_s and every level of g_s against the oracle on identical inputs, the
reconstructed colours with a stated tolerance, pack_batches' NaN / clip handling, and the end-to-end
decision mismatch counts SURVEY.md 7.4.7 asks for (oracle in bf16 mode and in fp32 mode).

Reference call sites: receiver/decoder/codec_parallel.py:376 (h_s), :469 (g_s(y_hat, k=ks)), :474-502
(pack_batches); sender/encoder/codec_pipeline.py:354 (h_s on the encoder), :397-437 (symbols).

Stated tolerances (the product multiplies bf16 operands, accumulates in fp32 and stores bf16 between layers;
the oracle's bf16 mode does the same with another fp32 summation order, so a value may land on the
neighbouring bf16 — 2^-8 relative — after a layer, and such flips propagate through the following layers):

    T_FEAT  bf16 feature rows after a chain of <= 6 layers on identical inputs:
            mean |d| <= 4e-3 * mean|ref|   and   max |d| <= 0.08 * max(1, max|ref|)
    T_F32   fp32 outputs (occupancy logits, Gaussian parameters) of such a chain: the same two bounds
    T_Y     the latents y after the 17 chained layers of g_a: mean |d| <= 8e-3 * mean|ref|, max |d| <= 0.15 * max(1, max|ref|)
    T_HEAD  one fp32-output convolution on IDENTICAL bf16 inputs (colour head stage-wise): |d| <= 2e-3 + 2e-3 |ref|
    T_COL   reconstructed colours (after the [0,1] clip) of the whole synthesis, both sides computing all ~16 layers on
            their own from the same y_hat, the oracle following the product's top-k decisions so that every point is
            common: mean |d| <= 4e-3, max |d| <= 0.06
"""
import copy

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import sparse_ref as S  # noqa: E402
from oracle import entropy_ref as E  # noqa: E402
from oracle.model_ref import RefModel  # noqa: E402
from oracle.codec_ref import RefCodec  # noqa: E402
from helpers import to_canonical, stack_frames, report  # noqa: E402

T_FEAT_MEAN, T_FEAT_MAX = 4e-3, 0.08
T_COL_MEAN, T_COL_MAX = 4e-3, 0.06


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
    return [synth.small_frame(seed=31, n_target=18000), synth.small_frame(seed=32, n_target=8000, offset=(-2, 5, 1))]


@pytest.fixture(scope="module")
def fast_oracle():
    prev = S.use_fast(True)
    yield
    S.use_fast(prev)


def _chain_bounds(got, ref, what, t_mean=T_FEAT_MEAN, t_max=T_FEAT_MAX):
    err = np.abs(got - ref)
    scale_mean, scale_max = np.abs(ref).mean(), max(1.0, float(np.abs(ref).max()))
    assert err.mean() <= t_mean * scale_mean, (what, "mean", err.mean(), scale_mean)
    assert err.max() <= t_max * scale_max, (what, "max", err.max(), scale_max)
    return {"mean_err": float(err.mean()), "max_err": float(err.max()), "mean_ref": float(scale_mean),
            "max_ref": float(np.abs(ref).max())}


def _coded(model, dev, frames, settings=((1.0, 1.0),)):
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    from b200pcc import container
    enc = CompressionPipeline([list(s) for s in settings], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    enc.keep_debug = dec.keep_debug = True
    comp, _ = enc.compress({"frames": copy.deepcopy(frames), "timestamps": {}})
    out, _ = dec.decompress(comp[1])
    ks = container.read_bitstream_batched(comp[1])[5]
    return enc, dec, comp, out, ks


def test_h_s_vs_oracle(model, dev, frames):
    """Row a6: the product's z_hat through RefModel.h_s -- generated coordinates exact, the 2*Cy Gaussian
    parameters within T_F32 (two bf16 generative up-convolutions + the fp32 1x1 head)."""
    enc, dec, _, _, _ = _coded(model, dev, frames)
    z_hat, gp = enc.debug["z_hat"], enc.debug["gaussian_params"]
    Cz, zh = to_canonical(z_hat.C.cpu().numpy(), z_hat.F.cpu().numpy())
    ref = RefModel(model.config, model.state_dict(), bf16=True)
    Cp_ref, p_ref = ref.h_s(Cz, zh)
    Cp, pg = to_canonical(gp.C.cpu().numpy(), gp.F.cpu().numpy())
    assert np.array_equal(Cp, Cp_ref)
    assert gp.tensor_stride == [8, 8, 8]
    r = _chain_bounds(pg, p_ref, "gaussian params")
    # the decoder rebuilds the same tensor from the decoded z (SURVEY 3 invariant 3), so the comparison holds for it too
    assert torch.equal(dec.debug["gaussian_params"]._feats, gp._feats)
    report("h_s_vs_oracle", rows=int(len(Cp)), **r)


def _stagewise(model, dev, frames, tag):
    """Every up-stage of g_s on identical inputs (the product's kept rows of stage i-1 go to the oracle's stage i)."""
    from b200pcc import ops
    enc, dec, comp, out, ks = _coded(model, dev, frames)
    y_hat = dec.debug["y_hat"]
    x_hat, inter = model.g_s(y_hat, k=ks, return_intermediates=True)
    ref = RefModel(model.config, model.state_dict(), bf16=True)
    C_in, yh = to_canonical(y_hat.C.cpu().numpy(), y_hat.F.cpu().numpy())
    Cy0 = C_in
    F_in = ref.g_s_pre(yh)
    stride = 8
    masks = []
    for i in range(3):
        Cc, Fc, lg = ref.g_s_stage(i, C_in, F_in, stride)
        stride //= 2
        lv = inter[i]
        Cg = ops.coords_from_keys(lv["keys"]).cpu().numpy()
        Fg_m = lv["F"].float().cpu().numpy()
        lg_m = lv["logits"].cpu().numpy()
        Cg_c, Fg, lgg = to_canonical(Cg, Fg_m, lg_m)
        assert np.array_equal(Cg_c, Cc), f"candidate coordinates of level {i}"
        rf = _chain_bounds(Fg, Fc, f"level {i} features")
        rl = _chain_bounds(lgg, lg, f"level {i} logits")
        # top-k: exact on equal logits (the product's own logits through the oracle's rule) ...
        keep_rule = S.topk_mask(lg_m[:, 0], Cg, ks[i])
        keep_got = lv["keep"].cpu().numpy()
        assert np.array_equal(np.nonzero(keep_rule)[0], keep_got), f"top-k rule at level {i}"
        # ... and against the oracle's own logits: count the flipped decisions and show they are near-ties
        keep_ref = S.topk_mask(lg[:, 0], Cc, ks[i])
        got_mask_m = np.zeros(len(Cg), bool)
        got_mask_m[keep_got] = True
        _, got_mask = to_canonical(Cg, got_mask_m)
        masks.append(got_mask)
        flips = np.nonzero(keep_ref != got_mask)[0]
        margin = 0.0
        if len(flips):
            # a flipped row's oracle logit lies within the logit tolerance of the frame's oracle threshold
            for b in range(len(ks[i])):
                rows = np.nonzero(Cc[:, 0] == b)[0]
                fr = flips[np.isin(flips, rows)]
                if len(fr) == 0:
                    continue
                kept = lg[rows, 0][keep_ref[rows]]
                thr = kept.min() if len(kept) else np.inf
                margin = max(margin, float(np.abs(lg[fr, 0] - thr).max()))
            assert margin <= 2 * rl["max_err"] + 1e-6, (i, margin, rl)
        report(f"g_s_stagewise_{tag}", level=i, rows=int(len(Cc)), kept=int(keep_ref.sum()), topk_flips=int(len(flips)),
               flip_margin=margin, feat=rf, logits=rl)
        # next stage runs on the PRODUCT's kept rows
        C_in, F_in = to_canonical(Cg[keep_got], Fg_m[keep_got])
    # colour head on identical inputs (T_HEAD), before the clip
    col_ref = ref.g_s_color(C_in, F_in)
    Cx, colg = to_canonical(x_hat.C.cpu().numpy(), x_hat.F.cpu().numpy())
    assert np.array_equal(Cx, C_in)
    np.testing.assert_allclose(colg, col_ref, rtol=2e-3, atol=2e-3)
    report(f"g_s_colour_head_{tag}", rows=int(len(Cx)), max_err=float(np.abs(colg - col_ref).max()))
    # T_COL: the oracle runs the WHOLE synthesis on its own features (bf16 rounding differences accumulate through
    # ~16 layers) but with the product's top-k decisions, so both sides reconstruct the same points and every
    # reconstructed colour is compared, after pack_batches' clip (codec_parallel.py:489-491)
    C_chain, col_chain = ref.g_s(Cy0, yh, ks, keeps=masks)
    assert np.array_equal(C_chain, Cx)
    want = np.clip(np.nan_to_num(col_chain, nan=0.0) * 255.0, 0, 255) / 255
    got = np.concatenate([f["colors"] for f in out])           # decompress() output: frames in order, Morton rows
    Pg = np.concatenate([np.concatenate([np.full((len(f["points"]), 1), b, np.int32), f["points"].astype(np.int32)], axis=1)
                         for b, f in enumerate(out)])
    Pg_c, got_c = to_canonical(Pg, got)
    assert np.array_equal(Pg_c, Cx)
    d = np.abs(got_c - want)
    report(f"g_s_colours_same_decisions_{tag}", points=int(len(Cx)), colour_mean_err=float(d.mean()), colour_max_err=float(d.max()))
    assert d.mean() <= T_COL_MEAN and d.max() <= T_COL_MAX, (d.mean(), d.max())
    return enc, dec, comp, out, ks, ref


def test_g_s_all_levels_stagewise_and_colours(model, dev, frames):
    """Row a14: stage-wise parity of all three levels and the colour head, the reconstructed colours of the whole
    synthesis within T_COL (inside _stagewise); then, reported, what happens when the oracle follows its own
    top-k decisions."""
    enc, dec, comp, out, ks, ref = _stagewise(model, dev, frames, "small")
    y_hat = dec.debug["y_hat"]
    Cy, yh = to_canonical(y_hat.C.cpu().numpy(), y_hat.F.cpu().numpy())
    C_ref, col_ref = ref.g_s(Cy, yh, ks)
    for b, f_out in enumerate(out):
        m = C_ref[:, 0] == b
        P_ref = C_ref[m][:, 1:]
        c_ref = np.clip(np.nan_to_num(col_ref[m], nan=0.0) * 255.0, 0, 255) / 255          # codec_parallel.py:489-491
        assert f_out["points"].shape == P_ref.shape
        key = lambda P: (P[:, 0].astype(np.int64) + 32768 << 32) | (P[:, 1].astype(np.int64) + 32768 << 16) | (P[:, 2].astype(np.int64) + 32768)
        kg, kr = key(f_out["points"]), key(P_ref)
        common, ig, ir = np.intersect1d(kg, kr, return_indices=True)
        agree = len(common) / len(kr)
        d = np.abs(f_out["colors"][ig] - c_ref[ir])
        report("g_s_own_decisions_small", frame=b, points=int(len(kr)), geometry_agreement=agree,
               differing_points=int(len(kr) - len(common)), colour_mean_err=float(d.mean()))
        # Reported, not bounded tightly: with the oracle following its OWN decisions a flipped near-tie (2 per level
        # here) changes which rows exist around it, and with the seeded, untrained weights of the stand-in model that
        # change spreads through the receptive field of the following ~10 convolutions -- on a frame this small a
        # handful of flips reaches a fifth of the points.  The bars are the stage-wise ones above: identical
        # candidates, flips only at near-ties, colours within T_COL on identical decisions.
        assert agree >= 0.5, agree


def test_pack_batches_nan_and_clip_bit_exact(dev):
    """Row a16: pcc_unpack_output == np.clip(np.nan_to_num(c, nan=0.0) * 255.0, 0, 255) / 255
    (codec_parallel.py:489-491) and pcc_pack_visualizer == the client's packing (receiver/client/client.py:139-146),
    bit for bit, on NaN, infinities, negative, > 1, exactly 1, -0, denormal and just-off-grid colours."""
    from b200pcc import ops
    rng = np.random.default_rng(5)
    special = np.array([np.nan, -np.nan, np.inf, -np.inf, -0.2, 1.7, 1.0, 0.0, -0.0, 1e-45, -1e-45, 1.17e-38,
                        0.5, 1.0 - 2 ** -24, 1.0 + 2 ** -23, 254.5 / 255, 255.5 / 255, 0.00390625, 3.4e38, -3.4e38,
                        1 / 255, 2 / 255, 127.5 / 255, 128 / 255], dtype=np.float32)
    rand = np.concatenate([rng.uniform(-0.5, 1.5, 5000), rng.integers(0, 256, 2000) / 255.0]).astype(np.float32)
    c = np.concatenate([special, rand])
    c = np.concatenate([c, np.zeros((-len(c)) % 3, np.float32)]).reshape(-1, 3)
    n = len(c)
    pts = np.concatenate([np.zeros((n, 1)), rng.integers(-3000, 3000, (n, 3))], axis=1).astype(np.int32)
    keys = ops.keys_from_coords(torch.from_numpy(pts).to(dev))
    feats = torch.from_numpy(c).to(dev)
    p_d, c_d = ops.unpack_output(keys, feats)
    with np.errstate(over="ignore", invalid="ignore"):
        want = np.clip(np.nan_to_num(c, nan=0.0) * 255.0, 0, 255) / 255
    got = c_d.cpu().numpy()
    assert want.dtype == np.float32 and got.dtype == np.float32
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(p_d.cpu().numpy(), pts[:, 1:])
    vp, vc = ops.pack_visualizer(keys, feats, 100.0)
    points = pts[:, 1:] + 100
    colors = 255 * want
    assert vp.cpu().numpy().tobytes() == np.array(points, dtype=np.float32).tobytes()
    assert vc.cpu().numpy().tobytes() == np.array(colors, dtype=np.uint8).tobytes()


def _end_to_end_counts(model, dev, frames, tag, modes=(True, False)):
    """the whole codec on the product vs the oracle running on its own (no stage-wise feeding):
    latent coordinates / k / counts exact; symbol and top-k decision mismatch counts per oracle mode."""
    if isinstance(frames, dict):
        frames = [frames]
    enc, dec, comp, out, ks = _coded(model, dev, frames)
    y = enc.debug["y"]
    Cy_g = y.C.cpu().numpy()
    order = np.argsort(S.lex_key(Cy_g), kind="stable")
    sym_g = enc.debug["y_sym"][0].cpu().numpy()
    key = lambda P: (P[:, 0].astype(np.int64) + 32768 << 32) | (P[:, 1].astype(np.int64) + 32768 << 16) | (P[:, 2].astype(np.int64) + 32768)
    res = {}
    for bf16 in modes:
        codec = RefCodec(model.config, model.state_dict(), bf16=bf16)
        blob = codec.compress(frames, [[1.0, 1.0]])
        assert np.array_equal(Cy_g[order], blob["y_coords"]), "latent coordinates"
        assert blob["k"] == ks, "k"
        sym_r = blob["debug"]["y_sym"][0]
        assert sym_r.shape == sym_g.shape
        mism = int((sym_r != sym_g).sum())
        off_by_more = int((np.abs(sym_r.astype(np.int64) - sym_g) > 1).sum())
        yerr = np.abs(y.F.cpu().numpy()[order] - blob["debug"]["y"])
        frames_r, _ = codec.decompress(blob, 0)
        n_pts = n_common = 0
        d = []
        for f_g, f_r in zip(out, frames_r):
            assert f_r["points"].shape == f_g["points"].shape
            common, ig, ir = np.intersect1d(key(f_g["points"]), key(f_r["points"]), return_indices=True)
            d.append(np.abs(f_g["colors"][ig] - f_r["colors"][ir]).reshape(-1))
            n_pts, n_common = n_pts + len(f_r["points"]), n_common + len(common)
        d = np.concatenate(d)
        res["bf16" if bf16 else "fp32"] = r = {
            "symbols": int(sym_r.size), "symbol_mismatches": mism, "symbol_mismatch_rate": mism / sym_r.size,
            "symbols_off_by_more_than_one": off_by_more,
            "y_mean_rel_err": float(yerr.mean() / np.abs(blob["debug"]["y"]).mean()),
            "points": int(n_pts), "points_differing": int(n_pts - n_common),
            "geometry_agreement": n_common / n_pts,
            "colour_mean_err": float(d.mean()), "colour_max_err": float(d.max())}
        report(f"end_to_end_{tag}", oracle_mode="bf16" if bf16 else "fp32", **r)
    return res


def test_end_to_end_mismatch_counts_bf16_and_fp32(model, dev, fast_oracle):
    """Both oracle modes on a 40 k-point frame.  In bf16 mode the oracle rounds exactly where the product rounds,
    so the few symbol flips come from fp32 summation order only; fp32 mode adds the bf16 rounding itself."""
    from b200pcc import synth
    frame = synth.small_frame(seed=33, n_target=40_000)
    res = _end_to_end_counts(model, dev, frame, "40k")
    b, f = res["bf16"], res["fp32"]
    # Both sides code and decode THEIR OWN symbols here.  y differs by ~0.4 % (bf16 rounding flips through 17 layers),
    # so ~1 % of the symbols land on the other side of a rounding boundary (never by more than one step); with the
    # seeded, untrained weights of the stand-in model g_s is not smooth in y_hat, so those +-1 steps move ~10 % of the
    # top-k decisions.  On IDENTICAL symbols (the stage-wise tests above: what a decoder sees in deployment, where it
    # decodes the encoder's stream) the decisions differ in < 1e-4 of the rows.
    assert b["symbol_mismatch_rate"] <= 0.02 and b["symbols_off_by_more_than_one"] == 0, b
    assert f["symbol_mismatch_rate"] <= 0.03 and f["symbols_off_by_more_than_one"] == 0, f
    assert f["symbol_mismatch_rate"] >= b["symbol_mismatch_rate"]
    assert b["geometry_agreement"] >= 0.8 and f["geometry_agreement"] >= 0.8, (b, f)


def test_cfg2_full_frame_vs_oracle(model, dev, fast_oracle):
    """BASELINE.json cfg-2 at full size (798 223 points) against the oracle (C core, seconds): latent coordinates,
    k and counts exact; y within the chain bound; h_a, h_s, symbols and indexes stage-wise; all three g_s levels
    stage-wise with flip counts; then the end-to-end counts in bf16 mode."""
    from b200pcc import synth, ops
    frame = synth.frame_8ivfb(seed=1)
    res = _full_frame_vs_oracle(model, dev, frame, "cfg2_800k")
    assert res["bf16"]["symbol_mismatch_rate"] <= 0.02 and res["bf16"]["symbols_off_by_more_than_one"] == 0


def test_cfg3_zed_frame_vs_oracle(model, dev, fast_oracle):
    """One cfg-3 frame (ZED-shaped sheet, ~300 k points, negative coordinates), same checks."""
    from b200pcc import synth
    frame = synth.frame_zed(seed=2)
    res = _full_frame_vs_oracle(model, dev, frame, "cfg3_300k")
    assert res["bf16"]["symbol_mismatch_rate"] <= 0.02 and res["bf16"]["symbols_off_by_more_than_one"] == 0


def _full_frame_vs_oracle(model, dev, frames, tag):
    if isinstance(frames, dict):
        frames = [frames]
    enc, dec, comp, out, ks, ref = _stagewise(model, dev, frames, tag)
    # ---- encoder side, stage-wise
    C, F = stack_frames(frames)
    Co, Fo = S.sparse_tensor(C.astype(np.float64), F)
    Cy, y_ref, k_ref = ref.g_a(Co, Fo, len(frames))
    y = enc.debug["y"]
    Cg, yg = to_canonical(y.C.cpu().numpy(), y.F.cpu().numpy())
    assert np.array_equal(Cg, Cy) and ks == k_ref
    ry = _chain_bounds(yg, y_ref, "y", 8e-3, 0.15)
    z = enc.debug["z"]
    Cz, z_ref = ref.h_a(Cy, yg)
    Czg, zg = to_canonical(z.C.cpu().numpy(), z.F.cpu().numpy())
    assert np.array_equal(Czg, Cz)
    rz = _chain_bounds(zg, z_ref, "z")
    z_hat, gp = enc.debug["z_hat"], enc.debug["gaussian_params"]
    Czh, zh = to_canonical(z_hat.C.cpu().numpy(), z_hat.F.cpu().numpy())
    Cp_ref, p_ref = ref.h_s(Czh, zh)
    Cp, pg = to_canonical(gp.C.cpu().numpy(), gp.F.cpu().numpy())
    assert np.array_equal(Cp, Cp_ref)
    rp = _chain_bounds(pg, p_ref, "gaussian params")
    # symbols / indexes on the product's own y and parameters: bit-exact
    rows = S.lookup_rows(Cp, Cg)
    P = np.where(rows[:, None] >= 0, pg[np.maximum(rows, 0)], 0.0).astype(np.float32)
    cy = model.config["c_y"]
    s_ref, i_ref = E.gaussian_encode_symbols(yg.T, P[:, :cy].T, P[:, cy:].T, enc.debug["scale"][0].cpu().numpy())
    assert np.array_equal(enc.debug["y_sym"][0].cpu().numpy(), s_ref)
    assert np.array_equal(enc.debug["y_idx"][0].cpu().numpy().astype(np.int32), i_ref)
    report(f"full_frame_{tag}", points=int(sum(len(f["points"]) for f in frames)), latent_points=int(len(Cy)), y=ry, z=rz,
           params=rp, symbols=int(s_ref.size))
    return _end_to_end_counts(model, dev, frames, tag, modes=(True,))


def recorded_frames():
    """The reference's recorded capturer frames (tests/golden/recorded_frames.npz, made by make_recorded.py from
    evaluation/data/test_sequence/frame_0000{0,1,2}.pkl)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "recorded_frames.npz"))
    return [{"points": g[f"points_{i}"], "colors": g[f"colors_{i}"]} for i in g["frames"]]


def test_recorded_frames_coordinates_and_kernel_maps(dev):
    """on a real recorded ZED frame (19 679 points, negative coordinates) the output coordinates at
    strides 2 / 4 / 8 / 16 / 32 and the k3 kernel maps equal the oracle's exactly (sizes 19 679 -> 8 874 -> 2 831 ->
    815 -> 229 -> 63), as do the k2/s2 maps and the generative children."""
    from b200pcc import ops
    frame = recorded_frames()[0]
    C, _ = stack_frames([frame])
    keys = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C).to(dev)))[0]
    sizes, Cl = [len(C)], S.sparse_tensor(C.astype(np.float64))[0]
    for l in range(6):
        t = 1 << l
        Cg = ops.coords_from_keys(keys).cpu().numpy()
        assert np.array_equal(to_canonical(Cg)[0], Cl)
        if l <= 3:
            ref = S.kernel_map(Cg, Cg, S.region_offsets(3, t))
            got = ops.kmap_k3_blocks(keys, l, ops.blockmap_build(keys, l)).cpu().numpy()
            assert np.array_equal(got, ref), l
            if l == 0:
                assert abs((ref >= 0).sum() / len(Cg) - 7.2) < 0.1          # pairs per point measured in SURVEY 7.3
        if l == 5:
            break
        pk, parent = ops.coords_downsample(keys, l + 1)
        Cp = ops.coords_from_keys(pk).cpu().numpy()
        refd = S.kernel_map(Cg, Cp, S.region_offsets(2, t))          # oracle on the product's row orders
        assert np.array_equal(ops.kmap_down2(keys, parent, l, pk.shape[0]).cpu().numpy(), refd), l
        ch = ops.coords_from_keys(ops.coords_children(pk, l)).cpu().numpy()
        assert np.array_equal(to_canonical(ch)[0], S.generative_coords(Cp, t))
        keys, Cl = pk, S.stride_coords(Cl, 2 * t)
        sizes.append(len(Cl))
    assert sizes == [19679, 8874, 2831, 815, 229, 63], sizes


def test_recorded_gop_vs_oracle(model, dev, fast_oracle):
    """A GOP of three recorded frames through every stage-wise and end-to-end check of the synthetic configurations."""
    res = _full_frame_vs_oracle(model, dev, recorded_frames(), "recorded_3x19k")
    assert res["bf16"]["symbol_mismatch_rate"] <= 0.02 and res["bf16"]["symbols_off_by_more_than_one"] == 0
