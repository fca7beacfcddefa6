"""GPU parity: tcgen05 sparse convolution vs the CPU oracle and vs the CUDA-core comparator."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import sparse_ref as S  # noqa: E402
from helpers import stack_frames  # noqa: E402

# bf16 operands, fp32 accumulation on both sides; only the summation order differs
RTOL, ATOL = 2e-3, 2e-3


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def cloud(dev):
    from b200pcc import synth, ops
    from b200pcc.sparse import CoordMap
    C, _ = stack_frames([synth.small_frame(seed=5, n_target=12000)])
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C).to(dev)))
    cm = CoordMap(keys, 0)
    return cm, ops.coords_from_keys(keys).cpu().numpy()


def _bf16_t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev).to(torch.bfloat16).contiguous()


def _run_umma(ops, x, nbr, n_out, w, bias, relu=False, residual=None, f32=True):
    kvol, cin, cout = w.shape
    dev = x.device
    wp = ops.pack_weights_umma(w, x.shape[1]).to(dev)
    bd = torch.from_numpy(bias).to(dev)
    ob = torch.empty((n_out, cout), dtype=torch.bfloat16, device=dev)
    of = torch.empty((n_out, cout), dtype=torch.float32, device=dev) if f32 else None
    ops.spconv_umma(x, nbr, n_out, kvol, wp, bd, cout, residual, relu, ob, of)
    torch.cuda.synchronize()
    return ob, of


@pytest.mark.parametrize("cin,cout", [(16, 32), (32, 32), (64, 64), (128, 128), (64, 256), (32, 16)])
def test_k3_conv_vs_oracle(cloud, dev, cin, cout):
    from b200pcc import ops
    cm, Cg = cloud
    n = cm.n
    rng = np.random.default_rng(cin * 1000 + cout)
    F = rng.normal(0, 1, (n, cin)).astype(np.float32)
    W = (rng.normal(0, 1, (27, cin, cout)) / np.sqrt(cin * 9)).astype(np.float32)
    b = rng.normal(0, 0.1, cout).astype(np.float32)
    R = rng.normal(0, 1, (n, cout)).astype(np.float32)
    x = _bf16_t(F, dev)
    r = _bf16_t(R, dev)
    nbr = cm.nbr3()
    ob, of = _run_umma(ops, x, nbr, n, W, b, relu=True, residual=r)
    nbr_np = nbr.cpu().numpy()
    ref = S.conv_forward(S.bf16_round(F), nbr_np, W, b, residual=S.bf16_round(R), relu=True, bf16=True)
    got = of.cpu().numpy()
    np.testing.assert_allclose(got, ref, rtol=RTOL, atol=ATOL)
    # the bf16 output is the rounding of the fp32 output
    assert torch.equal(ob, of.to(torch.bfloat16))
    # independent CUDA-core comparator
    wb = _bf16_t(W, dev)
    of2 = torch.empty_like(of)
    ops.spconv_simt(x, nbr, n, 27, wb, torch.from_numpy(b).to(dev), cout, r, True, None, of2)
    np.testing.assert_allclose(got, of2.cpu().numpy(), rtol=RTOL, atol=ATOL)


def test_conv_deterministic_and_row_order_invariant(cloud, dev):
    """SURVEY 3.2 invariant 3: same result run to run and for any row order of the input set."""
    from b200pcc import ops
    cm, Cg = cloud
    n, c = cm.n, 64
    rng = np.random.default_rng(7)
    F = rng.normal(0, 1, (n, c)).astype(np.float32)
    W = (rng.normal(0, 1, (27, c, c)) / np.sqrt(c * 9)).astype(np.float32)
    b = np.zeros(c, np.float32)
    x = _bf16_t(F, dev)
    nbr = cm.nbr3()
    _, a = _run_umma(ops, x, nbr, n, W, b)
    _, a2 = _run_umma(ops, x, nbr, n, W, b)
    assert torch.equal(a, a2)
    # permute the rows: features move with their coordinates, kernel map re-indexed accordingly
    perm = torch.from_numpy(rng.permutation(n)).to(dev)            # new row i holds old row perm[i]
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(n, device=dev)
    xp = x[perm].contiguous()
    nbr_p = nbr[:, perm].long()
    nbr_p = torch.where(nbr_p >= 0, inv[nbr_p.clamp_min(0)], nbr_p).int().contiguous()
    _, ap = _run_umma(ops, xp, nbr_p, n, W, b)
    assert torch.equal(ap, a[perm])


def test_down_conv_and_1x1(cloud, dev):
    from b200pcc import ops
    cm, Cg = cloud
    coarse = cm.downsample()
    nbr = cm.nbr_down()
    rng = np.random.default_rng(11)
    cin, cout = 32, 64
    F = rng.normal(0, 1, (cm.n, cin)).astype(np.float32)
    W = (rng.normal(0, 1, (8, cin, cout)) / np.sqrt(cin * 4)).astype(np.float32)
    b = rng.normal(0, 0.1, cout).astype(np.float32)
    x = _bf16_t(F, dev)
    _, of = _run_umma(ops, x, nbr, coarse.n, W, b, relu=True)
    ref = S.conv_forward(S.bf16_round(F), nbr.cpu().numpy(), W, b, relu=True, bf16=True)
    np.testing.assert_allclose(of.cpu().numpy(), ref, rtol=RTOL, atol=ATOL)
    # 1x1 (identity rows), tail tile not a multiple of 128
    W1 = (rng.normal(0, 1, (1, cin, 48)) / np.sqrt(cin)).astype(np.float32)
    b1 = rng.normal(0, 0.1, 48).astype(np.float32)
    _, o1 = _run_umma(ops, x, None, cm.n, W1, b1)
    ref1 = S.conv_forward(S.bf16_round(F), None, W1, b1, bf16=True)
    np.testing.assert_allclose(o1.cpu().numpy(), ref1, rtol=RTOL, atol=ATOL)


def test_generative_transpose_dense_form(cloud, dev):
    """GenUp (dense [n, 8*cout] GEMM) == oracle transposed-convolution kernel map semantics."""
    from b200pcc import ops
    from b200pcc.model import GenUp
    from b200pcc.sparse import CoordMap
    from helpers import to_canonical
    cm, Cg = cloud
    C8 = S.stride_coords(Cg, 8)
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C8).to(dev)))
    pm = CoordMap(keys, 3)
    Cp = ops.coords_from_keys(keys).cpu().numpy()
    rng = np.random.default_rng(13)
    for cin, cout in [(64, 32), (128, 64)]:
        g = GenUp("t", cin, cout)
        W = (rng.normal(0, 1, (8, cin, cout)) / np.sqrt(cin)).astype(np.float32)
        b = rng.normal(0, 0.1, cout).astype(np.float32)
        g.load({"t.kernel": W, "t.bias": b})
        g.to(dev)
        F = rng.normal(0, 1, (pm.n, cin)).astype(np.float32)
        out = g(_bf16_t(F, dev), relu=True).float().cpu().numpy()
        ch = pm.children()
        Cc = ops.coords_from_keys(ch.keys).cpu().numpy()
        # oracle: canonical order on both sides
        Cp_c, F_c = to_canonical(Cp, F)
        Co = S.generative_coords(Cp_c, 4)
        nbr = S.kernel_map(Cp_c, Co, S.region_offsets(2, 4), transposed=True)
        ref = S.bf16_round(S.conv_forward(S.bf16_round(F_c), nbr, W, b, relu=True, bf16=True))
        Cc_c, out_c = to_canonical(Cc, out)
        assert np.array_equal(Cc_c, Co)
        np.testing.assert_allclose(out_c, ref, rtol=1e-2, atol=1e-2)


def test_narrow_heads(cloud, dev):
    from b200pcc import ops
    from b200pcc.model import Conv
    cm, Cg = cloud
    rng = np.random.default_rng(17)
    n, cin = cm.n, 32
    F = rng.normal(0, 1, (n, cin)).astype(np.float32)
    x = _bf16_t(F, dev)
    nbr = cm.nbr3()
    for cout in (1, 3):
        W = (rng.normal(0, 1, (27, cin, cout)) / np.sqrt(cin * 9)).astype(np.float32)
        b = rng.normal(0, 0.1, cout).astype(np.float32)
        out = torch.empty((n, cout), dtype=torch.float32, device=dev)
        ops.spconv_simt(x, nbr, n, 27, _bf16_t(W, dev), torch.from_numpy(b).to(dev), cout, None, False, None, out)
        ref = S.conv_forward(S.bf16_round(F), nbr.cpu().numpy(), W, b, bf16=True)
        np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=RTOL, atol=ATOL)
        # product path: the same head on the tensor-core kernel, output channels zero-padded to N=16
        conv = Conv(f"head{cout}", 27, cin, cout)
        conv.load({f"head{cout}.kernel": W, f"head{cout}.bias": b})
        conv.to(dev)
        out2 = conv(x, nbr, n)
        assert out2.shape == (n, cout) and out2.dtype == torch.float32
        np.testing.assert_allclose(out2.cpu().numpy(), ref, rtol=RTOL, atol=ATOL)


def test_persistent_kernel_many_tiles_and_sparse_offsets(dev):
    """More tiles than SMs (every CTA walks several tiles through both TMEM accumulators) on a cloud
    thin enough that whole kernel offsets are absent from some 128-row tiles (offset skipping)."""
    from b200pcc import ops
    from b200pcc.sparse import CoordMap
    rng = np.random.default_rng(23)
    # a single-voxel-thick plane z = 0 plus a few scattered points: 18 of 27 offsets are mostly empty
    xy = np.stack(np.meshgrid(np.arange(-200, 200), np.arange(-160, 160), indexing="ij"), -1).reshape(-1, 2)
    C = np.concatenate([np.zeros((len(xy), 1), np.int64), xy, np.zeros((len(xy), 1), np.int64)], axis=1)
    extra = np.concatenate([np.zeros((500, 1), np.int64), rng.integers(-200, 200, (500, 3))], axis=1)
    C = np.concatenate([C, extra]).astype(np.int32)
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C).to(dev)))
    cm = CoordMap(keys, 0)
    n = cm.n
    assert n > 128 * 148 * 3
    for cin, cout in ((32, 32), (64, 64)):
        F = rng.normal(0, 1, (n, cin)).astype(np.float32)
        W = (rng.normal(0, 1, (27, cin, cout)) / np.sqrt(cin * 9)).astype(np.float32)
        b = rng.normal(0, 0.1, cout).astype(np.float32)
        x = _bf16_t(F, dev)
        nbr = cm.nbr3()
        _, of = _run_umma(ops, x, nbr, n, W, b, relu=False)
        ref = S.conv_forward(S.bf16_round(F), nbr.cpu().numpy(), W, b, bf16=True)
        np.testing.assert_allclose(of.cpu().numpy(), ref, rtol=RTOL, atol=ATOL)
        _, of2 = _run_umma(ops, x, nbr, n, W, b, relu=False)
        assert torch.equal(of, of2)


def test_topk_select(cloud, dev):
    from b200pcc import ops
    cm, Cg = cloud
    rng = np.random.default_rng(19)
    C2 = np.concatenate([Cg, Cg * np.array([1, 1, 1, 1]) + np.array([1, 0, 0, 0])]).astype(np.int32)  # 2 frames
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C2).to(dev)))
    Ck = ops.coords_from_keys(keys).cpu().numpy()
    n = len(Ck)
    logits = rng.normal(0, 1, n).astype(np.float32)
    logits[::7] = 0.25          # many exact ties
    logits[5::31] = -0.0
    logits[6::31] = 0.0
    ks = [n // 5, n // 3]
    off = ops.frame_offsets(keys, 2)
    keep = ops.topk_select(torch.from_numpy(logits).to(dev), keys, off,
                           torch.tensor(ks, dtype=torch.int32, device=dev), 2).cpu().numpy()
    ref = np.nonzero(S.topk_mask(logits, Ck, ks))[0]
    assert np.array_equal(keep, ref)
    assert len(keep) == sum(ks)


@pytest.mark.parametrize("case", ["k0", "kbig", "allequal", "nan_negzero", "emptyframe", "large"])
def test_topk_select_edge_cases(dev, case):
    """Radix select vs the oracle's sort on the cases a sort handles implicitly."""
    from b200pcc import ops
    rng = np.random.default_rng(len(case))
    n0, n1 = (400_000, 300_000) if case == "large" else (3000, 2500)
    pts = np.unique(rng.integers(-60 if case != "large" else -400, 60 if case != "large" else 400, (n0 + n1, 3)), axis=0)
    nf = 3
    b = np.where(np.arange(len(pts)) < len(pts) // 2, 0, 2 if case == "emptyframe" else 1)
    C = np.concatenate([b[:, None], pts], axis=1).astype(np.int32)
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C).to(dev)))
    Ck = ops.coords_from_keys(keys).cpu().numpy()
    n = len(Ck)
    logits = rng.normal(0, 2, n).astype(np.float32)
    cnt = np.bincount(Ck[:, 0], minlength=nf)
    ks = [int(cnt[0] // 3), int(cnt[1] // 2), int(cnt[2] // 4)]
    if case == "k0":
        ks = [0, int(cnt[1]), 0]
    elif case == "kbig":
        ks = [int(cnt[0]) + 5, 10 * n, 1]
    elif case == "allequal":
        logits[:] = 1.5
    elif case == "nan_negzero":
        logits[::3] = np.nan
        logits[1::5] = -0.0
        logits[2::5] = 0.0
        ks = [int(cnt[0] * 0.8), int(cnt[1] * 0.8), 0]
    elif case == "large":
        logits = np.round(logits, 1)          # heavy ties at every threshold
    off = ops.frame_offsets(keys, nf)
    keep = ops.topk_select(torch.from_numpy(logits).to(dev), keys, off,
                           torch.tensor(ks, dtype=torch.int32, device=dev), nf).cpu().numpy()
    ref = np.nonzero(S.topk_mask(logits, Ck, ks))[0]
    assert np.array_equal(keep, ref)
    assert len(keep) == sum(min(k, int(c)) for k, c in zip(ks, cnt))


@pytest.mark.parametrize("n_target,stride_log2", [(3000, 1), (40000, 1), (40000, 2), (150000, 1)])
def test_octet_conv_vs_oracle_and_general_kernel(dev, n_target, stride_log2):
    """pcc_spconv_octet (k3 convolution on complete octets, driven by the PARENTS' kernel map) against the oracle
    on the children's own kernel map and against pcc_spconv_umma: residual layer (bf16 out), plain layer, and the
    narrow occupancy head (one fp32 column); several frames, tiles with a ragged tail, repeated launches identical."""
    from b200pcc import synth, ops
    from b200pcc.sparse import CoordMap
    frames = [synth.small_frame(seed=61, n_target=n_target), synth.small_frame(seed=62, n_target=max(n_target // 3, 500), offset=(40, 3, -7))]
    C, _ = stack_frames(frames)
    t = 1 << stride_log2
    Cp = S.stride_coords(C, t)                                             # parents at tensor stride t
    pk, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(np.ascontiguousarray(Cp)).to(dev)))
    parent = CoordMap(pk, stride_log2)
    child = parent.children()                                              # rows 8p + o at stride t / 2
    n_par, n = parent.n, child.n
    assert n == 8 * n_par
    Cc = ops.coords_from_keys(child.keys).cpu().numpy()
    rng = np.random.default_rng(n_target + stride_log2)
    F = rng.normal(0, 1, (n, 32)).astype(np.float32)
    R = rng.normal(0, 1, (n, 32)).astype(np.float32)
    W = (rng.normal(0, 1, (27, 32, 32)) / np.sqrt(32 * 14)).astype(np.float32)
    Wh = (rng.normal(0, 1, (27, 32, 1)) / np.sqrt(32 * 14)).astype(np.float32)
    b = rng.normal(0, 0.1, 32).astype(np.float32)
    bh = np.array([0.3], np.float32)
    x, r = _bf16_t(F, dev), _bf16_t(R, dev)
    nbp = parent.nbr3()
    nbc = child.nbr3()
    nbc_np = nbc.cpu().numpy()
    assert np.array_equal(nbc_np, S.kernel_map(Cc, Cc, S.region_offsets(3, t // 2)))
    wo = ops.pack_weights_octet(W, 32).to(dev)
    bd = torch.from_numpy(b).to(dev)
    # residual + ReLU layer, bf16 output
    got = torch.empty((n, 32), dtype=torch.bfloat16, device=dev)
    ops.spconv_octet(x, nbp, n_par, wo, bd, 32, r, True, got, None)
    ref = S.conv_forward(S.bf16_round(F), nbc_np, W, b, residual=S.bf16_round(R), relu=True, bf16=True)
    g = got.float().cpu().numpy()
    # bf16 output: half a bf16 step on top of the fp32 tolerance
    np.testing.assert_allclose(g, ref, rtol=RTOL + 2.0 ** -8, atol=ATOL)
    ob, of = _run_umma(ops, x, nbc, n, W, b, relu=True, residual=r)
    assert (ob != got).float().mean().item() < 0.02                        # same values up to summation order (rare bf16 flips)
    np.testing.assert_allclose(g, ob.float().cpu().numpy(), rtol=2.0 ** -7, atol=ATOL)
    # plain layer, fp32 output
    of2 = torch.empty((n, 32), dtype=torch.float32, device=dev)
    ops.spconv_octet(x, nbp, n_par, wo, bd, 32, None, False, None, of2)
    ref2 = S.conv_forward(S.bf16_round(F), nbc_np, W, b, bf16=True)
    np.testing.assert_allclose(of2.cpu().numpy(), ref2, rtol=RTOL, atol=ATOL)
    # narrow head: one fp32 column
    wh = ops.pack_weights_octet(Wh, 16).to(dev)
    bh16 = torch.zeros(16, dtype=torch.float32, device=dev)
    bh16[0] = float(bh[0])
    lg = torch.full((n, 1), float("nan"), dtype=torch.float32, device=dev)
    ops.spconv_octet(x, nbp, n_par, wh, bh16, 16, None, False, None, lg, cstore=1)
    refh = S.conv_forward(S.bf16_round(F), nbc_np, Wh, bh, bf16=True)
    np.testing.assert_allclose(lg.cpu().numpy(), refh, rtol=RTOL, atol=ATOL)
    # deterministic run to run
    again = torch.empty_like(got)
    ops.spconv_octet(x, nbp, n_par, wo, bd, 32, r, True, again, None)
    assert torch.equal(again, got)
    ops.verify_deferred(device=dev)                                        # no fault word left behind
