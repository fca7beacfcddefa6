"""CPU suite: the oracle against properties and committed golden vectors, the host logic, and that the
C-ABI library loads and exports every declared symbol (no kernel is launched here)."""
import json
import os
import re

import numpy as np
import pytest

from oracle import sparse_ref as S
from oracle import entropy_ref as E
from oracle import rans_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_library_exports_every_declared_symbol():
    from b200pcc import _lib
    hdr = open(os.path.join(ROOT, "include", "pcc_b200.h")).read()
    declared = set(re.findall(r"\b(pcc_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(_lib.lib, name)
    assert _lib.lib.pcc_version() >= 100


def test_abi_host_functions_without_gpu():
    from b200pcc import ops, _lib
    pmf = np.array([0.1, 0.2, 0.4, 0.2, 0.1, 1e-9], np.float32)
    got = ops.pmf_to_quantized_cdf(pmf)
    assert np.array_equal(got, E.pmf_to_quantized_cdf(pmf))
    assert got[-1] == 65536 and np.all(np.diff(got) > 0)
    with pytest.raises(_lib.PccError):
        ops.pmf_to_quantized_cdf(np.array([np.nan, 1.0], np.float32))
    # weight packing: UMMA K-major core-matrix layout  byte = kc*cout*16 + co*16 + e*2
    rng = np.random.default_rng(0)
    w = rng.normal(0, 1, (2, 20, 16)).astype(np.float32)
    packed = ops.pack_weights_umma(w, 32).numpy().view(np.uint16).reshape(2, 4, 16, 8)
    ref = S.bf16_round(np.concatenate([w, np.zeros((2, 12, 16), np.float32)], axis=1))
    ref16 = (ref.view(np.uint32) >> 16).astype(np.uint16)            # [j, ci, co]
    assert np.array_equal(packed, ref16.reshape(2, 4, 8, 16).transpose(0, 1, 3, 2))


def test_product_has_no_oracle_import_and_no_cpu_path():
    pkg = os.path.join(ROOT, "demo-learned-point-cloud-compression_b200", "b200pcc")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("CPU oracle", ""), fn


def test_canonical_sort_equals_reference_weights():
    rng = np.random.default_rng(0)
    C = np.concatenate([rng.integers(0, 4, (5000, 1)), rng.integers(-400, 400, (5000, 3))], axis=1)
    C = np.unique(C, axis=0)
    rng.shuffle(C)
    assert np.array_equal(np.argsort(S.lex_key(C), kind="stable"), S.sort_order_reference(C))


def test_sparse_tensor_floor_dedup_first_wins():
    C = np.array([[0, 1.5, -0.5, 2.0], [0, 1.2, -0.9, 2.9], [0, -3.0, 0.0, 0.0]])
    F = np.array([[1.0], [2.0], [3.0]], np.float32)
    Co, Fo = S.sparse_tensor(C, F)
    assert Co.tolist() == [[0, -3, 0, 0], [0, 1, -1, 2]] and Fo.ravel().tolist() == [3.0, 1.0]


def test_stride_and_kernel_map_conventions():
    C = np.array([[0, -1, 0, 3], [0, -2, 1, 2], [0, 2, 2, 2], [1, 0, 0, 0]], np.int32)
    assert S.stride_coords(C, 2).tolist() == [[0, -2, 0, 2], [0, 2, 2, 2], [1, 0, 0, 0]]
    o3 = S.region_offsets(3, 2)
    assert o3[0].tolist() == [-2, -2, -2] and o3[1].tolist() == [0, -2, -2] and o3[13].tolist() == [0, 0, 0]
    o2 = S.region_offsets(2, 4)
    assert o2.tolist()[:3] == [[0, 0, 0], [4, 0, 0], [0, 4, 0]]
    # forward/transposed maps are mirror images
    Cs = S.sparse_tensor(np.random.default_rng(1).integers(-6, 6, (300, 4)).astype(float) * [0, 1, 1, 1])[0]
    Cc = S.stride_coords(Cs, 2)
    fwd = S.kernel_map(Cs, Cc, S.region_offsets(2, 1))
    bwd = S.kernel_map(Cc, Cs, S.region_offsets(2, 1), transposed=True)
    for j in range(8):
        o = np.nonzero(fwd[j] >= 0)[0]
        assert np.array_equal(bwd[j][fwd[j][o]], o)
    # every fine voxel has exactly one parent/offset
    assert np.array_equal((bwd >= 0).sum(axis=0), np.ones(len(Cs), int))
    # generative coordinates: 8 disjoint children per voxel
    assert len(S.generative_coords(Cc, 1)) == 8 * len(Cc)


def test_conv_linearity_and_bf16_rounding():
    rng = np.random.default_rng(2)
    Cs = S.sparse_tensor(np.concatenate([np.zeros((400, 1)), rng.integers(-5, 5, (400, 3))], axis=1).astype(float))[0]
    nbr = S.kernel_map(Cs, Cs, S.region_offsets(3, 1))
    W = rng.normal(0, 1, (27, 8, 5)).astype(np.float32)
    A, B = rng.normal(0, 1, (2, len(Cs), 8)).astype(np.float32)
    lhs = S.conv_forward(A + B, nbr, W)
    rhs = S.conv_forward(A, nbr, W) + S.conv_forward(B, nbr, W)
    np.testing.assert_allclose(lhs, rhs, rtol=1e-4, atol=1e-4)
    x = np.array([1.0, 1.00390625, 1.01171875, -3.3895314e38, 0.0], np.float32)
    r = S.bf16_round(x)
    assert r[0] == 1.0 and r[1] == 1.0 and r[2] == np.float32(1.015625)     # ties to even
    assert np.all((r.view(np.uint32) & 0xFFFF) == 0)


def test_topk_tie_rule_and_counts():
    C = np.array([[0, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 1, 1, 0], [1, 0, 0, 0]], np.int32)
    logits = np.array([0.5, 0.5, 0.5, 0.9, -1.0], np.float32)
    keep = S.topk_mask(logits, C, [2, 1])
    assert keep.tolist() == [True, False, False, True, True]   # tie -> lowest Morton key (0,0,0)


def test_gaussian_tables_shape_facts():
    rows, lens, offs = E.gaussian_tables()
    assert len(rows) == 64 and sum(len(r) for r in rows) == 27256 and max(lens) == 3133
    assert all(r[0] == 0 and r[-1] == 65536 and np.all(np.diff(r) > 0) for r in rows)
    t = E.scale_table()
    assert abs(t[0] - 0.11) < 1e-6 and abs(t[-1] - 256) < 1e-3
    idx = E.build_indexes(np.array([0.0, 0.11, 0.12, 1.0, 300.0], np.float32))
    assert idx[0] == 0 and idx[1] == 0 and idx[-1] == 63 and np.all(np.diff(idx) >= 0)
    assert E.quantize_symbols(np.array([0.5, 1.5, 2.5, -0.5], np.float32), 0).tolist() == [0, 2, 2, 0]


@pytest.mark.parametrize("n,seg", [(1, 32), (31, 32), (1000, 64), (50_000, 4096)])
def test_rans_oracles_roundtrip_any_int32(n, seg):
    rows, lens, offs = E.gaussian_tables()
    t = R.Tables(rows, offs)
    rng = np.random.default_rng(n)
    idx = rng.integers(0, 64, n)
    sym = np.rint(rng.normal(0, 1, n) * E.scale_table()[idx]).astype(np.int32)
    sym[:: max(n // 7, 1)] = rng.integers(-2 ** 30, 2 ** 30, len(sym[:: max(n // 7, 1)]))
    b = R.ref64_encode(sym, idx, t)
    assert len(b) % 4 == 0 and np.array_equal(R.ref64_decode(b, idx, t), sym)
    c = R.pcl1_encode(sym, idx, 1, t, seg)
    assert c[:4] == b"PCL1" and np.array_equal(R.pcl1_decode(c, idx, 1, n, t), sym)
    with pytest.raises(RuntimeError):
        R.pcl1_decode(c[: len(c) // 2], idx, 1, n, t)


def test_geometry_stream_oracle_roundtrip():
    """OCT2 coordinate streams: the numpy restatement round-trips (any order in, Morton order out) and
    the product's host-side header helpers agree with it (the coder itself is CUDA: test_gpu_octree)."""
    from oracle import octree_ref as O
    from b200pcc import geometry, synth
    p = synth.small_frame(seed=9, n_target=8000)["points"].astype(np.int64)
    p8 = np.unique(np.floor_divide(p, 8) * 8, axis=0)
    s = O.encode_points(p8[::-1], 3)
    q = O.decode_points(s, 3)
    assert sorted(map(tuple, q)) == sorted(map(tuple, p8))
    k = S.morton_key(np.concatenate([np.zeros((len(q), 1), np.int64), q], axis=1))
    assert np.all(k[1:] > k[:-1])                   # decoded in ascending Morton order
    assert len(s) < 3 * len(p8)                     # ~1-2 bytes per point
    assert geometry.stream_points(s) == len(p8)
    empty = O.encode_points(np.zeros((0, 3), int))
    assert len(empty) == 20 and O.decode_points(empty).shape == (0, 3)
    one = np.array([[-16, 8, 0]])
    assert len(O.encode_points(one)) == 20 and O.decode_points(O.encode_points(one)).tolist() == one.tolist()
    for bad in (s[: len(s) - 5], b"OCT1" + s[4:], s[:24] + bytes([s[24] ^ 1]) + s[25:]):
        with pytest.raises(ValueError):
            O.decode_points(bad, 3)
    blob, offs, meta32, counts = geometry.pack_streams([s, empty, O.encode_points(one)])
    assert counts == [len(p8), 0, 1] and offs.tolist() == [0, (len(s) + 3) // 4 * 4, (len(s) + 3) // 4 * 4 + 20]
    assert meta32.tolist() == [len(s), 20, 20, 0, len(p8), len(p8), len(p8) + 1]
    assert bytes(blob[offs[2]:offs[2] + 20]) == O.encode_points(one)
    with pytest.raises(ValueError):
        geometry.stream_points(b"nope")


def test_container_roundtrip_and_layout():
    from b200pcc import container
    import struct
    ks = [[3, 4], [30, 40], [300, 400]]
    blob = container.make_bitstream_batched(b"YYYY", b"ZZ", 7, 2, [b"p0", b"pp1"], ks, (1.0, 0.25))
    # App. B.1: int32 num_frames, float64 q_geo, float64 q_att, big endian
    assert struct.unpack(">idd", blob[:20]) == (2, 1.0, 0.25)
    assert struct.unpack(">iiii", blob[20:36]) == (7, 2, 4, 2)
    y, z, ysh, zsh, pts, ks2, q = container.read_bitstream_batched(blob)
    assert (y, z, ysh, zsh, pts, ks2, q) == ([b"YYYY"], [b"ZZ"], 7, 2, [b"p0", b"pp1"], ks, (1.0, 0.25))
    with pytest.raises(ValueError):
        container.read_bitstream_batched(blob[:-1])
    b2 = container.make_bitstream([b"y0", b"y1"], [b"z0", b"z1"], [5, 6], [1, 2], [b"a", b"bb"], ks, (0.0, 1.0))
    out = container.read_bitstream(b2)
    assert out[0] == [b"y0", b"y1"] and out[1] == [b"z0", b"z1"] and out[2] == [5, 6] and out[5] == ks


def test_container_c_abi_equals_struct_version():
    """include/pcc_b200.h section 6: the C reader/writer and container.py agree byte
    for byte on both layouts, including empty strings, and reject every truncation."""
    from b200pcc import container
    rng = np.random.default_rng(5)
    for nf in (0, 1, 3, 17):
        pts = [rng.integers(0, 256, int(rng.integers(0, 90)), dtype=np.uint8).tobytes() for _ in range(nf)]
        ks = [[int(rng.integers(0, 2 ** 31 - 1)) for _ in range(nf)] for _ in range(3)]
        y = rng.integers(0, 256, 1000, dtype=np.uint8).tobytes()
        z = b"" if nf == 1 else rng.integers(0, 256, 33, dtype=np.uint8).tobytes()
        q = (float(rng.random()), -0.5)
        blob = container.make_bitstream_batched(y, z, 12345, -7, pts, ks, q)
        assert container.make_bitstream_batched_native(y, z, 12345, -7, pts, ks, q) == blob
        assert container.read_bitstream_batched_native(blob, max_frames=32) == container.read_bitstream_batched(blob)
        ys = [rng.integers(0, 256, int(rng.integers(0, 50)), dtype=np.uint8).tobytes() for _ in range(nf)]
        zs = [rng.integers(0, 256, int(rng.integers(0, 9)), dtype=np.uint8).tobytes() for _ in range(nf)]
        ysh, zsh = list(range(10, 10 + nf)), list(range(nf))
        b2 = container.make_bitstream(ys, zs, ysh, zsh, pts, ks, q)
        assert container.make_bitstream_native(ys, zs, ysh, zsh, pts, ks, q) == b2
        assert container.read_bitstream_native(b2, max_frames=32) == container.read_bitstream(b2)
        for cut in sorted({1, 19, 35, len(blob) // 2, len(blob) - 1} & set(range(len(blob)))):
            with pytest.raises(ValueError):
                container.read_bitstream_batched_native(blob[:cut], max_frames=32)
        if nf:
            with pytest.raises(ValueError):
                container.read_bitstream_native(b2[:-1], max_frames=32)
    with pytest.raises(ValueError):     # more frames than the caller has room for
        container.read_bitstream_batched_native(container.make_bitstream_batched(b"", b"", 0, 0, [b"a"] * 5, [[0] * 5] * 3, (0, 0)),
                                                max_frames=4)


def test_scale_nn_host_matches_oracle():
    from b200pcc import config as cfgmod
    from b200pcc.model import ScaleNN
    sd = cfgmod.make_state_dict()
    nn = ScaleNN(64, 64)
    nn.load(sd)
    w1, b1 = sd["entropy_model.scale_nn.w1"].astype(np.float64), sd["entropy_model.scale_nn.b1"].astype(np.float64)
    w2, b2 = sd["entropy_model.scale_nn.w2"].astype(np.float64), sd["entropy_model.scale_nn.b2"].astype(np.float64)
    for q in ([1.0, 0.0], [0.0, 1.0], [1, 1], [0.3, 0.7]):
        ref = E.scale_nn(q, sd["entropy_model.scale_nn.w1"], sd["entropy_model.scale_nn.b1"],
                         sd["entropy_model.scale_nn.w2"], sd["entropy_model.scale_nn.b2"])
        assert np.array_equal(nn.numpy([q])[0], ref) and np.all(ref > 0)
        # independent statement (float64 matrix form): softplus(relu(q W1 + b1) W2 + b2)
        o = np.maximum(np.asarray(q, np.float64) @ w1 + b1, 0) @ w2 + b2
        np.testing.assert_allclose(ref, np.log1p(np.exp(o)), rtol=2e-6)


def test_bin_offset_is_the_laplacian_bin_centroid():
    """E.bin_offset (and its twin EntropyModel.get_offsets) against an independent statement: for a Laplacian of
    standard deviation sd the mean of x over the unit bin [k - 1/2, k + 1/2], k >= 1, is k - offset(sd), by
    numerical quadrature in float64."""
    for sd in (0.11, 0.3, 0.8, 1.5, 3.0, 5.0, 5.7, 9.0, 40.0, 256.0):
        b = sd / np.sqrt(2.0)
        k = 3.0
        x = np.linspace(k - 0.5, k + 0.5, 200001)
        w = np.exp(-(x - k) / b)                 # density up to a constant factor (x > 0)
        centroid = np.trapezoid(x * w, x) / np.trapezoid(w, x)
        want = min(max(k - centroid, 0.0), 0.5)
        got = float(E.bin_offset(np.array([sd], np.float32))[0])
        assert abs(got - want) <= 2e-5 + 2e-4 * want, (sd, got, want)


def test_fast_oracle_core_equals_numpy_statement():
    """oracle/sparse_ref.c (OpenMP core used for the full-size frames) vs the numpy statements: kernel maps
    (k3 stride 1 and 2, k2 strided, k2 generative/transposed) identical, convolutions equal to fp32 summation
    order, bf16 rounding identical."""
    rng = np.random.default_rng(3)
    raw = np.concatenate([rng.integers(0, 2, (6000, 1)), rng.integers(-24, 24, (6000, 3))], axis=1)
    C, _ = S.sparse_tensor(raw.astype(np.float64))
    C2 = S.stride_coords(C, 2)
    Cc = S.generative_coords(C2, 1)
    cases = [(C, C, S.region_offsets(3, 1), False), (C2, C2, S.region_offsets(3, 2), False),
             (C, C2, S.region_offsets(2, 1), False), (C2, Cc, S.region_offsets(2, 1), True)]
    F = rng.standard_normal((len(C), 32)).astype(np.float32) * 3
    W = (rng.standard_normal((27, 32, 48)) * 0.1).astype(np.float32)
    b = rng.standard_normal(48).astype(np.float32)
    R_ = rng.standard_normal((len(C), 48)).astype(np.float32)
    a = rng.standard_normal(10000).astype(np.float32) * np.float32(10.0) ** rng.integers(-30, 30, 10000)
    a[:4] = [np.inf, -np.inf, 0.0, 3.4e38]
    slow_maps = [S.kernel_map(ci, co, offs, transposed=t) for ci, co, offs, t in cases]
    slow_conv = S.conv_forward(F, slow_maps[0], W, b, R_, relu=True, bf16=True)
    slow_id = S.conv_forward(F, None, W[:1], b)
    slow_r = S.bf16_round(a)
    prev = S.use_fast(True)
    try:
        for (ci, co, offs, t), want in zip(cases, slow_maps):
            assert np.array_equal(S.kernel_map(ci, co, offs, transposed=t), want)
        np.testing.assert_allclose(S.conv_forward(F, slow_maps[0], W, b, R_, relu=True, bf16=True), slow_conv,
                                   rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(S.conv_forward(F, None, W[:1], b), slow_id, rtol=1e-5, atol=1e-5)
        assert np.array_equal(S.bf16_round(a).view(np.uint32), slow_r.view(np.uint32))
    finally:
        S.use_fast(prev)
    assert S.count_pairs(slow_maps[0]) == int((slow_maps[0] >= 0).sum()) > len(C)


def _cdf_diff(A, B):
    total = sum(len(a) for a in A)
    differ = sum(int((np.asarray(a) != np.asarray(b)).sum()) for a, b in zip(A, B))
    worst = max(int(np.abs(np.asarray(a, np.int64) - np.asarray(b, np.int64)).max()) for a, b in zip(A, B))
    return total, differ, worst


def test_gaussian_tables_independent_builders_and_float32_convention():
    """GaussianConditional.update(): the product builds the pmf with torch float32 ops in CompressAI 1.2.4's order,
    the oracle with numpy float32 + scipy's erfc -- two independent routes (their pmfs differ in the last bit of
    ~6 % of the entries) that must give IDENTICAL quantised tables.  The float64 evaluation is a different
    convention (what round 1 used): it changes ~1 % of the table entries, which is why App. A says to name it."""
    from b200pcc.model import GaussianConditional
    from b200pcc import ops
    gc = GaussianConditional()
    prows, center = gc.gaussian_pmf_f32()
    r32, c32 = E.gaussian_pmf(dtype=np.float32)
    assert np.array_equal(center, c32) and all(a.dtype == np.float32 for a in prows)
    ulp = max(float(np.abs(a.astype(np.float64) - b).max()) for a, b in zip(prows, r32))
    assert ulp <= 2.0 ** -23           # values are < 1: one unit in the last place at most
    product = [ops.pmf_to_quantized_cdf(r) for r in prows]      # C++ quantiser of the library (host function)
    oracle = E.gaussian_tables()[0]                               # Python quantiser on the oracle's own pmf
    assert _cdf_diff(product, oracle)[1] == 0
    total, differ, worst = _cdf_diff(oracle, E.gaussian_tables(dtype=np.float64)[0])
    assert total == 27256 and 0 < differ < 0.03 * total and worst <= 3, (differ, worst)


def test_factorized_tables_independent_builders():
    """EntropyBottleneck.update(force=True): torch float32 (product, CompressAI's ops) vs numpy float32 (oracle).
    softplus / matmul / tanh / sigmoid come from different libraries, so a pmf value may differ in its last bits
    and a table entry by one count: lengths, offsets and medians exact, <= 2 % of the entries off by one."""
    from b200pcc import config as cfgmod, ops
    from b200pcc.model import EntropyBottleneck
    sd = cfgmod.make_state_dict()
    eb = EntropyBottleneck(64)
    eb.load(sd)
    prows, minima, pmf_length, medians = eb.factorized_pmf_f32()
    rows, lens, offs, med = E.eb_tables(eb.matrices, eb.biases, eb.factors, eb.quantiles)
    assert np.array_equal(lens, pmf_length + 2) and np.array_equal(offs, -minima) and np.array_equal(med, medians)
    total, differ, worst = _cdf_diff([ops.pmf_to_quantized_cdf(r) for r in prows], rows)
    assert differ <= 0.02 * total and worst <= 1, (total, differ, worst)
    for r in rows:
        assert r[0] == 0 and r[-1] == 65536 and np.all(np.diff(r) > 0)


def test_golden_vectors():
    """Known-answer fixtures generated by tests/golden/make_golden.py (oracle output, committed)."""
    path = os.path.join(ROOT, "tests", "golden", "oracle_small.npz")
    g = np.load(path)
    from b200pcc import config as cfgmod, synth
    from oracle.codec_ref import RefCodec
    frames = [synth.small_frame(seed=int(g["seed"]), n_target=int(g["n_target"]))]
    assert np.array_equal(frames[0]["points"], g["points"])
    codec = RefCodec(cfgmod.SYNTH_SMALL, cfgmod.make_state_dict(), bf16=True)
    blob = codec.compress(frames, [[1.0, 1.0]])
    assert np.array_equal(blob["y_coords"], g["y_coords"])
    assert blob["k"] == g["k"].tolist()
    assert np.array_equal(blob["debug"]["y_sym"][0], g["y_sym"])
    assert np.frombuffer(blob["y_strings"][0], np.uint8).tobytes() == g["y_string"].tobytes()
    assert np.frombuffer(blob["z_string"], np.uint8).tobytes() == g["z_string"].tobytes()
    out, _ = codec.decompress(blob, 0)
    assert np.array_equal(out[0]["points"], g["dec_points"])
    np.testing.assert_allclose(out[0]["colors"], g["dec_colors"], atol=1e-6)


def test_recorded_frame_coordinate_sizes():
    """The reference's recorded frame 0 (tests/golden/recorded_frames.npz) through the oracle's coordinate
    conventions reproduces the sizes SURVEY 7.3 measured: 19 679 -> 8 874 -> 2 831 -> 815 -> 229 -> 63 and 7.2
    kernel-map pairs per point at stride 1."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "recorded_frames.npz"))
    p = g["points_0"]
    assert p.dtype == np.int16 and g["colors_0"].dtype == np.float64 and p.min() < 0
    C = np.concatenate([np.zeros((len(p), 1), np.int32), p.astype(np.int32)], axis=1)
    Cl = S.sparse_tensor(C.astype(np.float64))[0]
    assert abs(S.count_pairs(S.kernel_map(Cl, Cl, S.region_offsets(3, 1))) / len(Cl) - 7.2) < 0.1
    sizes = [len(Cl)]
    for l in range(5):
        Cl = S.stride_coords(Cl, 2 << l)
        sizes.append(len(Cl))
    assert sizes == [19679, 8874, 2831, 815, 229, 63]
