"""GPU parity: fused entropy-model kernels and the PCL1 lane coder vs the CPU oracle (bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import entropy_ref as E  # noqa: E402
from oracle import rans_ref as R  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def model(dev):
    from b200pcc.pipeline import load_model
    return load_model(dev)


def test_tables_match_oracle(model):
    em = model.entropy_model
    rows, lens, offs = E.gaussian_tables(em.gaussian_conditional.scale_table)
    assert np.array_equal(em.gaussian_conditional.scale_table, E.scale_table())
    assert all(np.array_equal(a, b) for a, b in zip(rows, em.gaussian_conditional._quantized_cdf))
    assert np.array_equal(lens, em.gaussian_conditional._cdf_length)
    assert np.array_equal(offs, em.gaussian_conditional._offset)
    eb = em.entropy_bottleneck
    # the factorized pmf goes through softplus / matmul / tanh / sigmoid of two different libraries (torch float32
    # in the product like CompressAI, numpy float32 in the oracle): an entry may differ by one count
    # (tests/test_cpu_oracle.py::test_factorized_tables_independent_builders states the bar)
    rz, lz, oz, med = E.eb_tables(eb.matrices, eb.biases, eb.factors, eb.quantiles)
    diff = [np.abs(np.asarray(a, np.int64) - np.asarray(b, np.int64)) for a, b in zip(rz, eb._quantized_cdf)]
    assert max(int(d.max()) for d in diff) <= 1 and sum(int((d > 0).sum()) for d in diff) <= 0.02 * sum(map(len, rz))
    assert np.array_equal(lz, eb._cdf_length) and np.array_equal(oz, eb._offset)
    assert np.array_equal(med, eb.medians)


def _latents(n, c, seed):
    rng = np.random.default_rng(seed)
    y = (rng.normal(0, 4, (n, c)) * rng.uniform(0.1, 3, (1, c))).astype(np.float32)
    y[::97] = np.rint(y[::97]) + 0.5                      # exact .5 ties: round half to even
    params = np.concatenate([rng.normal(1.5, 2.0, (n + 50, c)), rng.normal(0, 2, (n + 50, c))], axis=1).astype(np.float32)
    prow = rng.permutation(n + 50)[:n].astype(np.int32)
    prow[::53] = -1                                        # absent rows -> zeros
    perm = rng.permutation(n).astype(np.int32)
    return y, params, prow, perm


@pytest.mark.parametrize("n,c,q", [(3000, 64, 3), (1237, 48, 1), (31, 64, 2)])
def test_gauss_encode_bit_exact(model, dev, n, c, q):
    from b200pcc import ops
    em = model.entropy_model
    gc = em.gaussian_conditional
    y, params, prow, perm = _latents(n, c, n)
    rng = np.random.default_rng(5)
    scale = rng.uniform(0.3, 3.0, (q, c)).astype(np.float32)
    t = lambda a: torch.from_numpy(a).to(dev)
    sym, idx = ops.gauss_encode(t(y), t(params), t(prow), t(perm), t(scale), gc.table_d, gc.scale_bound)
    # oracle in canonical column order
    P = np.where(prow[:, None] >= 0, params[np.maximum(prow, 0)], 0).astype(np.float32)
    yl, Pl = y[perm], P[perm]
    for qi in range(q):
        s_ref, i_ref = E.gaussian_encode_symbols(yl.T, Pl[:, :c].T, Pl[:, c:].T, scale[qi])
        assert np.array_equal(sym[qi].cpu().numpy(), s_ref)
        assert np.array_equal(idx[qi].cpu().numpy().astype(np.int32), i_ref)
    # decoder-side index kernel agrees with the encoder's
    idx_d = ops.gauss_indexes(t(params), t(prow), t(perm), t(scale[:1].copy()), gc.table_d, gc.scale_bound, n, c)
    assert torch.equal(idx_d, idx[0])
    # dequantisation (fp tolerance: tanhf differs between libm and CUDA)
    yb, yf = ops.gauss_dequant(sym[0].contiguous(), t(params), t(prow), t(perm), t(scale[0].copy()), em.offset_gain_d[:c].contiguous(),
                               gc.scale_bound, want_f32=True)
    ref = E.gaussian_dequant(sym[0].cpu().numpy(), Pl[:, :c].T, Pl[:, c:].T, scale[0], em.offset_gain[:c])
    got = yf.cpu().numpy()[perm].T
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-5)
    assert torch.equal(yb, yf.to(torch.bfloat16))


def test_factorized_quant_bit_exact(model, dev):
    from b200pcc import ops
    eb = model.entropy_model.entropy_bottleneck
    n, c = 777, eb.channels
    rng = np.random.default_rng(3)
    z = rng.normal(0, 4, (n, c)).astype(np.float32)
    z[::13] = np.rint(z[::13]) + 0.5
    perm = rng.permutation(n).astype(np.int32)
    sym, zb, zf = ops.factorized_quant(torch.from_numpy(z).to(dev), torch.from_numpy(perm).to(dev), eb.medians_d, want_f32=True)
    s_ref, zh_ref = E.factorized_quantize(z[perm].T, eb.medians)
    assert np.array_equal(sym.cpu().numpy(), s_ref)
    assert np.array_equal(zf.cpu().numpy()[perm].T, zh_ref)
    zb2, zf2 = ops.factorized_dequant(sym, torch.from_numpy(perm).to(dev), eb.medians_d, want_f32=True)
    assert torch.equal(zf2, zf) and torch.equal(zb2, zb)


def _tables_ref(m):
    return R.Tables(m._quantized_cdf, m._offset)


@pytest.mark.parametrize("n,seg", [(200_000, 1 << 15), (70_001, 4096), (33, 32), (1, 64)])
def test_rans_gaussian_bytes_match_oracle_and_roundtrip(model, dev, n, seg):
    from b200pcc import ops
    gc = model.entropy_model.gaussian_conditional
    rng = np.random.default_rng(n)
    idx = rng.integers(0, 64, n).astype(np.int32)
    sym = np.rint(rng.normal(0, 1, n) * gc.scale_table[idx]).astype(np.int32)
    if n > 100:
        sym[::501] += 70000          # escapes, both signs
        sym[3::733] -= 123456
        sym[5] = 2 ** 30
        sym[6] = -(2 ** 30)
    ts, ti = torch.from_numpy(sym).to(dev), torch.from_numpy(idx.astype(np.int16)).to(dev)
    buf, nb = ops.rans_encode(ts, ti, 1, gc.tables, seg)
    nbytes = int(nb.item())
    got = buf[:nbytes].cpu().numpy().tobytes()
    ref = R.pcl1_encode(sym, idx, 1, _tables_ref(gc), seg)
    assert got == ref
    # CPU oracle decodes the GPU stream; GPU decodes it too
    assert np.array_equal(R.pcl1_decode(got, idx, 1, n, _tables_ref(gc)), sym)
    cont = torch.from_numpy(np.frombuffer(got, np.uint8).copy()).to(dev)
    dec, status = ops.rans_decode(cont, seg, ti, 1, n, gc.tables)
    assert int(status.item()) == 0
    assert np.array_equal(dec.cpu().numpy(), sym)
    # same stream without the search-window table (plain binary search over the whole CDF row)
    import copy
    from b200pcc._lib import CdfTables
    plain = copy.copy(gc.tables)
    plain.struct = CdfTables(gc.tables.cdf.data_ptr(), gc.tables.cdf_off.data_ptr(), gc.tables.cdf_size.data_ptr(),
                             gc.tables.offset.data_ptr(), None, gc.tables.num_tables, 0)
    dec2, status2 = ops.rans_decode(cont, seg, ti, 1, n, plain)
    assert int(status2.item()) == 0 and torch.equal(dec2, dec)
    if n >= 200_000:
        # bit-rate: within 2 % of the CompressAI-style single 64-bit stream
        single = R.ref64_encode(sym, idx, _tables_ref(gc))
        assert len(got) <= 1.02 * len(single), (len(got), len(single))


def test_rans_encode_batch_equals_per_stream_and_the_oracle(model, dev):
    """All quality levels in one coder call (codec_pipeline.py:414-432): every stream of the batch is byte-identical
    to coding it alone, and to the scalar restatement."""
    from b200pcc import ops
    gc = model.entropy_model.gaussian_conditional
    Q, n, seg = 5, 40_000 + 17, 4096
    rng = np.random.default_rng(77)
    idx = rng.integers(0, 64, (Q, n)).astype(np.int32)
    sym = np.rint(rng.normal(0, 1, (Q, n)) * gc.scale_table[idx] * (1 + np.arange(Q))[:, None]).astype(np.int32)
    sym[1, ::301] += 90000
    sym[3, 7::911] -= 2 ** 29
    ts, ti = torch.from_numpy(sym).to(dev), torch.from_numpy(idx.astype(np.int16)).to(dev)
    pend = ops.rans_encode_batch(ts, ti, 1, gc.tables, seg)
    assert len(pend) == Q
    for q, (buf, nb) in enumerate(pend):
        got = buf[: int(nb.item())].cpu().numpy().tobytes()
        b1, n1 = ops.rans_encode(ts[q].contiguous(), ti[q].contiguous(), 1, gc.tables, seg)
        assert got == b1[: int(n1.item())].cpu().numpy().tobytes()
        assert got == R.pcl1_encode(sym[q], idx[q], 1, _tables_ref(gc), seg)
    # the strings of a batch travel through fetch_strings like any others
    from b200pcc.model import fetch_strings
    strings = fetch_strings(pend)
    assert [len(x) for x in strings] == [int(nb.item()) for _, nb in pend]
    # factorized model (idx = None, table = channel) batched over two "levels"
    eb = model.entropy_model.entropy_bottleneck
    c, m = eb.channels, 333
    zs = np.rint(rng.normal(0, 4, (2, c, m))).astype(np.int32)
    pz = ops.rans_encode_batch(torch.from_numpy(zs.reshape(2, -1)).to(dev), None, m, eb.tables, 4096)
    for q, (buf, nb) in enumerate(pz):
        assert buf[: int(nb.item())].cpu().numpy().tobytes() == R.pcl1_encode(zs[q].reshape(-1), None, m, _tables_ref(eb), 4096)


def test_rans_factorized_channel_index(model, dev):
    from b200pcc import ops
    eb = model.entropy_model.entropy_bottleneck
    c, n = eb.channels, 901
    rng = np.random.default_rng(9)
    sym = np.rint(rng.normal(0, 4, (c, n))).astype(np.int32)     # beyond +-10 -> escapes
    buf, nb = ops.rans_encode(torch.from_numpy(sym.reshape(-1)).to(dev), None, n, eb.tables, 4096)
    got = buf[: int(nb.item())].cpu().numpy().tobytes()
    assert got == R.pcl1_encode(sym.reshape(-1), None, n, _tables_ref(eb), 4096)
    cont = torch.from_numpy(np.frombuffer(got, np.uint8).copy()).to(dev)
    dec, status = ops.rans_decode(cont, 4096, None, n, c * n, eb.tables)
    assert int(status.item()) == 0 and np.array_equal(dec.cpu().numpy().reshape(c, n), sym)
    # CompressAI-shaped surface
    x = torch.from_numpy(rng.normal(0, 3, (1, c, 50)).astype(np.float32)).to(dev)
    strings = eb.compress(x)
    xh = eb.decompress(strings, [50])
    ref = np.rint(x.cpu().numpy() - eb.medians[None, :, None]) + eb.medians[None, :, None]
    assert np.array_equal(xh.cpu().numpy(), ref)


def test_rans_corrupt_stream_is_reported(model, dev):
    from b200pcc import ops
    gc = model.entropy_model.gaussian_conditional
    n = 5000
    rng = np.random.default_rng(1)
    idx = rng.integers(0, 64, n).astype(np.int16)
    sym = np.zeros(n, np.int32)
    buf, nb = ops.rans_encode(torch.from_numpy(sym).to(dev), torch.from_numpy(idx).to(dev), 1, gc.tables, 4096)
    good = buf[: int(nb.item())].clone()
    trunc = good[: good.numel() // 2 // 4 * 4].contiguous()
    _, status = ops.rans_decode(trunc, 4096, torch.from_numpy(idx).to(dev), 1, n, gc.tables)
    assert int(status.item()) != 0
    _, status = ops.rans_decode(good, 4096, torch.from_numpy(idx).to(dev), 1, n + 1, gc.tables)
    assert int(status.item()) == 3


def test_pack_strings_one_buffer(dev):
    """pcc_pack_strings: n strings with device-side lengths -> lengths + 16-byte aligned payloads in one
    buffer; more than one launch batch (n > 64), empty strings, unaligned sources, capacity overflow."""
    import ctypes as C
    from b200pcc._lib import lib, check
    rng = np.random.default_rng(9)
    n = 150
    lens = [0, 1, 15, 16, 17, 100_003] + [int(v) for v in rng.integers(0, 3000, n - 6)]
    srcs = [rng.integers(0, 256, l + 3, dtype=np.uint8) for l in lens]
    # odd-indexed strings start at an unaligned device address
    bufs_d = [torch.from_numpy(s).to(dev)[(i & 1) * 3:] for i, s in enumerate(srcs)]
    lens_d = [torch.tensor([l], dtype=torch.int32, device=dev) for l in lens]
    hdr = lib.pcc_pack_strings_header_bytes(n)
    assert hdr % 16 == 0 and hdr >= 4 * (n + 3)
    need = hdr + sum((l + 15) // 16 * 16 for l in lens)
    for cap in (need, need - 32):
        out = torch.zeros(cap, dtype=torch.uint8, device=dev)
        bp = (C.c_void_p * n)(*[b.data_ptr() for b in bufs_d])
        lp = (C.c_void_p * n)(*[l.data_ptr() for l in lens_d])
        check(lib.pcc_pack_strings(bp, lp, n, out.data_ptr(), cap, torch.cuda.current_stream().cuda_stream))
        h = out.cpu().numpy()
        h32 = h[:hdr].view(np.uint32)
        assert h32[:n].tolist() == lens and int(h32[n + 2]) == need          # end[] says what a full copy needs
        pos = hdr
        for i, l in enumerate(lens):
            if pos + l <= cap:
                assert np.array_equal(h[pos:pos + l], srcs[i][(i & 1) * 3:(i & 1) * 3 + l]), i
            pos += (l + 15) // 16 * 16


def test_fetch_strings_guess_grows_and_shrinks(dev):
    from b200pcc.model import fetch_strings
    rng = np.random.default_rng(10)
    for sizes in ([10, 0, 5000], [2_000_000, 3], [7], [300_000] * 3):
        data = [rng.integers(0, 256, s, dtype=np.uint8) for s in sizes]
        pending = [(torch.from_numpy(np.concatenate([d, np.zeros(64, np.uint8)])).to(dev),
                    torch.tensor([len(d)], dtype=torch.int32, device=dev)) for d in data]
        got = fetch_strings(pending)
        assert [bytes(g) for g in got] == [d.tobytes() for d in data]
        views = fetch_strings(pending, as_bytes=False)
        assert [bytes(v) for v in views] == [d.tobytes() for d in data]
