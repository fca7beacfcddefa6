"""GPU parity: coordinate hash / kernel maps vs the CPU oracle (bit-exact integer work)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import sparse_ref as S  # noqa: E402
from helpers import to_canonical, stack_frames  # noqa: E402


@pytest.fixture(scope="module")
def frames():
    from b200pcc import synth
    return [synth.small_frame(seed=3, n_target=20000), synth.small_frame(seed=4, n_target=6000, offset=(5, -3, 2))]


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _keys(C, dev):
    from b200pcc import ops
    k = ops.keys_from_coords(torch.from_numpy(np.ascontiguousarray(C, np.int32)).to(dev))
    return ops.sort_unique(k)


def test_keys_roundtrip_and_dedup(frames, dev):
    from b200pcc import ops
    C, F = stack_frames(frames)
    rng = np.random.default_rng(0)
    # duplicate 500 rows at the end (different features) and shuffle nothing: first occurrence must win
    dup = rng.integers(0, len(C), 500)
    C2 = np.concatenate([C, C[dup]])
    F2 = np.concatenate([F, F[dup] + 7.0])
    keys, src = _keys(C2, dev)
    Cg = ops.coords_from_keys(keys).cpu().numpy()
    Co, Fo = S.sparse_tensor(C2.astype(np.float64), F2)
    Cg_c, src_c = to_canonical(Cg, src.cpu().numpy())
    assert np.array_equal(Cg_c, Co)
    assert np.array_equal(F2[src_c], Fo)
    # keys ascending and unique
    k = keys.cpu().numpy().view(np.uint64)
    assert np.all(k[1:] > k[:-1])
    # float coordinates are floored (ME.SparseTensor on float input)
    Cf = C.astype(np.float32) + 0.4
    kf = ops.keys_from_coords(torch.from_numpy(Cf).to(dev))
    ki = ops.keys_from_coords(torch.from_numpy(C).to(dev))
    assert torch.equal(kf, ki)


def test_lex_order_matches_reference_sort(frames, dev):
    from b200pcc import ops
    C, _ = stack_frames(frames)
    keys, _ = _keys(C, dev)
    Cg = ops.coords_from_keys(keys).cpu().numpy()
    perm = ops.lex_order(keys).cpu().numpy()
    ref = S.sort_order_reference(Cg)           # shared/utils.py:131-133 weights
    assert np.array_equal(perm, ref)


@pytest.mark.parametrize("log2s", [0, 1, 3])
def test_downsample_and_kmap_down2(frames, dev, log2s):
    from b200pcc import ops
    C, _ = stack_frames(frames)
    t = 1 << log2s
    C = S.stride_coords(C, t) if t > 1 else S.sparse_tensor(C.astype(np.float64))[0]
    keys, _ = _keys(C, dev)
    pk, parent = ops.coords_downsample(keys, log2s + 1)
    Cc = ops.coords_from_keys(keys).cpu().numpy()
    Cp = ops.coords_from_keys(pk).cpu().numpy()
    Co = S.stride_coords(C, 2 * t)
    assert np.array_equal(to_canonical(Cp)[0], Co)
    # parent_row maps every child to floor(c / 2t) * 2t
    par = parent.cpu().numpy()
    exp = Cc.copy()
    exp[:, 1:] = np.floor_divide(exp[:, 1:], 2 * t) * 2 * t
    assert np.array_equal(Cp[par], exp)
    # kernel map of the k2/s2 conv: same (offset, in-coord, out-coord) triples as the oracle
    nbr = ops.kmap_down2(keys, parent, log2s, pk.shape[0]).cpu().numpy()
    ref = S.kernel_map(C, Co, S.region_offsets(2, t))
    got = set()
    for j in range(8):
        o = np.nonzero(nbr[j] >= 0)[0]
        for oo, ii in zip(o, nbr[j][o]):
            got.add((j, tuple(Cc[ii]), tuple(Cp[oo])))
    exp_set = set()
    for j in range(8):
        o = np.nonzero(ref[j] >= 0)[0]
        for oo, ii in zip(o, ref[j][o]):
            exp_set.add((j, tuple(C[ii]), tuple(Co[oo])))
    assert got == exp_set


@pytest.mark.parametrize("log2s", [0, 2])
def test_kmap_k3_and_hash_query(frames, dev, log2s):
    from b200pcc import ops
    C, _ = stack_frames(frames)
    t = 1 << log2s
    C = S.stride_coords(C, t) if t > 1 else S.sparse_tensor(C.astype(np.float64))[0]
    keys, _ = _keys(C, dev)
    table = ops.hash_build(keys)
    nbr = ops.kmap_k3(keys, log2s, table).cpu().numpy()
    Cg = ops.coords_from_keys(keys).cpu().numpy()
    ref = S.kernel_map(Cg, Cg, S.region_offsets(3, t))   # oracle on the product's row order
    assert np.array_equal(nbr, ref)
    # product path: 4x4x4 block table (<= 8 probes per row) must give the identical map
    nbr_b = ops.kmap_k3_blocks(keys, log2s, ops.blockmap_build(keys, log2s)).cpu().numpy()
    assert np.array_equal(nbr_b, ref)
    pairs = (nbr >= 0).sum() / len(Cg)
    assert 3.0 < pairs < 27.0
    # queries: half present, half absent
    rng = np.random.default_rng(1)
    q = Cg[rng.integers(0, len(Cg), 4000)].copy()
    q[2000:, 1] += 1 if t == 1 else 1   # off-grid / probably absent
    rows = ops.hash_query(ops.keys_from_coords(torch.from_numpy(q).to(dev)), table).cpu().numpy()
    assert np.array_equal(rows, S.lookup_rows(Cg, q))


def test_block_kmap_edges_of_the_coordinate_range(dev):
    """Voxels on the int16 borders, several frames, isolated voxels and dense 8-children slabs."""
    from b200pcc import ops
    rng = np.random.default_rng(5)
    pts = [np.array([[0, -32768, -32768, -32768], [0, 32767, 32767, 32767], [0, -32768, 32767, 0],
                     [1, -32768, -32768, -32768], [2, 0, 0, 0], [2, 3, 3, 3], [2, 4, 4, 4], [2, -1, -1, -1]])]
    dense = np.stack(np.meshgrid(np.arange(-9, 7), np.arange(30, 41), np.arange(-3, 2), indexing="ij"), -1).reshape(-1, 3)
    pts.append(np.concatenate([np.full((len(dense), 1), 1), dense], axis=1))
    iso = rng.integers(-3000, 3000, (3000, 3))
    pts.append(np.concatenate([np.full((len(iso), 1), 3), iso], axis=1))
    C = np.unique(np.concatenate(pts).astype(np.int32), axis=0)
    for log2s in (0, 1, 3):
        t = 1 << log2s
        Ct = C.copy()
        Ct[:, 1:] = (Ct[:, 1:] // t) * t          # floor to the stride-t grid (borders stay in int16 range)
        Ct = np.unique(Ct, axis=0)
        keys, _ = _keys(Ct, dev)
        Cg = ops.coords_from_keys(keys).cpu().numpy()
        ref = S.kernel_map(Cg, Cg, S.region_offsets(3, t))
        nbr_b = ops.kmap_k3_blocks(keys, log2s, ops.blockmap_build(keys, log2s)).cpu().numpy()
        assert np.array_equal(nbr_b, ref), log2s


def test_children_and_frame_offsets(frames, dev):
    from b200pcc import ops
    C, _ = stack_frames(frames)
    C8 = S.stride_coords(C, 8)
    keys, _ = _keys(C8, dev)
    ck = ops.coords_children(keys, 2)
    Cch = ops.coords_from_keys(ck).cpu().numpy()
    k = ck.cpu().numpy().view(np.uint64)
    assert np.all(k[1:] > k[:-1])                  # children of sorted parents are sorted
    assert np.array_equal(to_canonical(Cch)[0], S.generative_coords(C8, 4))
    # child row 8p + j sits at parent + (dx,dy,dz)*4 with j = dx + 2dy + 4dz
    Cp = ops.coords_from_keys(keys).cpu().numpy()
    offs = S.region_offsets(2, 4)
    for j in range(8):
        assert np.array_equal(Cch[j::8, 1:], Cp[:, 1:] + offs[j])
    off = ops.frame_offsets(keys, 2).cpu().numpy()
    assert list(np.diff(off)) == [int((C8[:, 0] == b).sum()) for b in range(2)]


def test_empty_inputs(dev):
    from b200pcc import ops
    e = torch.empty((0, 4), dtype=torch.int32, device=dev)
    k = ops.keys_from_coords(e)
    u, src = ops.sort_unique(k)
    assert u.numel() == 0 and src.numel() == 0


def test_sort_and_lex_with_known_zero_bits_equal_the_general_calls(frames, dev):
    """pcc_sort_unique_bits / pcc_lex_order_bits skip the radix passes of bits the caller knows to be zero: same results
    as the 64-bit calls on multi-frame sets with negative coordinates, at strides 1 and 8."""
    from b200pcc import ops
    C, _ = stack_frames(frames)                      # two frames (b = 0, 1), coordinates straddle zero
    k_all = ops.keys_from_coords(torch.from_numpy(np.ascontiguousarray(C, np.int32)).to(dev))
    perm = torch.randperm(k_all.shape[0], device=dev)
    shuffled = k_all[perm].contiguous()
    u0, s0 = ops.sort_unique(shuffled)
    for bits in (49, 50, 56, 64):                    # 2 frames need 48 + 1 bits
        u1, s1 = ops.sort_unique(shuffled, key_bits=bits)
        assert torch.equal(u0, u1) and torch.equal(s0, s1), bits
    # canonical order at stride 1 and at stride 8 (coordinates are multiples of 8 there)
    p0 = ops.lex_order(u0)
    for fb in (1, 2, 16):
        assert torch.equal(ops.lex_order(u0, 0, fb), p0), fb
    C8 = C.copy()
    C8[:, 1:] = np.floor_divide(C8[:, 1:], 8) * 8
    k8, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(np.ascontiguousarray(C8, np.int32)).to(dev)))
    q0 = ops.lex_order(k8)
    for l2, fb in ((3, 1), (3, 16), (2, 1), (0, 1)):
        assert torch.equal(ops.lex_order(k8, l2, fb), q0), (l2, fb)
    ref = S.sort_order_reference(ops.coords_from_keys(k8).cpu().numpy())
    assert np.array_equal(q0.cpu().numpy(), ref)


def test_zero_unkept_rows_matches_numpy(dev):
    """pcc_zero_unkept_rows: rows not listed in keep (ascending) become zero, the others stay bit for bit; edge cases
    keep = nothing, keep = everything, first / last row kept or not, long gaps."""
    from b200pcc import ops
    rng = np.random.default_rng(5)
    n, c = 5003, 32
    x = torch.randn((n, c), device=dev).to(torch.bfloat16)
    cases = [np.array([], np.int32), np.arange(n, dtype=np.int32), np.array([0], np.int32), np.array([n - 1], np.int32),
             np.sort(rng.choice(n, 1777, replace=False)).astype(np.int32),
             np.concatenate([np.arange(10), np.arange(4000, 4003)]).astype(np.int32)]
    for keep in cases:
        y = x.clone()
        ops.zero_unkept_rows(y, torch.from_numpy(keep).to(dev))
        want = torch.zeros_like(x)
        if len(keep):
            idx = torch.from_numpy(keep.astype(np.int64)).to(dev)
            want[idx] = x[idx]
        assert torch.equal(y.view(torch.int16), want.view(torch.int16)), len(keep)
