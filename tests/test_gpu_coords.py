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
