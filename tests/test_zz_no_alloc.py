"""GPU: pcc_coords_downsample_static (scratch in a static device array, no allocation) against
pcc_coords_downsample (caller-owned workspace) through the C ABI.  Integer work: results must be identical.
The file is named so that pytest collects it last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _sorted_keys(n_target, seed, dev, offset=(0, 0, 0)):
    from b200pcc import ops, synth
    f = synth.small_frame(seed=seed, n_target=n_target, offset=offset)
    C = np.concatenate([np.zeros((len(f["points"]), 1), np.int32), f["points"].astype(np.int32)], axis=1)
    keys, _ = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(np.ascontiguousarray(C)).to(dev)))
    return keys


def _call_workspace(keys, log2_stride):
    from b200pcc import ops
    out, parent = ops.coords_downsample(keys, log2_stride)
    return out.clone(), parent.clone()


def _call_static(keys, log2_stride, out, parent, cnt):
    """Outputs are preallocated by the caller; the call itself allocates nothing."""
    from b200pcc._lib import lib, check
    from b200pcc.ops import _ptr, _stream
    check(lib.pcc_coords_downsample_static(_ptr(keys), keys.shape[0], log2_stride, _ptr(out), _ptr(parent), _ptr(cnt),
                                           _stream()))


def test_static_scratch_downsample_equals_workspace_variant_twice():
    from b200pcc._lib import lib
    dev = torch.device("cuda:0")
    assert lib.pcc_coords_downsample_static_max_rows() == 1 << 20
    # first call: the larger input at stride 2; second call: a smaller, different input at stride 4, so the
    # second call runs over whatever the first one left in the shared array
    cases = [(_sorted_keys(60000, 11, dev), 1), (_sorted_keys(9000, 12, dev, offset=(-40, 7, 3)), 2)]
    expect = [_call_workspace(k, l) for k, l in cases]
    bufs = [(torch.full((k.shape[0],), -1, dtype=torch.int64, device=dev),
             torch.full((k.shape[0],), -1, dtype=torch.int32, device=dev),
             torch.full((1,), -1, dtype=torch.int32, device=dev)) for k, _ in cases]
    torch.cuda.synchronize()
    free_before = None
    for rep in range(2):            # rep 0 also loads the kernels' module; rep 1 is checked for allocations
        if rep == 1:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
        for (k, l), (out, parent, cnt) in zip(cases, bufs):
            _call_static(k, l, out, parent, cnt)
        torch.cuda.synchronize()
        if rep == 1:
            assert torch.cuda.mem_get_info(dev)[0] == free_before, "device memory changed across the static calls"
        for (k, l), (out, parent, cnt), (e_out, e_parent) in zip(cases, bufs, expect):
            m = int(cnt.item())
            assert m == e_out.shape[0]
            assert torch.equal(out[:m], e_out)
            assert torch.equal(parent, e_parent)
    # empty input: count 0, nothing else touched
    cnt = torch.full((1,), -1, dtype=torch.int32, device=dev)
    e = torch.empty(0, dtype=torch.int64, device=dev)
    _call_static(e, 1, e, torch.empty(0, dtype=torch.int32, device=dev), cnt)
    assert int(cnt.item()) == 0


def test_static_scratch_rejects_oversize_with_a_clear_error():
    from b200pcc._lib import lib, PccError
    dev = torch.device("cuda:0")
    n = (1 << 20) + 1
    keys = torch.arange(n, dtype=torch.int64, device=dev) << 3
    out = torch.empty(n, dtype=torch.int64, device=dev)
    parent = torch.empty(n, dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    with pytest.raises(PccError, match="static scratch limit"):
        _call_static(keys, 1, out, parent, cnt)
    # exactly the limit is accepted and equals the workspace variant
    keys = keys[:1 << 20].contiguous()
    e_out, e_parent = _call_workspace(keys, 1)
    _call_static(keys, 1, out, parent, cnt)
    m = int(cnt.item())
    assert m == e_out.shape[0] and torch.equal(out[:m], e_out) and torch.equal(parent[:1 << 20], e_parent)
