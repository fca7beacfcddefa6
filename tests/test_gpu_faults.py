"""GPU: nothing in the library traps or faults on bad input or on its own failure paths.  A trap / misaligned
address is a sticky error that ends the CUDA context of the whole process -- in the reference's long-lived
encoder / decoder services (sender/encoder/encoder.py:50-75, receiver/decoder/decoder.py:41-62) that is the
process, not the call.  Covered here: the device fault word (undersized block table, forced pipeline-barrier
timeout in the convolution kernel), byte-flip fuzzing of the PCL1 (rANS) and OCT2 (coordinate) streams including
their size tables, back-to-back stage_frames() on one pinned buffer, and two devices in one process."""
import copy

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import octree_ref as O  # noqa: E402
from oracle import sparse_ref as S  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def model(dev):
    from b200pcc.pipeline import load_model
    return load_model(dev)


def _still_alive(dev):
    torch.cuda.synchronize()          # a sticky fault would surface here
    a = torch.arange(1000, device=dev)
    assert int(a.sum().item()) == 499500


def test_undersized_block_table_is_a_fault_code_not_a_trap(dev):
    from b200pcc import ops, synth
    from b200pcc._lib import lib, check, PccError
    frame = synth.small_frame(seed=41, n_target=20000)
    C = np.concatenate([np.zeros((len(frame["points"]), 1), np.int32), frame["points"].astype(np.int32)], axis=1)
    keys = ops.sort_unique(ops.keys_from_coords(torch.from_numpy(C).to(dev)))[0]
    ops.verify_deferred(device=dev)                      # clears anything earlier tests left behind
    cap = 8                                              # far fewer slots than occupied 4x4x4 blocks
    tk = torch.empty(cap, dtype=torch.int64, device=dev)
    tv = torch.empty((cap, 2), dtype=torch.int64, device=dev)
    check(lib.pcc_blockmap_build(keys.data_ptr(), keys.shape[0], 0, tk.data_ptr(), tv.data_ptr(), cap, ops._stream()))
    with pytest.raises(PccError, match="block table"):
        ops.verify_deferred(device=dev)
    _still_alive(dev)
    ops.verify_deferred(device=dev)                      # the fetch cleared the word
    # a table of the documented size works afterwards
    nbr = ops.kmap_k3_blocks(keys, 0, ops.blockmap_build(keys, 0))
    Cg = ops.coords_from_keys(keys).cpu().numpy()
    assert np.array_equal(nbr.cpu().numpy(), S.kernel_map(Cg, Cg, S.region_offsets(3, 1)))
    ops.verify_deferred(device=dev)


def test_forced_pipeline_timeout_raises_and_leaves_the_context_usable(model, dev):
    """pcc_debug_inject(1): the next convolution launch drops one hand-off; its gather warps run out of patience,
    the CTA aborts, the kernel ends, the fault word says so -- through verify_deferred (decoder route) and through
    the packed-strings header (encoder route).  The same calls work again right after."""
    from b200pcc import ops, synth
    from b200pcc._lib import lib, check, PccError
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline
    frame = synth.small_frame(seed=42, n_target=30000)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    good, _ = enc.compress({"frames": [copy.deepcopy(frame)]})
    want, _ = dec.decompress(good[1])
    ops.verify_deferred(device=dev)
    # encoder route
    check(lib.pcc_debug_inject(1))
    with pytest.raises(PccError, match="pipeline barrier"):
        enc.compress({"frames": [copy.deepcopy(frame)]})
    _still_alive(dev)
    again, _ = enc.compress({"frames": [copy.deepcopy(frame)]})
    assert again[1] == good[1]
    # decoder route
    check(lib.pcc_debug_inject(1))
    with pytest.raises(PccError, match="pipeline barrier"):
        dec.decompress(good[1])
    _still_alive(dev)
    out, _ = dec.decompress(good[1])
    assert np.array_equal(out[0]["points"], want[0]["points"]) and np.array_equal(out[0]["colors"], want[0]["colors"])


def test_rans_stream_fuzz_never_faults(model, dev):
    """Byte flips everywhere in a PCL1 container -- header, segment-size table (a size that is not a multiple of
    four used to become a misaligned 32-bit load), per-segment word counts, lane states, payload: the decoder either
    reports a status or returns n symbols; the device stays usable and the clean stream still decodes."""
    from b200pcc import ops
    gc = model.entropy_model.gaussian_conditional
    n, seg = 40_000, 4096
    rng = np.random.default_rng(2)
    idx = rng.integers(0, 64, n).astype(np.int16)
    sym = np.rint(rng.normal(0, 2, n)).astype(np.int32)
    ti = torch.from_numpy(idx).to(dev)
    buf, nb = ops.rans_encode(torch.from_numpy(sym).to(dev), ti, 1, gc.tables, seg)
    good = buf[: int(nb.item())].cpu().numpy().copy()
    nseg = int(good[12:16].view(np.uint32)[0])
    assert nseg == (n + seg - 1) // seg
    table = slice(16, 16 + 4 * nseg)
    seg0 = 16 + 4 * nseg
    cases = []
    for k in range(nseg):                                  # odd / unaligned / huge segment sizes
        for delta in (1, 2, 3, -1, 5, 0x7FFFFFF1):
            b = good.copy()
            v = b[16 + 4 * k: 20 + 4 * k].view(np.uint32)
            v[0] = np.uint32((int(v[0]) + delta) & 0xFFFFFFFF)
            cases.append(b)
    for pos in list(range(0, 16)) + list(range(seg0, seg0 + 140)):      # header bytes, wpos + lane states of segment 0
        b = good.copy()
        b[pos] ^= 0xFF
        cases.append(b)
    for _ in range(60):                                    # random flips anywhere
        b = good.copy()
        for pos in rng.integers(0, len(b), rng.integers(1, 6)):
            b[pos] ^= rng.integers(1, 256)
        cases.append(b)
    reported = 0
    for b in cases:
        pad = (-len(b)) % 4
        cont = torch.from_numpy(np.concatenate([b, np.zeros(pad, np.uint8)])).to(dev)
        out, status = ops.rans_decode(cont, seg, ti, 1, n, gc.tables)
        reported += int(status.item()) != 0
        assert out.shape[0] == n
    _still_alive(dev)
    assert reported >= 6 * nseg // 2                       # every misaligned size is reported
    out, status = ops.rans_decode(torch.from_numpy(good).to(dev), seg, ti, 1, n, gc.tables)
    assert int(status.item()) == 0 and np.array_equal(out.cpu().numpy(), sym)


def test_octree_stream_fuzz_never_faults(dev):
    from b200pcc import geometry
    rng = np.random.default_rng(6)
    pts = np.unique(rng.integers(-400, 400, (6000, 3)) // 8 * 8, axis=0)
    good = O.encode_points(pts, 3)
    n = int(np.frombuffer(good[4:8], np.uint32)[0])
    cases = []
    for pos in range(4, min(len(good), 20 + 4 * 16)):       # n, depth, root, node counts per level
        for flip in (0x01, 0x80, 0xFF):
            b = bytearray(good)
            b[pos] ^= flip
            cases.append(bytes(b))
    for _ in range(80):
        b = bytearray(good)
        for pos in rng.integers(4, len(b), rng.integers(1, 5)):
            b[pos] ^= int(rng.integers(1, 256))
        cases.append(bytes(b))
    for bad in cases:
        try:
            blob, offs, meta32, counts = geometry.pack_streams([bad])
        except ValueError:
            continue
        if counts[0] > 4 * n + 1000:                        # an absurd n is the host's job (it sizes the output)
            continue
        t = lambda a: torch.from_numpy(a).to(dev)
        keys, status = geometry.decode_device(t(blob), t(offs), t(meta32), counts)
        assert keys.shape[0] == counts[0]
        int(status.item())
    _still_alive(dev)


def test_stage_frames_back_to_back_keeps_both_frames_intact(model, dev):
    """stage_frames() twice in a row reuses one pinned buffer: the second call must not overwrite bytes the first
    call's DMA is still reading (frame 0 of the resident bench arm could be silently corrupted)."""
    from b200pcc import synth
    from b200pcc.pipeline import CompressionPipeline
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    a = synth.small_frame(seed=43, n_target=400_000)
    b = synth.small_frame(seed=44, n_target=300_000)       # not larger than a: would land on a's bytes
    for _ in range(3):
        sa = enc.stage_frames([a])
        sb = enc.stage_frames([b])
        torch.cuda.synchronize()
        assert np.array_equal(sa["points"][0].cpu().numpy(), a["points"])
        assert np.array_equal(sa["colors"][0].cpu().numpy(), a["colors"])
        assert np.array_equal(sb["points"][0].cpu().numpy(), b["points"])
        assert np.array_equal(sb["colors"][0].cpu().numpy(), b["colors"])


@pytest.mark.skipif(torch.cuda.is_available() and torch.cuda.device_count() < 2, reason="needs two devices")
def test_two_devices_in_one_process(model):
    """The shared-memory opt-in of the convolution / decoder kernels and the SM count are per device: a frame
    coded on cuda:0 and then on cuda:1 in the same process gives identical bytes and reconstructions."""
    from b200pcc import synth
    from b200pcc.pipeline import CompressionPipeline, DecompressionPipeline, load_model
    frame = synth.small_frame(seed=45, n_target=25000)
    res = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        m = model if d == 0 else load_model(dev)
        enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=m)
        dec = DecompressionPipeline(device=dev, model=m)
        comp, _ = enc.compress({"frames": [copy.deepcopy(frame)]})
        out, _ = dec.decompress(comp[1])
        res.append((comp[1], out))
    assert res[0][0] == res[1][0]
    assert np.array_equal(res[0][1][0]["points"], res[1][1][0]["points"])
    assert np.array_equal(res[0][1][0]["colors"], res[1][1][0]["colors"])
