"""Frame-parallel sharding on CPU: world_size-2 gloo process group.  Covers the frame -> rank
arithmetic, the bitstream gather (ragged and empty payloads, more ranks than frames) and the statistics
reduction.  The per-frame codec itself needs a GPU (tests/test_gpu_*.py); here the encoder is a stub."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _StubEncoder:
    """Deterministic stand-in with the CompressionPipeline.compress() shape."""

    def compress(self, gop):
        f = gop["frames"][0]
        n = len(f["points"])
        payload = bytes((int(f["points"][:, 0].sum()) + k) % 251 for k in range(n % 97))
        return {0: gop["frames"], 1: payload}, {}


def _frames(n):
    rng = np.random.default_rng(3)
    return [{"points": rng.integers(-50, 50, (40 + 13 * i, 3)).astype(np.int16),
             "colors": rng.random((40 + 13 * i, 3))} for i in range(n)]


def _worker(rank, world, port, num_frames, ret):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "demo-learned-point-cloud-compression_b200"))
    from b200pcc import shard
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames = _frames(num_frames)
        codec = shard.ShardedSequenceCodec(encoder=_StubEncoder())
        streams, points = codec.encode(frames)
        expect = [_StubEncoder().compress({"frames": [f]})[0][1] for f in frames]
        assert streams == expect, "gathered sequence differs from the single-process sequence"
        mine = shard.frames_for_rank(num_frames, rank, world)
        assert points == sum(len(frames[i]["points"]) for i in mine)
        tot = shard.reduce_stats([points, float(rank)], "sum")
        assert tot[0] == sum(len(f["points"]) for f in frames) and tot[1] == sum(range(world))
        assert shard.reduce_stats([float(rank)], "max") == [float(world - 1)]
        # wrong ownership is rejected
        try:
            shard.gather_bitstreams({0: b"x"} if rank else {1: b"x"}, 2)
            ok = False
        except ValueError:
            ok = True
        assert ok
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("num_frames", [7, 1, 0])
def test_frame_parallel_gather_gloo_world2(num_frames):
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, num_frames, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert dict(ret) == {0: 1, 1: 1}


def test_partition_covers_every_frame_once():
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "demo-learned-point-cloud-compression_b200"))
    from b200pcc import shard
    for world in (1, 2, 4, 8):
        for n in (0, 1, 7, 300):
            seen = sorted(i for r in range(world) for i in shard.frames_for_rank(n, r, world))
            assert seen == list(range(n))
            assert all(shard.owner_of(i, world) == r for r in range(world) for i in shard.frames_for_rank(n, r, world))
    # single process (no process group): gather is the identity
    assert shard.gather_bitstreams({0: b"a", 1: b"", 2: b"ccc"}, 3) == [b"a", b"", b"ccc"]
    with pytest.raises(ValueError):
        shard.frames_for_rank(3, 2, 2)
