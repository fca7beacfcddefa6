"""Frame-parallel sharding of a sequence over the GPUs of one box.

Frames are independent units of the codec: the batch index is only a coordinate column
(shared/utils.py:31-35), no layer mixes batch entries and top-k is per frame
(codec_pipeline.py:505-507).  So frame i goes to rank i mod G, every rank holds a replica of the model
and the CDF tables, and there is NO collective on the data path.  The one exchange step is assembling
the sequence's bitstreams (and reducing statistics) after the timed path: an all_gather of byte lengths
followed by an all_gather of padded byte tensors -- NCCL over NVLink when the process group is NCCL
(tensors on the rank's GPU), gloo on CPU tensors in the tests.

One process per GPU, launched by torch.distributed.run; RANK / LOCAL_RANK / WORLD_SIZE from the env.
"""
from typing import Dict, Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def frames_for_rank(num_frames: int, rank: int, world_size: int) -> List[int]:
    """Indices of the frames rank `rank` codes: i = rank (mod world_size)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    return list(range(rank, num_frames, world_size))


def owner_of(frame_index: int, world_size: int) -> int:
    return frame_index % world_size


def _world(group) -> int:
    return dist.get_world_size(group) if dist.is_initialized() else 1


def _rank(group) -> int:
    return dist.get_rank(group) if dist.is_initialized() else 0


def gather_bitstreams(local: Dict[int, bytes], num_frames: int, group=None,
                      device: Optional[torch.device] = None) -> List[bytes]:
    """Every rank passes {frame index: bitstream} for the frames it owns and receives the whole
    sequence [bitstream of frame 0, 1, ...].  Two collectives: lengths, then padded payloads."""
    world, rank = _world(group), _rank(group)
    mine = frames_for_rank(num_frames, rank, world)
    if sorted(local.keys()) != mine:
        raise ValueError(f"rank {rank} must provide frames {mine}, got {sorted(local.keys())}")
    if world == 1:
        return [local[i] for i in range(num_frames)]
    dev = device if device is not None else torch.device("cpu")
    per_rank = (num_frames + world - 1) // world
    lens = torch.zeros(per_rank, dtype=torch.int64, device=dev)
    for slot, i in enumerate(mine):
        lens[slot] = len(local[i])
    all_lens = [torch.zeros_like(lens) for _ in range(world)]
    dist.all_gather(all_lens, lens, group=group)
    total = [int(l.sum().item()) for l in all_lens]
    cap = max(max(total), 1)
    payload = torch.zeros(cap, dtype=torch.uint8, device=dev)
    blob = b"".join(local[i] for i in mine)
    if blob:
        payload[: len(blob)] = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
    all_payload = [torch.zeros_like(payload) for _ in range(world)]
    dist.all_gather(all_payload, payload, group=group)
    out: List[Optional[bytes]] = [None] * num_frames
    for r in range(world):
        buf = all_payload[r].cpu().numpy().tobytes()
        pos = 0
        for slot, i in enumerate(frames_for_rank(num_frames, r, world)):
            n = int(all_lens[r][slot].item())
            out[i] = buf[pos:pos + n]
            pos += n
    return out  # type: ignore[return-value]


def reduce_stats(values: Sequence[float], op: str = "sum", group=None,
                 device: Optional[torch.device] = None) -> List[float]:
    """all_reduce of a small float vector (points coded, seconds, bytes ...): op in sum|max|min."""
    t = torch.tensor(list(values), dtype=torch.float64, device=device if device is not None else "cpu")
    if _world(group) > 1:
        rop = {"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op]
        dist.all_reduce(t, op=rop, group=group)
    return [float(v) for v in t.cpu()]


class ShardedSequenceCodec:
    """Codes a sequence frame-parallel: each rank compresses (and optionally decompresses) the frames it
    owns with its own pipelines; `gather` assembles the sequence's bitstreams on every rank."""

    def __init__(self, encoder=None, decoder=None, group=None, device: Optional[torch.device] = None):
        self.encoder, self.decoder, self.group, self.device = encoder, decoder, group, device

    def encode(self, frames: Sequence[dict], quality_index: int = 1, gather: bool = True):
        world, rank = _world(self.group), _rank(self.group)
        local, points = {}, 0
        for i in frames_for_rank(len(frames), rank, world):
            compressed, _ = self.encoder.compress({"frames": [frames[i]], "timestamps": {}})
            local[i] = compressed[quality_index]
            points += len(frames[i]["points"])
        if not gather:
            return local, points
        return gather_bitstreams(local, len(frames), self.group, self.device), points

    def decode(self, bitstreams: Sequence[bytes]) -> Dict[int, list]:
        """Each rank decodes the frames it owns: {frame index: [{"points","colors"}]}."""
        world, rank = _world(self.group), _rank(self.group)
        out = {}
        for i in frames_for_rank(len(bitstreams), rank, world):
            out[i], _ = self.decoder.decompress(bitstreams[i])
        return out
