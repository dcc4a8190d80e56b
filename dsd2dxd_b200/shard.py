"""Shard maths for multi-GPU conversion: no collective on the data path.

Two levels, both pure host logic on top of the C ABI's planning calls:
  * files of a batch are balanced over ranks (largest first onto the least loaded rank) -- the reference
    hands one file to each Rayon worker (src/main.rs:280-300); here a rank is a GPU;
  * one long payload is cut in OUTPUT-FRAME space into time blocks; d2p_block_range() gives each block
    the input slice it reads (filter halo included), so blocks are independent and bit-identical to a
    single pass.
Nothing here computes samples.
"""
from dataclasses import dataclass

from . import block_range, total_frames


def assign_files(sizes, world_size):
    """-> list (per rank) of file indices; deterministic on every rank."""
    order = sorted(range(len(sizes)), key=lambda i: (-sizes[i], i))
    load = [0] * world_size
    out = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += sizes[i]
    return [sorted(x) for x in out]


@dataclass(frozen=True)
class TimeBlock:
    index: int
    first_frame: int
    frame_count: int
    in_offset: int      # byte offset of the input slice in the container-order payload
    in_bytes: int


def split_payload(cfg, payload_bytes, nblocks, align_frames=1):
    """Cut a payload into `nblocks` contiguous time blocks of (nearly) equal frame counts."""
    total = total_frames(cfg, payload_bytes)
    per = -(-total // nblocks)
    per = -(-per // align_frames) * align_frames
    blocks, f = [], 0
    for b in range(nblocks):
        n = max(0, min(per, total - f))
        off, ln = block_range(cfg, payload_bytes, f, n) if n else (0, 0)
        blocks.append(TimeBlock(b, f, n, off, ln))
        f += n
    return blocks


def blocks_of_rank(blocks, rank, world_size):
    """Contiguous run of blocks per rank (keeps each rank's output one contiguous frame range)."""
    per = -(-len(blocks) // world_size)
    return blocks[rank * per:(rank + 1) * per]
