"""Multi-GPU sharding logic on CPU: world_size 2 over gloo.

There is no collective on the data path, so what needs covering is the partition itself:
  * every rank derives the same file assignment and the same time blocks;
  * the blocks tile the output frames exactly once;
  * the input slice d2p_block_range() hands a block (halo included) really is all the block reads:
    the oracle, run on a payload whose bytes OUTSIDE the slice are replaced by other random bytes,
    reproduces the block's frames bit for bit (the oracle is only the checker here);
  * the pieces gathered on rank 0 equal one pass over the whole payload.
"""
import os
import socket

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import shard
from oracle import orclib

CASES = [
    dict(dsd_rate=4, out_rate=352800, bits=24, channels=2, layout="P", lsb_first=True),
    dict(dsd_rate=4, out_rate=192000, bits=32, channels=2, layout="P", lsb_first=True),
    dict(dsd_rate=8, out_rate=352800, bits=24, channels=6, layout="I", lsb_first=False),
    dict(dsd_rate=1, out_rate=96000, bits=16, channels=2, layout="I", lsb_first=False),
]


def payload(case, blocks=24, ragged=1000):
    n = case["channels"] * (blocks * 4096 + ragged)
    return np.random.default_rng(case["dsd_rate"] * 100 + case["channels"]).integers(0, 256, size=n, dtype=np.uint8)


def block_frames_via_oracle(case, raw, blk, bps):
    """Frames of one time block computed from its input slice only."""
    other = np.random.default_rng(blk.index + 99).integers(0, 256, size=raw.size, dtype=np.uint8)
    other[blk.in_offset:blk.in_offset + blk.in_bytes] = raw[blk.in_offset:blk.in_offset + blk.in_bytes]
    r = orclib.convert(other, want_f64=False, **case)
    fb = case["channels"] * bps
    return r.pcm[blk.first_frame * fb:(blk.first_frame + blk.frame_count) * fb]


def test_assign_files_is_balanced_and_complete():
    sizes = [10, 3, 8, 8, 1, 7, 2, 9]
    parts = shard.assign_files(sizes, 3)
    assert sorted(i for p in parts for i in p) == list(range(len(sizes)))
    loads = [sum(sizes[i] for i in p) for p in parts]
    assert max(loads) - min(loads) <= max(sizes)
    assert shard.assign_files([5] * 512, 8) == [list(range(r, 512, 8)) for r in range(8)]   # config 5: round robin


@pytest.mark.parametrize("case", CASES)
def test_blocks_tile_the_frames_and_slices_are_sufficient(case):
    raw = payload(case)
    cfg = d2p.make_config(**{("filt" if k == "filter" else k): v for k, v in case.items()})
    blocks = shard.split_payload(cfg, raw.size, 5)
    total = d2p.total_frames(cfg, raw.size)
    assert sum(b.frame_count for b in blocks) == total
    assert [b.first_frame for b in blocks] == list(np.cumsum([0] + [b.frame_count for b in blocks[:-1]]))
    whole = orclib.convert(raw, want_f64=False, **case)
    assert whole.frames[0] == total
    bps = d2p.plan(cfg)["bytes_per_sample"]
    pieces = [block_frames_via_oracle(case, raw, b, bps) for b in blocks]
    assert np.array_equal(np.concatenate(pieces), whole.pcm)
    for b in blocks[1:]:
        assert b.in_bytes < raw.size // 2          # a block ships its slice + halo, not the payload


def _worker(rank, world, port, case_idx, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        case = CASES[case_idx]
        raw = payload(case)
        cfg = d2p.make_config(**case)
        bps = d2p.plan(cfg)["bytes_per_sample"]
        blocks = shard.split_payload(cfg, raw.size, 4)
        mine = shard.blocks_of_rank(blocks, rank, world)
        out = [(b.index, block_frames_via_oracle(case, raw, b, bps).tobytes()) for b in mine]
        files = shard.assign_files([7, 7, 3, 9, 1], world)
        gathered = [None] * world
        dist.all_gather_object(gathered, (out, files))
        if rank == 0:
            assert all(g[1] == files for g in gathered)                       # same partition on every rank
            pieces = dict(p for g in gathered for p in g[0])
            got = b"".join(pieces[i] for i in range(len(blocks)))
            whole = orclib.convert(raw, want_f64=False, **case)
            q.put(got == whole.pcm.tobytes() and len(pieces) == len(blocks))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case_idx", [0, 1])
def test_world_size_2_gloo_time_block_shards_reassemble(case_idx):
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, case_idx, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
