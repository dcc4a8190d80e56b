"""d2p_run_noalloc -- the one entry point of the library that never calls the CUDA allocator -- against
d2p_batch_run on the same device-resident payloads.

Who needs it: a host program that owns all of the GPU's memory itself (its own arena, or a fixed budget it keeps
the books for), one that cannot stand the device-wide stall of a cudaMalloc / cudaFree while its other streams are
playing, or one that has to keep converting after the device has filled up.  For them the job table, the filter
look-up table and the statistics record sit in a 256 KiB array compiled into the library instead of in
allocations.  That array is shared by every call on the device and keeps what the last call left in it, so this
test makes several calls IN A ROW on one stream -- two different filters (an integer-ratio one and an L/M one,
whose tables differ in size and layout), a loud payload followed by quiet ones, dither on -- and checks each
against the allocating entry point: same bytes (the bar for integer and f32 PCM everywhere in this suite is
bit-exact), same frame counts, same clip count and peak.  A stale table or a statistics record that was not
cleared between calls shows up as a difference.

The file name sorts last on purpose: pytest -x stops at the first failure, and a failure here must not hide the
parity tests.  Nothing here depends on what ran before it in the process (the first call does the library's
one-time look-up of its array, wherever it comes), so it passes alone as well as at the end of the whole run.
"""
import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi

pytestmark = pytest.mark.gpu


def payload(seed, channels, blocks, loud=False):
    rng = np.random.default_rng(seed)
    raw = rng.integers(0, 256, size=channels * blocks * 4096, dtype=np.uint8)
    if loud:                       # long runs of ones: a full-scale low-frequency square wave, clips at +9 dB
        raw.reshape(-1, 4096)[:, 1024:3072] = 0xFF
    return raw


def out_bytes(cfg, raw):
    bps = 3 if cfg.bits == 20 else cfg.bits // 8
    return d2p.frames_for_bytes(cfg, raw.size // cfg.channels) * cfg.channels * bps


def both_ways(torch, cfg, raws, seeds):
    """-> (pcm list, frames, stats) from d2p_batch_run and from d2p_run_noalloc, same inputs, same stream."""
    stream = torch.cuda.current_stream().cuda_stream
    d_in = [torch.from_numpy(r).cuda() for r in raws]
    sizes = [out_bytes(cfg, r) for r in raws]

    def items(outs):
        return [dict(d_in=i.data_ptr(), in_bytes=i.numel(), d_out=o.data_ptr(), out_cap_bytes=o.numel(), dither_seed=s)
                for i, o, s in zip(d_in, outs, seeds)]

    want_out = [torch.zeros(n + 16, dtype=torch.uint8, device="cuda") for n in sizes]
    batch = d2p.Batch(cfg)
    want_frames = batch.run(batch.make_jobs(items(want_out)), stream)
    want_stats = batch.stats(stream)
    batch.close()

    # 0xEE behind the last frame must survive: the static-scratch path writes no more than the allocating one
    got_out = [torch.full((n + 16,), 0xEE, dtype=torch.uint8, device="cuda") for n in sizes]
    got_frames, got_stats = d2p.run_noalloc(cfg, items(got_out), stream)
    torch.cuda.synchronize()
    want = [o.cpu().numpy()[:n] for o, n in zip(want_out, sizes)]
    got = [o.cpu().numpy() for o in got_out]
    for g, n in zip(got, sizes):
        assert np.all(g[n:] == 0xEE)
    return want, [g[:n] for g, n in zip(got, sizes)], want_frames, got_frames, want_stats, got_stats


def test_calls_in_a_row_match_the_allocating_entry_point():
    torch = pytest.importorskip("torch")
    calls = [
        # integer-ratio filter, loud input at +9 dB: many clipped samples and a large peak are left in the scratch
        (d2p.make_config(dsd_rate=4, out_rate=352800, bits=24, level_db=9.0), [payload(1, 2, 6, loud=True)], [0]),
        # L/M filter (another table, another layout), quiet input: statistics must not carry over
        (d2p.make_config(dsd_rate=4, out_rate=192000, bits=32), [payload(2, 2, 7), payload(3, 2, 3)], [0, 0]),
        # the first configuration again, now with TPDF dither and three ragged files in one call
        (d2p.make_config(dsd_rate=4, out_rate=352800, bits=24, dither="T"),
         [payload(4, 2, 5), payload(5, 2, 1)[:5000], payload(6, 2, 9)], [11, 12, 13]),
        # a small filter (DSD64 8:1, 16-bit) after the large ones: the tail of the old table stays behind it in the array
        (d2p.make_config(dsd_rate=1, out_rate=352800, bits=16), [payload(7, 2, 4)], [0]),
    ]
    clipped_seen = []
    for cfg, raws, seeds in calls:
        want, got, want_frames, got_frames, want_stats, got_stats = both_ways(torch, cfg, raws, seeds)
        assert got_frames == want_frames
        for w, g in zip(want, got):
            assert w.size > 0 and np.array_equal(w, g)            # bit-exact, integer and f32 alike
        assert got_stats == want_stats
        clipped_seen.append(got_stats["clipped_samples"])
    assert clipped_seen[0] > 0 and clipped_seen[1] == 0           # the loud call clipped; nothing leaked into the next


def test_limits_are_refused_with_a_reason():
    torch = pytest.importorskip("torch")
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800)
    raw = payload(8, 2, 2)
    d_in = torch.from_numpy(raw).cuda()
    d_out = torch.zeros(out_bytes(cfg, raw) + 16, dtype=torch.uint8, device="cuda")
    item = dict(d_in=d_in.data_ptr(), in_bytes=d_in.numel(), d_out=d_out.data_ptr(), out_cap_bytes=d_out.numel())
    with pytest.raises(d2p.D2PError) as e:
        d2p.run_noalloc(cfg, [item] * (_capi.NOALLOC_MAX_JOBS + 1))
    assert e.value.status == _capi.ERR_CAPACITY and "64" in str(e.value)
    with pytest.raises(d2p.D2PError) as e:
        d2p.run_noalloc(d2p.make_config(dsd_rate=4, out_rate=352800, bits=32, dither="F"), [item])
    assert e.value.status == _capi.ERR_INVALID_ARG and "dither F" in str(e.value)
    # the full table is fine, and an empty call is a no-op that reports clean statistics
    outs = [torch.zeros(d_out.numel(), dtype=torch.uint8, device="cuda") for _ in range(_capi.NOALLOC_MAX_JOBS)]
    frames, _ = d2p.run_noalloc(cfg, [dict(item, d_out=o.data_ptr()) for o in outs])
    torch.cuda.synchronize()
    assert len(set(frames)) == 1 and all(torch.equal(outs[0], o) for o in outs[1:]) and int(outs[0].count_nonzero()) > 0
    frames, stats = d2p.run_noalloc(cfg, [])
    assert frames == [] and stats == {"clipped_samples": 0, "peak_abs": 0.0}
