"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same inputs.

Bit-exact for the pre-scale f64 samples and for integer / f32 PCM with dither off; exact against the
oracle's Philox dither with dither on, and within +-1 LSB of the undithered output.
"""
import hashlib

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi
from oracle import orclib

pytestmark = pytest.mark.gpu

# every (dsd_rate, out_rate, filter) the reference resolves on the integer path (A2) ...
INT_COMBOS = [
    (1, 352800, "D"), (1, 352800, "X"), (1, 352800, "E"), (2, 705600, "E"), (4, 1411200, "E"),
    (1, 176400, "X"), (1, 176400, "E"), (2, 352800, "E"), (2, 352800, "C"), (2, 352800, "X"), (4, 705600, "E"),
    (1, 88200, "X"), (1, 88200, "E"), (2, 176400, "E"), (2, 176400, "C"), (4, 352800, "E"), (8, 705600, "E"),
    (2, 88200, "E"), (2, 88200, "C"), (4, 176400, "E"), (8, 352800, "E"), (1, 44100, "E"),
    (4, 88200, "E"),
]
# ... and every L/M topology (A6), including the unguarded DSD512 rows (F8)
LM_COMBOS = [(r, o) for r in (1, 2, 4, 8) for o in (96000, 192000, 384000)]


def run_gpu(raw, chunks=None, **kw):
    cfg = d2p.make_config(keep_f64=True, **kw)
    conv = d2p.Converter(cfg)
    pcm, f64 = [], [[] for _ in range(cfg.channels)]
    pieces = [raw] if chunks is None else np.split(raw, chunks)
    for i, piece in enumerate(pieces):
        if piece.size:
            conv.push(piece)
        if i == len(pieces) - 1:
            conv.finish()
        for c in range(cfg.channels):
            f64[c].append(conv.peek_f64(c))
        pcm.append(conv.pull())
    st = conv.stats()
    conv.close()
    if cfg.out_format == _capi.OUT_PLANAR_32:
        pcm = np.concatenate(pcm, axis=1)
    else:
        pcm = np.concatenate(pcm)
    return pcm, np.stack([np.concatenate(x) for x in f64]), st


def rand_payload(seed, channels, blocks, block=4096):
    return np.random.default_rng(seed).integers(0, 256, size=channels * blocks * block, dtype=np.uint8)


@pytest.mark.parametrize("dsd_rate,out_rate,filt", INT_COMBOS)
def test_integer_path_bit_exact(dsd_rate, out_rate, filt):
    raw = rand_payload(dsd_rate * 1000 + out_rate % 997, 2, 5)
    ref = orclib.convert(raw, dsd_rate=dsd_rate, out_rate=out_rate, filt=filt, bits=24, channels=2)
    pcm, f64, _ = run_gpu(raw, dsd_rate=dsd_rate, out_rate=out_rate, filt=filt, bits=24, channels=2)
    assert f64.shape == ref.f64.shape
    assert np.array_equal(f64, ref.f64)
    assert np.array_equal(pcm, ref.pcm)


@pytest.mark.parametrize("dsd_rate,out_rate", LM_COMBOS)
def test_lm_path_bit_exact(dsd_rate, out_rate):
    raw = rand_payload(dsd_rate * 7 + out_rate % 991, 2, 6)
    ref = orclib.convert(raw, dsd_rate=dsd_rate, out_rate=out_rate, bits=24, channels=2)
    pcm, f64, _ = run_gpu(raw, dsd_rate=dsd_rate, out_rate=out_rate, bits=24, channels=2)
    assert f64.shape == ref.f64.shape
    assert np.array_equal(f64, ref.f64)
    assert np.array_equal(pcm, ref.pcm)


@pytest.mark.parametrize("bits", [16, 20, 24, 32])
@pytest.mark.parametrize("out_rate", [352800, 192000])
@pytest.mark.parametrize("level_db", [0.0, -4.0, 9.0])
def test_bit_depths_levels_and_clipping(bits, out_rate, level_db):
    raw = rand_payload(bits + out_rate % 13, 2, 4)
    # a loud low-frequency square wave on top of noise makes +9 dB clip in runs
    raw[: raw.size // 2] = np.repeat(np.array([0xFF, 0xFF, 0x00, 0x00] * 8, dtype=np.uint8), raw.size // 64)[: raw.size // 2]
    ref = orclib.convert(raw, dsd_rate=1, out_rate=out_rate, bits=bits, level_db=level_db, channels=2)
    pcm, f64, st = run_gpu(raw, dsd_rate=1, out_rate=out_rate, bits=bits, level_db=level_db, channels=2)
    assert np.array_equal(pcm, ref.pcm)
    if bits != 32:
        assert st["clips"] == ref.clips
    assert st["peak_abs"] == np.abs(ref.f64 * ref.scale).max()


@pytest.mark.parametrize("out_format", [_capi.OUT_INTERLEAVED_BE, _capi.OUT_PLANAR_32])
@pytest.mark.parametrize("bits", [16, 20, 24, 32])
def test_output_containers(out_format, bits):
    raw = rand_payload(99 + bits, 2, 3)
    pcm, _, _ = run_gpu(raw, dsd_rate=1, out_rate=352800, bits=bits, channels=2, out_format=out_format)
    ref = orclib.convert(raw, dsd_rate=1, out_rate=352800, bits=bits, channels=2, stdout_mode=False)
    if out_format == _capi.OUT_PLANAR_32:
        assert np.array_equal(pcm, ref.planar)          # AudioFile samples: 20-bit values NOT shifted
    else:
        # AIFF/AIFC payload: big-endian, 20-bit left-aligned in 3 bytes (rd/audio_file.rs:257-275)
        q = np.ascontiguousarray(ref.planar.T)
        if bits == 32:
            exp = np.ascontiguousarray(q, dtype=">f4").view(np.uint8).reshape(-1)
        elif bits == 16:
            exp = q.astype(">i2").view(np.uint8).reshape(-1)
        else:
            v = (q << 4 if bits == 20 else q).astype(np.uint32)
            exp = np.stack([(v >> 16) & 0xFF, (v >> 8) & 0xFF, v & 0xFF], axis=-1).astype(np.uint8).reshape(-1)
        assert np.array_equal(pcm, exp)


@pytest.mark.parametrize("channels,layout,lsb_first", [(1, "P", True), (2, "I", False), (6, "I", False), (3, "P", False), (2, "I", True)])
@pytest.mark.parametrize("out_rate", [352800, 96000])
def test_input_layouts(channels, layout, lsb_first, out_rate):
    raw = rand_payload(channels * 31 + out_rate % 7, channels, 4)
    ref = orclib.convert(raw, dsd_rate=1, out_rate=out_rate, channels=channels, layout=layout, lsb_first=lsb_first)
    pcm, f64, _ = run_gpu(raw, dsd_rate=1, out_rate=out_rate, channels=channels, layout=layout, lsb_first=lsb_first)
    assert np.array_equal(f64, ref.f64)
    assert np.array_equal(pcm, ref.pcm)


@pytest.mark.parametrize("out_rate", [352800, 88200, 96000, 384000])
@pytest.mark.parametrize("layout", ["P", "I"])
def test_streaming_chunking_equals_one_shot(out_rate, layout):
    """State persists across pushes (8b): ragged pushes == one push == oracle."""
    raw = rand_payload(5 + out_rate % 11, 2, 6)
    ref = orclib.convert(raw, dsd_rate=1, out_rate=out_rate, channels=2, layout=layout)
    cuts = [1, 17, 4096, 8191, 8192 * 2 + 5, 40000]
    pcm, f64, _ = run_gpu(raw, chunks=cuts, dsd_rate=1, out_rate=out_rate, channels=2, layout=layout)
    assert np.array_equal(f64, ref.f64)
    assert np.array_equal(pcm, ref.pcm)


@pytest.mark.parametrize("nbytes", [0, 2, 100, 4096 * 2 - 2, 4096 * 2 * 3 + 2 * 1000])
@pytest.mark.parametrize("out_rate", [352800, 96000])
def test_empty_and_ragged_inputs(nbytes, out_rate):
    raw = rand_payload(3, 2, 4)[:nbytes]
    ref = orclib.convert(raw, dsd_rate=1, out_rate=out_rate, channels=2)
    pcm, f64, _ = run_gpu(raw, dsd_rate=1, out_rate=out_rate, channels=2)
    assert f64.shape[1] == ref.frames[0]
    assert np.array_equal(pcm, ref.pcm)


@pytest.mark.parametrize("dither", ["T", "R"])
@pytest.mark.parametrize("out_rate", [88200, 192000])
def test_dither_exact_and_within_one_lsb(dither, out_rate):
    raw = rand_payload(77, 2, 5)
    kw = dict(dsd_rate=2, out_rate=out_rate, bits=24, channels=2)
    clean = orclib.convert(raw, dither="X", stdout_mode=False, **kw).planar
    ref = orclib.convert(raw, dither=dither, dither_seed=1234, stdout_mode=False, **kw).planar
    got, _, _ = run_gpu(raw, dither=dither, dither_seed=1234, out_format=_capi.OUT_PLANAR_32, **kw)
    assert np.array_equal(got, ref)                                  # same counter-based noise definition
    assert np.abs(got.astype(np.int64) - clean).max() <= 1           # F5: |dithered - undithered| <= 1 LSB
    other, _, _ = run_gpu(raw, dither=dither, dither_seed=99, out_format=_capi.OUT_PLANAR_32, **kw)
    assert not np.array_equal(other, got)


def test_dither_scale_env_equivalent():
    raw = rand_payload(78, 2, 3)
    kw = dict(dsd_rate=1, out_rate=352800, bits=16, channels=2, dither="T", dither_seed=5, dither_scale_db=6.0)
    ref = orclib.convert(raw, **kw)
    pcm, _, _ = run_gpu(raw, **kw)
    assert np.array_equal(pcm, ref.pcm)


def test_fixture_excerpts_match_oracle_and_golden_digests(fixtures, golden_digests):
    for name, (raw, meta) in fixtures.items():
        for out_rate, bits in ((352800, 24), (176400, 16), (96000, 24), (192000, 32)):
            kw = dict(dsd_rate=meta["dsd_rate"], out_rate=out_rate, bits=bits, channels=meta["channels"],
                      layout=meta["layout"], lsb_first=meta["lsb_first"])
            pcm, f64, _ = run_gpu(raw, **kw)
            g = golden_digests[f"{name}/{out_rate}/{bits}"]
            assert f64.shape[1] == g["frames"]
            assert hashlib.sha256(pcm.tobytes()).hexdigest() == g["pcm_sha256"], (name, out_rate, bits)
            assert hashlib.sha256(f64.tobytes()).hexdigest() == g["f64_sha256"], (name, out_rate, bits)


def test_impulse_known_answer(fixtures):
    """KAT: one +2 impulse on the idle pattern reproduces 2*h of the tap table; idle gives exactly 0.0."""
    raw, meta = fixtures["impulse_mono_toggle"]
    _, f64, _ = run_gpu(raw, dsd_rate=1, out_rate=352800, channels=1)
    half = d2p.tap_table("HTAPS_DSD64_8TO1_EQ")
    nz = np.nonzero(f64[0])[0]
    assert nz.min() == 1995 and nz.max() == 2007
    full = np.concatenate([half[::-1], half])
    # byte 2048 bit 0 flips -1 -> +1: frame n sees it at tap index 8*(n+h-1+1) - 8*nt... -> decimated taps
    got = f64[0][1995:2008]
    best = min(np.abs(got - 2 * full[o::8][:13]).max() for o in range(8) if full[o::8].size >= 13)
    assert best < 1e-16


def test_convert_host_pipelined_matches_streaming():
    raw = rand_payload(11, 2, 40)
    for out_rate in (352800, 192000):
        cfg = d2p.make_config(dsd_rate=1, out_rate=out_rate, channels=2)
        pcm, frames, _ = d2p.convert_host(cfg, raw)
        ref = orclib.convert(raw, dsd_rate=1, out_rate=out_rate, channels=2, want_f64=False)
        assert frames == ref.frames[0]
        assert np.array_equal(pcm, ref.pcm)


def test_batch_of_device_resident_files_matches_oracle():
    torch = pytest.importorskip("torch")
    nfiles = 5
    raws = [rand_payload(200 + i, 2, 3 + i) for i in range(nfiles)]
    cfg = d2p.make_config(dsd_rate=4, out_rate=352800, channels=2)
    batch = d2p.Batch(cfg)
    d_in = [torch.from_numpy(r).cuda() for r in raws]
    caps = [d2p.frames_for_bytes(cfg, r.size // 2) * 6 for r in raws]
    d_out = [torch.zeros(max(c, 16), dtype=torch.uint8, device="cuda") for c in caps]
    jobs = batch.make_jobs(dict(d_in=i.data_ptr(), in_bytes=i.numel(), d_out=o.data_ptr(), out_cap_bytes=o.numel())
                           for i, o in zip(d_in, d_out))
    frames = batch.run(jobs, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert batch.launches() == 1
    for r, o, f, cap in zip(raws, d_out, frames, caps):
        ref = orclib.convert(r, dsd_rate=4, out_rate=352800, channels=2, want_f64=False)
        assert f == ref.frames[0]
        assert np.array_equal(o.cpu().numpy()[:cap], ref.pcm)
    batch.close()


def test_time_block_split_is_exact():
    """8e: a long stream cut into halo-padded time blocks gives the same bytes as one pass
    (size-independent property; 64 MiB-scale input is exercised by bench.py)."""
    raw = rand_payload(321, 2, 64)
    cfg = d2p.make_config(dsd_rate=4, out_rate=176400, bits=20, channels=2)
    whole, frames, _ = d2p.convert_host(cfg, raw)
    conv = d2p.Converter(cfg)
    for piece in np.array_split(raw, 13):
        conv.push(piece)
    conv.finish()
    parts = conv.pull()
    conv.close()
    assert np.array_equal(whole, parts)


@pytest.mark.parametrize("kw", [
    dict(dsd_rate=4, out_rate=352800, bits=24, channels=2, layout="P", lsb_first=True),
    dict(dsd_rate=4, out_rate=192000, bits=32, channels=2, layout="P", lsb_first=True),
    dict(dsd_rate=8, out_rate=352800, bits=24, channels=6, layout="I", lsb_first=False),
    dict(dsd_rate=2, out_rate=96000, bits=20, channels=3, layout="I", lsb_first=True, dither="T"),
])
def test_time_blocks_on_device_equal_one_pass(kw):
    """8e through the C ABI: d2p_block_range + d2p_batch_run_blocks; every block gets ONLY its slice."""
    torch = pytest.importorskip("torch")
    from dsd2dxd_b200 import shard
    C = kw["channels"]
    raw = np.random.default_rng(17 + C).integers(0, 256, size=C * (30 * 4096 + 777), dtype=np.uint8)
    cfg = d2p.make_config(dither_seed=4242, **kw)
    bps = d2p.plan(cfg)["bytes_per_sample"]
    ref = orclib.convert(raw, want_f64=False, dither_seed=4242, **kw)
    blocks = shard.split_payload(cfg, raw.size, 7)
    batch = d2p.Batch(cfg)
    d_in = [torch.from_numpy(raw[b.in_offset:b.in_offset + b.in_bytes].copy()).cuda() for b in blocks]
    d_out = [torch.zeros(b.frame_count * C * bps + 16, dtype=torch.uint8, device="cuda") for b in blocks]
    batch.run_blocks([dict(d_in=i.data_ptr(), in_offset=b.in_offset, in_bytes=b.in_bytes, payload_bytes=raw.size,
                           first_frame=b.first_frame, frame_count=b.frame_count, d_out=o.data_ptr(),
                           out_cap_bytes=o.numel(), dither_seed=4242) for b, i, o in zip(blocks, d_in, d_out)],
                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = np.concatenate([o.cpu().numpy()[:b.frame_count * C * bps] for b, o in zip(blocks, d_out)])
    assert np.array_equal(got, ref.pcm)
    batch.close()


@pytest.mark.parametrize("out_rate,bits", [(352800, 32), (352800, 24), (88200, 16), (192000, 32), (96000, 24), (384000, 32)])
@pytest.mark.parametrize("layout,channels", [("P", 2), ("I", 3)])
def test_fpd_dither_matches_reference_order(out_rate, bits, layout, channels):
    """Dither F (A15): one xorshift32 stream per file in reader order (block, channel, frame) -- the GPU jumps
    to every run's state; one-shot, ragged streaming pushes and the pipelined host path all agree with the oracle."""
    raw = rand_payload(out_rate % 89 + bits, channels, 5)[: channels * (4 * 4096 + 1234)]
    kw = dict(dsd_rate=1, out_rate=out_rate, bits=bits, channels=channels, layout=layout, lsb_first=(layout == "P"),
              dither="F", dither_seed=0xABCDEF12345)
    ref = orclib.convert(raw, want_f64=False, **kw)
    pcm, _, _ = run_gpu(raw, **kw)
    assert np.array_equal(pcm, ref.pcm)
    cuts = [5, 4096 * channels + 7, 9000 * channels, 3 * 4096 * channels - 1]
    pcm2, _, _ = run_gpu(raw, chunks=cuts, **kw)
    assert np.array_equal(pcm2, ref.pcm)
    cfg = d2p.make_config(**kw)
    pcm3, frames, _ = d2p.convert_host(cfg, raw)
    assert frames == ref.frames[0] and np.array_equal(pcm3, ref.pcm)
    clean = orclib.convert(raw, want_f64=False, **dict(kw, dither="X")).pcm
    assert not np.array_equal(clean, ref.pcm) or bits != 32        # the float dither really changes f32 output


def test_fpd_dither_time_blocks_and_long_jump():
    """Blocks of one file converted independently still see the file's single noise stream."""
    torch = pytest.importorskip("torch")
    from dsd2dxd_b200 import shard
    kw = dict(dsd_rate=4, out_rate=352800, bits=32, channels=2, layout="P", lsb_first=True, dither="F", dither_seed=99)
    raw = rand_payload(4242, 2, 70)
    cfg = d2p.make_config(**kw)
    ref = orclib.convert(raw, want_f64=False, **kw)
    blocks = shard.split_payload(cfg, raw.size, 5)
    batch = d2p.Batch(cfg)
    d_in = [torch.from_numpy(raw[b.in_offset:b.in_offset + b.in_bytes].copy()).cuda() for b in blocks]
    d_out = [torch.zeros(b.frame_count * 8 + 16, dtype=torch.uint8, device="cuda") for b in blocks]
    batch.run_blocks([dict(d_in=i.data_ptr(), in_offset=b.in_offset, in_bytes=b.in_bytes, payload_bytes=raw.size,
                           first_frame=b.first_frame, frame_count=b.frame_count, d_out=o.data_ptr(),
                           out_cap_bytes=o.numel(), dither_seed=99) for b, i, o in zip(blocks, d_in, d_out)],
                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = np.concatenate([o.cpu().numpy()[:b.frame_count * 8] for b, o in zip(blocks, d_out)])
    assert np.array_equal(got, ref.pcm)
    batch.close()


@pytest.mark.parametrize("dsd_rate,out_rate,channels,layout,lsb_first", [
    (1, 352800, 2, "P", True), (4, 352800, 2, "P", True), (1, 96000, 2, "P", True), (4, 192000, 1, "P", True),
    (8, 352800, 6, "I", False), (2, 88200, 2, "I", True)])
def test_level_check_matches_process_levels(dsd_rate, out_rate, channels, layout, lsb_first):
    """dsd_levels path: peak over all channels, no PCM written (D2P_OUT_NONE)."""
    raw = rand_payload(900 + dsd_rate + channels, channels, 40)
    raw[: raw.size // 3] = 0x69                                    # a quiet stretch in front, peaks later
    cfg = d2p.make_config(dsd_rate=dsd_rate, out_rate=out_rate, channels=channels, layout=layout, lsb_first=lsb_first,
                          bits=16, dither="T", level_db=-3.0, filt="E")     # all ignored, as in new_level_check
    peak, db = d2p.check_level(cfg, raw)
    ref_peak, ref_db = orclib.check_level(raw, dsd_rate=dsd_rate, out_rate=out_rate, channels=channels, layout=layout,
                                          lsb_first=lsb_first)
    assert peak == ref_peak
    assert abs(float(db) - float(ref_db)) <= abs(float(np.spacing(ref_db)))


def test_level_check_device_resident_batch_and_silence():
    import torch
    cfg = d2p.make_config(dsd_rate=1, out_rate=352800, bits=32, dither="X", out_format=_capi.OUT_NONE)
    raws = [rand_payload(950 + i, 2, 12) for i in range(3)]
    batch = d2p.Batch(cfg)
    d_in = [torch.from_numpy(r).cuda() for r in raws]
    jobs = batch.make_jobs(dict(d_in=t.data_ptr(), in_bytes=t.numel(), d_out=None, out_cap_bytes=0) for t in d_in)
    frames = batch.run(jobs, torch.cuda.current_stream().cuda_stream)
    st = batch.stats(torch.cuda.current_stream().cuda_stream)
    batch.close()
    refs = [orclib.check_level(r)[0] for r in raws]
    assert frames == [d2p.frames_for_bytes(cfg, r.size // 2) for r in raws]
    assert st["peak_abs"] == max(refs)
    # nothing in, nothing out: -inf like process_levels
    peak, db = d2p.check_level(cfg, np.zeros(0, dtype=np.uint8))
    assert peak == 0.0 and np.isneginf(db)
    with pytest.raises(_capi.D2PError):
        d2p.Converter(cfg)                                         # the streaming context has no level-only mode


@pytest.mark.parametrize("threshold", [0, 20000, None])
def test_block_by_block_pushes_are_coalesced_and_exact(threshold):
    """The reference's own call pattern: one 4096-byte block per channel per call (process_blocks).  Small pushes
    are gathered in pinned memory and shipped per threshold / flush / finish; the frames are the same."""
    raw = rand_payload(77, 2, 40)
    ref = orclib.convert(raw, dsd_rate=1, out_rate=176400, bits=24, channels=2)
    cfg = d2p.make_config(dsd_rate=1, out_rate=176400, bits=24, channels=2)
    conv = d2p.Converter(cfg)
    if threshold is not None:
        conv.set_push_threshold(threshold)
    pcm, ready_after_push = [], []
    for off in range(0, raw.size, 4096):
        conv.push(raw[off:off + 4096])
        ready_after_push.append(conv.frames_ready())
        if off == 20 * 8192:
            conv.flush()
            assert conv.frames_ready() > 0 or pcm                  # everything pushed so far is on its way
        pcm.append(conv.pull())
    conv.finish()
    pcm.append(conv.pull())
    conv.close()
    assert np.array_equal(np.concatenate(pcm), ref.pcm)
    if threshold == 0:
        assert all(r > 0 for r in ready_after_push[3::2])          # every completed block group yields frames at once
    if threshold is None:
        assert sum(1 for r in ready_after_push if r > 0) <= 2      # 320 KiB < 8 MiB: only the explicit flush ships


def test_randomised_configurations_match_oracle():
    """Seeded sweep over the option space (rate pair, filter, bits, dither, level, layout, bit order, channel count,
    block size, payload length incl. ragged tails, push chunking): one-shot host conversion and the streaming context
    both equal the oracle byte for byte."""
    rng = np.random.default_rng(20261104)
    valid = [(r, o, f) for (r, o, f) in INT_COMBOS] + [(r, o, "E") for (r, o) in LM_COMBOS]
    done = 0
    for case in range(48):
        dsd_rate, out_rate, filt = valid[rng.integers(len(valid))]
        layout = "PI"[rng.integers(2)]
        channels = int(rng.integers(1, 7))
        block = int(rng.choice([4096, 4096, 2048, 8192])) if layout == "P" else 4096
        bits = int(rng.choice([16, 20, 24, 32]))
        dither = str(rng.choice(["X", "X", "T", "R", "F"]))
        lsb = bool(rng.integers(2))
        level = float(rng.choice([0.0, 0.0, -6.0, 3.5]))
        group = channels * (block if layout == "P" else 1)
        nbytes = int(rng.integers(2, 9)) * channels * block + int(rng.integers(0, 3)) * int(rng.integers(0, group))
        raw = rng.integers(0, 256, size=nbytes, dtype=np.uint8)
        kw = dict(dsd_rate=dsd_rate, out_rate=out_rate, filt=filt, bits=bits, dither=dither, level_db=level,
                  channels=channels, layout=layout, lsb_first=lsb, block_size=block, dither_seed=1000 + case)
        try:
            ref = orclib.convert(raw, want_f64=False, **kw)
        except ValueError:
            continue                                                 # a combination the reference rejects as well
        cfg = d2p.make_config(**kw)
        pcm, frames, _ = d2p.convert_host(cfg, raw)
        assert frames == ref.frames[0], kw
        assert np.array_equal(pcm, ref.pcm), kw
        conv = d2p.Converter(cfg)
        conv.set_push_threshold(int(rng.choice([0, 5000, 1 << 20])))
        cuts = np.sort(rng.integers(0, nbytes + 1, size=int(rng.integers(0, 6))))
        got = []
        for piece in np.split(raw, cuts):
            if piece.size:
                conv.push(piece)
            got.append(conv.pull())
        conv.finish()
        got.append(conv.pull())
        conv.close()
        assert np.array_equal(np.concatenate(got), ref.pcm), kw
        done += 1
    assert done >= 40


@pytest.mark.parametrize("dither", ["X", "T"])
def test_fast_writer_statistics_ignore_lanes_past_a_ragged_tile(dither):
    """The specialised 3-byte writer runs whole frame sets; frames past the end of the last (ragged) tile are computed
    along but must reach neither the PCM nor the peak / clip statistics."""
    raw = rand_payload(4321, 2, 301)[: 2 * (4096 * 300 + 1234)]
    kw = dict(dsd_rate=4, out_rate=176400, bits=24, channels=2, dither=dither, dither_seed=3)
    cfg = d2p.make_config(**kw)
    conv = d2p.Converter(cfg)                                    # no f64 side output: Emitter<1, .>
    conv.push(raw)
    conv.finish()
    pcm = conv.pull()
    st = conv.stats()
    conv.close()
    ref = orclib.convert(raw, **kw)
    assert np.array_equal(pcm, ref.pcm)
    assert st["peak_abs"] == np.abs(ref.f64 * ref.scale).max() and st["clipped_samples"] == 0 and st["clips"] == ref.clips


@pytest.mark.parametrize("in_off,out_off", [(0, 0), (1, 0), (0, 2), (7, 9), (16, 16)])
@pytest.mark.parametrize("kw", [dict(dsd_rate=4, out_rate=352800, bits=24, channels=2, dither="T"),
                                dict(dsd_rate=1, out_rate=96000, bits=16, channels=2)])
def test_unaligned_device_buffers_take_the_register_paths(in_off, out_off, kw):
    """TMA bulk copies need 16-byte aligned addresses on both sides: a payload or an output buffer that is not aligned
    must fall back to the per-thread loads / stores (per job, inside the same launch as aligned jobs) and give the same bytes."""
    torch = pytest.importorskip("torch")
    raws = [rand_payload(700 + i, 2, 40 + 3 * i) for i in range(3)]
    cfg = d2p.make_config(dither_seed=5, **kw)
    bps = d2p.plan(cfg)["bytes_per_sample"]
    batch = d2p.Batch(cfg)
    offs = [(in_off, out_off), (0, 0), (in_off, 0)]                  # aligned and unaligned jobs side by side
    d_in = [torch.zeros(r.size + 32, dtype=torch.uint8, device="cuda") for r in raws]
    for t, r, (a, _) in zip(d_in, raws, offs):
        t[a:a + r.size] = torch.from_numpy(r).cuda()
    nbytes = [d2p.total_frames(cfg, r.size) * 2 * bps for r in raws]
    d_out = [torch.zeros(n + 64, dtype=torch.uint8, device="cuda") for n in nbytes]
    jobs = batch.make_jobs(dict(d_in=t.data_ptr() + a, in_bytes=r.size, d_out=o.data_ptr() + b, out_cap_bytes=n + 16, dither_seed=5)
                           for t, r, o, n, (a, b) in zip(d_in, raws, d_out, nbytes, offs))
    batch.run(jobs, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for r, o, n, (_, b) in zip(raws, d_out, nbytes, offs):
        ref = orclib.convert(r, want_f64=False, dither_seed=5, **kw)
        assert np.array_equal(o.cpu().numpy()[b:b + n], ref.pcm)
    batch.close()
