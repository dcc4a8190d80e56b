"""FLAC back end: d2p_write_flac_file against an independent decoder (oracle/flac_decode.py).

The reference (rd/audio_file.rs:454-573) delegates the bitstream to the flac-codec crate, so the parity contract is
decode equality: the integers that come back are the ones save_flac_file would have fed its encoder -- 16/24-bit as
is, 20-bit unshifted at bits_per_sample 20, float32 as (clamp(x,-1,1) * 8388607.0f) truncated -- plus a structurally
valid stream (sync, frame numbers, CRC-8, CRC-16, STREAMINFO totals and MD5)."""
import os

import re

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from oracle import flac_decode


def pack(values, bits):
    """[frames, channels] integers (or floats for bits == 32) -> the little-endian interleaved bytes the kernels emit."""
    v = np.asarray(values)
    if bits == 32:
        return v.astype("<f4").tobytes()
    if bits == 16:
        return v.astype("<i2").tobytes()
    w = v.astype(np.int64) << (4 if bits == 20 else 0)
    b = np.stack([w & 0xFF, (w >> 8) & 0xFF, (w >> 16) & 0xFF], axis=-1).astype(np.uint8)
    return b.tobytes()


def expected(values, bits):
    v = np.asarray(values)
    if bits == 32:
        return (np.clip(v.astype(np.float32), np.float32(-1), np.float32(1)) * np.float32(8388607.0)).astype(np.int32)
    return v.astype(np.int64)


def signal(kind, frames, channels, bits, rng):
    sbits = 24 if bits == 32 else bits
    full = (1 << (sbits - 1)) - 1
    t = np.arange(frames)[:, None] + np.arange(channels)[None, :] * 17
    if kind == "sine":
        x = np.round(0.6 * full * np.sin(2 * np.pi * t / 97.3) + 0.1 * full * np.sin(2 * np.pi * t / 11.1))
    elif kind == "noise":
        x = rng.integers(-full - 1, full + 1, size=(frames, channels))
    elif kind == "walk":
        x = np.clip(np.cumsum(rng.integers(-300, 301, size=(frames, channels)), axis=0), -full - 1, full)
    elif kind == "const":
        x = np.full((frames, channels), -12345 % full)
    elif kind == "silence":
        x = np.zeros((frames, channels))
    else:                                               # rails: alternating extremes, the largest residuals there are
        x = np.where((t // 3) % 2 == 0, full, -full - 1)
    x = x.astype(np.int64)
    if bits == 32:
        return (x / 8388607.0).astype(np.float32) * np.float32(1.0)
    return x


def roundtrip(tmp_path, values, bits, channels, rate, comments=()):
    path = tmp_path / "t.flac"
    frames = len(values)
    d2p.write_flac_file(path, bits, channels, rate, np.frombuffer(pack(values, bits), dtype=np.uint8), frames, comments)
    raw = path.read_bytes()
    dec = flac_decode.decode(raw)
    want = expected(values, bits)
    assert dec["rate"] == rate and dec["channels"] == channels and dec["bits"] == (24 if bits == 32 else bits)
    assert dec["total"] == frames and dec["md5_ok"]
    got = np.array(dec["samples"], dtype=np.int64).T
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    assert dec["frames"] == (frames + 4095) // 4096
    assert dec["min_frame"] == min(dec["frame_sizes"]) and dec["max_frame"] == max(dec["frame_sizes"])
    # the reference's vendor string (rd/pcm_writer.rs:431-436): crate version + the time of writing
    assert re.fullmatch(r"dsd2dxd v0\.3\.0 Unix datetime \d{9,}", dec["vendor"]) and dec["comments"] == list(comments)
    return raw, dec


@pytest.mark.parametrize("bits", [16, 20, 24, 32])
@pytest.mark.parametrize("kind", ["sine", "noise", "walk", "const", "silence", "rails"])
def test_decode_equality(tmp_path, bits, kind):
    rng = np.random.default_rng(bits * 131 + len(kind))
    for frames, channels in ((4096 + 777, 2), (300, 1)):
        roundtrip(tmp_path, signal(kind, frames, channels, bits, rng), bits, channels, 88200)


@pytest.mark.parametrize("frames", [1, 2, 4, 5, 255, 256, 257, 4095, 4096, 4097, 8192, 3 * 4096 + 1])
def test_block_edges(tmp_path, frames):
    rng = np.random.default_rng(frames)
    roundtrip(tmp_path, signal("walk", frames, 2, 24, rng), 24, 2, 352800)


@pytest.mark.parametrize("channels", [1, 2, 3, 6, 8])
def test_channel_counts(tmp_path, channels):
    rng = np.random.default_rng(channels)
    roundtrip(tmp_path, signal("sine", 5000, channels, 24, rng), 24, channels, 176400)


@pytest.mark.parametrize("rate", [44100, 48000, 88200, 96000, 176400, 192000, 352800, 384000, 705600, 768000, 12345])
def test_sample_rates(tmp_path, rate):
    rng = np.random.default_rng(rate)
    roundtrip(tmp_path, signal("walk", 1000, 2, 16, rng), 16, 2, rate)


def test_many_frames_number_coding(tmp_path):
    """Frame numbers past 127 and 2047 use 2- and 3-byte UTF-8 style codes."""
    rng = np.random.default_rng(5)
    raw, dec = roundtrip(tmp_path, signal("silence", 4096 * 2100, 1, 16, rng), 16, 1, 44100)
    assert dec["frames"] == 2100


def test_comments_and_compression(tmp_path):
    rng = np.random.default_rng(9)
    x = signal("sine", 40000, 2, 24, rng)
    raw, _ = roundtrip(tmp_path, x, 24, 2, 352800, comments=("TITLE=1 kHz", "ARTIST=nobody", "COMMENT=ünïcode"))
    assert len(raw) < 0.8 * x.size * 3                  # a two-tone sine must come out smaller than the raw PCM


def test_float_conversion_is_the_references(tmp_path):
    """audio_file.rs:647-649: (x.clamp(-1, 1) * 8388607.0) as i32 -- out-of-range floats saturate, NaN gives 0."""
    x = np.array([[0.0, 1.0], [-1.0, 2.5], [-7.0, 0.5], [np.nan, 1e-9], [0.99999994, -0.99999994], [1.1920929e-07, -1.1920929e-07]],
                 dtype=np.float32)
    path = tmp_path / "f.flac"
    d2p.write_flac_file(path, 32, 2, 96000, np.frombuffer(x.astype("<f4").tobytes(), dtype=np.uint8), len(x))
    got = np.array(flac_decode.decode(path.read_bytes())["samples"]).T
    want = np.array([[0, 8388607], [-8388607, 8388607], [-8388607, 4194303], [0, 0], [8388606, -8388606], [0, 0]])
    assert np.array_equal(got, want)


def test_write_pcm_file_takes_the_flac_letter(tmp_path):
    rng = np.random.default_rng(2)
    x = signal("walk", 5000, 2, 20, rng)
    path = tmp_path / "p.flac"
    d2p.write_pcm_file(path, "F", 20, 2, 176400, np.frombuffer(pack(x, 20), dtype=np.uint8), len(x))
    dec = flac_decode.decode(path.read_bytes())
    assert dec["bits"] == 20 and np.array_equal(np.array(dec["samples"]).T, x)


def test_reference_error_cases(tmp_path):
    one = np.zeros(6, dtype=np.uint8)
    with pytest.raises(d2p.D2PError, match="No samples"):
        d2p.write_flac_file(tmp_path / "e.flac", 24, 2, 88200, one, 0)
    with pytest.raises(d2p.D2PError, match="No channels"):
        d2p.write_flac_file(tmp_path / "e.flac", 24, 0, 88200, one, 1)
    with pytest.raises(d2p.D2PError, match="only 16, 20, or 24-bit"):
        d2p.write_flac_file(tmp_path / "e.flac", 8, 2, 88200, one, 1)


def test_decoder_rejects_corruption(tmp_path):
    """The checker itself: a flipped payload bit must trip a CRC (otherwise the equality tests prove little)."""
    rng = np.random.default_rng(3)
    raw, _ = roundtrip(tmp_path, signal("walk", 6000, 2, 24, rng), 24, 2, 88200)
    bad = bytearray(raw)
    bad[len(bad) // 2] ^= 0x10
    with pytest.raises(flac_decode.FlacError):
        flac_decode.decode(bytes(bad))


@pytest.mark.gpu
@pytest.mark.parametrize("bits,out_rate", [(24, 352800), (20, 176400), (16, 88200), (32, 96000)])
def test_convert_file_to_flac_equals_wav(tmp_path, bits, out_rate):
    """File in -> FLAC out carries exactly the samples of file in -> WAV out (same kernels, same seed)."""
    from oracle import containers
    rng = np.random.default_rng(bits)
    payload = rng.integers(0, 256, size=2 * 4096 * 5, dtype=np.uint8)
    src = tmp_path / "in.dsf"
    src.write_bytes(containers.make_dsf(payload.tobytes(), 2, 2))
    rw = d2p.convert_file(src, tmp_path / "o.wav", out_rate=out_rate, bits=bits, dither="T", dither_seed=11, container="W")
    rf = d2p.convert_file(src, tmp_path / "o.flac", out_rate=out_rate, bits=bits, dither="T", dither_seed=11, container="F")
    assert rw["frames"] == rf["frames"] and rf["out_bytes"] == os.path.getsize(tmp_path / "o.flac")
    wav = (tmp_path / "o.wav").read_bytes()[44:]
    sb = {16: 2, 20: 3, 24: 3, 32: 4}[bits]
    n = rw["frames"] * 2
    if bits == 32:
        vals = np.frombuffer(wav, dtype="<f4")[:n].reshape(-1, 2)
    elif bits == 16:
        vals = np.frombuffer(wav, dtype="<i2")[:n].reshape(-1, 2)
    else:
        b = np.frombuffer(wav, dtype=np.uint8)[:n * sb].reshape(-1, 3).astype(np.int64)
        v = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
        v = np.where(v >= 1 << 23, v - (1 << 24), v) >> (4 if bits == 20 else 0)
        vals = v.reshape(-1, 2)
    dec = flac_decode.decode((tmp_path / "o.flac").read_bytes())
    assert dec["md5_ok"] and dec["total"] == rw["frames"] and dec["rate"] == out_rate
    assert np.array_equal(np.array(dec["samples"], dtype=np.int64).T, expected(vals, bits))
