"""Every DSD fixture the reference holds (test/*.dsf, test/*.dsd, id3_test/**), WHOLE, through the CUDA path and the
oracle (north_star: "reproduce rdsd2pcm output on all test/*.dsf and *.dsd inputs").

The files travel as tests/golden/ref_fixtures.tar.xz (inputs only).  For each file, every (output rate, filter) the
reference resolves for its DSD rate (A2 / A6) x {16, 20, 24, 32f} with dither off is compared bit for bit (pre-scale f64
and packed PCM); dithers T / R / F are compared exactly against the oracle's counter-keyed stand-in and must stay within
1 LSB (1 f32 ulp-scale step for F) of the undithered result; whole-file conversion goes through d2p_convert_file and the
streaming context as well.

CPU part (not gpu): the tarball holds what SURVEY.md section 4 lists and the container parser reads every file.
"""
import os

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi
from oracle import containers as orc
from oracle import orclib

# raw .dsd fixtures carry no header: format as the reference's scripts pass it (run_all_tests.sh, build_test_*.sh)
RAW_FORMATS = {
    "test/1kHz_mono_p.dsd": dict(channels=1, layout="P", lsb_first=True, dsd_rate=1),
    "test/1kHz_stereo_i.dsd": dict(channels=2, layout="I", lsb_first=False, dsd_rate=1),
    "test/impulse_mono_toggle.dsd": dict(channels=1, layout="P", lsb_first=True, dsd_rate=1),
}
EXPECTED = ["id3_test/1kHz_mono_brokenid3.dsf", "id3_test/dff/1kHz_stereo_i.dff", "id3_test/dff/1kHz_stereo_i_brokenid3.dff",
            "test/1kHz_mono_128.dsf", "test/1kHz_mono_p.dsd", "test/1kHz_mono_p.dsf", "test/1kHz_stereo_128.dsf",
            "test/1kHz_stereo_i.dsd", "test/impulse_mono_toggle.dsd", "test/pinknoise_mono_128.dsf",
            "test/pinknoise_stereo_128.dsf"]

# A2: (out_rate, filter) per DSD rate on the integer path; A6: the three 48k multiples on the L/M path
INT_BY_RATE = {
    1: [(352800, "D"), (352800, "X"), (352800, "E"), (176400, "X"), (176400, "E"), (88200, "X"), (88200, "E"), (44100, "E")],
    2: [(705600, "E"), (352800, "E"), (352800, "C"), (352800, "X"), (176400, "E"), (176400, "C"), (88200, "E"), (88200, "C")],
}
LM_RATES = [96000, 192000, 384000]


def load(ref_files, rel):
    """-> (payload bytes as the reader hands them over, format kwargs for both arms)."""
    path = ref_files[rel]
    if rel in RAW_FORMATS:
        return np.fromfile(path, dtype=np.uint8), dict(RAW_FORMATS[rel])
    info = d2p.container_open(path)
    data = np.fromfile(path, dtype=np.uint8)
    payload = data[info["data_offset"]: info["data_offset"] + info["data_bytes"]]
    layout = chr(info["in_layout"])
    fmt = dict(channels=info["channels"], layout=layout, lsb_first=bool(info["lsb_first"]), dsd_rate=info["dsd_rate"])
    if layout == "P":
        fmt["block_size"] = info["block_size"]
    return payload, fmt


def gpu_stream(raw, **kw):
    cfg = d2p.make_config(keep_f64=True, **kw)
    conv = d2p.Converter(cfg)
    conv.push(raw)
    conv.finish()
    f64 = np.stack([conv.peek_f64(c) for c in range(cfg.channels)])
    pcm = conv.pull()
    conv.close()
    return pcm, f64


def test_tarball_holds_every_reference_fixture(ref_files):
    assert sorted(ref_files) == EXPECTED
    for rel in EXPECTED:
        if rel in RAW_FORMATS:
            continue
        data = open(ref_files[rel], "rb").read()
        want = (orc.parse_dsf if rel.endswith(".dsf") else orc.parse_dff)(data)
        got = d2p.container_open(ref_files[rel])
        for k in ("channels", "dsd_rate", "lsb_first", "block_size", "data_offset", "data_bytes", "sample_count"):
            assert got[k] == want[k], (rel, k)


@pytest.mark.gpu
@pytest.mark.parametrize("rel", EXPECTED)
def test_whole_fixture_every_rate_filter_and_depth_bit_exact(ref_files, rel):
    raw, fmt = load(ref_files, rel)
    combos = INT_BY_RATE[fmt["dsd_rate"]] + [(r, "E") for r in LM_RATES]
    for out_rate, filt in combos:
        for bits in (16, 20, 24, 32):
            kw = dict(out_rate=out_rate, filt=filt, bits=bits, dither="X", **fmt)
            ref = orclib.convert(raw, want_f64=(bits == 24), **kw)
            if bits == 24:
                pcm, f64 = gpu_stream(raw, **kw)
                assert f64.shape == ref.f64.shape, (rel, kw)
                assert np.array_equal(f64, ref.f64), (rel, kw)
            else:
                pcm, frames, _ = d2p.convert_host(d2p.make_config(**kw), raw)
                assert frames == ref.frames[0]
            assert np.array_equal(pcm, ref.pcm), (rel, kw)


@pytest.mark.gpu
@pytest.mark.parametrize("rel", [r for r in EXPECTED if "impulse" not in r])
@pytest.mark.parametrize("dither", ["T", "R", "F"])
def test_whole_fixture_dithers_exact_and_within_one_lsb(ref_files, rel, dither):
    raw, fmt = load(ref_files, rel)
    for out_rate, bits in ((352800, 24), (88200, 16), (192000, 20 if dither != "F" else 32)):
        if dither == "F" and bits != 32:
            bits = 32                                        # the float dither is the 32-bit float writer's (CLI default)
        kw = dict(out_rate=out_rate, bits=bits, dither_seed=0xD17 + bits, **fmt)
        ref = orclib.convert(raw, want_f64=False, dither=dither, stdout_mode=False, **kw)
        clean = orclib.convert(raw, want_f64=False, dither="X", stdout_mode=False, **kw)
        conv = d2p.Converter(d2p.make_config(dither=dither, out_format=_capi.OUT_PLANAR_32, **kw))
        conv.push(raw)
        conv.finish()
        got = conv.pull()                                     # [channels, frames] int32 / float32 (AudioFile samples)
        conv.close()
        assert got.shape[1] == ref.frames[0]
        assert np.array_equal(got, ref.planar), (rel, kw)
        if bits == 32:      # one step of the 2^(e-24)-scaled noise: a few f32 ulps of the sample
            assert np.all(np.abs(got.astype(np.float64) - clean.planar) <= 4 * np.spacing(np.abs(clean.planar).astype(np.float32)) + 1e-30)
        else:
            assert np.abs(got.astype(np.int64) - clean.planar).max() <= 1
        if bits != 32:      # and the packed little-endian stream of the same run (fast 3-byte writer for 20 / 24 bit)
            ref_le = orclib.convert(raw, want_f64=False, dither=dither, **kw)
            le, _, _ = d2p.convert_host(d2p.make_config(dither=dither, **kw), raw)
            assert np.array_equal(le, ref_le.pcm), (rel, kw)


@pytest.mark.gpu
@pytest.mark.parametrize("rel", [r for r in EXPECTED if r not in RAW_FORMATS])
def test_whole_fixture_file_to_file(ref_files, rel, tmp_path):
    """d2p_convert_file on the real containers (incl. the broken-ID3 and wrong-FRM8 ones) == oracle pipeline."""
    raw, fmt = load(ref_files, rel)
    for container, bits, out_rate in (("W", 24, 352800), ("C", 32, 96000), ("A", 20, 176400)):
        dst = tmp_path / "o.bin"
        res = d2p.convert_file(ref_files[rel], dst, out_rate=out_rate, bits=bits, container=container)
        r = orclib.convert(raw, want_f64=False, out_rate=out_rate, bits=bits, stdout_mode=False, **fmt)
        want = orc.WRITERS[container](r.planar, bits, out_rate)
        assert res["frames"] == r.frames[0]
        assert dst.read_bytes() == want, (rel, container)


@pytest.mark.gpu
def test_raw_fixtures_file_to_file(ref_files, tmp_path):
    for rel, fmt in RAW_FORMATS.items():
        raw = np.fromfile(ref_files[rel], dtype=np.uint8)
        dst = tmp_path / "o.wav"
        res = d2p.convert_file(ref_files[rel], dst, out_rate=352800, bits=24, container="W", raw_channels=fmt["channels"],
                               raw_layout=fmt["layout"], raw_lsb_first=fmt["lsb_first"], raw_dsd_rate=fmt["dsd_rate"])
        r = orclib.convert(raw, want_f64=False, out_rate=352800, bits=24, stdout_mode=False, **fmt)
        assert res["frames"] == r.frames[0]
        assert dst.read_bytes() == orc.WRITERS["W"](r.planar, 24, 352800), rel
