"""Container front / back end: DSF / DSDIFF readers and WAV / AIFF / AIFC writers of the C ABI
(include/dsd2pcm_b200_files.h) against the Python restatement in oracle/containers.py.

CPU part: header bytes, whole files, container parsing (synthetic containers, committed heads of the reference's
fixtures, and the fixtures themselves when /root/reference is present).
GPU part (-m gpu): d2p_convert_file end to end == container parse -> oracle conversion -> oracle writer, byte for byte.
"""
import json
import os

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from dsd2dxd_b200 import _capi
from oracle import containers as orc

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
KEYS = ("channels", "dsd_rate", "sample_rate", "lsb_first", "block_size", "data_offset", "data_bytes", "sample_count",
        "id3_offset", "id3_bytes")


def _samples(bits, channels, frames, seed=3):
    rng = np.random.default_rng(seed)
    if bits == 32:
        return rng.standard_normal((channels, frames)).astype(np.float32)
    peak = 1 << (bits - 1)
    x = rng.integers(-peak, peak, size=(channels, frames), dtype=np.int64)
    if frames:
        x[0, 0], x[-1, -1] = -peak, peak - 1
    return x.astype(np.int32)


@pytest.mark.parametrize("container", ["W", "A", "C"])
@pytest.mark.parametrize("bits", [16, 20, 24, 32])
@pytest.mark.parametrize("channels,frames", [(1, 0), (1, 1), (2, 1001), (6, 77)])
def test_pcm_headers_and_files_match_reference_writers(tmp_path, container, bits, channels, frames):
    rate = {16: 44100, 20: 176400, 24: 352800, 32: 192000}[bits]
    smp = _samples(bits, channels, frames)
    if container == "A" and bits == 32:                      # "32 bit float only allowed with wav, aifc, or stdout"
        assert d2p.pcm_header(container, bits, channels, rate, frames) == b""
        with pytest.raises(_capi.D2PError) as e:
            d2p.write_pcm_file(tmp_path / "x.aiff", container, bits, channels, rate, np.zeros(4, np.uint8), 0)
        assert e.value.status == _capi.ERR_BIT_DEPTH
        return
    want = orc.WRITERS[container](smp, bits, rate)
    nbytes = frames * channels * (3 if bits == 20 else bits // 8)
    header, payload = want[: len(want) - nbytes], want[len(want) - nbytes:]
    assert d2p.pcm_header(container, bits, channels, rate, frames) == header
    if container == "C":
        assert len(header) % 4096 == 0                       # FLLR: sample data starts on a page boundary
    path = tmp_path / "out.bin"
    d2p.write_pcm_file(path, container, bits, channels, rate, np.frombuffer(payload, dtype=np.uint8), frames)
    assert path.read_bytes() == want


def test_extended_known_answers():
    # the classic AIFF constants: 44100 Hz = 400E AC44 0000 0000 0000, 48000 = 400E BB80 ..., 352800 = 4011 AC44 ...
    assert orc.extended(44100.0).hex() == "400eac44000000000000"
    for rate, hx in ((44100, "400eac44000000000000"), (48000, "400ebb80000000000000"), (352800, "4011ac44000000000000")):
        assert d2p.pcm_header("A", 16, 2, rate, 10)[28:38].hex() == hx
    for rate in (88200, 96000, 176400, 192000, 384000, 705600, 768000, 1411200, 1, 3, 4294967295):
        assert d2p.pcm_header("A", 24, 1, rate, 0)[28:38] == orc.extended(float(rate))


@pytest.mark.parametrize("channels,dsd_rate,id3", [(1, 1, b""), (2, 2, b"ID3\x04" + bytes(60)), (6, 8, b""), (2, 4, b"ID3")])
def test_parse_synthetic_containers(channels, dsd_rate, id3):
    rng = np.random.default_rng(channels)
    payload = rng.integers(0, 256, size=channels * 4096 * 3, dtype=np.uint8).tobytes()
    dsf = orc.make_dsf(payload, channels, dsd_rate, sample_count=3 * 4096 * 8 - 1000, id3=id3)
    got, want = d2p.container_parse(dsf), orc.parse_dsf(dsf)
    assert got["format"] == ord("S") and got["in_layout"] == ord("P")
    assert {k: got[k] for k in KEYS} == {k: want[k] for k in KEYS}
    assert dsf[got["data_offset"]: got["data_offset"] + got["data_bytes"]] == payload
    for wrong in (False, True):
        dff = orc.make_dff(payload[: channels * 5001], channels, dsd_rate, id3=id3, wrong_frm8=wrong)
        got, want = d2p.container_parse(dff), orc.parse_dff(dff)
        assert got["format"] == ord("F") and got["in_layout"] == ord("I") and got["lsb_first"] == 0
        assert {k: got[k] for k in KEYS} == {k: want[k] for k in KEYS}
        assert dff[got["data_offset"]: got["data_offset"] + got["data_bytes"]] == payload[: channels * 5001]
    # a tag chunk cut short by the end of the file is clipped, not an error (id3_test/*brokenid3*)
    if len(id3) > 10:
        cut = dff[: len(dff) - 20]
        assert d2p.container_parse(cut)["id3_bytes"] == orc.parse_dff(cut)["id3_bytes"] == len(id3) - 20


def test_rejects_what_is_not_dsd():
    for junk in (b"RIFF" + bytes(100), b"DSD " + bytes(20), b"FRM8" + bytes(8) + b"DST " + bytes(50), b""):
        with pytest.raises(_capi.D2PError) as e:
            d2p.container_parse(junk if junk else b"\0")
        assert e.value.status == _capi.ERR_INVALID_ARG


def test_committed_fixture_heads():
    """First 256 bytes + length of every container fixture of the reference (tests/golden/container_heads.json,
    made by tests/golden/make_container_heads.py): the parser must report what the oracle parser reported there."""
    heads = json.load(open(os.path.join(HERE, "golden", "container_heads.json")))
    assert len(heads) >= 8
    for name, h in heads.items():
        got = d2p.container_parse(bytes.fromhex(h["head_hex"]), h["file_bytes"])
        for k in ("channels", "dsd_rate", "sample_rate", "lsb_first", "block_size", "data_offset", "data_bytes", "sample_count"):
            assert got[k] == h["expect"][k], (name, k)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference fixtures are only mounted in the build container")
def test_reference_fixtures_parse_like_the_oracle():
    n = 0
    for sub in ("test", "id3_test", "id3_test/dff"):
        for fn in sorted(os.listdir(os.path.join(REF, sub))):
            p = os.path.join(REF, sub, fn)
            if not fn.endswith((".dsf", ".dff")):
                continue
            data = open(p, "rb").read()
            want = (orc.parse_dsf if fn.endswith(".dsf") else orc.parse_dff)(data)
            got = d2p.container_open(p)
            assert {k: got[k] for k in KEYS} == {k: want[k] for k in KEYS}, fn
            assert d2p.container_parse(data) == got
            n += 1
    assert n >= 8
    # the DSDIFF fixture carries the same sound data as the raw interleaved fixture
    i = d2p.container_open(os.path.join(REF, "id3_test/dff/1kHz_stereo_i.dff"))
    dff = open(os.path.join(REF, "id3_test/dff/1kHz_stereo_i.dff"), "rb").read()
    assert dff[i["data_offset"]: i["data_offset"] + i["data_bytes"]] == open(os.path.join(REF, "test/1kHz_stereo_i.dsd"), "rb").read()


# ---- GPU: whole-file conversion -------------------------------------------------------------------------------
def _oracle_file(payload, info, container, bits, out_rate, **kw):
    from oracle import orclib
    r = orclib.convert(np.frombuffer(payload, dtype=np.uint8), dsd_rate=info["dsd_rate"], out_rate=out_rate, bits=bits,
                       channels=info["channels"], layout=chr(info["in_layout"]), lsb_first=bool(info["lsb_first"]),
                       block_size=info["block_size"] if chr(info["in_layout"]) == "P" else 4096, stdout_mode=False, **kw)
    return orc.WRITERS[container](r.planar, bits, out_rate), r


@pytest.mark.gpu
@pytest.mark.parametrize("container,bits,out_rate", [("W", 24, 352800), ("W", 16, 88200), ("W", 32, 96000), ("A", 20, 176400),
                                                     ("A", 24, 192000), ("C", 32, 352800), ("C", 16, 176400), ("C", 24, 88200)])
def test_convert_file_dsf_matches_oracle_pipeline(tmp_path, fixtures, container, bits, out_rate):
    payload = fixtures["pinknoise_stereo_128"][0].tobytes()
    src = tmp_path / "in.dsf"
    src.write_bytes(orc.make_dsf(payload, 2, 2, id3=b"ID3\x03" + bytes(40)))
    dst = tmp_path / "out.bin"
    res = d2p.convert_file(src, dst, out_rate=out_rate, bits=bits, container=container)
    want, r = _oracle_file(payload, res["in"], container, bits, out_rate)
    assert res["frames"] == r.frames[0] and res["out_bytes"] == len(want)
    assert dst.read_bytes() == want


@pytest.mark.gpu
def test_convert_file_dff_raw_and_sample_count(tmp_path, fixtures):
    # DSDIFF (byte-interleaved, MSB first) with a wrong FRM8 size and a tag chunk behind the sound data
    payload = fixtures["1kHz_stereo_i"][0].tobytes()
    src = tmp_path / "in.dff"
    src.write_bytes(orc.make_dff(payload, 2, 1, id3=b"ID3\x04" + bytes(30), wrong_frm8=True))
    dst = tmp_path / "o.wav"
    res = d2p.convert_file(src, dst, out_rate=176400, bits=24, container="W")
    want, _ = _oracle_file(payload, res["in"], "W", 24, 176400)
    assert dst.read_bytes() == want
    # the same bytes as a raw .dsd file: format comes from the options
    raw = tmp_path / "in.dsd"
    raw.write_bytes(payload)
    res2 = d2p.convert_file(raw, dst, out_rate=176400, bits=24, container="W", raw_channels=2, raw_layout="I",
                            raw_lsb_first=False, raw_dsd_rate=1)
    assert res2["in"]["format"] == ord("R") and dst.read_bytes() == want
    # DSF whose fmt.sample_count ends inside the last block: stop there when asked to
    pay = fixtures["1kHz_mono_p"][0].tobytes()
    count = (len(pay) - 1500) * 8
    src = tmp_path / "m.dsf"
    src.write_bytes(orc.make_dsf(pay, 1, 1, sample_count=count))
    whole = d2p.convert_file(src, tmp_path / "w.wav", out_rate=352800, bits=24)
    cut = d2p.convert_file(src, tmp_path / "c.wav", out_rate=352800, bits=24, dsf_stop_at_sample_count=True)
    assert whole["frames"] - cut["frames"] == 1500
    a, b = (tmp_path / "w.wav").read_bytes(), (tmp_path / "c.wav").read_bytes()
    assert a[44: 44 + cut["frames"] * 3] == b[44:]
    # stereo: the ragged last block group reaches the converter as [ch0 tail | ch1 tail], like a short final read
    pay2 = fixtures["pinknoise_stereo_128"][0]
    groups, tail = pay2.size // 8192 - 1, 2600
    src2 = tmp_path / "s.dsf"
    src2.write_bytes(orc.make_dsf(pay2.tobytes(), 2, 2, sample_count=(groups * 4096 + tail) * 8))
    res3 = d2p.convert_file(src2, tmp_path / "s.wav", out_rate=176400, bits=24, dsf_stop_at_sample_count=True)
    last = pay2[groups * 8192: (groups + 1) * 8192]
    ragged = np.concatenate([pay2[: groups * 8192], last[:tail], last[4096: 4096 + tail]])
    want3, _ = _oracle_file(ragged.tobytes(), res3["in"], "W", 24, 176400)
    assert (tmp_path / "s.wav").read_bytes() == want3
    # errors keep the reference's wording
    with pytest.raises(_capi.D2PError, match="32 bit float only allowed with wav, aifc, or stdout"):
        d2p.convert_file(src, tmp_path / "x.aiff", bits=32, container="A")
    with pytest.raises(_capi.D2PError, match="Taps not found"):
        d2p.convert_file(src, tmp_path / "x.wav", out_rate=352800, filt="C")
