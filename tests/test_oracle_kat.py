"""CPU tests that pin the oracle (test infrastructure) as far as the reference allows.

The reference has no golden vectors and cannot be built here (F1), so the oracle is
pinned by self-evident known answers and by an independent second restatement:
  * literal streaming state machines (oracle/oracle.cpp) == index-based numpy closed forms
    (tests/closed_forms.py), bit for bit, on every topology;
  * impulse fixture -> the tap table; idle pattern -> exactly 0.0; DC gain; 1 kHz tone level;
  * the committed digests of the oracle's own outputs on the fixture excerpts.
"""
import hashlib
import json
import os
import re
import struct

import numpy as np
import pytest

import closed_forms as CF
from oracle import orclib as O

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


@pytest.fixture(scope="module")
def taps():
    return {n: O.table(n) for n in O.table_names()}


def test_tap_tables_match_committed_digests(taps):
    idx = json.load(open(os.path.join(HERE, "golden", "taps_index.json")))
    assert len(idx) == 33 and set(taps) == {e["name"] for e in idx}
    for e in idx:
        t = taps[e["name"]]
        assert t.size == e["len"]
        assert hashlib.sha256(struct.pack("<%dd" % t.size, *t)).hexdigest() == e["sha256"], e["name"]
        assert 0.99 < 2 * t.sum() < 1.03          # unity DC gain of the full symmetric filter


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference mount only exists in the build container")
def test_tap_tables_equal_reference_source(taps):
    """Independent parse of rd/filters*.rs (line-based, unlike tools/extract_taps.py's regex)."""
    for rel in ("src/rdsd2pcm/src/filters.rs", "src/rdsd2pcm/src/filters_lm.rs"):
        name, vals = None, []
        for line in open(os.path.join(REF, rel)):
            m = re.match(r"pub const (HTAPS_\w+): \[f64; (\d+)\]", line)
            if m:
                name, n, vals = m.group(1), int(m.group(2)), []
                continue
            if name and line.strip().startswith("];"):
                if name in taps:
                    assert len(vals) == n
                    assert np.array_equal(np.array(vals), taps[name]), name
                name = None
                continue
            if name:
                tok = line.split("//")[0].strip().rstrip(",")
                if tok:
                    vals.extend(float(x) for x in tok.split(",") if x.strip())


SELECT = {  # (filter, dsd_rate, decim) -> table, transcribed from SURVEY.md A2
    ("D", 1, 8): "HTAPS_D2P", ("X", 1, 8): "HTAPS_XLD", ("E", 1, 8): "HTAPS_DSD64_8TO1_EQ",
    ("E", 2, 8): "HTAPS_DDR_8TO1_EQ", ("E", 4, 8): "HTAPS_DSD256_8TO1_EQ", ("E", 8, 8): None, ("C", 2, 8): None,
    ("X", 1, 16): "HTAPS_16TO1_XLD", ("X", 4, 16): "HTAPS_16TO1_XLD", ("E", 1, 16): "HTAPS_DSD64_16TO1_EQ",
    ("E", 2, 16): "HTAPS_DDR_16TO1_EQ", ("E", 4, 16): "HTAPS_DSD256_16TO1_EQ", ("E", 8, 16): None,
    ("C", 2, 16): "HTAPS_DDR_16TO1_CHEB", ("C", 1, 16): None, ("D", 1, 16): None,
    ("X", 2, 32): "HTAPS_32TO1", ("E", 1, 32): "HTAPS_DSD64_32TO1_EQ", ("E", 4, 32): "HTAPS_DSD256_32TO1_EQ",
    ("E", 2, 32): "HTAPS_DDR_32TO1_EQ", ("E", 8, 32): "HTAPS_DDR_32TO1_EQ", ("C", 1, 32): "HTAPS_DDR_32TO1_CHEB",
    ("E", 4, 64): "HTAPS_DSD256_64TO1_EQ", ("E", 8, 64): "HTAPS_DSD512_64TO1_EQ", ("E", 2, 64): "HTAPS_DDR_64TO1_EQ",
    ("E", 1, 64): "HTAPS_DDR_64TO1_EQ", ("C", 2, 64): "HTAPS_DDR_64TO1_CHEB", ("X", 2, 64): None, ("D", 2, 64): None,
    ("E", 4, 128): "HTAPS_DSD256_128TO1_EQ", ("E", 2, 128): None, ("C", 4, 128): None, ("E", 1, 4): None,
}


def test_select_precalc_taps_table():
    for (f, r, d), want in SELECT.items():
        assert O.select_precalc_taps(f, r, d) == want, (f, r, d)


def test_compute_decim_and_upsample():
    want = {(1, 96000): (147, 5), (1, 192000): (147, 10), (1, 384000): (147, 20), (2, 96000): (294, 5),
            (2, 192000): (147, 5), (2, 384000): (147, 10), (4, 96000): (588, 5), (4, 192000): (294, 5),
            (4, 384000): (147, 5), (8, 384000): (147, 5), (1, 352800): (8, 1), (1, 88200): (32, 1),
            (4, 88200): (128, 1), (8, 352800): (64, 1), (2, 705600): (8, 1), (1, 100000): (64, 1)}
    for k, v in want.items():
        assert O.compute_decim_and_upsample(*k) == v, k


def test_calc_frames_cap():
    assert [O.calc_frames_cap(4096, d, 1) for d in (8, 16, 32, 64, 128)] == [4096, 2048, 1024, 512, 256]
    assert O.calc_frames_cap(4096, 147, 5) == 1131 and O.calc_frames_cap(4096, 147, 10) == 2246
    assert O.calc_frames_cap(4096, 147, 20) == 4475 and O.calc_frames_cap(4096, 294, 5) == 574
    assert O.calc_frames_cap(4096, 588, 5) == 295


@pytest.mark.parametrize("name,decim", [("HTAPS_D2P", 8), ("HTAPS_DSD64_8TO1_EQ", 8), ("HTAPS_16TO1_XLD", 16),
                                        ("HTAPS_DSD64_16TO1_EQ", 16), ("HTAPS_DSD256_32TO1_EQ", 32),
                                        ("HTAPS_DDR_64TO1_CHEB", 64), ("HTAPS_DDR_64TO1_EQ", 64),
                                        ("HTAPS_DSD256_128TO1_EQ", 128)])
def test_bpd_state_machine_equals_closed_form(taps, name, decim):
    B = np.random.default_rng(decim).integers(0, 256, size=4096 * 5, dtype=np.uint8)
    d = O.BytePrecalcDecimator(taps[name], decim)
    assert np.array_equal(d.tables(), CF.bpd_tables(taps[name]))
    o = np.concatenate([d.process_bytes(B[i:i + 4096], 4096 * 8 // decim) for i in range(0, B.size, 4096)])
    assert o.size == CF.bpd_frames(B.size, taps[name].size, decim)
    assert np.array_equal(o, CF.bpd_closed(B, taps[name], decim))


@pytest.mark.parametrize("dsd_rate", [1, 2, 4, 8])
@pytest.mark.parametrize("out_rate", [96000, 192000, 384000])
def test_lm_state_machine_equals_closed_form(taps, dsd_rate, out_rate):
    m, l = O.compute_decim_and_upsample(dsd_rate, out_rate)
    B = np.random.default_rng(m + l).integers(0, 256, size=4096 * 3, dtype=np.uint8)
    cap = O.calc_frames_cap(4096, m, l)
    r = O.LMResampler(l, m, out_rate)
    o = np.concatenate([r.process_bytes_lm(B[i:i + 4096], cap) for i in range(0, B.size, 4096)])
    y, y1, _ = CF.lm_closed(B, taps, l, m, out_rate)
    s1, _, _ = CF.lm_topology(l, m, out_rate)
    assert np.array_equal(O.stage1_run(s1[0], l, s1[1], B), y1)
    assert np.array_equal(o, y)                      # no frame is dropped by the budget logic (A10)
    assert 0 <= B.size * 8 * l // m - o.size <= 10   # start-up loss of 3..10 frames


def test_stage1_ring_aliasing_quirk_is_real(taps):
    """DDRX10_21TO1 with L=5 has 5 LUT groups but a 32-entry byte ring: group 4 re-reads the newest
    window.  A restatement without the modulo differs -- so the quirk must be (and is) reproduced."""
    B = np.random.default_rng(1).integers(0, 256, size=2048, dtype=np.uint8)
    y1 = O.stage1_run("HTAPS_DDRX10_21TO1_EQ", 5, 21, B)
    assert np.array_equal(y1, CF.stage1_closed(B, taps["HTAPS_DDRX10_21TO1_EQ"], 5, 21))
    bits = np.unpackbits(B, bitorder="little")
    lut = CF.stage1_lut(taps["HTAPS_DDRX10_21TO1_EQ"], 5)
    n_hi = CF.stage1_schedule(86, 5, 21, bits.size)
    t, p = n_hi // 5, n_hi % 5
    ideal = np.zeros(n_hi.size)
    for i in range(n_hi.size):
        s = 0.0
        for g in range(lut[p[i]].shape[0]):
            s += lut[p[i]][g][CF.window8(bits, np.array([t[i] - 8 * g]))[0]]
        ideal[i] = s
    assert not np.array_equal(ideal, y1)
    assert np.abs(ideal - y1).max() < 1e-3           # the aliased taps are the tiny tail of the filter


def test_impulse_and_idle_known_answers(fixtures, taps):
    raw, _ = fixtures["impulse_mono_toggle"]          # 0xAA idle, byte 2048 = 0xAB
    half = taps["HTAPS_DSD64_8TO1_EQ"]
    out = O.BytePrecalcDecimator(half, 8).process_bytes(raw, 4096)
    nz = np.nonzero(out)[0]
    assert nz.min() == 1995 and nz.max() == 2007      # 13 consecutive frames, everything else exactly 0.0
    full = np.concatenate([half[::-1], half])
    # bit 0 of byte 2048 flipped -1 -> +1 (delta +2) at stream bit m = 16384; frame k has x[c+r] on
    # full tap (h+r) with c = 8(k+1) - 8nt, so the impulse meets tap index h + m - c
    h, nt = 54, 7
    for n in nz:
        k = n + h - 1
        c = 8 * (k + 1) - 8 * nt
        tap = h + 16384 - c
        assert 0 <= tap < 2 * h
        assert abs(out[n] - 2.0 * full[tap]) < 1e-16


@pytest.mark.skipif(not os.path.isdir(REF), reason="full fixture only in the build container")
def test_1khz_tone_level_full_fixture():
    raw = np.fromfile(os.path.join(REF, "test/1kHz_mono_p.dsf"), dtype=np.uint8)[92:]
    r = O.convert(raw, dsd_rate=1, out_rate=88200, channels=1)
    x = r.f64[0][2000:2000 + 88200]
    spec = np.abs(np.fft.rfft(x * np.hanning(x.size))) / (x.size / 4)
    k = int(np.argmax(spec))
    assert abs(k - 1000) <= 1                          # 1 kHz at 1 Hz/bin
    level_db = 20 * np.log10(spec[k - 2:k + 3].max())
    assert -9.6 < level_db < -8.8                      # -9.18 dB tone, LSB-first reading


def test_golden_digests_of_oracle_outputs(fixtures, golden_digests):
    for key, g in golden_digests.items():
        name, out_rate, bits = key.split("/")
        raw, meta = fixtures[name]
        r = O.convert(raw, dsd_rate=meta["dsd_rate"], out_rate=int(out_rate), bits=int(bits), channels=meta["channels"],
                      layout=meta["layout"], lsb_first=meta["lsb_first"])
        assert r.frames[0] == g["frames"]
        assert hashlib.sha256(r.pcm.tobytes()).hexdigest() == g["pcm_sha256"], key
        assert hashlib.sha256(r.f64.tobytes()).hexdigest() == g["f64_sha256"], key


def test_my_round_quirk_and_scale():
    # F9: negatives land one LSB below round-half-up
    for x, want in ((-0.3, -1), (-1.0, -2), (-1.6, -3), (0.0, 0), (0.49, 0), (0.5, 1), (2.5, 3), (-0.0, 0)):
        assert O.lib().orc_my_round(x) == want
        assert CF.my_round(np.array([x]))[0] == want
    for bits, lvl, l in ((16, 0.0, 1), (20, -3.0, 5), (24, 4.0, 20), (32, 0.0, 10), (32, -6.0, 1)):
        assert O.lib().orc_scale_factor(bits, lvl, l) == CF.scale_factor(bits, lvl, l)


@pytest.mark.parametrize("bits", [16, 20, 24, 32])
@pytest.mark.parametrize("out_rate", [352800, 96000])
def test_writer_quantise_pack_matches_numpy(taps, bits, out_rate):
    raw = np.random.default_rng(bits).integers(0, 256, size=2 * 4096 * 3, dtype=np.uint8)
    r = O.convert(raw, dsd_rate=1, out_rate=out_rate, bits=bits, level_db=7.0, channels=2)
    s = r.f64 * r.scale
    q = s.astype(np.float32) if bits == 32 else CF.quantize(s, bits)
    assert np.array_equal(r.pcm, CF.pack_le(q.T, bits))
    f = O.convert(raw, dsd_rate=1, out_rate=out_rate, bits=bits, level_db=7.0, channels=2, stdout_mode=False)
    assert np.array_equal(f.planar, q)
    if bits != 32:
        assert r.clips == f.clips
        if out_rate == 352800:
            assert r.clips > 0                          # +7 dB on full-scale noise clips in runs


def test_reference_errors():
    raw = np.zeros(8192, dtype=np.uint8)
    for filt in ("X", "D"):                             # F11: DSD128 -> 88.2k resolves only E and C
        with pytest.raises(ValueError, match="Taps not found for decimation ratio 64"):
            O.convert(raw, dsd_rate=2, out_rate=88200, filt=filt)
    with pytest.raises(ValueError, match="Unsupported bit depth"):
        O.convert(raw, bits=8)
    O.convert(raw, dsd_rate=2, out_rate=88200, filt="C")


def test_layouts_and_bit_order_are_equivalent_views():
    rng = np.random.default_rng(4)
    chans = rng.integers(0, 256, size=(2, 4096 * 3), dtype=np.uint8)
    planar = chans.reshape(2, 3, 4096).transpose(1, 0, 2).reshape(-1)
    inter = chans.T.reshape(-1)
    a = O.convert(planar, channels=2, layout="P", lsb_first=True)
    b = O.convert(inter, channels=2, layout="I", lsb_first=True)
    c = O.convert(CF.REV8[inter], channels=2, layout="I", lsb_first=False)
    assert np.array_equal(a.pcm, b.pcm) and np.array_equal(a.pcm, c.pcm)
    assert np.array_equal(a.f64[0], CF.bpd_closed(chans[0], O.table("HTAPS_DSD64_8TO1_EQ"), 8))


def test_dither_statistics_and_bounds():
    raw = np.random.default_rng(9).integers(0, 256, size=2 * 4096 * 8, dtype=np.uint8)
    clean = O.convert(raw, dsd_rate=1, out_rate=352800, bits=16, level_db=-20, stdout_mode=False)
    for d in ("T", "R"):
        x = O.convert(raw, dsd_rate=1, out_rate=352800, bits=16, level_db=-20, dither=d, dither_seed=3, stdout_mode=False)
        diff = x.planar.astype(np.int64) - clean.planar
        assert np.abs(diff).max() <= 1 and np.abs(diff).sum() > 0
        assert abs(diff.mean()) < 0.02
    f = O.convert(raw, dsd_rate=1, out_rate=352800, bits=32, dither="F", dither_seed=3, stdout_mode=False)
    c32 = O.convert(raw, dsd_rate=1, out_rate=352800, bits=32, stdout_mode=False)
    ulp = np.spacing(np.abs(c32.planar))
    assert (np.abs(f.planar - c32.planar) <= ulp).all() and (f.planar != c32.planar).any()
