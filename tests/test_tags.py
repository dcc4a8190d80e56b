"""Metadata carry-over (the last row of the scope table): ID3v2 -> Vorbis comments / PICTURE blocks
for FLAC, the tag as an "ID3 " chunk in WAV / AIFF / AIFC, " [rate]" on the album title with -a, the reference's output
naming through the native CLI.  The C++ (dsd2dxd_b200/csrc/d2p_tags.cpp) against oracle/tags.py, a Python restatement of
PcmWriter::id3_to_flac_meta (rd/pcm_writer.rs:30-96, 425-605) -- on the tags the reference's own fixtures carry
(id3_test/*, two of them deliberately truncated) and on synthetic v2.3 / v2.4 tags in all four text encodings.
"""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

import dsd2dxd_b200 as d2p
from oracle import containers as orc
from oracle import flac_decode
from oracle import orclib
from oracle import tags as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def png(w, h):
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d))
    raw = b"".join(b"\0" + bytes(3 * w) for _ in range(h))
    return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")


def rich_tag(major, enc):
    frames = [
        T.text_frame("TIT2", "Größe – 標題", enc if enc != 0 else 3), T.text_frame("TPE1", "Artist A\0Artist B" if major == 4 else "Artist A", enc),
        T.text_frame("TALB", "The Album", enc), T.text_frame("TRCK", "03/12", enc), T.text_frame("TPOS", "1/2", enc),
        T.text_frame("TDRC" if major == 4 else "TYER", "2021-05-04" if major == 4 else "1999", enc), T.text_frame("TDOR" if major == 4 else "TORY", "1984", enc),
        T.text_frame("MVIN", "2/4", enc), T.text_frame("TCON", "Test Tones", enc), T.text_frame("TBPM", "120", enc),
        T.comm("COMM", "a comment", "desc", enc), T.comm("USLT", "la la la", "", enc),
        T.people_frame("TIPL" if major == 4 else "IPLS", [("producer", "P. Roducer"), ("Engineer", "E. Ngineer"), ("cook", "nobody")], enc),
        T.people_frame("TMCL", [("violin", "V. Player")], enc),
        ("UFID", b"http://musicbrainz.org\0" + b"0123-4567"), ("UFID", b"http://other.example\0xyz"),
        ("WCOP", b"https://license.example/by"), ("WOAR", b"https://artist.example"),
        T.txxx("MusicBrainz Album Id", "mb-album", enc), T.txxx("musicbrainz album type", "album", enc),
        T.txxx("REPLAYGAIN_TRACK_GAIN", "-3.2 dB", enc), T.txxx("", "ignored", enc),
        T.apic(3, "image/png", "front", png(7, 5), enc), T.apic(4, "image/png", "", png(3, 2), enc), T.apic(0, "image/png", "other", png(1, 1), enc),
    ]
    return T.make_tag(frames, major=major, padding=64)


def test_malformed_text_still_gives_utf8_comments():
    """FLAC wants UTF-8 Vorbis comments: ill-formed UTF-8 and surrogates without a partner become U+FFFD (one per maximal
    ill-formed run, as String::from_utf8_lossy and Python's errors="replace" put them)."""
    u16 = lambda units, big: b"".join(struct.pack(">H" if big else "<H", u) for u in units)
    frames = [
        ("TIT2", b"\x03" + b"ok \xff\xfe stray, \xe2\x82 cut short, \xed\xa0\x80 surrogate, \xc0\xaf overlong, \xf4\x90\x80\x80 too big, fine: \xe6\xa8\x99"),
        ("TALB", b"\x01\xff\xfe" + u16([0x41, 0xD800, 0x42, 0xDC00, 0x43, 0xD83D, 0xDE00, 0xD800], False)),
        ("TPE1", b"\x02" + u16([0xDC00, 0x61, 0xD800, 0xD800, 0xDC00], True)),
        ("TCON", b"\x03tail \xe6\xa8"),
        ("UFID", b"http://musicbrainz.org\0" + b"id-\xff-\xe2\x82"),
    ]
    blob = T.make_tag(frames, major=4, padding=16)
    want = T.id3_to_vorbis(blob)
    got = d2p.id3_to_vorbis(blob)                               # the wrapper decodes strictly: invalid bytes would raise
    assert got == want
    assert "ALBUM=A\ufffdB\ufffdC\U0001F600\ufffd" in got and "ARTIST=\ufffda\ufffd\U00010000" in got
    assert any(c.startswith("TITLE=ok \ufffd\ufffd stray, \ufffd cut short, \ufffd\ufffd\ufffd surrogate") for c in got)
    assert "GENRE=tail \ufffd" in got and "MUSICBRAINZ_TRACKID=id-\ufffd-\ufffd" in got


@pytest.mark.parametrize("major", [3, 4])
@pytest.mark.parametrize("enc", [0, 1, 3] + [2])
def test_id3_to_vorbis_mapping_matches_the_reference_table(major, enc):
    if major == 3 and enc in (2, 3):
        pytest.skip("ID3v2.3 knows Latin-1 and UTF-16 with BOM only")
    blob = rich_tag(major, enc)
    want = T.id3_to_vorbis(blob)
    got = d2p.id3_to_vorbis(blob)
    assert got == want
    keys = [c.split("=")[0] for c in got]
    for k in ("ARTIST", "ALBUM", "TITLE", "TRACKNUMBER", "TOTALTRACKS", "TRACKTOTAL", "DISCNUMBER", "TOTALDISCS", "DISCTOTAL", "DATE",
              "ORIGINALDATE", "MOVEMENTTOTAL", "MOVEMENT", "COMMENT", "LYRICS", "PRODUCER", "ENGINEER", "PERFORMER",
              "MUSICBRAINZ_TRACKID", "LICENSE", "WEBSITE", "MUSICBRAINZ_ALBUMID", "RELEASETYPE", "REPLAYGAIN_TRACK_GAIN"):
        assert k in keys, k
    assert "TRACKNUMBER=3" in got and "PERFORMER=V. Player (violin)" in got and "MUSICBRAINZ_TRACKID=0123-4567" in got
    assert keys.index("ARTIST") < keys.index("ALBUM") < keys.index("TITLE") < keys.index("TRACKNUMBER")     # the table's order


def test_fixture_tags_and_truncated_tags(ref_files):
    for rel, ok in (("id3_test/dff/1kHz_stereo_i.dff", True), ("id3_test/dff/1kHz_stereo_i_brokenid3.dff", False),
                    ("id3_test/1kHz_mono_brokenid3.dsf", False)):
        info = d2p.container_open(ref_files[rel])
        blob = open(ref_files[rel], "rb").read()[info["id3_offset"]: info["id3_offset"] + info["id3_bytes"]]
        assert blob[:3] == b"ID3"
        got, want = d2p.id3_to_vorbis(blob), T.id3_to_vorbis(blob)
        assert got == want
        if ok:
            assert got == ["ARTIST=clone206", "ALBUM=clone206's Greatest Test Tone Hits", "TITLE=1kHz Test Tone DSD64"]
        else:
            assert got is None                                   # a tag cut off by the end of the file is no tag
    assert d2p.id3_to_vorbis(b"") == [] and d2p.id3_to_vorbis(b"ID3\x02\x00\x00\x00\x00\x00\x00") is None


def _dsf_with_tag(tmp_path, fixtures, blob, name="tagged.dsf"):
    pay = fixtures["pinknoise_stereo_128"][0].tobytes()
    src = tmp_path / name
    src.write_bytes(orc.make_dsf(pay, 2, 2, id3=blob))
    return src, pay


@pytest.mark.gpu
def test_flac_gets_comments_and_cover_pictures(tmp_path, fixtures):
    blob = rich_tag(4, 3)
    src, pay = _dsf_with_tag(tmp_path, fixtures, blob)
    res = d2p.convert_file(src, tmp_path / "o.flac", out_rate=88200, bits=24, container="F", copy_tags=True, append_rate=True)
    assert res["tag_bytes"] == len(blob) + len(" [88.2K]")
    dec = flac_decode.decode((tmp_path / "o.flac").read_bytes())
    want = [c if not c.startswith("ALBUM=") else c + " [88.2K]" for c in T.id3_to_vorbis(blob)]     # append_album_suffix
    assert dec["comments"] == want and dec["md5_ok"]
    pics = T.pictures(blob)
    assert [(p["type"], p["mime"], p["description"], p["data"]) for p in dec["pictures"]] == pics and len(pics) == 2
    assert (dec["pictures"][0]["width"], dec["pictures"][0]["height"], dec["pictures"][0]["depth"]) == (7, 5, 24)
    ref = orclib.convert(np.frombuffer(pay, dtype=np.uint8), dsd_rate=2, out_rate=88200, bits=24, channels=2, stdout_mode=False, want_f64=False)
    assert np.array_equal(np.array(dec["samples"], dtype=np.int64), ref.planar)
    # switch off (the default): the bare file of before
    d2p.convert_file(src, tmp_path / "bare.flac", out_rate=88200, bits=24, container="F")
    bare = flac_decode.decode((tmp_path / "bare.flac").read_bytes())
    assert bare["comments"] == [] and bare["pictures"] == []


@pytest.mark.gpu
@pytest.mark.parametrize("container", ["W", "A", "C"])
def test_pcm_files_carry_the_tag_as_an_id3_chunk(tmp_path, fixtures, container):
    blob = rich_tag(3, 1)
    src, pay = _dsf_with_tag(tmp_path, fixtures, blob)
    dst = tmp_path / "o.bin"
    res = d2p.convert_file(src, dst, out_rate=176400, bits=24, container=container, copy_tags=True)
    ref = orclib.convert(np.frombuffer(pay, dtype=np.uint8), dsd_rate=2, out_rate=176400, bits=24, channels=2, stdout_mode=False, want_f64=False)
    bare = orc.WRITERS[container](ref.planar, 24, 176400)
    data = dst.read_bytes()
    assert res["out_bytes"] == len(data) == len(bare) + 8 + len(blob) + len(blob) % 2
    assert data[8: len(bare)] == bare[8:]                                   # everything but the container size is the bare file
    size_field = int.from_bytes(data[4:8], "little" if container == "W" else "big")
    assert size_field == int.from_bytes(bare[4:8], "little" if container == "W" else "big") + 8 + len(blob) + len(blob) % 2
    chunk = data[len(bare):]
    assert chunk[:4] == b"ID3 " and int.from_bytes(chunk[4:8], "little" if container == "W" else "big") == len(blob)
    assert chunk[8: 8 + len(blob)] == blob                                  # a reader finds exactly the source tag
    assert T.id3_to_vorbis(chunk[8: 8 + len(blob)]) == T.id3_to_vorbis(blob)
    # -a: only the album frame changes, in its own encoding
    d2p.convert_file(src, dst, out_rate=176400, bits=24, container=container, copy_tags=True, append_rate=True)
    chunk = dst.read_bytes()[len(bare):]
    n = int.from_bytes(chunk[4:8], "little" if container == "W" else "big")
    got = T.id3_to_vorbis(chunk[8: 8 + n])
    assert got == [c + " [176.4K]" if c.startswith("ALBUM=") else c for c in T.id3_to_vorbis(blob)]


def _walk_chunks(data, little, ssnd_short=False):
    """Chunks of a RIFF / FORM file the way a conformant reader finds them: every chunk starts on an even offset.
    ssnd_short: the reference's AIFF writer leaves the 8 offset / blockSize bytes out of SSND's size field."""
    assert int.from_bytes(data[4:8], "little" if little else "big") == len(data) - 8
    at, out = 12, {}
    while at + 8 <= len(data):
        n = int.from_bytes(data[at + 4: at + 8], "little" if little else "big")
        out[bytes(data[at: at + 4])] = (at + 8, n)
        if ssnd_short and data[at: at + 4] == b"SSND":
            n += 8
        at += 8 + n + (n & 1)
    assert at in (len(data), len(data) + 1)      # the reference leaves out the pad behind a LAST odd chunk (bare files)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("container", ["W", "A", "C"])
def test_tag_chunk_behind_an_odd_sized_payload_starts_on_an_even_offset(tmp_path, fixtures, container):
    """Mono 24-bit with an odd frame count: the sample chunk has an odd size and no pad of its own."""
    blob = rich_tag(3, 1)
    pay = fixtures["1kHz_mono_p"][0].tobytes()
    src = tmp_path / "mono.dsf"
    src.write_bytes(orc.make_dsf(pay, 1, 1, id3=blob))
    dst = tmp_path / "o.bin"
    res = d2p.convert_file(src, dst, out_rate=352800, bits=24, container=container, copy_tags=True)
    assert res["frames"] % 2 == 1
    ref = orclib.convert(np.frombuffer(pay, dtype=np.uint8), dsd_rate=1, out_rate=352800, bits=24, channels=1, stdout_mode=False, want_f64=False)
    bare = orc.WRITERS[container](ref.planar, 24, 352800)
    data = dst.read_bytes()
    assert len(bare) % 2 == 1 and data[8: len(bare)] == bare[8:]
    assert res["out_bytes"] == len(data) == len(bare) + 1 + 8 + len(blob) + len(blob) % 2
    chunks = _walk_chunks(data, container == "W", container == "A")
    at, n = chunks[b"ID3 "]
    assert at % 2 == 0 and data[at: at + n] == blob
    sample_chunk = b"data" if container == "W" else b"SSND"
    assert sample_chunk in chunks and _walk_chunks(bare, container == "W", container == "A").keys() == chunks.keys() - {b"ID3 "}


@pytest.mark.gpu
def test_album_suffix_leaves_frames_with_body_changing_flags_alone(tmp_path, fixtures):
    """-a rewrites TALB in place; a compressed / encrypted / grouped frame (bits 0x80 / 0x40 / 0x20 of the second flag
    byte in v2.3; 0x08 / 0x04 / 0x40, and 0x02 / 0x01, in v2.4) does not hold plain text: the tag is copied unchanged."""
    def chunk_of(blob, name):
        src, _ = _dsf_with_tag(tmp_path, fixtures, blob, name)
        dst = tmp_path / "o.wav"
        d2p.convert_file(src, dst, out_rate=176400, bits=16, container="W", copy_tags=True, append_rate=True)
        data = dst.read_bytes()
        at, n = _walk_chunks(data, True)[b"ID3 "]
        return data[at: at + n]
    for major, bits in ((3, (0x80, 0x40, 0x20)), (4, (0x08, 0x04, 0x40, 0x02, 0x01))):
        blob = T.make_tag([T.text_frame("TALB", "The Album", 0), T.text_frame("TIT2", "t", 0)], major=major, padding=8)
        assert blob[10:14] == b"TALB"
        plain = chunk_of(blob, f"p{major}.dsf")
        assert plain != blob and b"The Album [176.4K]" in plain
        for bit in bits:
            flagged = blob[:19] + bytes([bit]) + blob[20:]
            assert chunk_of(flagged, f"f{major}_{bit}.dsf") == flagged, (major, hex(bit))


@pytest.mark.gpu
def test_reference_fixture_with_a_broken_tag_converts_without_one(ref_files, tmp_path):
    for rel in ("id3_test/dff/1kHz_stereo_i_brokenid3.dff", "id3_test/1kHz_mono_brokenid3.dsf"):
        a = d2p.convert_file(ref_files[rel], tmp_path / "a.wav", copy_tags=True)
        b = d2p.convert_file(ref_files[rel], tmp_path / "b.wav", copy_tags=False)
        assert a["tag_bytes"] == 0 and (tmp_path / "a.wav").read_bytes() == (tmp_path / "b.wav").read_bytes()
    ok = d2p.convert_file(ref_files["id3_test/dff/1kHz_stereo_i.dff"], tmp_path / "c.flac", container="F", copy_tags=True)
    assert ok["tag_bytes"] == 1129
    assert flac_decode.decode((tmp_path / "c.flac").read_bytes())["comments"] == [
        "ARTIST=clone206", "ALBUM=clone206's Greatest Test Tone Hits", "TITLE=1kHz Test Tone DSD64"]


@pytest.mark.gpu
def test_cli_names_outputs_and_copies_tags_like_the_reference(ref_files, tmp_path):
    """get_out_filename_path / derive_output_dir / copy_artwork (rd/conversion_context.rs:400-533): <stem>[_<rate>].<ext>,
    -p output directory, artwork files of the input directory copied next to the output."""
    cli = os.path.join(ROOT, "tools", "d2p_cli")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tools"), "-s", "d2p_cli"])
    album = tmp_path / "in" / "album"
    album.mkdir(parents=True)
    src = album / "tone.dff"
    src.write_bytes(open(ref_files["id3_test/dff/1kHz_stereo_i.dff"], "rb").read())
    (album / "cover.jpg").write_bytes(b"\xff\xd8not really a jpeg")
    (album / "notes.txt").write_text("not artwork")
    out = tmp_path / "out"
    out.mkdir()
    r = subprocess.run([cli, "-o", "F", "-r", "96000", "-b", "24", "-d", "X", "-a", "-p", str(out), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    flac = out / "tone_96K.flac"
    assert flac.exists() and (out / "cover.jpg").read_bytes() == (album / "cover.jpg").read_bytes() and not (out / "notes.txt").exists()
    dec = flac_decode.decode(flac.read_bytes())
    assert dec["comments"] == ["ARTIST=clone206", "ALBUM=clone206's Greatest Test Tone Hits [96K]", "TITLE=1kHz Test Tone DSD64"]
