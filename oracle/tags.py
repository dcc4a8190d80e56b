"""Test infrastructure: Python restatement of the reference's metadata carry-over, written from
the cited reference lines, NOT from the C++ under test.  Only tests/ may import it.

  id3_to_vorbis(blob)   PcmWriter::id3_to_flac_meta                      rd/pcm_writer.rs:30-96, 425-580
  pictures(blob)        ... its picture loop (front / back cover only)   rd/pcm_writer.rs:583-603
  album_with_suffix     ConversionContext::append_album_suffix           rd/conversion_context.rs:592-621
  make_tag / frame builders: synthetic ID3v2.3 / 2.4 tags for the tests
The reference reads tags with the `id3` crate (1.16.3, not in the mount): frame layouts follow the published ID3v2.3 /
2.4 structure documents.  **Parity unpinned** for this piece: no reference-run output exists.
"""
import struct

ID3_VORBIS_MAP = [
    ("TPE1", "ARTIST"), ("TPE2", "ALBUMARTIST"), ("TSO2", "ALBUMARTISTSORT"), ("TSOA", "ALBUMSORT"), ("TSOP", "ARTISTSORT"),
    ("TALB", "ALBUM"), ("TIT2", "TITLE"), ("TBPM", "BPM"), ("TCOM", "COMPOSER"), ("TSOC", "COMPOSERSORT"), ("TPE3", "CONDUCTOR"),
    ("TCOP", "COPYRIGHT"), ("TCON", "GENRE"), ("TIT1", "GROUPING"), ("TKEY", "KEY"), ("TSRC", "ISRC"), ("TPUB", "LABEL"),
    ("TLAN", "LANGUAGE"), ("TEXT", "LYRICIST"), ("TMED", "MEDIA"), ("TMOO", "MOOD"), ("MVNM", "MOVEMENTNAME"),
    ("TOFN", "ORIGINALFILENAME"), ("TDRL", "RELEASEDATE"), ("TPE4", "REMIXER"), ("TIT3", "SUBTITLE"), ("TSOT", "TITLESORT"),
    ("TENC", "ENCODEDBY"), ("TSSE", "ENCODERSETTINGS"), ("TCMP", "COMPILATION"), ("TSST", "DISCSUBTITLE"),
]
ID3_TXXX_VORBIS_MAP = [
    ("MusicBrainz Artist Id", "MUSICBRAINZ_ARTISTID"), ("MusicBrainz Album Id", "MUSICBRAINZ_ALBUMID"),
    ("MusicBrainz Album Artist Id", "MUSICBRAINZ_ALBUMARTISTID"), ("MusicBrainz Release Group Id", "MUSICBRAINZ_RELEASEGROUPID"),
    ("MusicBrainz Release Track Id", "MUSICBRAINZ_RELEASETRACKID"), ("MusicBrainz TRM Id", "MUSICBRAINZ_TRMID"),
    ("MusicBrainz Work Id", "MUSICBRAINZ_WORKID"), ("MusicBrainz Disc Id", "MUSICBRAINZ_DISCID"),
    ("MusicBrainz Original Artist Id", "MUSICBRAINZ_ORIGINALARTISTID"), ("MusicBrainz Original Album Id", "MUSICBRAINZ_ORIGINALALBUMID"),
    ("MusicBrainz Album Release Country", "RELEASECOUNTRY"), ("MusicBrainz Album Status", "RELEASESTATUS"),
    ("MusicBrainz Album Type", "RELEASETYPE"), ("Acoustid Id", "ACOUSTID_ID"), ("Acoustid Fingerprint", "ACOUSTID_FINGERPRINT"),
    ("MusicIP PUID", "MUSICIP_PUID"), ("Writer", "WRITER"),
]
ID3_TIPL_VORBIS_MAP = [("arranger", "ARRANGER"), ("engineer", "ENGINEER"), ("dj-mix", "DJMIXER"), ("mix", "MIXER"), ("producer", "PRODUCER")]
CODECS = {0: "latin-1", 1: "utf-16", 2: "utf-16-be", 3: "utf-8"}


def _syncsafe(b):
    return (b[0] & 0x7F) << 21 | (b[1] & 0x7F) << 14 | (b[2] & 0x7F) << 7 | (b[3] & 0x7F)


def parse(blob):
    """-> (major version, [(frame id, body bytes)]) or None when the tag is unreadable (truncated, unknown version)."""
    blob = bytes(blob)
    if len(blob) < 10 or blob[:3] != b"ID3" or blob[3] not in (3, 4) or any(x & 0x80 for x in blob[6:10]):
        return None
    major, flags, size = blob[3], blob[5], _syncsafe(blob[6:10])
    if 10 + size > len(blob):
        return None
    body = blob[10:10 + size]
    if flags & 0x80 and major == 3:
        body = body.replace(b"\xff\x00", b"\xff")
    q = 0
    if flags & 0x40:
        q = _syncsafe(body[:4]) if major == 4 else 4 + struct.unpack(">I", body[:4])[0]
    frames = []
    while q + 10 <= len(body) and body[q] != 0:
        fid = body[q:q + 4]
        if not all(65 <= c <= 90 or 48 <= c <= 57 for c in fid):
            break
        raw = _syncsafe(body[q + 4:q + 8]) if major == 4 else struct.unpack(">I", body[q + 4:q + 8])[0]
        ff = body[q + 9]
        if q + 10 + raw > len(body):
            return None
        data = body[q + 10:q + 10 + raw]
        q += 10 + raw
        if major == 4:
            if ff & 0x0C:
                continue
            if ff & 0x01:
                data = data[4:]
            if ff & 0x02 or flags & 0x80:
                data = data.replace(b"\xff\x00", b"\xff")
        elif ff & 0xC0:
            continue
        frames.append((fid.decode("ascii"), data))
    return major, frames


def _dec(enc, data):
    if enc == 1 and data[:2] not in (b"\xff\xfe", b"\xfe\xff"):
        return data[: len(data) // 2 * 2].decode("utf-16-le", errors="replace")
    if enc in (1, 2):
        data = data[: len(data) // 2 * 2]
    return data.decode(CODECS[enc], errors="replace")


def _split_term(enc, data, start=0):
    """(decoded string up to the terminator, index just behind the terminator)"""
    if enc in (1, 2):
        i = start
        while i + 1 < len(data) and not (data[i] == 0 and data[i + 1] == 0):
            i += 2
        return _dec(enc, data[start:i]), min(len(data), i + 2)
    i = data.find(b"\0", start)
    if i < 0:
        return _dec(enc, data[start:]), len(data)
    return _dec(enc, data[start:i]), i + 1


def _text(frames, fid):
    for f, d in frames:
        if f == fid and d:
            return _dec(d[0], d[1:]).rstrip("\0")
    return None


def _pair(s):
    a, _, b = s.partition("/")
    n = int(a) if a.isdigit() else None
    t = int(b) if b.isdigit() else None
    return n, t


def _people(d):
    enc, q, items = d[0], 1, []
    while q < len(d):
        s, q = _split_term(enc, d, q)
        items.append(s)
    return list(zip(items[0::2], items[1::2]))


def id3_to_vorbis(blob):
    p = parse(blob)
    if p is None:
        return None
    _, frames = p
    out = []
    ins = lambda k, v: out.append(f"{k}={v}")
    for fid, name in ID3_VORBIS_MAP:
        v = _text(frames, fid)
        if v is not None:
            for val in v.split("\0"):
                if val:
                    ins(name, val)
    for fid, a, b1, b2 in (("TRCK", "TRACKNUMBER", "TOTALTRACKS", "TRACKTOTAL"), ("TPOS", "DISCNUMBER", "TOTALDISCS", "DISCTOTAL")):
        v = _text(frames, fid)
        if v is not None:
            n, t = _pair(v)
            if n is not None:
                ins(a, n)
                if t is not None:
                    ins(b1, t)
                    ins(b2, t)
    d = _text(frames, "TDRC")
    if d:
        ins("DATE", d)
    else:
        y = _text(frames, "TYER")
        if y and y.lstrip("-").isdigit():
            ins("DATE", int(y))
    v = _text(frames, "TDOR")
    if v is None:
        v = _text(frames, "TORY")
    if v is not None:
        ins("ORIGINALDATE", v)
    v = _text(frames, "MVIN")
    if v is not None:
        ins("MOVEMENTTOTAL", v)
        ins("MOVEMENT", v)
    for fid, name in (("COMM", "COMMENT"), ("USLT", "LYRICS")):
        for f, dd in frames:
            if f == fid and len(dd) >= 4:
                _, q = _split_term(dd[0], dd, 4)
                ins(name, _dec(dd[0], dd[q:]).rstrip("\0"))
                break
    for f, dd in frames:
        if f in ("TIPL", "IPLS") and dd:
            for role, who in _people(dd):
                for r, name in ID3_TIPL_VORBIS_MAP:
                    if role.lower() == r:
                        ins(name, who)
    for f, dd in frames:
        if f == "TMCL" and dd:
            for role, who in _people(dd):
                ins("PERFORMER", f"{who} ({role})")
    for f, dd in frames:
        if f == "UFID" and b"\0" in dd:
            owner, _, ident = dd.partition(b"\0")
            if owner == b"http://musicbrainz.org":
                ins("MUSICBRAINZ_TRACKID", ident.decode("utf-8", errors="replace"))
    for fid, name in (("WCOP", "LICENSE"), ("WOAR", "WEBSITE")):
        for f, dd in frames:
            if f == fid:
                ins(name, dd.decode("latin-1").rstrip("\0"))
                break
    for f, dd in frames:
        if f == "TXXX" and dd:
            desc, q = _split_term(dd[0], dd, 1)
            if not desc:
                continue
            name = next((n for d_, n in ID3_TXXX_VORBIS_MAP if d_.lower() == desc.lower()), desc)
            for val in _dec(dd[0], dd[q:]).rstrip("\0").split("\0"):
                if val:
                    ins(name, val)
    return out


def pictures(blob):
    """-> [(type, mime, description, data)] for front (3) and back (4) covers, in tag order."""
    p = parse(blob)
    out = []
    if p is None:
        return out
    for f, d in p[1]:
        if f != "APIC" or len(d) < 4:
            continue
        enc = d[0]
        z = d.find(b"\0", 1)
        mime, kind = d[1:z].decode("latin-1"), d[z + 1]
        if kind not in (3, 4):
            continue
        desc, q = _split_term(enc, d, z + 2)
        out.append((kind, mime, desc, d[q:]))
    return out


RATE_DOT = {88200: "88.2K", 96000: "96K", 176400: "176.4K", 192000: "192K", 352800: "352.8K", 384000: "384K", 705600: "705.6K",
            1411200: "1411.2K"}


# ---- builders for synthetic tags ------------------------------------------------------------------------------------
def text_frame(fid, text, enc=3):
    data = text.encode(CODECS[enc])
    return fid, bytes([enc]) + data


def txxx(desc, value, enc=3):
    term = b"\0\0" if enc in (1, 2) else b"\0"
    return "TXXX", bytes([enc]) + desc.encode(CODECS[enc]) + term + value.encode(CODECS[enc])


def comm(fid, text, desc="", enc=3, lang=b"eng"):
    term = b"\0\0" if enc in (1, 2) else b"\0"
    return fid, bytes([enc]) + lang + desc.encode(CODECS[enc]) + term + text.encode(CODECS[enc])


def apic(kind, mime, desc, data, enc=3):
    term = b"\0\0" if enc in (1, 2) else b"\0"
    return "APIC", bytes([enc]) + mime.encode("latin-1") + b"\0" + bytes([kind]) + desc.encode(CODECS[enc]) + term + data


def people_frame(fid, pairs, enc=3):
    term = b"\0\0" if enc in (1, 2) else b"\0"
    body = b"".join(a.encode(CODECS[enc]) + term + b.encode(CODECS[enc]) + term for a, b in pairs)
    return fid, bytes([enc]) + body


def make_tag(frames, major=4, padding=0):
    def ss(v):
        return bytes([(v >> 21) & 0x7F, (v >> 14) & 0x7F, (v >> 7) & 0x7F, v & 0x7F])
    body = b""
    for fid, data in frames:
        body += fid.encode("ascii") + (ss(len(data)) if major == 4 else struct.pack(">I", len(data))) + b"\0\0" + data
    body += bytes(padding)
    return b"ID3" + bytes([major, 0, 0]) + ss(len(body)) + body
