"""CPU restatement of the reference's PCM file writers and of the DSF / DSDIFF layouts -- TEST INFRASTRUCTURE.

Only tests/ may import this module (it is the checker for dsd2dxd_b200/csrc/d2p_files.cpp, never the product).

Writers follow /root/reference/src/rdsd2pcm/src/audio_file.rs:
    wav_file   save_wave_file   :93-207        aiff_file  save_aiff_file  :209-292
    aifc_file  save_aifc_file   :294-452       extended   encode_extended :576-621
Input is what AudioFile holds: planar samples[channel][frame], int32 (20-bit NOT shifted) or float32.
Readers: the dsd-reader crate is absent from the reference mount -- parity unpinned; parse_dsf /
parse_dff restate the published DSF 1.01 / DSDIFF 1.5 layouts independently of the C++ (struct-based), and
make_dsf / make_dff build synthetic containers for the round-trip tests.
"""
import struct

import numpy as np


def extended(value):
    """encode_extended :576-621 -- 80-bit IEEE extended, integer bit explicit."""
    out = bytearray(10)
    if value <= 0.0:
        return bytes(out)
    exp = 0
    while value >= 2.0:
        value *= 0.5
        exp += 1
    while value < 1.0:
        value *= 2.0
        exp -= 1
    exp_field = (exp + 16383) & 0xFFFF
    mant = int(round(value * float(1 << 63)))
    if mant == 1 << 64:
        mant = 1 << 63
        exp_field = (exp_field + 1) & 0xFFFF
    return struct.pack(">HQ", exp_field, mant)


def _bps(bits):
    return 3 if bits == 20 else bits // 8


def _payload(samples, bits, big_endian):
    """The per-frame, per-channel packing loops of the three writers."""
    samples = np.asarray(samples)
    ch, n = samples.shape
    inter = samples.T.reshape(-1)                       # frame-major, channel-minor
    if bits == 32:
        return inter.astype(">f4" if big_endian else "<f4").tobytes()
    if bits == 16:
        return inter.astype(np.int16).astype(">i2" if big_endian else "<i2").tobytes()
    v = inter.astype(np.int64)
    if bits == 20:
        v = v << 4
    b = np.stack([(v >> 16) & 0xFF, (v >> 8) & 0xFF, v & 0xFF] if big_endian else [v & 0xFF, (v >> 8) & 0xFF, (v >> 16) & 0xFF],
                 axis=1)
    return b.astype(np.uint8).tobytes()


def wav_file(samples, bits, rate):
    ch, frames = np.asarray(samples).shape
    bps = _bps(bits)
    data_size = (frames * ch * bps) & 0xFFFFFFFF
    h = b"RIFF" + struct.pack("<I", (data_size + 36) & 0xFFFFFFFF) + b"WAVE"
    h += b"fmt " + struct.pack("<IHHIIHH", 16, 3 if bits == 32 else 1, ch, rate, (rate * ch * bps) & 0xFFFFFFFF, ch * bps, bits)
    h += b"data" + struct.pack("<I", data_size)
    return h + _payload(samples, bits, False)


def aiff_file(samples, bits, rate):
    if bits == 32:
        raise ValueError("Unsupported bit depth")       # :286-290
    ch, frames = np.asarray(samples).shape
    data_size = (frames * ch * _bps(bits)) & 0xFFFFFFFF
    h = b"FORM" + struct.pack(">I", (data_size + 46) & 0xFFFFFFFF) + b"AIFF"
    h += b"COMM" + struct.pack(">IHIH", 18, ch, frames & 0xFFFFFFFF, bits) + extended(float(rate))
    h += b"SSND" + struct.pack(">III", data_size, 0, 0)
    return h + _payload(samples, bits, True)


def aifc_file(samples, bits, rate):
    ch, frames = np.asarray(samples).shape
    data_size = (frames * ch * _bps(bits)) & 0xFFFFFFFF
    if bits == 32:
        ctype, name = b"fl32", b"Linear PCM, 32 bit big-endian floating point"
    elif bits == 16:
        ctype, name = b"twos", b"Linear PCM, 16 bit big-endian signed integer"
    else:
        ctype, name = b"NONE", b"Linear PCM, 24 bit big-endian signed integer"
    pname = bytes([len(name)]) + name
    comm_size = 22 + len(pname)
    comm_padded = (comm_size + 1) & ~1
    r = (32 + comm_padded + 24) % 4096
    fllr = 0 if r == 0 else 4096 - r
    fllr_chunk = 8 + fllr if fllr > 0 else 0
    h = b"FORM" + struct.pack(">I", (40 + comm_padded + fllr_chunk + data_size) & 0xFFFFFFFF) + b"AIFC"
    h += b"FVER" + struct.pack(">II", 4, 0xA2805140)
    h += b"COMM" + struct.pack(">IHIH", comm_padded, ch, frames & 0xFFFFFFFF, bits) + extended(float(rate)) + ctype + pname
    if comm_size % 2:
        h += b"\0"
    if fllr > 0:
        h += b"FLLR" + struct.pack(">I", fllr) + bytes(fllr)
    h += b"SSND" + struct.pack(">III", (data_size + 8) & 0xFFFFFFFF, 0, 0)
    return h + _payload(samples, bits, True)


WRITERS = {"W": wav_file, "A": aiff_file, "C": aifc_file}


# ---- containers ---------------------------------------------------------------------------------------------
def parse_dsf(data):
    data = bytes(data)
    magic, size, total, meta = struct.unpack_from("<4sQQQ", data, 0)
    assert magic == b"DSD " and size == 28
    fm, fsize, ver, fid, ctype, ch, fs, bps, count, block, _ = struct.unpack_from("<4sQIIIIIIQII", data, 28)
    assert fm == b"fmt " and ver == 1 and fid == 0
    off = 28 + fsize
    dm, dsize = struct.unpack_from("<4sQ", data, off)
    assert dm == b"data"
    start = off + 12
    end = min(len(data), start + dsize - 12)
    if start < meta < len(data):
        end = min(end, meta)
    nbytes = (end - start) // (ch * block) * (ch * block)
    return {"channels": ch, "sample_rate": fs, "dsd_rate": fs // 2822400, "lsb_first": 1 if bps == 1 else 0,
            "block_size": block, "sample_count": count, "data_offset": start, "data_bytes": nbytes,
            "id3_offset": meta if 0 < meta < len(data) else 0, "id3_bytes": len(data) - meta if 0 < meta < len(data) else 0}


def parse_dff(data):
    data = bytes(data)
    assert data[:4] == b"FRM8" and data[12:16] == b"DSD "
    out = {"lsb_first": 0, "block_size": 1, "id3_offset": 0, "id3_bytes": 0}
    pos = 16
    while pos + 12 <= len(data):
        cid, size = struct.unpack_from(">4sQ", data, pos)
        body = pos + 12
        clipped = min(size, len(data) - body)
        if cid == b"PROP" and data[body:body + 4] == b"SND ":
            q = body + 4
            while q + 12 <= body + clipped:
                sid, ssz = struct.unpack_from(">4sQ", data, q)
                if sid == b"FS  ":
                    out["sample_rate"] = struct.unpack_from(">I", data, q + 12)[0]
                elif sid == b"CHNL":
                    out["channels"] = struct.unpack_from(">H", data, q + 12)[0]
                elif sid == b"CMPR":
                    assert data[q + 12:q + 16] == b"DSD "
                q += 12 + ((ssz + 1) & ~1)
        elif cid == b"DSD ":
            out["data_offset"], out["data_bytes"] = body, clipped
        elif cid == b"ID3 ":
            out["id3_offset"], out["id3_bytes"] = body, clipped
        if size > len(data) - body:
            break
        pos = body + ((size + 1) & ~1)
    out["data_bytes"] = out["data_bytes"] // out["channels"] * out["channels"]
    out["dsd_rate"] = out["sample_rate"] // 2822400
    out["sample_count"] = out["data_bytes"] // out["channels"] * 8
    return out


def make_dsf(payload, channels, dsd_rate, sample_count=None, lsb_first=True, id3=b"", block=4096):
    payload = bytes(payload)
    per_ch = len(payload) // channels
    if sample_count is None:
        sample_count = per_ch * 8
    total = 92 + len(payload) + len(id3)
    h = struct.pack("<4sQQQ", b"DSD ", 28, total, 92 + len(payload) if id3 else 0)
    h += struct.pack("<4sQIIIIIIQII", b"fmt ", 52, 1, 0, 2 if channels == 2 else channels, channels, 2822400 * dsd_rate,
                     1 if lsb_first else 8, sample_count, block, 0)
    h += struct.pack("<4sQ", b"data", 12 + len(payload))
    return h + payload + id3


def make_dff(payload, channels, dsd_rate, id3=b"", wrong_frm8=False):
    payload = bytes(payload)
    ids = [b"SLFT", b"SRGT", b"C   ", b"LFE ", b"LS  ", b"RS  "][:channels] if channels != 1 else [b"C   "]
    snd = b"SND "
    snd += b"FS  " + struct.pack(">QI", 4, 2822400 * dsd_rate)
    snd += b"CHNL" + struct.pack(">QH", 2 + 4 * channels, channels) + b"".join(ids)
    cname = b"\x0enot compressed"                          # pstring, 15 bytes -> padded to 16 with the count byte
    cm = b"DSD " + cname + b"\0"
    snd += b"CMPR" + struct.pack(">Q", len(cm) - 1) + cm     # odd size + pad byte
    body = b"FVER" + struct.pack(">QI", 4, 0x01050000)
    body += b"PROP" + struct.pack(">Q", len(snd)) + snd
    body += b"DSD " + struct.pack(">Q", len(payload)) + payload + (b"\0" if len(payload) % 2 else b"")
    if id3:
        body += b"ID3 " + struct.pack(">Q", len(id3)) + id3 + (b"\0" if len(id3) % 2 else b"")
    size = 4 + len(body)
    return b"FRM8" + struct.pack(">Q", size * 2 if wrong_frm8 else size) + b"DSD " + body
