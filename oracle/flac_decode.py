"""Independent FLAC decoder -- TEST INFRASTRUCTURE (only tests/ may import it).

Checker for dsd2dxd_b200/csrc/d2p_flac.cpp: SURVEY.md 8(c) asks for "decoded PCM equality" for the FLAC back end
(the reference hands its samples to the flac-codec crate, rd/audio_file.rs:454-573; the bitstream is not part of
the parity contract).  Written from the published FLAC format: metadata blocks, frame header with UTF-8 coded
frame number, CONSTANT / VERBATIM / FIXED / LPC subframes, partitioned Rice (4- and 5-bit parameters, escapes),
wasted bits, the three stereo decorrelation modes, CRC-8 of every header, CRC-16 of every frame, and the MD5 of
the decoded audio against STREAMINFO.  Pure-Python bit reading: for small files.
"""
import hashlib
import struct


class FlacError(ValueError):
    pass


def _crc8(data):
    c = 0
    for b in data:
        c ^= b
        for _ in range(8):
            c = ((c << 1) ^ 0x07) & 0xFF if c & 0x80 else (c << 1) & 0xFF
    return c


_CRC16 = []
for _i in range(256):
    _c = _i << 8
    for _ in range(8):
        _c = ((_c << 1) ^ 0x8005) & 0xFFFF if _c & 0x8000 else (_c << 1) & 0xFFFF
    _CRC16.append(_c)


def _crc16(data):
    c = 0
    for b in data:
        c = ((c << 8) & 0xFFFF) ^ _CRC16[(c >> 8) ^ b]
    return c


class _Bits:
    def __init__(self, data, pos):
        self.d, self.byte, self.bit = data, pos, 0          # bit: bits already consumed of d[byte]

    def read(self, n):
        v = 0
        while n > 0:
            if self.byte >= len(self.d):
                raise FlacError("ran past the end of the stream")
            avail = 8 - self.bit
            take = min(avail, n)
            cur = self.d[self.byte]
            v = (v << take) | ((cur >> (avail - take)) & ((1 << take) - 1))
            self.bit += take
            n -= take
            if self.bit == 8:
                self.bit, self.byte = 0, self.byte + 1
        return v

    def signed(self, n):
        v = self.read(n)
        return v - (1 << n) if n and v >> (n - 1) else v

    def unary(self):
        q = 0
        while self.read(1) == 0:
            q += 1
        return q

    def align(self):
        if self.bit:
            if self.read(8 - self.bit) != 0:
                raise FlacError("non-zero padding bits")


_BLOCK = {1: 192, 2: 576, 3: 1152, 4: 2304, 5: 4608, 8: 256, 9: 512, 10: 1024, 11: 2048, 12: 4096, 13: 8192, 14: 16384, 15: 32768}
_RATE = {1: 88200, 2: 176400, 3: 192000, 4: 8000, 5: 16000, 6: 22050, 7: 24000, 8: 32000, 9: 44100, 10: 48000, 11: 96000}
_BITS = {1: 8, 2: 12, 4: 16, 5: 20, 6: 24, 7: 32}
_FIXED = {0: [], 1: [1], 2: [2, -1], 3: [3, -3, 1], 4: [4, -6, 4, -1]}


def _residual(b, n, order):
    method = b.read(2)
    if method > 1:
        raise FlacError("reserved residual coding method")
    pbits = 5 if method else 4
    porder = b.read(4)
    if n % (1 << porder):
        raise FlacError("partition order does not divide the block size")
    plen = n >> porder
    out = []
    for q in range(1 << porder):
        cnt = plen - order if q == 0 else plen
        if cnt < 0:
            raise FlacError("first partition shorter than the predictor order")
        k = b.read(pbits)
        if k == (1 << pbits) - 1:
            raw = b.read(5)
            out.extend(b.signed(raw) if raw else 0 for _ in range(cnt))
        else:
            for _ in range(cnt):
                u = (b.unary() << k) | b.read(k)
                out.append((u >> 1) ^ -(u & 1))
    return out


def _subframe(b, n, bps):
    if b.read(1):
        raise FlacError("subframe padding bit set")
    kind = b.read(6)
    wasted = 0
    if b.read(1):
        wasted = b.unary() + 1
        bps -= wasted
    if kind == 0:
        x = [b.signed(bps)] * n
    elif kind == 1:
        x = [b.signed(bps) for _ in range(n)]
    elif 8 <= kind <= 12:
        order = kind - 8
        x = [b.signed(bps) for _ in range(order)]
        co = _FIXED[order]
        for r in _residual(b, n, order):
            x.append(r + sum(c * x[-1 - i] for i, c in enumerate(co)))
    elif kind >= 32:
        order = kind - 31
        x = [b.signed(bps) for _ in range(order)]
        prec = b.read(4) + 1
        shift = b.signed(5)
        co = [b.signed(prec) for _ in range(order)]
        for r in _residual(b, n, order):
            x.append(r + (sum(c * x[-1 - i] for i, c in enumerate(co)) >> shift))
    else:
        raise FlacError(f"reserved subframe type {kind}")
    return [v << wasted for v in x] if wasted else x


def decode(data):
    """bytes of a .flac file -> dict(rate, channels, bits, total, samples=[[...] per channel], comments, vendor,
    md5_ok, min_block, max_block, frames).  Raises FlacError on any structural or CRC problem."""
    data = bytes(data)
    if data[:4] != b"fLaC":
        raise FlacError("no fLaC marker")
    pos = 4
    info = None
    vendor, comments, pictures = None, [], []
    while True:
        hdr = data[pos]
        size = int.from_bytes(data[pos + 1:pos + 4], "big")
        body = data[pos + 4:pos + 4 + size]
        if len(body) != size:
            raise FlacError("truncated metadata block")
        kind = hdr & 0x7F
        if kind == 0:
            if size != 34:
                raise FlacError("STREAMINFO must be 34 bytes")
            mn, mx = struct.unpack(">HH", body[:4])
            fmin, fmax = int.from_bytes(body[4:7], "big"), int.from_bytes(body[7:10], "big")
            packed = int.from_bytes(body[10:18], "big")
            info = {"min_block": mn, "max_block": mx, "min_frame": fmin, "max_frame": fmax, "rate": packed >> 44,
                    "channels": ((packed >> 41) & 7) + 1, "bits": ((packed >> 36) & 31) + 1, "total": packed & ((1 << 36) - 1),
                    "md5": body[18:34]}
        elif kind == 4:
            n = struct.unpack_from("<I", body, 0)[0]
            vendor = body[4:4 + n].decode()
            q = 4 + n
            cnt = struct.unpack_from("<I", body, q)[0]
            q += 4
            for _ in range(cnt):
                n = struct.unpack_from("<I", body, q)[0]
                comments.append(body[q + 4:q + 4 + n].decode())
                q += 4 + n
            if q != size:
                raise FlacError("VORBIS_COMMENT length mismatch")
        elif kind == 6:                                            # PICTURE: big-endian fields
            ptype, ml = struct.unpack_from(">II", body, 0)
            mime = body[8:8 + ml].decode("ascii")
            q = 8 + ml
            dl = struct.unpack_from(">I", body, q)[0]
            desc = body[q + 4:q + 4 + dl].decode("utf-8")
            q += 4 + dl
            w, h, depth, colors, n = struct.unpack_from(">IIIII", body, q)
            q += 20
            if q + n != size:
                raise FlacError("PICTURE length mismatch")
            pictures.append({"type": ptype, "mime": mime, "description": desc, "width": w, "height": h, "depth": depth,
                             "colors": colors, "data": body[q:q + n]})
        elif kind == 127:
            raise FlacError("invalid metadata block type")
        pos += 4 + size
        if hdr & 0x80:
            break
    if info is None:
        raise FlacError("no STREAMINFO")
    C, bps = info["channels"], info["bits"]
    chans = [[] for _ in range(C)]
    nframes, sizes = 0, []
    while pos < len(data):
        start = pos
        b = _Bits(data, pos)
        if b.read(14) != 0x3FFE or b.read(1):
            raise FlacError(f"bad frame sync at byte {pos}")
        variable = b.read(1)
        bs_code, sr_code, ch_code, ss_code = b.read(4), b.read(4), b.read(4), b.read(3)
        if b.read(1):
            raise FlacError("reserved header bit set")
        lead = b.read(8)
        number = lead
        if lead & 0x80:
            extra = 0
            while lead & (0x40 >> extra):
                extra += 1
            if extra == 0 or extra > 6:
                raise FlacError("bad UTF-8 lead byte in the frame number")
            number = lead & (0x3F >> extra)
            for _ in range(extra):
                cont = b.read(8)
                if cont & 0xC0 != 0x80:
                    raise FlacError("bad UTF-8 continuation in the frame number")
                number = (number << 6) | (cont & 0x3F)
        if bs_code == 6:
            n = b.read(8) + 1
        elif bs_code == 7:
            n = b.read(16) + 1
        elif bs_code in _BLOCK:
            n = _BLOCK[bs_code]
        else:
            raise FlacError("reserved block size code")
        if sr_code == 0:
            rate = info["rate"]
        elif sr_code == 12:
            rate = b.read(8) * 1000
        elif sr_code == 13:
            rate = b.read(16)
        elif sr_code == 14:
            rate = b.read(16) * 10
        elif sr_code in _RATE:
            rate = _RATE[sr_code]
        else:
            raise FlacError("invalid sample rate code")
        if rate != info["rate"]:
            raise FlacError("frame sample rate differs from STREAMINFO")
        fb = bps if ss_code == 0 else _BITS.get(ss_code)
        if fb != bps:
            raise FlacError("frame sample size differs from STREAMINFO")
        if _crc8(data[start:b.byte]) != b.read(8):
            raise FlacError("frame header CRC-8 mismatch")
        if not variable and number != nframes:
            raise FlacError(f"frame number {number}, expected {nframes}")
        if ch_code < 8:
            if ch_code + 1 != C:
                raise FlacError("frame channel count differs from STREAMINFO")
            sub = [_subframe(b, n, bps) for _ in range(C)]
        elif ch_code == 8:
            l, s = _subframe(b, n, bps), _subframe(b, n, bps + 1)
            sub = [l, [a - d for a, d in zip(l, s)]]
        elif ch_code == 9:
            s, r = _subframe(b, n, bps + 1), _subframe(b, n, bps)
            sub = [[a + d for a, d in zip(r, s)], r]
        elif ch_code == 10:
            m, s = _subframe(b, n, bps), _subframe(b, n, bps + 1)
            sub = [[(((a << 1) | (d & 1)) + d) >> 1 for a, d in zip(m, s)], [(((a << 1) | (d & 1)) - d) >> 1 for a, d in zip(m, s)]]
        else:
            raise FlacError("reserved channel assignment")
        b.align()
        if _crc16(data[start:b.byte]) != b.read(16):
            raise FlacError("frame CRC-16 mismatch")
        pos = b.byte
        sizes.append(pos - start)
        for c in range(C):
            chans[c].extend(sub[c])
        nframes += 1
    nbytes = (bps + 7) // 8
    h = hashlib.md5()
    inter = bytearray()
    for i in range(len(chans[0])):
        for c in range(C):
            inter += (chans[c][i] & ((1 << (8 * nbytes)) - 1)).to_bytes(nbytes, "little")
    h.update(bytes(inter))
    out = dict(info)
    out.update(samples=chans, vendor=vendor, comments=comments, pictures=pictures, md5_ok=h.digest() == info["md5"], frames=nframes,
               frame_sizes=sizes)
    return out
