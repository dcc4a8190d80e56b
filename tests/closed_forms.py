"""Closed-form (index-based, stateless) numpy restatement of the rdsd2pcm hot path.

Independent of oracle/oracle.cpp's streaming state machines: every output sample is written as a
pure function of absolute input positions (SURVEY.md section 8a "closed form" columns).  The CUDA
kernels implement exactly these formulas; tests/test_oracle_kat.py proves they equal the literal
state machines, which is what lets any time block be computed independently on the GPU.

All folds keep the reference's f64 operation order (numpy elementwise ops round each operation).
"""
import numpy as np

REV8 = np.array([int(f"{b:08b}"[::-1], 2) for b in range(256)], dtype=np.uint8)


def next_pow2(x):
    p = 1
    while p < x:
        p <<= 1
    return p


# ---------------------------------------------------------------------------------------------
# Integer-ratio LUT FIR (A3/A4)   rd/byte_precalc_decimator.rs:67-162
# ---------------------------------------------------------------------------------------------
def bpd_tables(half):
    """tables[i][b], i = 0..nt-1, covering taps t..t+7 with t = (nt-1-i)*8; left fold from 0.0."""
    half = np.asarray(half, dtype=np.float64)
    h = half.size
    nt = (h + 7) // 8
    b = np.arange(256)
    tables = np.zeros((nt, 256), dtype=np.float64)
    for i in range(nt):
        t = (nt - 1 - i) * 8
        acc = np.zeros(256, dtype=np.float64)
        for bit in range(min(8, h - t)):
            sign = (((b >> bit) & 1) * 2 - 1).astype(np.float64)
            acc = acc + sign * half[t + bit]
        tables[i] = acc
    return tables


def bpd_closed(chan_bytes, half, decim, first=0, count=None):
    """Emitted frames n = first..first+count-1 of one channel (LSB-first bytes, whole stream)."""
    B = np.asarray(chan_bytes, dtype=np.uint8)
    h = len(half)
    nt = (h + 7) // 8
    bpo = decim // 8
    total = max(0, B.size // bpo - (h - 1))
    if count is None:
        count = total - first
    count = max(0, min(count, total - first))
    T = bpd_tables(half)
    n = np.arange(first, first + count, dtype=np.int64)
    j = (n + (h - 1) + 1) * bpo - 1          # newest byte of the frame
    acc = np.zeros(count, dtype=np.float64)
    for i in range(nt):
        b1 = B[j - i]
        b2 = REV8[B[j - (2 * nt - 1 - i)]]
        acc = acc + T[i][b1]
        acc = acc + T[i][b2]
    return acc


def bpd_frames(nbytes, h, decim):
    return max(0, nbytes // (decim // 8) - (h - 1))


# ---------------------------------------------------------------------------------------------
# L/M resampler (A6-A10)   rd/lm_resampler.rs
# ---------------------------------------------------------------------------------------------
# (dsd_rate, out_rate) -> (L, M, [(table, L, m1)], [(table, D2)], [(table, D3)] or None)
def lm_topology(l, m, out_rate):
    if m == 588:
        if out_rate == 96000 and l == 5:
            return ("HTAPS_DDRX10_21TO1_EQ", 21), ("HTAPS_2_68MHZ_28TO1_EQ", 28), None
        raise ValueError("Unsupported 588: must be out_rate=96k and L=5")
    if m == 294:
        if out_rate == 192000 and l == 5:
            return ("HTAPS_DDRX10_21TO1_EQ", 21), ("HTAPS_2_68MHZ_14TO1_EQ", 14), None
        return ("HTAPS_DDRX5_14TO1_EQ", 14), ("HTAPS_2MHZ_7TO1_EQ", 7), ("HTAPS_288K_3TO1_EQ", 3)
    if m == 147:
        if out_rate == 384000:
            return ("HTAPS_DDRX10_21TO1_EQ", 21), ("HTAPS_2_68MHZ_7TO1_EQ", 7), None
        if out_rate == 192000:
            if l == 5:
                return ("HTAPS_DDRX5_7TO_1_EQ", 7), ("HTAPS_4MHZ_7TO1_EQ", 7), ("HTAPS_576K_3TO1_EQ", 3)
            if l == 10:
                return ("HTAPS_DSDX10_21TO1_EQ", 21), ("HTAPS_1_34MHZ_7TO1_EQ", 7), None
        return ("HTAPS_DSDX5_7TO1_EQ", 7), ("HTAPS_2MHZ_7TO1_EQ", 7), ("HTAPS_288K_3TO1_EQ", 3)
    raise ValueError("Unsupported L/M combination")


def stage1_lut(half, l):
    """LUT[p][g][b] with per-phase group counts (ragged list of arrays)."""
    half = np.asarray(half, dtype=np.float64)
    full = np.concatenate([half[::-1], half])
    b = np.arange(256)
    out = []
    for p in range(l):
        taps = full[p::l]
        groups = (taps.size + 7) // 8
        lut = np.zeros((groups, 256), dtype=np.float64)
        for g in range(groups):
            s = np.zeros(256, dtype=np.float64)
            for jj in range(8):
                c = taps[8 * g + jj] if 8 * g + jj < taps.size else 0.0
                sign = np.where(((b >> jj) & 1) == 1, 1.0, -1.0)
                s = s + c * sign
            lut[g] = s
        out.append(lut)
    return out


def stage1_schedule(h1, l, m, nbits):
    """High-rate indices n_hi of every stage-1 output produced from `nbits` input bits."""
    if m == 21:   # Accum mode
        input_delay = (h1 - 1 + l - 1) // l
        e0 = 0
        while -(-(e0 + 1) * 21 // l) < input_delay:   # ceil((e+1)*21/l) >= input_delay
            e0 += 1
        e_last = nbits * l // 21 - 1                    # (e+1)*21 <= nbits*l
        e = np.arange(e0, e_last + 1, dtype=np.int64)
        return (e + 1) * 21 - 1
    n_last = (nbits * l - 1 - (h1 - 1)) // m - 1         # (h1-1) + m(n+1) <= nbits*l - 1
    n = np.arange(0, n_last + 1, dtype=np.int64)
    return (h1 - 1) + m * (n + 1)


def window8(bits, t):
    """8-bit window ending at input bit t: bit j = x[t-j], pre-stream bits read as 0."""
    w = np.zeros(t.shape, dtype=np.int64)
    for jj in range(8):
        idx = t - jj
        ok = idx >= 0
        w |= (np.where(ok, bits[np.maximum(idx, 0)], 0).astype(np.int64) << jj)
    return w


def stage1_closed(chan_bytes, half, l, m):
    B = np.asarray(chan_bytes, dtype=np.uint8)
    bits = np.unpackbits(B, bitorder="little")
    lut = stage1_lut(half, l)
    gmax = max(x.shape[0] for x in lut)
    cap = max(next_pow2(gmax + 16), 32)       # byte_ring capacity (:425): offsets 8g alias modulo cap
    n_hi = stage1_schedule(len(half), l, m, bits.size)
    t = n_hi // l
    p = n_hi % l
    y = np.zeros(n_hi.size, dtype=np.float64)
    for ph in range(l):
        sel = np.nonzero(p == ph)[0]
        if sel.size == 0:
            continue
        s = np.zeros(sel.size, dtype=np.float64)
        for g in range(lut[ph].shape[0]):
            w = window8(bits, t[sel] - ((8 * g) % cap))
            s = s + lut[ph][g][w]
        y[sel] = s
    return y


def decimfir_closed(x, half, d):
    x = np.asarray(x, dtype=np.float64)
    half = np.asarray(half, dtype=np.float64)
    h = half.size
    full = np.concatenate([half[::-1], half])
    if x.size < h:
        return np.zeros(0)
    nout = (x.size - 1 - (h - 1)) // d + 1
    t = (h - 1) + d * np.arange(nout, dtype=np.int64)
    pad = 2 * h
    xp = np.concatenate([np.zeros(pad), x])
    total = np.zeros(nout, dtype=np.float64)
    for k in range(h):
        xf = xp[pad + t - k]
        xb = xp[pad + t - (2 * h - 1 - k)]
        total = total + full[k] * (xf + xb)
    return total


def lm_closed(chan_bytes, tables, l, m, out_rate):
    """tables: dict name -> half taps.  Returns (y_final, y1, y2)."""
    s1, s2, s3 = lm_topology(l, m, out_rate)
    y1 = stage1_closed(chan_bytes, tables[s1[0]], l, s1[1])
    y2 = decimfir_closed(y1, tables[s2[0]], s2[1])
    if s3 is None:
        return y2, y1, y2
    y3 = decimfir_closed(y2, tables[s3[0]], s3[1])
    return y3, y1, y2


# ---------------------------------------------------------------------------------------------
# scale / quantise / pack (A13, A16, A17)   rd/pcm_writer.rs
# ---------------------------------------------------------------------------------------------
def scale_factor(bits, level_db, l):
    s = 1.0
    if bits != 32:
        s = 2.0 ** (bits - 1)
    s *= 10.0 ** (level_db / 20.0)
    s *= float(l)
    return s


def my_round(x):
    x = np.asarray(x, dtype=np.float64)
    return np.where(x < 0.0, np.floor(x - 0.5), np.floor(x + 0.5)).astype(np.int64)


def quantize(x, bits):
    peak = 1 << (bits - 1)
    v = my_round(x).astype(np.int32)     # i64 -> i32 wrapping (values are tiny in practice)
    return np.clip(v, -peak, peak - 1)


def pack_le(q, bits):
    """q: int32 [frames, channels] (or f32 for bits == 32) -> interleaved little-endian bytes."""
    if bits == 32:
        return np.ascontiguousarray(q, dtype="<f4").view(np.uint8).reshape(-1)
    q = np.ascontiguousarray(q, dtype=np.int32)
    if bits == 16:
        return q.astype("<i2").view(np.uint8).reshape(-1)
    v = q << 4 if bits == 20 else q
    v = v.astype(np.uint32)
    out = np.stack([v & 0xFF, (v >> 8) & 0xFF, (v >> 16) & 0xFF], axis=-1).astype(np.uint8)
    return out.reshape(-1)
