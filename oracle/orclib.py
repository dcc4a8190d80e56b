"""ctypes binding of the CPU parity oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs -- never by the dsd2dxd_b200 package.  See oracle.cpp for what each entry restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")


class OrcConfig(C.Structure):
    _fields_ = [
        ("dsd_rate", C.c_int), ("out_rate", C.c_uint32), ("filter", C.c_int), ("bits", C.c_int),
        ("dither", C.c_int), ("level_db", C.c_double), ("channels", C.c_int), ("layout", C.c_int),
        ("lsb_first", C.c_int), ("block_size", C.c_uint32), ("dither_seed", C.c_uint64),
        ("has_dither_scale", C.c_int), ("dither_scale_db", C.c_double), ("stdout_mode", C.c_int),
    ]


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("oracle.cpp", "ref_taps.inc")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        u8p, f64p = C.POINTER(C.c_uint8), C.POINTER(C.c_double)
        L.orc_num_tables.restype = C.c_int
        L.orc_table_name.restype = C.c_char_p
        L.orc_table_name.argtypes = [C.c_int]
        L.orc_table_len.restype = C.c_int
        L.orc_table_len.argtypes = [C.c_int]
        L.orc_table_data.restype = f64p
        L.orc_table_data.argtypes = [C.c_int]
        L.orc_table_id.restype = C.c_int
        L.orc_table_id.argtypes = [C.c_char_p]
        L.orc_select_precalc_taps.restype = C.c_int
        L.orc_select_precalc_taps.argtypes = [C.c_int, C.c_int, C.c_int]
        L.orc_compute_decim_and_upsample.argtypes = [C.c_int, C.c_uint32, C.POINTER(C.c_int), C.POINTER(C.c_uint32)]
        L.orc_calc_frames_cap.restype = C.c_size_t
        L.orc_calc_frames_cap.argtypes = [C.c_size_t, C.c_int, C.c_uint32]
        L.orc_bpd_new.restype = C.c_void_p
        L.orc_bpd_new.argtypes = [f64p, C.c_size_t, C.c_uint32]
        L.orc_bpd_process.restype = C.c_size_t
        L.orc_bpd_process.argtypes = [C.c_void_p, u8p, C.c_size_t, f64p, C.c_size_t]
        L.orc_bpd_num_tables.restype = C.c_int
        L.orc_bpd_num_tables.argtypes = [C.c_void_p]
        L.orc_bpd_table.restype = f64p
        L.orc_bpd_table.argtypes = [C.c_void_p, C.c_int]
        L.orc_bpd_free.argtypes = [C.c_void_p]
        L.orc_lm_new.restype = C.c_void_p
        L.orc_lm_new.argtypes = [C.c_uint32, C.c_int, C.c_uint32, C.c_char_p, C.c_size_t]
        L.orc_lm_process.restype = C.c_size_t
        L.orc_lm_process.argtypes = [C.c_void_p, u8p, C.c_size_t, f64p, C.c_size_t]
        L.orc_lm_free.argtypes = [C.c_void_p]
        L.orc_stage1_run.restype = C.c_size_t
        L.orc_stage1_run.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, u8p, C.c_size_t, f64p, C.c_size_t]
        L.orc_decimfir_run.restype = C.c_size_t
        L.orc_decimfir_run.argtypes = [C.c_char_p, C.c_size_t, f64p, C.c_size_t, f64p, C.c_size_t]
        L.orc_my_round.restype = C.c_int64
        L.orc_my_round.argtypes = [C.c_double]
        L.orc_scale_factor.restype = C.c_double
        L.orc_scale_factor.argtypes = [C.c_int, C.c_double, C.c_uint32]
        L.orc_convert.restype = C.c_int
        L.orc_convert.argtypes = [C.POINTER(OrcConfig), u8p, C.c_size_t, u8p, C.c_size_t, C.POINTER(C.c_size_t),
                                  C.POINTER(C.c_int32), C.POINTER(C.c_float), f64p, C.c_size_t,
                                  C.POINTER(C.c_uint64), C.POINTER(C.c_int32), C.c_char_p, C.c_size_t]
        L.orc_bench.restype = C.c_double
        L.orc_bench.argtypes = [C.POINTER(OrcConfig), u8p, C.c_size_t, C.c_int, C.c_int]
        _lib = L
    return _lib


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _f64(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def table_names():
    L = lib()
    return [L.orc_table_name(i).decode() for i in range(L.orc_num_tables())]


def table(name):
    L = lib()
    i = L.orc_table_id(name.encode())
    if i < 0:
        raise KeyError(name)
    n = L.orc_table_len(i)
    return np.ctypeslib.as_array(L.orc_table_data(i), shape=(n,)).copy()


def select_precalc_taps(filt, dsd_rate, decim):
    """-> table name or None (rd/byte_precalc_decimator.rs:167)."""
    L = lib()
    i = L.orc_select_precalc_taps(ord(filt), dsd_rate, decim)
    return None if i < 0 else L.orc_table_name(i).decode()


def compute_decim_and_upsample(in_rate, out_rate):
    m, l = C.c_int(), C.c_uint32()
    lib().orc_compute_decim_and_upsample(in_rate, out_rate, C.byref(m), C.byref(l))
    return m.value, l.value


def calc_frames_cap(block, decim, up):
    return lib().orc_calc_frames_cap(block, decim, up)


class BytePrecalcDecimator:
    def __init__(self, half, decim):
        half = np.ascontiguousarray(half, dtype=np.float64)
        self._h = lib().orc_bpd_new(_f64(half), half.size, decim)
        if not self._h:
            raise ValueError("BytePrecalcDecimator::new returned None")

    def process_bytes(self, data, cap):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        out = np.zeros(cap, dtype=np.float64)
        n = lib().orc_bpd_process(self._h, _u8(data), data.size, _f64(out), cap)
        return out[:n]

    def tables(self):
        L = lib()
        nt = L.orc_bpd_num_tables(self._h)
        return np.stack([np.ctypeslib.as_array(L.orc_bpd_table(self._h, i), shape=(256,)).copy() for i in range(nt)])

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_bpd_free(self._h)
            self._h = None


class LMResampler:
    def __init__(self, l, m, out_rate):
        err = C.create_string_buffer(256)
        self._h = lib().orc_lm_new(l, m, out_rate, err, 256)
        if not self._h:
            raise ValueError(err.value.decode())

    def process_bytes_lm(self, data, cap):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        out = np.zeros(cap, dtype=np.float64)
        n = lib().orc_lm_process(self._h, _u8(data), data.size, _f64(out), cap)
        return out[:n]

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_lm_free(self._h)
            self._h = None


def stage1_run(table_name, l, m, data):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    cap = data.size * 8 * l // m + 64
    out = np.zeros(cap, dtype=np.float64)
    n = lib().orc_stage1_run(table_name.encode(), l, m, _u8(data), data.size, _f64(out), cap)
    assert n <= cap
    return out[:n]


def decimfir_run(table_name, decim, x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    cap = x.size // decim + 8
    out = np.zeros(cap, dtype=np.float64)
    n = lib().orc_decimfir_run(table_name.encode(), decim, _f64(x), x.size, _f64(out), cap)
    return out[:n]


def make_config(dsd_rate=1, out_rate=352800, filt="E", bits=24, dither="X", level_db=0.0, channels=2,
                layout="P", lsb_first=True, block_size=4096, dither_seed=0, dither_scale_db=None,
                stdout_mode=True):
    return OrcConfig(dsd_rate, out_rate, ord(filt), bits, ord(dither), level_db, channels, ord(layout),
                     1 if lsb_first else 0, block_size, dither_seed,
                     0 if dither_scale_db is None else 1, 0.0 if dither_scale_db is None else dither_scale_db,
                     1 if stdout_mode else 0)


class ConvertResult:
    pass


def convert(raw, want_f64=True, **kw):
    """Whole-file conversion through the restated process_blocks loop.  Returns ConvertResult with
    pcm (bytes, stdout mode), planar (int32/float32 [ch, frames], file mode), f64 [ch, frames], frames, clips.
    Raises ValueError(message) where the reference returns Err(message)."""
    cfg = make_config(**kw)
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    m, l = compute_decim_and_upsample(cfg.dsd_rate, cfg.out_rate)
    per_ch = raw.size // cfg.channels
    frames_cap = per_ch * 8 * l // abs(m) + 64
    bps = 3 if cfg.bits == 20 else cfg.bits // 8
    pcm = np.zeros(frames_cap * cfg.channels * bps, dtype=np.uint8)
    pcm_bytes = C.c_size_t()
    pi32 = np.zeros((cfg.channels, frames_cap), dtype=np.int32)
    pf32 = np.zeros((cfg.channels, frames_cap), dtype=np.float32)
    f64 = np.zeros((cfg.channels, frames_cap), dtype=np.float64) if want_f64 else None
    frames = (C.c_uint64 * cfg.channels)()
    clips = C.c_int32()
    err = C.create_string_buffer(256)
    rc = lib().orc_convert(C.byref(cfg), _u8(raw), raw.size, _u8(pcm), pcm.size, C.byref(pcm_bytes),
                           pi32.ctypes.data_as(C.POINTER(C.c_int32)), pf32.ctypes.data_as(C.POINTER(C.c_float)),
                           _f64(f64) if want_f64 else None, frames_cap, frames, C.byref(clips), err, 256)
    if rc:
        raise ValueError(err.value.decode())
    r = ConvertResult()
    r.frames = [int(x) for x in frames]
    n = r.frames[0] if r.frames else 0
    r.pcm = pcm[:pcm_bytes.value].copy()
    r.planar = (pf32 if cfg.bits == 32 else pi32)[:, :n].copy()
    r.f64 = f64[:, :n].copy() if want_f64 else None
    r.clips = clips.value
    r.scale = lib().orc_scale_factor(cfg.bits, cfg.level_db, l)
    return r


def check_level(raw, dsd_rate=1, out_rate=352800, channels=2, layout="P", lsb_first=True, block_size=4096):
    """Rdsd2Pcm::new_level_check + process_levels (rd/lib.rs:124-158, rd/conversion_context.rs:209-256):
    32-bit sink, no dither, 0 dB, equiripple; peak_abs = max |s| * scale_factor over every channel's samples,
    peak_dbfs = (20 * log10(peak_abs)) as f32, -inf when peak_abs <= 0.  -> (peak_abs, np.float32 dBFS)."""
    r = convert(raw, dsd_rate=dsd_rate, out_rate=out_rate, filt="E", bits=32, dither="X", level_db=0.0,
                channels=channels, layout=layout, lsb_first=lsb_first, block_size=block_size)
    peak = float(np.max(np.abs(r.f64) * r.scale)) if r.f64.size else 0.0
    return peak, (np.float32(-np.inf) if peak <= 0.0 else np.float32(20.0 * np.log10(peak)))


def bench(raw, nfiles, nthreads, **kw):
    cfg = make_config(**kw)
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    s = lib().orc_bench(C.byref(cfg), _u8(raw), raw.size, nfiles, nthreads)
    if s < 0:
        raise RuntimeError("oracle bench failed")
    return s
