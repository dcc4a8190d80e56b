"""dsd2dxd_b200 -- B200-native (sm_100a CUDA) implementation of the rdsd2pcm DSD->PCM conversion core.

Thin Python view of the C ABI (include/dsd2pcm_b200.h) used by the tests and bench.py.  The parameter
names follow Rdsd2Pcm::new (rd/lib.rs:75-91) and the CLI letters of src/main.rs.  All conversion work
happens inside libdsd2pcm_b200.so on the GPU; nothing here computes samples.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import Config, D2PError, Job, PlanInfo  # noqa: F401

__all__ = ["make_config", "plan", "frames_for_bytes", "tap_table", "tap_table_names", "device_count",
           "Converter", "convert_host", "Batch", "run_noalloc", "D2PError", "HostBuffer", "total_frames", "block_range"]


def make_config(dsd_rate=1, out_rate=352800, filt="E", bits=24, dither="X", level_db=0.0, channels=2,
                layout="P", lsb_first=True, block_size=4096, out_format=_capi.OUT_INTERLEAVED_LE,
                dither_seed=0, dither_scale_db=None, device=0, keep_f64=False):
    """Defaults are the reference CLI's (-c 2 -b 24 -t E -s 4096 -r 352800 -i 1, src/main.rs:40-133)
    except the input format, which defaults here to a DSF-like planar / LSB-first stream."""
    return Config(dsd_rate, out_rate, ord(filt), bits, ord(dither), channels, level_db, ord(layout),
                  1 if lsb_first else 0, block_size, out_format, dither_seed,
                  0 if dither_scale_db is None else 1, 0.0 if dither_scale_db is None else float(dither_scale_db),
                  device, 1 if keep_f64 else 0)


def device_count():
    return _capi.lib().d2p_device_count()


def plan(cfg):
    info = PlanInfo()
    _capi.check(_capi.lib().d2p_plan(C.byref(cfg), C.byref(info)))
    n = info.num_stages
    return {
        "decim_ratio": info.decim_ratio, "upsample_ratio": info.upsample_ratio, "num_stages": n,
        "tables": [bytes(info.stage_table[i]).split(b"\0")[0].decode() for i in range(n)],
        "half_taps": [info.stage_half_taps[i] for i in range(n)],
        "decims": [info.stage_decim[i] for i in range(n)],
        "scale_factor": info.scale_factor, "peak_level": info.peak_level,
        "bytes_per_sample": info.bytes_per_sample, "frames_cap": info.frames_cap, "lut_bytes": info.lut_bytes,
    }


def frames_for_bytes(cfg, bytes_per_channel):
    return int(_capi.lib().d2p_frames_for_bytes(C.byref(cfg), int(bytes_per_channel)))


def total_frames(cfg, payload_bytes):
    """Frames per channel a whole container-order payload of `payload_bytes` yields."""
    return int(_capi.lib().d2p_total_frames(C.byref(cfg), int(payload_bytes)))


def block_range(cfg, payload_bytes, first_frame, frame_count):
    """(in_offset, in_bytes): payload slice that frames [first_frame, first_frame+frame_count) read (halo included)."""
    off, n = C.c_uint64(), C.c_uint64()
    _capi.check(_capi.lib().d2p_block_range(C.byref(cfg), int(payload_bytes), int(first_frame), int(frame_count),
                                            C.byref(off), C.byref(n)))
    return int(off.value), int(n.value)


def tap_table_names():
    L = _capi.lib()
    return [L.d2p_tap_table_name(i).decode() for i in range(L.d2p_tap_table_count())]


def tap_table(name):
    p = C.POINTER(C.c_double)()
    n = _capi.lib().d2p_tap_table(name.encode(), C.byref(p))
    if n < 0:
        raise KeyError(name)
    return np.ctypeslib.as_array(p, shape=(n,)).copy()


def _sample_bytes(cfg):
    if cfg.out_format == _capi.OUT_PLANAR_32:
        return 4
    return 3 if cfg.bits == 20 else cfg.bits // 8


class Converter:
    """One input file: d2p_create / d2p_push / d2p_finish / d2p_pull (the process_blocks seam)."""

    def __init__(self, cfg):
        self.cfg = cfg
        self._h = C.c_void_p()
        _capi.check(_capi.lib().d2p_create(C.byref(cfg), C.byref(self._h)))

    def close(self):
        if self._h:
            _capi.lib().d2p_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def push(self, raw):
        raw = np.ascontiguousarray(raw, dtype=np.uint8)
        _capi.check(_capi.lib().d2p_push(self._h, raw.ctypes.data, raw.size))

    def flush(self):
        _capi.check(_capi.lib().d2p_flush(self._h))

    def set_push_threshold(self, nbytes):
        _capi.check(_capi.lib().d2p_set_push_threshold(self._h, nbytes))

    def finish(self):
        _capi.check(_capi.lib().d2p_finish(self._h))

    def frames_ready(self):
        return int(_capi.lib().d2p_frames_ready(self._h))

    def pull(self, max_frames=None):
        """-> uint8 array of packed interleaved PCM, or [channels, frames] int32/float32 for planar output."""
        n = self.frames_ready() if max_frames is None else min(max_frames, self.frames_ready())
        sb = _sample_bytes(self.cfg)
        buf = np.zeros(n * self.cfg.channels * sb, dtype=np.uint8)
        got = C.c_uint64()
        _capi.check(_capi.lib().d2p_pull(self._h, buf.ctypes.data, buf.size, C.byref(got)))
        assert got.value == n
        if self.cfg.out_format == _capi.OUT_PLANAR_32:
            dt = np.float32 if self.cfg.bits == 32 else np.int32
            return buf.view(dt).reshape(self.cfg.channels, n)
        return buf

    def peek_f64(self, chan):
        n = self.frames_ready()
        out = np.zeros(n, dtype=np.float64)
        got = C.c_uint64()
        _capi.check(_capi.lib().d2p_pull_f64(self._h, chan, out.ctypes.data_as(C.POINTER(C.c_double)), n, C.byref(got)))
        return out[:got.value]

    def stats(self):
        clips, peak, clipped = C.c_int64(), C.c_double(), C.c_int64()
        _capi.check(_capi.lib().d2p_stats(self._h, C.byref(clips), C.byref(peak), C.byref(clipped)))
        return {"clips": clips.value, "peak_abs": peak.value, "clipped_samples": clipped.value}


class HostBuffer:
    """Pinned host memory from d2p_host_alloc, viewed as a uint8 numpy array."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        self.ptr = _capi.lib().d2p_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise MemoryError("d2p_host_alloc failed")
        self.array = np.ctypeslib.as_array((C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr))[: self.nbytes]

    def free(self):
        if self.ptr:
            self.array = None
            _capi.lib().d2p_host_free(self.ptr)
            self.ptr = None

    __del__ = free


def convert_host(cfg, raw, out=None):
    """Whole payload in host memory -> packed interleaved PCM in host memory (d2p_convert_host).
    `raw`/`out` may be numpy arrays or HostBuffer.array views.  Returns (pcm_bytes_view, frames, stats)."""
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    per_ch = raw.size // cfg.channels
    sb = _sample_bytes(cfg)
    need = frames_for_bytes(cfg, per_ch) * cfg.channels * sb
    if out is None:
        out = np.zeros(need, dtype=np.uint8)
    frames, clips, peak = C.c_uint64(), C.c_int64(), C.c_double()
    _capi.check(_capi.lib().d2p_convert_host(C.byref(cfg), raw.ctypes.data, raw.size, out.ctypes.data, out.size,
                                             C.byref(frames), C.byref(clips), C.byref(peak)))
    return out[: frames.value * cfg.channels * sb], frames.value, {"clipped_samples": clips.value, "peak_abs": peak.value}


def check_level(cfg, raw):
    """Peak level of a whole payload (Rdsd2Pcm::new_level_check + check_level, rd/lib.rs:124-182):
    -> (peak_abs, peak_dbfs as float32)."""
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    peak, db = C.c_double(), C.c_float()
    _capi.check(_capi.lib().d2p_check_level(C.byref(cfg), raw.ctypes.data, raw.size, C.byref(peak), C.byref(db)))
    return peak.value, np.float32(db.value)


class Batch:
    """Many payloads already resident in HBM, converted by one launch set (d2p_batch_*)."""

    def __init__(self, cfg):
        self.cfg = cfg
        self._h = C.c_void_p()
        _capi.check(_capi.lib().d2p_batch_create(C.byref(cfg), C.byref(self._h)))

    def close(self):
        if self._h:
            _capi.lib().d2p_batch_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def make_jobs(self, items):
        """items: iterable of dicts(d_in, in_bytes, d_out, out_cap_bytes[, d_f64, f64_stride, dither_seed])."""
        items = list(items)
        arr = (Job * len(items))()
        for j, it in zip(arr, items):
            j.d_in, j.in_bytes = it["d_in"], it["in_bytes"]
            j.d_out, j.out_cap_bytes = it["d_out"], it["out_cap_bytes"]
            j.d_f64, j.f64_stride = it.get("d_f64", None), it.get("f64_stride", 0)
            j.dither_seed = it.get("dither_seed", 0)
        return arr

    def run(self, jobs, stream=0):
        _capi.check(_capi.lib().d2p_batch_run(self._h, jobs, len(jobs), C.c_void_p(stream)))
        return [int(j.frames) for j in jobs]

    def run_blocks(self, items, stream=0):
        """items: dicts(d_in, in_offset, in_bytes, payload_bytes, first_frame, frame_count, d_out, out_cap_bytes
        [, d_f64, f64_stride, dither_seed]) -- time blocks of long payloads (d2p_batch_run_blocks)."""
        items = list(items)
        arr = (_capi.BlockJob * len(items))()
        for j, it in zip(arr, items):
            for k in ("d_in", "in_offset", "in_bytes", "payload_bytes", "first_frame", "frame_count", "d_out", "out_cap_bytes"):
                setattr(j, k, it[k])
            j.d_f64, j.f64_stride = it.get("d_f64", None), it.get("f64_stride", 0)
            j.dither_seed = it.get("dither_seed", 0)
        _capi.check(_capi.lib().d2p_batch_run_blocks(self._h, arr, len(items), C.c_void_p(stream)))

    def convert_host(self, raw, out):
        """raw/out: uint8 numpy views of host memory (ideally HostBuffer.array).  Returns frames."""
        frames = C.c_uint64()
        _capi.check(_capi.lib().d2p_batch_convert_host(self._h, raw.ctypes.data, raw.size, out.ctypes.data, out.size,
                                                       C.byref(frames)))
        return int(frames.value)

    def job_stats(self, njobs, stream=0):
        """Per job of the most recent run: (peak_abs list, clipped_samples list) -- dsd_levels' one peak per file."""
        clipped = (C.c_int64 * max(njobs, 1))()
        peak = (C.c_double * max(njobs, 1))()
        _capi.check(_capi.lib().d2p_batch_job_stats(self._h, C.c_void_p(stream), njobs, clipped, peak))
        return [float(x) for x in peak[:njobs]], [int(x) for x in clipped[:njobs]]

    def stats_reset(self, stream=0):
        _capi.check(_capi.lib().d2p_batch_stats_reset(self._h, C.c_void_p(stream)))

    def set_control(self, cancel_flag=None, progress=None):
        """cancel_flag: ctypes.c_int32 the caller may raise at any time; progress(frames_done, frames_total)."""
        self._cancel = cancel_flag
        self._progress = _capi.PROGRESS_FN(lambda _u, done, total: progress(done, total)) if progress else _capi.PROGRESS_FN()
        _capi.check(_capi.lib().d2p_batch_set_control(self._h, C.byref(cancel_flag) if cancel_flag is not None else None,
                                                      self._progress, None))

    def launches(self):
        return int(_capi.lib().d2p_batch_launches(self._h))

    def stats(self, stream=0):
        clipped, peak = C.c_int64(), C.c_double()
        _capi.check(_capi.lib().d2p_batch_stats(self._h, C.c_void_p(stream), C.byref(clipped), C.byref(peak)))
        return {"clipped_samples": clipped.value, "peak_abs": peak.value}


def run_noalloc(cfg, items, stream=0, want_stats=True):
    """d2p_run_noalloc: Batch.run without a Batch, for hosts where the library must not allocate device memory (its job
    table, LUT and statistics sit in a 256 KiB array inside the library).  items as for Batch.make_jobs; at most
    _capi.NOALLOC_MAX_JOBS of them; no dither F.  Returns (frames per job, stats dict of this call or None).
    Asking for the stats waits for the stream."""
    items = list(items)
    arr = (Job * max(len(items), 1))()
    for j, it in zip(arr, items):
        j.d_in, j.in_bytes = it["d_in"], it["in_bytes"]
        j.d_out, j.out_cap_bytes = it["d_out"], it["out_cap_bytes"]
        j.d_f64, j.f64_stride = it.get("d_f64", None), it.get("f64_stride", 0)
        j.dither_seed = it.get("dither_seed", 0)
    clipped, peak = C.c_int64(), C.c_double()
    _capi.check(_capi.lib().d2p_run_noalloc(C.byref(cfg), arr, len(items), C.c_void_p(stream),
                                            C.byref(clipped) if want_stats else None, C.byref(peak) if want_stats else None))
    stats = {"clipped_samples": clipped.value, "peak_abs": peak.value} if want_stats else None
    return [int(j.frames) for j in arr[:len(items)]], stats


# ---- container front / back end (include/dsd2pcm_b200_files.h) ------------------------------------------
def _info_dict(ci):
    return {k: getattr(ci, k) for k, _ in _capi.ContainerInfo._fields_}


def container_parse(data, file_bytes=None):
    """Headers of a DSF / DSDIFF file held in memory -> dict (d2p_container_parse)."""
    data = np.ascontiguousarray(np.frombuffer(bytes(data), dtype=np.uint8))
    ci = _capi.ContainerInfo()
    _capi.check(_capi.lib().d2p_container_parse(data.ctypes.data, data.size, data.size if file_bytes is None else file_bytes,
                                                C.byref(ci)))
    return _info_dict(ci)


def container_open(path):
    ci = _capi.ContainerInfo()
    _capi.check(_capi.lib().d2p_container_open(str(path).encode(), C.byref(ci)))
    return _info_dict(ci)


def pcm_header(container, bits, channels, rate, frames):
    """Header bytes of a WAV ('W') / AIFF ('A') / AIFC ('C') file (d2p_pcm_header); b'' if unsupported."""
    n = _capi.lib().d2p_pcm_header(ord(container), bits, channels, rate, frames, None, 0)
    buf = np.zeros(max(n, 1), dtype=np.uint8)
    _capi.lib().d2p_pcm_header(ord(container), bits, channels, rate, frames, buf.ctypes.data, buf.size)
    return buf[:n].tobytes()


def write_pcm_file(path, container, bits, channels, rate, payload, frames):
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    _capi.check(_capi.lib().d2p_write_pcm_file(str(path).encode(), ord(container), bits, channels, rate,
                                               payload.ctypes.data, frames))


def write_flac_file(path, bits, channels, rate, payload, frames, comments=()):
    """FLAC file from packed little-endian interleaved PCM as the kernels emit it (d2p_write_flac_file;
    the reference's save_flac_file, rd/audio_file.rs:454-573).  `comments`: "KEY=value" strings."""
    payload = np.ascontiguousarray(payload, dtype=np.uint8)
    arr = (C.c_char_p * max(len(comments), 1))(*[c.encode() for c in comments])
    _capi.check(_capi.lib().d2p_write_flac_file(str(path).encode(), bits, channels, rate, payload.ctypes.data, frames,
                                                arr if comments else None, len(comments)))


def convert_file(in_path, out_path, out_rate=352800, filt="E", bits=24, dither="X", level_db=0.0, container="W",
                 dither_seed=0, device=0, dsf_stop_at_sample_count=False, raw_channels=2, raw_layout="P",
                 raw_lsb_first=True, raw_dsd_rate=1, raw_block_size=4096, cancel_flag=None, progress=None,
                 copy_tags=False, append_rate=False):
    """One DSF / DSDIFF / raw DSD file -> one PCM file on the GPU (d2p_convert_file[_ex]).  Returns a dict.
    cancel_flag: ctypes.c_int32 that may be raised while the call runs; progress(frames_done, frames_total)."""
    opt = _capi.FileOptions(out_rate, ord(filt), bits, ord(dither), level_db, ord(container), dither_seed, device,
                            1 if dsf_stop_at_sample_count else 0, raw_channels, ord(raw_layout),
                            1 if raw_lsb_first else 0, raw_dsd_rate, raw_block_size, 1 if copy_tags else 0,
                            1 if append_rate else 0)
    res = _capi.FileResult()
    if cancel_flag is None and progress is None:
        _capi.check(_capi.lib().d2p_convert_file(C.byref(opt), str(in_path).encode(), str(out_path).encode(), C.byref(res)))
    else:
        cb = _capi.PROGRESS_FN(lambda _u, done, total: progress(done, total)) if progress else _capi.PROGRESS_FN()
        _capi.check(_capi.lib().d2p_convert_file_ex(C.byref(opt), str(in_path).encode(), str(out_path).encode(), C.byref(res),
                                                    C.byref(cancel_flag) if cancel_flag is not None else None, cb, None))
    return {"in": _info_dict(res.info), "frames": int(res.frames), "out_bytes": int(res.out_bytes),
            "clipped_samples": int(res.clipped_samples), "peak_abs": float(res.peak_abs), "dither_seed": int(res.dither_seed),
            "tag_bytes": int(res.tag_bytes)}


def id3_to_vorbis(blob):
    """ID3v2.3 / 2.4 tag bytes -> list of "KEY=value" Vorbis comments (d2p_id3_to_vorbis); None if the tag does not parse."""
    blob = bytes(blob)
    need = C.c_size_t()
    n = _capi.lib().d2p_id3_to_vorbis(blob, len(blob), None, 0, C.byref(need))
    if n < 0:
        return None
    buf = C.create_string_buffer(need.value)
    _capi.lib().d2p_id3_to_vorbis(blob, len(blob), buf, need.value, C.byref(need))
    return [x.decode("utf-8") for x in buf.raw[:need.value].split(b"\0") if x]
