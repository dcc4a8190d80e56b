"""ctypes binding of libdsd2pcm_b200.so -- the C ABI declared in include/dsd2pcm_b200.h.

The library is the product: CUDA kernels for sm_100a behind plain-C entry points.  This module only
loads it and declares signatures; there is no Python or CPU implementation of the conversion, and
nothing here touches oracle/.  Import fails loudly if the library has not been built
(`python -c "import __graft_entry__ as g; g.build()"` or `make -C dsd2dxd_b200/csrc`).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdsd2pcm_b200.so")

OK = 0
ERR_INVALID_ARG, ERR_TAPS_NOT_FOUND, ERR_UNSUPPORTED_LM, ERR_BIT_DEPTH = 1, 2, 3, 4
ERR_NO_DEVICE, ERR_CUDA, ERR_NOMEM, ERR_STATE, ERR_CAPACITY, ERR_CANCELLED = 5, 6, 7, 8, 9, 10
PROGRESS_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_uint64)
OUT_INTERLEAVED_LE, OUT_INTERLEAVED_BE, OUT_PLANAR_32, OUT_NONE = 0, 1, 2, 3
NOALLOC_SCRATCH_BYTES, NOALLOC_MAX_JOBS = 256 * 1024, 64


class Config(C.Structure):
    _fields_ = [
        ("dsd_rate", C.c_int32), ("out_rate", C.c_uint32), ("filter", C.c_int32), ("bits", C.c_int32),
        ("dither", C.c_int32), ("channels", C.c_int32), ("level_db", C.c_double), ("in_layout", C.c_int32),
        ("lsb_first", C.c_int32), ("block_size", C.c_uint32), ("out_format", C.c_int32),
        ("dither_seed", C.c_uint64), ("has_dither_scale", C.c_int32), ("dither_scale_db", C.c_double),
        ("device", C.c_int32), ("keep_f64", C.c_int32),
    ]


class PlanInfo(C.Structure):
    _fields_ = [
        ("decim_ratio", C.c_int32), ("upsample_ratio", C.c_uint32), ("num_stages", C.c_int32),
        ("stage_table", (C.c_char * 32) * 3), ("stage_half_taps", C.c_int32 * 3), ("stage_decim", C.c_int32 * 3),
        ("scale_factor", C.c_double), ("peak_level", C.c_int32), ("bytes_per_sample", C.c_int32),
        ("frames_cap", C.c_uint64), ("lut_bytes", C.c_uint64),
    ]


class Job(C.Structure):
    _fields_ = [
        ("d_in", C.c_void_p), ("in_bytes", C.c_uint64), ("d_out", C.c_void_p), ("out_cap_bytes", C.c_uint64),
        ("d_f64", C.c_void_p), ("f64_stride", C.c_uint64), ("dither_seed", C.c_uint64), ("frames", C.c_uint64),
    ]


class BlockJob(C.Structure):
    _fields_ = [
        ("d_in", C.c_void_p), ("in_offset", C.c_uint64), ("in_bytes", C.c_uint64), ("payload_bytes", C.c_uint64),
        ("first_frame", C.c_uint64), ("frame_count", C.c_uint64), ("d_out", C.c_void_p), ("out_cap_bytes", C.c_uint64),
        ("d_f64", C.c_void_p), ("f64_stride", C.c_uint64), ("dither_seed", C.c_uint64),
    ]


class ContainerInfo(C.Structure):
    _fields_ = [
        ("format", C.c_int32), ("channels", C.c_int32), ("dsd_rate", C.c_int32), ("sample_rate", C.c_uint32),
        ("in_layout", C.c_int32), ("lsb_first", C.c_int32), ("block_size", C.c_uint32),
        ("data_offset", C.c_uint64), ("data_bytes", C.c_uint64), ("sample_count", C.c_uint64),
        ("id3_offset", C.c_uint64), ("id3_bytes", C.c_uint64),
    ]


class FileOptions(C.Structure):
    _fields_ = [
        ("out_rate", C.c_uint32), ("filter", C.c_int32), ("bits", C.c_int32), ("dither", C.c_int32),
        ("level_db", C.c_double), ("out_container", C.c_int32), ("dither_seed", C.c_uint64), ("device", C.c_int32),
        ("dsf_stop_at_sample_count", C.c_int32),
        ("raw_channels", C.c_int32), ("raw_layout", C.c_int32), ("raw_lsb_first", C.c_int32), ("raw_dsd_rate", C.c_int32),
        ("raw_block_size", C.c_uint32), ("copy_tags", C.c_int32), ("append_rate", C.c_int32),
    ]


class FileResult(C.Structure):
    _fields_ = [("info", ContainerInfo), ("frames", C.c_uint64), ("out_bytes", C.c_uint64),
                ("clipped_samples", C.c_int64), ("peak_abs", C.c_double), ("dither_seed", C.c_uint64),
                ("tag_bytes", C.c_uint64)]


class FlacPicture(C.Structure):
    _fields_ = [("type", C.c_int32), ("mime", C.c_char_p), ("description", C.c_char_p), ("width", C.c_uint32),
                ("height", C.c_uint32), ("depth", C.c_uint32), ("colors", C.c_uint32), ("data", C.c_void_p),
                ("data_bytes", C.c_uint64)]


# every symbol include/dsd2pcm_b200.h and include/dsd2pcm_b200_files.h declare
EXPORTS = [
    "d2p_abi_version", "d2p_strerror", "d2p_last_error", "d2p_device_count",
    "d2p_plan", "d2p_frames_for_bytes", "d2p_tap_table", "d2p_tap_table_count", "d2p_tap_table_name",
    "d2p_create", "d2p_destroy", "d2p_push", "d2p_flush", "d2p_set_push_threshold", "d2p_finish", "d2p_frames_ready", "d2p_pull", "d2p_pull_f64", "d2p_stats",
    "d2p_convert_host", "d2p_check_level", "d2p_host_alloc", "d2p_host_free",
    "d2p_batch_create", "d2p_batch_destroy", "d2p_batch_run", "d2p_batch_convert_host", "d2p_batch_launches",
    "d2p_batch_stats", "d2p_batch_stats_reset", "d2p_batch_job_stats", "d2p_batch_set_control", "d2p_total_frames", "d2p_block_range", "d2p_batch_run_blocks", "d2p_run_noalloc", "d2p_debug_fpd_exponent",
    "d2p_container_parse", "d2p_container_open", "d2p_pcm_header", "d2p_write_pcm_file", "d2p_write_flac_file", "d2p_write_flac_file_ex", "d2p_id3_to_vorbis", "d2p_convert_file", "d2p_convert_file_ex",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build the CUDA library first (there is no fallback path)")
    L = C.CDLL(LIB_PATH)
    cfgp, vp, u8p = C.POINTER(Config), C.c_void_p, C.POINTER(C.c_uint8)
    L.d2p_abi_version.restype = C.c_int
    L.d2p_strerror.restype = C.c_char_p
    L.d2p_strerror.argtypes = [C.c_int]
    L.d2p_last_error.restype = C.c_char_p
    L.d2p_device_count.restype = C.c_int
    L.d2p_plan.argtypes = [cfgp, C.POINTER(PlanInfo)]
    L.d2p_frames_for_bytes.restype = C.c_uint64
    L.d2p_frames_for_bytes.argtypes = [cfgp, C.c_uint64]
    L.d2p_tap_table.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_double))]
    L.d2p_tap_table_count.restype = C.c_int
    L.d2p_tap_table_name.restype = C.c_char_p
    L.d2p_tap_table_name.argtypes = [C.c_int]
    L.d2p_create.argtypes = [cfgp, C.POINTER(vp)]
    L.d2p_destroy.argtypes = [vp]
    L.d2p_destroy.restype = None
    L.d2p_push.argtypes = [vp, vp, C.c_size_t]
    L.d2p_finish.argtypes = [vp]
    L.d2p_flush.argtypes = [vp]
    L.d2p_set_push_threshold.argtypes = [vp, C.c_size_t]
    L.d2p_frames_ready.restype = C.c_uint64
    L.d2p_frames_ready.argtypes = [vp]
    L.d2p_pull.argtypes = [vp, vp, C.c_size_t, C.POINTER(C.c_uint64)]
    L.d2p_pull_f64.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.c_size_t, C.POINTER(C.c_uint64)]
    L.d2p_stats.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.d2p_convert_host.argtypes = [cfgp, vp, C.c_size_t, vp, C.c_size_t, C.POINTER(C.c_uint64),
                                   C.POINTER(C.c_int64), C.POINTER(C.c_double)]
    L.d2p_check_level.argtypes = [cfgp, vp, C.c_size_t, C.POINTER(C.c_double), C.POINTER(C.c_float)]
    L.d2p_host_alloc.restype = vp
    L.d2p_host_alloc.argtypes = [C.c_size_t]
    L.d2p_host_free.argtypes = [vp]
    L.d2p_host_free.restype = None
    L.d2p_batch_create.argtypes = [cfgp, C.POINTER(vp)]
    L.d2p_batch_destroy.argtypes = [vp]
    L.d2p_batch_destroy.restype = None
    L.d2p_batch_run.argtypes = [vp, C.POINTER(Job), C.c_int, vp]
    L.d2p_batch_convert_host.argtypes = [vp, vp, C.c_size_t, vp, C.c_size_t, C.POINTER(C.c_uint64)]
    L.d2p_batch_launches.restype = C.c_uint64
    L.d2p_batch_launches.argtypes = [vp]
    L.d2p_batch_stats.argtypes = [vp, vp, C.POINTER(C.c_int64), C.POINTER(C.c_double)]
    L.d2p_batch_stats_reset.argtypes = [vp, vp]
    L.d2p_debug_fpd_exponent.argtypes = [C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_int32), C.c_int]
    L.d2p_batch_job_stats.argtypes = [vp, vp, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_double)]
    L.d2p_batch_set_control.argtypes = [vp, C.POINTER(C.c_int32), PROGRESS_FN, vp]
    L.d2p_total_frames.restype = C.c_uint64
    L.d2p_total_frames.argtypes = [cfgp, C.c_uint64]
    L.d2p_block_range.argtypes = [cfgp, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.d2p_batch_run_blocks.argtypes = [vp, C.POINTER(BlockJob), C.c_int, vp]
    L.d2p_run_noalloc.argtypes = [cfgp, C.POINTER(Job), C.c_int, vp, C.POINTER(C.c_int64), C.POINTER(C.c_double)]
    cip = C.POINTER(ContainerInfo)
    L.d2p_container_parse.argtypes = [vp, C.c_size_t, C.c_uint64, cip]
    L.d2p_container_open.argtypes = [C.c_char_p, cip]
    L.d2p_pcm_header.restype = C.c_size_t
    L.d2p_pcm_header.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint64, vp, C.c_size_t]
    L.d2p_write_pcm_file.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_uint32, vp, C.c_uint64]
    L.d2p_write_flac_file.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_uint32, vp, C.c_uint64, C.POINTER(C.c_char_p), C.c_int]
    L.d2p_write_flac_file_ex.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_uint32, vp, C.c_uint64, C.POINTER(C.c_char_p), C.c_int,
                                         C.POINTER(FlacPicture), C.c_int]
    L.d2p_id3_to_vorbis.argtypes = [vp, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.d2p_convert_file.argtypes = [C.POINTER(FileOptions), C.c_char_p, C.c_char_p, C.POINTER(FileResult)]
    L.d2p_convert_file_ex.argtypes = [C.POINTER(FileOptions), C.c_char_p, C.c_char_p, C.POINTER(FileResult),
                                      C.POINTER(C.c_int32), PROGRESS_FN, vp]
    _lib = L
    return L


class D2PError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(message)
        self.status = status


def check(status):
    if status != OK:
        L = lib()
        detail = L.d2p_last_error().decode() or L.d2p_strerror(status).decode()
        raise D2PError(status, detail)
