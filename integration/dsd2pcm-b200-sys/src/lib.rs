//! Raw bindings, one item per declaration of `include/dsd2pcm_b200.h` and `include/dsd2pcm_b200_files.h`
//! (tests/test_capi_surface.py checks that the two stay in step).  Not compiled in this repository's CI: the
//! build image has no Rust toolchain.  The comments name the rdsd2pcm item each entry point replaces
//! (rd/ = src/rdsd2pcm/src/ of the reference).
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

pub const D2P_ABI_VERSION: c_int = 2;

// d2p_status
pub const D2P_OK: c_int = 0;
pub const D2P_ERR_INVALID_ARG: c_int = 1;
pub const D2P_ERR_TAPS_NOT_FOUND: c_int = 2; // rd/conversion_context.rs:155-163
pub const D2P_ERR_UNSUPPORTED_LM: c_int = 3; // rd/lm_resampler.rs:115,264
pub const D2P_ERR_BIT_DEPTH: c_int = 4; // rd/pcm_writer.rs:195-197
pub const D2P_ERR_NO_DEVICE: c_int = 5;
pub const D2P_ERR_CUDA: c_int = 6;
pub const D2P_ERR_NOMEM: c_int = 7;
pub const D2P_ERR_STATE: c_int = 8;
pub const D2P_ERR_CAPACITY: c_int = 9;
pub const D2P_ERR_CANCELLED: c_int = 10;

// FilterType (rd/lib.rs:313-323), DitherType (rd/lib.rs:301-311), FmtType
pub const D2P_FILTER_EQUIRIPPLE: i32 = b'E' as i32;
pub const D2P_FILTER_XLD: i32 = b'X' as i32;
pub const D2P_FILTER_DSD2PCM: i32 = b'D' as i32;
pub const D2P_FILTER_CHEBYSHEV: i32 = b'C' as i32;
pub const D2P_DITHER_TPDF: i32 = b'T' as i32;
pub const D2P_DITHER_RECT: i32 = b'R' as i32;
pub const D2P_DITHER_FPD: i32 = b'F' as i32;
pub const D2P_DITHER_NONE: i32 = b'X' as i32;
pub const D2P_LAYOUT_PLANAR: i32 = b'P' as i32;
pub const D2P_LAYOUT_INTERLEAVED: i32 = b'I' as i32;
pub const D2P_OUT_INTERLEAVED_LE: i32 = 0; // pack_int / pack_float, rd/pcm_writer.rs:365-396
pub const D2P_OUT_INTERLEAVED_BE: i32 = 1; // rd/audio_file.rs:257-275
pub const D2P_OUT_PLANAR_32: i32 = 2; // AudioFile.samples[channel][frame], rd/pcm_writer.rs:411-421
pub const D2P_OUT_NONE: i32 = 3; // PcmWriter::new_sink, rd/pcm_writer.rs:151-182
pub const D2P_CONTAINER_DSF: i32 = b'S' as i32;
pub const D2P_CONTAINER_DFF: i32 = b'F' as i32;
pub const D2P_CONTAINER_RAW: i32 = b'R' as i32;
pub const D2P_FILE_WAV: i32 = b'W' as i32;
pub const D2P_FILE_AIFF: i32 = b'A' as i32;
pub const D2P_FILE_AIFC: i32 = b'C' as i32;
pub const D2P_FILE_RAW: i32 = b'S' as i32;
pub const D2P_FILE_FLAC: i32 = b'F' as i32;

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_config {
    pub dsd_rate: i32,
    pub out_rate: u32,
    pub filter: i32,
    pub bits: i32,
    pub dither: i32,
    pub channels: i32,
    pub level_db: f64,
    pub in_layout: i32,
    pub lsb_first: i32,
    pub block_size: u32,
    pub out_format: i32,
    pub dither_seed: u64,
    pub has_dither_scale: i32,
    pub dither_scale_db: f64,
    pub device: i32,
    pub keep_f64: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_plan_info {
    pub decim_ratio: i32,
    pub upsample_ratio: u32,
    pub num_stages: i32,
    pub stage_table: [[c_char; 32]; 3],
    pub stage_half_taps: [i32; 3],
    pub stage_decim: [i32; 3],
    pub scale_factor: f64,
    pub peak_level: i32,
    pub bytes_per_sample: i32,
    pub frames_cap: u64,
    pub lut_bytes: u64,
}

#[repr(C)]
pub struct d2p_ctx {
    _private: [u8; 0],
}
#[repr(C)]
pub struct d2p_batch {
    _private: [u8; 0],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_job {
    pub d_in: *const c_void,
    pub in_bytes: u64,
    pub d_out: *mut c_void,
    pub out_cap_bytes: u64,
    pub d_f64: *mut f64,
    pub f64_stride: u64,
    pub dither_seed: u64,
    pub frames: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_block_job {
    pub d_in: *const c_void,
    pub in_offset: u64,
    pub in_bytes: u64,
    pub payload_bytes: u64,
    pub first_frame: u64,
    pub frame_count: u64,
    pub d_out: *mut c_void,
    pub out_cap_bytes: u64,
    pub d_f64: *mut f64,
    pub f64_stride: u64,
    pub dither_seed: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_container_info {
    pub format: i32,
    pub channels: i32,
    pub dsd_rate: i32,
    pub sample_rate: u32,
    pub in_layout: i32,
    pub lsb_first: i32,
    pub block_size: u32,
    pub data_offset: u64,
    pub data_bytes: u64,
    pub sample_count: u64,
    pub id3_offset: u64,
    pub id3_bytes: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_file_options {
    pub out_rate: u32,
    pub filter: i32,
    pub bits: i32,
    pub dither: i32,
    pub level_db: f64,
    pub out_container: i32,
    pub dither_seed: u64,
    pub device: i32,
    pub dsf_stop_at_sample_count: i32,
    pub raw_channels: i32,
    pub raw_layout: i32,
    pub raw_lsb_first: i32,
    pub raw_dsd_rate: i32,
    pub raw_block_size: u32,
    pub copy_tags: i32,
    pub append_rate: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_file_result {
    pub input: d2p_container_info, // `in` in the C header
    pub frames: u64,
    pub out_bytes: u64,
    pub clipped_samples: i64,
    pub peak_abs: f64,
    pub dither_seed: u64,
    pub tag_bytes: u64,
}

/// One FLAC PICTURE block (front cover = 3, back cover = 4)   rd/pcm_writer.rs:583-603
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct d2p_flac_picture {
    pub r#type: i32,
    pub mime: *const c_char,
    pub description: *const c_char,
    pub width: u32,
    pub height: u32,
    pub depth: u32,
    pub colors: u32,
    pub data: *const c_void,
    pub data_bytes: u64,
}

/// `progress` of d2p_batch_set_control / d2p_convert_file_ex: (user, frames_done, frames_total)
pub type d2p_progress_fn = Option<unsafe extern "C" fn(user: *mut c_void, frames_done: u64, frames_total: u64)>;

extern "C" {
    pub fn d2p_abi_version() -> c_int;
    pub fn d2p_strerror(status: c_int) -> *const c_char;
    pub fn d2p_last_error() -> *const c_char;
    pub fn d2p_device_count() -> c_int;

    // ConversionContext::new: filter / path selection, calc_frames_cap   rd/conversion_context.rs:41-195, 295-328
    pub fn d2p_plan(cfg: *const d2p_config, info: *mut d2p_plan_info) -> c_int;
    pub fn d2p_frames_for_bytes(cfg: *const d2p_config, bytes_per_channel: u64) -> u64;
    pub fn d2p_total_frames(cfg: *const d2p_config, payload_bytes: u64) -> u64;
    // rd/filters.rs, rd/filters_lm.rs
    pub fn d2p_tap_table(name: *const c_char, data: *mut *const f64) -> c_int;
    pub fn d2p_tap_table_count() -> c_int;
    pub fn d2p_tap_table_name(index: c_int) -> *const c_char;

    // streaming context = Dsd2Pcm / LMResampler per channel + PcmWriter   rd/conversion_context.rs:258-398
    pub fn d2p_create(cfg: *const d2p_config, out: *mut *mut d2p_ctx) -> c_int;
    pub fn d2p_destroy(ctx: *mut d2p_ctx);
    pub fn d2p_push(ctx: *mut d2p_ctx, raw: *const u8, nbytes: usize) -> c_int; // one read_next_block payload
    pub fn d2p_flush(ctx: *mut d2p_ctx) -> c_int;
    pub fn d2p_set_push_threshold(ctx: *mut d2p_ctx, nbytes: usize) -> c_int;
    pub fn d2p_finish(ctx: *mut d2p_ctx) -> c_int;
    pub fn d2p_frames_ready(ctx: *const d2p_ctx) -> u64;
    pub fn d2p_pull(ctx: *mut d2p_ctx, pcm: *mut c_void, cap_bytes: usize, frames: *mut u64) -> c_int;
    pub fn d2p_pull_f64(ctx: *mut d2p_ctx, chan: c_int, out: *mut f64, cap: usize, n: *mut u64) -> c_int;
    pub fn d2p_stats(ctx: *mut d2p_ctx, clips: *mut i64, peak_abs: *mut f64, clipped_samples: *mut i64) -> c_int;

    // whole payload in host memory: process_blocks over every block   rd/conversion_context.rs:258-293
    pub fn d2p_convert_host(cfg: *const d2p_config, raw: *const u8, nbytes: usize, pcm: *mut c_void, out_cap_bytes: usize,
                            frames: *mut u64, clips: *mut i64, peak_abs: *mut f64) -> c_int;
    // check_level / process_levels   rd/conversion_context.rs:196-256
    pub fn d2p_check_level(cfg: *const d2p_config, raw: *const u8, nbytes: usize, peak_abs: *mut f64, peak_dbfs: *mut f32) -> c_int;
    pub fn d2p_host_alloc(nbytes: usize) -> *mut c_void;
    pub fn d2p_host_free(p: *mut c_void);

    // many files / time blocks resident on the device: the per-file rayon loop of src/main.rs:283-301
    pub fn d2p_batch_create(cfg: *const d2p_config, out: *mut *mut d2p_batch) -> c_int;
    pub fn d2p_batch_destroy(b: *mut d2p_batch);
    pub fn d2p_batch_run(b: *mut d2p_batch, jobs: *mut d2p_job, njobs: c_int, cuda_stream: *mut c_void) -> c_int;
    // no counterpart in the reference: d2p_batch_run for hosts where this library may not allocate device memory
    // (scratch is a 256 KiB array inside the library; at most 64 jobs per call; calls that overlap on the GPU race)
    pub fn d2p_run_noalloc(cfg: *const d2p_config, jobs: *mut d2p_job, njobs: c_int, cuda_stream: *mut c_void,
                           clipped_samples: *mut i64, peak_abs: *mut f64) -> c_int;
    pub fn d2p_batch_convert_host(b: *mut d2p_batch, raw: *const u8, nbytes: usize, pcm: *mut c_void, out_cap_bytes: usize,
                                  frames: *mut u64) -> c_int;
    pub fn d2p_batch_launches(b: *const d2p_batch) -> u64;
    pub fn d2p_batch_stats(b: *mut d2p_batch, cuda_stream: *mut c_void, clipped_samples: *mut i64, peak_abs: *mut f64) -> c_int;
    pub fn d2p_batch_stats_reset(b: *mut d2p_batch, cuda_stream: *mut c_void) -> c_int;
    // one peak per file, as dsd_levels prints it (src/bin/dsd_levels/main.rs:207)
    pub fn d2p_batch_job_stats(b: *mut d2p_batch, cuda_stream: *mut c_void, njobs: c_int, clipped_samples: *mut i64,
                               peak_abs: *mut f64) -> c_int;
    // cancel_flag: &AtomicBool (as *const i32) and the progress channel of do_conversion   rd/lib.rs:275-281
    pub fn d2p_batch_set_control(b: *mut d2p_batch, cancel_flag: *const i32, progress: d2p_progress_fn, user: *mut c_void) -> c_int;
    pub fn d2p_block_range(cfg: *const d2p_config, payload_bytes: u64, first_frame: u64, frame_count: u64,
                           in_offset: *mut u64, in_bytes: *mut u64) -> c_int;
    pub fn d2p_batch_run_blocks(b: *mut d2p_batch, jobs: *const d2p_block_job, njobs: c_int, cuda_stream: *mut c_void) -> c_int;
    // diagnostic: Dither::fpdither's `log2().floor() as i32` as the kernels evaluate it   rd/dither.rs:102-104
    pub fn d2p_debug_fpd_exponent(samples: *const f64, n: c_int, exponents: *mut i32, device: c_int) -> c_int;

    // DsdReader::from_container (crate dsd-reader; call sites rd/lib.rs:92-99, 171-183)
    pub fn d2p_container_parse(bytes: *const u8, nbytes: usize, file_bytes: u64, info: *mut d2p_container_info) -> c_int;
    pub fn d2p_container_open(path: *const c_char, info: *mut d2p_container_info) -> c_int;
    // AudioFile::save_wave_file / save_aiff_file / save_aifc_file   rd/audio_file.rs:93-452, 576-621
    pub fn d2p_pcm_header(container: c_int, bits: c_int, channels: c_int, rate: u32, frames: u64, out: *mut u8, cap: usize) -> usize;
    pub fn d2p_write_pcm_file(path: *const c_char, container: c_int, bits: c_int, channels: c_int, rate: u32,
                              payload: *const c_void, frames: u64) -> c_int;
    // AudioFile::save_flac_file   rd/audio_file.rs:454-573
    pub fn d2p_write_flac_file(path: *const c_char, bits: c_int, channels: c_int, rate: u32, payload: *const c_void, frames: u64,
                               comments: *const *const c_char, n_comments: c_int) -> c_int;
    // Rdsd2Pcm::from_container + do_conversion + PcmWriter::save_file   rd/lib.rs:171-182, 275-281, rd/pcm_writer.rs:300-318
    pub fn d2p_write_flac_file_ex(path: *const c_char, bits: c_int, channels: c_int, rate: u32, payload: *const c_void, frames: u64,
                                  comments: *const *const c_char, n_comments: c_int, pictures: *const d2p_flac_picture,
                                  n_pictures: c_int) -> c_int;
    // PcmWriter::id3_to_flac_meta   rd/pcm_writer.rs:425-580
    pub fn d2p_id3_to_vorbis(id3: *const u8, nbytes: usize, out: *mut c_char, cap: usize, needed: *mut usize) -> c_int;
    pub fn d2p_convert_file(opt: *const d2p_file_options, in_path: *const c_char, out_path: *const c_char,
                            res: *mut d2p_file_result) -> c_int;
    pub fn d2p_convert_file_ex(opt: *const d2p_file_options, in_path: *const c_char, out_path: *const c_char,
                               res: *mut d2p_file_result, cancel_flag: *const i32, progress: d2p_progress_fn,
                               user: *mut c_void) -> c_int;
}
