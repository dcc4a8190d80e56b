/*
 * dsd2pcm_b200.h -- C ABI of libdsd2pcm_b200.so: the rdsd2pcm conversion core (DSD unpack ->
 * LUT-FIR decimate / L-M resample -> scale / dither / quantise / pack) on NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for clone206/dsd2dxd's `rdsd2pcm` library.  The reference has no
 * FFI of its own; the seam these entry points replace is the body of
 *     ConversionContext::process_blocks          rd/conversion_context.rs:295-328
 *         process_channel(chan, &chan_bufs[chan])   :376-398  (BytePrecalcDecimator::process_bytes
 *                                                   rd/byte_precalc_decimator.rs:112 |
 *                                                   LMResampler::process_bytes_lm rd/lm_resampler.rs:270)
 *         PcmWriter::write_to_buffer(n, chan)       rd/pcm_writer.rs:608
 * plus the constructor parameter set of Rdsd2Pcm::new / from_container (rd/lib.rs:75-91, 171-182).
 * (rd/ = src/rdsd2pcm/src/ of the reference.)  INTEGRATION.md shows the Rust `extern "C"` block
 * and the call-site patch.
 *
 * Plain C types only.  All functions return a d2p_status (0 = success) unless noted; none aborts.
 * There is no CPU fallback: without a CUDA device every compute entry returns D2P_ERR_NO_DEVICE.
 * A d2p_ctx is owned by one thread at a time (like one Rdsd2Pcm per Rayon worker).
 */
#ifndef DSD2PCM_B200_H
#define DSD2PCM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define D2P_ABI_VERSION 2

typedef enum d2p_status {
    D2P_OK = 0,
    D2P_ERR_INVALID_ARG = 1,
    D2P_ERR_TAPS_NOT_FOUND = 2,   /* "Taps not found for decimation ratio ..."  rd/conversion_context.rs:155-163 */
    D2P_ERR_UNSUPPORTED_LM = 3,   /* "Unsupported L/M combination" / "Unsupported 588: ..."  rd/lm_resampler.rs:115,264 */
    D2P_ERR_BIT_DEPTH = 4,        /* "Unsupported bit depth"  rd/pcm_writer.rs:195-197 */
    D2P_ERR_NO_DEVICE = 5,        /* no usable CUDA device: the library has no CPU path */
    D2P_ERR_CUDA = 6,             /* a CUDA call failed; d2p_last_error() has the detail */
    D2P_ERR_NOMEM = 7,
    D2P_ERR_STATE = 8,            /* call sequence error (push after finish, ...) */
    D2P_ERR_CAPACITY = 9,         /* output buffer too small */
    D2P_ERR_CANCELLED = 10        /* the caller's cancel flag was raised (cancel_flag, rd/conversion_context.rs:222, 304-308) */
} d2p_status;

/* Values are the reference CLI's own option letters (src/main.rs:171-215). */
enum { D2P_FILTER_EQUIRIPPLE = 'E', D2P_FILTER_XLD = 'X', D2P_FILTER_DSD2PCM = 'D', D2P_FILTER_CHEBYSHEV = 'C' };
enum { D2P_DITHER_TPDF = 'T', D2P_DITHER_RECT = 'R', D2P_DITHER_FPD = 'F', D2P_DITHER_NONE = 'X' };
enum { D2P_LAYOUT_PLANAR = 'P', D2P_LAYOUT_INTERLEAVED = 'I' };
/* Output sample container */
enum {
    D2P_OUT_INTERLEAVED_LE = 0,   /* stdout / WAV payload: pack_int / pack_float  rd/pcm_writer.rs:365-396 */
    D2P_OUT_INTERLEAVED_BE = 1,   /* AIFF / AIFC payload  rd/audio_file.rs:257-275 */
    D2P_OUT_PLANAR_32 = 2,        /* AudioFile<i32|f32>.samples[channel][frame]  rd/pcm_writer.rs:411-421 (20-bit NOT shifted) */
    D2P_OUT_NONE = 3              /* level check: no PCM is written, only max |sample*scale| is kept
                                     (PcmWriter::new_sink rd/pcm_writer.rs:151-182, process_levels rd/conversion_context.rs:209-256) */
};

/* Mirrors the parameters of Rdsd2Pcm::new (rd/lib.rs:75-91) that reach the hot path. */
typedef struct d2p_config {
    int32_t  dsd_rate;         /* DsdRate: 1 (DSD64), 2, 4, 8 (DSD512)                               */
    uint32_t out_rate;         /* output PCM rate in Hz (CLI -r)                                     */
    int32_t  filter;           /* D2P_FILTER_*  (CLI -t)                                             */
    int32_t  bits;             /* 16, 20, 24 (integer) or 32 (float)  (CLI -b)                       */
    int32_t  dither;           /* D2P_DITHER_*  (CLI -d)                                             */
    int32_t  channels;         /* CLI -c                                                             */
    double   level_db;         /* CLI -l                                                             */
    int32_t  in_layout;        /* D2P_LAYOUT_*: planar blocks of block_size per channel, or byte-interleaved (CLI -f) */
    int32_t  lsb_first;        /* Endianness::LsbFirst (CLI -e L) = 1; MSB-first sources are bit-reversed on load */
    uint32_t block_size;       /* bytes per channel per block (CLI -s; DSF: 4096).  Also the reader chunk that
                                  orders the FPD dither stream and the clip counter, as in process_blocks.  */
    int32_t  out_format;       /* D2P_OUT_*                                                          */
    uint64_t dither_seed;      /* the reference seeds from the OS (irreproducible); here it is explicit */
    int32_t  has_dither_scale; /* env DSD2DXD_DITHERSCALE present  (rd/dither.rs:39-49)              */
    double   dither_scale_db;
    int32_t  device;           /* CUDA device ordinal                                                */
    int32_t  keep_f64;         /* also keep the pre-scale f64 decimator output per channel (parity)  */
} d2p_config;

/* What a configuration resolves to (host-only; needs no GPU).  Mirrors
 * compute_decim_and_upsample (rd/lm_resampler.rs:911), select_precalc_taps
 * (rd/byte_precalc_decimator.rs:167), LMResampler::new (rd/lm_resampler.rs:87), set_scaling
 * (rd/pcm_writer.rs:288) and calc_frames_cap (rd/lib.rs:259). */
typedef struct d2p_plan_info {
    int32_t  decim_ratio;      /* M */
    uint32_t upsample_ratio;   /* L (1 = integer path) */
    int32_t  num_stages;       /* 1 (LUT FIR), 2 or 3 (L/M) */
    char     stage_table[3][32];
    int32_t  stage_half_taps[3];
    int32_t  stage_decim[3];   /* integer path: decim; L/M: m1, D2, D3 */
    double   scale_factor;
    int32_t  peak_level;
    int32_t  bytes_per_sample;
    uint64_t frames_cap;       /* calc_frames_cap(block_size, M, L) */
    uint64_t lut_bytes;        /* shared-memory LUT footprint of stage 1 */
} d2p_plan_info;

typedef struct d2p_ctx d2p_ctx;

/* ---- library ---------------------------------------------------------------------------------- */
int         d2p_abi_version(void);
const char* d2p_strerror(int status);          /* reference-style message for a status */
const char* d2p_last_error(void);              /* thread-local detail of the last failure */
int         d2p_device_count(void);            /* 0 when no CUDA device is usable */

/* ---- planning (host only) ---------------------------------------------------------------------- */
int      d2p_plan(const d2p_config* cfg, d2p_plan_info* info);
/* Frames per channel the reference emits for `bytes_per_channel` input bytes fed in block_size chunks. */
uint64_t d2p_frames_for_bytes(const d2p_config* cfg, uint64_t bytes_per_channel);
/* Half-tap table by reference name ("HTAPS_DSD64_8TO1_EQ", ...): returns length, *data -> static f64. */
int      d2p_tap_table(const char* name, const double** data);
int      d2p_tap_table_count(void);
const char* d2p_tap_table_name(int index);

/* ---- streaming conversion: one context per input file ------------------------------------------- */
int  d2p_create(const d2p_config* cfg, d2p_ctx** out);
void d2p_destroy(d2p_ctx* ctx);
/* Feed container-order DSD payload bytes (any chunking; host memory).  Replaces the per-block
 * process_channel + write_to_buffer pair: all frames that become computable are produced on the GPU
 * (see d2p_flush for when). */
int  d2p_push(d2p_ctx* ctx, const uint8_t* raw, size_t nbytes);
/* Pushes smaller than the push threshold (default 8 MiB; the reference hands over one 4096-byte block per channel at
 * a time) are gathered in pinned memory and shipped as ONE copy + ONE launch when the threshold is reached, at
 * d2p_flush() or at d2p_finish(): frames therefore become ready in bursts.  Threshold 0 ships every push at once. */
int  d2p_flush(d2p_ctx* ctx);
int  d2p_set_push_threshold(d2p_ctx* ctx, size_t nbytes);
/* End of input: a trailing partial block group is split evenly between channels. */
int  d2p_finish(d2p_ctx* ctx);
/* Frames produced and not yet pulled. */
uint64_t d2p_frames_ready(const d2p_ctx* ctx);
/* Copy up to cap_bytes of packed PCM (out_format layout) to host memory; *frames = frames copied.
 * For D2P_OUT_PLANAR_32, `pcm` receives channel-major int32/f32 with `*frames` per channel. */
int  d2p_pull(d2p_ctx* ctx, void* pcm, size_t cap_bytes, uint64_t* frames);
/* Pre-scale f64 output of one channel (needs keep_f64): frames [first, first+cap) counted from the
 * oldest frame not yet pulled by d2p_pull.  Does not consume. */
int  d2p_pull_f64(d2p_ctx* ctx, int chan, double* out, size_t cap, uint64_t* n);
/* clips: PcmWriter::clips semantics (rd/pcm_writer.rs:658-675); peak_abs: max |sample*scale|
 * (process_levels, rd/conversion_context.rs:209-256); clipped_samples: plain count.
 * `clips` replays the reference's run-length counter in its processing order (block, channel, frame), which is only
 * settled for whole reader blocks: call it after d2p_finish (before that, clip events of the unfinished last block may
 * still be ordered differently from the final stream; peak_abs and clipped_samples are exact at any time). */
int  d2p_stats(d2p_ctx* ctx, int64_t* clips, double* peak_abs, int64_t* clipped_samples);

/* ---- whole-payload conversion with host buffers (pipelined H2D / kernel / D2H) ------------------- */
/* out_cap_bytes must hold d2p_frames_for_bytes(...) * channels * bytes_per_sample.
 * clips receives the plain count of clipped samples (the run-length statistic needs the streaming context). */
int  d2p_convert_host(const d2p_config* cfg, const uint8_t* raw, size_t nbytes,
                      void* pcm, size_t out_cap_bytes, uint64_t* frames, int64_t* clips, double* peak_abs);
/* Peak level of a whole payload in host memory: Rdsd2Pcm::new_level_check + check_level
 * (rd/lib.rs:124-158, 171-182; process_levels rd/conversion_context.rs:209-256).  Like the reference it
 * ignores cfg's bits / dither / level / out_format / filter (32-bit sink, no dither, 0 dB, equiripple filters):
 * *peak_abs = max over all channels of |sample| * L,  *peak_dbfs = (f32)(20*log10(peak_abs)), -inf if silent.
 * Runs the same kernels with the writer reduced to the level reduction; nothing is copied back but 8 bytes. */
int  d2p_check_level(const d2p_config* cfg, const uint8_t* raw, size_t nbytes, double* peak_abs, float* peak_dbfs);
/* Pinned host memory for the buffers above (plain malloc'ed memory also works, more slowly). */
void* d2p_host_alloc(size_t nbytes);
void  d2p_host_free(void* p);

/* ---- device-resident batch: many payloads already in HBM, one launch set ------------------------- */
typedef struct d2p_job {
    const void* d_in;          /* device pointer: container-order payload */
    uint64_t    in_bytes;      /* payload bytes (all channels) */
    void*       d_out;         /* device pointer: packed PCM, out_format layout */
    uint64_t    out_cap_bytes;
    double*     d_f64;         /* optional device pointer: channels * f64_stride doubles (may be NULL) */
    uint64_t    f64_stride;
    uint64_t    dither_seed;   /* per-file seed */
    uint64_t    frames;        /* out: frames per channel written */
} d2p_job;

typedef struct d2p_batch d2p_batch;
int  d2p_batch_create(const d2p_config* cfg, d2p_batch** out);
void d2p_batch_destroy(d2p_batch* b);
/* Enqueue the conversion of all jobs on `cuda_stream` (a cudaStream_t; NULL = default stream).
 * Asynchronous: returns after the launches; jobs[i].frames is filled before returning. */
int  d2p_batch_run(d2p_batch* b, d2p_job* jobs, int njobs, void* cuda_stream);
/* Whole payload in host memory -> packed interleaved PCM in host memory through this batch object's
 * persistent pipeline (3 streams; H2D / kernel / D2H of neighbouring 64 MiB chunks overlap).
 * d2p_convert_host() is this call on a temporary batch object. */
int  d2p_batch_convert_host(d2p_batch* b, const uint8_t* raw, size_t nbytes, void* pcm, size_t out_cap_bytes,
                            uint64_t* frames);
/* Cancellation and progress of d2p_batch_convert_host (and of what is built on it: d2p_convert_file_ex).
 * The reference polls `cancel_flag: &AtomicBool` once per reader block and reports through track_io
 * (rd/conversion_context.rs:222, 304-308, 331-371; rd/lib.rs:275-281 do_conversion).  Here *cancel_flag is read and
 * `progress` is called between pipeline chunks (64 MiB of input), from the calling thread; frames_done counts frames
 * that have arrived in the caller's buffer.  A raised flag ends the call with D2P_ERR_CANCELLED.  NULLs switch it off. */
typedef void (*d2p_progress_fn)(void* user, uint64_t frames_done, uint64_t frames_total);
int  d2p_batch_set_control(d2p_batch* b, const volatile int32_t* cancel_flag, d2p_progress_fn progress, void* user);
/* Number of kernel launches issued by this batch object so far. */
uint64_t d2p_batch_launches(const d2p_batch* b);
/* Sum over jobs of clipped samples / max |sample*scale| since creation or the last reset (synchronises the stream). */
int  d2p_batch_stats(d2p_batch* b, void* cuda_stream, int64_t* clipped_samples, double* peak_abs);
int  d2p_batch_stats_reset(d2p_batch* b, void* cuda_stream);
/* Per job of the most recent d2p_batch_run / d2p_batch_run_blocks (dsd_levels reports one peak per file,
 * src/bin/dsd_levels/main.rs:207): arrays of njobs entries, either may be NULL.  Synchronises the stream.
 * d2p_batch_convert_host keeps no per-job table: after it, njobs > 0 is refused with D2P_ERR_INVALID_ARG. */
int  d2p_batch_job_stats(d2p_batch* b, void* cuda_stream, int njobs, int64_t* clipped_samples, double* peak_abs);

/* ---- time-block split of one long payload -------------------------------------------- */
/* Every output frame is a pure function of a bounded input window, so a payload can be cut in
 * output-frame space and each block converted independently (another stream, another GPU).
 * d2p_total_frames: frames per channel the whole payload yields.
 * d2p_block_range : the slice [*in_offset, *in_offset + *in_bytes) of the container-order payload that
 *                   frames [first_frame, first_frame + frame_count) read, filter halo included and
 *                   aligned to block groups (host only; needs no GPU). */
uint64_t d2p_total_frames(const d2p_config* cfg, uint64_t payload_bytes);
int      d2p_block_range(const d2p_config* cfg, uint64_t payload_bytes, uint64_t first_frame, uint64_t frame_count,
                         uint64_t* in_offset, uint64_t* in_bytes);
typedef struct d2p_block_job {
    const void* d_in;          /* device pointer: payload bytes [in_offset, in_offset + in_bytes) */
    uint64_t    in_offset;     /* as returned by d2p_block_range */
    uint64_t    in_bytes;
    uint64_t    payload_bytes; /* length of the whole payload (all channels) */
    uint64_t    first_frame;   /* absolute frame index of the block's first output frame */
    uint64_t    frame_count;
    void*       d_out;         /* device pointer: the block's packed PCM (frame first_frame at offset 0) */
    uint64_t    out_cap_bytes;
    double*     d_f64;         /* optional: channels * f64_stride doubles */
    uint64_t    f64_stride;
    uint64_t    dither_seed;   /* per-file seed: blocks of one file share it (noise is keyed by absolute frame) */
} d2p_block_job;
int  d2p_batch_run_blocks(d2p_batch* b, const d2p_block_job* jobs, int njobs, void* cuda_stream);

/* ---- device-resident conversion that allocates nothing ------------------------------------------- */
/* No counterpart in the reference (a CPU program has no device allocator to stay away from); it is d2p_batch_run for
 * hosts in which this library is not allowed to obtain memory from the CUDA allocator:
 *   - an application that gave the whole of HBM to an arena of its own, or runs under a fixed device-memory budget it
 *     accounts for itself (a mastering or transcoding service sharing the GPU with other work): a cudaMalloc made
 *     behind its back either fails or breaks that accounting, and d2p_batch_create makes several;
 *   - a caller that keeps other streams busy on the same GPU (playback, metering) and cannot afford the device-wide
 *     stall cudaMalloc / cudaFree may cause, nor their unbounded latency;
 *   - a worker that has to finish its files after the device has run out of memory.
 * The caller passes device pointers for input and output exactly as for d2p_batch_run.  What a d2p_batch would have
 * allocated besides -- job table, filter look-up table, statistics record -- lives in an array of
 * D2P_NOALLOC_SCRATCH_BYTES (256 KiB) compiled into the library; every GPU that loads the library carries that array
 * until the process exits, whether this entry point is used or not.  The call makes no cudaMalloc, cudaFree,
 * cudaMallocHost or any other allocating CUDA call (ordinary host heap use, such as building the filter table,
 * remains).
 *
 * Limits, each refused with a message in d2p_last_error(): at most D2P_NOALLOC_MAX_JOBS jobs per call
 * (D2P_ERR_CAPACITY); a filter table larger than the scratch (D2P_ERR_CAPACITY; none of the reference's filters is);
 * dither F (D2P_ERR_INVALID_ARG: it needs per-sample generator state on the device).  Payload sizes are not limited.
 *
 * jobs[i].frames is filled as by d2p_batch_run and the launch is asynchronous on `cuda_stream`.  clipped_samples /
 * peak_abs (either may be NULL) receive what d2p_batch_stats would report for THIS call alone; asking for either makes
 * the call wait for the stream.
 *
 * First call on a device: the library looks up the device address of its array there (cudaGetSymbolAddress, which
 * also loads the module on that device).  Make that first call -- any small conversion will do -- before beginning a
 * stream capture on that device; later calls only enqueue work on the stream.
 *
 * Sharing: all calls on one device use the same array, and it keeps what the last call wrote.  The entry point
 * re-zeroes the statistics and rewrites job table and filter table on the caller's stream every time, so back-to-back
 * calls on ONE stream are ordered and safe, with the same or with different configurations.  Two calls whose GPU work
 * can overlap -- different streams without an event between them, or different host threads -- overwrite each
 * other's tables while kernels read them: that is a data race, and the PCM of both is undefined.  Serialise them. */
#define D2P_NOALLOC_SCRATCH_BYTES (256 * 1024)
#define D2P_NOALLOC_MAX_JOBS 64
int  d2p_run_noalloc(const d2p_config* cfg, d2p_job* jobs, int njobs, void* cuda_stream,
                     int64_t* clipped_samples, double* peak_abs);

/* ---- diagnostics --------------------------------------------------------------------------------- */
/* The octave step of the float dither, `s.abs().log2().floor() as i32` (Dither::fpdither, rd/dither.rs:102-104), as
 * the kernels evaluate it, on n host samples: exponents[i] receives the value for samples[i].  Exists so that the one
 * libm-dependent expression of the path can be compared with the host's libm at the octave boundaries. */
int  d2p_debug_fpd_exponent(const double* samples, int n, int32_t* exponents, int device);

#ifdef __cplusplus
}
#endif
#endif /* DSD2PCM_B200_H */
