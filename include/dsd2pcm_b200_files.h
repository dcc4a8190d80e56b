/*
 * dsd2pcm_b200_files.h -- container front end and back end around the conversion core
 * (SURVEY.md 8(f) rows 2 and 3): DSF / DSDIFF / raw DSD readers, WAV / AIFF / AIFC writers, and a
 * whole-file conversion call that strings them around d2p_convert_host().
 *
 * Reference interfaces these entry points stand in for:
 *   reader   DsdReader::from_container / DsdReader::new   (crate dsd-reader 0.2.0 -- source NOT in the
 *            reference mount; call sites rd/lib.rs:92-99, 171-183).  Re-created from the
 *            DSF (Sony "DSF File Format Specification" 1.01) and DSDIFF 1.5 layouts and the contract
 *            the call sites imply (per-channel block size, channel count, DSD rate, bit order, ID3 blob).
 *   writers  AudioFile::save_wave_file / save_aiff_file / save_aifc_file / encode_extended
 *            rd/audio_file.rs:93-207, 209-292, 294-452, 576-621   (byte for byte, quirks included)
 *   driver   Rdsd2Pcm::from_container + do_conversion + PcmWriter::save_file
 *            rd/lib.rs:171-182, 275-281, rd/pcm_writer.rs:300-318
 * (rd/ = src/rdsd2pcm/src/ of the reference.)  Host-side byte formatting only: every sample still
 * comes from the CUDA kernels; the payload reaches the file in the byte order the kernel packed it.
 *   flac     AudioFile::save_flac_file   rd/audio_file.rs:454-573   (which integers go in: 16 / 20 / 24 bit, float
 *            written as 24 bit; the bitstream is the encoder's own -- the reference delegates it to the
 *            flac-codec crate -- and is checked by decode equality)
 *   tags     PcmWriter::id3_to_flac_meta, ConversionContext::save_output / append_album_suffix
 *            rd/pcm_writer.rs:30-96, 425-605, rd/conversion_context.rs:565-621   (crates id3 1.16.3 and flac-codec are not
 *            in the mount: field mapping per the reference's tables, chunk bytes unpinned).  Output naming, output
 *            directories and artwork copying are the caller's business: tools/d2p_cli.cpp does them like the reference.
 */
#ifndef DSD2PCM_B200_FILES_H
#define DSD2PCM_B200_FILES_H

#include "dsd2pcm_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { D2P_CONTAINER_DSF = 'S', D2P_CONTAINER_DFF = 'F', D2P_CONTAINER_RAW = 'R' };
/* Output containers: the reference CLI's -o letters that carry PCM (src/main.rs:196-204) */
enum { D2P_FILE_WAV = 'W', D2P_FILE_AIFF = 'A', D2P_FILE_AIFC = 'C', D2P_FILE_RAW = 'S' /* stdout-style headerless LE */,
       D2P_FILE_FLAC = 'F' };

typedef struct d2p_container_info {
    int32_t  format;          /* D2P_CONTAINER_* */
    int32_t  channels;
    int32_t  dsd_rate;        /* sample_rate / 2 822 400: 1, 2, 4, 8 (0 if not a DSD64 multiple) */
    uint32_t sample_rate;     /* Hz, as stored */
    int32_t  in_layout;       /* D2P_LAYOUT_PLANAR (DSF) / D2P_LAYOUT_INTERLEAVED (DFF) */
    int32_t  lsb_first;       /* DSF bits_per_sample == 1 -> 1; DSF 8 and DFF -> 0 */
    uint32_t block_size;      /* bytes per channel per block (DSF: 4096; DFF: 1) */
    uint64_t data_offset;     /* file offset of the first payload byte */
    uint64_t data_bytes;      /* payload bytes present in the file (all channels, whole block groups for DSF) */
    uint64_t sample_count;    /* DSF: samples per channel from the fmt chunk; DFF: data_bytes * 8 / channels */
    uint64_t id3_offset;      /* ID3v2 blob (DSF metadata pointer / DFF "ID3 " chunk), 0 if none */
    uint64_t id3_bytes;       /* clipped to the file size (the reference tolerates truncated tags) */
} d2p_container_info;

/* Parse the headers of a DSF or DSDIFF file held in memory (`bytes` = what is available of the file's
 * beginning, `file_bytes` = the real file length; chunk sizes that point past it are clipped, as the
 * wrong FRM8 size of id3_test/dff/1kHz_stereo_i.dff demands).  Returns D2P_ERR_INVALID_ARG for
 * anything that is not DSD in a DSF / DSDIFF container (compressed DST included). */
int d2p_container_parse(const uint8_t* bytes, size_t nbytes, uint64_t file_bytes, d2p_container_info* info);
/* The same from a path (reads the header region and, for DSDIFF, walks the chunk list). */
int d2p_container_open(const char* path, d2p_container_info* info);

/* Header of a PCM file whose payload is `frames` frames of packed interleaved samples
 * (WAV: little-endian, AIFF / AIFC: big-endian; 20-bit samples left-aligned in 3 bytes).
 * Writes at most cap bytes to `out` and returns the header length (0 = unsupported combination:
 * AIFF with 32-bit float, bits outside 16/20/24/32).  AIFC places a FLLR chunk so that the sample
 * data starts on a 4096-byte boundary (rd/audio_file.rs:331-343); the header can therefore be 4 KiB. */
size_t d2p_pcm_header(int container, int bits, int channels, uint32_t rate, uint64_t frames, uint8_t* out, size_t cap);
/* header + payload -> file (D2P_FILE_FLAC: d2p_write_flac_file without comments) */
int d2p_write_pcm_file(const char* path, int container, int bits, int channels, uint32_t rate,
                       const void* payload, uint64_t frames);
/* FLAC file from `frames` frames of packed LITTLE-endian interleaved samples as the kernels emit them
 * (16: int16; 20: left-aligned in 3 bytes; 24: 3 bytes; 32: float32).  Like the reference: 20-bit samples are
 * stored unshifted at bits_per_sample 20, float becomes (clamp(x,-1,1) * 8388607.0f) truncated at 24 bits, zero
 * channels or zero frames is an error.  `comments` are n_comments "KEY=value" strings for the VORBIS_COMMENT
 * block (may be NULL / 0).  Lossless: decoding returns exactly those integers; STREAMINFO carries their MD5. */
int d2p_write_flac_file(const char* path, int bits, int channels, uint32_t rate, const void* payload, uint64_t frames,
                        const char* const* comments, int n_comments);
/* One FLAC PICTURE block (PcmWriter::add_picture, rd/pcm_writer.rs:583-603: front cover = 3, back cover = 4). */
typedef struct d2p_flac_picture {
    int32_t     type;
    const char* mime;
    const char* description;   /* UTF-8 */
    uint32_t    width, height, depth, colors;
    const void* data;
    uint64_t    data_bytes;
} d2p_flac_picture;
int d2p_write_flac_file_ex(const char* path, int bits, int channels, uint32_t rate, const void* payload, uint64_t frames,
                           const char* const* comments, int n_comments, const d2p_flac_picture* pictures, int n_pictures);
/* ID3v2.3 / 2.4 tag blob -> Vorbis comments, field for field like PcmWriter::id3_to_flac_meta (rd/pcm_writer.rs:30-96,
 * 425-580): "KEY=value" strings in the reference's insertion order, UTF-8, each NUL-terminated, the list closed by an
 * empty string.  Returns the number of comments (-1 if the blob does not parse, as for the truncated tags of the
 * *brokenid3* fixtures); *needed = bytes the list takes; nothing is written unless cap >= *needed. */
int d2p_id3_to_vorbis(const uint8_t* id3, size_t nbytes, char* out, size_t cap, size_t* needed);

typedef struct d2p_file_options {
    uint32_t out_rate;        /* -r */
    int32_t  filter;          /* -t  D2P_FILTER_* */
    int32_t  bits;            /* -b  16 / 20 / 24 / 32 */
    int32_t  dither;          /* -d  D2P_DITHER_* */
    double   level_db;        /* -l */
    int32_t  out_container;   /* -o  D2P_FILE_* */
    uint64_t dither_seed;     /* 0 with a dither other than X: drawn from the OS per file, like the reference's thread_rng
                                 (rd/dither.rs:64-76); the value used is reported in d2p_file_result.dither_seed */
    int32_t  device;
    int32_t  dsf_stop_at_sample_count;   /* 0: convert every whole block of the data chunk; 1: stop at fmt.sample_count */
    /* raw input (.dsd / anything that is not DSF or DSDIFF): the CLI's -c -f -e -s -i */
    int32_t  raw_channels, raw_layout, raw_lsb_first, raw_dsd_rate;
    uint32_t raw_block_size;
    /* metadata carry-over (ConversionContext::save_output, rd/conversion_context.rs:565-590).  The reference always does
     * it when the input has a readable ID3v2 tag; here it is a switch (0 = bare PCM file, as before):
     *   FLAC: Vorbis comments + front / back cover PICTURE blocks from the tag (id3_to_flac_meta);
     *   WAV / AIFF / AIFC: the tag copied into an "ID3 " chunk behind the sample data (tag.write_to_path). */
    int32_t  copy_tags;
    int32_t  append_rate;     /* -a: " [<rate>]" appended to the album title (append_album_suffix :592-601) */
} d2p_file_options;

typedef struct d2p_file_result {
    d2p_container_info in;
    uint64_t frames;          /* per channel */
    uint64_t out_bytes;       /* size of the file written */
    int64_t  clipped_samples;
    double   peak_abs;
    uint64_t dither_seed;     /* the seed the conversion used (see d2p_file_options.dither_seed) */
    uint64_t tag_bytes;       /* size of the ID3v2 tag carried over (0: none, unreadable, or copy_tags off) */
} d2p_file_result;

/* One input file -> one PCM file: parse the container, stream the payload through the GPU, write header + payload.
 * WAV / AIFF / AIFC / headerless output: 64 MiB slices through a streaming context (d2p_push / d2p_pull) with two
 * pinned input and two pinned output buffers, whatever the file length; reading slice n+1 and writing the PCM of slice
 * n-1 overlap with slice n on the GPU, while copy-in, kernel and copy-out of ONE slice run one after the other.
 * FLAC output: the whole payload and its PCM are held in pinned memory and go through d2p_batch_convert_host
 * (H2D / kernel / D2H of neighbouring chunks overlapped); the encoder needs the complete stream. */
int d2p_convert_file(const d2p_file_options* opt, const char* in_path, const char* out_path, d2p_file_result* res);
/* The same with the reference's cancellation flag and progress report (Rdsd2Pcm::do_conversion(&AtomicBool,
 * Option<Sender<ProgressUpdate>>), rd/lib.rs:275-281): see d2p_batch_set_control.  A cancelled or failed conversion
 * leaves no output file behind. */
int d2p_convert_file_ex(const d2p_file_options* opt, const char* in_path, const char* out_path, d2p_file_result* res,
                        const volatile int32_t* cancel_flag, d2p_progress_fn progress, void* user);

#ifdef __cplusplus
}
#endif
#endif /* DSD2PCM_B200_FILES_H */
