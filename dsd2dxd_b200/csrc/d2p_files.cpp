// d2p_files.cpp -- container readers (DSF / DSDIFF / raw), PCM file writers (WAV / AIFF / AIFC) and the
// whole-file conversion call of include/dsd2pcm_b200_files.h.  Host-side byte formatting only: every
// sample comes from the CUDA kernels through d2p_convert_host().
//
// Writers follow rd/audio_file.rs of the reference byte for byte (save_wave_file :93-207, save_aiff_file
// :209-292, save_aifc_file :294-452, encode_extended :576-621), including its quirks: no pad byte after an
// odd-sized data chunk, AIFF's SSND size field without the 8 offset/blockSize bytes, the bit-depth field 20
// for 20-bit samples stored in 3 bytes, compression type "NONE" for 24-bit AIFC.
// The reader is re-created from the DSF 1.01 and DSDIFF 1.5 layouts (the dsd-reader crate is absent from the
// reference mount).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <random>
#include <string>
#include <vector>

#include "../../include/dsd2pcm_b200_files.h"

namespace d2p { int api_set_error(int st, const char* msg);
// d2p_tags.cpp
bool tag_usable(const uint8_t* blob, size_t n);
bool tag_append_album(std::vector<uint8_t>* blob, const char* suffix);
void tag_to_flac(const uint8_t* blob, size_t n, std::vector<std::string>* comments, std::vector<d2p_flac_picture>* pics,
                 std::vector<std::vector<uint8_t>>* storage, std::vector<std::string>* strings); }   // d2p_api.cu: thread-local d2p_last_error()
using d2p::api_set_error;
using d2p::tag_usable;
using d2p::tag_append_album;
using d2p::tag_to_flac;

namespace {

uint32_t le32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
uint64_t le64(const uint8_t* p) { return (uint64_t)le32(p) | ((uint64_t)le32(p + 4) << 32); }
uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3]; }
uint64_t be64(const uint8_t* p) { return ((uint64_t)be32(p) << 32) | (uint64_t)be32(p + 4); }
uint16_t be16(const uint8_t* p) { return (uint16_t)(((uint16_t)p[0] << 8) | p[1]); }
bool tag(const uint8_t* p, const char* t) { return std::memcmp(p, t, 4) == 0; }

int rate_to_dsd(uint32_t fs) {
    if (fs == 0 || fs % 2822400u != 0) return 0;
    const uint32_t m = fs / 2822400u;
    return (m == 1 || m == 2 || m == 4 || m == 8) ? (int)m : 0;
}

// ---- DSF: "DSD " (28) + "fmt " (52) + "data" (12 + payload), everything little-endian -----------------
int parse_dsf(const uint8_t* b, size_t n, uint64_t file_bytes, d2p_container_info* o) {
    if (n < 92) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: header truncated");
    if (le64(b + 4) != 28) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: bad DSD chunk size");
    const uint64_t meta = le64(b + 20);
    const uint8_t* f = b + 28;
    if (!tag(f, "fmt ")) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: fmt chunk missing");
    const uint64_t fmt_size = le64(f + 4);
    if (fmt_size < 52 || fmt_size > n - 40) return        // n >= 92; no sum of file-supplied sizes can wrap
        api_set_error(D2P_ERR_INVALID_ARG, "DSF: bad fmt chunk size");
    if (le32(f + 12) != 1) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: format version is not 1");
    if (le32(f + 16) != 0) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: format id is not DSD raw");
    o->format = D2P_CONTAINER_DSF;
    o->channels = (int32_t)le32(f + 24);
    o->sample_rate = le32(f + 28);
    o->dsd_rate = rate_to_dsd(o->sample_rate);
    const uint32_t bps = le32(f + 32);
    if (bps != 1 && bps != 8) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: bits per sample must be 1 or 8");
    o->lsb_first = bps == 1 ? 1 : 0;
    o->sample_count = le64(f + 36);
    o->block_size = le32(f + 44);
    o->in_layout = D2P_LAYOUT_PLANAR;
    if (o->channels < 1 || o->channels > 8 || o->block_size == 0)
        return api_set_error(D2P_ERR_INVALID_ARG, "DSF: bad channel count or block size");
    const uint8_t* d = b + 28 + fmt_size;
    if (!tag(d, "data")) return api_set_error(D2P_ERR_INVALID_ARG, "DSF: data chunk missing");
    const uint64_t dsz = le64(d + 4);
    o->data_offset = 28 + fmt_size + 12;
    uint64_t avail = file_bytes > o->data_offset ? file_bytes - o->data_offset : 0;
    if (meta > o->data_offset && meta < file_bytes) avail = std::min<uint64_t>(avail, meta - o->data_offset);
    uint64_t bytes = dsz >= 12 ? dsz - 12 : 0;
    bytes = std::min(bytes, avail);
    const uint64_t group = (uint64_t)o->channels * o->block_size;
    o->data_bytes = bytes / group * group;                       // whole block groups only
    o->id3_offset = 0; o->id3_bytes = 0;
    if (meta != 0 && meta < file_bytes) { o->id3_offset = meta; o->id3_bytes = file_bytes - meta; }
    return D2P_OK;
}

// ---- DSDIFF: "FRM8" size "DSD " { chunk id(4) size(8, big-endian) data pad-to-even } ------------------
// `read` fetches [off, off+len) of the file into dst and returns the bytes it could deliver.
template <typename ReadFn>
int parse_dff(ReadFn read, uint64_t file_bytes, d2p_container_info* o) {
    uint8_t h[16];
    if (read(0, 16, h) < 16 || !tag(h, "FRM8") || !tag(h + 12, "DSD "))
        return api_set_error(D2P_ERR_INVALID_ARG, "DSDIFF: FRM8/DSD form header missing");
    // the FRM8 size field is not trusted (the reference's own fixture carries a wrong one): walk to the file's end
    o->format = D2P_CONTAINER_DFF;
    o->in_layout = D2P_LAYOUT_INTERLEAVED; o->lsb_first = 0; o->block_size = 1;
    o->channels = 0; o->sample_rate = 0; o->data_offset = 0; o->data_bytes = 0; o->id3_offset = 0; o->id3_bytes = 0;
    bool dsd_cmpr = false, have_data = false;
    uint64_t pos = 16;
    while (pos + 12 <= file_bytes) {
        uint8_t ch[12];
        if (read(pos, 12, ch) < 12) break;
        const uint64_t size = be64(ch + 4);
        const uint64_t body = pos + 12;
        const uint64_t avail = file_bytes - body;
        const uint64_t clipped = std::min(size, avail);
        if (tag(ch, "PROP")) {
            std::vector<uint8_t> p((size_t)std::min<uint64_t>(clipped, 1 << 16));
            if (read(body, p.size(), p.data()) < p.size()) return api_set_error(D2P_ERR_INVALID_ARG, "DSDIFF: PROP chunk truncated");
            if (p.size() >= 4 && tag(p.data(), "SND ")) {
                size_t q = 4;
                while (q + 12 <= p.size()) {
                    const uint64_t s2 = be64(p.data() + q + 4);
                    const uint8_t* d2 = p.data() + q + 12;
                    if (s2 > p.size() - q - 12) break;               // compare without adding: s2 comes from the file
                    if (tag(p.data() + q, "FS  ") && s2 >= 4) o->sample_rate = be32(d2);
                    else if (tag(p.data() + q, "CHNL") && s2 >= 2) o->channels = be16(d2);
                    else if (tag(p.data() + q, "CMPR") && s2 >= 4) dsd_cmpr = tag(d2, "DSD ");
                    q += 12 + (size_t)((s2 + 1) & ~(uint64_t)1);
                }
            }
        } else if (tag(ch, "DSD ")) {
            o->data_offset = body; o->data_bytes = clipped; have_data = true;
        } else if (tag(ch, "DST ")) {
            return api_set_error(D2P_ERR_INVALID_ARG, "DSDIFF: DST-compressed audio is not supported");
        } else if (tag(ch, "ID3 ")) {
            o->id3_offset = body; o->id3_bytes = clipped;
        }
        if (size > avail) break;                                   // a chunk that runs past the end is the last one
        pos = body + ((size + 1) & ~(uint64_t)1);
    }
    if (!have_data || o->channels < 1 || o->channels > 8 || o->sample_rate == 0)
        return api_set_error(D2P_ERR_INVALID_ARG, "DSDIFF: PROP/SND properties or DSD sound data chunk missing");
    if (!dsd_cmpr) return api_set_error(D2P_ERR_INVALID_ARG, "DSDIFF: compression type is not 'DSD '");
    o->dsd_rate = rate_to_dsd(o->sample_rate);
    o->data_bytes = o->data_bytes / o->channels * o->channels;
    o->sample_count = o->data_bytes / o->channels * 8;
    return D2P_OK;
}

// ---- AIFF 80-bit extended, as encode_extended (rd/audio_file.rs:576-621) -------------------------------
void encode_extended(double value, uint8_t out[10]) {
    std::memset(out, 0, 10);
    if (value <= 0.0) return;
    int exp = 0;
    while (value >= 2.0) { value *= 0.5; ++exp; }
    while (value < 1.0) { value *= 2.0; --exp; }
    uint16_t exp_field = (uint16_t)(exp + 16383);
    // value is in [1, 2): value * 2^63 is exact in f64 (a power-of-two scale) and an integer below 2^64, so the
    // reference's round() and its 2^64 overflow branch never change anything; both are kept for the letter of it
    const double scaled = std::round(value * 9223372036854775808.0);
    uint64_t mant;
    if (scaled >= 18446744073709551616.0) { mant = 1ull << 63; exp_field = (uint16_t)(exp_field + 1); }
    else mant = (uint64_t)scaled;
    out[0] = (uint8_t)(exp_field >> 8); out[1] = (uint8_t)exp_field;
    for (int i = 0; i < 8; ++i) out[2 + i] = (uint8_t)(mant >> (56 - 8 * i));
}

struct Bytes {
    std::vector<uint8_t> v;
    void raw(const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; v.insert(v.end(), b, b + n); }
    void str(const char* s) { raw(s, std::strlen(s)); }
    void u16le(uint16_t x) { uint8_t b[2] = {(uint8_t)x, (uint8_t)(x >> 8)}; raw(b, 2); }
    void u32le(uint32_t x) { uint8_t b[4] = {(uint8_t)x, (uint8_t)(x >> 8), (uint8_t)(x >> 16), (uint8_t)(x >> 24)}; raw(b, 4); }
    void u16be(uint16_t x) { uint8_t b[2] = {(uint8_t)(x >> 8), (uint8_t)x}; raw(b, 2); }
    void u32be(uint32_t x) { uint8_t b[4] = {(uint8_t)(x >> 24), (uint8_t)(x >> 16), (uint8_t)(x >> 8), (uint8_t)x}; raw(b, 4); }
    void zeros(size_t n) { v.insert(v.end(), n, 0); }
};

int sample_bytes(int bits) { return bits == 20 ? 3 : bits / 8; }

bool build_header(int container, int bits, int channels, uint32_t rate, uint64_t frames, Bytes* h) {
    if (bits != 16 && bits != 20 && bits != 24 && bits != 32) return false;
    if (channels < 1) return false;
    const bool is_float = bits == 32;
    // the reference computes every size in u32 (`as u32`): keep the wrap-around for > 4 GiB payloads
    const uint32_t data_size = (uint32_t)(frames * (uint64_t)channels * (uint64_t)sample_bytes(bits));
    const uint32_t block_align = (uint32_t)(channels * sample_bytes(bits));
    uint8_t ext[10];
    switch (container) {
    case D2P_FILE_WAV:                                             // save_wave_file :93-207
        h->str("RIFF"); h->u32le(data_size + 36); h->str("WAVE");
        h->str("fmt "); h->u32le(16);
        h->u16le(is_float ? 3 : 1); h->u16le((uint16_t)channels); h->u32le(rate);
        h->u32le(rate * block_align); h->u16le((uint16_t)block_align); h->u16le((uint16_t)bits);
        h->str("data"); h->u32le(data_size);
        return true;
    case D2P_FILE_AIFF:                                            // save_aiff_file :209-292 (integer only)
        if (is_float) return false;
        h->str("FORM"); h->u32be(data_size + 46); h->str("AIFF");
        h->str("COMM"); h->u32be(18);
        h->u16be((uint16_t)channels); h->u32be((uint32_t)frames); h->u16be((uint16_t)bits);
        encode_extended((double)rate, ext); h->raw(ext, 10);
        h->str("SSND"); h->u32be(data_size); h->u32be(0); h->u32be(0);
        return true;
    case D2P_FILE_AIFC: {                                          // save_aifc_file :294-452
        const char* name = is_float ? "Linear PCM, 32 bit big-endian floating point"
                         : bits == 16 ? "Linear PCM, 16 bit big-endian signed integer"
                                      : "Linear PCM, 24 bit big-endian signed integer";
        const char* ctype = is_float ? "fl32" : (bits == 16 ? "twos" : "NONE");
        const uint32_t name_len = (uint32_t)std::strlen(name);
        const uint32_t comm_size = 22 + 1 + name_len;
        const uint32_t comm_padded = (comm_size + 1) & ~1u;
        const uint32_t before_fllr = 32 + comm_padded;
        const uint32_t r = (before_fllr + 24) % 4096;
        const uint32_t fllr_data = r == 0 ? 0 : 4096 - r;
        const uint32_t fllr_chunk = fllr_data > 0 ? 8 + fllr_data : 0;
        h->str("FORM"); h->u32be(40 + comm_padded + fllr_chunk + data_size); h->str("AIFC");
        h->str("FVER"); h->u32be(4); h->u32be(0xA2805140u);
        h->str("COMM"); h->u32be(comm_padded);
        h->u16be((uint16_t)channels); h->u32be((uint32_t)frames); h->u16be((uint16_t)bits);
        encode_extended((double)rate, ext); h->raw(ext, 10);
        h->str(ctype);
        { uint8_t l = (uint8_t)name_len; h->raw(&l, 1); } h->str(name);
        if (comm_size % 2 != 0) h->zeros(1);
        if (fllr_data > 0) { h->str("FLLR"); h->u32be(fllr_data); h->zeros(fllr_data); }
        h->str("SSND"); h->u32be(data_size + 8); h->u32be(0); h->u32be(0);
        return true;
    }
    case D2P_FILE_RAW:
        return true;                                               // stdout mode: payload only
    default:
        return false;
    }
}

struct FileCloser { FILE* f; void close() { if (f) std::fclose(f); f = nullptr; } ~FileCloser() { close(); } };

}  // namespace

extern "C" {

int d2p_container_parse(const uint8_t* bytes, size_t nbytes, uint64_t file_bytes, d2p_container_info* info) {
    if (!bytes || !info) return api_set_error(D2P_ERR_INVALID_ARG, "null argument");
    std::memset(info, 0, sizeof *info);
    if (file_bytes < nbytes) file_bytes = nbytes;
    if (nbytes >= 4 && tag(bytes, "DSD ")) return parse_dsf(bytes, nbytes, file_bytes, info);
    if (nbytes >= 4 && tag(bytes, "FRM8")) {
        auto rd = [&](uint64_t off, size_t len, uint8_t* dst) -> size_t {
            if (off >= nbytes) return 0;
            const size_t n = (size_t)std::min<uint64_t>(len, nbytes - off);
            std::memcpy(dst, bytes + off, n);
            return n;
        };
        return parse_dff(rd, file_bytes, info);
    }
    return api_set_error(D2P_ERR_INVALID_ARG, "not a DSF or DSDIFF file");
}

int d2p_container_open(const char* path, d2p_container_info* info) {
    if (!path || !info) return api_set_error(D2P_ERR_INVALID_ARG, "null argument");
    std::memset(info, 0, sizeof *info);
    FileCloser fc{std::fopen(path, "rb")};
    if (!fc.f) return api_set_error(D2P_ERR_INVALID_ARG, (std::string("cannot open ") + path).c_str());
    std::fseek(fc.f, 0, SEEK_END);
    const uint64_t file_bytes = (uint64_t)std::ftell(fc.f);
    uint8_t head[256];
    std::fseek(fc.f, 0, SEEK_SET);
    const size_t got = std::fread(head, 1, sizeof head, fc.f);
    if (got >= 4 && tag(head, "DSD ")) return parse_dsf(head, got, file_bytes, info);
    if (got >= 4 && tag(head, "FRM8")) {
        FILE* f = fc.f;
        auto rd = [f](uint64_t off, size_t len, uint8_t* dst) -> size_t {
            if (std::fseek(f, (long)off, SEEK_SET) != 0) return 0;
            return std::fread(dst, 1, len, f);
        };
        return parse_dff(rd, file_bytes, info);
    }
    return api_set_error(D2P_ERR_INVALID_ARG, "not a DSF or DSDIFF file");
}

size_t d2p_pcm_header(int container, int bits, int channels, uint32_t rate, uint64_t frames, uint8_t* out, size_t cap) {
    Bytes h;
    if (!build_header(container, bits, channels, rate, frames, &h)) return 0;
    if (out) std::memcpy(out, h.v.data(), std::min(cap, h.v.size()));
    return h.v.size();
}

int d2p_write_pcm_file(const char* path, int container, int bits, int channels, uint32_t rate,
                       const void* payload, uint64_t frames) {
    if (!path || (!payload && frames)) return api_set_error(D2P_ERR_INVALID_ARG, "null argument");
    if (container == D2P_FILE_FLAC) return d2p_write_flac_file(path, bits, channels, rate, payload, frames, nullptr, 0);
    Bytes h;
    if (!build_header(container, bits, channels, rate, frames, &h)) {
        if (container == D2P_FILE_AIFF && bits == 32)
            return api_set_error(D2P_ERR_BIT_DEPTH, "32 bit float only allowed with wav, aifc, or stdout");
        return api_set_error(D2P_ERR_BIT_DEPTH, "Unsupported bit depth");
    }
    FileCloser fc{std::fopen(path, "wb")};
    if (!fc.f) return api_set_error(D2P_ERR_INVALID_ARG, (std::string("cannot create ") + path).c_str());
    const size_t bytes = (size_t)frames * channels * sample_bytes(bits);
    if (std::fwrite(h.v.data(), 1, h.v.size(), fc.f) != h.v.size() || (bytes && std::fwrite(payload, 1, bytes, fc.f) != bytes))
        return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
    return D2P_OK;
}

// The tag as an "ID3 " chunk behind the sample data of a RIFF / FORM file (what Tag::write_to_path does for WAV and
// AIFF files): chunk id, size (little-endian in RIFF, big-endian in FORM), blob, pad byte to an even length; the
// container's own size field grows by the same amount.  The PCM writers leave out the pad byte behind an odd-sized
// data / SSND chunk (the reference's quirk, see the file header) -- harmless while that chunk is the last one, but a
// chunk that follows must start on an even offset or RIFF / IFF readers, which skip the pad, land one byte into it:
// an odd-length file gets the missing pad first (mono or 3- / 5-channel 24-bit output with an odd frame count).
static int append_id3_chunk(const char* path, bool riff, const std::vector<uint8_t>& tag, uint64_t* added) {
    FileCloser f{std::fopen(path, "rb+")};
    if (!f.f) return api_set_error(D2P_ERR_INVALID_ARG, "cannot reopen the output file for the tag chunk");
    uint8_t head[8];
    if (std::fread(head, 1, 8, f.f) != 8) return api_set_error(D2P_ERR_INVALID_ARG, "short output file");
    if (std::fseek(f.f, 0, SEEK_END) != 0) return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
    const uint32_t pre = (uint32_t)(std::ftell(f.f) & 1);
    const uint32_t n = (uint32_t)tag.size(), padded = (n + 1) & ~1u;
    uint32_t total = riff ? le32(head + 4) : be32(head + 4);
    total += pre + 8 + padded;
    uint8_t ch[8] = {'I', 'D', '3', ' ', 0, 0, 0, 0}, sz[4];
    for (int i = 0; i < 4; ++i) { ch[4 + i] = (uint8_t)(riff ? n >> (8 * i) : n >> (8 * (3 - i))); sz[i] = (uint8_t)(riff ? total >> (8 * i) : total >> (8 * (3 - i))); }
    const uint8_t zero = 0;
    if ((pre && std::fwrite(&zero, 1, 1, f.f) != 1) || std::fwrite(ch, 1, 8, f.f) != 8 || std::fwrite(tag.data(), 1, n, f.f) != n ||
        (padded != n && std::fwrite(&zero, 1, 1, f.f) != 1) || std::fseek(f.f, 4, SEEK_SET) != 0 || std::fwrite(sz, 1, 4, f.f) != 4)
        return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
    if (added) *added = (uint64_t)pre + 8 + padded;
    return D2P_OK;
}

// The reference seeds its dither from the OS for every run (rand::thread_rng, rd/dither.rs:64-76).  A caller that
// leaves dither_seed at 0 gets the same behaviour: a seed drawn from the OS entropy source mixed with a hash of the
// input path, so that no two files share a noise sequence.  A non-zero seed is used as given (reproducible runs).
static uint64_t file_dither_seed(uint64_t given, int dither, const char* path) {
    if (given != 0 || dither == D2P_DITHER_NONE) return given;
    uint64_t h = 0xcbf29ce484222325ull;                           // FNV-1a of the path
    for (const char* c = path; *c; ++c) { h ^= (uint8_t)*c; h *= 0x100000001b3ull; }
    std::random_device rd;
    h ^= ((uint64_t)rd() << 32) | rd();
    return h ? h : 1;
}

int d2p_convert_file(const d2p_file_options* opt, const char* in_path, const char* out_path, d2p_file_result* res) {
    return d2p_convert_file_ex(opt, in_path, out_path, res, nullptr, nullptr, nullptr);
}

int d2p_convert_file_ex(const d2p_file_options* opt, const char* in_path, const char* out_path, d2p_file_result* res,
                        const volatile int32_t* cancel_flag, d2p_progress_fn progress, void* user) {
    if (!opt || !in_path || !out_path) return api_set_error(D2P_ERR_INVALID_ARG, "null argument");
    d2p_file_result local;
    d2p_file_result& R = res ? *res : local;
    std::memset(&R, 0, sizeof R);
    if (opt->out_container == D2P_FILE_AIFF && opt->bits == 32)
        return api_set_error(D2P_ERR_BIT_DEPTH, "32 bit float only allowed with wav, aifc, or stdout");
    if (opt->out_container != D2P_FILE_WAV && opt->out_container != D2P_FILE_AIFF && opt->out_container != D2P_FILE_AIFC &&
        opt->out_container != D2P_FILE_RAW && opt->out_container != D2P_FILE_FLAC)
        return api_set_error(D2P_ERR_INVALID_ARG, "output container must be W, A, C, S or F");

    d2p_container_info& ci = R.in;
    FileCloser fc{std::fopen(in_path, "rb")};
    if (!fc.f) return api_set_error(D2P_ERR_INVALID_ARG, (std::string("cannot open ") + in_path).c_str());
    uint8_t magic[4] = {0, 0, 0, 0};
    const size_t mg = std::fread(magic, 1, 4, fc.f);
    if (mg == 4 && (tag(magic, "DSD ") || tag(magic, "FRM8"))) {
        const int st = d2p_container_open(in_path, &ci);
        if (st != D2P_OK) return st;
        if (ci.dsd_rate == 0) return api_set_error(D2P_ERR_INVALID_ARG, "container sample rate is not DSD64/128/256/512");
    } else {                                                       // raw DSD: everything comes from the options
        std::fseek(fc.f, 0, SEEK_END);
        ci.format = D2P_CONTAINER_RAW;
        ci.channels = opt->raw_channels; ci.dsd_rate = opt->raw_dsd_rate; ci.sample_rate = 2822400u * (uint32_t)opt->raw_dsd_rate;
        ci.in_layout = opt->raw_layout; ci.lsb_first = opt->raw_lsb_first;
        ci.block_size = opt->raw_layout == D2P_LAYOUT_INTERLEAVED ? 1 : opt->raw_block_size;
        ci.data_offset = 0; ci.data_bytes = (uint64_t)std::ftell(fc.f);
        if (ci.channels < 1) return api_set_error(D2P_ERR_INVALID_ARG, "raw input needs a channel count");
        ci.data_bytes = ci.data_bytes / ci.channels * ci.channels;
        ci.sample_count = ci.data_bytes / ci.channels * 8;
    }

    d2p_config cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.dsd_rate = ci.dsd_rate; cfg.out_rate = opt->out_rate; cfg.filter = opt->filter; cfg.bits = opt->bits;
    cfg.dither = opt->dither; cfg.channels = ci.channels; cfg.level_db = opt->level_db;
    cfg.in_layout = ci.in_layout; cfg.lsb_first = ci.lsb_first;
    cfg.block_size = ci.in_layout == D2P_LAYOUT_INTERLEAVED ? 4096u : ci.block_size;   // interleaved: reader chunk only
    cfg.out_format = (opt->out_container == D2P_FILE_AIFF || opt->out_container == D2P_FILE_AIFC) ? D2P_OUT_INTERLEAVED_BE
                                                                                                  : D2P_OUT_INTERLEAVED_LE;
    cfg.dither_seed = file_dither_seed(opt->dither_seed, opt->dither, in_path); cfg.device = opt->device;
    R.dither_seed = cfg.dither_seed;
    d2p_plan_info pi;
    int st = d2p_plan(&cfg, &pi);
    if (st != D2P_OK) return st;

    // What the reader hands over: whole block groups read linearly from the file, then (DSF stopped at fmt.sample_count)
    // a ragged last group in which every channel's run is `tail` bytes long and the runs follow each other -- the layout
    // both conversion paths expect for a partial group.
    const uint64_t per_ch = ci.data_bytes / ci.channels;
    uint64_t linear_bytes = ci.data_bytes, tail = 0, per_ch_used = per_ch;
    const size_t group = (size_t)ci.channels * ci.block_size;
    if (ci.format == D2P_CONTAINER_DSF && opt->dsf_stop_at_sample_count) {
        per_ch_used = std::min<uint64_t>((ci.sample_count + 7) / 8, per_ch);
        linear_bytes = per_ch_used / ci.block_size * group;
        tail = per_ch_used % ci.block_size;
    }
    const uint64_t frames_max = d2p_frames_for_bytes(&cfg, per_ch_used);
    auto read_tail = [&](uint8_t* dst) -> bool {                   // the ragged group, compacted: [ch0 tail | ch1 tail | ...]
        std::vector<uint8_t> g(group);
        if (std::fseek(fc.f, (long)(ci.data_offset + linear_bytes), SEEK_SET) != 0 || std::fread(g.data(), 1, group, fc.f) != group) return false;
        for (int c = 0; c < ci.channels; ++c) std::memcpy(dst + (size_t)c * tail, g.data() + (size_t)c * ci.block_size, (size_t)tail);
        return true;
    };

    // ---- metadata carry-over (save_output, rd/conversion_context.rs:565-590) ----
    std::vector<uint8_t> id3;
    if (opt->copy_tags && ci.id3_offset && ci.id3_bytes >= 10 && ci.id3_bytes < (64u << 20)) {
        id3.resize((size_t)ci.id3_bytes);
        if (std::fseek(fc.f, (long)ci.id3_offset, SEEK_SET) != 0 || std::fread(id3.data(), 1, id3.size(), fc.f) != id3.size() ||
            !tag_usable(id3.data(), id3.size()))
            id3.clear();                                           // "Input file has no tag; skipping tag copy."
    }
    if (!id3.empty() && opt->append_rate) {
        const char* dot = nullptr;                                 // abbrev_rate_pair :603-621
        switch (opt->out_rate) {
        case 88200: dot = " [88.2K]"; break; case 96000: dot = " [96K]"; break; case 176400: dot = " [176.4K]"; break;
        case 192000: dot = " [192K]"; break; case 352800: dot = " [352.8K]"; break; case 384000: dot = " [384K]"; break;
        case 705600: dot = " [705.6K]"; break; case 1411200: dot = " [1411.2K]"; break; default: break;
        }
        if (dot) tag_append_album(&id3, dot);
    }
    auto fail = [&](int code) {                                    // no partly written file is left behind
        if (std::strncmp(out_path, "/dev/", 5) != 0) std::remove(out_path);
        return code;
    };

    uint64_t frames = 0;
    uint64_t tag_chunk_bytes = 0;                                  // "ID3 " chunk behind the samples, pads included
    if (opt->out_container == D2P_FILE_FLAC) {
        // The FLAC encoder wants the whole stream (STREAMINFO totals, MD5): payload and PCM are held in pinned memory and
        // go through the pipelined whole-payload conversion.
        const size_t out_cap = (size_t)frames_max * ci.channels * pi.bytes_per_sample;
        const size_t in_bytes = (size_t)(linear_bytes + tail * ci.channels);
        uint8_t* h_in = (uint8_t*)d2p_host_alloc(std::max<size_t>(in_bytes, 16));
        uint8_t* h_out = (uint8_t*)d2p_host_alloc(std::max<size_t>(out_cap, 16));
        struct Free { uint8_t* a; uint8_t* b; ~Free() { d2p_host_free(a); d2p_host_free(b); } } fr{h_in, h_out};
        if (!h_in || !h_out) return api_set_error(D2P_ERR_NOMEM, "pinned host allocation failed");
        if (std::fseek(fc.f, (long)ci.data_offset, SEEK_SET) != 0 || std::fread(h_in, 1, (size_t)linear_bytes, fc.f) != (size_t)linear_bytes ||
            (tail && !read_tail(h_in + linear_bytes)))
            return api_set_error(D2P_ERR_INVALID_ARG, "short read of the DSD payload");
        d2p_batch* b = nullptr;
        st = d2p_batch_create(&cfg, &b);
        if (st != D2P_OK) return st;
        d2p_batch_set_control(b, cancel_flag, progress, user);
        st = d2p_batch_convert_host(b, h_in, in_bytes, h_out, out_cap, &frames);
        if (st == D2P_OK) st = d2p_batch_stats(b, nullptr, &R.clipped_samples, &R.peak_abs);
        d2p_batch_destroy(b);
        if (st != D2P_OK) return st;                               // nothing has been written yet
        std::vector<std::string> comments, strings;
        std::vector<d2p_flac_picture> pics;
        std::vector<std::vector<uint8_t>> storage;
        if (!id3.empty()) tag_to_flac(id3.data(), id3.size(), &comments, &pics, &storage, &strings);
        std::vector<const char*> cptr;
        for (const std::string& c : comments) cptr.push_back(c.c_str());
        st = d2p_write_flac_file_ex(out_path, opt->bits, ci.channels, opt->out_rate, h_out, frames, cptr.data(), (int)cptr.size(),
                                    pics.data(), (int)pics.size());
        if (st != D2P_OK) return fail(st);
    } else {
        // WAV / AIFF / AIFC / headerless: the file is streamed through the GPU in 64 MiB slices with bounded pinned memory
        // (two input and two output buffers, whatever the length of the input: a 60-minute 6-channel DSD512 payload is
        // 61 GB).  The streaming context keeps the filter halo between slices, so the bytes equal the one-pass conversion.
        // Slice n+1 is read from the input file and the PCM of slice n-1 is written to the output file by helper threads
        // while slice n is on the GPU; D2P_FILE_SERIAL=1 switches the overlap off (for measuring it).
        Bytes h;
        if (opt->out_container != D2P_FILE_RAW && !build_header(opt->out_container, opt->bits, ci.channels, opt->out_rate, frames_max, &h))
            return api_set_error(D2P_ERR_BIT_DEPTH, "Unsupported bit depth");
        const size_t slice = std::max<size_t>(group, ((size_t)64 << 20) / group * group);
        const size_t fb = (size_t)ci.channels * pi.bytes_per_sample, out_cap = ((size_t)48 << 20) / fb * fb + fb;
        const size_t in_cap = std::max<size_t>(std::min<uint64_t>(slice, std::max<uint64_t>(linear_bytes, group)), 16);
        const char* serial_env = std::getenv("D2P_FILE_SERIAL");
        const bool overlap = !(serial_env && serial_env[0] == '1') && linear_bytes > slice;   // one slice: nothing to overlap
        uint8_t* h_in[2] = {(uint8_t*)d2p_host_alloc(in_cap), overlap ? (uint8_t*)d2p_host_alloc(in_cap) : nullptr};
        uint8_t* h_out[2] = {(uint8_t*)d2p_host_alloc(out_cap), overlap ? (uint8_t*)d2p_host_alloc(out_cap) : nullptr};
        struct Free { uint8_t** a; uint8_t** b; ~Free() { for (int i = 0; i < 2; ++i) { d2p_host_free(a[i]); d2p_host_free(b[i]); } } } fr{h_in, h_out};
        if (!h_in[0] || !h_out[0] || (overlap && (!h_in[1] || !h_out[1]))) return api_set_error(D2P_ERR_NOMEM, "pinned host allocation failed");
        d2p_ctx* ctx = nullptr;
        st = d2p_create(&cfg, &ctx);
        if (st != D2P_OK) return st;
        struct Ctx { d2p_ctx* c; ~Ctx() { d2p_destroy(c); } } guard{ctx};
        FileCloser of{std::fopen(out_path, "wb")};
        if (!of.f) return api_set_error(D2P_ERR_INVALID_ARG, (std::string("cannot create ") + out_path).c_str());
        if (!h.v.empty() && std::fwrite(h.v.data(), 1, h.v.size(), of.f) != h.v.size()) return fail(api_set_error(D2P_ERR_INVALID_ARG, "write failed"));
        // declared behind the buffers and files they touch: a pending helper is waited for before those go away
        std::future<bool> reading, writing;
        auto settle = [&]() -> bool {                              // true if every helper so far succeeded
            bool ok = true;
            if (reading.valid()) ok = reading.get() && ok;
            if (writing.valid()) ok = writing.get() && ok;
            return ok;
        };
        auto failw = [&](int code) { settle(); return fail(code); };
        int ob = 0;                                                // output buffer the next pull goes into
        auto drain = [&]() -> int {
            for (;;) {
                uint64_t n = 0;
                const int s2 = d2p_pull(ctx, h_out[ob], out_cap, &n);
                if (s2 != D2P_OK) return s2;
                if (n == 0) return D2P_OK;
                if (writing.valid() && !writing.get()) return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
                const uint8_t* src = h_out[ob];
                FILE* dst = of.f;
                auto put = [src, dst, fb, n]() { return std::fwrite(src, fb, (size_t)n, dst) == (size_t)n; };
                if (overlap) { writing = std::async(std::launch::async, put); ob ^= 1; }
                else if (!put()) return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
                frames += n;                                       // with the overlap on: handed to the writer
                if (progress) progress(user, frames, frames_max);
            }
        };
        if (std::fseek(fc.f, (long)ci.data_offset, SEEK_SET) != 0) return fail(api_set_error(D2P_ERR_INVALID_ARG, "short read of the DSD payload"));
        auto get = [&](uint8_t* dst, size_t n) { FILE* src = fc.f; return [src, dst, n]() { return std::fread(dst, 1, n, src) == n; }; };
        int ib = 0;
        if (overlap) reading = std::async(std::launch::async, get(h_in[0], (size_t)std::min<uint64_t>(slice, linear_bytes)));
        for (uint64_t done = 0; done < linear_bytes;) {
            if (cancel_flag && *cancel_flag) return failw(api_set_error(D2P_ERR_CANCELLED, "conversion cancelled"));
            const size_t n = (size_t)std::min<uint64_t>(slice, linear_bytes - done);
            const bool got = overlap ? reading.get() : get(h_in[ib], n)();
            if (!got) return failw(api_set_error(D2P_ERR_INVALID_ARG, "short read of the DSD payload"));
            const uint64_t next = std::min<uint64_t>(slice, linear_bytes - done - n);
            if (overlap && next) reading = std::async(std::launch::async, get(h_in[ib ^ 1], (size_t)next));
            if ((st = d2p_push(ctx, h_in[ib], n)) != D2P_OK || (st = drain()) != D2P_OK) return failw(st);
            done += n;
            if (overlap) ib ^= 1;
        }
        if (cancel_flag && *cancel_flag) return failw(api_set_error(D2P_ERR_CANCELLED, "conversion cancelled"));
        if (tail) {
            if (!read_tail(h_in[ib])) return failw(api_set_error(D2P_ERR_INVALID_ARG, "short read of the DSD payload"));
            if ((st = d2p_push(ctx, h_in[ib], (size_t)(tail * ci.channels))) != D2P_OK) return failw(st);
        }
        if ((st = d2p_finish(ctx)) != D2P_OK || (st = drain()) != D2P_OK) return failw(st);
        if (!settle()) return fail(api_set_error(D2P_ERR_INVALID_ARG, "write failed"));
        if (frames != frames_max) return fail(api_set_error(D2P_ERR_STATE, "internal: frame count differs from the plan"));
        int64_t clips = 0;
        if ((st = d2p_stats(ctx, &clips, &R.peak_abs, &R.clipped_samples)) != D2P_OK) return fail(st);
        if (std::fflush(of.f) != 0) return fail(api_set_error(D2P_ERR_INVALID_ARG, "write failed"));
        of.close();
        if (!id3.empty() && opt->out_container != D2P_FILE_RAW && (st = append_id3_chunk(out_path, opt->out_container == D2P_FILE_WAV, id3, &tag_chunk_bytes)) != D2P_OK)
            return fail(st);
    }
    R.frames = frames;
    R.tag_bytes = id3.size();
    if (opt->out_container == D2P_FILE_FLAC) {                     // compressed: the size is whatever the encoder produced
        FileCloser of{std::fopen(out_path, "rb")};
        if (of.f && std::fseek(of.f, 0, SEEK_END) == 0) R.out_bytes = (uint64_t)std::ftell(of.f);
        return D2P_OK;
    }
    Bytes h;
    build_header(opt->out_container, opt->bits, ci.channels, opt->out_rate, frames, &h);
    R.out_bytes = h.v.size() + frames * (uint64_t)ci.channels * (uint64_t)pi.bytes_per_sample;
    R.out_bytes += tag_chunk_bytes;
    return D2P_OK;
}

}  // extern "C"
