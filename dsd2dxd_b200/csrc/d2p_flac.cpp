// d2p_flac.cpp -- FLAC back end for the GPU-produced PCM ("FLAC stays a CPU encode of the
// GPU-produced samples; check by decode-equality").
//
// Stands in for AudioFile::save_flac_file, /root/reference/src/rdsd2pcm/src/audio_file.rs:454-573, which hands the
// samples to the third-party flac-codec 1.2.0 crate (not in the reference mount).  What is reproduced from the
// reference is everything that decides WHICH integers end up in the stream:
//   * 16 / 20 / 24 bit only, a 32-bit (float) AudioFile is written as 24 bit          (:477-491)
//   * 20-bit samples go in unshifted at bits_per_sample = 20                           (:553, no "<<= 4" here)
//   * float samples become (clamp(x, -1, 1) * 8388607.0f) as i32                       (:647-649)
//   * "No channels" / "No samples" are errors                                          (:463-475)
// The bitstream itself is this file's own: fixed-blocksize frames of 4096, independent channels, CONSTANT /
// VERBATIM / FIXED(0..4) subframes with partitioned Rice coding, STREAMINFO with the MD5 of the decoded audio and a
// VORBIS_COMMENT block.  It is a lossless container: tests decode it with an independent decoder and compare
// samples, MD5 and both CRCs (tests/test_flac.py).
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dsd2pcm_b200_files.h"

namespace d2p { int api_set_error(int st, const char* msg); }

namespace {

using d2p::api_set_error;

// ---- MD5 (RFC 1321) ------------------------------------------------------------------------------------------
struct Md5 {
    uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    uint64_t total = 0;
    uint8_t buf[64];
    size_t fill = 0;
    static uint32_t rol(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }
    void block(const uint8_t* p) {
        static const uint32_t K[64] = {
            0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af,
            0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa,
            0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8,
            0x676f02d9, 0x8d2a4c8a, 0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
            0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97,
            0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1,
            0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
        static const int S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                  4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
        uint32_t w[16];
        for (int i = 0; i < 16; ++i)
            w[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
        for (int i = 0; i < 64; ++i) {
            uint32_t f;
            int g;
            if (i < 16) { f = (b & c) | (~b & d); g = i; }
            else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
            else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
            else { f = c ^ (b | ~d); g = (7 * i) & 15; }
            const uint32_t t = d;
            d = c; c = b;
            b = b + rol(a + f + K[i] + w[g], S[i]);
            a = t;
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d;
    }
    void update(const uint8_t* p, size_t n) {
        total += n;
        if (fill) {
            const size_t take = std::min(n, 64 - fill);
            std::memcpy(buf + fill, p, take);
            fill += take; p += take; n -= take;
            if (fill < 64) return;
            block(buf);
            fill = 0;
        }
        for (; n >= 64; p += 64, n -= 64) block(p);
        if (n) { std::memcpy(buf, p, n); fill = n; }
    }
    void finish(uint8_t out[16]) {
        const uint64_t bits = total * 8;
        const uint8_t one = 0x80, zero = 0;
        update(&one, 1);
        while (fill != 56) update(&zero, 1);
        uint8_t len[8];
        for (int i = 0; i < 8; ++i) len[i] = (uint8_t)(bits >> (8 * i));
        update(len, 8);
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) out[4 * i + j] = (uint8_t)(h[i] >> (8 * j));
    }
};

// ---- CRCs of the frame layer ---------------------------------------------------------------------------------
struct CrcTables {
    uint8_t c8[256];
    uint16_t c16[256];
    CrcTables() {
        for (int i = 0; i < 256; ++i) {
            uint8_t a = (uint8_t)i;
            uint16_t b = (uint16_t)(i << 8);
            for (int k = 0; k < 8; ++k) {
                a = (uint8_t)((a & 0x80) ? (a << 1) ^ 0x07 : a << 1);          // x^8 + x^2 + x + 1
                b = (uint16_t)((b & 0x8000) ? (b << 1) ^ 0x8005 : b << 1);      // x^16 + x^15 + x^2 + 1
            }
            c8[i] = a;
            c16[i] = b;
        }
    }
};
const CrcTables& crc_tables() { static const CrcTables t; return t; }

// ---- MSB-first bit writer ------------------------------------------------------------------------------------
struct BitWriter {
    std::vector<uint8_t> v;
    uint64_t acc = 0;
    int nacc = 0;                                   // bits held in acc (< 8 after every call)
    void put(uint32_t value, int bits) {            // bits <= 32
        if (bits == 0) return;
        acc = (acc << bits) | (bits == 32 ? (uint64_t)value : (uint64_t)(value & ((1u << bits) - 1u)));
        nacc += bits;
        while (nacc >= 8) {
            v.push_back((uint8_t)(acc >> (nacc - 8)));
            nacc -= 8;
        }
        acc &= (1ull << nacc) - 1ull;
    }
    void put_signed(int32_t value, int bits) { put((uint32_t)value, bits); }
    void unary(uint32_t q) {                        // q zero bits, then a one
        while (q >= 32) { put(0, 32); q -= 32; }
        put(1, (int)q + 1);
    }
    void align() { if (nacc) put(0, 8 - nacc); }
};

constexpr int kBlock = 4096;

struct Plan {                                       // how one channel of one block is written
    int kind = 1;                                   // 0 CONSTANT, 1 VERBATIM, 2 FIXED
    int order = 0, porder = 0;
    bool rice2 = false;
    std::vector<uint8_t> k;                         // Rice parameter per partition
    uint64_t bits = 0;
};

inline uint32_t zigzag(int32_t r) { return ((uint32_t)r << 1) ^ (uint32_t)(r >> 31); }

void residual(const int32_t* x, int n, int order, std::vector<int32_t>& r) {
    r.resize((size_t)n);
    switch (order) {
        case 0: for (int i = 0; i < n; ++i) r[i] = x[i]; break;
        case 1: for (int i = 1; i < n; ++i) r[i] = x[i] - x[i - 1]; break;
        case 2: for (int i = 2; i < n; ++i) r[i] = x[i] - 2 * x[i - 1] + x[i - 2]; break;
        case 3: for (int i = 3; i < n; ++i) r[i] = x[i] - 3 * x[i - 1] + 3 * x[i - 2] - x[i - 3]; break;
        default: for (int i = 4; i < n; ++i) r[i] = x[i] - 4 * x[i - 1] + 6 * x[i - 2] - 4 * x[i - 3] + x[i - 4]; break;
    }
}

// cheapest Rice parameter for u[0..n): bits(k) = n (k + 1) + sum (u >> k)
uint64_t best_rice(const uint32_t* u, int n, int* kbest) {
    if (n == 0) { *kbest = 0; return 0; }
    uint64_t sum = 0;
    for (int i = 0; i < n; ++i) sum += u[i];
    int k0 = 0;
    while (k0 < 30 && ((uint64_t)n << (k0 + 1)) <= sum) ++k0;     // about log2(mean)
    uint64_t best = ~0ull;
    for (int k = std::max(0, k0 - 1); k <= std::min(30, k0 + 1); ++k) {
        uint64_t b = (uint64_t)n * (uint64_t)(k + 1);
        for (int i = 0; i < n; ++i) b += u[i] >> k;
        if (b < best) { best = b; *kbest = k; }
    }
    return best;
}

Plan plan_channel(const int32_t* x, int n, int bps, std::vector<int32_t>& r, std::vector<uint32_t>& u) {
    Plan p;
    p.kind = 1;
    p.bits = 8 + (uint64_t)bps * (uint64_t)n;
    bool constant = true;
    for (int i = 1; i < n && constant; ++i) constant = x[i] == x[0];
    if (constant) { p.kind = 0; p.bits = 8 + (uint64_t)bps; return p; }
    if (n <= 4) return p;
    // predictor order by the sum of magnitudes over the common range
    int order = 0;
    {
        uint64_t best = ~0ull;
        uint64_t s[5] = {0, 0, 0, 0, 0};
        for (int i = 4; i < n; ++i) {
            const int64_t a = x[i], b = x[i - 1], c = x[i - 2], d = x[i - 3], e = x[i - 4];
            const int64_t r0 = a, r1 = a - b, r2 = a - 2 * b + c, r3 = a - 3 * b + 3 * c - d, r4 = a - 4 * b + 6 * c - 4 * d + e;
            s[0] += (uint64_t)(r0 < 0 ? -r0 : r0); s[1] += (uint64_t)(r1 < 0 ? -r1 : r1); s[2] += (uint64_t)(r2 < 0 ? -r2 : r2);
            s[3] += (uint64_t)(r3 < 0 ? -r3 : r3); s[4] += (uint64_t)(r4 < 0 ? -r4 : r4);
        }
        for (int o = 0; o <= 4; ++o)
            if (s[o] < best) { best = s[o]; order = o; }
    }
    residual(x, n, order, r);
    u.resize((size_t)n);
    for (int i = order; i < n; ++i) u[i] = zigzag(r[i]);
    Plan f;
    f.kind = 2; f.order = order; f.bits = ~0ull;
    for (int po = 0; po <= 8; ++po) {
        if (n & ((1 << po) - 1)) break;
        const int plen = n >> po;
        if (plen <= order) break;
        std::vector<uint8_t> ks((size_t)1 << po);
        uint64_t bits = 0;
        int kmax = 0;
        for (int q = 0; q < (1 << po); ++q) {
            const int lo = q == 0 ? order : q * plen, hi = (q + 1) * plen;
            int k = 0;
            bits += best_rice(u.data() + lo, hi - lo, &k);
            ks[(size_t)q] = (uint8_t)k;
            kmax = std::max(kmax, k);
        }
        const bool rice2 = kmax > 14;
        bits += (uint64_t)(rice2 ? 5 : 4) << po;
        if (bits < f.bits) { f.bits = bits; f.porder = po; f.k = ks; f.rice2 = rice2; }
    }
    f.bits += 8 + (uint64_t)order * (uint64_t)bps + 6;
    return f.bits < p.bits ? f : p;
}

void write_subframe(BitWriter& w, const int32_t* x, int n, int bps, const Plan& p, const std::vector<int32_t>& r) {
    if (p.kind == 0) { w.put(0x00, 8); w.put_signed(x[0], bps); return; }
    if (p.kind == 1) {
        w.put(0x02, 8);                              // 0 | 000001 | 0
        for (int i = 0; i < n; ++i) w.put_signed(x[i], bps);
        return;
    }
    w.put((uint32_t)((0x08 | p.order) << 1), 8);     // 0 | 001ooo | 0
    for (int i = 0; i < p.order; ++i) w.put_signed(x[i], bps);
    w.put(p.rice2 ? 1 : 0, 2);
    w.put((uint32_t)p.porder, 4);
    const int plen = n >> p.porder;
    for (int q = 0; q < (1 << p.porder); ++q) {
        const int k = p.k[(size_t)q];
        w.put((uint32_t)k, p.rice2 ? 5 : 4);
        const int lo = q == 0 ? p.order : q * plen, hi = (q + 1) * plen;
        for (int i = lo; i < hi; ++i) {
            const uint32_t u = zigzag(r[(size_t)i]);
            w.unary(u >> k);
            w.put(u, k);
        }
    }
}

int sample_rate_code(uint32_t rate, int* extra_bits, uint32_t* extra) {
    *extra_bits = 0; *extra = 0;
    switch (rate) {
        case 88200: return 1; case 176400: return 2; case 192000: return 3; case 8000: return 4; case 16000: return 5;
        case 22050: return 6; case 24000: return 7; case 32000: return 8; case 44100: return 9; case 48000: return 10;
        case 96000: return 11; default: break;
    }
    if (rate % 10 == 0 && rate / 10 <= 0xFFFF) { *extra_bits = 16; *extra = rate / 10; return 14; }
    if (rate <= 0xFFFF) { *extra_bits = 16; *extra = rate; return 13; }
    return 0;                                       // "see STREAMINFO" (705.6 k, 768 k)
}

// one frame: `n` samples per channel of planar int32 data ch[c][0..n)
void encode_frame(std::vector<uint8_t>& out, const std::vector<std::vector<int32_t>>& ch, int n, int bps, uint32_t rate,
                  uint64_t frame_no) {
    const int C = (int)ch.size();
    BitWriter w;
    w.v.reserve((size_t)n * C * 3 + 64);
    w.put(0xFFF8, 16);
    int bs_code, bs_extra = 0;
    if (n == kBlock) bs_code = 12;
    else if (n <= 256) { bs_code = 6; bs_extra = 8; }
    else { bs_code = 7; bs_extra = 16; }
    int sr_bits; uint32_t sr_extra;
    const int sr_code = sample_rate_code(rate, &sr_bits, &sr_extra);
    w.put((uint32_t)bs_code, 4);
    w.put((uint32_t)sr_code, 4);
    w.put((uint32_t)(C - 1), 4);
    w.put(bps == 16 ? 4u : bps == 20 ? 5u : 6u, 3);
    w.put(0, 1);
    // frame number, UTF-8 style (up to 31 bits here)
    if (frame_no < 0x80) w.put((uint32_t)frame_no, 8);
    else {
        int extra = 1;
        while (extra < 5 && (frame_no >> (6 * extra + (6 - extra))) != 0) ++extra;   // payload bits: 6*extra + (6 - extra)
        w.put((uint32_t)(((0xFF00u >> (extra + 1)) & 0xFF) | (uint32_t)(frame_no >> (6 * extra))), 8);
        for (int i = extra - 1; i >= 0; --i) w.put(0x80u | (uint32_t)((frame_no >> (6 * i)) & 0x3F), 8);
    }
    if (bs_extra) w.put((uint32_t)(n - 1), bs_extra);
    if (sr_bits) w.put(sr_extra, sr_bits);
    const CrcTables& T = crc_tables();
    uint8_t c8 = 0;
    for (uint8_t b : w.v) c8 = T.c8[c8 ^ b];
    w.put(c8, 8);
    std::vector<int32_t> r;
    std::vector<uint32_t> u;
    for (int c = 0; c < C; ++c) {
        const Plan p = plan_channel(ch[(size_t)c].data(), n, bps, r, u);
        write_subframe(w, ch[(size_t)c].data(), n, bps, p, r);
    }
    w.align();
    uint16_t c16 = 0;
    for (uint8_t b : w.v) c16 = (uint16_t)((c16 << 8) ^ T.c16[(c16 >> 8) ^ b]);
    w.put(c16, 16);
    out.swap(w.v);
}

struct FileCloser { FILE* f; ~FileCloser() { if (f) std::fclose(f); } };

// packed little-endian interleaved payload (what d2p_convert_host returns for D2P_OUT_INTERLEAVED_LE) -> the integers the
// reference would hand to its encoder, planar
void unpack(const uint8_t* p, int bits, int C, size_t first, int n, std::vector<std::vector<int32_t>>& ch) {
    const int sb = bits == 16 ? 2 : bits == 32 ? 4 : 3;
    for (int c = 0; c < C; ++c) ch[(size_t)c].resize((size_t)n);
    for (int i = 0; i < n; ++i)
        for (int c = 0; c < C; ++c) {
            const uint8_t* s = p + ((first + (size_t)i) * (size_t)C + (size_t)c) * (size_t)sb;
            int32_t v;
            if (bits == 16) v = (int16_t)((uint16_t)s[0] | ((uint16_t)s[1] << 8));
            else if (bits == 32) {
                float x;
                std::memcpy(&x, s, 4);
                x = x != x ? 0.0f : std::min(1.0f, std::max(-1.0f, x));   // Rust: NaN.clamp stays NaN, `as i32` gives 0
                v = (int32_t)(x * 8388607.0f);                            // audio_file.rs:647-649
            } else {
                v = (int32_t)(((uint32_t)s[0] << 8) | ((uint32_t)s[1] << 16) | ((uint32_t)s[2] << 24)) >> 8;
                if (bits == 20) v >>= 4;                                  // the kernel left-aligns 20-bit samples; FLAC takes them as they are
            }
            ch[(size_t)c][(size_t)i] = v;
        }
}

}  // namespace

extern "C" int d2p_write_flac_file(const char* path, int bits, int channels, uint32_t rate, const void* payload, uint64_t frames,
                                   const char* const* comments, int n_comments) {
    return d2p_write_flac_file_ex(path, bits, channels, rate, payload, frames, comments, n_comments, nullptr, 0);
}

extern "C" int d2p_write_flac_file_ex(const char* path, int bits, int channels, uint32_t rate, const void* payload, uint64_t frames,
                                      const char* const* comments, int n_comments, const d2p_flac_picture* pictures, int n_pictures) {
    if (!path || (!payload && frames) || n_comments < 0 || (n_comments && !comments) || n_pictures < 0 || (n_pictures && !pictures))
        return api_set_error(D2P_ERR_INVALID_ARG, "null argument");
    if (channels <= 0) return api_set_error(D2P_ERR_INVALID_ARG, "No channels");
    if (channels > 8) return api_set_error(D2P_ERR_INVALID_ARG, "FLAC: at most 8 channels");
    if (frames == 0) return api_set_error(D2P_ERR_INVALID_ARG, "No samples");
    if (bits != 16 && bits != 20 && bits != 24 && bits != 32)
        return api_set_error(D2P_ERR_BIT_DEPTH, "FLAC: only 16, 20, or 24-bit supported");
    if (rate == 0 || rate >= (1u << 20)) return api_set_error(D2P_ERR_INVALID_ARG, "FLAC: sample rate does not fit STREAMINFO");
    if (frames >= (1ull << 36)) return api_set_error(D2P_ERR_INVALID_ARG, "FLAC: too many samples for STREAMINFO");
    const int bps = bits > 24 ? 24 : bits;
    const int md5_bytes = bps == 16 ? 2 : 3;
    FileCloser fc{std::fopen(path, "wb+")};
    if (!fc.f) return api_set_error(D2P_ERR_INVALID_ARG, (std::string("cannot create ") + path).c_str());

    // "fLaC", STREAMINFO (rewritten at the end), VORBIS_COMMENT, PICTURE blocks; the last one carries the end flag
    std::vector<uint8_t> head;
    auto be = [&](uint64_t x, int nbytes) { for (int i = nbytes - 1; i >= 0; --i) head.push_back((uint8_t)(x >> (8 * i))); };
    auto le32 = [&](uint32_t x) { for (int i = 0; i < 4; ++i) head.push_back((uint8_t)(x >> (8 * i))); };
    head.insert(head.end(), {'f', 'L', 'a', 'C'});
    be(0x00, 1); be(34, 3);
    const size_t si = head.size();
    head.insert(head.end(), 34, 0);
    // vendor string in the reference's format, `dsd2dxd v<rdsd2pcm crate version> Unix datetime <seconds>`
    // (PcmWriter::id3_to_flac_meta, rd/pcm_writer.rs:426-437; the crate is at 0.3.0).  SOURCE_DATE_EPOCH, when set,
    // replaces the clock so that two runs can give identical files.
    long long stamp = (long long)std::time(nullptr);
    if (const char* sde = std::getenv("SOURCE_DATE_EPOCH")) stamp = std::atoll(sde);
    const std::string vendor = "dsd2dxd v0.3.0 Unix datetime " + std::to_string(stamp);
    size_t vc = 4 + vendor.size() + 4;
    for (int i = 0; i < n_comments; ++i) vc += 4 + std::strlen(comments[i]);
    if (vc >= (1u << 24)) return api_set_error(D2P_ERR_INVALID_ARG, "FLAC: comment block too large");
    be((n_pictures ? 0x00 : 0x80) | 4, 1); be(vc, 3);
    le32((uint32_t)vendor.size()); head.insert(head.end(), vendor.begin(), vendor.end());
    le32((uint32_t)n_comments);
    for (int i = 0; i < n_comments; ++i) {
        const size_t L = std::strlen(comments[i]);
        le32((uint32_t)L); head.insert(head.end(), comments[i], comments[i] + L);
    }
    for (int i = 0; i < n_pictures; ++i) {                         // METADATA_BLOCK_PICTURE: all fields big-endian
        const d2p_flac_picture& p = pictures[i];
        const size_t ml = p.mime ? std::strlen(p.mime) : 0, dl = p.description ? std::strlen(p.description) : 0;
        const uint64_t sz = 32 + ml + dl + p.data_bytes;
        if (sz >= (1u << 24)) return api_set_error(D2P_ERR_INVALID_ARG, "FLAC: picture block too large");
        be((i == n_pictures - 1 ? 0x80 : 0x00) | 6, 1); be(sz, 3);
        be((uint32_t)p.type, 4);
        be(ml, 4); head.insert(head.end(), p.mime, p.mime + ml);
        be(dl, 4); head.insert(head.end(), p.description, p.description + dl);
        be(p.width, 4); be(p.height, 4); be(p.depth, 4); be(p.colors, 4);
        be(p.data_bytes, 4);
        head.insert(head.end(), (const uint8_t*)p.data, (const uint8_t*)p.data + p.data_bytes);
    }
    if (std::fwrite(head.data(), 1, head.size(), fc.f) != head.size()) return api_set_error(D2P_ERR_INVALID_ARG, "write failed");

    const uint64_t nblocks = (frames + kBlock - 1) / kBlock;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const int T = (int)std::min<uint64_t>(std::min(hw, 32u), nblocks);
    const uint64_t wave = (uint64_t)T * 16;                         // blocks encoded between two writes
    std::vector<std::vector<uint8_t>> enc((size_t)wave);
    Md5 md5;
    uint32_t min_frame = 0xFFFFFF, max_frame = 0;
    const uint8_t* src = (const uint8_t*)payload;
    std::vector<uint8_t> mbuf;
    for (uint64_t b0 = 0; b0 < nblocks; b0 += wave) {
        const uint64_t nb = std::min(wave, nblocks - b0);
        auto work = [&](int t) {
            std::vector<std::vector<int32_t>> ch((size_t)channels);
            for (uint64_t j = (uint64_t)t; j < nb; j += (uint64_t)T) {
                const uint64_t blk = b0 + j;
                const int n = (int)std::min<uint64_t>(kBlock, frames - blk * kBlock);
                unpack(src, bits, channels, (size_t)(blk * kBlock), n, ch);
                encode_frame(enc[(size_t)j], ch, n, bps, rate, blk);
            }
        };
        std::vector<std::thread> pool;
        for (int t = 1; t < T; ++t) pool.emplace_back(work, t);
        // the calling thread hashes this wave's decoded audio while the others start, then takes its share of blocks
        {
            const uint64_t f0 = b0 * kBlock, f1 = std::min<uint64_t>(frames, (b0 + nb) * kBlock);
            if (bits == 16 || bits == 24) {
                md5.update(src + f0 * (uint64_t)channels * md5_bytes, (size_t)((f1 - f0) * (uint64_t)channels * md5_bytes));
            } else {                                                 // 20-bit (shifted in the payload) and float: hash what is encoded
                std::vector<std::vector<int32_t>> ch((size_t)channels);
                for (uint64_t f = f0; f < f1; f += kBlock) {
                    const int n = (int)std::min<uint64_t>(kBlock, f1 - f);
                    unpack(src, bits, channels, (size_t)f, n, ch);
                    mbuf.resize((size_t)n * channels * 3);
                    for (int i = 0; i < n; ++i)
                        for (int c = 0; c < channels; ++c) {
                            const int32_t v = ch[(size_t)c][(size_t)i];
                            uint8_t* d = mbuf.data() + ((size_t)i * channels + c) * 3;
                            d[0] = (uint8_t)v; d[1] = (uint8_t)(v >> 8); d[2] = (uint8_t)(v >> 16);
                        }
                    md5.update(mbuf.data(), mbuf.size());
                }
            }
        }
        work(0);
        for (auto& th : pool) th.join();
        for (uint64_t j = 0; j < nb; ++j) {
            const std::vector<uint8_t>& e = enc[(size_t)j];
            min_frame = std::min<uint32_t>(min_frame, (uint32_t)e.size());
            max_frame = std::max<uint32_t>(max_frame, (uint32_t)e.size());
            if (std::fwrite(e.data(), 1, e.size(), fc.f) != e.size()) return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
        }
    }

    // STREAMINFO
    uint8_t s[34];
    s[0] = (uint8_t)(kBlock >> 8); s[1] = (uint8_t)kBlock; s[2] = s[0]; s[3] = s[1];
    s[4] = (uint8_t)(min_frame >> 16); s[5] = (uint8_t)(min_frame >> 8); s[6] = (uint8_t)min_frame;
    s[7] = (uint8_t)(max_frame >> 16); s[8] = (uint8_t)(max_frame >> 8); s[9] = (uint8_t)max_frame;
    const uint64_t packed = ((uint64_t)rate << 44) | ((uint64_t)(channels - 1) << 41) | ((uint64_t)(bps - 1) << 36) | frames;
    for (int i = 0; i < 8; ++i) s[10 + i] = (uint8_t)(packed >> (8 * (7 - i)));
    md5.finish(s + 18);
    if (std::fseek(fc.f, (long)si, SEEK_SET) != 0 || std::fwrite(s, 1, 34, fc.f) != 34)
        return api_set_error(D2P_ERR_INVALID_ARG, "write failed");
    return D2P_OK;
}
