// d2p_tags.cpp -- metadata carry-over of the file front end, host-side byte formatting only:
//   * ID3v2.3 / 2.4 tag blob of a DSF / DSDIFF file  ->  FLAC Vorbis comments + PICTURE blocks, field for field like
//     PcmWriter::id3_to_flac_meta (rd/pcm_writer.rs:30-96, 425-605; "the same field mapping used by MusicBrainz Picard"),
//   * the same blob copied into WAV / AIFF / AIFC output as an "ID3 " chunk (ConversionContext::save_output,
//     rd/conversion_context.rs:565-590: tag.write_to_path after the file is saved),
//   * "-a": " [<rate>]" appended to the album title (append_album_suffix, rd/conversion_context.rs:592-621).
// The reference does this through the `id3` 1.16.3 and `flac-codec` crates, neither of which is in the mount: the frame
// layouts below are the published ID3v2.3 / 2.4 ones, the mapping is the reference's own table, and the byte layout of
// the embedded chunk (a verbatim copy here; the crate re-serialises the parsed frames) is NOT pinned -- what is checked
// is that a reader gets the same fields back (tests/test_tags.py against oracle/tags.py).
// A tag that does not parse (truncated: both *brokenid3* fixtures) is treated as absent, like a failed Tag::read_from.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/dsd2pcm_b200_files.h"

namespace d2p {
int api_set_error(int st, const char* msg);

namespace {

struct Frame { std::string id; std::vector<uint8_t> body; };
struct Tag { int major = 0; std::vector<Frame> frames; };

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
uint32_t syncsafe(const uint8_t* p) { return ((uint32_t)(p[0] & 0x7F) << 21) | ((uint32_t)(p[1] & 0x7F) << 14) | ((uint32_t)(p[2] & 0x7F) << 7) | (p[3] & 0x7F); }

std::vector<uint8_t> unsync(const uint8_t* p, size_t n) {           // FF 00 -> FF
    std::vector<uint8_t> o;
    o.reserve(n);
    for (size_t i = 0; i < n; ++i) {
        o.push_back(p[i]);
        if (p[i] == 0xFF && i + 1 < n && p[i + 1] == 0x00) ++i;
    }
    return o;
}

bool parse_id3(const uint8_t* b, size_t n, Tag* t) {
    if (n < 10 || std::memcmp(b, "ID3", 3) != 0) return false;
    const int major = b[3];
    if (major != 3 && major != 4) return false;
    const uint8_t flags = b[5];
    if ((b[6] | b[7] | b[8] | b[9]) & 0x80) return false;
    const size_t size = syncsafe(b + 6);
    if (10 + size > n) return false;                               // truncated tag: the id3 crate gives up, so do we
    std::vector<uint8_t> body(b + 10, b + 10 + size);
    if ((flags & 0x80) && major == 3) body = unsync(body.data(), body.size());
    size_t q = 0;
    if (flags & 0x40) {                                            // extended header
        if (body.size() < 4) return false;
        const size_t ext = major == 4 ? syncsafe(body.data()) : 4 + (size_t)be32(body.data());
        if (ext > body.size()) return false;
        q = ext;
    }
    t->major = major;
    while (q + 10 <= body.size()) {
        const uint8_t* h = body.data() + q;
        if (h[0] == 0) break;                                      // padding
        for (int i = 0; i < 4; ++i)
            if (!((h[i] >= 'A' && h[i] <= 'Z') || (h[i] >= '0' && h[i] <= '9'))) return !t->frames.empty();
        size_t fs = major == 4 ? syncsafe(h + 4) : be32(h + 4);
        const uint8_t ff = h[9];
        if (q + 10 + fs > body.size()) return false;
        Frame f;
        f.id.assign((const char*)h, 4);
        const uint8_t* d = h + 10;
        bool skip = false;
        if (major == 4) {
            if (ff & 0x0C) skip = true;                            // compressed / encrypted
            if ((ff & 0x01) && fs >= 4) { d += 4; fs -= 4; }        // data length indicator
            if ((ff & 0x02) || (flags & 0x80)) f.body = unsync(d, fs); else f.body.assign(d, d + fs);
        } else {
            if (ff & 0xC0) skip = true;
            f.body.assign(d, d + fs);
        }
        q += 10 + (major == 4 ? syncsafe(h + 4) : be32(h + 4));
        if (!skip) t->frames.push_back(std::move(f));
    }
    return true;
}

void put_utf8(std::string& o, uint32_t c) {
    if (c < 0x80) o.push_back((char)c);
    else if (c < 0x800) { o.push_back((char)(0xC0 | (c >> 6))); o.push_back((char)(0x80 | (c & 0x3F))); }
    else if (c < 0x10000) { o.push_back((char)(0xE0 | (c >> 12))); o.push_back((char)(0x80 | ((c >> 6) & 0x3F))); o.push_back((char)(0x80 | (c & 0x3F))); }
    else { o.push_back((char)(0xF0 | (c >> 18))); o.push_back((char)(0x80 | ((c >> 12) & 0x3F))); o.push_back((char)(0x80 | ((c >> 6) & 0x3F))); o.push_back((char)(0x80 | (c & 0x3F))); }
}

// Bytes that claim to be UTF-8 -> valid UTF-8: every maximal ill-formed subsequence becomes one U+FFFD (what
// String::from_utf8_lossy and Python's errors="replace" do).  FLAC requires Vorbis comments to be UTF-8.
std::string utf8_lossy(const uint8_t* p, size_t n) {
    std::string o;
    size_t i = 0;
    while (i < n) {
        const uint8_t b = p[i];
        if (b < 0x80) { o.push_back((char)b); ++i; continue; }
        int need = 0; uint8_t lo = 0x80, hi = 0xBF;                // bounds of the second byte (Unicode table 3-7)
        if (b >= 0xC2 && b <= 0xDF) need = 1;
        else if (b == 0xE0) { need = 2; lo = 0xA0; }
        else if (b == 0xED) { need = 2; hi = 0x9F; }               // no surrogates
        else if (b >= 0xE1 && b <= 0xEF) need = 2;
        else if (b == 0xF0) { need = 3; lo = 0x90; }
        else if (b >= 0xF1 && b <= 0xF3) need = 3;
        else if (b == 0xF4) { need = 3; hi = 0x8F; }
        size_t j = i + 1;
        int k = 0;
        for (; k < need && j < n; ++k, ++j)
            if (p[j] < (k == 0 ? lo : 0x80) || p[j] > (k == 0 ? hi : 0xBF)) break;
        if (need && k == need) o.append((const char*)p + i, j - i);
        else put_utf8(o, 0xFFFD);
        i = j;
    }
    return o;
}

// one encoded string [p, p+n) -> UTF-8 (embedded NULs kept: multi-value frames are split by the caller)
std::string decode(int enc, const uint8_t* p, size_t n) {
    std::string o;
    if (enc == 0) { for (size_t i = 0; i < n; ++i) put_utf8(o, p[i]); return o; }
    if (enc == 3) return utf8_lossy(p, n);
    bool be = enc == 2;
    size_t i = 0;
    if (enc == 1 && n >= 2) {
        if (p[0] == 0xFE && p[1] == 0xFF) { be = true; i = 2; }
        else if (p[0] == 0xFF && p[1] == 0xFE) { be = false; i = 2; }
    }
    for (; i + 1 < n; i += 2) {
        uint32_t c = be ? ((uint32_t)p[i] << 8) | p[i + 1] : ((uint32_t)p[i + 1] << 8) | p[i];
        if (c >= 0xD800 && c < 0xDC00 && i + 3 < n) {
            const uint32_t d = be ? ((uint32_t)p[i + 2] << 8) | p[i + 3] : ((uint32_t)p[i + 3] << 8) | p[i + 2];
            if (d >= 0xDC00 && d < 0xE000) { c = 0x10000 + ((c - 0xD800) << 10) + (d - 0xDC00); i += 2; }
        }
        if (c >= 0xD800 && c < 0xE000) c = 0xFFFD;                 // a surrogate without its partner
        put_utf8(o, c);
    }
    return o;
}

// position just behind the terminator of an encoded string starting at q (or n if there is none)
size_t skip_term(int enc, const uint8_t* p, size_t n, size_t q, size_t* end) {
    if (enc == 1 || enc == 2) {
        for (size_t i = q; i + 1 < n; i += 2) if (p[i] == 0 && p[i + 1] == 0) { *end = i; return i + 2; }
    } else {
        for (size_t i = q; i < n; ++i) if (p[i] == 0) { *end = i; return i + 1; }
    }
    *end = n;
    return n;
}

std::string strip_nuls(std::string s) {
    while (!s.empty() && s.back() == '\0') s.pop_back();
    return s;
}
std::vector<std::string> split_nul(const std::string& s) {         // `.split('\0').filter(|s| !s.is_empty())`
    std::vector<std::string> o;
    size_t a = 0;
    while (a <= s.size()) {
        const size_t z = s.find('\0', a);
        const std::string part = s.substr(a, z == std::string::npos ? std::string::npos : z - a);
        if (!part.empty()) o.push_back(part);
        if (z == std::string::npos) break;
        a = z + 1;
    }
    return o;
}

const Frame* find(const Tag& t, const char* id) {
    for (const Frame& f : t.frames) if (f.id == id) return &f;
    return nullptr;
}
bool text_of(const Frame* f, std::string* out) {                   // Content::Text
    if (!f || f->body.empty()) return false;
    *out = strip_nuls(decode(f->body[0], f->body.data() + 1, f->body.size() - 1));
    return true;
}
bool ieq(const std::string& a, const char* b) {
    if (a.size() != std::strlen(b)) return false;
    for (size_t i = 0; i < a.size(); ++i) if (std::tolower((unsigned char)a[i]) != std::tolower((unsigned char)b[i])) return false;
    return true;
}
// "n/total" -> (n, total) as the id3 crate parses TRCK / TPOS: unsigned integers, total optional
bool pair_of(const std::string& s, bool* has_n, unsigned long* n, bool* has_t, unsigned long* tot) {
    *has_n = *has_t = false;
    const size_t slash = s.find('/');
    const std::string a = s.substr(0, slash), b = slash == std::string::npos ? "" : s.substr(slash + 1);
    char* e = nullptr;
    if (!a.empty()) { *n = std::strtoul(a.c_str(), &e, 10); *has_n = e && *e == 0; }
    if (!b.empty()) { *tot = std::strtoul(b.c_str(), &e, 10); *has_t = e && *e == 0; }
    return *has_n;
}

const char* const kMap[][2] = {    // ID3_VORBIS_MAP  rd/pcm_writer.rs:33-65
    {"TPE1", "ARTIST"}, {"TPE2", "ALBUMARTIST"}, {"TSO2", "ALBUMARTISTSORT"}, {"TSOA", "ALBUMSORT"}, {"TSOP", "ARTISTSORT"},
    {"TALB", "ALBUM"}, {"TIT2", "TITLE"}, {"TBPM", "BPM"}, {"TCOM", "COMPOSER"}, {"TSOC", "COMPOSERSORT"}, {"TPE3", "CONDUCTOR"},
    {"TCOP", "COPYRIGHT"}, {"TCON", "GENRE"}, {"TIT1", "GROUPING"}, {"TKEY", "KEY"}, {"TSRC", "ISRC"}, {"TPUB", "LABEL"},
    {"TLAN", "LANGUAGE"}, {"TEXT", "LYRICIST"}, {"TMED", "MEDIA"}, {"TMOO", "MOOD"}, {"MVNM", "MOVEMENTNAME"},
    {"TOFN", "ORIGINALFILENAME"}, {"TDRL", "RELEASEDATE"}, {"TPE4", "REMIXER"}, {"TIT3", "SUBTITLE"}, {"TSOT", "TITLESORT"},
    {"TENC", "ENCODEDBY"}, {"TSSE", "ENCODERSETTINGS"}, {"TCMP", "COMPILATION"}, {"TSST", "DISCSUBTITLE"}};
const char* const kTxxx[][2] = {   // ID3_TXXX_VORBIS_MAP  :69-87
    {"MusicBrainz Artist Id", "MUSICBRAINZ_ARTISTID"}, {"MusicBrainz Album Id", "MUSICBRAINZ_ALBUMID"},
    {"MusicBrainz Album Artist Id", "MUSICBRAINZ_ALBUMARTISTID"}, {"MusicBrainz Release Group Id", "MUSICBRAINZ_RELEASEGROUPID"},
    {"MusicBrainz Release Track Id", "MUSICBRAINZ_RELEASETRACKID"}, {"MusicBrainz TRM Id", "MUSICBRAINZ_TRMID"},
    {"MusicBrainz Work Id", "MUSICBRAINZ_WORKID"}, {"MusicBrainz Disc Id", "MUSICBRAINZ_DISCID"},
    {"MusicBrainz Original Artist Id", "MUSICBRAINZ_ORIGINALARTISTID"}, {"MusicBrainz Original Album Id", "MUSICBRAINZ_ORIGINALALBUMID"},
    {"MusicBrainz Album Release Country", "RELEASECOUNTRY"}, {"MusicBrainz Album Status", "RELEASESTATUS"},
    {"MusicBrainz Album Type", "RELEASETYPE"}, {"Acoustid Id", "ACOUSTID_ID"}, {"Acoustid Fingerprint", "ACOUSTID_FINGERPRINT"},
    {"MusicIP PUID", "MUSICIP_PUID"}, {"Writer", "WRITER"}};
const char* const kTipl[][2] = {{"arranger", "ARRANGER"}, {"engineer", "ENGINEER"}, {"dj-mix", "DJMIXER"}, {"mix", "MIXER"},
                                {"producer", "PRODUCER"}};   // ID3_TIPL_VORBIS_MAP  :90-96

// involved-people frames (TIPL / TMCL / IPLS): encoding byte, then alternating NUL-terminated role / name strings
std::vector<std::pair<std::string, std::string>> people(const Frame& f) {
    std::vector<std::pair<std::string, std::string>> o;
    if (f.body.empty()) return o;
    const int enc = f.body[0];
    const uint8_t* p = f.body.data();
    const size_t n = f.body.size();
    size_t q = 1;
    std::vector<std::string> items;
    while (q < n) {
        size_t end;
        const size_t nx = skip_term(enc, p, n, q, &end);
        items.push_back(decode(enc, p + q, end - q));
        q = nx;
    }
    for (size_t i = 0; i + 1 < items.size(); i += 2) o.emplace_back(items[i], items[i + 1]);
    return o;
}

void to_vorbis(const Tag& t, std::vector<std::string>* out) {
    auto ins = [&](const std::string& k, const std::string& v) { out->push_back(k + "=" + v); };
    std::string s;
    for (const auto& m : kMap)                                                         // :439-448
        if (text_of(find(t, m[0]), &s)) for (const std::string& v : split_nul(s)) ins(m[1], v);
    bool hn, ht; unsigned long n = 0, tot = 0;
    if (text_of(find(t, "TRCK"), &s)) {                                                // :450-458
        pair_of(s, &hn, &n, &ht, &tot);
        if (hn) ins("TRACKNUMBER", std::to_string(n));
        if (hn && ht) { ins("TOTALTRACKS", std::to_string(tot)); ins("TRACKTOTAL", std::to_string(tot)); }
    }
    if (text_of(find(t, "TPOS"), &s)) {                                                // :460-468
        pair_of(s, &hn, &n, &ht, &tot);
        if (hn) ins("DISCNUMBER", std::to_string(n));
        if (hn && ht) { ins("TOTALDISCS", std::to_string(tot)); ins("DISCTOTAL", std::to_string(tot)); }
    }
    if (text_of(find(t, "TDRC"), &s) && !s.empty()) ins("DATE", s);                    // :470-475
    else if (text_of(find(t, "TYER"), &s)) { char* e = nullptr; const long y = std::strtol(s.c_str(), &e, 10); if (e && *e == 0 && !s.empty()) ins("DATE", std::to_string(y)); }
    if (text_of(find(t, "TDOR") ? find(t, "TDOR") : find(t, "TORY"), &s)) ins("ORIGINALDATE", s);   // :477-484
    if (text_of(find(t, "MVIN"), &s)) { ins("MOVEMENTTOTAL", s); ins("MOVEMENT", s); }              // :486-492
    for (const char* id : {"COMM", "USLT"}) {                                          // :494-506  enc, lang[3], description\0, text
        const Frame* f = find(t, id);
        if (f && f->body.size() >= 4) {
            const int enc = f->body[0];
            size_t end;
            const size_t q = skip_term(enc, f->body.data(), f->body.size(), 4, &end);
            ins(id[0] == 'C' ? "COMMENT" : "LYRICS", strip_nuls(decode(enc, f->body.data() + q, f->body.size() - q)));
        }
    }
    for (const Frame& f : t.frames)                                                    // :508-521
        if (f.id == "TIPL" || f.id == "IPLS")
            for (const auto& it : people(f))
                for (const auto& m : kTipl) if (ieq(it.first, m[0])) ins(m[1], it.second);
    for (const Frame& f : t.frames)                                                    // :523-535
        if (f.id == "TMCL") for (const auto& it : people(f)) ins("PERFORMER", it.second + " (" + it.first + ")");
    for (const Frame& f : t.frames)                                                    // :537-550  owner\0 identifier
        if (f.id == "UFID") {
            const size_t z = std::find(f.body.begin(), f.body.end(), (uint8_t)0) - f.body.begin();
            if (z < f.body.size() && std::string((const char*)f.body.data(), z) == "http://musicbrainz.org")
                ins("MUSICBRAINZ_TRACKID", utf8_lossy(f.body.data() + z + 1, f.body.size() - z - 1));
        }
    for (const auto& w : {std::make_pair("WCOP", "LICENSE"), std::make_pair("WOAR", "WEBSITE")}) {   // :552-560
        const Frame* f = find(t, w.first);
        if (f) ins(w.second, strip_nuls(decode(0, f->body.data(), f->body.size())));
    }
    for (const Frame& f : t.frames)                                                    // :562-579  enc, description\0, value
        if (f.id == "TXXX" && !f.body.empty()) {
            const int enc = f.body[0];
            size_t end;
            const size_t q = skip_term(enc, f.body.data(), f.body.size(), 1, &end);
            const std::string desc = decode(enc, f.body.data() + 1, end - 1);
            if (desc.empty()) continue;
            std::string name = desc;
            for (const auto& m : kTxxx) if (ieq(desc, m[0])) { name = m[1]; break; }
            for (const std::string& v : split_nul(strip_nuls(decode(enc, f.body.data() + q, f.body.size() - q)))) ins(name, v);
        }
}

struct Picture { int type; std::string mime, desc; std::vector<uint8_t> data; uint32_t w = 0, h = 0, depth = 0, colors = 0; };

void image_geometry(Picture* p) {                                   // what flac_codec's Picture::new reads out of the image
    const std::vector<uint8_t>& d = p->data;
    if (d.size() >= 26 && std::memcmp(d.data(), "\x89PNG\r\n\x1a\n", 8) == 0) {
        p->w = be32(d.data() + 16); p->h = be32(d.data() + 20);
        const int bit = d[24], ct = d[25];
        p->depth = (uint32_t)bit * (ct == 2 ? 3 : ct == 6 ? 4 : ct == 4 ? 2 : 1);
        if (ct == 3) p->depth = 24;                                // palette: the colours are RGB triples
        return;
    }
    if (d.size() >= 4 && d[0] == 0xFF && d[1] == 0xD8) {           // JPEG: first SOFn segment
        size_t q = 2;
        while (q + 9 < d.size() && d[q] == 0xFF) {
            const int m = d[q + 1];
            const size_t len = ((size_t)d[q + 2] << 8) | d[q + 3];
            if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
                p->h = ((uint32_t)d[q + 5] << 8) | d[q + 6]; p->w = ((uint32_t)d[q + 7] << 8) | d[q + 8];
                p->depth = (uint32_t)d[q + 4] * d[q + 9];
                return;
            }
            q += 2 + len;
        }
    }
}

void pictures_of(const Tag& t, std::vector<Picture>* out) {         // :583-603: front and back cover only
    for (const Frame& f : t.frames) {
        if (f.id != "APIC" || f.body.size() < 4) continue;
        const int enc = f.body[0];
        const uint8_t* p = f.body.data();
        const size_t n = f.body.size();
        size_t end;
        size_t q = skip_term(0, p, n, 1, &end);                    // MIME type, Latin-1
        Picture pic;
        pic.mime.assign((const char*)p + 1, end - 1);
        if (q >= n) continue;
        const int type = p[q++];
        if (type != 3 && type != 4) continue;
        const size_t d0 = q;
        q = skip_term(enc, p, n, d0, &end);
        pic.desc = decode(enc, p + d0, end - d0);
        pic.type = type;
        pic.data.assign(p + q, p + n);
        image_geometry(&pic);
        out->push_back(std::move(pic));
    }
}

// " [352.8K]" on the album title: rebuild the TALB frame inside a copy of the blob (same encoding), fix the tag size
bool append_album(std::vector<uint8_t>* blob, const char* suffix) {
    Tag t;
    if (!parse_id3(blob->data(), blob->size(), &t)) return false;
    const uint8_t* b = blob->data();
    if (b[5] & 0xC0) return false;                                 // unsynchronised / extended-header tags are left alone
    const size_t size = syncsafe(b + 6);
    size_t q = 10;
    while (q + 10 <= 10 + size && b[q] != 0) {
        const size_t fs = t.major == 4 ? syncsafe(b + q + 4) : be32(b + q + 4);
        // format flags that change the frame body (compression, encryption, grouping byte, unsynchronisation, data
        // length indicator) sit in different bits in v2.3 and v2.4: such a frame is left as it is
        if (std::memcmp(b + q, "TALB", 4) == 0 && fs >= 1 && !(b[q + 9] & (t.major == 4 ? 0x4F : 0xE0))) {
            const int enc = b[q + 10];
            std::vector<uint8_t> body(b + q + 10, b + q + 10 + fs);
            while (body.size() > 1 && body.back() == 0 && (enc == 0 || enc == 3)) body.pop_back();
            while (body.size() > 2 && (enc == 1 || enc == 2) && body[body.size() - 1] == 0 && body[body.size() - 2] == 0 && (body.size() - 1) % 2 == 0) { body.pop_back(); body.pop_back(); }
            bool be = enc == 2;
            if (enc == 1 && body.size() >= 3) be = body[1] == 0xFE;
            for (const char* c = suffix; *c; ++c) {
                if (enc == 0 || enc == 3) body.push_back((uint8_t)*c);
                else if (be) { body.push_back(0); body.push_back((uint8_t)*c); }
                else { body.push_back((uint8_t)*c); body.push_back(0); }
            }
            const size_t nfs = body.size(), delta = nfs - fs;
            std::vector<uint8_t> o(b, b + q + 10);
            o.insert(o.end(), body.begin(), body.end());
            o.insert(o.end(), b + q + 10 + fs, b + blob->size());
            auto put_size = [&](size_t at, size_t v, bool safe) {
                if (safe) { o[at] = (v >> 21) & 0x7F; o[at + 1] = (v >> 14) & 0x7F; o[at + 2] = (v >> 7) & 0x7F; o[at + 3] = v & 0x7F; }
                else { o[at] = (uint8_t)(v >> 24); o[at + 1] = (uint8_t)(v >> 16); o[at + 2] = (uint8_t)(v >> 8); o[at + 3] = (uint8_t)v; }
            };
            put_size(q + 4, nfs, t.major == 4);
            put_size(6, size + delta, true);
            blob->swap(o);
            return true;
        }
        q += 10 + fs;
    }
    return false;
}

}  // namespace

// ---- used by d2p_files.cpp -------------------------------------------------------------------------------------------
bool tag_usable(const uint8_t* blob, size_t n) { Tag t; return parse_id3(blob, n, &t); }
bool tag_append_album(std::vector<uint8_t>* blob, const char* suffix) { return append_album(blob, suffix); }
void tag_to_flac(const uint8_t* blob, size_t n, std::vector<std::string>* comments, std::vector<d2p_flac_picture>* pics,
                 std::vector<std::vector<uint8_t>>* storage, std::vector<std::string>* strings) {
    Tag t;
    if (!parse_id3(blob, n, &t)) return;
    to_vorbis(t, comments);
    std::vector<Picture> ps;
    pictures_of(t, &ps);
    strings->reserve(strings->size() + 2 * ps.size());
    for (Picture& p : ps) {
        storage->push_back(std::move(p.data));
        strings->push_back(p.mime); const char* mime = strings->back().c_str();
        strings->push_back(p.desc); const char* desc = strings->back().c_str();
        d2p_flac_picture fp;
        fp.type = p.type; fp.mime = mime; fp.description = desc; fp.width = p.w; fp.height = p.h; fp.depth = p.depth; fp.colors = p.colors;
        fp.data = storage->back().data(); fp.data_bytes = storage->back().size();
        pics->push_back(fp);
    }
}

}  // namespace d2p

extern "C" {

// "KEY=value" strings, NUL-separated, double-NUL terminated, into out (UTF-8); returns the number of comments, or -1
int d2p_id3_to_vorbis(const uint8_t* id3, size_t nbytes, char* out, size_t cap, size_t* needed) {
    if (!id3 && nbytes) return -1;
    d2p::Tag t;
    std::vector<std::string> c;
    if (nbytes && d2p::parse_id3(id3, nbytes, &t)) d2p::to_vorbis(t, &c);
    else if (nbytes) { if (needed) *needed = 1; if (out && cap) out[0] = 0; return -1; }
    size_t need = 1;
    for (const std::string& s : c) need += s.size() + 1;
    if (needed) *needed = need;
    if (out && cap >= need) {
        size_t q = 0;
        for (const std::string& s : c) { std::memcpy(out + q, s.data(), s.size()); q += s.size(); out[q++] = 0; }
        out[q] = 0;
    }
    return (int)c.size();
}

}  // extern "C"
