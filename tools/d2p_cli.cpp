// d2p_cli.cpp -- a native caller of the C ABI (include/dsd2pcm_b200.h, include/dsd2pcm_b200_files.h) that mirrors
// the reference command line (clap struct in src/main.rs:40-133 of clone206/dsd2dxd) for the conversion path:
//
//     d2p_cli [-c channels] [-f I|P] [-b 16|20|24|32] [-t E|X|D|C] [-e M|L] [-s block] [-d T|R|F|X] [-r rate]
//             [-i 1|2|4|8] [-o S|A|C|W|F] [-l dB | --level=dB] [-a] [-p outdir] [-q] [--seed n] [--device n] FILES... | -
//
//   * FILES: every input goes through d2p_convert_file_ex (container parse -> GPU -> header + payload); the output
//     name follows ConversionContext::get_out_filename_path (rd/conversion_context.rs:400-440): <stem>[_<rate>].<ext>
//     in the input's directory (or -p), extensions wav / aif / aifc / flac, "-a" appends _88_2K ... _1411_2K.
//   * "-" : raw DSD from stdin, converted through the STREAMING seam exactly like process_blocks
//     (rd/conversion_context.rs:295-328): one block group (block_size bytes per channel) per d2p_push, PCM pulled
//     and written to stdout (-o S) as it becomes ready.  This is the call pattern a Rust host would drive.
//   * Ctrl-C raises the cancel flag the conversion polls (CANCEL_FLAG, src/main.rs:37, 239-246).
//   * prints the reference's own summary line "DSP speed for <file>: <x>x, End-to-end: <y>x"
//     (report_timing, rd/conversion_context.rs:641-679) on stderr unless -q.
//
// It links nothing but libdsd2pcm_b200.so: no Python, no CUDA headers.  Build: see the Makefile in this directory or
//   g++ -O2 -std=c++17 -I include tools/d2p_cli.cpp -o tools/d2p_cli -L dsd2dxd_b200 -ldsd2pcm_b200 -Wl,-rpath,'$ORIGIN/../dsd2dxd_b200'
#include <cctype>
#include <chrono>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <string>
#include <vector>

#include "dsd2pcm_b200.h"
#include "dsd2pcm_b200_files.h"

static volatile int32_t g_cancel = 0;
static void on_sigint(int) { g_cancel = 1; }

struct Cli {
    int channels = 2;
    char format = 'I';
    int bits = 24;
    char filter = 'E';
    char endian = 'M';
    unsigned block = 4096;
    char dither = 0;            // 0 = default: F for 32 bit, T otherwise (src/main.rs:165-181)
    unsigned rate = 352800;
    int in_rate = 1;
    char output = 'S';
    double level = 0.0;
    bool append = false, quiet = false;
    std::string out_dir;
    unsigned long long seed = 0;
    int device = 0;
    std::vector<std::string> files;
};

static int usage(const char* msg) {
    if (msg) std::fprintf(stderr, "d2p_cli: %s\n", msg);
    std::fprintf(stderr, "usage: d2p_cli [-c N] [-f I|P] [-b 16|20|24|32] [-t E|X|D|C] [-e M|L] [-s bytes] [-d T|R|F|X] [-r Hz] "
                         "[-i 1|2|4|8] [-o S|A|C|W|F] [-l dB] [-a] [-p dir] [-q] [--seed n] [--device n] FILES...|-\n");
    return 2;
}

static char upper(const char* s) { return (char)std::toupper((unsigned char)s[0]); }

static bool parse(int argc, char** argv, Cli* c) {
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto val = [&](const char* longname) -> const char* {
            const std::string eq = std::string("--") + longname + "=";
            if (a.rfind(eq, 0) == 0) return argv[i] + eq.size();
            if (i + 1 < argc) return argv[++i];
            return nullptr;
        };
        auto is = [&](const char* s, const char* l) { return a == s || a == std::string("--") + l || a.rfind(std::string("--") + l + "=", 0) == 0; };
        const char* v = nullptr;
        if (a == "-") c->files.push_back(a);
        else if (is("-c", "channels")) { if (!(v = val("channels"))) return false; c->channels = std::atoi(v); }
        else if (is("-f", "fmt")) { if (!(v = val("fmt"))) return false; c->format = upper(v); }
        else if (is("-b", "bitdepth")) { if (!(v = val("bitdepth"))) return false; c->bits = std::atoi(v); }
        else if (is("-t", "filttype")) { if (!(v = val("filttype"))) return false; c->filter = upper(v); }
        else if (is("-e", "endianness")) { if (!(v = val("endianness"))) return false; c->endian = upper(v); }
        else if (is("-s", "bs")) { if (!(v = val("bs"))) return false; c->block = (unsigned)std::strtoul(v, nullptr, 10); }
        else if (is("-d", "dither")) { if (!(v = val("dither"))) return false; c->dither = upper(v); }
        else if (is("-r", "rate")) { if (!(v = val("rate"))) return false; c->rate = (unsigned)std::strtoul(v, nullptr, 10); }
        else if (is("-i", "inrate")) { if (!(v = val("inrate"))) return false; c->in_rate = std::atoi(v); }
        else if (is("-o", "output")) { if (!(v = val("output"))) return false; c->output = upper(v); }
        else if (is("-l", "level")) { if (!(v = val("level"))) return false; c->level = std::atof(v); }
        else if (is("-p", "path")) { if (!(v = val("path"))) return false; c->out_dir = v; }
        else if (a == "--seed" || a.rfind("--seed=", 0) == 0) { if (!(v = val("seed"))) return false; c->seed = std::strtoull(v, nullptr, 0); }
        else if (a == "--device" || a.rfind("--device=", 0) == 0) { if (!(v = val("device"))) return false; c->device = std::atoi(v); }
        else if (a == "-a" || a == "--append") c->append = true;
        else if (a == "-q" || a == "--quiet") c->quiet = true;
        else if (a == "-v" || a == "--verbose") {}
        else if (a.size() > 1 && a[0] == '-') return false;
        else c->files.push_back(a);
    }
    return true;
}

static const char* rate_suffix(unsigned rate) {       // abbrev_rate_pair, rd/conversion_context.rs:603-621
    switch (rate) {
    case 88200: return "88_2K"; case 96000: return "96K"; case 176400: return "176_4K"; case 192000: return "192K";
    case 352800: return "352_8K"; case 384000: return "384K"; case 705600: return "705_6K"; case 1411200: return "1411_2K";
    default: return nullptr;
    }
}

// Base directory of main() (src/main.rs:262-273): one input -> its parent; several -> the parent of their longest
// common path.  Inputs are canonicalised first, as the reference does.
static std::filesystem::path base_dir_of(const std::vector<std::string>& files) {
    namespace fs = std::filesystem;
    std::vector<fs::path> paths;
    std::error_code ec;
    for (const std::string& f : files) {
        if (f == "-") continue;
        fs::path p = fs::canonical(f, ec);
        if (!ec) paths.push_back(p);
    }
    if (paths.empty()) return fs::path("/");
    if (paths.size() == 1) return paths[0].has_parent_path() ? paths[0].parent_path() : fs::path("/");
    fs::path common;
    for (auto it = paths[0].begin(); it != paths[0].end(); ++it) {
        const fs::path cand = common / *it;
        bool all = true;
        for (const fs::path& q : paths) {
            auto a = cand.begin(); auto b = q.begin();
            for (; a != cand.end() && b != q.end() && *a == *b; ++a, ++b) {}
            if (a != cand.end()) { all = false; break; }
        }
        if (!all) break;
        common = cand;
    }
    if (common.empty()) common = "/";
    return common.has_parent_path() && common != common.root_path() ? common.parent_path() : fs::path("/");
}

// get_out_filename_path (rd/conversion_context.rs:400-445) + derive_output_dir (:447-477): <stem>[_<rate>].<ext>, in the
// input's directory, or with -p in <outdir>/<input directory relative to the base directory> (created when missing);
// stdin: "output[_<rate>].<ext>" in <outdir> or the working directory.
static std::string out_name(const Cli& c, const std::string& in, const std::filesystem::path& base_dir) {
    namespace fs = std::filesystem;
    const char* ext = c.output == 'W' ? "wav" : c.output == 'A' ? "aif" : c.output == 'C' ? "aifc" : c.output == 'F' ? "flac" : "out";
    std::string stem = "output";
    fs::path dir;
    std::error_code ec;
    if (in == "-") {
        if (!c.out_dir.empty()) dir = c.out_dir;
    } else {
        fs::path full = fs::canonical(in, ec);
        if (ec) full = fs::absolute(in);
        if (full.has_stem()) stem = full.stem().string();
        const fs::path parent = full.parent_path();
        if (c.out_dir.empty()) {
            const size_t slash = in.find_last_of('/');                 // as given on the command line
            dir = slash == std::string::npos ? "" : in.substr(0, slash);
            if (slash == 0) dir = "/";
        } else {
            const fs::path rel = parent.lexically_relative(base_dir);
            const bool inside = !rel.empty() && *rel.begin() != "..";
            dir = fs::path(c.out_dir) / (inside ? (rel == "." ? fs::path() : rel) : parent);   // strip_prefix(..).unwrap_or(parent)
        }
    }
    if (!c.out_dir.empty() && !fs::exists(dir, ec)) fs::create_directories(dir, ec);
    std::string name = stem;
    if (c.append && rate_suffix(c.rate)) name += std::string("_") + rate_suffix(c.rate);
    return (dir / (name + "." + ext)).string();
}

// copy_artwork (rd/conversion_context.rs:479-533): with -p, every .jpg / .jpeg / .png of the input's directory is copied
// next to the output unless a file of that name is already there
static void copy_artwork(const Cli& c, const std::string& in, const std::string& out) {
    namespace fs = std::filesystem;
    if (c.out_dir.empty() || in == "-") return;
    std::error_code ec;
    const fs::path src = fs::path(in).parent_path().empty() ? fs::path(".") : fs::path(in).parent_path();
    const fs::path dst = fs::path(out).parent_path();
    for (const auto& e : fs::directory_iterator(src, ec)) {
        if (!e.is_regular_file(ec)) continue;
        std::string ext = e.path().extension().string();
        for (char& ch : ext) ch = (char)std::tolower((unsigned char)ch);
        if (ext != ".jpg" && ext != ".png" && ext != ".jpeg") continue;
        const fs::path to = dst / e.path().filename();
        if (!fs::exists(to, ec)) fs::copy_file(e.path(), to, fs::copy_options::none, ec);
    }
}

static void report(const Cli& c, const std::string& file, double audio_s, double dsp_s, double total_s) {
    if (c.quiet) return;
    const size_t slash = file.find_last_of('/');
    std::fprintf(stderr, "DSP speed for %s: %.2fx, End-to-end: %.2fx\n", (slash == std::string::npos ? file : file.substr(slash + 1)).c_str(),
                 audio_s / (dsp_s > 1e-9 ? dsp_s : 1e-9), audio_s / (total_s > 1e-9 ? total_s : 1e-9));
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// stdin -> container file "output[_<rate>].<ext>" (convert_stdin src/main.rs:316-346, file name :418-425; the reference's
// run_all_tests.sh does `-o w -p out . -`).  A container header needs the frame count before the first sample: stdin is
// spooled to a temporary raw file beside the output and converted like any headerless input, in bounded memory.
static int run_stdin_to_file(const Cli& c, char dither) {
    namespace fs = std::filesystem;
    const std::string out = out_name(c, "-", fs::path("/"));
    const std::string spool = out + ".stdin.tmp";
    struct Rm { const std::string& p; ~Rm() { std::error_code ec; std::filesystem::remove(p, ec); } } rm{spool};
    unsigned long long bytes = 0;
    {
        std::FILE* f = std::fopen(spool.c_str(), "wb");
        if (!f) { std::fprintf(stderr, "d2p_cli: cannot create %s\n", spool.c_str()); return 1; }
        std::vector<uint8_t> buf((size_t)4 << 20);
        size_t n;
        bool ok = true;
        while (!g_cancel && (n = std::fread(buf.data(), 1, buf.size(), stdin)) > 0) { ok = std::fwrite(buf.data(), 1, n, f) == n; bytes += n; if (!ok) break; }
        ok = std::fclose(f) == 0 && ok;
        if (!ok) { std::fprintf(stderr, "d2p_cli: write failed: %s\n", spool.c_str()); return 1; }
        if (g_cancel) { std::fprintf(stderr, "d2p_cli: cancelled\n"); return 1; }
    }
    d2p_file_options o;
    std::memset(&o, 0, sizeof o);
    o.out_rate = c.rate; o.filter = c.filter; o.bits = c.bits; o.dither = dither; o.level_db = c.level; o.out_container = c.output;
    o.dither_seed = c.seed; o.device = c.device;
    o.raw_channels = c.channels; o.raw_layout = c.format; o.raw_lsb_first = c.endian == 'L'; o.raw_dsd_rate = c.in_rate;
    o.raw_block_size = c.block;
    d2p_file_result r;
    const double t0 = now();
    const int st = d2p_convert_file_ex(&o, spool.c_str(), out.c_str(), &r, &g_cancel, nullptr, nullptr);
    const double dt = now() - t0;
    if (st != D2P_OK) { std::fprintf(stderr, "d2p_cli: stdin: %s\n", st == D2P_ERR_CANCELLED ? "cancelled" : d2p_last_error()); return 1; }
    report(c, "stdin", (double)bytes / c.channels * 8.0 / (2822400.0 * c.in_rate), dt, dt);
    if (!c.quiet && r.clipped_samples) std::fprintf(stderr, "%lld samples clipped\n", (long long)r.clipped_samples);
    return 0;
}

// stdin -> stdout through d2p_push / d2p_pull, one reader block group per call (process_blocks)
static int run_stdin(const Cli& c, char dither) {
    d2p_config cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.dsd_rate = c.in_rate; cfg.out_rate = c.rate; cfg.filter = c.filter; cfg.bits = c.bits; cfg.dither = dither;
    cfg.channels = c.channels; cfg.level_db = c.level; cfg.in_layout = c.format; cfg.lsb_first = c.endian == 'L';
    cfg.block_size = c.block; cfg.out_format = D2P_OUT_INTERLEAVED_LE; cfg.dither_seed = c.seed; cfg.device = c.device;
    if (c.output != 'S') return run_stdin_to_file(c, dither);
    d2p_plan_info pi;
    d2p_ctx* ctx = nullptr;
    int st = d2p_plan(&cfg, &pi);
    if (st == D2P_OK) st = d2p_create(&cfg, &ctx);
    if (st != D2P_OK) { std::fprintf(stderr, "d2p_cli: %s\n", d2p_last_error()); return 1; }
    const char* thr = std::getenv("D2P_CLI_PUSH_THRESHOLD");
    if (thr) d2p_set_push_threshold(ctx, (size_t)std::strtoull(thr, nullptr, 10));
    const size_t group = (size_t)c.block * c.channels;              // what one dsd_iter() step hands over
    std::vector<uint8_t> in(group), out((size_t)4 << 20);
    const size_t fb = (size_t)c.channels * pi.bytes_per_sample;
    const double t0 = now();
    double dsp = 0.0;
    unsigned long long bytes = 0, calls = 0;
    auto drain = [&](bool all) -> int {
        for (;;) {
            uint64_t n = 0;
            const int s = d2p_pull(ctx, out.data(), out.size(), &n);
            if (s != D2P_OK) return s;
            if (n) std::fwrite(out.data(), fb, (size_t)n, stdout);
            if (n == 0 || !all) return D2P_OK;
        }
    };
    for (;;) {
        if (g_cancel) { st = D2P_ERR_CANCELLED; break; }           // the reference polls once per block (:304-308)
        const size_t n = std::fread(in.data(), 1, group, stdin);
        if (n == 0) break;
        const double a = now();
        st = d2p_push(ctx, in.data(), n);
        if (st == D2P_OK && d2p_frames_ready(ctx) * fb >= out.size() / 2) st = drain(false);
        dsp += now() - a;
        bytes += n; ++calls;
        if (st != D2P_OK || n < group) break;
    }
    if (st == D2P_OK) { const double a = now(); st = d2p_finish(ctx); if (st == D2P_OK) st = drain(true); dsp += now() - a; }
    std::fflush(stdout);
    if (st != D2P_OK) std::fprintf(stderr, "d2p_cli: %s\n", st == D2P_ERR_CANCELLED ? "cancelled" : d2p_last_error());
    else {
        const double audio = (double)bytes / c.channels * 8.0 / (2822400.0 * c.in_rate);
        report(c, "stdin", audio, dsp, now() - t0);
        if (!c.quiet) std::fprintf(stderr, "seam: %llu pushes of %zu bytes, %.3g DSD samples/s through d2p_push/d2p_pull\n", calls, group,
                                   (double)bytes * 8.0 / (dsp > 1e-9 ? dsp : 1e-9));
    }
    d2p_destroy(ctx);
    return st == D2P_OK ? 0 : 1;
}

int main(int argc, char** argv) {
    Cli c;
    if (!parse(argc, argv, &c)) return usage("bad arguments");
    if (c.files.empty()) return usage("no input");
    if (c.format != 'I' && c.format != 'P') return usage("Invalid format; must be I (interleaved) or P (planar)");
    if (c.endian != 'M' && c.endian != 'L') return usage("Invalid endianness; must be M or L");
    const char dither = c.dither ? c.dither : (c.bits == 32 ? 'F' : 'T');
    std::signal(SIGINT, on_sigint);
    if (d2p_device_count() < 1) { std::fprintf(stderr, "d2p_cli: no CUDA device (the library has no CPU path)\n"); return 1; }
    int failed = 0;
    const std::filesystem::path base_dir = base_dir_of(c.files);
    for (const std::string& f : c.files) {
        if (f == "-") { failed += run_stdin(c, dither) != 0; continue; }
        d2p_file_options o;
        std::memset(&o, 0, sizeof o);
        o.out_rate = c.rate; o.filter = c.filter; o.bits = c.bits; o.dither = dither; o.level_db = c.level;
        o.out_container = c.output == 'S' ? D2P_FILE_RAW : c.output;
        o.dither_seed = c.seed; o.device = c.device;
        o.raw_channels = c.channels; o.raw_layout = c.format; o.raw_lsb_first = c.endian == 'L'; o.raw_dsd_rate = c.in_rate;
        o.raw_block_size = c.block;
        o.copy_tags = 1;                        // the reference always carries a readable tag over (save_output :565-590)
        o.append_rate = c.append ? 1 : 0;
        d2p_file_result r;
        const std::string out = c.output == 'S' ? "/dev/stdout" : out_name(c, f, base_dir);
        const double t0 = now();
        const int st = d2p_convert_file_ex(&o, f.c_str(), out.c_str(), &r, &g_cancel, nullptr, nullptr);
        const double dt = now() - t0;
        if (st != D2P_OK) { std::fprintf(stderr, "d2p_cli: %s: %s\n", f.c_str(), st == D2P_ERR_CANCELLED ? "cancelled" : d2p_last_error()); ++failed; continue; }
        if (c.output != 'S') copy_artwork(c, f, out);
        const double audio = (double)r.in.data_bytes / r.in.channels * 8.0 / (double)r.in.sample_rate;
        report(c, f, audio, dt, dt);
        if (!c.quiet && r.clipped_samples) std::fprintf(stderr, "%lld samples clipped\n", (long long)r.clipped_samples);
    }
    return failed ? 1 : 0;
}
