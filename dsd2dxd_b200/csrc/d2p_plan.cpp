// d2p_plan.cpp -- see d2p_plan.h.  Host only; compiled with FP contraction off.
#include "d2p_plan.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "d2p_taps.inc"

namespace d2p {

#define D2P_ROW(NAME, LEN) {#NAME, D2P_##NAME, LEN},
static const TapTable kTables[] = {D2P_TAP_TABLE_LIST(D2P_ROW)};
#undef D2P_ROW
static const int kNumTables = (int)(sizeof(kTables) / sizeof(kTables[0]));

const TapTable* find_table(const char* name) {
    for (int i = 0; i < kNumTables; ++i)
        if (std::strcmp(kTables[i].name, name) == 0) return &kTables[i];
    return nullptr;
}
int table_count() { return kNumTables; }
const TapTable* table_at(int i) { return (i >= 0 && i < kNumTables) ? &kTables[i] : nullptr; }

static size_t next_pow2(size_t x) { size_t p = 1; while (p < x) p <<= 1; return p; }

// (filter, dsd_rate, decim) -> table name.  Data form of select_precalc_taps.
struct PrecalcRow { int decim; int filter; int dsd_rate; /* 0 = any */ const char* table; };
static const PrecalcRow kPrecalc[] = {
    {128, 'E', 4, "HTAPS_DSD256_128TO1_EQ"},
    {8, 'D', 1, "HTAPS_D2P"}, {8, 'X', 1, "HTAPS_XLD"}, {8, 'E', 1, "HTAPS_DSD64_8TO1_EQ"},
    {8, 'E', 2, "HTAPS_DDR_8TO1_EQ"}, {8, 'E', 4, "HTAPS_DSD256_8TO1_EQ"},
    {16, 'X', 0, "HTAPS_16TO1_XLD"},
    {16, 'E', 1, "HTAPS_DSD64_16TO1_EQ"}, {16, 'E', 2, "HTAPS_DDR_16TO1_EQ"}, {16, 'E', 4, "HTAPS_DSD256_16TO1_EQ"},
    {16, 'C', 2, "HTAPS_DDR_16TO1_CHEB"},
    {32, 'X', 0, "HTAPS_32TO1"},
    {32, 'E', 1, "HTAPS_DSD64_32TO1_EQ"}, {32, 'E', 4, "HTAPS_DSD256_32TO1_EQ"}, {32, 'E', 0, "HTAPS_DDR_32TO1_EQ"},
    {32, 'C', 0, "HTAPS_DDR_32TO1_CHEB"},
    {64, 'E', 4, "HTAPS_DSD256_64TO1_EQ"}, {64, 'E', 8, "HTAPS_DSD512_64TO1_EQ"}, {64, 'E', 0, "HTAPS_DDR_64TO1_EQ"},
    {64, 'C', 0, "HTAPS_DDR_64TO1_CHEB"},
};

static const TapTable* select_precalc(int filter, int dsd_rate, int decim) {
    // rows are ordered so that a specific dsd_rate is met before the "any rate" row of the same key
    for (const PrecalcRow& r : kPrecalc)
        if (r.decim == decim && r.filter == filter && (r.dsd_rate == dsd_rate || r.dsd_rate == 0)) {
            // the 16:1 'E' entry has no catch-all and the 8:1 family none at all: handled by absence of rows
            return find_table(r.table);
        }
    return nullptr;
}

// L/M topology rows: LMResampler::new as a table.  out_rate 0 / l 0 = wildcard, first match wins.
struct LmRow { int m; uint32_t out_rate; uint32_t l; const char* t1; int m1; const char* t2; int d2; const char* t3; int d3; };
static const LmRow kLm[] = {
    {588, 96000, 5, "HTAPS_DDRX10_21TO1_EQ", 21, "HTAPS_2_68MHZ_28TO1_EQ", 28, nullptr, 0},
    {294, 192000, 5, "HTAPS_DDRX10_21TO1_EQ", 21, "HTAPS_2_68MHZ_14TO1_EQ", 14, nullptr, 0},
    {294, 0, 0, "HTAPS_DDRX5_14TO1_EQ", 14, "HTAPS_2MHZ_7TO1_EQ", 7, "HTAPS_288K_3TO1_EQ", 3},
    {147, 384000, 0, "HTAPS_DDRX10_21TO1_EQ", 21, "HTAPS_2_68MHZ_7TO1_EQ", 7, nullptr, 0},
    {147, 192000, 5, "HTAPS_DDRX5_7TO_1_EQ", 7, "HTAPS_4MHZ_7TO1_EQ", 7, "HTAPS_576K_3TO1_EQ", 3},
    {147, 192000, 10, "HTAPS_DSDX10_21TO1_EQ", 21, "HTAPS_1_34MHZ_7TO1_EQ", 7, nullptr, 0},
    {147, 0, 0, "HTAPS_DSDX5_7TO1_EQ", 7, "HTAPS_2MHZ_7TO1_EQ", 7, "HTAPS_288K_3TO1_EQ", 3},
};

static void decim_and_upsample(int in_rate, uint32_t out_rate, int* m_out, uint32_t* l_out) {
    int m;
    if (out_rate == 96000 && in_rate == 4) m = 588;
    else if ((out_rate == 96000 && in_rate == 2) || (out_rate == 192000 && in_rate == 4)) m = 294;
    else if (out_rate == 96000 || out_rate == 192000 || out_rate == 384000) m = 147;
    else {
        const uint32_t base = 2822400u * (uint32_t)in_rate;      // DSD_64_RATE * in_rate
        m = (out_rate > 0 && base % out_rate == 0) ? (int)(base / out_rate) : 64;
    }
    uint32_t l = 5;
    if (m < 147) l = 1;
    else if (out_rate == 384000) l = in_rate == 1 ? 20 : (in_rate == 2 ? 10 : 5);
    else if (out_rate == 192000) l = in_rate == 1 ? 10 : 5;
    *m_out = m;
    *l_out = l;
}

static long long fir_count(long long n_in, int h, int d) { return n_in >= h ? (n_in - h) / d + 1 : 0; }

long long Plan::stage1_count(long long bits) const {
    const long long hi = bits * (long long)L;
    long long n;
    if (accum) n = hi / 21 - 1 - e0 + 1;
    else n = hi >= h1 ? (hi - h1) / m1 : 0;
    return n > 0 ? n : 0;
}

long long Plan::frames_for_bytes(long long bytes) const {
    if (bytes <= 0) return 0;
    if (nstages == 1) {
        long long f = bytes / bpo - (h - 1);
        return f > 0 ? f : 0;
    }
    long long n1 = stage1_count(bytes * 8);
    long long n2 = fir_count(n1, h2, D2);
    return nstages == 3 ? fir_count(n2, h3, D3) : n2;
}

long long Plan::bytes_needed_for_frames(long long nframes) const {
    if (nframes <= 0) return 0;
    // frames_for_bytes is monotone: binary search the smallest byte count giving >= nframes
    long long lo = 0, hi = 1;
    while (frames_for_bytes(hi) < nframes) hi *= 2;
    while (lo + 1 < hi) {
        long long mid = lo + (hi - lo) / 2;
        if (frames_for_bytes(mid) >= nframes) hi = mid; else lo = mid;
    }
    return hi;
}

// Oldest input bit, counted back from the newest one, that a stage-1 output reads: group g looks at the 8-bit window
// ending 8g bits back, taken modulo the reference's byte ring (Stage1Poly, rd/lm_resampler.rs:408-450, 603-640).
int Plan::lm_reach() const {
    int reach = 7;
    for (int g = 0; g < gmax; ++g) reach = std::max(reach, (8 * g) % ring_cap + 7);
    return reach;
}

long long Plan::first_byte_needed(long long f) const {
    long long q;
    if (nstages == 1) {
        q = (f + (h - 1) + 1) * bpo - 2LL * nt;
    } else {
        long long k2 = nstages == 3 ? (h3 - 1) + f * D3 - (2LL * h3 - 1) : f;
        if (k2 < 0) k2 = 0;
        long long n = (h2 - 1) + k2 * D2 - (2LL * h2 - 1);
        if (n < 0) n = 0;
        const long long t = stage1_nhi(n) / (long long)L;
        q = (t - std::max(lm_reach(), 31)) >> 3;                // the kernels always stage a 32-bit window per output
    }
    return q > 0 ? q : 0;
}

int make_plan(const d2p_config& cfg, Plan* out, std::string* msg) {
    Plan p;
    p.cfg = cfg;
    auto fail = [&](int st, const std::string& m) { if (msg) *msg = m; return st; };
    if (!(cfg.dsd_rate == 1 || cfg.dsd_rate == 2 || cfg.dsd_rate == 4 || cfg.dsd_rate == 8))
        return fail(D2P_ERR_INVALID_ARG, "dsd_rate must be 1, 2, 4 or 8");
    if (cfg.channels < 1 || cfg.channels > 16) return fail(D2P_ERR_INVALID_ARG, "channels must be 1..16");
    if (cfg.in_layout != D2P_LAYOUT_PLANAR && cfg.in_layout != D2P_LAYOUT_INTERLEAVED)
        return fail(D2P_ERR_INVALID_ARG, "Invalid format; must be I (interleaved) or P (planar)");
    if (cfg.block_size == 0) return fail(D2P_ERR_INVALID_ARG, "block_size must be > 0");
    if (cfg.block_size > 0x7FFFFFFFu) return fail(D2P_ERR_INVALID_ARG, "block_size must be below 2^31");
    if (!(cfg.dither == 'T' || cfg.dither == 'R' || cfg.dither == 'F' || cfg.dither == 'X'))
        return fail(D2P_ERR_INVALID_ARG, "Invalid dither type; must be T, R, F, or X");
    if (cfg.out_format < 0 || cfg.out_format > 3) return fail(D2P_ERR_INVALID_ARG, "invalid out_format");
    if (!(cfg.bits == 16 || cfg.bits == 20 || cfg.bits == 24 || cfg.bits == 32))
        return fail(D2P_ERR_BIT_DEPTH, "Unsupported bit depth");

    decim_and_upsample(cfg.dsd_rate, cfg.out_rate, &p.M, &p.L);

    if (p.L > 1) {
        const LmRow* row = nullptr;
        for (const LmRow& r : kLm)
            if (r.m == p.M && (r.out_rate == 0 || r.out_rate == cfg.out_rate) && (r.l == 0 || r.l == p.L)) { row = &r; break; }
        if (!row) return fail(D2P_ERR_UNSUPPORTED_LM, p.M == 588 ? "Unsupported 588: must be out_rate=96k and L=5"
                                                                 : "Unsupported L/M combination");
        p.tab[0] = find_table(row->t1); p.sdecim[0] = row->m1;
        p.tab[1] = find_table(row->t2); p.sdecim[1] = row->d2;
        p.nstages = 2;
        if (row->t3) { p.tab[2] = find_table(row->t3); p.sdecim[2] = row->d3; p.nstages = 3; }
        p.h1 = p.tab[0]->len; p.m1 = row->m1;
        p.h2 = p.tab[1]->len; p.D2 = row->d2;
        if (p.nstages == 3) { p.h3 = p.tab[2]->len; p.D3 = row->d3; }
        if (p.h2 > kMaxFirTaps || p.h3 > kMaxFir3Taps || (int)p.L > kMaxPhases)
            return fail(D2P_ERR_INVALID_ARG, "internal: tap table larger than kernel parameter space");
        p.accum = (p.m1 == 21);
        const long long n = 2LL * p.h1;
        const long long delay_high = (n - 1) / 2;
        const long long input_delay = (delay_high + (p.L - 1)) / p.L;
        p.e0 = 0;
        if (p.accum)
            while (((p.e0 + 1) * 21 + p.L - 1) / p.L < input_delay) ++p.e0;
        p.gmax = 0;
        for (uint32_t ph = 0; ph < p.L; ++ph) {
            long long taps = (n - ph + p.L - 1) / p.L;     // entries of full[ph::L]
            p.groups[ph] = (int)((taps + 7) / 8);
            if (p.groups[ph] > p.gmax) p.gmax = p.groups[ph];
        }
        size_t cap = next_pow2((size_t)p.gmax + 16);
        p.ring_cap = (int)(cap < 32 ? 32 : cap);
    } else {
        const TapTable* t = select_precalc(cfg.filter, cfg.dsd_rate, p.M);
        if (!t || p.M % 8 != 0) {
            char buf[200];
            const char* fname = cfg.filter == 'E' ? "Equiripple" : cfg.filter == 'X' ? "XLD" : cfg.filter == 'D' ? "Dsd2Pcm"
                              : cfg.filter == 'C' ? "Chebyshev" : "?";                 // FilterType's Debug names
            std::snprintf(buf, sizeof buf, "Taps not found for decimation ratio %d / filter '%s' (dsd_rate %d). ",
                          p.M, fname, cfg.dsd_rate);
            return fail(D2P_ERR_TAPS_NOT_FOUND, buf);
        }
        p.tab[0] = t; p.sdecim[0] = p.M; p.nstages = 1;
        p.h = t->len; p.nt = (p.h + 7) / 8; p.bpo = p.M / 8;
        // process_bytes restarts its byte-in-frame counter on every call (rd/byte_precalc_decimator.rs:121-123): with a
        // block that is not a whole number of output frames the reference drops the partial frame of every block and
        // its frame phase drifts.  The closed forms here assume continuous frames, so such a block size is refused
        // rather than converted differently (4096 and every power of two >= 16 are fine).
        if (cfg.block_size % (uint32_t)p.bpo != 0) {
            char buf[160];
            std::snprintf(buf, sizeof buf, "block_size %u is not a multiple of the %d input bytes per output frame", cfg.block_size, p.bpo);
            return fail(D2P_ERR_INVALID_ARG, buf);
        }
    }

    // set_scaling  rd/pcm_writer.rs:288-298
    double scale = 1.0;
    const double vol_scale = std::pow(10.0, cfg.level_db / 20.0);
    if (cfg.bits != 32) scale = std::ldexp(1.0, cfg.bits - 1);
    p.peak = (int32_t)std::floor(scale);
    scale *= vol_scale;
    scale *= (double)p.L;
    p.scale = scale;
    p.bps = cfg.bits == 20 ? 3 : cfg.bits / 8;
    if (cfg.has_dither_scale) {                                  // rd/dither.rs:39-49
        p.neg_scale = std::pow(10.0, -cfg.dither_scale_db / 20.0);
        p.pos_scale = std::pow(10.0, cfg.dither_scale_db / 20.0);
    }
    *out = p;
    return D2P_OK;
}

void build_bpd_lut(const Plan& p, std::vector<double>* lut) {
    const double* half = p.tab[0]->data;
    lut->assign((size_t)p.nt * 256, 0.0);
    for (int i = 0; i < p.nt; ++i) {
        const int t = (p.nt - 1 - i) * 8;                        // tables are stored last-segment-first
        const int nbits = p.h - t < 8 ? p.h - t : 8;
        for (int b = 0; b < 256; ++b) {
            double acc = 0.0;
            for (int bit = 0; bit < nbits; ++bit) {
                const double sign = (double)(((b >> bit) & 1) * 2 - 1);
                acc = acc + sign * half[t + bit];
            }
            (*lut)[(size_t)i * 256 + b] = acc;
        }
    }
}

void build_full_taps(const TapTable* t, std::vector<double>* full) {
    full->resize((size_t)t->len * 2);
    for (int i = 0; i < t->len; ++i) {
        (*full)[i] = t->data[t->len - 1 - i];
        (*full)[t->len + i] = t->data[i];
    }
}

void build_stage1_lut(const Plan& p, std::vector<double>* lut) {
    std::vector<double> full;
    build_full_taps(p.tab[0], &full);
    lut->assign((size_t)p.L * p.gmax * 256, 0.0);
    for (uint32_t ph = 0; ph < p.L; ++ph) {
        std::vector<double> taps;
        for (size_t i = ph; i < full.size(); i += p.L) taps.push_back(full[i]);
        for (int g = 0; g < p.groups[ph]; ++g) {
            double coeffs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int j = 0; j < 8; ++j)
                if ((size_t)(g * 8 + j) < taps.size()) coeffs[j] = taps[g * 8 + j];
            for (int b = 0; b < 256; ++b) {
                double sum = 0.0;
                for (int j = 0; j < 8; ++j) {
                    const double sign = ((b >> j) & 1) ? 1.0 : -1.0;
                    sum += coeffs[j] * sign;
                }
                (*lut)[((size_t)ph * p.gmax + g) * 256 + b] = sum;
            }
        }
    }
}

}  // namespace d2p
