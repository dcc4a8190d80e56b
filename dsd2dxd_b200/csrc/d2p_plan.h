// d2p_plan.h -- host-side resolution of a d2p_config into a conversion plan.
//
// Mirrors, as data and closed forms, the reference's set-up logic:
//   compute_decim_and_upsample   rd/lm_resampler.rs:911-969
//   select_precalc_taps          rd/byte_precalc_decimator.rs:167-262
//   LMResampler::new             rd/lm_resampler.rs:87-266   (topology table)
//   Stage1Poly::new              rd/lm_resampler.rs:408-450  (mode, delays, ring capacity)
//   PcmWriter::set_scaling       rd/pcm_writer.rs:288-298
//   calc_frames_cap              rd/lib.rs:259-270
// and turns the streaming state machines into index arithmetic (SURVEY.md 8a closed forms), so that
// any block of output frames can be computed from absolute input positions.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/dsd2pcm_b200.h"

namespace d2p {

struct TapTable { const char* name; const double* data; int len; };
const TapTable* find_table(const char* name);
int table_count();
const TapTable* table_at(int i);

static const int kMaxPhases = 20;     // largest L in the reference topologies (DSD64 -> 384k)
static const int kMaxFirTaps = 176;   // stage-2 half taps: <= 173 (HTAPS_2_68MHZ_28TO1_EQ)
static const int kMaxFir3Taps = 32;   // stage-3 half taps: <= 26

struct Plan {
    d2p_config cfg;
    int M = 0;                 // decim_ratio
    uint32_t L = 1;            // upsample_ratio
    int nstages = 1;
    const TapTable* tab[3] = {nullptr, nullptr, nullptr};
    int sdecim[3] = {0, 0, 0}; // integer path: {decim}; L/M: {m1, D2, D3}

    // integer path (BytePrecalcDecimator)
    int h = 0, nt = 0, bpo = 0;

    // L/M stage 1 (Stage1Poly)
    int h1 = 0, m1 = 0;
    bool accum = false;        // Stage1Mode::Accum (m == 21) vs SlotSim
    long long e0 = 0;          // first emitted event (Accum)
    int groups[kMaxPhases];    // LUT groups per phase
    int gmax = 0;
    int ring_cap = 32;         // byte_ring capacity: window offsets 8g alias modulo this
    // stages 2 / 3 (DecimFIRSym)
    int h2 = 0, D2 = 0, h3 = 0, D3 = 0;

    // writer
    double scale = 1.0, neg_scale = 1.0, pos_scale = 1.0;
    int32_t peak = 0;
    int bps = 3;

    // ---- closed forms ----
    // stage-1 outputs produced from `bits` input bits of one channel
    long long stage1_count(long long bits) const;
    // high-rate index of emitted stage-1 output n
    long long stage1_nhi(long long n) const { return accum ? (e0 + n + 1) * 21 - 1 : (long long)(h1 - 1) + (long long)m1 * (n + 1); }
    // frames per channel emitted for `bytes` input bytes of one channel
    long long frames_for_bytes(long long bytes) const;
    // smallest per-channel byte count whose frames_for_bytes() is >= nframes
    long long bytes_needed_for_frames(long long nframes) const;
    // L/M: how many bits behind its newest input bit a stage-1 output reads (31 for every table in use)
    int lm_reach() const;
    // oldest per-channel byte position (>= 0) that the computation of frame f reads
    long long first_byte_needed(long long f) const;
};

// Resolve cfg; on failure returns a d2p_status and fills msg with the reference-style message.
int make_plan(const d2p_config& cfg, Plan* out, std::string* msg);

// Table builders with the reference's f64 fold order.
//   integer path: lut[i*256 + b], i = 0..nt-1            rd/byte_precalc_decimator.rs:79-100
//   stage 1     : lut[(p*gmax + g)*256 + b]              rd/lm_resampler.rs:847-887 (missing groups = +0.0)
void build_bpd_lut(const Plan& p, std::vector<double>* lut);
void build_stage1_lut(const Plan& p, std::vector<double>* lut);
// full symmetric taps reverse(half) || half, first `h` entries are what convolve_direct walks
void build_full_taps(const TapTable* t, std::vector<double>* full);

}  // namespace d2p
