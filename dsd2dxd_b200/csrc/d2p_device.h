// d2p_device.h -- structures shared between the host launcher and the sm_100a kernels.
#pragma once
#include <cstdint>

namespace d2p {

// One contiguous run of output frames of one input payload ("file").  All positions are absolute
// per-channel stream positions, so a job can describe a whole file or one streamed block with halo.
struct DevJob {
    const uint8_t* in;        // container-order bytes; byte 0 of the buffer is per-channel position base_q
    long long base_q;         // per-channel byte index of the first byte held in `in` (block-group aligned)
    long long end_q;          // per-channel bytes of the stream available (absolute)
    long long tail_q;         // planar only: position where a ragged final block group starts (else LLONG_MAX)
    long long tail_len;       // per-channel bytes of that group (each channel's run is tail_len long)
    long long f_begin;        // first absolute frame to produce
    long long f_count;        // frames to produce
    uint8_t* out;             // interleaved: where frame f_begin goes; planar32: channel 0 / frame f_begin
    long long out_stride;     // planar32: elements between channels
    double* f64;              // optional pre-scale output, channel-major, frame f_begin at index 0
    long long f64_stride;
    unsigned long long seed;  // dither key
    unsigned int tile_begin;  // first global tile index of this job
    int job_id;
    // Airwindows float dither (-d F): the reference's xorshift32 state is sequential over the samples in reader
    // order (block, channel, frame).  fpd_kernel materialises the state of every sample of the job first.
    uint32_t* fpd;            // [channel * fpd_stride + (frame - f_begin)], or null
    long long fpd_stride;
    long long total_q;        // per-channel bytes of the whole stream if the job reaches its last block, else LLONG_MAX
    long long blk_lo;         // reader block holding frame f_begin
    unsigned int fpd_item_begin;   // first (block, channel) work item of this job; nblk * C items
    int nblk;
    uint32_t fpd0;            // state after Dither::init_rand
    int pad0;
};

struct DevStats {
    unsigned long long peak_bits;   // max |sample * scale| as f64 bits (non-negative doubles order as integers)
    unsigned long long clip_lo, clip_hi;
    unsigned int n_events;          // clip events appended (may exceed capacity: then events were dropped)
    unsigned int pad;
};

struct DevJobStats { unsigned long long peak_bits, clipped; };   // per job_id (device-resident batches)

struct ClipEvent { int job; int ch_kind; long long frame; };   // ch_kind = channel | (1<<16 if high clip)

static const int kMaxPhasesDev = 20;
static const int kMaxFirTapsDev = 176;
static const int kMaxFir3TapsDev = 32;

struct KParams {
    const DevJob* jobs;
    int njobs;
    unsigned int total_tiles;
    const double* lut;        // global copy of the stage-1 LUT (copied to shared memory per CTA)
    int lut_doubles;

    // input format (A1)
    int C, layout, lsb_first, block_size;
    int block_shift;          // log2(block_size) when it is a power of two, else -1
    // writer (A13-A17)
    int bits, bps, out_format, dither, has_ds, peak;
    int shl20;                // 4 for 20-bit output (left-aligned in its 3-byte container), else 0
    double scale, neg_scale, pos_scale;

    int TF;                   // output frames per tile
    int tile_stride;          // shared-memory bytes per channel input tile

    // integer path (A3/A4)
    int h, nt, bpo;
    // skewed, bank-conflict-free LUT layout of the integer path (DESIGN.md "LUT layout"):
    //   entry (replica r, table i, byte b) lives at  b * lut_row_bytes + (lut_st[r] + i) * 8
    // lanes of a half-warp are split into R = 16/lutK replicas x lutK skews; lane (r, c) runs table i at
    // step u = i + c, so the 16 lanes of a half-warp always address 16 distinct 8-byte bank pairs.
    int lutK, lut_row_bytes;
    int lut_st[16];
    int ngroups, gp_lo, gp_hi;   // 4-step groups per frame; [gp_lo, gp_hi) need no activity guard
    int last_steps;              // steps of the last group that any lane still needs (1..4)
    int zpad_bytes;              // zero region in front of the LUT (idle skew steps read 0.0 from it)
    // frame sets of the integer path: a lane runs setF neighbouring frames (setDW tile words apart) in lockstep;
    // TF == setF * SL (SL frame sets per channel tile, a multiple of 32), sl_inv = ceil(2^32 / SL),
    // fpw_shift = log2(frames per 32-bit stream word)
    int setF, setDW, SL, fpw_shift;
    unsigned int sl_inv;
    int nc;                      // consumer threads (the remaining threads of the CTA stage tiles)
    int in_buf_bytes, ostage_bytes;   // per buffer (both are double-buffered)
    // L/M path (A6-A10)
    int nstages, L, m1, accum, h1, gmax, ring_mask;
    long long e0;
    int h2, D2, h3, D3;
    int sigma2, sigma3;       // element shift of the phase-major tiles: -h (mod D), aligns the FIR walkers' wraps
    int fir2_rf, fir2_q, fir3_rf, fir3_q;   // (2h-1+sigma) mod D and div D: row / column of tap 0's front element
    int inv_d2, inv_d3;       // ceil(65536 / D): (x * inv) >> 16 == x / D for the small x the kernels divide
    int s1_cap, s2_cap;       // doubles per channel slot of the stage-1 / stage-2 tiles
    int NB;                   // channel slots kept in flight together
    int pitch2, pitch3;       // doubles per column (one decimation period, padded to an odd count) of the stage-1 / stage-2 tiles
    int lm_cons;              // consumer threads of the lm_kernel instantiation in use (448 or 896)
    int lmA, lmGp, lm_cyc;    // active lanes per half-warp, LUT block pitch (slots, odd), phase patterns (1 or 2)
    long long lm_base_hi;     // high-rate index of stage-1 output 0
    long long lm_step_t;      // input-time step of one half-warp iteration: m1 * A * (threads/16) / L
    int lm_step_q, lm_step_ph; // the same step in stage-1 outputs, as whole stage-2 periods and the remainder
    unsigned char groups[kMaxPhasesDev];

    const uint32_t* fpd_jump; // xorshift32 jump tables: [level 0..3][digit 0..255][byte 0..3][256] (4 MiB), or null
    unsigned int fpd_items;

    DevStats* stats;
    DevJobStats* job_stats;   // optional table indexed by DevJob::job_id
    ClipEvent* events;
    unsigned int ev_cap;

    double c2[kMaxFirTapsDev];   // first h2 entries of reverse(half2): the taps convolve_direct walks
    double c3[kMaxFir3TapsDev];
};

static_assert(sizeof(KParams) <= 4096, "kernel parameter block is limited to 4 KB");

}  // namespace d2p
