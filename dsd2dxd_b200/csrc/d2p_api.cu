// d2p_api.cu -- C ABI of libdsd2pcm_b200.so (include/dsd2pcm_b200.h) on top of the sm_100a kernels.
//
// Host-side responsibilities, all stateless with respect to the signal (every output frame is a pure
// function of absolute input positions -- d2p_plan.h):
//   * resolve the configuration (Plan) and keep the filter LUTs resident in HBM per device,
//   * cut the requested frame range into tiles and describe payloads as jobs (DevJob),
//   * streaming: keep the filter halo of already-pushed input on the device between pushes,
//   * whole-payload: pipeline H2D / kernel / D2H over three streams,
//   * replay the reference's clip counter over the sparse clip events the kernels report.
// There is no CPU compute path: without a CUDA device the compute entries fail with D2P_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dsd2pcm_b200.h"
#include "d2p_launch.h"
#include "d2p_plan.h"

namespace d2p {

static thread_local std::string g_last_error;
static int set_err(int st, const std::string& m) { g_last_error = m; return st; }
static int cuda_err(cudaError_t e, const char* what) {
    char buf[256];
    std::snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
    return set_err((e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? D2P_ERR_NO_DEVICE : D2P_ERR_CUDA, buf);
}
int api_set_error(int st, const char* msg) { return set_err(st, msg ? msg : ""); }   // for d2p_files.cpp
#define CK(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return cuda_err(e__, #call); } while (0)

static const unsigned int kEventCap = 1u << 16;

// ------------------------------------------------------------------------------------------------
// DevicePlan: Plan + device-resident tables + tile geometry.  Cached per (device, config key).
// ------------------------------------------------------------------------------------------------
struct DevicePlan {
    Plan plan;
    bool lm = false;
    double* d_lut = nullptr;
    KParams kp;              // template: jobs / stats filled per launch
    int threads = kMaxThreads;
    size_t smem = 0;
    int ctas_per_sm = 1, num_sms = 0;
    bool fast_writer = false;  // plan-level conditions of the specialised 3-byte-LE writer (Emitter<1, .>)
    ~DevicePlan() { if (d_lut) cudaFree(d_lut); }
};

static size_t round16(size_t x) { return (x + 15) & ~(size_t)15; }

// Bank-aware LUT layout of the integer path (see d2p_kernels.cu): R = 16/K replicas of the nt tables in
// one row per byte value; replica r starts at slot st[r] with st[r] = 0 (mod K) and st[r]/K distinct
// modulo R, so that lanes (r, c) at table u - c always fall into 16 different 8-byte bank pairs.
struct LutLayout { int K = 16, R = 1, slots = 0; int st[16] = {0}; };

static LutLayout make_lut_layout(int nt, int K) {
    LutLayout L;
    L.K = K; L.R = 16 / K;
    bool used[16] = {false};
    int next = 0;
    for (int r = 0; r < L.R; ++r) {
        int s = (next + K - 1) / K * K;
        while (used[(s / K) % L.R]) s += K;
        used[(s / K) % L.R] = true;
        L.st[r] = s;
        next = s + nt;
    }
    L.slots = (next + 15) / 16 * 16;
    return L;
}

// Zero region in front of the LUT: idle skew steps of a lane read 0.0 at offsets [-15, 15 + 4*ngroups) slots
// around its first 128 bytes (see LaneLut in d2p_kernels.cu).
static int lutfir_zpad(int nt, int K) {
    const int ng = (nt + K - 1 + 3) / 4;
    return (128 + (15 + 4 * ng) * 8 + 127) / 128 * 128;
}

static void lutfir_geometry(const Plan& p, const LutLayout& L, int TF, int* tile_stride, int* in_buf, int* ostage, size_t* smem) {
    const int C = p.cfg.channels;
    *tile_stride = (int)round16((size_t)TF * p.bpo + 2 * p.nt + 80) + 16;
    *in_buf = C * *tile_stride;
    *ostage = (p.cfg.out_format == D2P_OUT_PLANAR_32 || p.cfg.out_format == D2P_OUT_NONE) ? 0 : (int)round16((size_t)TF * C * p.bps);
    *smem = (size_t)lutfir_zpad(p.nt, L.K) + (size_t)L.slots * 2048 + 2 * (size_t)*in_buf + 2 * (size_t)*ostage;
}

// L/M tile geometry (see lm_kernel): NB channel slots of phase-major stage-1 / stage-2 tiles.
struct LmGeom { int tile_stride, s1_cap, s2_cap, pitch2, pitch3, ostage; size_t smem; };
struct LmLut { int A, Gp, cyc, slots; };       // active lanes per half-warp, block pitch, phase patterns, row slots

static LmLut lm_lut_layout(const Plan& p) {
    LmLut l;
    const int L = (int)p.L;
    l.A = L <= 16 ? L * (16 / L) : 10;          // L = 5 -> 15 lanes, 10 -> 10, 20 -> 10 (two alternating phase patterns)
    l.cyc = (p.m1 * l.A) % L == 0 ? 1 : 2;
    l.Gp = p.gmax | 1;
    l.slots = (l.A * l.cyc * l.Gp + 15) / 16 * 16;
    return l;
}

// Stage tiles are stored column by column: a column holds the D samples of one decimation period in Dp = D | 1
// doubles (odd), so a tap walk is contiguous and the columns of neighbouring outputs fall into different bank pairs.
static int lm_col_pitch(int D) { return D | 1; }

static LmGeom lm_geometry(const Plan& p, const LmLut& ll, int TF, int NB) {
    LmGeom g;
    const int C = p.cfg.channels;
    const long long k2span = p.nstages == 3 ? (long long)(TF - 1) * p.D3 + (2 * p.h3 - 1) : (TF - 1);
    const long long cnt1 = k2span * p.D2 + (2 * p.h2 - 1) + 1;
    g.pitch2 = lm_col_pitch(p.D2);
    g.s1_cap = (int)((((cnt1 + 2 * p.D2 - 1) / p.D2 + 2) * g.pitch2 + 1) & ~1LL);     // + one column for the sigma shift
    g.pitch3 = 0; g.s2_cap = 0;
    if (p.nstages == 3) {
        g.pitch3 = lm_col_pitch(p.D3);
        g.s2_cap = (int)((((k2span + 1 + 2 * p.D3 - 1) / p.D3 + 2) * g.pitch3 + 1) & ~1LL);
    }
    const long long bits = cnt1 * p.m1 / p.L + 64 + p.m1;
    g.tile_stride = (int)round16((size_t)(bits / 8) + 16) + 48;
    const size_t ostage = (p.cfg.out_format == D2P_OUT_PLANAR_32 || p.cfg.out_format == D2P_OUT_NONE) ? 0 : round16((size_t)TF * C * p.bps);
    g.ostage = (int)ostage;
    g.smem = (size_t)ll.slots * 2048 + (size_t)NB * (g.s1_cap + g.s2_cap) * 8 + 2 * (size_t)C * g.tile_stride + 2 * ostage;
    return g;
}

// host_lut: when given, the LUT (in its shared-memory layout) is handed back in host memory instead of being placed in
// a device allocation of its own, and kp.lut stays null for the caller to fill (d2p_run_noalloc).
static int build_device_plan(const d2p_config& cfg, DevicePlan** out, std::vector<double>* host_lut = nullptr) {
    std::string msg;
    std::unique_ptr<DevicePlan> dp(new DevicePlan());
    int st = make_plan(cfg, &dp->plan, &msg);
    if (st != D2P_OK) return set_err(st, msg);
    const Plan& p = dp->plan;
    dp->lm = p.nstages > 1;

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return set_err(D2P_ERR_NO_DEVICE, "no CUDA device available (this library has no CPU path)");
    if (cfg.device < 0 || cfg.device >= ndev) return set_err(D2P_ERR_INVALID_ARG, "invalid device ordinal");
    CK(cudaSetDevice(cfg.device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg.device));
    dp->num_sms = prop.multiProcessorCount;

    std::vector<double> lut;
    if (dp->lm) build_stage1_lut(p, &lut); else build_bpd_lut(p, &lut);

    KParams& k = dp->kp;
    std::memset(&k, 0, sizeof k);
    k.C = cfg.channels; k.layout = cfg.in_layout; k.lsb_first = cfg.lsb_first ? 1 : 0; k.block_size = (int)cfg.block_size;
    k.block_shift = -1;
    if ((cfg.block_size & (cfg.block_size - 1)) == 0) { k.block_shift = 0; while ((1u << k.block_shift) < cfg.block_size) ++k.block_shift; }
    k.bits = cfg.bits; k.bps = p.bps; k.out_format = cfg.out_format; k.dither = cfg.dither;
    k.has_ds = cfg.has_dither_scale ? 1 : 0; k.peak = p.peak; k.shl20 = cfg.bits == 20 ? 4 : 0;
    k.scale = p.scale; k.neg_scale = p.neg_scale; k.pos_scale = p.pos_scale;
    k.h = p.h; k.nt = p.nt; k.bpo = p.bpo;
    k.nstages = p.nstages; k.L = (int)p.L; k.m1 = p.m1; k.accum = p.accum ? 1 : 0; k.h1 = p.h1; k.gmax = p.gmax;
    k.ring_mask = p.ring_cap - 1; k.e0 = p.e0;
    k.h2 = p.h2; k.D2 = p.D2; k.h3 = p.h3; k.D3 = p.D3;
    for (uint32_t ph = 0; ph < p.L && ph < (uint32_t)kMaxPhasesDev; ++ph) k.groups[ph] = (unsigned char)p.groups[ph];
    if (dp->lm) {
        std::vector<double> full;
        build_full_taps(p.tab[1], &full);
        for (int i = 0; i < p.h2; ++i) k.c2[i] = full[i];
        if (p.nstages == 3) {
            build_full_taps(p.tab[2], &full);
            for (int i = 0; i < p.h3; ++i) k.c3[i] = full[i];
        }
    }

    const size_t budget = (size_t)kMaxSmem - 1024;
    if (dp->lm) {
        if (p.gmax > 5 || p.ring_cap != 32 || p.lm_reach() > 31 || (int)p.L > kMaxPhasesDev)   // lm_tile_header's window
            return set_err(D2P_ERR_INVALID_ARG, "internal: stage-1 filter outside the kernel's group / ring assumptions");
        const LmLut ll = lm_lut_layout(p);
        { int a = p.m1 % (int)p.L, b = (int)p.L; while (b) { const int t = a % b; a = b; b = t; }
          if (a != 1) return set_err(D2P_ERR_INVALID_ARG, "internal: m1 and L are not coprime"); }
        // Tile geometry per build (448 / 896 consumers, d2p_launch.h): keep the whole CTA busy in stages 2/3 (NB * TF a
        // multiple of the consumer count) with the largest tile that fits.
        const int C = cfg.channels;
        const double halo = p.nstages == 3 ? (2.0 * p.h3 / p.D3 + 2.0 * p.h2 / (p.D2 * p.D3)) : 2.0 * p.h2 / p.D2;
        struct Pick { int cons, tf, nb; double score; };
        auto pick = [&](int cons_) {
            Pick b{cons_, 0, 0, 0.0};
            const int nhw_ = cons_ / 16;
            if (((long long)p.m1 * ll.A * nhw_) % p.L != 0 || (ll.cyc == 2 && ((long long)p.m1 * ll.A * 2) % p.L != 0)) return b;
            for (int tf = 1024; tf >= 32; tf -= 32)
                for (int nb = 1; nb <= C; ++nb) {
                    if (C % nb != 0 && nb != C) continue;
                    const LmGeom g = lm_geometry(p, ll, tf, nb);
                    if (g.smem > budget) continue;
                    // stages 2/3 hand one output to each consumer thread per round: count the idle threads of the last round
                    const double items = (double)nb * tf;
                    const double busy = items / (std::ceil(items / cons_) * cons_);
                    const double score = busy * tf / (tf + halo);
                    if (score > b.score * 1.0001) { b.score = score; b.tf = tf; b.nb = nb; }
                }
            return b;
        };
        const Pick small = pick(kLmConsumersSmall), large = pick(kLmConsumers);
        // Which build (measured over all twelve topologies, tools/lm_cons_sweep.sh): the three-stage chains and the short
        // two-stage filters (stage 2 decimating by 7) run best on 896 consumers (64 registers) -- the latter only while
        // their tiles keep that many threads busy (6-channel input does not) -- the long two-stage filters walked in whole
        // columns of 14 or 28 taps (DSD256 -> 96 / 192 kHz) want the 128 registers of the 448-consumer build.
        Pick best = small;
        if (p.nstages == 3) best = large;
        else if (p.D2 < 14 && large.tf && large.score >= 0.9 * small.score) best = large;
        if (!best.tf) best = small.tf ? small : large;
        if (const char* e = std::getenv("D2P_LM_CONS")) {                  // tuning knob
            const Pick forced = std::atoi(e) > 448 ? large : small;
            if (forced.tf) best = forced;
        }
        const int cons = best.cons, nhw = cons / 16;
        const int best_tf = best.tf, best_nb = best.nb;
        if (!best_tf) return set_err(D2P_ERR_NOMEM, "filter LUT and tiles do not fit in shared memory");
        const LmGeom g = lm_geometry(p, ll, best_tf, best_nb);
        k.TF = best_tf; k.NB = best_nb;
        k.tile_stride = g.tile_stride; k.s1_cap = g.s1_cap; k.s2_cap = g.s2_cap; k.pitch2 = g.pitch2; k.pitch3 = g.pitch3;
        dp->smem = g.smem;
        k.lmA = ll.A; k.lmGp = ll.Gp; k.lm_cyc = ll.cyc; k.lut_row_bytes = ll.slots * 8;
        k.lm_base_hi = p.stage1_nhi(0);
        k.lm_step_t = (long long)p.m1 * ll.A * nhw / p.L;
        k.lm_step_q = (ll.A * nhw) / p.D2; k.lm_step_ph = (ll.A * nhw) % p.D2;
        k.ostage_bytes = g.ostage;
        dp->threads = cons + kLmProducers;
        k.lm_cons = cons;
        k.sigma2 = (p.D2 - p.h2 % p.D2) % p.D2;
        k.sigma3 = p.nstages == 3 ? (p.D3 - p.h3 % p.D3) % p.D3 : 0;
        k.fir2_rf = (2 * p.h2 - 1 + k.sigma2) % p.D2; k.fir2_q = (2 * p.h2 - 1 + k.sigma2) / p.D2;
        k.fir3_rf = p.nstages == 3 ? (2 * p.h3 - 1 + k.sigma3) % p.D3 : 0;
        k.fir3_q = p.nstages == 3 ? (2 * p.h3 - 1 + k.sigma3) / p.D3 : 0;
        k.inv_d2 = (65536 + p.D2 - 1) / p.D2;
        k.inv_d3 = p.nstages == 3 ? (65536 + p.D3 - 1) / p.D3 : 0;
        // row per byte value; block (u + A*cyc) holds the tables of phase (m1 * (A*cyc + u)) mod L
        std::vector<double> sk((size_t)256 * ll.slots, 0.0);
        for (int b = 0; b < 256; ++b)
            for (int cy = 0; cy < ll.cyc; ++cy)
                for (int u = 0; u < ll.A; ++u) {
                    const int ph = (int)(((long long)p.m1 * (ll.A * cy + u)) % p.L);
                    for (int gq = 0; gq < p.gmax; ++gq)
                        sk[(size_t)b * ll.slots + (size_t)ll.Gp * (u + ll.A * cy) + gq] = lut[((size_t)ph * p.gmax + gq) * 256 + b];
                }
        if (!host_lut) {
            CK(cudaMalloc(&dp->d_lut, sk.size() * sizeof(double)));
            CK(cudaMemcpy(dp->d_lut, sk.data(), sk.size() * sizeof(double), cudaMemcpyHostToDevice));
        }
        k.lut = dp->d_lut;
        k.lut_doubles = (int)sk.size();
        if (host_lut) host_lut->swap(sk);
    } else {
        // integer path: pick the replication (skew depth K) and the tile size together.  More replicas mean
        // fewer idle skew steps per frame (nt / (nt + K - 1)) but a larger LUT; one persistent CTA per SM.
        double best_score = 0.0; LutLayout bestL; int best_tf = 0;
        int setF = 1, setDW = 1;
        lutfir_set_shape(p.bpo, &setF, &setDW);                   // a lane runs setF neighbouring frames in lockstep
        for (int K = 1; K <= 16; K <<= 1) {
            if (K > 1 && K - 1 > p.nt) continue;
            const LutLayout L = make_lut_layout(p.nt, K);
            for (int sl = 1024; sl >= 32; sl -= 32) {              // frame sets per channel tile: any multiple of a warp
                const int tf = setF * sl;
                int ts, ib, os; size_t sm;
                lutfir_geometry(p, L, tf, &ts, &ib, &os, &sm);
                if (sm > budget) continue;
                const double steps = 4.0 * ((p.nt + K - 1 + 3) / 4);
                const double spt = (double)sl * cfg.channels / kMaxThreads;     // frame sets per consumer thread per tile
                const double busy = spt / std::ceil(spt);                       // idle threads of the last round
                const double score = (p.nt / steps) * busy * (spt / (spt + 0.12));
                if (score > best_score) { best_score = score; bestL = L; best_tf = tf; }
            }
        }
        if (!best_tf) return set_err(D2P_ERR_NOMEM, "filter LUT and tiles do not fit in shared memory");
        const LutLayout& L = bestL;
        k.TF = best_tf;
        k.setF = setF; k.setDW = setDW;
        k.SL = best_tf / setF;
        k.sl_inv = (unsigned int)((0x100000000ull + (unsigned)k.SL - 1) / (unsigned)k.SL);
        k.fpw_shift = p.bpo >= 4 ? 0 : (p.bpo == 2 ? 1 : 2);
        lutfir_geometry(p, L, best_tf, &k.tile_stride, &k.in_buf_bytes, &k.ostage_bytes, &dp->smem);
        k.lutK = L.K; k.lut_row_bytes = L.slots * 8;
        for (int r = 0; r < 16; ++r) k.lut_st[r] = r < L.R ? L.st[r] : 0;
        k.ngroups = (p.nt + L.K - 1 + 3) / 4;
        k.last_steps = p.nt + L.K - 1 - 4 * (k.ngroups - 1);
        k.zpad_bytes = lutfir_zpad(p.nt, L.K);
        k.gp_lo = (L.K - 1 + 3) / 4;
        k.gp_hi = p.nt / 4;
        if (k.gp_hi < k.gp_lo) k.gp_hi = k.gp_lo;
        if (k.gp_hi > k.ngroups) k.gp_hi = k.ngroups;
        k.nc = kMaxThreads;
        dp->threads = kLutfirThreads;
        std::vector<double> sk((size_t)256 * L.slots, 0.0);
        for (int b = 0; b < 256; ++b)
            for (int r = 0; r < L.R; ++r)
                for (int i = 0; i < p.nt; ++i)
                    sk[(size_t)b * L.slots + L.st[r] + i] = lut[(size_t)i * 256 + b];
        if (!host_lut) {
            CK(cudaMalloc(&dp->d_lut, sk.size() * sizeof(double)));
            CK(cudaMemcpy(dp->d_lut, sk.data(), sk.size() * sizeof(double), cudaMemcpyHostToDevice));
        }
        k.lut = dp->d_lut;
        k.lut_doubles = (int)sk.size();
        if (host_lut) host_lut->swap(sk);
    }
    if (!dp->lm) {
        // |y| <= sum |H| for a +-1 input, so |y*scale| + dither + rounding stays far inside i32 for every sane level
        double sum_abs = 0.0;
        for (int i = 0; i < p.tab[0]->len; ++i) sum_abs += 2.0 * std::fabs(p.tab[0]->data[i]);
        dp->fast_writer = p.bps == 3 && cfg.bits != 32 && cfg.out_format == D2P_OUT_INTERLEAVED_LE && !cfg.has_dither_scale &&
                          !cfg.keep_f64 && (cfg.dither == 'X' || cfg.dither == 'T' || cfg.dither == 'R') &&
                          sum_abs * std::fabs(p.scale) < 2147483000.0;
    }
    if (std::getenv("D2P_DEBUG_GEOM"))
        std::fprintf(stderr, "[d2p] %s: tile %d frames, channel slots in flight %d, K %d, LUT %d B/row x 256, smem %zu B, %d threads\n",
                     dp->lm ? "lm_kernel" : "lutfir_kernel", k.TF, k.NB, k.lutK, k.lut_row_bytes, dp->smem, dp->threads);
    dp->ctas_per_sm = max_ctas_per_sm(dp->lm, dp->threads, dp->smem);
    if (dp->ctas_per_sm < 1) return set_err(D2P_ERR_CUDA, "kernel cannot be resident with the chosen tile geometry");
    *out = dp.release();
    return D2P_OK;
}

static void free_device_plan(DevicePlan* dp) { delete dp; }

// ------------------------------------------------------------------------------------------------
// Dither F support: Dither::init_rand seed and the xorshift32 jump tables used by fpd_kernel.
// ------------------------------------------------------------------------------------------------
// init_rand (rd/dither.rs:71-76) draws u32s until one is >= 16386; the draws come from Philox4x32-10 with
// counter (k, 0, 0, 0xF9D) and the 64-bit seed as key -- the same stand-in for thread_rng as oracle/oracle.cpp.
static uint32_t fpd_init_state(uint64_t seed) {
    uint32_t fpd = 1;
    for (uint32_t k = 0; fpd < 16386; ++k) {
        uint32_t c0 = k, c1 = 0, c2 = 0, c3 = 0xF9Du, k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
        for (int r = 0; r < 10; ++r) {
            const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
            const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
            c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        fpd = c0;
    }
    return fpd;
}

struct Gf2Mat { uint32_t col[32]; };
static uint32_t gf2_apply(const Gf2Mat& m, uint32_t v) {
    uint32_t r = 0;
    for (int i = 0; v; ++i, v >>= 1) if (v & 1u) r ^= m.col[i];
    return r;
}
static Gf2Mat gf2_mul(const Gf2Mat& a, const Gf2Mat& b) {          // a after b
    Gf2Mat r;
    for (int i = 0; i < 32; ++i) r.col[i] = gf2_apply(a, b.col[i]);
    return r;
}
// [level][digit][byte][256]: M^(digit * 256^level) applied to (w << 8*byte); M = one xorshift32 (13, 17, 5) step
static const std::vector<uint32_t>& fpd_jump_tables() {
    static std::vector<uint32_t> T;
    static std::once_flag once;
    std::call_once(once, [] {
        T.resize((size_t)4 * 256 * 4 * 256);
        Gf2Mat base;
        for (int i = 0; i < 32; ++i) { uint32_t x = 1u << i; x ^= x << 13; x ^= x >> 17; x ^= x << 5; base.col[i] = x; }
        for (int lv = 0; lv < 4; ++lv) {
            Gf2Mat pw;
            for (int i = 0; i < 32; ++i) pw.col[i] = 1u << i;       // identity
            for (int dg = 0; dg < 256; ++dg) {
                uint32_t* out = &T[((size_t)lv * 256 + dg) << 10];
                for (int by = 0; by < 4; ++by)
                    for (uint32_t w = 0; w < 256; ++w) out[by * 256 + w] = gf2_apply(pw, w << (8 * by));
                pw = gf2_mul(base, pw);
            }
            base = pw;                                              // base^256
        }
    });
    return T;
}
static int fpd_jump_device(int device, const uint32_t** out) {
    static std::mutex mu;
    static std::map<int, uint32_t*> per_dev;
    std::lock_guard<std::mutex> lk(mu);
    auto it = per_dev.find(device);
    if (it == per_dev.end()) {
        const std::vector<uint32_t>& T = fpd_jump_tables();
        uint32_t* d = nullptr;
        CK(cudaMalloc(&d, T.size() * sizeof(uint32_t)));
        CK(cudaMemcpy(d, T.data(), T.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        it = per_dev.emplace(device, d).first;
    }
    *out = it->second;
    return D2P_OK;
}

// ------------------------------------------------------------------------------------------------
// Launcher: uploads the job table and starts one persistent kernel over all tiles of all jobs.
// ------------------------------------------------------------------------------------------------
struct Launcher {
    static const int kRing = 4;
    DevJob* h_jobs[kRing] = {nullptr, nullptr, nullptr, nullptr};   // pinned staging ring
    DevJob* d_jobs[kRing] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev[kRing] = {nullptr, nullptr, nullptr, nullptr};
    int cap[kRing] = {0, 0, 0, 0};
    int next = 0;
    DevStats* d_stats = nullptr;
    ClipEvent* d_events = nullptr;
    uint32_t* d_fpd[kRing] = {nullptr, nullptr, nullptr, nullptr};   // dither F: per-sample xorshift32 states of a launch
    size_t fpd_cap[kRing] = {0, 0, 0, 0};
    uint64_t launches = 0;
    bool want_jobstats = false;                                      // batch objects: peak / clipped samples per job
    bool use_jobstats = false;                                       // ... for the run in progress
    DevJobStats* d_jobstats = nullptr;
    int jobstats_cap = 0, jobstats_n = 0;

    int init(bool want_events) {
        CK(cudaMalloc(&d_stats, sizeof(DevStats)));
        CK(cudaMemset(d_stats, 0, sizeof(DevStats)));
        if (want_events) CK(cudaMalloc(&d_events, sizeof(ClipEvent) * kEventCap));
        for (int i = 0; i < kRing; ++i) CK(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
        return D2P_OK;
    }
    void destroy() {
        for (int i = 0; i < kRing; ++i) {
            if (h_jobs[i]) cudaFreeHost(h_jobs[i]);
            if (d_jobs[i]) cudaFree(d_jobs[i]);
            if (ev[i]) cudaEventDestroy(ev[i]);
            if (d_fpd[i]) cudaFree(d_fpd[i]);
        }
        if (d_stats) cudaFree(d_stats);
        if (d_events) cudaFree(d_events);
        if (d_jobstats) cudaFree(d_jobstats);
    }

    // Dither F needs 4 bytes of xorshift state per output sample of a launch: long jobs are cut into pieces of at
    // most kFpdSamples samples, launched one group after the other on the same stream (the buffer is reused).
    static const long long kFpdSamples = 256ll << 20;
    // per_job: keep the per-job table for this run (d2p_batch_run / d2p_batch_run_blocks).  The pipelined host
    // conversion drives one launcher on three streams and must not: a memset of the shared table on one stream would
    // race with the kernels of the other two, and nobody reads per-job figures of pipeline chunks.
    int run(const DevicePlan& dp, std::vector<DevJob>& jobs, cudaStream_t stream, bool per_job = true) {
        for (size_t i = 0; i < jobs.size(); ++i) jobs[i].job_id = (int)i;
        use_jobstats = want_jobstats && per_job;
        if (!use_jobstats) jobstats_n = 0;                           // a later d2p_batch_job_stats fails cleanly
        if (use_jobstats) {                                          // table of the most recent run, indexed by job
            if (jobstats_cap < (int)jobs.size()) {
                if (d_jobstats) cudaFree(d_jobstats);
                jobstats_cap = (int)jobs.size() + 16;
                CK(cudaMalloc(&d_jobstats, sizeof(DevJobStats) * jobstats_cap));
            }
            jobstats_n = (int)jobs.size();
            if (jobstats_n) CK(cudaMemsetAsync(d_jobstats, 0, sizeof(DevJobStats) * jobstats_n, stream));
        }
        if (dp.plan.cfg.dither != D2P_DITHER_FPD) return run_once(dp, jobs, stream, false);
        const d2p_config& c = dp.plan.cfg;
        const long long cap_frames = std::max<long long>(kFpdSamples / c.channels, dp.kp.TF);
        const bool planar_out = c.out_format == D2P_OUT_PLANAR_32;
        std::vector<DevJob> group;
        long long used = 0;
        for (const DevJob& j : jobs) {
            for (long long off = 0; off < j.f_count;) {
                const long long take = std::min(j.f_count - off, cap_frames - used);
                DevJob pc = j;
                pc.f_begin = j.f_begin + off; pc.f_count = take;
                pc.out = j.out + (size_t)off * (planar_out ? 4 : (size_t)c.channels * dp.plan.bps);
                if (j.f64) pc.f64 = j.f64 + off;
                group.push_back(pc);
                off += take; used += take;
                if (used >= cap_frames) {
                    int st = run_once(dp, group, stream, true);
                    if (st != D2P_OK) return st;
                    group.clear(); used = 0;
                }
            }
        }
        return group.empty() ? D2P_OK : run_once(dp, group, stream, true);
    }

    // jobs[i].tile_begin is filled here.  Asynchronous on `stream`.
    int run_once(const DevicePlan& dp, std::vector<DevJob>& jobs, cudaStream_t stream, bool fpd) {
        unsigned long long tiles = 0;
        std::vector<DevJob> live;
        for (size_t i = 0; i < jobs.size(); ++i) {
            if (jobs[i].f_count <= 0) continue;
            jobs[i].tile_begin = (unsigned int)tiles;
            tiles += (unsigned long long)((jobs[i].f_count + dp.kp.TF - 1) / dp.kp.TF);
            live.push_back(jobs[i]);
        }
        if (tiles == 0) return D2P_OK;
        if (tiles > 0xFFFFFFF0ull) return set_err(D2P_ERR_INVALID_ARG, "too many tiles in one launch");
        const int slot = next; next = (next + 1) % kRing;
        CK(cudaEventSynchronize(ev[slot]));                      // staging slot free again?
        if (cap[slot] < (int)live.size()) {
            if (h_jobs[slot]) cudaFreeHost(h_jobs[slot]);
            if (d_jobs[slot]) cudaFree(d_jobs[slot]);
            cap[slot] = (int)live.size() + 16;
            CK(cudaMallocHost(&h_jobs[slot], sizeof(DevJob) * cap[slot]));
            CK(cudaMalloc(&d_jobs[slot], sizeof(DevJob) * cap[slot]));
        }
        KParams k = dp.kp;
        k.fpd_jump = nullptr; k.fpd_items = 0;
        if (fpd) {
            const Plan& p = dp.plan;
            const long long B = p.cfg.block_size;
            size_t samples = 0;
            unsigned long long items = 0;
            for (DevJob& j : live) samples += (size_t)j.f_count * p.cfg.channels;
            if (fpd_cap[slot] < samples) {
                if (d_fpd[slot]) cudaFree(d_fpd[slot]);
                fpd_cap[slot] = samples + samples / 8 + 1024;
                CK(cudaMalloc(&d_fpd[slot], fpd_cap[slot] * sizeof(uint32_t)));
            }
            size_t at = 0;
            for (DevJob& j : live) {
                j.fpd = d_fpd[slot] + at; j.fpd_stride = j.f_count;
                at += (size_t)j.f_count * p.cfg.channels;
                if (j.total_q <= 0) j.total_q = LLONG_MAX;
                j.blk_lo = (p.bytes_needed_for_frames(j.f_begin + 1) - 1) / B;
                const long long blk_hi = (p.bytes_needed_for_frames(j.f_begin + j.f_count) - 1) / B;
                j.nblk = (int)(blk_hi - j.blk_lo + 1);
                j.fpd_item_begin = (unsigned int)items;
                items += (unsigned long long)j.nblk * p.cfg.channels;
                j.fpd0 = fpd_init_state(j.seed);
            }
            if (items > 0xFFFFFFF0ull) return set_err(D2P_ERR_INVALID_ARG, "too many dither work items in one launch");
            k.fpd_items = (unsigned int)items;
            int st = fpd_jump_device(p.cfg.device, &k.fpd_jump);
            if (st != D2P_OK) return st;
        }
        std::memcpy(h_jobs[slot], live.data(), sizeof(DevJob) * live.size());
        CK(cudaMemcpyAsync(d_jobs[slot], h_jobs[slot], sizeof(DevJob) * live.size(), cudaMemcpyHostToDevice, stream));
        k.jobs = d_jobs[slot];
        k.njobs = (int)live.size();
        k.total_tiles = (unsigned int)tiles;
        k.stats = d_stats;
        k.job_stats = use_jobstats ? d_jobstats : nullptr;
        k.events = d_events;
        k.ev_cap = d_events ? kEventCap : 0;
        const unsigned long long want = (unsigned long long)dp.num_sms * dp.ctas_per_sm;
        const int grid = (int)std::min<unsigned long long>(tiles, want);
        if (fpd) CK(launch_fpd(k, stream));
        bool fast = dp.fast_writer;
        for (const DevJob& j : live) if (j.f64) fast = false;
        CK(launch_kernels(k, kernel_variant(k, dp.lm, fast), grid, dp.threads, dp.smem, stream));
        CK(cudaEventRecord(ev[slot], stream));
        ++launches;
        return D2P_OK;
    }
};

// PcmWriter::update_clip_stats (rd/pcm_writer.rs:658-675) replayed over the sparse clip events.
// Samples are visited in the reference's order: reader block, then channel, then frame.
struct SeqEvent { unsigned long long seq; int high; };
static int64_t replay_clips(std::vector<SeqEvent>& ev) {
    std::sort(ev.begin(), ev.end(), [](const SeqEvent& a, const SeqEvent& b) { return a.seq < b.seq; });
    int64_t clips = 0, last_low = 0, last_high = 0;
    unsigned long long last_seq = 0;
    bool have_last = false;
    for (const SeqEvent& e : ev) {
        if (!have_last || e.seq != last_seq + 1) { last_low = 0; last_high = 0; }   // an unclipped sample in between
        if (!e.high) { if (last_low == 1) ++clips; last_low += 1; }                  // a low clip leaves last_high alone
        else { last_low = 0; if (last_high == 1) ++clips; last_high += 1; }
        last_seq = e.seq; have_last = true;
    }
    return clips;
}

// Position of (channel, frame) in the reference's processing order for a stream fed in block_size chunks.
static unsigned long long seq_index(const Plan& p, int ch, long long f, long long total_q) {
    const long long B = p.cfg.block_size;
    const long long need = p.bytes_needed_for_frames(f + 1);       // frame f appears in the block containing byte need-1
    const long long blk = (need - 1) / B;
    const long long f_before = p.frames_for_bytes(blk * B);
    const long long f_after = p.frames_for_bytes(std::min((blk + 1) * B, total_q));
    return (unsigned long long)(p.cfg.channels * f_before + (long long)ch * (f_after - f_before) + (f - f_before));
}


// Geometry of a container-order payload: block groups, ragged tail, per-channel length.
struct PayloadGeom { size_t G; long long unit, groups, tail_per, total_q; bool planar; };
static PayloadGeom payload_geom(const d2p_config& c, uint64_t payload_bytes) {
    PayloadGeom g;
    g.planar = c.in_layout == D2P_LAYOUT_PLANAR;
    g.G = g.planar ? (size_t)c.channels * c.block_size : (size_t)c.channels;
    g.unit = g.planar ? (long long)c.block_size : 1;
    g.groups = (long long)(payload_bytes / g.G);
    g.tail_per = (long long)((payload_bytes - (size_t)g.groups * g.G) / c.channels);
    g.total_q = g.groups * g.unit + g.tail_per;
    return g;
}

// Slice of the payload that frames [f0, f0+cnt) read: groups [gk, g1) (+ the ragged tail if with_tail).
static void block_slice(const Plan& p, const PayloadGeom& g, long long f0, long long cnt, long long* gk, long long* g1,
                        bool* with_tail) {
    long long k = p.first_byte_needed(f0) / g.unit;
    if (!g.planar) k = (k / 16) * 16;                           // keep 16-byte vector alignment per channel run
    if (k > g.groups) k = g.groups;
    const long long need_q = p.bytes_needed_for_frames(f0 + cnt);
    long long e = (need_q + g.unit - 1) / g.unit;
    *with_tail = need_q > g.groups * g.unit;
    if (e > g.groups) e = g.groups;
    if (e < k) e = k;
    *gk = k; *g1 = e;
}

}  // namespace d2p

using namespace d2p;

// ================================================================================================
// streaming context
// ================================================================================================
struct OutChunk {
    uint8_t* d_pcm = nullptr;       // packed frames (interleaved) or planar32 block
    double* d_f64 = nullptr;
    long long frames = 0;           // frames in the chunk
    long long consumed = 0;         // frames already pulled
};

struct d2p_ctx {
    DevicePlan* dp = nullptr;
    Launcher launcher;
    cudaStream_t stream = nullptr;
    std::vector<uint8_t> pending;   // host bytes not yet forming a whole block group
    // small pushes (the reference hands over one 4096-byte block per channel at a time) are gathered in pinned memory
    // and shipped as one copy + one launch when `push_threshold` bytes have accumulated (or at d2p_flush / d2p_finish)
    uint8_t* h_stage = nullptr;
    size_t staged = 0, push_threshold = (size_t)8 << 20;
    uint8_t* d_in[2] = {nullptr, nullptr};
    size_t d_in_cap[2] = {0, 0};
    int cur = 0;
    long long base_q = 0;           // per-channel position of d_in[cur][0]
    long long have_q = 0;           // per-channel bytes resident/consumed so far (absolute)
    long long frames_done = 0;      // frames launched so far
    long long frames_pulled = 0;
    bool finished = false;
    std::deque<OutChunk> ready;
    // clip events reported by the kernels so far (replayed in d2p_stats)
    std::vector<ClipEvent> clip_events;
    bool clips_inexact = false;
};

static size_t group_bytes(const d2p_config& c) {
    return c.in_layout == D2P_LAYOUT_PLANAR ? (size_t)c.channels * c.block_size : (size_t)c.channels;
}

static int ctx_collect_events(d2p_ctx* ctx) {
    DevStats st;
    CK(cudaMemcpyAsync(&st, ctx->launcher.d_stats, sizeof st, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const unsigned int n = st.n_events;
    if (n > kEventCap) ctx->clips_inexact = true;
    const unsigned int take = std::min(n, kEventCap);
    if (ctx->clip_events.size() + take > (8u << 20)) ctx->clips_inexact = true;     // heavy clipping: give up on the exact count
    else if (take) {
        const size_t old = ctx->clip_events.size();
        ctx->clip_events.resize(old + take);
        CK(cudaMemcpy(ctx->clip_events.data() + old, ctx->launcher.d_events, sizeof(ClipEvent) * take, cudaMemcpyDeviceToHost));
    }
    if (n) CK(cudaMemsetAsync(&ctx->launcher.d_stats->n_events, 0, sizeof(unsigned int), ctx->stream));
    return D2P_OK;
}

// Launch everything computable from the bytes now resident: [frames_done, frames_for_bytes(end_q)).
static int ctx_process(d2p_ctx* ctx, long long end_q, long long tail_q, long long tail_len, bool final) {
    const Plan& p = ctx->dp->plan;
    // Dither F numbers the samples in reader order (block, channel, frame): a block's frames can only be placed
    // once the block is complete (or the stream has ended), so an unfinished block is held back.
    long long upto_q = end_q;
    if (p.cfg.dither == D2P_DITHER_FPD && !final) upto_q = end_q / p.cfg.block_size * p.cfg.block_size;
    const long long f_new = p.frames_for_bytes(upto_q);
    const long long cnt = f_new - ctx->frames_done;
    if (cnt <= 0) return D2P_OK;
    OutChunk oc;
    oc.frames = cnt;
    const size_t pcm_bytes = (size_t)cnt * p.cfg.channels * (p.cfg.out_format == D2P_OUT_PLANAR_32 ? 4 : p.bps);
    CK(cudaMallocAsync((void**)&oc.d_pcm, pcm_bytes + 16, ctx->stream));
    if (p.cfg.keep_f64) CK(cudaMallocAsync((void**)&oc.d_f64, (size_t)cnt * p.cfg.channels * 8, ctx->stream));
    std::vector<DevJob> jobs(1);
    DevJob& j = jobs[0];
    std::memset(&j, 0, sizeof j);
    j.in = ctx->d_in[ctx->cur];
    j.base_q = ctx->base_q; j.end_q = end_q; j.tail_q = tail_q; j.tail_len = tail_len;
    j.f_begin = ctx->frames_done; j.f_count = cnt;
    j.out = oc.d_pcm; j.out_stride = cnt;
    j.f64 = oc.d_f64; j.f64_stride = cnt;
    j.seed = p.cfg.dither_seed;
    j.total_q = final ? end_q : LLONG_MAX;
    int st = ctx->launcher.run(*ctx->dp, jobs, ctx->stream);
    if (st != D2P_OK) return st;
    ctx->frames_done = f_new;
    ctx->ready.push_back(oc);
    if (p.cfg.bits != 32) { st = ctx_collect_events(ctx); if (st != D2P_OK) return st; }
    return D2P_OK;
}

// Make room for `add` more container bytes after the halo of already-consumed input.
static int ctx_prepare_input(d2p_ctx* ctx, size_t add, uint8_t** dst) {
    const Plan& p = ctx->dp->plan;
    const d2p_config& c = p.cfg;
    const bool planar = c.in_layout == D2P_LAYOUT_PLANAR;
    const long long unit = planar ? (long long)c.block_size : 16;          // keep whole blocks (planar) / 16-byte runs
    long long keep_from = p.first_byte_needed(ctx->frames_done);
    keep_from = (keep_from / unit) * unit;
    if (keep_from < ctx->base_q) keep_from = ctx->base_q;
    if (keep_from > ctx->have_q) keep_from = (ctx->have_q / unit) * unit;
    const size_t carry = (size_t)(ctx->have_q - keep_from) * c.channels;
    const size_t need = carry + add + 64;
    const int nxt = ctx->cur ^ 1;
    if (ctx->d_in_cap[nxt] < need) {
        if (ctx->d_in[nxt]) CK(cudaFreeAsync(ctx->d_in[nxt], ctx->stream));
        ctx->d_in_cap[nxt] = need + need / 4;
        CK(cudaMallocAsync((void**)&ctx->d_in[nxt], ctx->d_in_cap[nxt], ctx->stream));
    }
    if (carry)
        CK(cudaMemcpyAsync(ctx->d_in[nxt], ctx->d_in[ctx->cur] + (size_t)(keep_from - ctx->base_q) * c.channels, carry,
                           cudaMemcpyDeviceToDevice, ctx->stream));
    ctx->cur = nxt;
    ctx->base_q = keep_from;
    *dst = ctx->d_in[nxt] + carry;
    return D2P_OK;
}

extern "C" {

int d2p_abi_version(void) { return D2P_ABI_VERSION; }

const char* d2p_strerror(int status) {
    switch (status) {
    case D2P_OK: return "ok";
    case D2P_ERR_INVALID_ARG: return "invalid argument";
    case D2P_ERR_TAPS_NOT_FOUND: return "Taps not found for decimation ratio / filter / dsd_rate";
    case D2P_ERR_UNSUPPORTED_LM: return "Unsupported L/M combination";
    case D2P_ERR_BIT_DEPTH: return "Unsupported bit depth";
    case D2P_ERR_NO_DEVICE: return "no CUDA device available (this library has no CPU path)";
    case D2P_ERR_CUDA: return "CUDA error";
    case D2P_ERR_NOMEM: return "out of memory";
    case D2P_ERR_STATE: return "invalid call sequence";
    case D2P_ERR_CAPACITY: return "output buffer too small";
    case D2P_ERR_CANCELLED: return "conversion cancelled";
    default: return "unknown status";
    }
}
const char* d2p_last_error(void) { return g_last_error.c_str(); }

int d2p_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int d2p_plan(const d2p_config* cfg, d2p_plan_info* info) {
    if (!cfg || !info) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    Plan p; std::string msg;
    int st = make_plan(*cfg, &p, &msg);
    if (st != D2P_OK) return set_err(st, msg);
    std::memset(info, 0, sizeof *info);
    info->decim_ratio = p.M; info->upsample_ratio = p.L; info->num_stages = p.nstages;
    for (int s = 0; s < p.nstages; ++s) {
        std::snprintf(info->stage_table[s], sizeof info->stage_table[s], "%s", p.tab[s]->name);
        info->stage_half_taps[s] = p.tab[s]->len;
        info->stage_decim[s] = p.sdecim[s];
    }
    info->scale_factor = p.scale; info->peak_level = p.peak; info->bytes_per_sample = p.bps;
    const uint64_t bits = (uint64_t)cfg->block_size * 8, d = (uint64_t)(p.M < 0 ? -p.M : p.M);
    info->frames_cap = (bits * p.L + (d - 1)) / d + (p.L > 1 ? 16 : 0);          // calc_frames_cap rd/lib.rs:259-270
    info->lut_bytes = p.nstages == 1 ? (uint64_t)p.nt * 2048 : (uint64_t)p.L * p.gmax * 2048;
    return D2P_OK;
}

uint64_t d2p_frames_for_bytes(const d2p_config* cfg, uint64_t bytes_per_channel) {
    if (!cfg) return 0;
    Plan p; std::string msg;
    if (make_plan(*cfg, &p, &msg) != D2P_OK) { g_last_error = msg; return 0; }
    return (uint64_t)p.frames_for_bytes((long long)bytes_per_channel);
}

int d2p_tap_table(const char* name, const double** data) {
    const TapTable* t = name ? find_table(name) : nullptr;
    if (!t) return -1;
    if (data) *data = t->data;
    return t->len;
}
int d2p_tap_table_count(void) { return table_count(); }
const char* d2p_tap_table_name(int index) { const TapTable* t = table_at(index); return t ? t->name : nullptr; }

// ---- streaming -----------------------------------------------------------------------------------
int d2p_create(const d2p_config* cfg, d2p_ctx** out) {
    if (!cfg || !out) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    if (cfg->out_format == D2P_OUT_NONE)
        return set_err(D2P_ERR_INVALID_ARG, "D2P_OUT_NONE (level check) is served by d2p_check_level / d2p_batch_*, not by a streaming context");
    DevicePlan* dp = nullptr;
    int st = build_device_plan(*cfg, &dp);
    if (st != D2P_OK) return st;
    d2p_ctx* ctx = new d2p_ctx();
    ctx->dp = dp;
    cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { d2p_destroy(ctx); return cuda_err(e, "cudaStreamCreate"); }
    st = ctx->launcher.init(cfg->bits != 32);
    if (st != D2P_OK) { d2p_destroy(ctx); return st; }
    *out = ctx;
    return D2P_OK;
}

void d2p_destroy(d2p_ctx* ctx) {
    if (!ctx) return;
    if (ctx->dp) cudaSetDevice(ctx->dp->plan.cfg.device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (OutChunk& oc : ctx->ready) { if (oc.d_pcm) cudaFree(oc.d_pcm); if (oc.d_f64) cudaFree(oc.d_f64); }
    for (int i = 0; i < 2; ++i) if (ctx->d_in[i]) cudaFree(ctx->d_in[i]);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    ctx->launcher.destroy();
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    free_device_plan(ctx->dp);
    delete ctx;
}

// Copy `nbytes` to the device and launch every frame that became computable.
static int ctx_ship(d2p_ctx* ctx, const uint8_t* raw, size_t nbytes) {
    if (nbytes == 0) return D2P_OK;
    const d2p_config& c = ctx->dp->plan.cfg;
    const size_t G = group_bytes(c);
    // whole groups available = pending + raw
    const size_t total = ctx->pending.size() + nbytes;
    const size_t whole = (total / G) * G;
    if (whole == 0) { ctx->pending.insert(ctx->pending.end(), raw, raw + nbytes); return D2P_OK; }
    uint8_t* dst = nullptr;
    int st = ctx_prepare_input(ctx, whole, &dst);
    if (st != D2P_OK) return st;
    size_t from_pending = ctx->pending.size();
    if (from_pending) CK(cudaMemcpyAsync(dst, ctx->pending.data(), from_pending, cudaMemcpyHostToDevice, ctx->stream));
    const size_t from_raw = whole - from_pending;
    CK(cudaMemcpyAsync(dst + from_pending, raw, from_raw, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));            // `raw` and `pending` may be reused by the caller after return
    ctx->pending.assign(raw + from_raw, raw + nbytes);
    ctx->have_q += (long long)(whole / c.channels);
    return ctx_process(ctx, ctx->have_q, LLONG_MAX, 0, false);
}

static int ctx_flush(d2p_ctx* ctx) {
    if (ctx->staged == 0) return D2P_OK;
    const size_t n = ctx->staged;
    ctx->staged = 0;
    return ctx_ship(ctx, ctx->h_stage, n);
}

int d2p_push(d2p_ctx* ctx, const uint8_t* raw, size_t nbytes) {
    if (!ctx || (!raw && nbytes)) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (ctx->finished) return set_err(D2P_ERR_STATE, "push after finish");
    if (nbytes == 0) return D2P_OK;
    CK(cudaSetDevice(ctx->dp->plan.cfg.device));
    const size_t T = ctx->push_threshold;
    if (T == 0 || nbytes >= T) {                       // large pushes go straight to the device
        int st = ctx_flush(ctx);
        return st != D2P_OK ? st : ctx_ship(ctx, raw, nbytes);
    }
    if (!ctx->h_stage) CK(cudaMallocHost((void**)&ctx->h_stage, 2 * T));
    std::memcpy(ctx->h_stage + ctx->staged, raw, nbytes);          // staged < T and nbytes < T: fits in 2T
    ctx->staged += nbytes;
    return ctx->staged >= T ? ctx_flush(ctx) : D2P_OK;
}

int d2p_flush(d2p_ctx* ctx) {
    if (!ctx) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (ctx->finished) return D2P_OK;
    CK(cudaSetDevice(ctx->dp->plan.cfg.device));
    return ctx_flush(ctx);
}

int d2p_set_push_threshold(d2p_ctx* ctx, size_t nbytes) {
    if (!ctx) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (nbytes > ((size_t)1 << 30)) return set_err(D2P_ERR_INVALID_ARG, "push threshold above 1 GiB");
    CK(cudaSetDevice(ctx->dp->plan.cfg.device));
    int st = ctx_flush(ctx);
    if (st != D2P_OK) return st;
    if (ctx->h_stage) { cudaFreeHost(ctx->h_stage); ctx->h_stage = nullptr; }
    ctx->push_threshold = nbytes;
    return D2P_OK;
}

int d2p_finish(d2p_ctx* ctx) {
    if (!ctx) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (ctx->finished) return D2P_OK;
    const d2p_config& c = ctx->dp->plan.cfg;
    CK(cudaSetDevice(c.device));
    { int fs = ctx_flush(ctx); if (fs != D2P_OK) return fs; }
    ctx->finished = true;
    const size_t per = ctx->pending.size() / c.channels;      // trailing partial group: split evenly (A1)
    if (per == 0) return ctx_process(ctx, ctx->have_q, LLONG_MAX, 0, true);   // frames held back for dither F
    const size_t take = per * c.channels;
    uint8_t* dst = nullptr;
    int st = ctx_prepare_input(ctx, take, &dst);
    if (st != D2P_OK) return st;
    CK(cudaMemcpyAsync(dst, ctx->pending.data(), take, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const long long tail_q = ctx->have_q;
    ctx->have_q += (long long)per;
    ctx->pending.clear();
    if (c.in_layout == D2P_LAYOUT_PLANAR) return ctx_process(ctx, ctx->have_q, tail_q, (long long)per, true);
    return ctx_process(ctx, ctx->have_q, LLONG_MAX, 0, true);
}

uint64_t d2p_frames_ready(const d2p_ctx* ctx) { return ctx ? (uint64_t)(ctx->frames_done - ctx->frames_pulled) : 0; }

int d2p_pull(d2p_ctx* ctx, void* pcm, size_t cap_bytes, uint64_t* frames) {
    if (!ctx || !pcm || !frames) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    const Plan& p = ctx->dp->plan;
    const d2p_config& c = p.cfg;
    CK(cudaSetDevice(c.device));
    *frames = 0;
    if (c.out_format == D2P_OUT_PLANAR_32) {
        // channel-major result: all ready chunks are merged, `n` frames per channel
        const uint64_t avail = d2p_frames_ready(ctx);
        const uint64_t n = std::min<uint64_t>(avail, cap_bytes / ((size_t)c.channels * 4));
        if (n == 0) return D2P_OK;
        uint64_t done = 0;
        while (done < n) {
            OutChunk& oc = ctx->ready.front();
            const uint64_t take = std::min<uint64_t>(n - done, (uint64_t)(oc.frames - oc.consumed));
            for (int ch = 0; ch < c.channels; ++ch)
                CK(cudaMemcpyAsync((uint8_t*)pcm + ((size_t)ch * n + done) * 4,
                                   oc.d_pcm + ((size_t)ch * oc.frames + oc.consumed) * 4, take * 4,
                                   cudaMemcpyDeviceToHost, ctx->stream));
            oc.consumed += (long long)take;
            done += take;
            if (oc.consumed == oc.frames) {
                CK(cudaStreamSynchronize(ctx->stream));
                cudaFreeAsync(oc.d_pcm, ctx->stream);
                if (oc.d_f64) cudaFreeAsync(oc.d_f64, ctx->stream);
                ctx->ready.pop_front();
            }
        }
        CK(cudaStreamSynchronize(ctx->stream));
        ctx->frames_pulled += (long long)n;
        *frames = n;
        return D2P_OK;
    }
    const size_t fb = (size_t)c.channels * p.bps;
    uint64_t room = cap_bytes / fb, done = 0;
    while (room && !ctx->ready.empty()) {
        OutChunk& oc = ctx->ready.front();
        const uint64_t take = std::min<uint64_t>(room, (uint64_t)(oc.frames - oc.consumed));
        CK(cudaMemcpyAsync((uint8_t*)pcm + done * fb, oc.d_pcm + (size_t)oc.consumed * fb, take * fb,
                           cudaMemcpyDeviceToHost, ctx->stream));
        oc.consumed += (long long)take; done += take; room -= take;
        if (oc.consumed == oc.frames) {
            CK(cudaStreamSynchronize(ctx->stream));
            cudaFreeAsync(oc.d_pcm, ctx->stream);
            if (oc.d_f64) cudaFreeAsync(oc.d_f64, ctx->stream);
            ctx->ready.pop_front();
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->frames_pulled += (long long)done;
    *frames = done;
    return D2P_OK;
}

int d2p_pull_f64(d2p_ctx* ctx, int chan, double* out, size_t cap, uint64_t* n) {
    if (!ctx || !out || !n) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    const d2p_config& c = ctx->dp->plan.cfg;
    if (!c.keep_f64) return set_err(D2P_ERR_STATE, "context was created without keep_f64");
    if (chan < 0 || chan >= c.channels) return set_err(D2P_ERR_INVALID_ARG, "bad channel");
    CK(cudaSetDevice(c.device));
    uint64_t done = 0;
    for (const OutChunk& oc : ctx->ready) {
        if (done >= cap) break;
        const uint64_t take = std::min<uint64_t>(cap - done, (uint64_t)(oc.frames - oc.consumed));
        CK(cudaMemcpyAsync(out + done, oc.d_f64 + (size_t)chan * oc.frames + oc.consumed, take * 8,
                           cudaMemcpyDeviceToHost, ctx->stream));
        done += take;
    }
    CK(cudaStreamSynchronize(ctx->stream));
    *n = done;
    return D2P_OK;
}

int d2p_stats(d2p_ctx* ctx, int64_t* clips, double* peak_abs, int64_t* clipped_samples) {
    if (!ctx) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    CK(cudaSetDevice(ctx->dp->plan.cfg.device));
    DevStats st;
    CK(cudaMemcpyAsync(&st, ctx->launcher.d_stats, sizeof st, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (clips) {
        if (ctx->clips_inexact) *clips = -1;
        else {
            const Plan& p = ctx->dp->plan;
            std::vector<SeqEvent> se(ctx->clip_events.size());
            for (size_t i = 0; i < se.size(); ++i) {
                se[i].seq = seq_index(p, ctx->clip_events[i].ch_kind & 0xFFFF, ctx->clip_events[i].frame, ctx->have_q);
                se[i].high = (ctx->clip_events[i].ch_kind >> 16) & 1;
            }
            *clips = replay_clips(se);
        }
    }
    if (peak_abs) { double d; std::memcpy(&d, &st.peak_bits, 8); *peak_abs = d; }
    if (clipped_samples) *clipped_samples = (int64_t)(st.clip_lo + st.clip_hi);
    return D2P_OK;
}

// ---- pinned host memory ----------------------------------------------------------------------------
void* d2p_host_alloc(size_t nbytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, nbytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void d2p_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- whole payload, host buffers, pipelined ---------------------------------------------------------
// HostPipe: three streams, each with its own device input/output staging, reused across calls.
// Chunk i runs H2D -> kernel -> D2H on stream i%3, so the copies of neighbouring chunks overlap the
// kernel (PCIe is full duplex).  Every chunk ships its own filter halo, so chunks are independent.
namespace d2p {
struct HostPipe {
    static const int NS = 3;
    cudaStream_t streams[NS] = {nullptr, nullptr, nullptr};
    uint8_t* d_in[NS] = {nullptr, nullptr, nullptr};
    uint8_t* d_out[NS] = {nullptr, nullptr, nullptr};
    size_t in_cap = 0, out_cap = 0;
    cudaEvent_t done[NS] = {nullptr, nullptr, nullptr};            // chunk on stream s has landed in the caller's buffer
    bool ready = false;

    int init() {
        if (ready) return D2P_OK;
        for (int s = 0; s < NS; ++s) CK(cudaStreamCreateWithFlags(&streams[s], cudaStreamNonBlocking));
        for (int s = 0; s < NS; ++s) CK(cudaEventCreateWithFlags(&done[s], cudaEventDisableTiming));
        ready = true;
        return D2P_OK;
    }
    int reserve(size_t in_need, size_t out_need) {
        if (in_need > in_cap) {
            for (int s = 0; s < NS; ++s) { if (d_in[s]) cudaFree(d_in[s]); d_in[s] = nullptr; }
            for (int s = 0; s < NS; ++s) if (cudaMalloc(&d_in[s], in_need) != cudaSuccess) { cudaGetLastError(); in_cap = 0; return set_err(D2P_ERR_NOMEM, "device allocation failed"); }
            in_cap = in_need;
        }
        if (out_need > out_cap) {
            for (int s = 0; s < NS; ++s) { if (d_out[s]) cudaFree(d_out[s]); d_out[s] = nullptr; }
            for (int s = 0; s < NS; ++s) if (cudaMalloc(&d_out[s], out_need) != cudaSuccess) { cudaGetLastError(); out_cap = 0; return set_err(D2P_ERR_NOMEM, "device allocation failed"); }
            out_cap = out_need;
        }
        return D2P_OK;
    }
    void destroy() {
        for (int s = 0; s < NS; ++s) {
            if (streams[s]) { cudaStreamSynchronize(streams[s]); cudaStreamDestroy(streams[s]); }
            if (done[s]) cudaEventDestroy(done[s]);
            if (d_in[s]) cudaFree(d_in[s]);
            if (d_out[s]) cudaFree(d_out[s]);
        }
    }
};

// Cancellation and progress: the reference polls its cancel flag once per reader block and reports
// progress through track_io (rd/conversion_context.rs:222, 304-308, 331-371).  Here both are honoured between pipeline
// chunks: the host thread never runs more than NS chunks ahead of the device, so frames_done is what has really
// arrived in the caller's buffer and a raised flag stops the call within a few chunks.
struct HostControl {
    const volatile int32_t* cancel = nullptr;
    d2p_progress_fn progress = nullptr;
    void* user = nullptr;
};

static int host_convert(const DevicePlan& dp, Launcher& launcher, HostPipe& pipe, const uint8_t* raw, size_t nbytes,
                        void* pcm, size_t out_cap_bytes, uint64_t* frames, size_t chunk_bytes, const HostControl& ctl = HostControl()) {
    const Plan& p = dp.plan;
    const d2p_config& c = p.cfg;
    const bool planar = c.in_layout == D2P_LAYOUT_PLANAR;
    const size_t G = group_bytes(c);
    const long long unit = planar ? (long long)c.block_size : 1;          // per-channel bytes per group
    const long long groups = (long long)(nbytes / G);
    const long long tail_per = (long long)((nbytes - (size_t)groups * G) / c.channels);
    const long long total_q = groups * unit + tail_per;
    const long long total_frames = p.frames_for_bytes(total_q);
    const bool no_pcm = c.out_format == D2P_OUT_NONE;              // level check: nothing comes back
    const size_t fb = no_pcm ? 0 : (size_t)c.channels * p.bps;
    *frames = 0;
    if ((size_t)total_frames * fb > out_cap_bytes) return set_err(D2P_ERR_CAPACITY, "output buffer too small");
    int rc = pipe.init();
    if (rc != D2P_OK) return rc;

    const long long chunk_groups = std::max<long long>(1, (long long)(chunk_bytes / G));
    // halo of one chunk: what the first frame of the chunk reads before the chunk's first byte
    const long long halo_q = p.bytes_needed_for_frames(1) + 64;            // >= filter span in input bytes
    const long long halo_groups = (halo_q + unit - 1) / unit + (planar ? 1 : 16);
    const size_t in_need = (size_t)(chunk_groups + halo_groups) * G + (size_t)tail_per * c.channels + 256;
    const size_t out_need = (size_t)(p.frames_for_bytes((chunk_groups + halo_groups) * unit + tail_per) + 64) * fb + 256;
    if ((rc = pipe.reserve(in_need, out_need)) != D2P_OK) return rc;

    long long f_done = 0;
    int ci = 0;
    long long f_upto[HostPipe::NS] = {0, 0, 0};                       // frames complete once stream s has drained
    const bool controlled = ctl.cancel || ctl.progress;
    for (long long g0 = 0;;) {
        if (controlled && ci >= HostPipe::NS) {                       // chunk ci - NS shares this chunk's staging buffers
            const int s = ci % HostPipe::NS;
            cudaError_t e = cudaEventSynchronize(pipe.done[s]);
            if (e != cudaSuccess) { rc = cuda_err(e, "chunk sync"); break; }
            if (ctl.progress) ctl.progress(ctl.user, (uint64_t)f_upto[s], (uint64_t)total_frames);
        }
        if (ctl.cancel && *ctl.cancel) { rc = set_err(D2P_ERR_CANCELLED, "conversion cancelled"); break; }
        const long long g1 = std::min(groups, g0 + chunk_groups);
        const bool last = g1 >= groups;
        const long long end_q = g1 * unit + (last ? tail_per : 0);
        const long long f_new = p.frames_for_bytes(end_q);
        if (f_new > f_done) {
            const int s = ci % HostPipe::NS; ++ci;
            long long gk = p.first_byte_needed(f_done) / unit;             // first group to ship (halo included)
            if (!planar) gk = (gk / 16) * 16;
            const size_t off = (size_t)gk * G;
            const size_t len = (size_t)(g1 - gk) * G + (last ? (size_t)tail_per * c.channels : 0);
            if (len > pipe.in_cap) { rc = set_err(D2P_ERR_NOMEM, "internal: chunk larger than staging buffer"); break; }
            cudaError_t e = cudaMemcpyAsync(pipe.d_in[s], raw + off, len, cudaMemcpyHostToDevice, pipe.streams[s]);
            if (e != cudaSuccess) { rc = cuda_err(e, "H2D"); break; }
            std::vector<DevJob> jobs(1);
            DevJob& j = jobs[0];
            std::memset(&j, 0, sizeof j);
            j.in = pipe.d_in[s]; j.base_q = gk * unit; j.end_q = end_q;
            j.tail_q = (last && tail_per > 0 && planar) ? groups * unit : LLONG_MAX; j.tail_len = tail_per;
            j.f_begin = f_done; j.f_count = f_new - f_done;
            j.out = pipe.d_out[s]; j.seed = c.dither_seed; j.total_q = total_q;
            if ((size_t)j.f_count * fb > pipe.out_cap) { rc = set_err(D2P_ERR_NOMEM, "internal: chunk output larger than staging buffer"); break; }
            if ((rc = launcher.run(dp, jobs, pipe.streams[s], false)) != D2P_OK) break;
            if (!no_pcm) {
                e = cudaMemcpyAsync((uint8_t*)pcm + (size_t)f_done * fb, pipe.d_out[s], (size_t)j.f_count * fb, cudaMemcpyDeviceToHost, pipe.streams[s]);
                if (e != cudaSuccess) { rc = cuda_err(e, "D2H"); break; }
            }
            f_done = f_new;
            if (controlled) {
                f_upto[s] = f_done;
                e = cudaEventRecord(pipe.done[s], pipe.streams[s]);
                if (e != cudaSuccess) { rc = cuda_err(e, "event record"); break; }
            }
        }
        if (last) break;
        g0 = g1;
    }
    for (int s = 0; s < HostPipe::NS; ++s) {
        cudaError_t e = cudaStreamSynchronize(pipe.streams[s]);
        if (e != cudaSuccess && rc == D2P_OK) rc = cuda_err(e, "stream sync");
    }
    if (rc == D2P_OK) {
        *frames = (uint64_t)f_done;
        if (ctl.progress) ctl.progress(ctl.user, (uint64_t)f_done, (uint64_t)total_frames);
    }
    return rc;
}
}  // namespace d2p

// ---- device-resident batch --------------------------------------------------------------------------
struct d2p_batch {
    DevicePlan* dp = nullptr;
    Launcher launcher;
    HostPipe pipe;
    HostControl ctl;
};

int d2p_batch_create(const d2p_config* cfg, d2p_batch** out) {
    if (!cfg || !out) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    DevicePlan* dp = nullptr;
    int st = build_device_plan(*cfg, &dp);
    if (st != D2P_OK) return st;
    d2p_batch* b = new d2p_batch();
    b->dp = dp;
    b->launcher.want_jobstats = true;
    st = b->launcher.init(false);
    if (st != D2P_OK) { d2p_batch_destroy(b); return st; }
    *out = b;
    return D2P_OK;
}

void d2p_batch_destroy(d2p_batch* b) {
    if (!b) return;
    if (b->dp) cudaSetDevice(b->dp->plan.cfg.device);
    cudaDeviceSynchronize();
    b->pipe.destroy();
    b->launcher.destroy();
    free_device_plan(b->dp);
    delete b;
}

int d2p_batch_convert_host(d2p_batch* b, const uint8_t* raw, size_t nbytes, void* pcm, size_t out_cap_bytes,
                           uint64_t* frames) {
    if (!b || (!raw && nbytes) || !frames) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    const d2p_config& c = b->dp->plan.cfg;
    if (!pcm && c.out_format != D2P_OUT_NONE) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (c.out_format == D2P_OUT_PLANAR_32 || c.keep_f64)
        return set_err(D2P_ERR_INVALID_ARG, "host conversion produces interleaved packed PCM only");
    CK(cudaSetDevice(c.device));
    static const size_t chunk = [] {                       // D2P_HOST_CHUNK_MB: pipeline granularity (tuning knob)
        const char* e = std::getenv("D2P_HOST_CHUNK_MB");
        const long mb = e ? std::atol(e) : 0;
        return (size_t)(mb >= 1 && mb <= 1024 ? mb : 64) << 20;
    }();
    return host_convert(*b->dp, b->launcher, b->pipe, raw, nbytes, pcm, out_cap_bytes, frames, chunk, b->ctl);
}

int d2p_batch_set_control(d2p_batch* b, const volatile int32_t* cancel_flag, d2p_progress_fn progress, void* user) {
    if (!b) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    b->ctl.cancel = cancel_flag; b->ctl.progress = progress; b->ctl.user = user;
    return D2P_OK;
}

int d2p_convert_host(const d2p_config* cfg, const uint8_t* raw, size_t nbytes, void* pcm, size_t out_cap_bytes,
                     uint64_t* frames, int64_t* clips, double* peak_abs) {
    if (!cfg || !frames) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    d2p_batch* b = nullptr;
    int st = d2p_batch_create(cfg, &b);
    if (st != D2P_OK) return st;
    st = d2p_batch_convert_host(b, raw, nbytes, pcm, out_cap_bytes, frames);
    if (st == D2P_OK && (clips || peak_abs)) st = d2p_batch_stats(b, nullptr, clips, peak_abs);
    d2p_batch_destroy(b);
    return st;
}

// Rdsd2Pcm::new_level_check + check_level (rd/lib.rs:124-158, 171-182): 32-bit sink, no dither, 0 dB, equiripple.
int d2p_check_level(const d2p_config* cfg, const uint8_t* raw, size_t nbytes, double* peak_abs, float* peak_dbfs) {
    if (!cfg || (!raw && nbytes)) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    d2p_config c = *cfg;
    c.bits = 32; c.dither = D2P_DITHER_NONE; c.level_db = 0.0; c.filter = D2P_FILTER_EQUIRIPPLE;
    c.out_format = D2P_OUT_NONE; c.keep_f64 = 0; c.has_dither_scale = 0;
    d2p_batch* b = nullptr;
    int st = d2p_batch_create(&c, &b);
    if (st != D2P_OK) return st;
    uint64_t frames = 0;
    double peak = 0.0;
    st = d2p_batch_convert_host(b, raw, nbytes, nullptr, 0, &frames);
    if (st == D2P_OK) st = d2p_batch_stats(b, nullptr, nullptr, &peak);
    d2p_batch_destroy(b);
    if (st != D2P_OK) return st;
    if (peak_abs) *peak_abs = peak;
    // process_levels :247-251
    if (peak_dbfs) *peak_dbfs = peak <= 0.0 ? -INFINITY : (float)(20.0 * std::log10(peak));
    return D2P_OK;
}

int d2p_batch_run(d2p_batch* b, d2p_job* jobs, int njobs, void* cuda_stream) {
    if (!b || (!jobs && njobs)) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    const Plan& p = b->dp->plan;
    const d2p_config& c = p.cfg;
    CK(cudaSetDevice(c.device));
    const bool planar = c.in_layout == D2P_LAYOUT_PLANAR;
    const size_t G = group_bytes(c);
    const long long unit = planar ? (long long)c.block_size : 1;
    const size_t fb = c.out_format == D2P_OUT_NONE ? 0 : (size_t)c.channels * (c.out_format == D2P_OUT_PLANAR_32 ? 4 : p.bps);
    std::vector<DevJob> dj((size_t)njobs);
    for (int i = 0; i < njobs; ++i) {
        d2p_job& u = jobs[i];
        const long long groups = (long long)(u.in_bytes / G);
        const long long tail_per = (long long)((u.in_bytes - (size_t)groups * G) / c.channels);
        const long long total_q = groups * unit + tail_per;
        const long long nf = p.frames_for_bytes(total_q);
        if ((size_t)nf * fb > u.out_cap_bytes) return set_err(D2P_ERR_CAPACITY, "job output buffer too small");
        if (u.d_f64 && (long long)u.f64_stride < nf) return set_err(D2P_ERR_CAPACITY, "job f64 stride too small");
        DevJob& j = dj[i];
        std::memset(&j, 0, sizeof j);
        j.in = (const uint8_t*)u.d_in; j.base_q = 0; j.end_q = total_q;
        j.tail_q = (planar && tail_per > 0) ? groups * unit : LLONG_MAX; j.tail_len = tail_per;
        j.f_begin = 0; j.f_count = nf;
        j.out = (uint8_t*)u.d_out; j.out_stride = nf;
        j.f64 = u.d_f64; j.f64_stride = (long long)u.f64_stride;
        j.seed = u.dither_seed; j.total_q = total_q;
        u.frames = (uint64_t)nf;
    }
    return b->launcher.run(*b->dp, dj, (cudaStream_t)cuda_stream);
}


// ---- time-block split ----------------------------------------------------------------------------------
uint64_t d2p_total_frames(const d2p_config* cfg, uint64_t payload_bytes) {
    if (!cfg) return 0;
    Plan p; std::string msg;
    if (make_plan(*cfg, &p, &msg) != D2P_OK) { g_last_error = msg; return 0; }
    return (uint64_t)p.frames_for_bytes(payload_geom(*cfg, payload_bytes).total_q);
}

int d2p_block_range(const d2p_config* cfg, uint64_t payload_bytes, uint64_t first_frame, uint64_t frame_count,
                    uint64_t* in_offset, uint64_t* in_bytes) {
    if (!cfg || !in_offset || !in_bytes) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    Plan p; std::string msg;
    int st = make_plan(*cfg, &p, &msg);
    if (st != D2P_OK) return set_err(st, msg);
    const PayloadGeom g = payload_geom(*cfg, payload_bytes);
    const long long total = p.frames_for_bytes(g.total_q);
    if ((long long)(first_frame + frame_count) > total) return set_err(D2P_ERR_INVALID_ARG, "frame range beyond the payload");
    if (frame_count == 0) { *in_offset = 0; *in_bytes = 0; return D2P_OK; }
    long long gk, g1; bool tail;
    block_slice(p, g, (long long)first_frame, (long long)frame_count, &gk, &g1, &tail);
    *in_offset = (uint64_t)gk * g.G;
    *in_bytes = (uint64_t)(g1 - gk) * g.G + (tail ? (uint64_t)g.tail_per * cfg->channels : 0);
    return D2P_OK;
}

int d2p_batch_run_blocks(d2p_batch* b, const d2p_block_job* jobs, int njobs, void* cuda_stream) {
    if (!b || (!jobs && njobs)) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    const Plan& p = b->dp->plan;
    const d2p_config& c = p.cfg;
    CK(cudaSetDevice(c.device));
    const size_t fb = c.out_format == D2P_OUT_NONE ? 0 : (size_t)c.channels * (c.out_format == D2P_OUT_PLANAR_32 ? 4 : p.bps);
    std::vector<DevJob> dj((size_t)njobs);
    for (int i = 0; i < njobs; ++i) {
        const d2p_block_job& u = jobs[i];
        const PayloadGeom g = payload_geom(c, u.payload_bytes);
        const long long total = p.frames_for_bytes(g.total_q);
        if ((long long)(u.first_frame + u.frame_count) > total) return set_err(D2P_ERR_INVALID_ARG, "frame range beyond the payload");
        DevJob& j = dj[i];
        std::memset(&j, 0, sizeof j);
        if (u.frame_count == 0) continue;
        long long gk, g1; bool tail;
        block_slice(p, g, (long long)u.first_frame, (long long)u.frame_count, &gk, &g1, &tail);
        const uint64_t off = (uint64_t)gk * g.G;
        const uint64_t len = (uint64_t)(g1 - gk) * g.G + (tail ? (uint64_t)g.tail_per * c.channels : 0);
        if (u.in_offset != off || u.in_bytes < len) return set_err(D2P_ERR_INVALID_ARG, "block input slice does not match d2p_block_range");
        if ((size_t)u.frame_count * fb > u.out_cap_bytes) return set_err(D2P_ERR_CAPACITY, "job output buffer too small");
        if (u.d_f64 && u.f64_stride < u.frame_count) return set_err(D2P_ERR_CAPACITY, "job f64 stride too small");
        j.in = (const uint8_t*)u.d_in; j.base_q = gk * g.unit;
        j.end_q = tail ? g.total_q : g1 * g.unit;
        j.tail_q = (g.planar && tail && g.tail_per > 0) ? g.groups * g.unit : LLONG_MAX; j.tail_len = g.tail_per;
        j.f_begin = (long long)u.first_frame; j.f_count = (long long)u.frame_count;
        j.out = (uint8_t*)u.d_out; j.out_stride = (long long)u.frame_count;
        j.f64 = u.d_f64; j.f64_stride = (long long)u.f64_stride;
        j.seed = u.dither_seed; j.total_q = g.total_q;
    }
    return b->launcher.run(*b->dp, dj, (cudaStream_t)cuda_stream);
}

uint64_t d2p_batch_launches(const d2p_batch* b) { return b ? b->launcher.launches : 0; }

// ---- conversion without any allocation ------------------------------------------------------------------
// d2p_batch_run for callers that may not allocate device memory (see the header for who they are).  What the batch
// object would have cudaMalloc'ed -- statistics record, job table, LUT -- lives in the scratch array built into the
// module (d2p_kernels.cu); this function only writes it, on the caller's stream, in front of the launch.  Nothing in
// here, nor in what it calls, allocates or frees device or pinned memory; host-side work uses the ordinary heap.
int d2p_run_noalloc(const d2p_config* cfg, d2p_job* jobs, int njobs, void* cuda_stream,
                    int64_t* clipped_samples, double* peak_abs) {
    if (!cfg || (!jobs && njobs) || njobs < 0) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (njobs > kNoallocMaxJobs) {
        char buf[160];
        std::snprintf(buf, sizeof buf, "d2p_run_noalloc: %d jobs in one call, the built-in job table holds %d (D2P_NOALLOC_MAX_JOBS)",
                      njobs, kNoallocMaxJobs);
        return set_err(D2P_ERR_CAPACITY, buf);
    }
    if (cfg->dither == D2P_DITHER_FPD)
        return set_err(D2P_ERR_INVALID_ARG, "d2p_run_noalloc: dither F keeps 4 bytes of generator state per output sample and a "
                                            "4 MiB jump table on the device; the built-in scratch has no room for them");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    std::vector<double> lut;
    DevicePlan* raw_dp = nullptr;
    int st = build_device_plan(*cfg, &raw_dp, &lut);             // sets the device; places nothing on it
    if (st != D2P_OK) return st;
    std::unique_ptr<DevicePlan> dp(raw_dp);
    const Plan& p = dp->plan;
    const d2p_config& c = p.cfg;
    const size_t lut_bytes = lut.size() * sizeof(double);
    if (lut_bytes > kNoallocScratchBytes - kNoallocLutOffset) {
        char buf[160];
        std::snprintf(buf, sizeof buf, "d2p_run_noalloc: filter LUT of %zu bytes exceeds the %zu bytes of built-in scratch",
                      lut_bytes, kNoallocScratchBytes - kNoallocLutOffset);
        return set_err(D2P_ERR_CAPACITY, buf);
    }

    // the same job description as d2p_batch_run: one whole payload per job
    const bool planar = c.in_layout == D2P_LAYOUT_PLANAR;
    const size_t G = group_bytes(c);
    const long long unit = planar ? (long long)c.block_size : 1;
    const size_t fb = c.out_format == D2P_OUT_NONE ? 0 : (size_t)c.channels * (c.out_format == D2P_OUT_PLANAR_32 ? 4 : p.bps);
    DevJob live[kNoallocMaxJobs];
    int nlive = 0;
    unsigned long long tiles = 0;
    bool fast = dp->fast_writer;
    for (int i = 0; i < njobs; ++i) {
        d2p_job& u = jobs[i];
        const long long groups = (long long)(u.in_bytes / G);
        const long long tail_per = (long long)((u.in_bytes - (size_t)groups * G) / c.channels);
        const long long total_q = groups * unit + tail_per;
        const long long nf = p.frames_for_bytes(total_q);
        if ((size_t)nf * fb > u.out_cap_bytes) return set_err(D2P_ERR_CAPACITY, "job output buffer too small");
        if (u.d_f64 && (long long)u.f64_stride < nf) return set_err(D2P_ERR_CAPACITY, "job f64 stride too small");
        u.frames = (uint64_t)nf;
        if (nf <= 0) continue;
        DevJob& j = live[nlive++];
        std::memset(&j, 0, sizeof j);
        j.in = (const uint8_t*)u.d_in; j.base_q = 0; j.end_q = total_q;
        j.tail_q = (planar && tail_per > 0) ? groups * unit : LLONG_MAX; j.tail_len = tail_per;
        j.f_begin = 0; j.f_count = nf;
        j.out = (uint8_t*)u.d_out; j.out_stride = nf;
        j.f64 = u.d_f64; j.f64_stride = (long long)u.f64_stride;
        j.seed = u.dither_seed; j.total_q = total_q;
        j.job_id = i;
        j.tile_begin = (unsigned int)tiles;
        tiles += (unsigned long long)((nf + dp->kp.TF - 1) / dp->kp.TF);
        if (u.d_f64) fast = false;
    }
    if (tiles > 0xFFFFFFF0ull) return set_err(D2P_ERR_INVALID_ARG, "too many tiles in one launch");

    void* base = nullptr;
    CK(noalloc_scratch_address(&base));                          // first call on a device: looks the array up
    unsigned char* scratch = (unsigned char*)base;
    DevStats* d_stats = (DevStats*)scratch;
    // the array still holds whatever the previous call left: the statistics must start from zero, job table and LUT
    // are overwritten in full for the part this launch reads
    CK(cudaMemsetAsync(d_stats, 0, sizeof(DevStats), stream));
    if (tiles) {
        // pageable sources: the runtime has taken its copy of `live` and `lut` when these calls return
        CK(cudaMemcpyAsync(scratch + kNoallocJobsOffset, live, sizeof(DevJob) * nlive, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(scratch + kNoallocLutOffset, lut.data(), lut_bytes, cudaMemcpyHostToDevice, stream));
        KParams k = dp->kp;
        k.lut = (const double*)(scratch + kNoallocLutOffset);
        k.jobs = (const DevJob*)(scratch + kNoallocJobsOffset);
        k.njobs = nlive;
        k.total_tiles = (unsigned int)tiles;
        k.stats = d_stats;
        k.job_stats = nullptr; k.events = nullptr; k.ev_cap = 0;
        k.fpd_jump = nullptr; k.fpd_items = 0;
        const unsigned long long want = (unsigned long long)dp->num_sms * dp->ctas_per_sm;
        const int grid = (int)std::min<unsigned long long>(tiles, want);
        CK(launch_kernels(k, kernel_variant(k, dp->lm, fast), grid, dp->threads, dp->smem, stream));
    }
    if (clipped_samples || peak_abs) {
        DevStats hs;
        CK(cudaMemcpyAsync(&hs, d_stats, sizeof hs, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        if (clipped_samples) *clipped_samples = (int64_t)(hs.clip_lo + hs.clip_hi);
        if (peak_abs) std::memcpy(peak_abs, &hs.peak_bits, 8);
    }
    return D2P_OK;
}

int d2p_debug_fpd_exponent(const double* samples, int n, int32_t* exponents, int device) {
    if ((!samples || !exponents) && n > 0) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (n <= 0) return D2P_OK;
    CK(cudaSetDevice(device));
    double* d_s = nullptr; int* d_e = nullptr;
    CK(cudaMalloc(&d_s, sizeof(double) * n));
    if (cudaMalloc(&d_e, sizeof(int) * n) != cudaSuccess) { cudaFree(d_s); return cuda_err(cudaGetLastError(), "cudaMalloc"); }
    cudaError_t e = cudaMemcpy(d_s, samples, sizeof(double) * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = launch_fpd_exponent(d_s, n, d_e, nullptr);
    if (e == cudaSuccess) e = cudaMemcpy(exponents, d_e, sizeof(int) * n, cudaMemcpyDeviceToHost);
    cudaFree(d_s); cudaFree(d_e);
    return e == cudaSuccess ? D2P_OK : cuda_err(e, "fpd exponent");
}

int d2p_batch_job_stats(d2p_batch* b, void* cuda_stream, int njobs, int64_t* clipped_samples, double* peak_abs) {
    if (!b || njobs < 0) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    if (njobs > b->launcher.jobstats_n) return set_err(D2P_ERR_INVALID_ARG, "more jobs asked for than the last run had");
    CK(cudaSetDevice(b->dp->plan.cfg.device));
    std::vector<DevJobStats> st((size_t)njobs);
    if (njobs) CK(cudaMemcpyAsync(st.data(), b->launcher.d_jobstats, sizeof(DevJobStats) * njobs, cudaMemcpyDeviceToHost, (cudaStream_t)cuda_stream));
    CK(cudaStreamSynchronize((cudaStream_t)cuda_stream));
    for (int i = 0; i < njobs; ++i) {
        if (clipped_samples) clipped_samples[i] = (int64_t)st[i].clipped;
        if (peak_abs) std::memcpy(&peak_abs[i], &st[i].peak_bits, 8);
    }
    return D2P_OK;
}

int d2p_batch_stats_reset(d2p_batch* b, void* cuda_stream) {
    if (!b) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    CK(cudaSetDevice(b->dp->plan.cfg.device));
    CK(cudaMemsetAsync(b->launcher.d_stats, 0, sizeof(DevStats), (cudaStream_t)cuda_stream));
    return D2P_OK;
}

int d2p_batch_stats(d2p_batch* b, void* cuda_stream, int64_t* clipped_samples, double* peak_abs) {
    if (!b) return set_err(D2P_ERR_INVALID_ARG, "null argument");
    CK(cudaSetDevice(b->dp->plan.cfg.device));
    DevStats st;
    CK(cudaMemcpyAsync(&st, b->launcher.d_stats, sizeof st, cudaMemcpyDeviceToHost, (cudaStream_t)cuda_stream));
    CK(cudaStreamSynchronize((cudaStream_t)cuda_stream));
    if (clipped_samples) *clipped_samples = (int64_t)(st.clip_lo + st.clip_hi);
    if (peak_abs) { double d; std::memcpy(&d, &st.peak_bits, 8); *peak_abs = d; }
    return D2P_OK;
}

}  // extern "C"
