// d2p_kernels.cu -- sm_100a kernels of the rdsd2pcm conversion core.
//
//   lutfir_kernel : integer-ratio path.  Unpack + byte-LUT linear-phase FIR decimation
//                   (BytePrecalcDecimator::process_bytes, rd/byte_precalc_decimator.rs:112-162, as the
//                   closed form of SURVEY.md A4) fused with scale/dither/quantise/pack
//                   (PcmWriter::write_to_buffer, rd/pcm_writer.rs:608-705).
//   lm_kernel     : rational L/M path.  Polyphase byte-LUT stage 1 (Stage1Poly, rd/lm_resampler.rs:408-640)
//                   + direct symmetric FIR stages 2/3 (DecimFIRSym, :645-752) chained through shared
//                   memory per output tile, same fused writer epilogue.
//
// Both are persistent kernels: each CTA copies the filter's LUT into shared memory once and then walks
// tiles of output frames.  A tile is computed from absolute input positions (halo included), so tiles,
// files and time blocks are independent -- no state is carried between them.
//
// Bit-exactness: every f64 operation is an explicit __dadd_rn/__dmul_rn in the reference's order
// (left folds; no FMA contraction; the file is also compiled with -fmad=false).
#include <cuda_runtime.h>
#include <stdint.h>

#include "d2p_device.h"
#include "d2p_launch.h"

namespace d2p {

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t bitrev_each_byte(uint32_t w) {   // MSB-first -> LSB-first
    return __byte_perm(__brev(w), 0, 0x0123);
}

// Philox4x32-10 (Salmon et al. 2011): counter-based noise keyed by (seed; frame, channel).
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ double u53(uint32_t hi, uint32_t lo) {
    const unsigned long long u = ((unsigned long long)hi << 32) | lo;
    return __dmul_rn((double)(u >> 11), 1.0 / 9007199254740992.0);
}

__device__ __forceinline__ const DevJob& find_job(const KParams& P, unsigned int tile) {
    int lo = 0, hi = P.njobs - 1;
    while (lo < hi) {                      // last job with tile_begin <= tile
        const int mid = (lo + hi + 1) >> 1;
        if (P.jobs[mid].tile_begin <= tile) lo = mid; else hi = mid - 1;
    }
    return P.jobs[lo];
}

// Byte offset inside a job's container-order buffer of per-channel stream position q (planar blocks).
__device__ __forceinline__ long long planar_offset(const KParams& P, const DevJob& J, int c, long long q) {
    if (q >= J.tail_q) return (J.tail_q - J.base_q) * P.C + (long long)c * J.tail_len + (q - J.tail_q);
    const long long qr = q - J.base_q;
    const long long blk = qr / P.block_size;
    const long long within = qr - blk * P.block_size;
    return (blk * P.C + c) * (long long)P.block_size + within;
}

// ------------------------------------------------------------------------------------------------
// A1: unpack.  Gather the per-channel, LSB-first byte tiles [s_lo, s_lo + nbytes) of all channels
// into shared memory straight from the container layout (de-interleave / de-block / bit-reverse fused
// into the load).  Positions outside [0, end_q) read as 0.  s_lo and nbytes are multiples of 16.
// ------------------------------------------------------------------------------------------------
__device__ void stage_input(const KParams& P, const DevJob& J, unsigned char* tiles, long long s_lo, int nbytes) {
    const int C = P.C;
    if (P.layout == 'P') {
        const int nvec = nbytes >> 4;
        const bool fast = ((P.block_size & 15) == 0) && ((((uintptr_t)J.in) & 15) == 0) && ((J.base_q & 15) == 0);
        for (int v = threadIdx.x; v < nvec * C; v += blockDim.x) {
            const int c = v / nvec, idx = v - c * nvec;
            const long long q = s_lo + 16LL * idx;
            uint4 val = make_uint4(0, 0, 0, 0);
            if (fast && q >= J.base_q && q + 16 <= J.end_q && q + 16 <= J.tail_q) {
                val = ld_nc_v4(J.in + planar_offset(P, J, c, q));
            } else if (q + 16 > 0 && q < J.end_q) {
                uint32_t w[4] = {0, 0, 0, 0};
                for (int b = 0; b < 16; ++b) {
                    const long long qq = q + b;
                    if (qq >= J.base_q && qq >= 0 && qq < J.end_q)
                        w[b >> 2] |= (uint32_t)J.in[planar_offset(P, J, c, qq)] << ((b & 3) * 8);
                }
                val = make_uint4(w[0], w[1], w[2], w[3]);
            }
            if (!P.lsb_first) {
                val.x = bitrev_each_byte(val.x); val.y = bitrev_each_byte(val.y);
                val.z = bitrev_each_byte(val.z); val.w = bitrev_each_byte(val.w);
            }
            *reinterpret_cast<uint4*>(tiles + (size_t)c * P.tile_stride + 16 * idx) = val;
        }
    } else {
        // byte-interleaved: consecutive threads read consecutive global bytes (coalesced), scatter to C tiles
        const int total = nbytes * C;
        const long long g0 = (s_lo - J.base_q) * C;
        for (int g = threadIdx.x; g < total; g += blockDim.x) {
            const int qi = g / C, c = g - qi * C;
            const long long q = s_lo + qi;
            unsigned int b = 0;
            if (q >= 0 && q >= J.base_q && q < J.end_q) b = J.in[g0 + g];
            if (!P.lsb_first) b = __brev(b) >> 24;
            tiles[(size_t)c * P.tile_stride + qi] = (unsigned char)b;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// A13-A17: scale -> dither -> quantise -> pack, plus the level / clip statistics.
// ------------------------------------------------------------------------------------------------
struct Emitter {
    double peak;
    unsigned int clip_lo, clip_hi;

    __device__ __forceinline__ void init() { peak = 0.0; clip_lo = 0; clip_hi = 0; }

    __device__ __forceinline__ void clip_event(const KParams& P, const DevJob& J, int ch, long long f, int high) {
        if (high) ++clip_hi; else ++clip_lo;
        if (P.events) {
            const unsigned int slot = atomicAdd(&P.stats->n_events, 1u);
            if (slot < P.ev_cap) {
                ClipEvent e; e.job = J.job_id; e.ch_kind = ch | (high << 16); e.frame = f;
                P.events[slot] = e;
            }
        }
    }

    // y: decimator output of channel ch, absolute frame f; fl: frame index inside the tile
    __device__ __forceinline__ void emit(const KParams& P, const DevJob& J, unsigned char* ostage,
                                         int ch, long long f, int fl, double y) {
        const long long fj = f - J.f_begin;
        if (J.f64) J.f64[(long long)ch * J.f64_stride + fj] = y;
        double s = __dmul_rn(y, P.scale);                       // scale_and_dither  rd/pcm_writer.rs:651
        peak = fmax(peak, fabs(s));                             // process_levels    rd/conversion_context.rs:235-240
        if (P.has_ds) s = __dmul_rn(s, P.neg_scale);            // Dither::process_samp rd/dither.rs:86-95
        if (P.dither == 'T' || P.dither == 'R') {
            uint32_t r[4];
            philox4x32_10((uint32_t)f, (uint32_t)((unsigned long long)f >> 32), (uint32_t)ch, 0u,
                          (uint32_t)J.seed, (uint32_t)(J.seed >> 32), r);
            if (P.dither == 'T') {
                const double r1 = __dmul_rn(u53(r[0], r[1]), 0.5);
                const double r2 = __dmul_rn(u53(r[2], r[3]), 0.5);
                s = __dadd_rn(s, __dsub_rn(r1, r2));
            } else {
                s = __dadd_rn(s, __dsub_rn(u53(r[0], r[1]), 0.5));
            }
        }
        if (P.has_ds) s = __dmul_rn(s, P.pos_scale);

        if (P.bits == 32) {
            const float v = __double2float_rn(s);               // pack_float / push_samp(q as f32)
            if (P.out_format == D2P_DEV_OUT_PLANAR) {
                reinterpret_cast<float*>(J.out)[(long long)ch * J.out_stride + fj] = v;
            } else {
                uint32_t u = __float_as_uint(v);
                if (P.out_format == D2P_DEV_OUT_BE) u = __byte_perm(u, 0, 0x0123);
                *reinterpret_cast<uint32_t*>(ostage + ((size_t)fl * P.C + ch) * 4) = u;
            }
            return;
        }
        // quantize / my_round / clamp_value   rd/pcm_writer.rs:678-705
        const double r = s < 0.0 ? floor(__dsub_rn(s, 0.5)) : floor(__dadd_rn(s, 0.5));
        int v = (int)(long long)__double2ll_rz(r);              // f64 -> i64 saturating (NaN -> 0), then i64 -> i32 wrapping
        const int mn = -P.peak, mx = P.peak - 1;
        if (v < mn) { v = mn; clip_event(P, J, ch, f, 0); }
        else if (v > mx) { v = mx; clip_event(P, J, ch, f, 1); }
        if (P.out_format == D2P_DEV_OUT_PLANAR) {
            reinterpret_cast<int*>(J.out)[(long long)ch * J.out_stride + fj] = v;
            return;
        }
        unsigned char* o = ostage + ((size_t)fl * P.C + ch) * P.bps;
        if (P.bps == 2) {                                        // pack_int  rd/pcm_writer.rs:372-396
            const unsigned short u = (unsigned short)v;
            *reinterpret_cast<unsigned short*>(o) = P.out_format == D2P_DEV_OUT_BE ? (unsigned short)((u >> 8) | (u << 8)) : u;
        } else {
            const unsigned int u = (unsigned int)(P.bits == 20 ? (v << 4) : v);
            if (P.out_format == D2P_DEV_OUT_BE) { o[0] = (unsigned char)(u >> 16); o[1] = (unsigned char)(u >> 8); o[2] = (unsigned char)u; }
            else { o[0] = (unsigned char)u; o[1] = (unsigned char)(u >> 8); o[2] = (unsigned char)(u >> 16); }
        }
    }

    __device__ __forceinline__ void flush(const KParams& P) {
        // non-negative doubles compare like their bit patterns
        unsigned long long pb = (unsigned long long)__double_as_longlong(peak);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, pb, o);
            pb = other > pb ? other : pb;
            clip_lo += __shfl_xor_sync(0xffffffffu, clip_lo, o);
            clip_hi += __shfl_xor_sync(0xffffffffu, clip_hi, o);
        }
        if ((threadIdx.x & 31) == 0) {
            if (pb) atomicMax(&P.stats->peak_bits, pb);
            if (clip_lo) atomicAdd(&P.stats->clip_lo, (unsigned long long)clip_lo);
            if (clip_hi) atomicAdd(&P.stats->clip_hi, (unsigned long long)clip_hi);
        }
    }
};

// Interleaved staging -> global, 16-byte vectors when the destination allows.
__device__ void copy_out(unsigned char* dst, const unsigned char* src, int nbytes) {
    if ((((uintptr_t)dst) & 15) == 0) {
        const int nv = nbytes >> 4;
        for (int i = threadIdx.x; i < nv; i += blockDim.x)
            reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
        for (int i = (nv << 4) + threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = src[i];
    } else {
        for (int i = threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = src[i];
    }
}

__device__ __forceinline__ void load_lut(double* dst, const double* __restrict__ src, int n) {
    const int nv = n >> 1;   // n is a multiple of 256
    for (int i = threadIdx.x; i < nv; i += blockDim.x)
        reinterpret_cast<double2*>(dst)[i] = reinterpret_cast<const double2*>(src)[i];
}

// ------------------------------------------------------------------------------------------------
// Integer-ratio LUT FIR.  Frame k (internal index) ends at byte j = (k+1)*bpo - 1 and is
//   y_k = fold_{i=0..nt-1} (acc + T_i[B[j-i]]) + T_i[bitrev8(B[j-(2nt-1-i)])]
// Bytes are fetched as aligned 32-bit shared-memory words and funnel-shifted into 4-byte groups; the
// mirrored half is bit-reversed 4 bytes at a time with one BREV (which also flips the byte order, so
// both halves are consumed from the top byte down).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double lutfir_frame(const double* __restrict__ lut, const uint32_t* __restrict__ tw,
                                               int jr, int nt) {
    const int a = jr + 1 - 4;                 // first byte of the newest 4-byte group
    int fw = a >> 2;
    const int fsh = (a & 3) * 8;
    uint32_t f_hi = tw[fw + 1];
    const int b = jr - (2 * nt - 1);          // oldest byte of the window
    int bw = b >> 2;
    const int bsh = (b & 3) * 8;
    uint32_t b_lo = tw[bw];
    double acc = 0.0;
    const double* T = lut;
    int i = 0;
    for (; i + 4 <= nt; i += 4) {
        const uint32_t f_lo = tw[fw];
        const uint32_t qf = __funnelshift_r(f_lo, f_hi, fsh);
        f_hi = f_lo; --fw;
        const uint32_t b_hi = tw[bw + 1];
        const uint32_t qb = __brev(__funnelshift_r(b_lo, b_hi, bsh));
        b_lo = b_hi; ++bw;
        acc = __dadd_rn(__dadd_rn(acc, T[qf >> 24]), T[qb >> 24]);
        acc = __dadd_rn(__dadd_rn(acc, T[256 + ((qf >> 16) & 0xFF)]), T[256 + ((qb >> 16) & 0xFF)]);
        acc = __dadd_rn(__dadd_rn(acc, T[512 + ((qf >> 8) & 0xFF)]), T[512 + ((qb >> 8) & 0xFF)]);
        acc = __dadd_rn(__dadd_rn(acc, T[768 + (qf & 0xFF)]), T[768 + (qb & 0xFF)]);
        T += 1024;
    }
    if (i < nt) {
        const int rem = nt - i;
        const uint32_t f_lo = tw[fw];
        const uint32_t qf = __funnelshift_r(f_lo, f_hi, fsh);
        const uint32_t b_hi = tw[bw + 1];
        const uint32_t qb = __brev(__funnelshift_r(b_lo, b_hi, bsh));
        acc = __dadd_rn(__dadd_rn(acc, T[qf >> 24]), T[qb >> 24]);
        if (rem > 1) acc = __dadd_rn(__dadd_rn(acc, T[256 + ((qf >> 16) & 0xFF)]), T[256 + ((qb >> 16) & 0xFF)]);
        if (rem > 2) acc = __dadd_rn(__dadd_rn(acc, T[512 + ((qf >> 8) & 0xFF)]), T[512 + ((qb >> 8) & 0xFF)]);
    }
    return acc;
}

__global__ void __launch_bounds__(kMaxThreads) lutfir_kernel(const __grid_constant__ KParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    double* lut = reinterpret_cast<double*>(smem);
    unsigned char* tiles = smem + (size_t)P.lut_doubles * 8;
    unsigned char* ostage = tiles + (size_t)P.C * P.tile_stride;

    load_lut(lut, P.lut, P.lut_doubles);
    Emitter em; em.init();

    for (unsigned int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
        const DevJob& J = find_job(P, tile);
        const long long f0 = J.f_begin + (long long)(tile - J.tile_begin) * P.TF;
        const long long left = J.f_begin + J.f_count - f0;
        const int nf = left < P.TF ? (int)left : P.TF;
        const long long k0 = f0 + (P.h - 1);                       // delay_count = h-1 frames are discarded (A4)
        const long long q_lo = (k0 + 1) * P.bpo - 2 * P.nt;        // oldest byte any frame of the tile needs
        const long long s_lo = q_lo & ~15LL;
        const long long q_hi = (k0 + nf) * P.bpo;
        const int nbytes = (int)(((q_hi - s_lo) + 15) & ~15LL) + 16;

        __syncthreads();                                           // previous tile fully consumed
        stage_input(P, J, tiles, s_lo, nbytes);
        __syncthreads();

        const int items = nf * P.C;
        for (int item = threadIdx.x; item < items; item += blockDim.x) {
            const int c = item / nf, fl = item - c * nf;           // a warp walks consecutive frames of one channel
            const uint32_t* tw = reinterpret_cast<const uint32_t*>(tiles + (size_t)c * P.tile_stride);
            const int jr = (int)((k0 + fl + 1) * P.bpo - 1 - s_lo);
            const double y = lutfir_frame(lut, tw, jr, P.nt);
            em.emit(P, J, ostage, c, f0 + fl, fl, y);
        }
        if (P.out_format != D2P_DEV_OUT_PLANAR) {
            __syncthreads();
            const size_t fb = (size_t)P.C * P.bps;
            copy_out(J.out + (size_t)(f0 - J.f_begin) * fb, ostage, (int)(nf * fb));
        }
    }
    em.flush(P);
}

// ------------------------------------------------------------------------------------------------
// Rational L/M path.
//   stage 1: output n sits at high-rate index n_hi (Plan::stage1_nhi); t = n_hi div L, p = n_hi mod L
//            y1[n] = fold_g LUT[p][g][ window8(t - (8g mod ring_cap)) ],  window8(t) bit j = x[t-j]
//            (pre-stream bits read 0; the modulo reproduces the reference's 32-entry byte ring)
//   stage s: y_s[k] = fold_{j<h} total + F[j]*(y[t-j] + y[t-(2h-1-j)]),  t = (h-1) + k*D, pre-stream = 0.0
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double fir_sym(const double* __restrict__ taps, const double* __restrict__ front,
                                          const double* __restrict__ back, int h) {
    double total = 0.0;
#pragma unroll 4
    for (int k = 0; k < h; ++k)
        total = __dadd_rn(total, __dmul_rn(taps[k], __dadd_rn(front[-k], back[k])));
    return total;
}

__global__ void __launch_bounds__(kMaxThreads) lm_kernel(const __grid_constant__ KParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    double* lut = reinterpret_cast<double*>(smem);
    double* s1 = lut + P.lut_doubles;
    double* s2 = s1 + P.s1_cap;
    unsigned char* tiles = reinterpret_cast<unsigned char*>(s2 + P.s2_cap);
    unsigned char* ostage = tiles + (size_t)P.C * P.tile_stride;

    __shared__ unsigned char sgroups[kMaxPhasesDev];
    if (threadIdx.x < kMaxPhasesDev) sgroups[threadIdx.x] = P.groups[threadIdx.x];
    load_lut(lut, P.lut, P.lut_doubles);
    Emitter em; em.init();

    for (unsigned int tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
        const DevJob& J = find_job(P, tile);
        const long long f0 = J.f_begin + (long long)(tile - J.tile_begin) * P.TF;
        const long long left = J.f_begin + J.f_count - f0;
        const int nf = left < P.TF ? (int)left : P.TF;

        // stage-2 output range feeding the tile (for 2 stages it *is* the tile)
        long long k2_lo, k2_hi;
        if (P.nstages == 3) { k2_lo = (P.h3 - 1) + f0 * P.D3 - (2 * P.h3 - 1); k2_hi = (P.h3 - 1) + (f0 + nf - 1) * P.D3; }
        else { k2_lo = f0; k2_hi = f0 + nf - 1; }
        const long long k2c = k2_lo > 0 ? k2_lo : 0;
        // stage-1 output range feeding those
        const long long n_lo = (P.h2 - 1) + k2c * P.D2 - (2 * P.h2 - 1);
        const long long n_hi = (P.h2 - 1) + k2_hi * P.D2;
        const int cnt1 = (int)(n_hi - n_lo + 1);
        const int cnt2 = (int)(k2_hi - k2_lo + 1);
        const long long nc = n_lo > 0 ? n_lo : 0;
        // input bits feeding those
        const long long nhi_first = P.accum ? (P.e0 + nc + 1) * 21 - 1 : (long long)(P.h1 - 1) + (long long)P.m1 * (nc + 1);
        const long long nhi_last = P.accum ? (P.e0 + n_hi + 1) * 21 - 1 : (long long)(P.h1 - 1) + (long long)P.m1 * (n_hi + 1);
        const long long t_first = nhi_first / P.L;
        const int p_first = (int)(nhi_first - t_first * P.L);
        const long long t_last = nhi_last / P.L;
        long long byte_lo = (t_first - 31) >> 3;                    // arithmetic shift: floor for negatives
        const long long s_lo = byte_lo & ~15LL;
        const int nbytes = (int)((((t_last >> 3) + 1 - s_lo) + 15) & ~15LL) + 16;
        const int bit_base = (int)(t_first - 31 - 8 * s_lo);        // tile-relative bit of the first window's oldest bit

        __syncthreads();
        stage_input(P, J, tiles, s_lo, nbytes);
        __syncthreads();

        for (int c = 0; c < P.C; ++c) {
            const uint32_t* tw = reinterpret_cast<const uint32_t*>(tiles + (size_t)c * P.tile_stride);
            // ---- stage 1 ----
            for (int idx = threadIdx.x; idx < cnt1; idx += blockDim.x) {
                const long long n = n_lo + idx;
                double y = 0.0;                                     // pre-stream stage-1 samples are 0.0 (ring zero-init)
                if (n >= 0) {
                    const unsigned int x = (unsigned int)p_first + (unsigned int)P.m1 * (unsigned int)(n - nc);
                    const unsigned int dt = x / (unsigned int)P.L;
                    const int p = (int)(x - dt * (unsigned int)P.L);
                    const int d = bit_base + (int)dt;
                    const int w = d >> 5;
                    const uint32_t z = __brev(__funnelshift_r(tw[w], tw[w + 1], d & 31));   // bit j = x[t-j]
                    const double* T = lut + (size_t)p * P.gmax * 256;
                    const int G = sgroups[p];
                    y = T[z & 0xFF];
                    for (int g = 1; g < G; ++g)
                        y = __dadd_rn(y, T[g * 256 + ((z >> ((8 * g) & P.ring_mask)) & 0xFF)]);
                }
                s1[idx] = y;
            }
            __syncthreads();
            // ---- stage 2 ----
            for (int idx = threadIdx.x; idx < cnt2; idx += blockDim.x) {
                const long long k2 = k2_lo + idx;
                double tot = 0.0;
                if (k2 >= 0) {
                    const int rel = (int)((P.h2 - 1) + k2 * P.D2 - n_lo);
                    tot = fir_sym(P.c2, s1 + rel, s1 + rel - (2 * P.h2 - 1), P.h2);
                }
                if (P.nstages == 2) em.emit(P, J, ostage, c, f0 + idx, idx, tot);
                else s2[idx] = tot;
            }
            // ---- stage 3 ----
            if (P.nstages == 3) {
                __syncthreads();
                for (int idx = threadIdx.x; idx < nf; idx += blockDim.x) {
                    const int rel = (int)((P.h3 - 1) + (f0 + idx) * P.D3 - k2_lo);
                    const double y = fir_sym(P.c3, s2 + rel, s2 + rel - (2 * P.h3 - 1), P.h3);
                    em.emit(P, J, ostage, c, f0 + idx, idx, y);
                }
            }
            __syncthreads();                                        // s1/s2 are reused by the next channel
        }
        if (P.out_format != D2P_DEV_OUT_PLANAR) {
            const size_t fb = (size_t)P.C * P.bps;
            copy_out(J.out + (size_t)(f0 - J.f_begin) * fb, ostage, (int)(nf * fb));
        }
    }
    em.flush(P);
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
const char* kernel_name(bool lm) { return lm ? "d2p::lm_kernel" : "d2p::lutfir_kernel"; }

static cudaError_t ensure_smem_attr(bool lm) {
    static bool attr_done[2][64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const void* fn = lm ? (const void*)lm_kernel : (const void*)lutfir_kernel;
    if (dev < 64 && !attr_done[lm ? 1 : 0][dev]) {
        // the opt-in limit (227 KiB) covers static + dynamic shared memory: leave room for the kernels' statics
        e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem - 1024);
        if (e != cudaSuccess) { cudaGetLastError(); return e; }
        attr_done[lm ? 1 : 0][dev] = true;
    }
    return cudaSuccess;
}

int max_ctas_per_sm(bool lm, int threads, size_t smem) {
    if (ensure_smem_attr(lm) != cudaSuccess) return 0;
    int n = 0;
    const void* fn = lm ? (const void*)lm_kernel : (const void*)lutfir_kernel;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, threads, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

cudaError_t launch_kernels(const KParams& P, bool lm, int grid, int threads, size_t smem, cudaStream_t stream) {
    cudaError_t e = ensure_smem_attr(lm);
    if (e != cudaSuccess) return e;
    if (lm) lm_kernel<<<grid, threads, smem, stream>>>(P);
    else lutfir_kernel<<<grid, threads, smem, stream>>>(P);
    return cudaGetLastError();
}

}  // namespace d2p
