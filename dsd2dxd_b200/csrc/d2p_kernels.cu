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
// Both are persistent kernels (one CTA per SM): each CTA copies the filter's LUT into shared memory once and then
// walks tiles of output frames.  A tile is computed from absolute input positions (halo included), so tiles,
// files and time blocks are independent -- no state is carried between them.
//
// Bit-exactness: every f64 operation is an explicit __dadd_rn/__dmul_rn in the reference's order
// (left folds; no FMA contraction; the file is also compiled with -fmad=false).
#include <cuda_runtime.h>
#include <limits.h>
#include <stdint.h>

#include <mutex>

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
// 16-byte asynchronous global -> shared copy (LDGSTS): the issuing warp does not wait for HBM
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// ---- TMA bulk copies (cp.async.bulk, SASS: UBLKCP) completed on an mbarrier (SYNCS) ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
// global -> shared, 16-byte aligned on both sides, size a multiple of 16; completion is counted on `bar`
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global, same alignment rules; tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void tma_store_1d(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory made visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint32_t bitrev_each_byte(uint32_t w) {   // MSB-first -> LSB-first
    return __byte_perm(__brev(w), 0, 0x0123);
}

// Philox2x32-10 (Salmon et al., SC'11): counter-based noise keyed by (seed; frame, channel).  Two 32-bit
// words per sample are all TPDF / rectangular dither need (same definition as oracle/oracle.cpp).
__device__ __forceinline__ void philox2x32_10(unsigned long long frame, uint32_t chan, uint32_t key, uint32_t& a, uint32_t& b) {
    const uint32_t M = 0xD256D193u, W = 0x9E3779B9u;
    uint32_t c0 = (uint32_t)frame, c1 = ((uint32_t)(frame >> 32) << 8) ^ chan;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi = __umulhi(M, c0), lo = M * c0;
        c0 = hi ^ key ^ c1; c1 = lo;
        key += W;
    }
    a = c0; b = c1;
}
__device__ __forceinline__ uint32_t philox_key(unsigned long long seed) {
    return (uint32_t)seed ^ ((uint32_t)(seed >> 32) * 0x85EBCA6Bu);
}
// (double)u for a 32-bit u without a conversion instruction: 2^52 + u is exact, differences too
__device__ __forceinline__ double u32_as_2p52(uint32_t u) { return __hiloint2double(0x43300000, (int)u); }
__device__ __forceinline__ double dither_noise(int kind, unsigned long long frame, uint32_t chan, uint32_t key) {
    uint32_t a, b;
    philox2x32_10(frame, chan, key, a, b);
    if (kind == 'T')      // r1 - r2 with r = u * 2^-33 in [0, 0.5): the difference of the integers is exact
        return __dmul_rn(__dsub_rn(u32_as_2p52(a), u32_as_2p52(b)), 1.0 / 8589934592.0);
    const unsigned long long u = (((unsigned long long)a << 32) | b) >> 11;   // 53 bits -> [0,1)
    return __dsub_rn(__dmul_rn((double)u, 1.0 / 9007199254740992.0), 0.5);
}

// Dither::fpdither's octave (rd/dither.rs:102-104): `s.abs().log2().floor() as i32` -- NaN -> 0, saturating.
// Evaluated WITHOUT a device log2: libdevice's log2 is not exact at powers of two (log2(2^-59) comes out just below
// -59) and within an ulp everywhere else, so the floor lands in the wrong octave on and next to every octave edge.
// The reference's value is floor(RN(log2 |s|)) with the host libm's (correctly rounded near the edges) log2:
//   |s| = m * 2^k, 1 <= m < 2  ->  k, except that for |s| only j doubles below 2^(k+1) the logarithm
//   (k+1) - j * 2^-53 / ln 2 rounds UP to k+1 when it is closer to k+1 than to the double in front of it, i.e. when
//   j <= floor(2^q * ln 2) with 2^(q-52) the spacing of the doubles on that side of k+1.
__device__ __forceinline__ int fpd_exponent(double s) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(s) & 0x7FFFFFFFFFFFFFFFull;
    if (b > 0x7FF0000000000000ull) return 0;                       // NaN as i32
    if (b == 0x7FF0000000000000ull) return INT_MAX;                // log2(inf) = inf, saturates
    if (b == 0ull) return INT_MIN;                                 // log2(0) = -inf, saturates
    const int ex = (int)(b >> 52);
    unsigned long long man = b & 0xFFFFFFFFFFFFFull;
    int k;
    if (ex == 0) { const int lz = __clzll((long long)man) - 11; man = (man << lz) & 0xFFFFFFFFFFFFFull; k = -1022 - lz; }
    else k = ex - 1023;
    const unsigned long long j = 0x10000000000000ull - man;        // mantissa steps up to the next power of two
    const int n = k + 1;
    if (n == 0 || j > 709ull) return k;
    const unsigned int an = n < 0 ? (unsigned int)(-n) : (unsigned int)n;
    int q = 31 - __clz((int)an);
    if (n > 0 && (an & (an - 1)) == 0) --q;                        // below a positive power of two the doubles are twice as dense
    if (q < 0) return k;
    const unsigned int jmax = q == 0 ? 0u : q == 1 ? 1u : q == 2 ? 2u : q == 3 ? 5u : q == 4 ? 11u : q == 5 ? 22u : q == 6 ? 44u
                            : q == 7 ? 88u : q == 8 ? 177u : q == 9 ? 354u : 709u;       // floor(2^q * ln 2)
    return j <= jmax ? n : k;
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
    long long blk;
    if (P.block_shift >= 0) blk = qr >> P.block_shift;          // 4096 (DSF) and every power of two: no 64-bit division
    else blk = qr / P.block_size;
    const long long within = qr - blk * P.block_size;
    return (blk * P.C + c) * (long long)P.block_size + within;
}

// Byte-interleaved payload (DFF / raw -f I): frame-byte q of channel c sits at global byte (q - base_q)*C + c.
// One work item = 4 consecutive stream bytes of ALL channels = 4*C contiguous global bytes: loaded as C
// aligned words (consecutive threads -> consecutive chunks, coalesced), transposed in registers into one
// word per channel, bit-reversed if the source is MSB-first, stored as one word per channel tile.
// CT = 0: generic (any channel count / unaligned source), byte-wise.
template <int CT>
__device__ void stage_interleaved(const KParams& P, const DevJob& J, unsigned char* tiles, long long s_lo, int nbytes,
                                  int tid, int nthr) {
    const int C = P.C;
    const long long g0 = (s_lo - J.base_q) * C;
    if (CT > 0 && ((((uintptr_t)J.in) & 3) == 0) && ((g0 & 3) == 0)) {
        const int nquad = nbytes >> 2;
        for (int w = tid; w < nquad; w += nthr) {
            const long long q = s_lo + 4LL * w;
            uint32_t O[CT > 0 ? CT : 1];
            if (q >= J.base_q && q >= 0 && q + 4 <= J.end_q) {
                uint32_t W[CT > 0 ? CT : 1];
                const uint32_t* src = reinterpret_cast<const uint32_t*>(J.in + g0 + 4LL * w * CT);
#pragma unroll
                for (int i = 0; i < CT; ++i) W[i] = __ldg(src + i);
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                    uint32_t o = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int k = c + CT * j;
                        o |= ((W[k >> 2] >> (8 * (k & 3))) & 0xFFu) << (8 * j);
                    }
                    O[c] = o;
                }
            } else {
#pragma unroll
                for (int c = 0; c < CT; ++c) O[c] = 0;
                if (q + 4 > 0 && q + 4 > J.base_q && q < J.end_q) {
                    for (int j = 0; j < 4; ++j) {
                        const long long qq = q + j;
                        if (qq >= 0 && qq >= J.base_q && qq < J.end_q) {
#pragma unroll
                            for (int c = 0; c < CT; ++c) O[c] |= (uint32_t)J.in[(qq - J.base_q) * CT + c] << (8 * j);
                        }
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < CT; ++c) {
                const uint32_t o = P.lsb_first ? O[c] : bitrev_each_byte(O[c]);
                *reinterpret_cast<uint32_t*>(tiles + (size_t)c * P.tile_stride + 4 * w) = o;
            }
        }
        return;
    }
    const int total = nbytes * C;
    for (int g = tid; g < total; g += nthr) {
        const int qi = g / C, c = g - qi * C;
        const long long q = s_lo + qi;
        unsigned int b = 0;
        if (q >= 0 && q >= J.base_q && q < J.end_q) b = J.in[g0 + g];
        if (!P.lsb_first) b = __brev(b) >> 24;
        tiles[(size_t)c * P.tile_stride + qi] = (unsigned char)b;
    }
}

// ------------------------------------------------------------------------------------------------
// A1: unpack.  Gather the per-channel, LSB-first byte tiles [s_lo, s_lo + nbytes) of all channels
// into shared memory straight from the container layout (de-interleave / de-block / bit-reverse fused
// into the load).  Positions outside [0, end_q) read as 0.  s_lo and nbytes are multiples of 16.
// ------------------------------------------------------------------------------------------------
// bar != nullptr: the untransformed interior of a DSF-style payload is moved by TMA bulk copies (one per channel and
// 4096-byte block run) that complete on *bar; thread 0 arms the barrier with the byte count of this tile (0 if the
// layout needs the register path), so a waiter on *bar never hangs.  bar == nullptr: 16-byte cp.async copies.
__device__ void stage_input(const KParams& P, const DevJob& J, unsigned char* tiles, long long s_lo, int nbytes,
                            int tid, int nthr, unsigned long long* bar = nullptr) {
    const int C = P.C;
    if (P.layout != 'P' && bar && tid == 0) mbar_arrive_expect_tx(bar, 0u);
    if (P.layout == 'P') {
        const int nvec = nbytes >> 4;
        const bool fast = ((P.block_size & 15) == 0) && ((((uintptr_t)J.in) & 15) == 0) && ((J.base_q & 15) == 0);
        // Interior chunks [i_lo, i_hi) of an LSB-first payload with power-of-two blocks (DSF): whole 16-byte pieces of
        // the stream that need no transformation.  One 64-bit base per tile and channel, 32-bit arithmetic per chunk,
        // asynchronous copies.  Only the few edge chunks (stream start / end, ragged tail) take the general loop.
        int i_lo = 0, i_hi = 0;
        if (fast && P.lsb_first && P.block_shift >= 4) {
            const long long lo_q = J.base_q > 0 ? J.base_q : 0;
            const long long hi_q = J.end_q < J.tail_q ? J.end_q : J.tail_q;
            long long a = (lo_q - s_lo + 15) >> 4, b = (hi_q - s_lo) >> 4;
            a = a < 0 ? 0 : (a > nvec ? nvec : a);
            b = b < a ? a : (b > nvec ? nvec : b);
            i_lo = (int)a; i_hi = (int)b;
            if (bar && i_hi > i_lo) {
                const int nint = i_hi - i_lo, bs = P.block_shift;
                const long long qr0 = s_lo + 16LL * i_lo - J.base_q;
                const long long blk0 = qr0 >> bs;
                const int w0 = (int)(qr0 & (P.block_size - 1)), w1 = w0 + 16 * nint;
                const int nruns = ((w1 - 1) >> bs) + 1;               // block runs the interior touches
                for (int it = tid; it < nruns * C; it += nthr) {
                    const int c = it / nruns, k = it - c * nruns;
                    const int lo = k == 0 ? w0 : (k << bs), hi = w1 < ((k + 1) << bs) ? w1 : ((k + 1) << bs);
                    tma_load_1d(tiles + (size_t)c * P.tile_stride + 16 * i_lo + (lo - w0),
                                J.in + (((blk0 + k) * C + c) << bs) + (lo - (k << bs)), (uint32_t)(hi - lo), bar);
                }
            } else if (i_hi > i_lo) {
                const long long qr0 = s_lo + 16LL * i_lo - J.base_q;
                const long long blk0 = qr0 >> P.block_shift;
                const int w0 = (int)(qr0 & (P.block_size - 1)), mask = P.block_size - 1;
                for (int c = 0; c < C; ++c) {
                    const unsigned char* src0 = J.in + ((blk0 * C + c) << P.block_shift);
                    unsigned char* dst = tiles + (size_t)c * P.tile_stride + 16 * i_lo;
                    for (int i = tid; i < i_hi - i_lo; i += nthr) {
                        const int r = w0 + 16 * i;
                        cp_async16(dst + 16 * i, src0 + ((size_t)((r >> P.block_shift) * C) << P.block_shift) + (r & mask));
                    }
                }
            }
        }
        if (bar && tid == 0) mbar_arrive_expect_tx(bar, (uint32_t)((i_hi - i_lo) * 16 * C));
        const int nedge = nvec - (i_hi - i_lo);
        for (int v = tid; v < nedge * C; v += nthr) {
            const int c = v / nedge, e = v - c * nedge;
            const int idx = e < i_lo ? e : e + (i_hi - i_lo);
            const long long q = s_lo + 16LL * idx;
            uint4 val = make_uint4(0, 0, 0, 0);
            if (fast && q >= J.base_q && q + 16 <= J.end_q && q + 16 <= J.tail_q) {
                if (P.lsb_first) {         // no transformation needed: asynchronous copy, completed by cp_async_wait_all()
                    cp_async16(tiles + (size_t)c * P.tile_stride + 16 * idx, J.in + planar_offset(P, J, c, q));
                    continue;
                }
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
        switch (C) {
        case 1: stage_interleaved<1>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 2: stage_interleaved<2>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 3: stage_interleaved<3>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 4: stage_interleaved<4>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 5: stage_interleaved<5>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 6: stage_interleaved<6>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 7: stage_interleaved<7>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        case 8: stage_interleaved<8>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        default: stage_interleaved<0>(P, J, tiles, s_lo, nbytes, tid, nthr); break;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// A13-A17: scale -> dither -> quantise -> pack, plus the level / clip statistics.
// ------------------------------------------------------------------------------------------------
// WM (writer mode): 0 = generic (every option, decided at run time);
//   1 = the common file/stdout case decided on the host: 3-byte little-endian interleaved integers (20/24 bit),
//       no f64 side output, no DSD2DXD_DITHERSCALE, |sample*scale| provably < 2^31 - 1;  DITH = 0 / 'T' / 'R'
//       is then a compile-time constant too.
template <int WM, int DITH>
struct Emitter {
    unsigned long long peakb;          // max |sample*scale| as f64 bits (non-negative doubles order as integers)
    unsigned int clip_lo, clip_hi;
    uint32_t key;
    int job;                           // job whose statistics are being accumulated (per-job table only)

    __device__ __forceinline__ void init() { peakb = 0ull; clip_lo = 0; clip_hi = 0; key = 0; job = -1; }
    // called by every consumer thread of the CTA for every tile (warp-uniform): with a per-job table the
    // accumulators are handed over whenever the tile belongs to another job (dsd_levels: one peak per file)
    __device__ __forceinline__ void set_job(const KParams& P, const DevJob& J) {
        if (P.job_stats && J.job_id != job) { if (job >= 0) flush(P); job = J.job_id; }
        key = philox_key(J.seed);
    }

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

    // WM == 1 in two halves, so that a caller can run the pure arithmetic of several samples side by side:
    // quant = scale, level, dither, round (no memory access, no branch); put = clamp, clip statistics, pack.
    // live: the frame exists (lanes past the end of a ragged tile compute along but must not reach the statistics)
    __device__ __forceinline__ int quant(const KParams& P, int ch, long long f, double y, bool live) {
        double s = __dmul_rn(y, P.scale);                       // scale_and_dither  rd/pcm_writer.rs:651
        const unsigned long long ab = (unsigned long long)__double_as_longlong(fabs(s));
        peakb = (live && ab > peakb) ? ab : peakb;              // process_levels    rd/conversion_context.rs:235-240
        if (DITH) s = __dadd_rn(s, dither_noise(DITH, (unsigned long long)f, (uint32_t)ch, key));
        // quantize / my_round   rd/pcm_writer.rs:678-705 ; floor fused into the conversion
        return __double2int_rd(__dadd_rn(s, s < 0.0 ? -0.5 : 0.5));
    }
    __device__ __forceinline__ void put(const KParams& P, const DevJob& J, unsigned char* ostage, int ch, long long f, int fl, int v) {
        const int mn = -P.peak, mx = P.peak - 1;                // clamp_value  rd/pcm_writer.rs:685
        if (v < mn) { v = mn; clip_event(P, J, ch, f, 0); }
        else if (v > mx) { v = mx; clip_event(P, J, ch, f, 1); }
        const unsigned int u = (unsigned int)v << P.shl20;      // 20-bit: left-aligned in the 3-byte container
        unsigned char* o = ostage + (fl * P.C + ch) * 3;        // pack_int  rd/pcm_writer.rs:372-396
        o[0] = (unsigned char)u; o[1] = (unsigned char)(u >> 8); o[2] = (unsigned char)(u >> 16);
    }

    // y: decimator output of channel ch, absolute frame f; fl: frame index inside the tile
    __device__ __forceinline__ void emit(const KParams& P, const DevJob& J, unsigned char* ostage,
                                         int ch, long long f, int fl, double y) {
        if (WM == 1) { put(P, J, ostage, ch, f, fl, quant(P, ch, f, y, true)); return; }
        const long long fj = f - J.f_begin;
        if (J.f64) J.f64[(long long)ch * J.f64_stride + fj] = y;
        double s = __dmul_rn(y, P.scale);                       // scale_and_dither  rd/pcm_writer.rs:651
        const unsigned long long ab = (unsigned long long)__double_as_longlong(fabs(s));
        peakb = ab > peakb ? ab : peakb;
        if (P.out_format == D2P_DEV_OUT_NONE) return;           // level check (process_levels): nothing is written
        if (P.has_ds) s = __dmul_rn(s, P.neg_scale);            // Dither::process_samp rd/dither.rs:86-95
        if (P.dither == 'T' || P.dither == 'R')
            s = __dadd_rn(s, dither_noise(P.dither, (unsigned long long)f, (uint32_t)ch, key));
        else if (P.dither == 'F' && J.fpd) {                    // Dither::fpdither  rd/dither.rs:102-110
            const uint32_t fpd = J.fpd[(long long)ch * J.fpd_stride + fj];
            const int e = fpd_exponent(s);
            long long e2 = (long long)e + 62;
            e2 = e2 > 4096 ? 4096 : (e2 < -4096 ? -4096 : e2);   // 2^e2 is 0 / inf beyond the f64 range either way
            s = __dadd_rn(s, __dmul_rn(__dmul_rn((double)fpd, 3.4e-36), ldexp(1.0, (int)e2)));
            s = (double)__double2float_rn(s);
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
        const double r = floor(__dadd_rn(s, s < 0.0 ? -0.5 : 0.5));   // x - 0.5 == x + (-0.5) exactly
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
        unsigned long long pb = peakb;
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
            if (P.job_stats && job >= 0) {
                if (pb) atomicMax(&P.job_stats[job].peak_bits, pb);
                if (clip_lo | clip_hi) atomicAdd(&P.job_stats[job].clipped, (unsigned long long)clip_lo + clip_hi);
            }
        }
        peakb = 0ull; clip_lo = 0; clip_hi = 0;
    }
};

// Interleaved staging -> global, 16-byte vectors when the destination allows.
__device__ void copy_out(unsigned char* dst, const unsigned char* src, int nbytes, int tid, int nthr) {
    if ((((uintptr_t)dst) & 15) == 0) {
        const int nv = nbytes >> 4;
        for (int i = tid; i < nv; i += nthr)
            reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
        for (int i = (nv << 4) + tid; i < nbytes; i += nthr) dst[i] = src[i];
    } else {
        for (int i = tid; i < nbytes; i += nthr) dst[i] = src[i];
    }
}

// The packed tile is one contiguous run of the output stream: one TMA bulk store (shared -> global) issued by thread 0
// for the 16-byte multiple of it when the destination is 16-byte aligned; the other threads copy what is left.
__device__ void copy_out_tma(unsigned char* dst, const unsigned char* src, int nbytes, int tid, int nthr) {
    if ((((uintptr_t)dst) & 15) != 0 || nbytes < 16) { copy_out(dst, src, nbytes, tid, nthr); return; }
    const int nb16 = nbytes & ~15;
    if (tid == 0) { tma_store_1d(dst, src, (uint32_t)nb16); tma_store_commit(); }
    for (int i = nb16 + tid; i < nbytes; i += nthr) dst[i] = src[i];
}

__device__ __forceinline__ void load_lut(double* dst, const double* __restrict__ src, int n) {
    const int nv = n >> 1;   // n is a multiple of 256
    for (int i = threadIdx.x; i < nv; i += blockDim.x)
        reinterpret_cast<double2*>(dst)[i] = reinterpret_cast<const double2*>(src)[i];
}

// ------------------------------------------------------------------------------------------------
// Integer-ratio LUT FIR.  Frame k (internal index) ends at byte j = (k+1)*bpo - 1 and is
//   y_k = fold_{i=0..nt-1} (acc + T_i[B[j-i]]) + T_i[bitrev8(B[j-(2nt-1-i)])]
//
// The binding unit is the shared-memory crossbar (128 B/clk/SM): 2*nt random 8-byte look-ups per frame.
// A warp-wide LDS.64 is served as two half-warp wavefronts when its 16 lanes hit 16 distinct bank pairs,
// but random table indices give ~4.7 wavefronts (ncu, profiles/r01_lutfir_v1_*).  So the tables are laid
// out bank-aware and the lanes are skewed in time:
//   * entry (replica r, table i, byte b) sits at b*row + (st[r] + i)*8, i.e. its bank pair depends only on
//     (r, i), never on the data byte;
//   * lane (r, c) of a half-warp (R = 16/K replicas x K skews) runs table i at step u = i + c, so at any
//     step the 16 lanes address 16 different bank pairs: every look-up costs exactly 2 wavefronts;
//   * the fold order per frame is untouched (each lane still walks i = 0..nt-1), so results stay
//     bit-exact; the price is K-1 idle steps per frame set, which read 0.0 from a zero region in front of the LUT at
//     the bank pair the live look-up would have used (acc + 0.0 == acc exactly, no predicated arithmetic).
// Input bytes are fetched as aligned 32-bit words and funnel-shifted into 4-byte groups; the mirrored half
// is bit-reversed 4 bytes at a time with one BREV (which also flips the byte order, so both halves are
// consumed from the top byte down).
//
// The CTA is warp-specialised: P.nc consumer threads only touch shared memory (LUT, input tile, output
// staging); the remaining threads stage the next input tile -- TMA bulk copies completed on an mbarrier for the
// untransformed interior of DSF-style payloads, register transposes for MSB-first / byte-interleaved sources -- and
// write the previous tile's packed PCM out with one bulk store, over double-buffered tiles: one __syncthreads per tile.
// ------------------------------------------------------------------------------------------------
struct TileInfo {
    const DevJob* J;
    long long f0, s_lo;
    int nf, nbytes, jr_base;
};

__device__ __forceinline__ TileInfo lutfir_tile(const KParams& P, unsigned int tile) {
    TileInfo t;
    const DevJob& J = find_job(P, tile);
    t.J = &J;
    t.f0 = J.f_begin + (long long)(tile - J.tile_begin) * P.TF;
    const long long left = J.f_begin + J.f_count - t.f0;
    t.nf = left < P.TF ? (int)left : P.TF;
    const long long k0 = t.f0 + (P.h - 1);                        // delay_count = h-1 frames are discarded (A4)
    const long long q_lo = (k0 + 1) * P.bpo - 2 * P.nt - 16;      // oldest byte any (skewed) lane of the tile touches
    t.s_lo = q_lo & ~15LL;
    const long long q_hi = (k0 + t.nf) * P.bpo + 32;
    t.nbytes = (int)(((q_hi - t.s_lo) + 15) & ~15LL);
    t.jr_base = (int)(k0 * P.bpo - t.s_lo);                       // frame fl of the tile ends at byte (fl+1)*bpo-1+jr_base
    return t;
}

template <int OFF>
__device__ __forceinline__ double lds_f64(uint32_t saddr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(saddr), "n"(OFF));
    return v;
}

// Per-lane constants of the look-up schedule.
//   rlo / rhi : LUT row pitch in the low / high 16 bits: IDP.2A (dp2a) then extracts one byte of a data word,
//               multiplies it by the pitch and adds the lane's slot address in ONE instruction
//               (bytes 3,2 of the word: dp2a.hi with rhi / rlo; bytes 1,0: dp2a.lo with rhi / rlo);
//   dz        : distance from the lane's LUT slot to its slot in the zero region.  An idle (skew) step uses
//               pitch 0 and slot + dz: it reads 0.0 from the SAME bank pair the live look-up would have
//               hit, so idle lanes never add wavefronts, and acc + 0.0 == acc exactly (acc is never -0.0:
//               it starts at +0.0 and x + y is -0.0 only if both are).
struct LaneLut { uint32_t rlo, rhi; int dz, c; };

// F frames of one lane in lockstep ("frame set").  The frames are neighbours in time: frame f+1 ends 4*DW bytes
// after frame f (DW words; bpo < 4 packs 4/bpo frames into a word, those go to neighbouring lanes).  All frames
// of the set sit at the same table index at every step, so the skew guard, the slot address and the loop
// bookkeeping are paid once per set and the F dependent DADD chains interleave.  Because the frames are one (DW)
// word apart, their 4-byte groups are one sliding window over the byte stream: group g of frame f is the word
// that frame f-1 used at group g-DW (forward half), f+1 at g-DW (mirrored half).  Per group and half a lane
// therefore loads ONE tile word, funnel-shifts once (and bit-reverses once for the mirrored half), whatever F is.
//
// Where a lane's set starts in its channel tile: word pointers and funnel shifts of the two byte streams.
//   forward half : step u reads byte jr + c - u  (group g: bytes a-4g .. a-4g+3 with a = jr + c - 3)
//   mirrored half: step u reads byte jr - (2nt-1) - c + u
// jr: tile-relative index of the newest byte of the set's first frame; c: the lane's skew.
template <int F, int DW>
struct FrameSet {
    static constexpr int WS = (F - 1) * DW + 1;   // window: frame f reads logical entry f*DW
    // The windows are ring buffers in registers.  After n advances (rotation ROT = n mod WS) logical entry j of the
    // forward window is qf[(j - ROT) mod WS] and of the mirrored window qb[(j + ROT) mod WS]: an advance overwrites
    // exactly the entry that drops out, nothing moves.  ROT is a compile-time constant of every group, so the main
    // loop is unrolled WS groups deep and all indexing stays static.
    uint32_t qf[WS], qb[WS];                      // forward 4-byte groups; mirrored ones, bit-reversed
    uint32_t wf_lo, wf_nx, wb_hi, wb_nx;          // raw tile words at the window edges and the prefetched next ones
    const uint32_t* pf;
    const uint32_t* pb;
    int fsh, bsh;
    double acc[F];
    uint32_t tp, rlo, rhi, dz;
    int i0, nt;

    __device__ __forceinline__ void init(const KParams& P, const LaneLut& L, const uint32_t* tw, int jr, uint32_t tp0) {
        nt = P.nt;
        const int c = L.c;
        const int a = jr + c - 3, b = jr - (2 * nt - 1) - c;
        pf = tw + (a >> 2); pb = tw + (b >> 2);
        fsh = (a & 3) * 8; bsh = (b & 3) * 8;
        uint32_t w[WS + 1];
#pragma unroll
        for (int j = 0; j <= WS; ++j) w[j] = pf[j];
#pragma unroll
        for (int j = 0; j < WS; ++j) qf[j] = __funnelshift_r(w[j], w[j + 1], fsh);
        wf_lo = w[0]; wf_nx = pf[-1];
#pragma unroll
        for (int j = 0; j <= WS; ++j) w[j] = pb[j];
#pragma unroll
        for (int j = 0; j < WS; ++j) qb[j] = __brev(__funnelshift_r(w[j], w[j + 1], bsh));
        wb_hi = w[WS]; wb_nx = pb[WS + 1];
#pragma unroll
        for (int f = 0; f < F; ++f) acc[f] = 0.0;
        tp = tp0; rlo = L.rlo; rhi = L.rhi; dz = (uint32_t)L.dz;
        i0 = -c;
    }

    template <bool GUARD, int D, int ROT>
    __device__ __forceinline__ void step() {
        uint32_t rsel = (D & 1) ? rlo : rhi;
        uint32_t t = tp;
        if (GUARD) {
            const bool live = (unsigned)(i0 + D) < (unsigned)nt;
            rsel = live ? rsel : 0u;
            t = live ? tp : tp + dz;
        }
        double tf[F], tb[F];
#pragma unroll
        for (int f = 0; f < F; ++f) {
            const uint32_t vf = qf[(f * DW - ROT + WS) % WS], vb = qb[(f * DW + ROT) % WS];
            const uint32_t af = D < 2 ? __dp2a_hi(rsel, vf, t) : __dp2a_lo(rsel, vf, t);
            const uint32_t ab = D < 2 ? __dp2a_hi(rsel, vb, t) : __dp2a_lo(rsel, vb, t);
            tf[f] = lds_f64<8 * D>(af); tb[f] = lds_f64<8 * D>(ab);
        }
#pragma unroll
        for (int f = 0; f < F; ++f) acc[f] = __dadd_rn(__dadd_rn(acc[f], tf[f]), tb[f]);
    }

    // the words of the NEXT group are fetched one group ahead, so that a group never waits on its own tile loads
    template <int ROT>
    __device__ __forceinline__ void advance() {
        const uint32_t nqf = __funnelshift_r(wf_nx, wf_lo, fsh);
        const uint32_t nqb = __brev(__funnelshift_r(wb_hi, wb_nx, bsh));
        wf_lo = wf_nx; wb_hi = wb_nx;
        --pf; ++pb;
        wf_nx = pf[-1]; wb_nx = pb[WS + 1];
        qf[(WS - 1 - ROT + WS) % WS] = nqf;       // becomes logical entry 0 at rotation ROT + 1
        qb[ROT % WS] = nqb;                       // becomes logical entry WS - 1 at rotation ROT + 1
        tp += 32; i0 += 4;
    }

    template <bool GUARD, int ROT>
    __device__ __forceinline__ void group() {
        step<GUARD, 0, ROT>(); step<GUARD, 1, ROT>(); step<GUARD, 2, ROT>(); step<GUARD, 3, ROT>();
        advance<ROT>();
    }
    // WS - ROT groups starting at rotation ROT (ends at rotation 0)
    template <bool GUARD, int ROT>
    __device__ __forceinline__ void groups_to_wrap() {
        group<GUARD, ROT>();
        if constexpr (ROT + 1 < WS) groups_to_wrap<GUARD, ROT + 1>();
    }
    // n < WS groups starting at rotation 0, then the entries are moved back to rotation 0
    template <bool GUARD>
    __device__ __forceinline__ void groups_few(int n) {
        for (int k = 0; k < n; ++k) {
            group<GUARD, 0>();
            // rotation 1 -> 0: logical j sits at qf[(j - 1) mod WS], qb[(j + 1) mod WS]
            const uint32_t lf = qf[WS - 1], lb = qb[0];
#pragma unroll
            for (int j = WS - 1; j > 0; --j) qf[j] = qf[j - 1];
#pragma unroll
            for (int j = 0; j < WS - 1; ++j) qb[j] = qb[j + 1];
            qf[0] = lf; qb[WS - 1] = lb;
        }
    }

    __device__ __forceinline__ void run(const KParams& P, double (&y)[F]) {
        const int gp_lo = P.gp_lo, gp_hi = P.gp_hi, ng = P.ngroups;
        int g = 0;
        groups_few<true>(gp_lo); g = gp_lo;
#pragma unroll 1
        for (; g + WS <= gp_hi; g += WS) groups_to_wrap<false, 0>();
        groups_few<false>(gp_hi - g); g = gp_hi;
        if (ng - 1 > g) { groups_few<true>(ng - 1 - g); g = ng - 1; }
        if (g < ng) {
            // last group: steps beyond nt + K - 1 are idle for every lane and are not issued at all
            const int ls = P.last_steps;
            step<true, 0, 0>();
            if (ls > 1) step<true, 1, 0>();
            if (ls > 2) step<true, 2, 0>();
            if (ls > 3) step<true, 3, 0>();
        }
#pragma unroll
        for (int f = 0; f < F; ++f) y[f] = acc[f];
    }
};

template <int F, int DW>
__device__ __forceinline__ void lutfir_set(const KParams& P, const LaneLut& L, const uint32_t* tw, int jr,
                                           uint32_t tp0, double (&y)[F]) {
    FrameSet<F, DW> S;
    S.init(P, L, tw, jr, tp0);
    S.run(P, y);
}

template <int WM, int DITH, int F, int DW>
__global__ void __launch_bounds__(kLutfirThreads, 1) lutfir_kernel(const __grid_constant__ KParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    // [zero region | LUT | 2 input tiles | 2 output stages]
    const unsigned char* lutb = smem + P.zpad_bytes;
    unsigned char* inbuf = smem + P.zpad_bytes + (size_t)P.lut_doubles * 8;
    unsigned char* obuf = inbuf + 2 * (size_t)P.in_buf_bytes;

    const int tid = threadIdx.x;
    const int nc = P.nc;
    const bool producer = tid >= nc;
    const int ptid = tid - nc, np = blockDim.x - nc;

    for (int i = tid; i < (P.zpad_bytes >> 3); i += blockDim.x) reinterpret_cast<double*>(smem)[i] = 0.0;
    load_lut(reinterpret_cast<double*>(smem + P.zpad_bytes), P.lut, P.lut_doubles);

    // lane geometry of the skewed look-up schedule
    const int lane16 = tid & 15;
    const int c = lane16 & (P.lutK - 1);
    const int r = lane16 / P.lutK;
    const uint32_t tp0 = (uint32_t)__cvta_generic_to_shared(lutb) + (uint32_t)((P.lut_st[r] - c) * 8);
    LaneLut LL;
    LL.rlo = (uint32_t)P.lut_row_bytes;
    LL.rhi = (uint32_t)P.lut_row_bytes << 16;
    LL.c = c;
    LL.dz = 128 - P.zpad_bytes - (P.lut_st[r] & ~15) * 8;       // same bank pair, inside the zero region

    Emitter<WM, DITH> em; em.init();
    const bool staged_out = P.out_format == D2P_DEV_OUT_LE || P.out_format == D2P_DEV_OUT_BE;
    const size_t fb = (size_t)P.C * P.bps;

    // tile headers (job look-up, 64-bit position arithmetic) are worked out by the producer warps only, one tile
    // ahead, and handed to the consumers through shared memory
    __shared__ TileInfo hdr[2];
    __shared__ __align__(8) unsigned long long full[2];            // mbarriers: TMA bytes of input buffer 0 / 1 have landed
    if (tid == 0) {
        mbar_init(&full[0], 1); mbar_init(&full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    unsigned int tile = blockIdx.x;
    TileInfo cur, prev;
    cur.J = nullptr; prev.J = nullptr;
    if (tile < P.total_tiles && producer) {
        cur = lutfir_tile(P, tile);
        if (ptid == 0) hdr[0] = cur;
        stage_input(P, *cur.J, inbuf, cur.s_lo, cur.nbytes, ptid, np, &full[0]);
        cp_async_wait_all();
        mbar_wait(&full[0], 0u);
    }
    __syncthreads();

    int j = 0;
    bool have_prev = false;
    for (; tile < P.total_tiles; tile += gridDim.x, ++j) {
        const int bsel = j & 1;
        if (producer) {
            const unsigned int nxt = tile + gridDim.x;
            TileInfo tn;
            tn.J = nullptr;
            if (nxt < P.total_tiles) {
                tn = lutfir_tile(P, nxt);
                if (ptid == 0) hdr[bsel ^ 1] = tn;
                stage_input(P, *tn.J, inbuf + (size_t)(bsel ^ 1) * P.in_buf_bytes, tn.s_lo, tn.nbytes, ptid, np, &full[bsel ^ 1]);
            }
            if (have_prev && staged_out)
                copy_out_tma(prev.J->out + (size_t)(prev.f0 - prev.J->f_begin) * fb, obuf + (size_t)(bsel ^ 1) * P.ostage_bytes,
                             (int)(prev.nf * fb), ptid, np);
            cp_async_wait_all();
            if (nxt < P.total_tiles) mbar_wait(&full[bsel ^ 1], (uint32_t)(((j + 1) >> 1) & 1));   // that buffer's ((j+1)/2)-th fill
            if (ptid == 0) tma_store_wait_read();                  // the output stage may be rewritten after the barrier
            prev = cur; cur = tn;
        } else {
            const TileInfo t = hdr[bsel];
            const unsigned char* tiles = inbuf + (size_t)bsel * P.in_buf_bytes;
            unsigned char* ostage = obuf + (size_t)bsel * P.ostage_bytes;
            const int bpo = P.bpo;
            em.set_job(P, *t.J);
            // One item = one frame set: slot s of channel ch holds the frames fl0 + f*fpw, f < F, with
            // fl0 = (s / fpw) * F * fpw + s % fpw  (fpw = frames per 32-bit word of the stream, 1 unless bpo < 4).
            const int SL = P.SL, items = P.C * SL;                // SL frame sets per channel (any multiple of 32)
            const int fs = P.fpw_shift, fpw = 1 << fs;
            for (int it = tid; it < items; it += nc) {
                const int ch = (int)__umulhi((unsigned)it, P.sl_inv), slot = it - ch * SL;   // it / SL, it % SL
                // warp-uniform: the first lane of the warp holds the warp's earliest frame
                if (((((slot & ~31) >> fs) * F) << fs) >= t.nf) continue;
                const int fl0 = (((slot >> fs) * F) << fs) + (slot & (fpw - 1));
                const uint32_t* tw = reinterpret_cast<const uint32_t*>(tiles + (size_t)ch * P.tile_stride);
                double y[F];
                lutfir_set<F, DW>(P, LL, tw, (fl0 + 1) * bpo - 1 + t.jr_base, tp0, y);
                if (WM == 1) {                                  // the dither / rounding chains of the set side by side
                    int v[F];
#pragma unroll
                    for (int f = 0; f < F; ++f) v[f] = em.quant(P, ch, t.f0 + fl0 + (f << fs), y[f], fl0 + (f << fs) < t.nf);
#pragma unroll
                    for (int f = 0; f < F; ++f) {
                        const int fl = fl0 + (f << fs);
                        if (fl < t.nf) em.put(P, *t.J, ostage, ch, t.f0 + fl, fl, v[f]);
                    }
                } else {
#pragma unroll
                    for (int f = 0; f < F; ++f) {
                        const int fl = fl0 + (f << fs);
                        if (fl < t.nf) em.emit(P, *t.J, ostage, ch, t.f0 + fl, fl, y[f]);
                    }
                }
            }
            if (staged_out) fence_proxy_async();                   // this thread's packed PCM -> visible to the bulk store
        }
        have_prev = true;
        __syncthreads();
    }
    if (producer && have_prev && staged_out) {
        copy_out_tma(prev.J->out + (size_t)(prev.f0 - prev.J->f_begin) * fb, obuf + (size_t)((j - 1) & 1) * P.ostage_bytes,
                     (int)(prev.nf * fb), ptid, np);
        if (ptid == 0) tma_store_wait_all();
    }
    if (!producer) em.flush(P);
}

// ------------------------------------------------------------------------------------------------
// Rational L/M path.
//   stage 1: output n sits at high-rate index n_hi = base_hi + m1*n; t = n_hi div L, p = n_hi mod L
//            y1[n] = fold_g LUT[p][g][ window8(t - (8g mod ring_cap)) ],  window8(t) bit j = x[t-j]
//            (pre-stream bits read 0; the modulo reproduces the reference's 32-entry byte ring)
//   stage s: y_s[k] = fold_{j<h} total + F[j]*(y[t-j] + y[t-(2h-1-j)]),  t = (h-1) + k*D, pre-stream = 0.0
//
// Both stages are bound by the shared-memory crossbar, so the layouts are chosen to make every warp-wide
// 8-byte access conflict-free (profiles/r01_lm_v1_*: 56 % of the wavefronts of the first version were
// bank-conflict replays):
//   * stage-1 LUT: one row per byte value; inside a row, lane u (of the A active lanes of a half-warp) owns the
//     block of slots [Gp*u, Gp*u + G) holding the G tables of ITS phase.  Gp is odd, so the 16 blocks start in
//     16 different bank pairs.  Half-warps take A = L*floor(16/L) consecutive outputs starting at an output of
//     phase 0, hence lane u always has phase (m1*u) mod L: no per-output division, no phase look-up;
//   * stage-1 results are stored phase-major with respect to the stage-2 decimation: sample n of the tile goes
//     to s1[(n mod D2)*pitch + n div D2].  Output j of stage 2 reads, at tap k, row (T0-k) mod D2 at column
//     (T0-k) div D2 + j: lanes = consecutive outputs read consecutive doubles.  The per-tap (row, column)
//     offsets are the same for every tile and come from the parameter block (off2f/off2b);
//   * stage 3 (when present) reads stage-2 results laid out the same way with respect to D3.
// NB channels are kept in flight together so that stages 2/3, which have only TF-ish outputs per channel,
// still occupy the whole CTA.
// ------------------------------------------------------------------------------------------------
// Direct symmetric FIR (DecimFIRSym::convolve_direct, rd/lm_resampler.rs:717-751) over a tile of the previous stage.
// Tap k reads the front sample e = T0 - k and the back sample e = k.  sigma = -h (mod D) makes both walkers leave their
// columns at the same taps, so the taps split into a few straight runs; `col` points at this output's column 0.  All
// bookkeeping is warp-uniform.  N taps of one straight run, all operand loads ahead of the adds: the loads of a run are
// independent of the accumulator chain, so the compiler may issue them back to back.
template <int N>
__device__ __forceinline__ double fir_run(const double* __restrict__ taps, const double*& pf, const double*& pb, double total) {
    double f[N], b[N];
#pragma unroll
    for (int i = 0; i < N; ++i) { f[i] = pf[-i]; b[i] = pb[i]; }   // one address register per walker, immediate offsets
    pf -= N; pb += N;
#pragma unroll
    for (int i = 0; i < N; ++i) total = __dadd_rn(total, __dmul_rn(taps[i], __dadd_rn(f[i], b[i])));
    return total;
}

// Tile layout: sample e of the tile lives in column (e + sigma) div D at row (e + sigma) mod D, a column being Dp >= D
// consecutive doubles (Dp odd).  A tap walk is therefore CONTIGUOUS in memory -- the front walker goes down one double
// per tap, the back walker up -- and only skips the Dp - D pad slots when it leaves a column; the lanes of a half-warp
// (consecutive outputs = consecutive columns) are Dp doubles apart, an odd stride, i.e. 16 different bank pairs.
__device__ __forceinline__ double fir_sym_t(const double* __restrict__ taps, const double* __restrict__ col, int h, int D,
                                            int Dp, int sigma, int rf0, int q0) {
    // front element of tap 0 (T0 = 2h-1), shifted: ef = 2h-1+sigma, rf0 = ef mod D, q0 = ef div D (host-computed)
    const int rf = rf0;
    const double* pf = col + q0 * Dp + rf;
    const double* pb = col + sigma;                                 // back element of tap 0: row sigma == D-1-rf, column 0
    const int skip = Dp - D;
    double total = 0.0;
    // The first run has (h mod D, or D) = rf0 + 1 taps; with sigma = -h (mod D) every later run is a whole column of D
    // taps: D = 7, 14 or 28 is walked in sevens, D = 3 in one piece -- no per-run trip-count arithmetic.
    {
        const int run = rf + 1;
        int i = 0;
        for (; i + 7 <= run; i += 7) total = fir_run<7>(taps + i, pf, pb, total);
        for (; i + 3 <= run; i += 3) total = fir_run<3>(taps + i, pf, pb, total);
        for (; i < run; ++i) total = fir_run<1>(taps + i, pf, pb, total);
    }
    if (D == 3) {
        for (int k = rf + 1; k < h; k += 3) { pf -= skip; pb += skip; total = fir_run<3>(taps + k, pf, pb, total); }
    } else {
        for (int k = rf + 1; k < h; k += D) {
            pf -= skip; pb += skip;                                 // front: bottom row of the previous column; back: row 0 of the next
            for (int c = 0; c < D; c += 7) total = fir_run<7>(taps + k + c, pf, pb, total);
        }
    }
    return total;
}

struct LmTile {                 // per-tile header, computed by one thread
    const DevJob* J;            // the job record: in global memory for the producers; the consumers' copy of the header
                                // points at `job` below, so the writer's per-sample reads of job fields stay on chip
    DevJob job;
    long long f0, k2_lo, s_lo;
    int nf, cnt2, j2_0, last_rel, relz, back, dA, nbytes;
    int nch;                    // = P.C; read back from the header so that the channel-group loop keeps its own counter
};

__device__ __forceinline__ void lm_tile_header(const KParams& P, unsigned int tile, LmTile& H) {
    const DevJob& J = find_job(P, tile);
    H.J = &J;
    const long long f0 = J.f_begin + (long long)(tile - J.tile_begin) * P.TF;
    const long long left = J.f_begin + J.f_count - f0;
    const int nf = left < P.TF ? (int)left : P.TF;
    H.f0 = f0; H.nf = nf; H.nch = P.C;
    // stage-2 output range feeding the tile (for 2 stages it *is* the tile)
    long long k2_lo, k2_hi;
    if (P.nstages == 3) { k2_lo = (P.h3 - 1) + f0 * P.D3 - (2 * P.h3 - 1); k2_hi = (P.h3 - 1) + (f0 + nf - 1) * P.D3; }
    else { k2_lo = f0; k2_hi = f0 + nf - 1; }
    const long long k2c = k2_lo > 0 ? k2_lo : 0;
    H.k2_lo = k2_lo; H.cnt2 = (int)(k2_hi - k2_lo + 1); H.j2_0 = (int)(k2_lo - k2c);   // <= 0: column of stage-2 output idx is idx + j2_0
    // stage-1 output range feeding those
    const long long n_lo = (P.h2 - 1) + k2c * P.D2 - (2 * P.h2 - 1);
    const long long n_hi = (P.h2 - 1) + k2_hi * P.D2;
    H.last_rel = (int)(n_hi - n_lo);
    H.relz = n_lo < 0 ? (int)(-n_lo) : 0;                          // first tile-relative output that exists (n >= 0)
    // input bits feeding those
    const long long ncl = n_lo > 0 ? n_lo : 0;
    const long long t_first = (P.lm_base_hi + (long long)P.m1 * ncl) / P.L;
    const long long t_last = (P.lm_base_hi + (long long)P.m1 * n_hi) / P.L;
    const long long s_lo = ((t_first - 31) >> 3) & ~15LL;          // arithmetic shift: floor for negatives
    H.s_lo = s_lo;
    H.nbytes = (int)((((t_last >> 3) + 1 - s_lo) + 15) & ~15LL) + 16;
    // first output of phase 0 at or before n_lo: half-warps start there, lane u then has phase (m1*u) mod L
    const long long num = P.lm_base_hi + (long long)P.m1 * n_lo;
    long long tq = num >= 0 ? num / P.L : -((-num + P.L - 1) / P.L);
    int ph = (int)(num - tq * P.L), back = 0;
    const int dm = P.m1 % P.L;
    while (ph != 0) { ph -= dm; if (ph < 0) ph += P.L; ++back; }
    H.back = back;
    const long long numA = num - (long long)P.m1 * back;           // multiple of L
    const long long tA = numA >= 0 ? numA / P.L : -((-numA) / P.L);
    H.dA = (int)(tA - 31 - 8 * s_lo);                              // tile-relative bit of output nA's oldest window bit
}

// N consecutive stage-1 outputs of one lane (input-time step dt bits, store step sq doubles): y = fold_g LUT[p][g][window
// byte g] (Stage1Poly, rd/lm_resampler.rs:603-640), the windows' bit j = x[t - j]; group 4 re-reads byte 0 (8g mod 32).
template <int GMAX, int N>
__device__ __forceinline__ void stage1_outputs(const uint32_t* __restrict__ tw, int d, int dt, uint32_t rlo, uint32_t rhi,
                                               uint32_t tpl, double* pd, int sq) {
    uint32_t z[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
        const int dk = d + k * dt, w = dk >> 5;
        z[k] = __brev(__funnelshift_r(tw[w], tw[w + 1], dk & 31));
    }
    double y[N];
#pragma unroll
    for (int k = 0; k < N; ++k) y[k] = lds_f64<0>(__dp2a_lo(rlo, z[k], tpl));
    if (GMAX > 1) {
#pragma unroll
        for (int k = 0; k < N; ++k) y[k] = __dadd_rn(y[k], lds_f64<8>(__dp2a_lo(rhi, z[k], tpl)));
    }
    if (GMAX > 2) {
#pragma unroll
        for (int k = 0; k < N; ++k) y[k] = __dadd_rn(y[k], lds_f64<16>(__dp2a_hi(rlo, z[k], tpl)));
    }
    if (GMAX > 3) {
#pragma unroll
        for (int k = 0; k < N; ++k) y[k] = __dadd_rn(y[k], lds_f64<24>(__dp2a_hi(rhi, z[k], tpl)));
    }
    if (GMAX > 4) {
#pragma unroll
        for (int k = 0; k < N; ++k) y[k] = __dadd_rn(y[k], lds_f64<32>(__dp2a_lo(rlo, z[k], tpl)));
    }
#pragma unroll
    for (int k = 0; k < N; ++k) pd[k * sq] = y[k];
}

template <int NCONS>
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" :: "n"(NCONS) : "memory"); }

// kLmConsumers consumer threads run the stages; kLmProducers extra threads stage the next tile's input from
// HBM and copy the previous tile's packed PCM out (double-buffered), so no consumer ever waits on HBM.
template <int GMAX, int NCONS>
__global__ void __launch_bounds__(NCONS + kLmProducers, 1) lm_kernel(const __grid_constant__ KParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ LmTile hdr[2];
    double* lut = reinterpret_cast<double*>(smem);
    double* s1 = lut + P.lut_doubles;
    double* s2 = s1 + (size_t)P.NB * P.s1_cap;
    unsigned char* tiles0 = reinterpret_cast<unsigned char*>(s2 + (size_t)P.NB * P.s2_cap);
    const size_t tiles_bytes = (size_t)P.C * P.tile_stride;
    unsigned char* ostage0 = tiles0 + 2 * tiles_bytes;

    const int tid = threadIdx.x;
    const bool producer = tid >= NCONS;
    const int ptid = tid - NCONS, np = blockDim.x - NCONS;
    const bool staged_out = P.out_format == D2P_DEV_OUT_LE || P.out_format == D2P_DEV_OUT_BE;
    const size_t fb = (size_t)P.C * P.bps;

    load_lut(lut, P.lut, P.lut_doubles);
    Emitter<0, 0> em; em.init();

    const int u = tid & 15, hw = tid >> 4, nhw = NCONS >> 4;
    const int A = P.lmA;
    // block of this lane: u, or u + A for odd half-warps when the phase pattern alternates (lm_cyc == 2)
    const uint32_t tpl = (uint32_t)__cvta_generic_to_shared(lut) + (uint32_t)(P.lmGp * (u + A * (hw & (P.lm_cyc - 1))) * 8);
    const uint32_t rlo = (uint32_t)P.lut_row_bytes, rhi = (uint32_t)P.lut_row_bytes << 16;
    // one half-warp iteration advances A*nhw outputs: (row, column) step of the phase-major store
    const int step_n = A * nhw;
    const int step_q = P.lm_step_q, step_ph = P.lm_step_ph;         // step_n div / mod D2 (host-computed)
    const int step_t = (int)P.lm_step_t;                            // = m1 * A * nhw / L (exact)
    const int step_el = step_ph + step_q * P.pitch2, wrap_el = P.pitch2 - P.D2;   // next column: row -= D2
    const int own = A * hw + u;                                     // this thread's output offset from nA
    const int own_t = (P.m1 * own) / P.L;                           // ... and its input-time offset (nA has phase 0)

    LmTile Hp, Hc;                                                  // producers: previous / current tile
    Hp.J = nullptr; Hc.J = nullptr;
    unsigned int tile = blockIdx.x;
    if (producer && tile < P.total_tiles) {
        lm_tile_header(P, tile, Hc);
        if (ptid == 0) { hdr[0] = Hc; hdr[0].job = *Hc.J; hdr[0].J = &hdr[0].job; }
        stage_input(P, *Hc.J, tiles0, Hc.s_lo, Hc.nbytes, ptid, np);
        cp_async_wait_all();
    }
    __syncthreads();

    for (int j = 0; tile < P.total_tiles; tile += gridDim.x, ++j) {
        const int bsel = j & 1;
        if (producer) {
            const unsigned int nxt = tile + gridDim.x;
            LmTile Hn; Hn.J = nullptr;
            if (nxt < P.total_tiles) {
                lm_tile_header(P, nxt, Hn);
                if (ptid == 0) { hdr[bsel ^ 1] = Hn; hdr[bsel ^ 1].job = *Hn.J; hdr[bsel ^ 1].J = &hdr[bsel ^ 1].job; }
                stage_input(P, *Hn.J, tiles0 + (size_t)(bsel ^ 1) * tiles_bytes, Hn.s_lo, Hn.nbytes, ptid, np);
            }
            if (Hp.J && staged_out)
                copy_out(Hp.J->out + (size_t)(Hp.f0 - Hp.J->f_begin) * fb, ostage0 + (size_t)(bsel ^ 1) * P.ostage_bytes,
                         (int)(Hp.nf * fb), ptid, np);
            Hp = Hc; Hc = Hn;
            cp_async_wait_all();
        } else {
            const LmTile& H = hdr[bsel];
            const DevJob& J = *H.J;
            const unsigned char* tiles = tiles0 + (size_t)bsel * tiles_bytes;
            unsigned char* ostage = ostage0 + (size_t)bsel * P.ostage_bytes;
            em.set_job(P, J);
            const long long f0 = H.f0;
            const int nf = H.nf, cnt2 = H.cnt2, last_rel = H.last_rel, relz = H.relz;

            int rel0 = own - H.back, d0 = H.dA + own_t;
            if (rel0 < 0) { rel0 += step_n; d0 += step_t; }        // outputs before the tile's first one are not needed
            const int e0 = rel0 + P.sigma2;                         // shifted element index (see fir_sym_t)
            const int q0 = (e0 * P.inv_d2) >> 16, p0 = e0 - q0 * P.D2;

            // the 448-consumer build reads the channel count back from the tile header: an opaque bound keeps the
            // compiler from collapsing the tile and channel-group loops into one counter with divisions per iteration
            // (the 64-register 896-consumer build is better off without the extra live value)
            const int nch = NCONS == kLmConsumersSmall ? H.nch : P.C;
            for (int cb = 0; cb < nch; cb += P.NB) {
                const int nbc = nch - cb < P.NB ? nch - cb : P.NB;
                // ---- stage 1 (A1 + A7/A8) ----
                if (u < A) {
                    for (int sl = 0; sl < nbc; ++sl) {
                        const uint32_t* tw = reinterpret_cast<const uint32_t*>(tiles + (size_t)(cb + sl) * P.tile_stride);
                        double* pd = s1 + (size_t)sl * P.s1_cap + q0 * P.pitch2 + p0;
                        const int sq = step_q * P.pitch2;
                        if (step_ph == 0 && relz <= rel0) {
                            // the usual case: a round advances a whole number of stage-2 periods (the lane's row of the
                            // phase-major tile never changes) and every output of the tile exists
                            // several outputs per round trip: all windows are fetched and all look-ups issued before the
                            // first store (stores to the stage-1 tile may alias the input tile as far as the compiler knows)
                            int rel = rel0, d = d0;
                            // (two at a time: four in flight measured the same -- stage 1 does not wait on its own loads)
                            for (; rel + step_n <= last_rel; rel += 2 * step_n, d += 2 * step_t, pd += 2 * sq)
                                stage1_outputs<GMAX, 2>(tw, d, step_t, rlo, rhi, tpl, pd, sq);
                            if (rel <= last_rel) stage1_outputs<GMAX, 1>(tw, d, step_t, rlo, rhi, tpl, pd, sq);
                            continue;
                        }
                        int ph = p0, d = d0;
                        for (int rel = rel0; rel <= last_rel; rel += step_n, d += step_t) {
                            double y = 0.0;                         // pre-stream stage-1 samples are 0.0 (ring zero-init)
                            if (rel >= relz) {
                                const int w = d >> 5;
                                const uint32_t z = __brev(__funnelshift_r(tw[w], tw[w + 1], d & 31));   // bit j = x[t-j]
                                y = lds_f64<0>(__dp2a_lo(rlo, z, tpl));
                                if (GMAX > 1) y = __dadd_rn(y, lds_f64<8>(__dp2a_lo(rhi, z, tpl)));
                                if (GMAX > 2) y = __dadd_rn(y, lds_f64<16>(__dp2a_hi(rlo, z, tpl)));
                                if (GMAX > 3) y = __dadd_rn(y, lds_f64<24>(__dp2a_hi(rhi, z, tpl)));
                                if (GMAX > 4) y = __dadd_rn(y, lds_f64<32>(__dp2a_lo(rlo, z, tpl)));   // 8g mod 32: byte 0 again
                            }
                            *pd = y;
                            pd += step_el; ph += step_ph;
                            if (ph >= P.D2) { ph -= P.D2; pd += wrap_el; }
                        }
                    }
                }
                consumer_sync<NCONS>();
                // ---- stage 2 (A9) ----
                for (int item = tid; item < nbc * cnt2; item += NCONS) {
                    int sl = 0, idx = item;
                    while (idx >= cnt2) { idx -= cnt2; ++sl; }
                    double tot = 0.0;
                    if (H.k2_lo + idx >= 0)
                        tot = fir_sym_t(P.c2, s1 + (size_t)sl * P.s1_cap + (idx + H.j2_0) * P.pitch2, P.h2, P.D2, P.pitch2, P.sigma2, P.fir2_rf, P.fir2_q);
                    if (P.nstages == 2) em.emit(P, J, ostage, cb + sl, f0 + idx, idx, tot);
                    else {
                        const int e3 = idx + P.sigma3;
                        const int qq = (e3 * P.inv_d3) >> 16;
                        s2[(size_t)sl * P.s2_cap + qq * P.pitch3 + (e3 - qq * P.D3)] = tot;
                    }
                }
                // ---- stage 3 ----
                if (P.nstages == 3) {
                    consumer_sync<NCONS>();
                    for (int item = tid; item < nbc * nf; item += NCONS) {
                        int sl = 0, idx = item;
                        while (idx >= nf) { idx -= nf; ++sl; }
                        const double y = fir_sym_t(P.c3, s2 + (size_t)sl * P.s2_cap + idx * P.pitch3, P.h3, P.D3, P.pitch3, P.sigma3, P.fir3_rf, P.fir3_q);
                        em.emit(P, J, ostage, cb + sl, f0 + idx, idx, y);
                    }
                }
                if (cb + P.NB < nch) consumer_sync<NCONS>();               // s1/s2 are reused by the next channel group
            }
        }
        __syncthreads();
    }
    if (producer && Hc.J == nullptr && Hp.J && staged_out) {
        // after the loop Hp is the last tile processed (Hc, the one after it, does not exist)
        const int jl = (int)((P.total_tiles - 1 - blockIdx.x) / gridDim.x);   // index of this CTA's last tile
        copy_out(Hp.J->out + (size_t)(Hp.f0 - Hp.J->f_begin) * fb, ostage0 + (size_t)(jl & 1) * P.ostage_bytes,
                 (int)(Hp.nf * fb), ptid, np);
    }
    if (!producer) em.flush(P);
}

// ------------------------------------------------------------------------------------------------
// Airwindows float dither (-d F, the CLI default for 32-bit float output): Dither::fpd is ONE xorshift32
// stream per file, advanced once per sample in reader order -- block, then channel, then frame
// (rd/conversion_context.rs:303-315, rd/pcm_writer.rs:608-648).  Sample number n of the file (0-based) sees the
// state after n+1 steps.  xorshift32 is linear over GF(2), so a work item jumps straight to the state in
// front of its run of samples with four byte-sliced matrix applications (P.fpd_jump), then walks the run.
// One work item = the frames one channel gets out of one reader block.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ long long dev_frames_for_bytes(const KParams& P, long long bytes) {   // Plan::frames_for_bytes
    if (bytes <= 0) return 0;
    if (P.nstages == 1) { const long long f = bytes / P.bpo - (P.h - 1); return f > 0 ? f : 0; }
    const long long hi = bytes * 8 * P.L;
    long long n1 = P.accum ? hi / 21 - P.e0 : (hi >= P.h1 ? (hi - P.h1) / P.m1 : 0);
    if (n1 < 0) n1 = 0;
    const long long n2 = n1 >= P.h2 ? (n1 - P.h2) / P.D2 + 1 : 0;
    if (P.nstages == 2) return n2;
    return n2 >= P.h3 ? (n2 - P.h3) / P.D3 + 1 : 0;
}

__device__ __forceinline__ uint32_t fpd_jump(const uint32_t* __restrict__ T, uint32_t s, unsigned long long steps) {
    const uint32_t n = (uint32_t)(steps % 4294967295ull);           // the period of xorshift32 is 2^32 - 1
#pragma unroll
    for (int lv = 0; lv < 4; ++lv) {
        const uint32_t dg = (n >> (8 * lv)) & 255u;
        if (dg) {
            const uint32_t* M = T + ((size_t)(lv * 256 + dg) << 10);
            s = M[s & 255u] ^ M[256 + ((s >> 8) & 255u)] ^ M[512 + ((s >> 16) & 255u)] ^ M[768 + (s >> 24)];
        }
    }
    return s;
}

__global__ void __launch_bounds__(256) fpd_kernel(const __grid_constant__ KParams P) {
    const unsigned int item = blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= P.fpd_items) return;
    int lo = 0, hi = P.njobs - 1;
    while (lo < hi) {                      // last job with fpd_item_begin <= item
        const int mid = (lo + hi + 1) >> 1;
        if (P.jobs[mid].fpd_item_begin <= item) lo = mid; else hi = mid - 1;
    }
    const DevJob& J = P.jobs[lo];
    if (!J.fpd) return;
    const unsigned int li = item - J.fpd_item_begin;
    const int c = (int)(li % (unsigned int)P.C);
    const long long blk = J.blk_lo + (long long)(li / (unsigned int)P.C);
    const long long B = P.block_size;
    const long long f_before = dev_frames_for_bytes(P, blk * B);
    long long q_end = (blk + 1) * B;
    if (q_end > J.total_q) q_end = J.total_q;
    const long long f_after = dev_frames_for_bytes(P, q_end);
    const long long run = f_after - f_before;
    long long fs = f_before > J.f_begin ? f_before : J.f_begin;
    long long fe = f_after < J.f_begin + J.f_count ? f_after : J.f_begin + J.f_count;
    if (fs >= fe) return;
    const unsigned long long n0 = (unsigned long long)P.C * f_before + (unsigned long long)c * run + (fs - f_before);
    uint32_t s = fpd_jump(P.fpd_jump, J.fpd0, n0);
    uint32_t* dst = J.fpd + (long long)c * J.fpd_stride + (fs - J.f_begin);
    for (long long f = fs; f < fe; ++f) {
        s ^= s << 13; s ^= s >> 17; s ^= s << 5;
        *dst++ = s;
    }
}

// diagnostic: the octave step of the float dither on a vector of samples (tests/test_fpd_octave.py)
__global__ void fpd_exponent_kernel(const double* __restrict__ s, int n, int* __restrict__ e) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) e[i] = fpd_exponent(s[i]);
}
cudaError_t launch_fpd_exponent(const double* d_s, int n, int* d_e, cudaStream_t stream) {
    if (n > 0) fpd_exponent_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_s, n, d_e);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
const char* kernel_name(bool lm) { return lm ? "d2p::lm_kernel" : "d2p::lutfir_kernel"; }

// variant: 10..25 = lm_kernel instantiations; 100 + 10*shape + writer = lutfir_kernel, with
//   shape  0: sets of 3 frames one word apart (bpo <= 4), 1: 3 frames two words apart (bpo 8), 2: 2 frames four words apart (bpo 16)
//   writer 1: generic, 2/3/4: fast 3-byte-LE writer with dither X/T/R
template <int F, int DW>
static const void* lutfir_fn(int writer) {
    switch (writer) {
    case 2: return (const void*)lutfir_kernel<1, 0, F, DW>;
    case 3: return (const void*)lutfir_kernel<1, 'T', F, DW>;
    case 4: return (const void*)lutfir_kernel<1, 'R', F, DW>;
    default: return (const void*)lutfir_kernel<0, 0, F, DW>;
    }
}
static const void* kernel_fn(int variant) {
    if (variant >= 100) {
        const int shape = (variant - 100) / 10, writer = (variant - 100) % 10;
        if (shape == 1) return lutfir_fn<3, 2>(writer);
        if (shape == 2) return lutfir_fn<2, 4>(writer);
        return lutfir_fn<3, 1>(writer);
    }
    switch (variant) {
    case 11: return (const void*)lm_kernel<1, kLmConsumersSmall>;
    case 12: return (const void*)lm_kernel<2, kLmConsumersSmall>;
    case 13: return (const void*)lm_kernel<3, kLmConsumersSmall>;
    case 14: return (const void*)lm_kernel<4, kLmConsumersSmall>;
    case 21: return (const void*)lm_kernel<1, kLmConsumers>;
    case 22: return (const void*)lm_kernel<2, kLmConsumers>;
    case 23: return (const void*)lm_kernel<3, kLmConsumers>;
    case 24: return (const void*)lm_kernel<4, kLmConsumers>;
    case 25: return (const void*)lm_kernel<5, kLmConsumers>;
    default: return (const void*)lm_kernel<5, kLmConsumersSmall>;
    }
}

void lutfir_set_shape(int bpo, int* F, int* DW) {
    *DW = bpo <= 4 ? 1 : bpo / 4;
    *F = *DW >= 4 ? 2 : 3;
}

int kernel_variant(const KParams& P, bool lm, bool fast_writer_ok) {
    if (lm) return (P.lm_cons == kLmConsumers ? 20 : 10) + (P.gmax < 1 ? 1 : (P.gmax > 5 ? 5 : P.gmax));
    const int shape = P.setDW == 1 ? 0 : (P.setDW == 2 ? 1 : 2);
    const int writer = !fast_writer_ok ? 1 : (P.dither == 'T' ? 3 : (P.dither == 'R' ? 4 : 2));
    return 100 + 10 * shape + writer;
}

static cudaError_t ensure_smem_attr(int variant) {
    static bool attr_done[160][64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 64 && !attr_done[variant][dev]) {
        // the opt-in limit (227 KiB) covers static + dynamic shared memory: leave room for the kernels' statics
        e = cudaFuncSetAttribute(kernel_fn(variant), cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem - 1024);
        if (e != cudaSuccess) { cudaGetLastError(); return e; }
        attr_done[variant][dev] = true;
    }
    return cudaSuccess;
}

int max_ctas_per_sm(bool lm, int threads, size_t smem) {
    const int variant = lm ? (threads > kLmConsumersSmall + kLmProducers ? 25 : 15) : 101;
    if (ensure_smem_attr(variant) != cudaSuccess) return 0;
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel_fn(variant), threads, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

cudaError_t launch_fpd(const KParams& P, cudaStream_t stream) {
    if (!P.fpd_items) return cudaSuccess;
    fpd_kernel<<<(P.fpd_items + 255) / 256, 256, 0, stream>>>(P);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// built-in scratch of d2p_run_noalloc (include/dsd2pcm_b200.h)
// ------------------------------------------------------------------------------------------------
// The kernels read three things from global memory that are neither the caller's input nor its output: the job table,
// the filter LUT (copied into shared memory by every CTA) and the statistics record.  Everywhere else the host side
// cudaMallocs them.  This array is the same storage reserved inside the module image instead, for callers that must
// not allocate: it exists on every device that loads this library, from module load until the process ends, and all
// calls of d2p_run_noalloc on a device share it.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocScratchBytes];

// Device address of g_noalloc_scratch on the CURRENT device.  In host code the name g_noalloc_scratch is only a handle
// for the runtime's symbol table -- a kernel handed `g_noalloc_scratch` itself would not be looking at this array -- so
// the address is asked for once per device (cudaGetSymbolAddress also makes sure the module is loaded there) and kept.
cudaError_t noalloc_scratch_address(void** ptr) {
    static std::mutex mu;
    static void* per_device[64] = {};
    *ptr = nullptr;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(mu);
    if (!per_device[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        per_device[dev] = p;
    }
    *ptr = per_device[dev];
    return cudaSuccess;
}

cudaError_t launch_kernels(const KParams& P, int variant, int grid, int threads, size_t smem, cudaStream_t stream) {
    cudaError_t e = ensure_smem_attr(variant);
    if (e != cudaSuccess) return e;
    void* args[] = {(void*)&P};
    return cudaLaunchKernel(kernel_fn(variant), dim3(grid), dim3(threads), args, smem, stream);
}

}  // namespace d2p
