// d2p_launch.h -- host-visible launch interface of d2p_kernels.cu.
#pragma once
#include <cuda_runtime.h>

#include "d2p_device.h"

#define D2P_DEV_OUT_LE 0
#define D2P_DEV_OUT_BE 1
#define D2P_DEV_OUT_PLANAR 2
#define D2P_DEV_OUT_NONE 3

namespace d2p {

static const int kMaxThreads = 512;          // CTA size of lm_kernel; consumer threads of lutfir_kernel
static const int kLutfirProducers = 128;     // extra threads of lutfir_kernel that stage tiles and copy results out
static const int kLutfirThreads = kMaxThreads + kLutfirProducers;
static const int kLmProducers = 64;          // extra threads of lm_kernel (same role)
// Consumer threads of lm_kernel: two instantiations.  448 = 28 half-warps: with 512 threads per CTA each thread may keep
// 128 registers, which the unrolled FIR runs of the long two-stage chains use to put all loads of a run ahead of the
// adds.  896 = 56 half-warps (64 registers): the three-stage chains, whose short runs (D3 = 3) gain nothing from
// registers, prefer the extra warps.  Both make one stage-1 round (A * half-warps outputs) a multiple of every stage-2
// decimation in use (7, 14, 28), so a lane's row in the phase-major tile never changes.
static const int kLmConsumersSmall = 448;
static const int kLmConsumers = 896;
static const int kLmThreads = kLmConsumers + kLmProducers;
static const int kMaxSmem = 227 * 1024;      // opt-in dynamic shared memory per CTA on sm_100a

// Scratch built into the module for d2p_run_noalloc (d2p_kernels.cu: g_noalloc_scratch), cut up as
//   [0, 256)                      DevStats
//   [256, kNoallocLutOffset)      job table: kNoallocMaxJobs DevJob records
//   [kNoallocLutOffset, end)      stage-1 LUT in its shared-memory layout (never more than kMaxSmem)
static const size_t kNoallocScratchBytes = 256 * 1024;
static const int    kNoallocMaxJobs = 64;
static const size_t kNoallocJobsOffset = 256;
static const size_t kNoallocLutOffset = (kNoallocJobsOffset + kNoallocMaxJobs * sizeof(DevJob) + 255) / 256 * 256;
static_assert(kNoallocLutOffset + kMaxSmem <= kNoallocScratchBytes, "every LUT that fits shared memory fits the scratch");
cudaError_t noalloc_scratch_address(void** ptr);         // device address of the scratch on the current device

const char* kernel_name(bool lm);
void lutfir_set_shape(int bpo, int* F, int* DW);          // frames per lane set / word distance the kernels are built for
int max_ctas_per_sm(bool lm, int threads, size_t smem);   // occupancy of the kernel at this geometry (0 on error)
int kernel_variant(const KParams& P, bool lm, bool fast_writer_ok);   // which instantiation a launch uses
cudaError_t launch_fpd_exponent(const double* d_s, int n, int* d_e, cudaStream_t stream);   // diagnostic
cudaError_t launch_fpd(const KParams& P, cudaStream_t stream);     // per-sample xorshift32 states for dither F
cudaError_t launch_kernels(const KParams& P, int variant, int grid, int threads, size_t smem, cudaStream_t stream);

}  // namespace d2p
