// microbench.cu -- measures the two per-SM rates that bound the LUT-FIR path on this GPU
// (SURVEY.md 8d "to be measured"): conflict-free 8-byte shared-memory look-ups and dependent-free DADDs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
// Prints one JSON line.  Not part of the product library.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

static const int kThreads = 1024, kIters = 4096;

// every lane of a half-warp reads its own 8-byte bank pair (slot = lane & 15) of a data-dependent row:
// exactly the access pattern of lutfir_kernel's skewed LUT.  MODE 0: look-ups only (xor-reduced),
// MODE 1: look-up + dependent DADD chain (4 independent chains per thread), as the kernel does.
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1) lds_kernel(double* out, int rows, uint32_t seed) {
    extern __shared__ __align__(16) unsigned char smem[];
    double* lut = reinterpret_cast<double*>(smem);
    for (int i = threadIdx.x; i < rows * 16; i += blockDim.x) lut[i] = (double)(i % 97) * 1e-3;
    __syncthreads();
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(lut) + (threadIdx.x & 15) * 8;
    uint32_t x = seed ^ (threadIdx.x * 2654435761u) ^ (blockIdx.x * 40503u);
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    unsigned long long xr = 0;
    const uint32_t rmask = rows - 1;
    for (int it = 0; it < kIters; ++it) {
        x = x * 1664525u + 1013904223u;
        double v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t row = (x >> (4 * j)) & rmask;
            asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v[j]) : "r"(base + row * 128));
        }
        if (MODE == 1) {
            a0 = __dadd_rn(__dadd_rn(a0, v[0]), v[4]); a1 = __dadd_rn(__dadd_rn(a1, v[1]), v[5]);
            a2 = __dadd_rn(__dadd_rn(a2, v[2]), v[6]); a3 = __dadd_rn(__dadd_rn(a3, v[3]), v[7]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) xr ^= (unsigned long long)__double_as_longlong(v[j]);
        }
    }
    if (MODE == 1) out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
    else out[blockIdx.x * blockDim.x + threadIdx.x] = __longlong_as_double((long long)xr);
}

__global__ void __launch_bounds__(kThreads, 1) dadd_kernel(double* out, double s) {
    double a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = threadIdx.x * 1e-9 + j;
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = __dadd_rn(a[j], s);
    }
    double t = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

__global__ void __launch_bounds__(kThreads, 1) idp_kernel(uint32_t* out, uint32_t r, uint32_t q) {
    uint32_t a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = threadIdx.x + j;
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = __dp2a_lo(r, q + j, a[j]);
    }
    uint32_t t = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) t ^= a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

template <typename F>
static float time_ms(F launch, int reps) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    launch(); launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { std::printf("{\"error\": \"no CUDA device\"}\n"); return 1; }
    const int sms = prop.multiProcessorCount;
    double* d_out; cudaMalloc(&d_out, sizeof(double) * sms * kThreads);
    const int rows = 256;
    const size_t smem = (size_t)rows * 128;
    cudaFuncSetAttribute(lds_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(lds_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const double n = (double)sms * kThreads * kIters * 8;
    const float t0 = time_ms([&] { lds_kernel<0><<<sms, kThreads, smem>>>(d_out, rows, 1u); }, 5);
    const float t1 = time_ms([&] { lds_kernel<1><<<sms, kThreads, smem>>>(d_out, rows, 1u); }, 5);
    const float t2 = time_ms([&] { dadd_kernel<<<sms, kThreads>>>(d_out, 1e-7); }, 5);
    const float t3 = time_ms([&] { idp_kernel<<<sms, kThreads>>>((uint32_t*)d_out, 0x00000080u, 0x01020304u); }, 5);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    std::printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d, \"lds64_lookups_per_s\": %.4e, "
                "\"lds64_plus_dadd_lookups_per_s\": %.4e, \"dadd_per_s\": %.4e, \"idp2a_per_s\": %.4e, "
                "\"lds64_lookups_per_clk_sm_at_max_clock\": %.3f, \"dadd_per_clk_sm_at_max_clock\": %.3f}\n",
                prop.name, sms, clk, n / (t0 * 1e-3), n / (t1 * 1e-3), n / (t2 * 1e-3), n / (t3 * 1e-3),
                n / (t0 * 1e-3) / sms / (clk * 1e3), n / (t2 * 1e-3) / sms / (clk * 1e3));
    cudaFree(d_out);
    return 0;
}
