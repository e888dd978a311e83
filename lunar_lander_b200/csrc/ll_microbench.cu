// ll_microbench.cu — roofline denominators measured on the device the benchmark runs on.
// MEASURED_PEAKS.json only carries the HBM copy bandwidth and the bf16 GEMM rate; the episode kernel is bound by the
// FP32 / FP64 FMA pipes, so the FMA peak is measured here with dependency-free register-only FMA chains at the clock
// the GPU actually sustains.  Codes: 32 FFMA, 322 packed FFMA2, 64 DFMA, 197 mma.sync.m16n8k16.f16 at one warp
// per SM sub-partition with rotating operands (the legacy tensor pipe the mma / mmat episode kernels run on).
// The result is the MEAN over the timed loop; the SM clock is cycles / nanoseconds of ONE thread (clock64 and %globaltimer
// read by the same thread around the same loop), so it does not depend on how blocks share an SM.
#include <algorithm>
#include <vector>

#include "ll_internal.h"

namespace
{
    constexpr int CHAINS = 16;
    constexpr int INNER  = 256;

    __device__ __forceinline__ unsigned long long gtimer()
    {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        return t;
    }
    struct span
    {
        long long c0;
        unsigned long long n0;
        __device__ __forceinline__ void begin() { n0 = gtimer(), c0 = clock64(); }
        __device__ __forceinline__ void end(long long* out) const
        {
            const long long c1          = clock64();
            const unsigned long long n1 = gtimer();
            if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = c1 - c0, out[1] = (long long)(n1 - n0);
        }
    };

    __global__ void __launch_bounds__(256) ffma_kernel(float* out, int outer, long long* cycles)
    {
        float acc[CHAINS];
        const float a = 1.0f + 1e-7f * threadIdx.x, b = 1e-9f * (blockIdx.x + 1);
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) acc[i] = (float)i;
        span sp;
        sp.begin();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < INNER / CHAINS; ++r)
#pragma unroll
                for (int i = 0; i < CHAINS; ++i) acc[i] = fmaf(acc[i], a, b);
        }
        sp.end(cycles);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) s += acc[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }

    // packed two-wide FP32 FMA (sm_100: fma.rn.f32x2 on 64-bit register pairs)
    __global__ void __launch_bounds__(256) ffma2_kernel(float* out, int outer, long long* cycles)
    {
        unsigned long long acc[CHAINS];
        const float af = 1.0f + 1e-7f * threadIdx.x, bf = 1e-9f * (blockIdx.x + 1);
        unsigned long long a, b;
        asm("mov.b64 %0, {%1, %1};" : "=l"(a) : "f"(af));
        asm("mov.b64 %0, {%1, %1};" : "=l"(b) : "f"(bf));
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
        {
            const float v = (float)i;
            asm("mov.b64 %0, {%1, %1};" : "=l"(acc[i]) : "f"(v));
        }
        span sp;
        sp.begin();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < INNER / CHAINS; ++r)
#pragma unroll
                for (int i = 0; i < CHAINS; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(a), "l"(b));
        }
        sp.end(cycles);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
        {
            float lo, hi;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
            s += lo + hi;
        }
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }

    __global__ void __launch_bounds__(256) dfma_kernel(double* out, int outer, long long* cycles)
    {
        double acc[CHAINS];
        const double a = 1.0 + 1e-12 * threadIdx.x, b = 1e-15 * (blockIdx.x + 1);
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) acc[i] = (double)i;
        span sp;
        sp.begin();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < INNER / CHAINS; ++r)
#pragma unroll
                for (int i = 0; i < CHAINS; ++i) acc[i] = fma(acc[i], a, b);
        }
        sp.end(cycles);
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) s += acc[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }

    // mma.sync.m16n8k16 f16 with realistic operand traffic: 24 accumulator tiles, 8 B fragments and 2 A fragments in rotation
    // (no register reuse between consecutive instructions), one warp per sub-partition
    __global__ void __launch_bounds__(128) hmma_f16_kernel(float* out, int outer, long long* cycles)
    {
        float acc[24][4];
        uint32_t a[2][4], b[8][2];
#pragma unroll
        for (int i = 0; i < 24; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = (float)(i + j);
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) a[i][j] = 0x3c003c00u + threadIdx.x + 8 * i + j;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) b[i][j] = 0x14001400u + blockIdx.x + 4 * i + j;
        span sp;
        sp.begin();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < 48; ++r)
            {
                float(&c)[4]          = acc[r % 24];
                const uint32_t(&x)[4] = a[r & 1];
                const uint32_t(&y)[2] = b[r & 7];
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                             : "r"(x[0]), "r"(x[1]), "r"(x[2]), "r"(x[3]), "r"(y[0]), "r"(y[1]));
            }
        }
        sp.end(cycles);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 24; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) s += acc[i][j];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }

    __global__ void copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n)
    {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
    }
} // namespace

#define MB_CUDA(call)                                                                                                \
    do {                                                                                                             \
        cudaError_t e__ = (call);                                                                                    \
        if (e__ != cudaSuccess)                                                                                      \
        {                                                                                                            \
            rc = ll_fail(nullptr, LL_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__));                    \
            goto done;                                                                                               \
        }                                                                                                            \
    } while (0)

extern "C" int ll_microbench_fma(int32_t device, int32_t precision, double seconds, double* tflops, double* sm_mhz_est)
{
    if (!tflops || (precision != 32 && precision != 322 && precision != 64 && precision != 197)) return ll_fail(nullptr, LL_EINVAL, "ll_microbench_fma: bad argument");
    int rc = LL_OK, n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    {
        cudaGetLastError();
        return ll_fail(nullptr, LL_ENODEVICE, "ll_microbench_fma: no CUDA device");
    }
    void* out        = nullptr;
    long long* d_cyc = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaDeviceProp prop;
    double sum_fmas = 0.0, sum_ms = 0.0, sum_cyc = 0.0, sum_ns = 0.0, elapsed = 0.0;
    int blocks = 0, outer = 2000, threads = 256, iter = 0;
    if (device >= 0) MB_CUDA(cudaSetDevice(device));
    MB_CUDA(cudaGetDevice(&device));
    MB_CUDA(cudaGetDeviceProperties(&prop, device));
    blocks = prop.multiProcessorCount * 8; // 8 x 256 threads = 64 warps per SM
    if (precision == 197) blocks = prop.multiProcessorCount, threads = 128; // one warp per SM sub-partition
    MB_CUDA(cudaMalloc(&out, (size_t)blocks * threads * 8));
    MB_CUDA(cudaMalloc(&d_cyc, 16));
    MB_CUDA(cudaEventCreate(&e0));
    MB_CUDA(cudaEventCreate(&e1));
    while (elapsed < seconds * 1e3 || iter < 3)
    {
        MB_CUDA(cudaEventRecord(e0));
        if (precision == 32) ffma_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
        else if (precision == 322) ffma2_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
        else if (precision == 197) hmma_f16_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
        else dfma_kernel<<<blocks, threads>>>(static_cast<double*>(out), outer, d_cyc);
        MB_CUDA(cudaEventRecord(e1));
        MB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        MB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        elapsed += ms;
        double fmas = (double)blocks * threads * (double)outer * INNER * (precision == 322 ? 2.0 : 1.0);
        if (precision == 197) fmas = (double)blocks * (threads / 32) * (double)outer * 48.0 * 2048.0;
        if (iter > 0)
        { // the first launch is warm-up; everything after it counts (mean over the loop, not the best iteration)
            long long cyc[2] = {0, 0};
            MB_CUDA(cudaMemcpy(cyc, d_cyc, 16, cudaMemcpyDeviceToHost));
            sum_fmas += fmas, sum_ms += ms, sum_cyc += (double)cyc[0], sum_ns += (double)cyc[1];
        }
        if (ms < 20.f && outer < (1 << 22)) outer *= 2;
        ++iter;
        if (iter > 10000) break;
    }
    *tflops = sum_ms > 0.0 ? 2.0 * sum_fmas / (sum_ms * 1e-3) / 1e12 : 0.0;
    if (sm_mhz_est) *sm_mhz_est = sum_ns > 0.0 ? sum_cyc / sum_ns * 1e3 : 0.0;
done:
    if (out) cudaFree(out);
    if (d_cyc) cudaFree(d_cyc);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return rc;
}

extern "C" int ll_microbench_copy(int32_t device, uint64_t bytes, int32_t iters, double* gbps)
{
    if (!gbps || bytes < (1u << 20) || iters < 1) return ll_fail(nullptr, LL_EINVAL, "ll_microbench_copy: bad argument");
    int rc = LL_OK, n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    {
        cudaGetLastError();
        return ll_fail(nullptr, LL_ENODEVICE, "ll_microbench_copy: no CUDA device");
    }
    void *src = nullptr, *dst = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaDeviceProp prop;
    double best    = 0.0;
    const size_t n = bytes / 16;
    if (device >= 0) MB_CUDA(cudaSetDevice(device));
    MB_CUDA(cudaGetDevice(&device));
    MB_CUDA(cudaGetDeviceProperties(&prop, device));
    MB_CUDA(cudaMalloc(&src, n * 16));
    MB_CUDA(cudaMalloc(&dst, n * 16));
    MB_CUDA(cudaMemset(src, 1, n * 16));
    MB_CUDA(cudaEventCreate(&e0));
    MB_CUDA(cudaEventCreate(&e1));
    for (int i = 0; i < iters + 1; ++i)
    {
        MB_CUDA(cudaEventRecord(e0));
        copy_kernel<<<prop.multiProcessorCount * 16, 512>>>(static_cast<const uint4*>(src), static_cast<uint4*>(dst), n);
        MB_CUDA(cudaEventRecord(e1));
        MB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        MB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (i > 0) best = std::max(best, 2.0 * (double)n * 16 / (ms * 1e-3) / 1e9);
    }
    *gbps = best;
done:
    if (src) cudaFree(src);
    if (dst) cudaFree(dst);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return rc;
}
