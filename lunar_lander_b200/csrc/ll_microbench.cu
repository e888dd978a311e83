// ll_microbench.cu — roofline denominators measured on the device the benchmark runs on.
// MEASURED_PEAKS.json only carries the HBM copy bandwidth and the bf16 GEMM rate; the episode kernel is bound by the
// FP32 / FP64 FMA pipes, so the FMA peak is measured here with dependency-free register-only FMA chains at the clock
// the GPU actually sustains.
#include <algorithm>
#include <vector>

#include "ll_internal.h"

namespace
{
    constexpr int CHAINS = 16;
    constexpr int INNER  = 256;

    __global__ void __launch_bounds__(256) ffma_kernel(float* out, int outer, long long* cycles)
    {
        float acc[CHAINS];
        const float a = 1.0f + 1e-7f * threadIdx.x, b = 1e-9f * (blockIdx.x + 1);
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) acc[i] = (float)i;
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < INNER / CHAINS; ++r)
#pragma unroll
                for (int i = 0; i < CHAINS; ++i) acc[i] = fmaf(acc[i], a, b);
        }
        const long long t1 = clock64();
        float s            = 0.f;
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) s += acc[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
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
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < INNER / CHAINS; ++r)
#pragma unroll
                for (int i = 0; i < CHAINS; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(a), "l"(b));
        }
        const long long t1 = clock64();
        float s            = 0.f;
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
        {
            float lo, hi;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
            s += lo + hi;
        }
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
    }

    __global__ void __launch_bounds__(256) dfma_kernel(double* out, int outer, long long* cycles)
    {
        double acc[CHAINS];
        const double a = 1.0 + 1e-12 * threadIdx.x, b = 1e-15 * (blockIdx.x + 1);
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) acc[i] = (double)i;
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < INNER / CHAINS; ++r)
#pragma unroll
                for (int i = 0; i < CHAINS; ++i) acc[i] = fma(acc[i], a, b);
        }
        const long long t1 = clock64();
        double s           = 0.0;
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) s += acc[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
    }


    // legacy warp-level tensor-core MMA (mma.sync.m16n8k8 tf32, fp32 accumulate): 8 independent accumulator tiles per warp.
    // One instruction = 16*8*8 = 1024 MAC.  Used to decide whether the multi-layout forward pass belongs on the tensor
    // cores (north_star: "only if ncu shows a win") and as the denominator of that kernel's tensor-pipe utilisation.
    __global__ void __launch_bounds__(256) mma_tf32_kernel(float* out, int outer, long long* cycles)
    {
        float acc[8][4];
        const uint32_t a0 = __float_as_uint(1.0f + 1e-3f * threadIdx.x), a1 = a0 ^ 0x2000u, a2 = a0 ^ 0x4000u, a3 = a0 ^ 0x6000u;
        const uint32_t b0 = __float_as_uint(1e-3f * (blockIdx.x + 1)), b1 = b0 ^ 0x2000u;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = (float)(i + j);
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                 : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
        const long long t1 = clock64();
        float s            = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) s += acc[i][j];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
    }


    // dependent chain: every mma accumulates into the same tile (issue-to-issue latency of back-to-back dependent HMMAs)
    __global__ void __launch_bounds__(256) mma_tf32_chain_kernel(float* out, int outer, long long* cycles)
    {
        float acc[4] = {0.f, 1.f, 2.f, 3.f};
        const uint32_t a0 = __float_as_uint(1.0f + 1e-3f * threadIdx.x), a1 = a0 ^ 0x2000u, a2 = a0 ^ 0x4000u, a3 = a0 ^ 0x6000u;
        const uint32_t b0 = __float_as_uint(1e-3f * (blockIdx.x + 1)), b1 = b0 ^ 0x2000u;
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < 32; ++r)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(acc[0]), "+f"(acc[1]), "+f"(acc[2]), "+f"(acc[3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
        const long long t1 = clock64();
        out[blockIdx.x * blockDim.x + threadIdx.x] = acc[0] + acc[1] + acc[2] + acc[3];
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
    }


    // realistic operand traffic: 24 accumulator tiles, 8 B fragments and 2 A fragments in rotation (no register reuse
    // between consecutive instructions), one warp per sub-partition.  F16 = false: m16n8k8 tf32; true: m16n8k16 f16.
    template <bool F16> __global__ void __launch_bounds__(128) mma_rotate_kernel(float* out, int outer, long long* cycles)
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
            for (int j = 0; j < 4; ++j) a[i][j] = F16 ? (0x3c003c00u + threadIdx.x + 8 * i + j) : __float_as_uint(1.0f + 1e-3f * (threadIdx.x + 8 * i + j));
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) b[i][j] = F16 ? (0x14001400u + blockIdx.x + 4 * i + j) : __float_as_uint(1e-3f * (blockIdx.x + 4 * i + j + 1));
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < 48; ++r)
            {
                float(&c)[4] = acc[r % 24];
                const uint32_t(&x)[4] = a[r & 1];
                const uint32_t(&y)[2] = b[r & 7];
                if (F16)
                    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                                 : "r"(x[0]), "r"(x[1]), "r"(x[2]), "r"(x[3]), "r"(y[0]), "r"(y[1]));
                else
                    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                                 : "r"(x[0]), "r"(x[1]), "r"(x[2]), "r"(x[3]), "r"(y[0]), "r"(y[1]));
            }
        }
        const long long t1 = clock64();
        float s            = 0.f;
#pragma unroll
        for (int i = 0; i < 24; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) s += acc[i][j];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
    }


    // does other work issue in the shadow of an HMMA?  K independent FFMAs (K = 0, 2, 4, 8) after every tf32 HMMA,
    // rotating operands.  Reported as HMMA TFLOP/s: unchanged rate = the FFMAs were free.
    template <int K> __global__ void __launch_bounds__(128) mma_mix_kernel(float* out, int outer, long long* cycles)
    {
        float acc[24][4], f[8];
        uint32_t a[2][4], b[8][2];
#pragma unroll
        for (int i = 0; i < 24; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = (float)(i + j);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = (float)i;
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) a[i][j] = __float_as_uint(1.0f + 1e-3f * (threadIdx.x + 8 * i + j));
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) b[i][j] = __float_as_uint(1e-3f * (blockIdx.x + 4 * i + j + 1));
        const float fa = 1.0f + 1e-7f * threadIdx.x, fb = 1e-9f * (blockIdx.x + 1);
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < 48; ++r)
            {
                float(&c)[4] = acc[r % 24];
                const uint32_t(&x)[4] = a[r & 1];
                const uint32_t(&y)[2] = b[r & 7];
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                             : "r"(x[0]), "r"(x[1]), "r"(x[2]), "r"(x[3]), "r"(y[0]), "r"(y[1]));
#pragma unroll
                for (int k = 0; k < K; ++k) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[k]) : "f"(fa), "f"(fb));
            }
        }
        const long long t1 = clock64();
        float s            = 0.f;
#pragma unroll
        for (int i = 0; i < 24; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) s += acc[i][j];
#pragma unroll
        for (int i = 0; i < 8; ++i) s += f[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
    }

    // same with bf16 operands (m16n8k16, 2048 MAC per instruction)
    __global__ void __launch_bounds__(256) mma_bf16_kernel(float* out, int outer, long long* cycles)
    {
        float acc[8][4];
        const uint32_t a0 = 0x3f803f80u + threadIdx.x, a1 = a0 ^ 0x20002u, a2 = a0 ^ 0x40004u, a3 = a0 ^ 0x60006u;
        const uint32_t b0 = 0x3a003a00u + blockIdx.x, b1 = b0 ^ 0x20002u;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = (float)(i + j);
        const long long t0 = clock64();
        for (int o = 0; o < outer; ++o)
        {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                 : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
        const long long t1 = clock64();
        float s            = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) s += acc[i][j];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
        if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
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
    if (!tflops || (precision != 32 && precision != 322 && precision != 64 && precision != 19 && precision != 191 && precision != 194 && precision != 195 && precision != 196 && precision != 197 && precision != 198 && precision != 16 && !(precision >= 2000 && precision < 2100))) return ll_fail(nullptr, LL_EINVAL, "ll_microbench_fma: bad argument");
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
    double best = 0.0, mhz = 0.0;
    int blocks = 0, outer = 2000;
    int threads = 256;
    if (device >= 0) MB_CUDA(cudaSetDevice(device));
    MB_CUDA(cudaGetDevice(&device));
    MB_CUDA(cudaGetDeviceProperties(&prop, device));
    blocks = prop.multiProcessorCount * (precision == 191 ? 2 : 8); // 8 x 256 threads = 64 warps per SM (191: 16 warps per SM)
    if (precision >= 194 && precision <= 197) blocks = prop.multiProcessorCount, threads = 128; // one warp per SM sub-partition
    if (precision == 198) blocks = 2 * prop.multiProcessorCount, threads = 128; // two warps per SM sub-partition
    if (precision >= 2000) blocks = (precision % 10) * prop.multiProcessorCount, threads = 128; // 20KW: K FFMAs per HMMA, W warps per sub-partition
    MB_CUDA(cudaMalloc(&out, (size_t)blocks * threads * 8));
    MB_CUDA(cudaMalloc(&d_cyc, 8));
    MB_CUDA(cudaEventCreate(&e0));
    MB_CUDA(cudaEventCreate(&e1));
    {
        double elapsed = 0.0;
        int iter       = 0;
        while (elapsed < seconds * 1e3 || iter < 3)
        {
            MB_CUDA(cudaEventRecord(e0));
            if (precision == 32) ffma_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision == 322) ffma2_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision >= 2000 && (precision / 10) % 10 == 0) mma_mix_kernel<0><<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision >= 2000 && (precision / 10) % 10 == 2) mma_mix_kernel<2><<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision >= 2000 && (precision / 10) % 10 == 4) mma_mix_kernel<4><<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision >= 2000) mma_mix_kernel<8><<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision == 196 || precision == 198) mma_rotate_kernel<false><<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision == 197) mma_rotate_kernel<true><<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision == 195) mma_tf32_chain_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision == 19 || precision == 191 || precision == 194) mma_tf32_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else if (precision == 16) mma_bf16_kernel<<<blocks, threads>>>(static_cast<float*>(out), outer, d_cyc);
            else dfma_kernel<<<blocks, threads>>>(static_cast<double*>(out), outer, d_cyc);
            MB_CUDA(cudaEventRecord(e1));
            MB_CUDA(cudaEventSynchronize(e1));
            float ms = 0.f;
            MB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            elapsed += ms;
            double fmas = (double)blocks * threads * (double)outer * INNER * (precision == 322 ? 2.0 : 1.0);
            if (precision == 19 || precision == 191 || precision == 194 || precision == 195) fmas = (double)blocks * (threads / 32) * (double)outer * 32.0 * 1024.0;
            if (precision == 196 || precision == 198 || precision >= 2000) fmas = (double)blocks * (threads / 32) * (double)outer * 48.0 * 1024.0;
            if (precision == 197) fmas = (double)blocks * (threads / 32) * (double)outer * 48.0 * 2048.0;
            if (precision == 16) fmas = (double)blocks * (threads / 32) * (double)outer * 32.0 * 2048.0;
            const double tf   = 2.0 * fmas / (ms * 1e-3) / 1e12;
            if (iter > 0 && tf > best)
            {
                best = tf;
                long long cyc = 0;
                MB_CUDA(cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost));
                mhz = (double)cyc / (ms * 1e3);
            }
            if (ms < 20.f && outer < (1 << 22)) outer *= 2;
            ++iter;
            if (iter > 10000) break;
        }
    }
    *tflops = best;
    if (sm_mhz_est) *sm_mhz_est = mhz;
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
