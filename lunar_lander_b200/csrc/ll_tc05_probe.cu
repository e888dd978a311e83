// ll_tc05_probe.cu — known-answer tests and microbenchmarks of the tcgen05 path used by ll_episode_tc05.cu.
//   test 0  D[128 x N] = A[128 x K] * B[N x K]^T on ONE tcgen05.mma chain; operands in the no-swizzle canonical layouts
//           (K-major or MN-major each), D read back with tcgen05.ld.32x32b — pins the descriptor conventions of ll_tc05.cuh
//           against a host product (small integers: exact in f16 and in the fp32 accumulator).
//   test 1  issue rate: `iters` back-to-back tcgen05.mma (M, N, K = 16) from one thread, accumulators rotating over TMEM,
//           optionally next to warps running an FFMA + MUFU.TANH stream (the epilogue work of the episode kernel) — the number
//           DESIGN.md used to assert without measuring: what the legacy HMMA blocks (its sub-partition's issue port) the
//           asynchronous tensor pipe does not.
//   test 2  hop latency: mma -> commit -> mbarrier wait -> tcgen05.ld -> wait::ld -> barrier, repeated (one dependent layer).
#include <cuda_fp16.h>

#include <algorithm>
#include <vector>

#include "ll_internal.h"
#include "ll_tc05.cuh"

namespace
{
    struct probe_args
    {
        int test, m_instr, n, k, a_mn, b_mn, swap, iters, n_acc, busy_warps, busy_iters, mufu_every, group_threads, ts, lane_off;
    };

    __device__ __forceinline__ float tanh_approx(float x)
    {
        float y;
        asm volatile("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    }

    __global__ void __launch_bounds__(1024, 1) tc05_probe_kernel(const __grid_constant__ probe_args p, const __half* __restrict__ A, const __half* __restrict__ B,
                                                                  float* __restrict__ D, unsigned long long* __restrict__ out)
    {
        extern __shared__ __align__(1024) unsigned char smem[];
        __shared__ __align__(8) unsigned long long mbar;
        __shared__ uint32_t tmem_base_s;
        __shared__ int timeout_s;
        const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
        const uint32_t mbar_a = tc05::smem_u32(&mbar);
        unsigned char* sA = smem;
        unsigned char* sB = smem + 64 * 1024;
        const int K = p.k, N = p.n;

        if (tid == 0)
        {
            tc05::mbar_init(mbar_a, 1);
            tc05::fence_mbar_init();
            timeout_s = 0;
        }
        if (warp == 0) tc05::tmem_alloc(tc05::smem_u32(&tmem_base_s), 512);
        // operands: canonical no-swizzle layouts (ll_tc05.cuh)
        const uint32_t a_lbo = p.a_mn ? 128u : 128u * 16u, a_sbo = p.a_mn ? (uint32_t)(K / 8) * 128u : 128u;       // A has 128 rows
        const uint32_t b_lbo = p.b_mn ? 128u : 128u * (uint32_t)(N / 8), b_sbo = p.b_mn ? (uint32_t)(K / 8) * 128u : 128u;
        for (int i = tid; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sA)[i] = 0u;
        for (int i = tid; i < 32 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sB)[i] = 0u;
        __syncthreads();
        if (p.test == 0)
        {
            for (int i = tid; i < 128 * K; i += blockDim.x)
            {
                const int r = i / K, k = i % K;
                const uint32_t off = p.a_mn ? (r / 8) * a_sbo + (k / 8) * a_lbo + (k % 8) * 16 + (r % 8) * 2 : (r / 8) * a_sbo + (k / 8) * a_lbo + (r % 8) * 16 + (k % 8) * 2;
                *reinterpret_cast<__half*>(sA + off) = A[i];
            }
            for (int i = tid; i < N * K; i += blockDim.x)
            {
                const int r = i / K, k = i % K;
                const uint32_t off = p.b_mn ? (r / 8) * b_sbo + (k / 8) * b_lbo + (k % 8) * 16 + (r % 8) * 2 : (r / 8) * b_sbo + (k / 8) * b_lbo + (r % 8) * 16 + (k % 8) * 2;
                *reinterpret_cast<__half*>(sB + off) = B[i];
            }
        }
        tc05::fence_proxy_async();
        tc05::fence_before_sync();
        __syncthreads();
        tc05::fence_after_sync();
        const uint32_t tbase = tmem_base_s;
        const uint32_t idesc = tc05::idesc_f16(p.m_instr, N, p.a_mn, p.b_mn);
        const uint32_t a_l = p.swap ? a_sbo : a_lbo, a_s = p.swap ? a_lbo : a_sbo, b_l = p.swap ? b_sbo : b_lbo, b_s = p.swap ? b_lbo : b_sbo;
        const uint32_t a_base = tc05::smem_u32(sA), b_base = tc05::smem_u32(sB);

        if (p.test == 0)
        {
            if (warp < 4)
            { // sentinel in every lane / column the dump reads
                uint32_t s[4];
                for (int j = 0; j < 4; ++j) s[j] = __float_as_uint(-12345.f);
                for (int c = 0; c < N; c += 4) tc05::st_x4(tbase + ((uint32_t)(32 * warp) << 16) + c, s);
                tc05::wait_st();
            }
            const uint32_t a_col0 = 256u; // TS mode: A lives in TMEM columns [256, 256 + K / 2), lane = row, column = (k even | k odd << 16)
            if (p.ts && warp < 4)
            {
                int row = -1;
                if (p.m_instr == 128) row = tid;
                else if (lane >= p.lane_off && lane < p.lane_off + 16) row = warp * 16 + lane - p.lane_off;
                for (int c = 0; c < K / 2; c += 8)
                {
                    uint32_t w[8];
                    for (int j = 0; j < 8; ++j)
                    {
                        const int k0 = 2 * (c + j);
                        const uint32_t lo16 = row >= 0 ? (uint32_t)__half_as_ushort(A[row * K + k0]) : 0u, hi16 = row >= 0 ? (uint32_t)__half_as_ushort(A[row * K + k0 + 1]) : 0u;
                        w[j] = lo16 | (hi16 << 16);
                    }
                    tc05::st_x8(tbase + ((uint32_t)(32 * warp) << 16) + a_col0 + c, w);
                }
                tc05::wait_st();
            }
            tc05::fence_before_sync();
            __syncthreads();
            tc05::fence_after_sync();
            if (tid == 0)
            {
                const uint32_t loff = (p.m_instr == 64 ? (uint32_t)p.lane_off : 0u) << 16;
                for (int k16 = 0; k16 < K / 16; ++k16)
                {
                    if (p.ts) tc05::mma_f16_ts(tbase + loff, tbase + loff + a_col0 + 8 * k16, tc05::smem_desc(b_base + 2 * k16 * b_lbo, b_l, b_s), idesc, k16 > 0);
                    else tc05::mma_f16_ss(tbase + loff, tc05::smem_desc(a_base + 2 * k16 * a_lbo, a_l, a_s), tc05::smem_desc(b_base + 2 * k16 * b_lbo, b_l, b_s), idesc, k16 > 0);
                }
                tc05::commit(mbar_a);
            }
            if (!tc05::mbar_wait(mbar_a, 0)) timeout_s = 1;
            tc05::fence_after_sync();
            if (warp < 4)
            {
                for (int c = 0; c < N; c += 16)
                {
                    uint32_t v[16];
                    tc05::ld_x16(tbase + ((uint32_t)(32 * warp) << 16) + c, v);
                    tc05::wait_ld();
                    for (int j = 0; j < 16; ++j) D[(size_t)(32 * warp + lane) * N + c + j] = __uint_as_float(v[j]);
                }
            }
        }
        else if (p.test == 1)
        {
            long long t0 = 0, t_issue = 0, t1 = 0, tb0 = 0, tb1 = 0;
            float sink = 0.f;
            __syncthreads();
            if (warp == 0 && p.iters > 0)
            { // warp-uniform issue loop, one elected lane issues (n_acc is a power of two)
                t0 = clock64();
                const uint64_t ad = tc05::smem_desc(a_base, a_l, a_s), bd = tc05::smem_desc(b_base, b_l, b_s);
                const int mask = p.n_acc - 1;
                for (int i = 0; i < p.iters; i += 4)
                {
                    if (tc05::elect_one())
                    {
                        if (p.ts)
                        { // A from TMEM columns [480, 488)
                            tc05::mma_f16_ts(tbase + (uint32_t)(((i + 0) & mask) * N), tbase + 480u, bd, idesc, 1u);
                            tc05::mma_f16_ts(tbase + (uint32_t)(((i + 1) & mask) * N), tbase + 480u, bd, idesc, 1u);
                            tc05::mma_f16_ts(tbase + (uint32_t)(((i + 2) & mask) * N), tbase + 480u, bd, idesc, 1u);
                            tc05::mma_f16_ts(tbase + (uint32_t)(((i + 3) & mask) * N), tbase + 480u, bd, idesc, 1u);
                        }
                        else
                        {
                            tc05::mma_f16_ss(tbase + (uint32_t)(((i + 0) & mask) * N), ad, bd, idesc, 1u);
                            tc05::mma_f16_ss(tbase + (uint32_t)(((i + 1) & mask) * N), ad, bd, idesc, 1u);
                            tc05::mma_f16_ss(tbase + (uint32_t)(((i + 2) & mask) * N), ad, bd, idesc, 1u);
                            tc05::mma_f16_ss(tbase + (uint32_t)(((i + 3) & mask) * N), ad, bd, idesc, 1u);
                        }
                    }
                    __syncwarp();
                }
                t_issue = clock64();
                if (tc05::elect_one()) tc05::commit(mbar_a);
                __syncwarp();
                if (!tc05::mbar_wait(mbar_a, 0)) timeout_s = 1;
                t1 = clock64();
            }
            else if (warp >= 4 && warp < 4 + p.busy_warps)
            { // the epilogue-like stream: dependent-free FFMA chains with a MUFU.TANH every `mufu_every` FFMAs
                float a0 = 1.0f + 1e-6f * tid, a1 = 0.5f, a2 = 0.25f, a3 = 0.125f, x = 1e-3f * lane;
                const float m = 0.999f, c = 1e-4f;
                tb0 = clock64();
                for (int i = 0; i < p.busy_iters; ++i)
                {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                    {
                        a0 = fmaf(a0, m, c), a1 = fmaf(a1, m, c), a2 = fmaf(a2, m, c), a3 = fmaf(a3, m, c);
                        if (p.mufu_every > 0 && (j % p.mufu_every) == 0) x = tanh_approx(x + a0);
                    }
                }
                tb1  = clock64();
                sink = a0 + a1 + a2 + a3 + x;
            }
            if (tid == 0)
            {
                out[blockIdx.x * 8 + 0] = (unsigned long long)(t1 - t0);
                out[blockIdx.x * 8 + 1] = (unsigned long long)(t_issue - t0);
            }
            if (warp == 4 && lane == 0)
            {
                out[blockIdx.x * 8 + 2] = (unsigned long long)(tb1 - tb0);
                out[blockIdx.x * 8 + 3] = (unsigned long long)__float_as_uint(sink);
            }
        }
        else if (p.test == 2)
        { // one dependent layer hop, `group_threads` (32..128) threads taking part
            const int G = p.group_threads;
            long long t0 = 0, t1 = 0, seg_mma = 0, seg_ld = 0, seg_bar = 0;
            uint32_t acc = 0;
            if (tid < G)
            {
                t0 = clock64();
                const uint64_t ad = tc05::smem_desc(a_base, a_l, a_s), bd = tc05::smem_desc(b_base, b_l, b_s);
                for (int i = 0; i < p.iters && !timeout_s; ++i)
                {
                    const long long ta = clock64();
                    if (warp == 0)
                    {
                        tc05::fence_after_sync();
                        if (tc05::elect_one())
                        {
                            for (int j = 0; j < p.n_acc; ++j) // n_acc = MMAs per hop here
                            {
                                if (p.ts) tc05::mma_f16_ts(tbase, tbase + 480u, bd, idesc, (uint32_t)(j > 0));
                                else tc05::mma_f16_ss(tbase, ad, bd, idesc, (uint32_t)(j > 0));
                            }
                            tc05::commit(mbar_a);
                        }
                        __syncwarp();
                    }
                    if (!tc05::mbar_wait(mbar_a, (uint32_t)(i & 1))) timeout_s = 1;
                    tc05::fence_after_sync();
                    const long long tb = clock64();
                    uint32_t v[16];
                    tc05::ld_x16(tbase + ((uint32_t)(32 * warp) << 16), v);
                    tc05::wait_ld();
                    acc += v[0] + v[15];
                    const long long tc = clock64();
                    tc05::fence_proxy_async();
                    tc05::fence_before_sync();
                    tc05::named_bar_sync(1, G);
                    const long long td = clock64();
                    seg_mma += tb - ta, seg_ld += tc - tb, seg_bar += td - tc;
                }
                t1 = clock64();
            }
            if (tid == 0)
            {
                out[blockIdx.x * 8 + 1] = (unsigned long long)seg_mma, out[blockIdx.x * 8 + 2] = (unsigned long long)seg_ld;
                out[blockIdx.x * 8 + 4] = (unsigned long long)seg_bar;
            }
            if (tid == 0) out[blockIdx.x * 8 + 0] = (unsigned long long)(t1 - t0), out[blockIdx.x * 8 + 3] = acc;
        }
        tc05::fence_before_sync();
        __syncthreads();
        if (tid == 0) out[blockIdx.x * 8 + 7] = (unsigned long long)timeout_s;
        if (warp == 0) tc05::tmem_dealloc(tbase, 512);
    }
} // namespace

#define PB_CUDA(call)                                                                                                 \
    do {                                                                                                              \
        cudaError_t e__ = (call);                                                                                     \
        if (e__ != cudaSuccess)                                                                                       \
        {                                                                                                             \
            rc = ll_fail(nullptr, LL_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__));                     \
            goto done;                                                                                                \
        }                                                                                                             \
    } while (0)

// args[16]: test, m_instr, n, k, a_mn, b_mn, swap, iters, n_acc, busy_warps, busy_iters, mufu_every, group_threads, all_sms,
//           ts (1: A operand from TMEM), lane_off (M = 64: 0 or 16, the half of each 32-lane sub-partition the tile lives in)
// d_out: test 0 -> [128][n] floats (the raw TMEM dump, lane-major); result[16]:
//   test 0: [0] max |D - ref| over rows < m_instr, [1] mismatching elements, [2] time-out flag
//   test 1: [0] cycles per mma (issue -> completion), [1] cycles per mma (issue loop only), [2] busy-stream cycles, [3] ms, [4] time-out
//   test 2: [0] cycles per hop, [1] of which issue -> mbarrier completion, [2] tcgen05.ld + wait::ld, [3] fences + named barrier, [4] time-out
//           (n_acc = tcgen05.mma per hop)
extern "C" int ll_tc05_probe(int32_t device, const int32_t* args, float* d_out, double* result)
{
    if (!args || !result) return ll_fail(nullptr, LL_EINVAL, "ll_tc05_probe: null argument");
    int rc = LL_OK, n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    {
        cudaGetLastError();
        return ll_fail(nullptr, LL_ENODEVICE, "ll_tc05_probe: no CUDA device");
    }
    probe_args p{};
    p.test = args[0], p.m_instr = args[1], p.n = args[2], p.k = args[3], p.a_mn = args[4], p.b_mn = args[5], p.swap = args[6], p.iters = args[7];
    p.n_acc = std::max(1, args[8]), p.busy_warps = args[9], p.busy_iters = args[10], p.mufu_every = args[11], p.group_threads = args[12];
    p.ts = args[14], p.lane_off = args[15];
    const int all_sms = args[13];
    if ((p.m_instr != 64 && p.m_instr != 128) || p.n < 8 || p.n > 256 || (p.n % 8) || p.k < 16 || p.k > 128 || (p.k % 16) || (p.test != 2 && p.n_acc * p.n > 480) || (p.lane_off != 0 && p.lane_off != 16) || p.n * p.k * 2 > 32 * 1024 || p.busy_warps < 0 ||
        p.busy_warps > 12 || (p.test == 2 && (p.group_threads < 32 || p.group_threads > 128 || (p.group_threads % 32))))
        return ll_fail(nullptr, LL_EINVAL, "ll_tc05_probe: bad shape");
    if (p.test == 0 && (!d_out || (p.n % 16))) return ll_fail(nullptr, LL_EINVAL, "ll_tc05_probe: test 0 needs d_out and n % 16 == 0");
    __half *dA = nullptr, *dB = nullptr;
    float* dD = nullptr;
    unsigned long long* dO = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaDeviceProp prop;
    std::vector<__half> hA(128 * p.k), hB((size_t)p.n * p.k);
    std::vector<float> fA(128 * p.k), fB((size_t)p.n * p.k), hD((size_t)128 * p.n);
    std::vector<unsigned long long> hO;
    int blocks = 1;
    const int threads = 128 + 32 * p.busy_warps;
    const size_t smem = 96 * 1024;
    float ms = 0.f;
    for (int i = 0; i < 16; ++i) result[i] = 0.0;
    if (device >= 0) PB_CUDA(cudaSetDevice(device));
    PB_CUDA(cudaGetDevice(&device));
    PB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) { rc = ll_fail(nullptr, LL_ENODEVICE, "ll_tc05_probe: needs an sm_100 device"); goto done; }
    if (all_sms) blocks = prop.multiProcessorCount;
    for (int r = 0; r < 128; ++r)
        for (int k = 0; k < p.k; ++k) fA[r * p.k + k] = (float)(((r * 7 + k * 3) % 13) - 6), hA[r * p.k + k] = __float2half(fA[r * p.k + k]);
    for (int r = 0; r < p.n; ++r)
        for (int k = 0; k < p.k; ++k) fB[r * p.k + k] = (float)(((r * 5 + k * 11) % 9) - 4), hB[r * p.k + k] = __float2half(fB[r * p.k + k]);
    PB_CUDA(cudaMalloc(&dA, hA.size() * 2));
    PB_CUDA(cudaMalloc(&dB, hB.size() * 2));
    PB_CUDA(cudaMalloc(&dD, hD.size() * 4));
    PB_CUDA(cudaMalloc(&dO, (size_t)blocks * 8 * 8));
    PB_CUDA(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    PB_CUDA(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    PB_CUDA(cudaMemset(dD, 0, hD.size() * 4));
    PB_CUDA(cudaMemset(dO, 0, (size_t)blocks * 8 * 8));
    PB_CUDA(cudaFuncSetAttribute(tc05_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PB_CUDA(cudaEventCreate(&e0));
    PB_CUDA(cudaEventCreate(&e1));
    PB_CUDA(cudaEventRecord(e0));
    tc05_probe_kernel<<<blocks, threads, smem>>>(p, dA, dB, dD, dO);
    PB_CUDA(cudaGetLastError());
    PB_CUDA(cudaEventRecord(e1));
    PB_CUDA(cudaEventSynchronize(e1));
    PB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    hO.resize((size_t)blocks * 8);
    PB_CUDA(cudaMemcpy(hO.data(), dO, hO.size() * 8, cudaMemcpyDeviceToHost));
    if (p.test == 0)
    {
        PB_CUDA(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
        double worst = 0.0, bad = 0.0;
        for (int r = 0; r < p.m_instr; ++r)
            for (int c = 0; c < p.n; ++c)
            {
                float ref = 0.f;
                for (int k = 0; k < p.k; ++k) ref += fA[r * p.k + k] * fB[c * p.k + k];
                const int dlane = p.m_instr == 128 ? r : (r / 16) * 32 + (r % 16) + p.lane_off; // M = 64: 16 lanes per sub-partition
                const double d = std::abs((double)hD[(size_t)dlane * p.n + c] - (double)ref);
                worst = std::max(worst, d), bad += d != 0.0;
            }
        if (d_out) std::copy(hD.begin(), hD.end(), d_out);
        result[0] = worst, result[1] = bad, result[2] = (double)hO[7];
    }
    else if (p.test == 1)
    {
        double c_all = 0, c_issue = 0, c_busy = 0, to = 0;
        for (int b = 0; b < blocks; ++b) c_all += (double)hO[b * 8 + 0], c_issue += (double)hO[b * 8 + 1], c_busy += (double)hO[b * 8 + 2], to += (double)hO[b * 8 + 7];
        result[0] = p.iters ? c_all / blocks / p.iters : 0.0, result[1] = p.iters ? c_issue / blocks / p.iters : 0.0, result[2] = c_busy / blocks, result[3] = ms, result[4] = to;
    }
    else
    {
        double c_all = 0, to = 0;
        for (int b = 0; b < blocks; ++b) c_all += (double)hO[b * 8 + 0], to += (double)hO[b * 8 + 7];
        const double it = (double)std::max(1, p.iters);
        result[0] = c_all / blocks / it, result[1] = (double)hO[1] / it, result[2] = (double)hO[2] / it, result[3] = (double)hO[4] / it, result[4] = to;
    }
done:
    if (dA) cudaFree(dA);
    if (dB) cudaFree(dB);
    if (dD) cudaFree(dD);
    if (dO) cudaFree(dO);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return rc;
}
