// ll_ops.cuh — arithmetic policies of the episode kernels.
//
// LL_MODE_EXACT reproduces the CPU path operation by operation: every +,-,*,/ and sqrt is an individually rounded
// IEEE operation (the __f*_rn / __d*_rn intrinsics are never contracted into FMAs by nvcc), dot products run in
// ascending k with a separate multiply and add, and the activation is the deterministic ll_math.h routine evaluated
// in double — exactly what the checker in oracle/ does (oracle/shim/ann_mlp_ga_v1.h feedforward()).
// LL_MODE_FAST (float only) uses fused multiply-add, MUFU tanh and reciprocal multiplies.
#ifndef LL_OPS_CUH_
#define LL_OPS_CUH_

#include "ll_internal.h"
#include "ll_math.h"

template <typename T, int MODE> struct ll_ops;

template <> struct ll_ops<float, LL_MODE_EXACT>
{
    static constexpr bool fast = false;
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
    static __device__ __forceinline__ float madd(float acc, float w, float a) { return __fadd_rn(acc, __fmul_rn(w, a)); }
    static __device__ __forceinline__ float act(float z, int id)
    {
        const double zd = (double)z;
        return (float)(id == LL_ACT_TANH ? ll_tanh(zd) : ll_sigmoid(zd));
    }
    template <int N> static __device__ __forceinline__ void act_n(float (&v)[N], int id)
    {
        if (id == LL_ACT_TANH)
        {
            constexpr int B = N > 8 ? 8 : N; // 8 double chains at a time (registers)
#pragma unroll
            for (int o = 0; o < N; o += B)
            {
                double x[B], y[B];
#pragma unroll
                for (int i = 0; i < B; ++i) x[i] = (double)v[o + i];
                ll_tanh_n<B>(x, y);
#pragma unroll
                for (int i = 0; i < B; ++i) v[o + i] = (float)y[i];
            }
        }
        else
        {
#pragma unroll
            for (int i = 0; i < N; ++i) v[i] = act(v[i], id);
        }
    }
};

template <> struct ll_ops<double, LL_MODE_EXACT>
{
    static constexpr bool fast = false;
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
    static __device__ __forceinline__ double madd(double acc, double w, double a) { return __dadd_rn(acc, __dmul_rn(w, a)); }
    static __device__ __forceinline__ double act(double z, int id) { return id == LL_ACT_TANH ? ll_tanh(z) : ll_sigmoid(z); }
    template <int N> static __device__ __forceinline__ void act_n(double (&v)[N], int id)
    { // the N activations of a tile of neurons: tanh chains side by side (ll_tanh_n), bit-identical to act() on each
        if (id == LL_ACT_TANH)
        {
            double y[N];
            ll_tanh_n<N>(v, y);
#pragma unroll
            for (int i = 0; i < N; ++i) v[i] = y[i];
        }
        else
        {
#pragma unroll
            for (int i = 0; i < N; ++i) v[i] = act(v[i], id);
        }
    }
};

__device__ __forceinline__ float ll_tanh_approx(float x)
{
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <> struct ll_ops<float, LL_MODE_FAST>
{
    static constexpr bool fast = true;
    static __device__ __forceinline__ float add(float a, float b) { return a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return a - b; }
    static __device__ __forceinline__ float mul(float a, float b) { return a * b; }
    static __device__ __forceinline__ float div(float a, float b) { return __fdividef(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return sqrtf(a); }
    static __device__ __forceinline__ float madd(float acc, float w, float a) { return fmaf(w, a, acc); }
    static __device__ __forceinline__ float act(float z, int id)
    {
        // sigmoid(z) = 0.5 * tanh(z / 2) + 0.5
        return id == LL_ACT_TANH ? ll_tanh_approx(z) : fmaf(0.5f, ll_tanh_approx(0.5f * z), 0.5f);
    }
    template <int N> static __device__ __forceinline__ void act_n(float (&v)[N], int id)
    {
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] = act(v[i], id);
    }
};

// ---- TMA (bulk async copy) + mbarrier wrappers: global -> shared staging of member rows ------------------------
__device__ __forceinline__ uint32_t ll_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void ll_mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ll_smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void ll_mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ll_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ll_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     ll_smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(ll_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void ll_mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(ll_smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!done);
}

#endif
