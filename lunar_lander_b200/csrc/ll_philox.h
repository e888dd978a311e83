// ll_philox.h — counter-based RNG for the GA kernels (population init, parent pick, Gaussian mutation).
//
// Replaces the std::mt19937 stream inside ma-libs' nn::ANN_MLP_GA<T>::CreatePopulation (reference call
// sites src/nn_engine_trainer.cpp:94, 405; RNG type recovered from the generator line of the
// sample_runs weight files).  A sequential Mersenne stream cannot be sharded over
// GPUs, so every random number is a pure function of (seed, generation, global member id, parameter
// pair index): the same children are produced for any GPU count and any launch geometry.
//
// Philox4x32-10 as published by Salmon, Moraes, Dror, Shaw, "Parallel random numbers: as easy as
// 1, 2, 3" (SC'11).  Host and device results are bit-identical (integer arithmetic; the Gaussian
// transform uses only ll_math.h primitives).
#ifndef LL_PHILOX_H_
#define LL_PHILOX_H_

#include "ll_math.h"

struct ll_u32x4
{
    uint32_t v[4];
};

LL_HD void ll_mulhilo32(uint32_t a, uint32_t b, uint32_t* hi, uint32_t* lo)
{
#if defined(__CUDA_ARCH__)
    *lo = a * b;
    *hi = __umulhi(a, b);
#else
    uint64_t p = (uint64_t)a * (uint64_t)b;
    *lo        = (uint32_t)p;
    *hi        = (uint32_t)(p >> 32);
#endif
}

template <int ROUNDS> LL_HD ll_u32x4 ll_philox4x32_r(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
LL_UNROLL
    for (int round = 0; round < ROUNDS; ++round)
    {
        uint32_t hi0, lo0, hi1, lo1;
        ll_mulhilo32(0xD2511F53u, c0, &hi0, &lo0);
        ll_mulhilo32(0xCD9E8D57u, c2, &hi1, &lo1);
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n1 = lo1;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        uint32_t n3 = lo0;
        c0 = n0, c1 = n1, c2 = n2, c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    ll_u32x4 out;
    out.v[0] = c0, out.v[1] = c1, out.v[2] = c2, out.v[3] = c3;
    return out;
}
LL_HD ll_u32x4 ll_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
    return ll_philox4x32_r<10>(c0, c1, c2, c3, k0, k1);
}

// Stream ids (counter word 3, high byte) so that the different consumers never share counters.
#define LL_STREAM_NORMAL 0x00000000u // low 24 bits: rejection attempt number
#define LL_STREAM_PARENT 0x01000000u
#define LL_KEY1 0x4C4C4742u          // second key word, "LLGB"

// uniform in (-1, 1), 32-bit resolution, never exactly 0 or +-1: (int32(w) + 0.5) * 2^-31
LL_HD double ll_u32_to_sym(uint32_t w) { return LL_MUL(LL_ADD((double)(int32_t)w, 0.5), 4.656612873077392578125e-10); }

// Two independent N(0,1) variates for parameter pair `pair` of global member `member` in generation
// `gen` (Marsaglia polar method; each Philox block carries two attempts).
LL_HD void ll_normal_pair(uint32_t seed, uint32_t gen, uint32_t member, uint32_t pair, double* n0, double* n1)
{
    for (uint32_t attempt = 0;; ++attempt)
    {
        ll_u32x4 w = ll_philox4x32_10(pair, member, gen, LL_STREAM_NORMAL | (attempt & 0x00ffffffu), seed, LL_KEY1);
LL_UNROLL
        for (int h = 0; h < 2; ++h)
        {
            double u = ll_u32_to_sym(w.v[2 * h]);
            double v = ll_u32_to_sym(w.v[2 * h + 1]);
            double s = LL_ADD(LL_MUL(u, u), LL_MUL(v, v));
            if (s < 1.0)
            { // s > 0 always: |u|,|v| >= 2^-32
                double f = LL_SQRT(LL_DIV(LL_MUL(-2.0, ll_log(s)), s));
                *n0      = LL_MUL(u, f);
                *n1      = LL_MUL(v, f);
                return;
            }
        }
    }
}

// ---- fp32 populations: four N(0,1) variates per Philox block, Box-Muller in float ------------------------------------------------
// The fp64 transform above (polar method, ll_log in double) is what makes K4 ALU-bound at 10 % of the HBM roofline for a float
// population; the float build needs no more than float normals.  Deterministic like the rest of this header: IEEE float +,-,*,/
// and sqrt in a fixed order, no FMA, integer bit manipulation — bit-identical on host and device.
//   block (quad, member, gen, LL_STREAM_NORMAL32): words (w0, w1) -> two variates, (w2, w3) -> two more; no rejection
//   radius r = sqrt(-2 ln u), u = ((w >> 9) + 0.5) 2^-23 in (0, 1);  angle 2 pi t: quadrant = top two bits of the second word,
//   t' = ((w >> 7 & 0x7fffff) + 0.5) 2^-23 of a quarter turn, folded onto [0, pi/4] where short Taylor polynomials are exact to 3e-8
#define LL_STREAM_NORMAL32 0x02000000u
#if defined(__CUDA_ARCH__)
#define LL_FMUL(a, b) __fmul_rn((a), (b))
#define LL_FADD(a, b) __fadd_rn((a), (b))
#define LL_FSUB(a, b) __fsub_rn((a), (b))
#define LL_FDIV(a, b) __fdiv_rn((a), (b))
#define LL_FSQRT(a) __fsqrt_rn((a))
#else
#define LL_FMUL(a, b) ((a) * (b))
#define LL_FADD(a, b) ((a) + (b))
#define LL_FSUB(a, b) ((a) - (b))
#define LL_FDIV(a, b) ((a) / (b))
#define LL_FSQRT(a) sqrtf((a))
#endif

LL_HD uint32_t ll_f2u(float x)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(x);
#else
    uint32_t u;
    memcpy(&u, &x, 4);
    return u;
#endif
}
LL_HD float ll_u2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float x;
    memcpy(&x, &u, 4);
    return x;
#endif
}

// ln(x) for a normal float x in (0, 1]: x = m 2^k with m in [sqrt(1/2), sqrt(2)), f = m - 1, ln m = f - f^2/2 + f^3 P(f) with the
// degree-8 polynomial of the Cephes logf (|error| < 1e-7 relative on the interval); no division
LL_HD float ll_logf_det(float x)
{
    uint32_t ux   = ll_f2u(x);
    int k         = (int)(ux >> 23) - 127;
    uint32_t mant = ux & 0x007fffffu;
    if (mant >= 0x003504f3u) k += 1, ux = mant | 0x3f000000u; // m >= sqrt(2): halve
    else ux = mant | 0x3f800000u;
    const float f = LL_FSUB(ll_u2f(ux), 1.0f);
    const float z = LL_FMUL(f, f);
    float q       = 7.0376836292e-2f;
    q             = LL_FADD(LL_FMUL(q, f), -1.1514610310e-1f);
    q             = LL_FADD(LL_FMUL(q, f), 1.1676998740e-1f);
    q             = LL_FADD(LL_FMUL(q, f), -1.2420140846e-1f);
    q             = LL_FADD(LL_FMUL(q, f), 1.4249322787e-1f);
    q             = LL_FADD(LL_FMUL(q, f), -1.6668057665e-1f);
    q             = LL_FADD(LL_FMUL(q, f), 2.0000714765e-1f);
    q             = LL_FADD(LL_FMUL(q, f), -2.4999993993e-1f);
    q             = LL_FADD(LL_FMUL(q, f), 3.3333331174e-1f);
    const float y = LL_FADD(LL_FMUL(LL_FMUL(f, z), q), LL_FMUL(-0.5f, z));
    return LL_FADD(LL_FMUL((float)k, 0.69314718f), LL_FADD(f, y));
}

// (r cos, r sin) of angle 2 pi (quadrant + t) / 4
LL_HD void ll_box_muller_f32(uint32_t w_radius, uint32_t w_angle, float* n0, float* n1)
{
    const float u = LL_FMUL(LL_FADD((float)(w_radius >> 9), 0.5f), 1.1920929e-07f); // 2^-23
    const float r = LL_FSQRT(LL_FMUL(-2.0f, ll_logf_det(u)));
    const uint32_t quad = w_angle >> 30;
    float t             = LL_FMUL(LL_FADD((float)((w_angle >> 7) & 0x007fffffu), 0.5f), 1.1920929e-07f); // (0, 1) of a quarter turn
    const bool fold     = t > 0.5f;
    if (fold) t = LL_FSUB(1.0f, t);
    const float x = LL_FMUL(t, 1.5707964f), x2 = LL_FMUL(x, x); // [0, pi/4]
    float sp      = 2.7557319e-06f;                                 // 1/9!
    sp            = LL_FADD(LL_FMUL(sp, x2), -1.9841270e-04f);      // -1/7!
    sp            = LL_FADD(LL_FMUL(sp, x2), 8.3333333e-03f);       // 1/5!
    sp            = LL_FADD(LL_FMUL(sp, x2), -1.6666667e-01f);      // -1/3!
    float sn      = LL_FADD(x, LL_FMUL(LL_FMUL(x, x2), sp));
    float cp      = 2.4801587e-05f;                                 // 1/8!
    cp            = LL_FADD(LL_FMUL(cp, x2), -1.3888889e-03f);      // -1/6!
    cp            = LL_FADD(LL_FMUL(cp, x2), 4.1666667e-02f);       // 1/4!
    cp            = LL_FADD(LL_FMUL(cp, x2), -0.5f);
    float cs      = LL_FADD(1.0f, LL_FMUL(x2, cp));
    if (fold)
    { // sin(pi/2 - x) = cos x
        const float tmp = sn;
        sn = cs, cs = tmp;
    }
    float c = cs, s_ = sn; // rotate by the quadrant
    if (quad == 1u) c = -sn, s_ = cs;
    else if (quad == 2u) c = -cs, s_ = -sn;
    else if (quad == 3u) c = sn, s_ = -cs;
    *n0 = LL_FMUL(r, c);
    *n1 = LL_FMUL(r, s_);
}

// Four N(0,1) variates for DEVICE-order positions [4 quad, 4 quad + 4) of global member `member` (the device row layout is a
// function of the topology alone, so the draw does not depend on the GPU count or the launch geometry).
LL_HD void ll_normal4_f32(uint32_t seed, uint32_t gen, uint32_t member, uint32_t quad, float* g)
{
    const ll_u32x4 w = ll_philox4x32_10(quad, member, gen, LL_STREAM_NORMAL32, seed, LL_KEY1);
    ll_box_muller_f32(w.v[0], w.v[1], &g[0], &g[1]);
    ll_box_muller_f32(w.v[2], w.v[3], &g[2], &g[3]);
}

#if defined(__CUDACC__)
// ---- LL_MODE_FAST, fp32 populations: the same draw structure (one Philox block per device-order quad -> two Box-Muller pairs) on the
// special-function unit: lg2.approx, sqrt.approx, sin.approx / cos.approx instead of the individually rounded polynomials above, and
// Philox4x32-7 (the smallest round count Salmon et al. report as passing BigCrush; 10 is their default with a safety margin) on its own
// stream.  115 instead of 288 instructions per four parameters.  Deterministic on the device and independent of the GPU count like every
// other draw, but not reproducible bit for bit on the host: fast mode is judged statistically, exact mode keeps
// ll_normal4_f32.
#define LL_STREAM_NORMAL32_FAST 0x03000000u
__device__ __forceinline__ void ll_box_muller_f32_fast(uint32_t w_radius, uint32_t w_angle, float* n0, float* n1)
{
    const float u = __uint_as_float((w_radius >> 9) | 0x3f800000u) - 0.99999994f; // ((w >> 9) + 0.5) 2^-23, exact, in (0, 1)
    const float v = __uint_as_float((w_angle >> 9) | 0x3f800000u) - 1.5f;         // [-0.5, 0.5) of a turn
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(-1.3862944f * __log2f(u))); // sqrt(-2 ln u)
    float sn, cs;
    __sincosf(6.2831855f * v, &sn, &cs);
    *n0 = r * cs;
    *n1 = r * sn;
}
__device__ __forceinline__ void ll_normal4_f32_fast(uint32_t seed, uint32_t gen, uint32_t member, uint32_t quad, float* g)
{
    const ll_u32x4 w = ll_philox4x32_r<7>(quad, member, gen, LL_STREAM_NORMAL32_FAST, seed, LL_KEY1);
    ll_box_muller_f32_fast(w.v[0], w.v[1], &g[0], &g[1]);
    ll_box_muller_f32_fast(w.v[2], w.v[3], &g[2], &g[3]);
}
#endif

// Parent (elite slot) of a mutated child: uniform in [0, top).
LL_HD uint32_t ll_parent_pick(uint32_t seed, uint32_t gen, uint32_t member, uint32_t top)
{
    ll_u32x4 w = ll_philox4x32_10(0u, member, gen, LL_STREAM_PARENT, seed, LL_KEY1);
    return (uint32_t)(((uint64_t)w.v[0] * (uint64_t)top) >> 32);
}

// Which child slots are fresh random individuals (MIXED_WITH_RANDOM_INJECTION): n_inject of the n_children slots, spread
// evenly (child c is fresh iff the running count floor((c + 1) n_inject / n_children) steps at c) instead of bunched at
// the end, so every contiguous shard of the population carries the same mix of long-lived elite children and short-lived
// random individuals (2 GPUs: 1.36 s vs 1.02 s per generation for the shard that held all the elite children).
LL_HD bool ll_is_injected(uint64_t child, uint64_t n_children, uint64_t n_inject)
{
    if (n_inject >= n_children) return true;
    return ((child + 1) * n_inject) / n_children != (child * n_inject) / n_children;
}

#endif // LL_PHILOX_H_
