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

LL_HD ll_u32x4 ll_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
LL_UNROLL
    for (int round = 0; round < 10; ++round)
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
