// ll_math.h — deterministic transcendental functions shared by the CUDA kernels (nvcc, sm_100a)
// and by the CPU checker (g++).  Every function here is built from IEEE-754 +,-,*,/ and sqrt on
// doubles plus integer bit manipulation, evaluated in a fixed order with NO fused multiply-add,
// so host and device produce bit-identical results.
//
// Why this exists: the policy network of the reference lives in the un-vendored
// ma-libs submodule and uses the C library's tanh; glibc's tanh (x86-64 ifunc/FMA variants) and
// CUDA's tanh differ in the last ulp, and one ulp can flip an argmax (reference call site:
// src/nn_engine_trainer.cpp:501-502).  Both sides of every parity test therefore use ll_tanh.
//
// Build rules: device code must be compiled with -fmad=false (or use the LL_MUL/LL_ADD wrappers
// below, which are contraction-proof on the device); host code with -ffp-contract=off.
#ifndef LL_MATH_H_
#define LL_MATH_H_

#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define LL_HD __host__ __device__ __forceinline__
#else
#define LL_HD inline
#endif
#if defined(__CUDA_ARCH__)
#define LL_UNROLL _Pragma("unroll")
#else
#define LL_UNROLL
#endif

#if defined(__CUDA_ARCH__)
#define LL_MUL(a, b) __dmul_rn((a), (b))
#define LL_ADD(a, b) __dadd_rn((a), (b))
#define LL_SUB(a, b) __dsub_rn((a), (b))
#define LL_DIV(a, b) __ddiv_rn((a), (b))
#define LL_SQRT(a) __dsqrt_rn((a))
#else
#include <math.h>
#define LL_MUL(a, b) ((a) * (b))
#define LL_ADD(a, b) ((a) + (b))
#define LL_SUB(a, b) ((a) - (b))
#define LL_DIV(a, b) ((a) / (b))
#define LL_SQRT(a) sqrt((a))
#endif

LL_HD uint64_t ll_d2u(double x)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t u;
    memcpy(&u, &x, 8);
    return u;
#endif
}

LL_HD double ll_u2d(uint64_t u)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double x;
    memcpy(&x, &u, 8);
    return x;
#endif
}

// 2^k for -1022 <= k <= 1023, exact.
LL_HD double ll_pow2i(int k) { return ll_u2d((uint64_t)(k + 1023) << 52); }

// expm1 on the reduced range |r| <= ln2/2: r + r^2 * (1/2 + r/6 + ... + r^12/14!), Horner, no FMA.
// Truncation error r^15/15! < 1e-19.
LL_HD double ll_expm1_reduced(double r)
{
    double q = 1.1470745597729725e-11;                  // 1/14!
    q        = LL_ADD(LL_MUL(q, r), 1.6059043836821613e-10); // 1/13!
    q        = LL_ADD(LL_MUL(q, r), 2.08767569878681e-09);   // 1/12!
    q        = LL_ADD(LL_MUL(q, r), 2.505210838544172e-08);  // 1/11!
    q        = LL_ADD(LL_MUL(q, r), 2.755731922398589e-07);  // 1/10!
    q        = LL_ADD(LL_MUL(q, r), 2.7557319223985893e-06); // 1/9!
    q        = LL_ADD(LL_MUL(q, r), 2.48015873015873e-05);   // 1/8!
    q        = LL_ADD(LL_MUL(q, r), 0.0001984126984126984);  // 1/7!
    q        = LL_ADD(LL_MUL(q, r), 0.001388888888888889);   // 1/6!
    q        = LL_ADD(LL_MUL(q, r), 0.008333333333333333);   // 1/5!
    q        = LL_ADD(LL_MUL(q, r), 0.041666666666666664);   // 1/4!
    q        = LL_ADD(LL_MUL(q, r), 0.16666666666666666);    // 1/3!
    q        = LL_ADD(LL_MUL(q, r), 0.5);                    // 1/2!
    return LL_ADD(r, LL_MUL(LL_MUL(r, r), q));
}

// Range reduction y = k*ln2 + r with |r| <= ln2/2 (Cody-Waite, two-part ln2).  |y| must be < 700.
LL_HD double ll_reduce_ln2(double y, int* k_out)
{
    const double invln2 = 1.4426950408889634;       // 0x1.71547652b82fep+0
    const double ln2_hi = 6.93147180369123816490e-01; // 0x1.62e42fee00000p-1 (low 21 bits zero)
    const double ln2_lo = 1.90821492927058770002e-10; // 0x1.a39ef35793c76p-33
    double t            = LL_MUL(y, invln2);
    int k               = (int)(t < 0.0 ? LL_SUB(t, 0.5) : LL_ADD(t, 0.5)); // round half away, via truncation
    double kd           = (double)k;
    double r            = LL_SUB(LL_SUB(y, LL_MUL(kd, ln2_hi)), LL_MUL(kd, ln2_lo));
    *k_out              = k;
    return r;
}

// tanh(x) = em1 / (em1 + 2), em1 = expm1(2|x|) = 2^k * p + (2^k - 1), p = expm1(r).
// Agrees with glibc 2.39 tanh to <= 2 ulp (measured in tests/test_ll_math.py).
LL_HD double ll_tanh(double x)
{
    uint64_t ux   = ll_d2u(x);
    uint64_t sign = ux & 0x8000000000000000ull;
    uint64_t uax  = ux & 0x7fffffffffffffffull;
    double ax     = ll_u2d(uax);
    if (uax > 0x7ff0000000000000ull) return LL_ADD(x, x); // NaN
    if (uax >= 0x4036000000000000ull) return ll_u2d(sign | 0x3ff0000000000000ull); // |x| >= 22 (incl. inf): +-1
    if (uax < 0x3e30000000000000ull) return x;            // |x| < 2^-28: tanh(x) = x to double precision
    int k;
    double r   = ll_reduce_ln2(LL_ADD(ax, ax), &k);
    double p   = ll_expm1_reduced(r);
    double em1 = p;
    if (k != 0)
    {
        double s = ll_pow2i(k); // k in [0, 64)
        em1      = LL_ADD(LL_MUL(s, p), LL_SUB(s, 1.0));
    }
    double z = LL_DIV(em1, LL_ADD(em1, 2.0));
    return ll_u2d(ll_d2u(z) | sign);
}

// ll_tanh of N values at once, stage by stage and without branches: a scalar ll_tanh is a chain of ~45 dependent operations and
// a division, N chains side by side fill the pipeline (the exact episode kernel calls this with the 8 or 16 neurons of a tile).
// Every element goes through exactly the operations of ll_tanh in the same order, so the results are bit-identical; elements in
// the special ranges (NaN, |x| >= 22, |x| < 2^-28) run the main path on a harmless stand-in and are replaced at the end.
template <int N> LL_HD void ll_tanh_n(const double (&x)[N], double (&y)[N])
{
    const double invln2 = 1.4426950408889634, ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
    const double c[13] = {1.1470745597729725e-11, 1.6059043836821613e-10, 2.08767569878681e-09,  2.505210838544172e-08,  2.755731922398589e-07,
                          2.7557319223985893e-06, 2.48015873015873e-05,   0.0001984126984126984, 0.001388888888888889,   0.008333333333333333,
                          0.041666666666666664,   0.16666666666666666,    0.5};
    uint64_t sign[N], uax[N];
    double r[N], q[N], kd[N];
    int k[N];
LL_UNROLL
    for (int i = 0; i < N; ++i)
    {
        const uint64_t ux = ll_d2u(x[i]);
        sign[i]           = ux & 0x8000000000000000ull;
        uax[i]            = ux & 0x7fffffffffffffffull;
        const bool main_r = uax[i] < 0x4036000000000000ull && uax[i] >= 0x3e30000000000000ull;
        const double ax   = main_r ? ll_u2d(uax[i]) : 1.0;
        r[i]              = LL_ADD(ax, ax); // y = 2|x|
    }
LL_UNROLL
    for (int i = 0; i < N; ++i) q[i] = LL_MUL(r[i], invln2);
LL_UNROLL
    for (int i = 0; i < N; ++i) k[i] = (int)(q[i] < 0.0 ? LL_SUB(q[i], 0.5) : LL_ADD(q[i], 0.5)), kd[i] = (double)k[i];
LL_UNROLL
    for (int i = 0; i < N; ++i) r[i] = LL_SUB(LL_SUB(r[i], LL_MUL(kd[i], ln2_hi)), LL_MUL(kd[i], ln2_lo));
LL_UNROLL
    for (int i = 0; i < N; ++i) q[i] = c[0];
LL_UNROLL
    for (int j = 1; j < 13; ++j)
LL_UNROLL
        for (int i = 0; i < N; ++i) q[i] = LL_ADD(LL_MUL(q[i], r[i]), c[j]);
LL_UNROLL
    for (int i = 0; i < N; ++i) q[i] = LL_ADD(r[i], LL_MUL(LL_MUL(r[i], r[i]), q[i])); // p = expm1(r)
LL_UNROLL
    for (int i = 0; i < N; ++i)
    {
        const double s = ll_pow2i(k[i]);
        const double e = LL_ADD(LL_MUL(s, q[i]), LL_SUB(s, 1.0));
        q[i]           = k[i] != 0 ? e : q[i]; // em1
    }
LL_UNROLL
    for (int i = 0; i < N; ++i) r[i] = LL_ADD(q[i], 2.0);
LL_UNROLL
    for (int i = 0; i < N; ++i) q[i] = LL_DIV(q[i], r[i]);
LL_UNROLL
    for (int i = 0; i < N; ++i)
    {
        double z = ll_u2d(ll_d2u(q[i]) | sign[i]);
        if (uax[i] < 0x3e30000000000000ull) z = x[i];
        if (uax[i] >= 0x4036000000000000ull) z = ll_u2d(sign[i] | 0x3ff0000000000000ull);
        if (uax[i] > 0x7ff0000000000000ull) z = LL_ADD(x[i], x[i]);
        y[i] = z;
    }
}

// exp(x) for |x| < 700, used by ll_sigmoid.
LL_HD double ll_exp(double x)
{
    int k;
    double r = ll_reduce_ln2(x, &k);
    double e = LL_ADD(ll_expm1_reduced(r), 1.0);
    return LL_MUL(e, ll_pow2i(k));
}

// logistic 1/(1+exp(-x)) (activation id 0, config.txt:40 "SIGMOID=0, TANH=1").
LL_HD double ll_sigmoid(double x)
{
    if (x != x) return x;
    if (x > 40.0) return 1.0;
    if (x < -700.0) return 0.0;
    return LL_DIV(1.0, LL_ADD(1.0, ll_exp(-x)));
}

// Natural log for finite x > 0 (normal numbers): x = 2^k * m, m in [sqrt(1/2), sqrt(2)),
// s = (m-1)/(m+1), log m = 2 atanh(s) = 2s (1 + s^2/3 + s^4/5 + ... + s^24/25).
LL_HD double ll_log(double x)
{
    const double ln2_hi = 6.93147180369123816490e-01;
    const double ln2_lo = 1.90821492927058770002e-10;
    uint64_t ux         = ll_d2u(x);
    int k               = (int)(ux >> 52) - 1023;
    uint64_t mant       = ux & 0x000fffffffffffffull;
    if (mant >= 0x6a09e667f3bcdull)
    { // m >= sqrt(2): halve
        k += 1;
        ux = mant | 0x3fe0000000000000ull;
    }
    else { ux = mant | 0x3ff0000000000000ull; }
    double m  = ll_u2d(ux);
    double f  = LL_SUB(m, 1.0);
    double s  = LL_DIV(f, LL_ADD(m, 1.0));
    double z  = LL_MUL(s, s);
    double q  = 0.04;                                      // 1/25
    q         = LL_ADD(LL_MUL(q, z), 0.043478260869565216); // 1/23
    q         = LL_ADD(LL_MUL(q, z), 0.047619047619047616); // 1/21
    q         = LL_ADD(LL_MUL(q, z), 0.05263157894736842);  // 1/19
    q         = LL_ADD(LL_MUL(q, z), 0.058823529411764705); // 1/17
    q         = LL_ADD(LL_MUL(q, z), 0.06666666666666667);  // 1/15
    q         = LL_ADD(LL_MUL(q, z), 0.07692307692307693);  // 1/13
    q         = LL_ADD(LL_MUL(q, z), 0.09090909090909091);  // 1/11
    q         = LL_ADD(LL_MUL(q, z), 0.1111111111111111);   // 1/9
    q         = LL_ADD(LL_MUL(q, z), 0.14285714285714285);  // 1/7
    q         = LL_ADD(LL_MUL(q, z), 0.2);                  // 1/5
    q         = LL_ADD(LL_MUL(q, z), 0.3333333333333333);   // 1/3
    double t  = LL_MUL(LL_MUL(s, z), q);                    // s^3/3 + s^5/5 + ...
    double lm = LL_MUL(2.0, LL_ADD(s, t));
    double kd = (double)k;
    return LL_ADD(LL_MUL(kd, ln2_hi), LL_ADD(lm, LL_MUL(kd, ln2_lo)));
}

#endif // LL_MATH_H_
