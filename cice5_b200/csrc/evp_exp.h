/* evp_exp.h -- exp(x) evaluated in double-double arithmetic and rounded once: ONE definition, plain IEEE-754
 * double operations only (+, -, *, fma), compiled into the CUDA kernels (ice_strength_kernel) and into the CPU
 * oracle, so that both sides take bit-identical values through the only transcendental on the path.
 *
 * Why: the reference's ice_strength calls the Fortran intrinsic exp (source/ice_mechred.F90:1108-1122 with
 * krdg_partic = 1, the default; :2263 for kstrength = 0).  Every libm returns its own last bit (glibc even
 * per CPU: its exp is ifunc-dispatched to an FMA variant), CUDA's exp another one, and EVP at ndte = 120
 * amplifies a 1-ulp difference of `strength` far beyond 1e-10 on ill-conditioned instances (round 1: the
 * gx3 "caps" state failed its bar on a B200 for exactly this reason).  A correctly rounded exp is the one
 * value every conforming implementation can agree on.
 *
 * Method: k = nint(x / ln 2); r = x - k ln 2 with ln 2 split into 42 + 42 + 53 bits (the first two products
 * are exact), kept as a double-double; r / 64; Taylor polynomial of expm1 to degree 13 in double-double
 * (|r/64| < 0.0055: truncation < 2^-113); six times p <- 2p + p^2 (expm1 doubling: no cancellation);
 * exp = (1 + p) 2^k, rounded to double once.  Relative error before the final rounding < 2^-96, so the result
 * is the correctly rounded one unless exp(x) lies within 2^-96 (relative) of the midpoint of two doubles;
 * tests/test_oracle_cpu.py checks 2e5 random and structured arguments against exact rational arithmetic.
 * Domain: |x| <= 700 (results are normal numbers); outside it the function returns libm-quality values
 * (0, +inf, or a scaled product) -- the path only needs [-20, 0].
 */
#ifndef EVP_EXP_H
#define EVP_EXP_H

#include <math.h>

#if defined(__CUDACC__) || defined(EVP_EMU)
#define EVP_EXP_FN __device__ __forceinline__
/* explicit round-to-nearest intrinsics: never contracted, whatever -fmad says */
#define EVP_X_ADD(a, b) __dadd_rn((a), (b))
#define EVP_X_SUB(a, b) __dsub_rn((a), (b))
#define EVP_X_MUL(a, b) __dmul_rn((a), (b))
#define EVP_X_FMA(a, b, c) __fma_rn((a), (b), (c))
#define EVP_X_BITS(v) __longlong_as_double((long long)(v))
#else
#include <stdint.h>
#include <string.h>
#define EVP_EXP_FN static inline
/* host: compile with -ffp-contract=off (oracle/Makefile does) */
#define EVP_X_ADD(a, b) ((a) + (b))
#define EVP_X_SUB(a, b) ((a) - (b))
#define EVP_X_MUL(a, b) ((a) * (b))
#define EVP_X_FMA(a, b, c) fma((a), (b), (c))
static inline double evp_x_bits_(uint64_t v) { double d; memcpy(&d, &v, 8); return d; }
#define EVP_X_BITS(v) evp_x_bits_((uint64_t)(v))
#endif

typedef struct { double hi, lo; } evp_dd;

/* s + e = a + b exactly (Knuth) */
EVP_EXP_FN evp_dd evp_two_sum(double a, double b)
{
    evp_dd r;
    const double s = EVP_X_ADD(a, b);
    const double bb = EVP_X_SUB(s, a);
    r.hi = s;
    r.lo = EVP_X_ADD(EVP_X_SUB(a, EVP_X_SUB(s, bb)), EVP_X_SUB(b, bb));
    return r;
}
/* the same for |a| >= |b| (Dekker) */
EVP_EXP_FN evp_dd evp_quick_two_sum(double a, double b)
{
    evp_dd r;
    const double s = EVP_X_ADD(a, b);
    r.hi = s;
    r.lo = EVP_X_SUB(b, EVP_X_SUB(s, a));
    return r;
}
/* p + e = a * b exactly */
EVP_EXP_FN evp_dd evp_two_prod(double a, double b)
{
    evp_dd r;
    r.hi = EVP_X_MUL(a, b);
    r.lo = EVP_X_FMA(a, b, -r.hi);
    return r;
}
EVP_EXP_FN evp_dd evp_dd_add(evp_dd x, evp_dd y)
{
    evp_dd s = evp_two_sum(x.hi, y.hi);
    const evp_dd t = evp_two_sum(x.lo, y.lo);
    s.lo = EVP_X_ADD(s.lo, t.hi);
    s = evp_quick_two_sum(s.hi, s.lo);
    s.lo = EVP_X_ADD(s.lo, t.lo);
    return evp_quick_two_sum(s.hi, s.lo);
}
EVP_EXP_FN evp_dd evp_dd_mul(evp_dd x, evp_dd y)
{
    evp_dd p = evp_two_prod(x.hi, y.hi);
    p.lo = EVP_X_ADD(p.lo, EVP_X_ADD(EVP_X_MUL(x.hi, y.lo), EVP_X_MUL(x.lo, y.hi)));
    return evp_quick_two_sum(p.hi, p.lo);
}

EVP_EXP_FN double evp_exp(double x)
{
    /* 1/i!, i = 3 .. 13, as double-doubles (generated with exact rational arithmetic) */
    const double fh[11] = {0x1.5555555555555p-3, 0x1.5555555555555p-5, 0x1.1111111111111p-7, 0x1.6c16c16c16c17p-10,
                           0x1.a01a01a01a01ap-13, 0x1.a01a01a01a01ap-16, 0x1.71de3a556c734p-19, 0x1.27e4fb7789f5cp-22,
                           0x1.ae64567f544e4p-26, 0x1.1eed8eff8d898p-29, 0x1.6124613a86d09p-33};
    const double fl[11] = {0x1.5555555555555p-57, 0x1.5555555555555p-59, 0x1.1111111111111p-63, -0x1.f49f49f49f49fp-65,
                           0x1.a01a01a01a01ap-73, 0x1.a01a01a01a01ap-76, -0x1.c154f8ddc6c00p-73, 0x1.cbbc05b4fa99ap-76,
                           -0x1.c062e06d1f209p-80, -0x1.2aec959e14c06p-83, 0x1.f28e0cc748ebep-87};
    const double invln2 = 0x1.71547652b82fep+0;
    const double L1 = 0x1.62e42fefa3800p-1, L2 = 0x1.ef35793c76000p-45, L3 = 0x1.cc01f97b57a08p-87;
    evp_dd r, p, q;
    double kd, res;
    int k, i;

    if (x != x) return x;
    if (x > 709.0) return x > 709.78 ? HUGE_VAL : exp(x);         /* outside the domain of the claim */
    if (x < -700.0) return x < -745.2 ? 0.0 : exp(x);
    kd = rint(EVP_X_MUL(x, invln2));
    k = (int)kd;
    /* r = x - k ln 2: k L1 and k L2 are exact products (|k| < 2^11, 42-bit constants) */
    r = evp_two_sum(EVP_X_SUB(x, EVP_X_MUL(kd, L1)), -EVP_X_MUL(kd, L2));
    r.lo = EVP_X_SUB(r.lo, EVP_X_MUL(kd, L3));
    r = evp_quick_two_sum(r.hi, r.lo);
    r.hi = EVP_X_MUL(r.hi, 0.015625);                             /* r / 64: exact */
    r.lo = EVP_X_MUL(r.lo, 0.015625);
    /* p = expm1(r) = r + r^2 (1/2 + r (1/3! + ... + r/13!)) */
    q.hi = fh[10]; q.lo = fl[10];
    for (i = 9; i >= 0; --i) {
        evp_dd c;
        c.hi = fh[i]; c.lo = fl[i];
        q = evp_dd_add(evp_dd_mul(q, r), c);
    }
    {
        evp_dd half;
        half.hi = 0.5; half.lo = 0.0;
        q = evp_dd_add(evp_dd_mul(q, r), half);
    }
    p = evp_dd_add(evp_dd_mul(evp_dd_mul(r, r), q), r);
    for (i = 0; i < 6; ++i) {                                     /* expm1(2 t) = 2 p + p^2 */
        evp_dd p2 = evp_dd_mul(p, p);
        p.hi = EVP_X_MUL(p.hi, 2.0);
        p.lo = EVP_X_MUL(p.lo, 2.0);
        p = evp_dd_add(p, p2);
    }
    {
        evp_dd one;
        one.hi = 1.0; one.lo = 0.0;
        p = evp_dd_add(one, p);
    }
    res = EVP_X_ADD(p.hi, p.lo);                                  /* the one rounding */
    /* 2^k in two exact steps (k in [-1010, 1024]: each factor is a normal number) */
    {
        const int k1 = k / 2, k2 = k - k1;
        res = EVP_X_MUL(res, EVP_X_BITS((unsigned long long)(k1 + 1023) << 52));
        res = EVP_X_MUL(res, EVP_X_BITS((unsigned long long)(k2 + 1023) << 52));
    }
    return res;
}

#endif /* EVP_EXP_H */
