/* evp_trig.h -- sin, cos and atan2 for the EAP rheology: ONE definition, plain IEEE-754 double operations with
 * explicit round-to-nearest (no contraction whatever the compiler flags), compiled into the CUDA kernels
 * (evp_eap.cuh) and into the CPU oracle (oracle/eap_oracle.inc), so that both sides take bit-identical values.
 *
 * Why: update_stress_rdg and calc_ffrac (source/ice_dyn_eap.F90:1474-1658, 1795-1864) go through the Fortran
 * intrinsics atan2, cos, sin and then TRUNCATE angles to table indices (`kx = int((x-piq-pi)*invdx) + 1`, :1583-1585):
 * a last-bit difference between two maths libraries can move a cell to another entry of the s11r ... s22s tables, a
 * finite jump in the stress.  Every libm returns its own last bit (glibc, Intel's, CUDA's all differ), so "the
 * reference's value" is defined only up to the intrinsic; the device and the oracle here share these three routines,
 * and the reference text is executed with the same three bound as its intrinsics (tests/golden/make_eap_vectors.py).
 *
 * Accuracy (tests/test_oracle_cpu.py, against 400-bit arithmetic): sin, cos < 1 ulp for |x| <= 1e5; atan2 < 1.5 ulp.
 * That is the quality of a production libm (CUDA documents 2 ulp for atan2, glibc 1 ulp for all three).
 *
 * Method.  sin / cos: k = nint(x 2/pi); r = x - k pi/2 with pi/2 split into 33 + 33 + 53 bits (k P1, k P2 exact for
 * |k| < 2^20), kept as head + tail; Taylor polynomials to r^19 / r^20 on |r| <= pi/4 (truncation < 2^-63), evaluated in
 * the fdlibm arrangement (the tail enters to first order).  atan2: t = min(|y|,|x|) / max(|y|,|x|) in [0, 1]; for
 * t > sqrt(2) - 1 the argument (t - 1)/(t + 1) = (a - b)/(a + b) and the offset pi/4, the quotient's rounding error
 * carried to first order (one fma; two divisions per call in all); odd Taylor polynomial to u^43 on |u| <= sqrt(2) - 1
 * (truncation < 2^-58 relative); pi/2 - r and pi - r with two-part constants; signs and zeros as C's atan2
 * (atan2(+-0, x >= +0) = +-0, atan2(+-0, x <= -0) = +-pi, atan2(y, +-0) = +-pi/2).
 */
#ifndef EVP_TRIG_H
#define EVP_TRIG_H

#include "evp_exp.h" /* EVP_EXP_FN, EVP_X_ADD / SUB / MUL */

/* Polynomial coefficients: file-scope tables.  On the device they live in __constant__ memory, so that DADD / DMUL take
 * them as c[bank][offset] operands; as local constant arrays every use cost two extra instructions that build the 64-bit
 * immediate (a third of eap_yield_kernel's issue slots, profiles/r02_eap.md).  Same values, same operations on every side. */
#if defined(__CUDACC__) && !defined(EVP_EMU)
#define EVP_TRIG_TABLE static __constant__ double
#else
#define EVP_TRIG_TABLE static const double
#endif
/* sin: (-1)^k / (2k+1)!, k = 1 .. 9;  cos: (-1)^k / (2k)!, k = 2 .. 10;  atan: (-1)^k / (2k+1), k = 1 .. 21 */
EVP_TRIG_TABLE evp_trig_S[9] = {-0x1.5555555555555p-3, 0x1.1111111111111p-7, -0x1.a01a01a01a01ap-13, 0x1.71de3a556c734p-19,
                                -0x1.ae64567f544e4p-26, 0x1.6124613a86d09p-33, -0x1.ae7f3e733b81fp-41, 0x1.952c77030ad4ap-49,
                                -0x1.2f49b46814157p-57};
EVP_TRIG_TABLE evp_trig_C[9] = {0x1.5555555555555p-5, -0x1.6c16c16c16c17p-10, 0x1.a01a01a01a01ap-16, -0x1.27e4fb7789f5cp-22,
                                0x1.1eed8eff8d898p-29, -0x1.93974a8c07c9dp-37, 0x1.ae7f3e733b81fp-45, -0x1.6827863b97d97p-53,
                                0x1.e542ba4020225p-62};
EVP_TRIG_TABLE evp_trig_A[21] = {-0x1.5555555555555p-2, 0x1.999999999999ap-3, -0x1.2492492492492p-3, 0x1.c71c71c71c71cp-4,
                                 -0x1.745d1745d1746p-4, 0x1.3b13b13b13b14p-4, -0x1.1111111111111p-4, 0x1.e1e1e1e1e1e1ep-5,
                                 -0x1.af286bca1af28p-5, 0x1.8618618618618p-5, -0x1.642c8590b2164p-5, 0x1.47ae147ae147bp-5,
                                 -0x1.2f684bda12f68p-5, 0x1.1a7b9611a7b96p-5, -0x1.0842108421084p-5, 0x1.f07c1f07c1f08p-6,
                                 -0x1.d41d41d41d41dp-6, 0x1.bacf914c1bad0p-6, -0x1.a41a41a41a41ap-6, 0x1.8f9c18f9c18fap-6,
                                 -0x1.7d05f417d05f4p-6};

#define EVP_X_DIV(a, b) evp_x_div_((a), (b))
#if defined(__CUDACC__) || defined(EVP_EMU)
EVP_EXP_FN double evp_x_div_(double a, double b) { return __ddiv_rn(a, b); }
#else
EVP_EXP_FN double evp_x_div_(double a, double b) { return a / b; }
#endif

/* x = k pi/2 + (rh + rl), |rh + rl| <= pi/4 (+ rounding); returns k mod 4 */
EVP_EXP_FN int evp_rem_pio2(double x, double *rh, double *rl)
{
    const double invpio2 = 0x1.45f306dc9c883p-1;
    const double P1 = 0x1.921fb54400000p+0, P2 = 0x1.0b4611a600000p-34, P3 = 0x1.3198a2e037073p-69;
    const double kd = rint(EVP_X_MUL(x, invpio2));
    const double t = EVP_X_SUB(x, EVP_X_MUL(kd, P1));
    const double w = EVP_X_MUL(kd, P2);
    const evp_dd s = evp_two_sum(t, -w);
    const double e = EVP_X_SUB(s.lo, EVP_X_MUL(kd, P3));
    const double h = EVP_X_ADD(s.hi, e);
    *rh = h;
    *rl = EVP_X_ADD(EVP_X_SUB(s.hi, h), e);
    return (int)((long long)kd & 3LL);
}

/* sin(r + rl), |r| <= pi/4 */
EVP_EXP_FN double evp_sin_k(double r, double rl)
{
    const double z = EVP_X_MUL(r, r);
    double p = evp_trig_S[8];
    int i;
    for (i = 7; i >= 0; --i) p = EVP_X_ADD(EVP_X_MUL(p, z), evp_trig_S[i]);
    /* r + (r z p + rl (1 - z/2)) */
    return EVP_X_ADD(r, EVP_X_ADD(EVP_X_MUL(EVP_X_MUL(r, z), p),
                                  EVP_X_MUL(rl, EVP_X_SUB(1.0, EVP_X_MUL(0.5, z)))));
}

/* cos(r + rl), |r| <= pi/4 */
EVP_EXP_FN double evp_cos_k(double r, double rl)
{
    const double z = EVP_X_MUL(r, r);
    double p = evp_trig_C[8];
    double hz, w;
    int i;
    for (i = 7; i >= 0; --i) p = EVP_X_ADD(EVP_X_MUL(p, z), evp_trig_C[i]);
    hz = EVP_X_MUL(0.5, z);
    w = EVP_X_SUB(1.0, hz);
    /* w + (((1 - w) - hz) + (z z p - r rl)) */
    return EVP_X_ADD(w, EVP_X_ADD(EVP_X_SUB(EVP_X_SUB(1.0, w), hz),
                                  EVP_X_SUB(EVP_X_MUL(EVP_X_MUL(z, z), p), EVP_X_MUL(r, rl))));
}

/* both at once (every caller on the path needs the pair of one angle) */
EVP_EXP_FN void evp_sincos(double x, double *s, double *c)
{
    double r, rl, sk, ck;
    int n;
    if (!(fabs(x) <= 1.0e5)) {         /* outside the claim (NaN, infinities, huge arguments): NaN, on every side alike */
        *s = *c = EVP_X_DIV(EVP_X_SUB(x, x), EVP_X_SUB(x, x));
        return;
    }
    if (x == 0.0) {                    /* sin(+-0) = +-0 */
        *s = x;
        *c = 1.0;
        return;
    }
    n = evp_rem_pio2(x, &r, &rl);
    sk = evp_sin_k(r, rl);
    ck = evp_cos_k(r, rl);
    *s = (n == 0) ? sk : (n == 1) ? ck : (n == 2) ? -sk : -ck;
    *c = (n == 0) ? ck : (n == 1) ? -sk : (n == 2) ? -ck : sk;
}
EVP_EXP_FN double evp_sin(double x) { double s, c; evp_sincos(x, &s, &c); return s; }
EVP_EXP_FN double evp_cos(double x) { double s, c; evp_sincos(x, &s, &c); return c; }

EVP_EXP_FN double evp_atan2(double y, double x)
{
    const double PI_H = 0x1.921fb54442d18p+1, PI_L = 0x1.1a62633145c07p-53;
    const double PIO2_H = 0x1.921fb54442d18p+0, PIO2_L = 0x1.1a62633145c07p-54;
    const double PIO4_H = 0x1.921fb54442d18p-1, PIO4_L = 0x1.1a62633145c07p-55;
    const double T0 = 0x1.a827999fcef32p-2;   /* sqrt(2) - 1 */
    double a, b, u, ul, z, p, r;
    int swap, shifted, i;

    if (x != x || y != y) return EVP_X_ADD(x, y);
    if (y == 0.0) return (x > 0.0 || (x == 0.0 && !signbit(x))) ? y : copysign(PI_H, y);
    if (x == 0.0) return copysign(PIO2_H, y);
    a = fabs(y);
    b = fabs(x);
    if (isinf(a) || isinf(b)) {               /* never on the path; the C conventions */
        if (isinf(a) && isinf(b)) r = (x > 0.0) ? PIO4_H : EVP_X_MUL(3.0, PIO4_H);
        else if (isinf(a)) r = PIO2_H;
        else r = (x > 0.0) ? 0.0 : PI_H;
        return copysign(r, y);
    }
    swap = a > b;
    if (swap) { const double q = a; a = b; b = q; }                          /* a <= b: a / b in [0, 1] */
    shifted = a > EVP_X_MUL(T0, b);                                           /* a / b > sqrt(2) - 1, without dividing */
    if (shifted) {            /* (t - 1)/(t + 1) = (a - b)/(a + b), numerator and denominator as head + tail */
        const evp_dd n = evp_two_sum(a, -b), d = evp_two_sum(a, b);
        u = EVP_X_DIV(n.hi, d.hi);
        ul = EVP_X_DIV(EVP_X_ADD(EVP_X_FMA(-u, d.hi, n.hi), EVP_X_SUB(n.lo, EVP_X_MUL(u, d.lo))), d.hi);
    } else {
        u = EVP_X_DIV(a, b);
        ul = EVP_X_DIV(EVP_X_FMA(-u, b, a), b);                               /* a / b = u + ul */
    }
    z = EVP_X_MUL(u, u);
    p = evp_trig_A[20];
    for (i = 19; i >= 0; --i) p = EVP_X_ADD(EVP_X_MUL(p, z), evp_trig_A[i]);
    /* atan(u + ul) = u + (u z p + ul / (1 + z)); 1 / (1 + z) = 1 - z up to z^2 <= 0.03 of a term below 2^-53 u */
    p = EVP_X_ADD(u, EVP_X_ADD(EVP_X_MUL(EVP_X_MUL(u, z), p), EVP_X_MUL(ul, EVP_X_SUB(1.0, z))));
    r = shifted ? EVP_X_ADD(PIO4_H, EVP_X_ADD(p, PIO4_L)) : p;
    if (swap) r = EVP_X_SUB(PIO2_H, EVP_X_SUB(r, PIO2_L));
    if (x < 0.0) r = EVP_X_SUB(PI_H, EVP_X_SUB(r, PI_L));
    return copysign(r, y);
}

#endif /* EVP_TRIG_H */
