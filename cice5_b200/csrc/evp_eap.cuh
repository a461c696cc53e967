// evp_eap.cuh -- sm_100a FP64 kernels of the elastic-anisotropic-plastic dynamics (kdyn = 2).
//
// Reference semantics (path:line in the reference checkout):
//   stress_eap         source/ice_dyn_eap.F90:1052-1467     update_stress_rdg  :1474-1658
//   stepa              source/ice_dyn_eap.F90:1664-1787     calc_ffrac         :1795-1864
//   eap driver         source/ice_dyn_eap.F90:66-486 (the structure-tensor reset :274-289, stepa every 10th subcycle :408)
//
// EAP shares evp_prep1 / evp_prep2 / stepu / evp_finish and the halo pattern with EVP; what is new is the stress update
// (a table look-up per corner instead of the elliptical yield curve), the structure tensor (a11, a12 at the four corners,
// persistent state) and its evolution.  A subcycle runs in two stages:
//   1. eap_yield_kernel: one thread per (T cell, corner) evaluates update_stress_rdg -- two atan2, one sin / cos pair
//      (cice5_b200/csrc/evp_trig.h: the one definition shared with the CPU oracle), 6 gathers from 1.7 MB of L2-resident
//      tables -- and stores the three yield stresses of its corner.  FP64-pipe bound; 54 registers keep 36 warps per SM in
//      flight.  (Measured first with this arithmetic inside the sweep: 140 registers, 2.7 warps per scheduler, the FP64
//      pipe 38 % active; profiles/r02_eap.md.)
//   2. eap_subcycle_kernel: the register-sweep kernel of evp_kernels.cuh with the rest of stress_eap as its cell function
//      (relaxation towards the yield stresses, stress combinations) fused with stepu: `str` never reaches memory.
// The part of update_stress_rdg that depends on the structure tensor alone (gamma, Q11, Q12, ka) is evaluated where the
// tensor changes -- eap_stepa_kernel on the subcycles the reference calls stepa, eap_reset_kernel at the start of a call --
// and stored.  The tensor is read-only inside the subcycle kernels.
#pragma once
#include "evp_kernels.cuh"
#include "evp_trig.h"

namespace evp {

constexpr int kEapNx = 41, kEapNy = 41, kEapNa = 21;                      // nx_yield, ny_yield, na_yield (:31-34)
constexpr int kEapTableElems = kEapNx * kEapNy * kEapNa;

struct EapArgs {
    const double *tab[6];              // s11r s12r s22r s11s s12s s22s, Fortran (nx_yield, ny_yield, na_yield)
    double *a11c[4], *a12c[4];         // structure tensor at the corners ne nw sw se (planes of the arena's pitch)
    double *a11, *a12;                 // cell averages (stepa)
    double *out[9];                    // e11 e12 e22 s11 s12 s22 yieldstress11 yieldstress12 yieldstress22
    double *rot[16];                   // per corner q: rot[4q] gamma, [4q+1] Q11, [4q+2] Q12, [4q+3] ka (as a double)
    double *yld[12];                   // per corner q: yld[3q] stressptmp, [3q+1] stressmtmp, [3q+2] stress12tmp
    double *alphar[4];                 // per corner: ridging function of the last subcycle
    double dtei;                       // 1 / dte (stepa)
};

// First part of update_stress_rdg (:1527-1548 and ka of :1601): what depends on the structure tensor alone -- the angle
// gamma of its principal axis, the rotation (Q11, Q12) and the table index ka.  a11, a12 change only in stepa (every
// 10th subcycle) and in the reset at the start of a call, so these four values are computed there and stored; the
// statements and their operands are the reference's, only the moment of evaluation differs.
__device__ __forceinline__ void eap_rotation(double a11, double a12, double &gamma, double &Q11, double &Q12, double &kaf)
{
    const double a22 = 1.0 - a11;
    gamma = 0.5 * evp_atan2((2.0 * a12), (a11 - a22));
    evp_sincos(gamma, &Q12, &Q11);
    const double Q11Q11 = Q11 * Q11, Q11Q12 = Q11 * Q12, Q12Q12 = Q12 * Q12;
    double atempprime = Q11Q11 * a11 + 2.0 * Q11Q12 * a12 + Q12Q12 * a22;
    atempprime = fmax(atempprime, 1.0 - atempprime);
    const double da = 0.5 / (double)(kEapNa - 1), invda = 1.0 / da;
    // (a tensor with an eigenvalue outside [0, 1] -- possible only outside the model's regime, e.g. dte of the order of
    // 1 / kfrac -- would index past the tables, which the reference does not check: the nearest entry is taken instead)
    kaf = (double)min(max((int)((atempprime - 0.5) * invda) + 1, 1), kEapNa);
}

struct EapCorner {
    double sp, sm, s12;                // the yield stresses stressptmp, stressmtmp, stress12tmp
    double alphar;                     // ridging function (last subcycle only)
};

// The rest of update_stress_rdg (:1550-1656), every statement the reference's; alphas is not computed (its only use,
// rdg_shear, is commented out at :1283-1285).  invsin = c1/sin(pi2/c12) * invstressconviso (:1523-1525).
template <bool LAST>
__device__ __forceinline__ EapCorner eap_yield(const EapArgs &e, double invsin, double divu, double tension, double shear,
                                               double gamma, double Q11, double Q12, int ka, double strength)
{
    const double pi = 3.14159265358979323846, pi2 = 2.0 * pi, piq = 0.5 * (0.5 * pi);
    const double kfriction = 0.45, puny = 1.0e-11;
    EapCorner o;
    double Qd11, Qd12;
    const double Q11Q11 = Q11 * Q11, Q11Q12 = Q11 * Q12, Q12Q12 = Q12 * Q12;

    const double dtemp11 = 0.5 * (divu + tension), dtemp12 = shear * 0.5, dtemp22 = 0.5 * (divu - tension);
    double alpha = 0.5 * evp_atan2((2.0 * dtemp12), (dtemp11 - dtemp22));
    if (alpha > gamma) alpha = alpha - pi;
    if (alpha < gamma - pi) alpha = alpha + pi;
    const double y = gamma - alpha;

    evp_sincos(alpha, &Qd12, &Qd11);
    double dtemp1 = Qd11 * (Qd11 * dtemp11 + 2.0 * Qd12 * dtemp12) + Qd12 * Qd12 * dtemp22;
    double dtemp2 = Qd12 * (Qd12 * dtemp11 - 2.0 * Qd11 * dtemp12) + Qd11 * Qd11 * dtemp22;
    double x = 0.0;
    if ((fabs(dtemp1) > puny) || (fabs(dtemp2) > puny)) {
        const double invleng = 1.0 / sqrt(dtemp1 * dtemp1 + dtemp2 * dtemp2);
        dtemp1 = dtemp1 * invleng;
        dtemp2 = dtemp2 * invleng;
        x = evp_atan2(dtemp2, dtemp1);
    }
    if (x < piq) x = x + pi2;

    const double dx = pi / (double)(kEapNx - 1), dy = pi / (double)(kEapNy - 1);
    const double invdx = 1.0 / dx, invdy = 1.0 / dy;
    const int kx = min(max((int)((x - piq - pi) * invdx) + 1, 1), kEapNx);   // clamped: see eap_rotation
    const int ky = min(max((int)(y * invdy) + 1, 1), kEapNy);
    const int q = ((ka - 1) * kEapNy + (ky - 1)) * kEapNx + (kx - 1);
    const double stemp11r = e.tab[0][q], stemp12r = e.tab[1][q], stemp22r = e.tab[2][q];
    const double stemp11s = e.tab[3][q], stemp12s = e.tab[4][q], stemp22s = e.tab[5][q];

    const double stressp = strength * (stemp11r + kfriction * stemp11s + stemp22r + kfriction * stemp22s) * invsin;
    const double stress12 = strength * (stemp12r + kfriction * stemp12s) * invsin;
    const double stressm = strength * (stemp11r + kfriction * stemp11s - stemp22r - kfriction * stemp22s) * invsin;

    const double sig11 = 0.5 * (stressp + stressm), sig12 = stress12, sig22 = 0.5 * (stressp - stressm);
    const double sgprm11 = Q11Q11 * sig11 + Q12Q12 * sig22 - 2.0 * Q11Q12 * sig12;
    const double sgprm12 = Q11Q12 * sig11 - Q11Q12 * sig22 + (Q11Q11 - Q12Q12) * sig12;
    const double sgprm22 = Q12Q12 * sig11 + Q11Q11 * sig22 + 2.0 * Q11Q12 * sig12;
    o.sp = sgprm11 + sgprm22;
    o.s12 = sgprm12;
    o.sm = sgprm11 - sgprm22;
    o.alphar = 0.0;
    if (LAST) {
        const double rotstemp11r = Q11Q11 * stemp11r - 2.0 * Q11Q12 * stemp12r + Q12Q12 * stemp22r;
        const double rotstemp12r = Q11Q11 * stemp12r + Q11Q12 * (stemp11r - stemp22r) - Q12Q12 * stemp12r;
        const double rotstemp22r = Q12Q12 * stemp11r + 2.0 * Q11Q12 * stemp12r + Q11Q11 * stemp22r;
        o.alphar = rotstemp11r * dtemp11 + 2.0 * rotstemp12r * dtemp12 + rotstemp22r * dtemp22;
    }
    return o;
}

// the stress combinations for the momentum equation (:1357-1463; the same statements as ice_dyn_evp.F90:752-845)
__device__ __forceinline__ void stress_combine(const Consts &k, const double (&s)[12], double dxt, double dyt, double dxhy,
                                               double dyhx, double (&str)[8])
{
    const double sp1 = s[0], sp2 = s[1], sp3 = s[2], sp4 = s[3];
    const double sm1 = s[4], sm2 = s[5], sm3 = s[6], sm4 = s[7];
    const double s121 = s[8], s122 = s[9], s123 = s[10], s124 = s[11];
    const double ssigpn = sp1 + sp2, ssigps = sp3 + sp4, ssigpe = sp1 + sp4, ssigpw = sp2 + sp3;
    const double ssigp1 = (sp1 + sp3) * k.p055, ssigp2 = (sp2 + sp4) * k.p055;
    const double ssigmn = sm1 + sm2, ssigms = sm3 + sm4, ssigme = sm1 + sm4, ssigmw = sm2 + sm3;
    const double ssigm1 = (sm1 + sm3) * k.p055, ssigm2 = (sm2 + sm4) * k.p055;
    const double ssig12n = s121 + s122, ssig12s = s123 + s124;
    const double ssig12e = s121 + s124, ssig12w = s122 + s123;
    const double ssig121 = (s121 + s123) * k.p111, ssig122 = (s122 + s124) * k.p111;

    const double csigpne = k.p111 * sp1 + ssigp2 + k.p027 * sp3;
    const double csigpnw = k.p111 * sp2 + ssigp1 + k.p027 * sp4;
    const double csigpsw = k.p111 * sp3 + ssigp2 + k.p027 * sp1;
    const double csigpse = k.p111 * sp4 + ssigp1 + k.p027 * sp2;
    const double csigmne = k.p111 * sm1 + ssigm2 + k.p027 * sm3;
    const double csigmnw = k.p111 * sm2 + ssigm1 + k.p027 * sm4;
    const double csigmsw = k.p111 * sm3 + ssigm2 + k.p027 * sm1;
    const double csigmse = k.p111 * sm4 + ssigm1 + k.p027 * sm2;
    const double csig12ne = k.p222 * s121 + ssig122 + k.p055 * s123;
    const double csig12nw = k.p222 * s122 + ssig121 + k.p055 * s124;
    const double csig12sw = k.p222 * s123 + ssig122 + k.p055 * s121;
    const double csig12se = k.p222 * s124 + ssig121 + k.p055 * s122;

    const double str12ew = 0.5 * dxt * (k.p333 * ssig12e + k.p166 * ssig12w);
    const double str12we = 0.5 * dxt * (k.p333 * ssig12w + k.p166 * ssig12e);
    const double str12ns = 0.5 * dyt * (k.p333 * ssig12n + k.p166 * ssig12s);
    const double str12sn = 0.5 * dyt * (k.p333 * ssig12s + k.p166 * ssig12n);

    double strp_tmp = 0.25 * dyt * (k.p333 * ssigpn + k.p166 * ssigps);
    double strm_tmp = 0.25 * dyt * (k.p333 * ssigmn + k.p166 * ssigms);
    str[0] = -strp_tmp - strm_tmp - str12ew + dxhy * (-csigpne + csigmne) + dyhx * csig12ne;
    str[1] = strp_tmp + strm_tmp - str12we + dxhy * (-csigpnw + csigmnw) + dyhx * csig12nw;
    strp_tmp = 0.25 * dyt * (k.p333 * ssigps + k.p166 * ssigpn);
    strm_tmp = 0.25 * dyt * (k.p333 * ssigms + k.p166 * ssigmn);
    str[2] = -strp_tmp - strm_tmp + str12ew + dxhy * (-csigpse + csigmse) + dyhx * csig12se;
    str[3] = strp_tmp + strm_tmp + str12we + dxhy * (-csigpsw + csigmsw) + dyhx * csig12sw;
    strp_tmp = 0.25 * dxt * (k.p333 * ssigpe + k.p166 * ssigpw);
    strm_tmp = 0.25 * dxt * (k.p333 * ssigme + k.p166 * ssigmw);
    str[4] = -strp_tmp + strm_tmp - str12ns - dyhx * (csigpne + csigmne) + dxhy * csig12ne;
    str[5] = strp_tmp - strm_tmp - str12sn - dyhx * (csigpse + csigmse) + dxhy * csig12se;
    strp_tmp = 0.25 * dxt * (k.p333 * ssigpw + k.p166 * ssigpe);
    strm_tmp = 0.25 * dxt * (k.p333 * ssigmw + k.p166 * ssigme);
    str[6] = -strp_tmp + strm_tmp + str12ns - dyhx * (csigpnw + csigmnw) + dxhy * csig12nw;
    str[7] = strp_tmp - strm_tmp + str12sn - dyhx * (csigpsw + csigmsw) + dxhy * csig12sw;
}

// ---- stage 1 of a subcycle: the yield stresses, one thread per (T cell, corner) ---------------------------------------
// stress_eap's strain rates of ONE corner (:1194-1232) and update_stress_rdg for it (:1239-1262).  A warp works on one
// corner of 32 consecutive cells of a row (the corner is warp-uniform: no divergence in the choice of coefficients; loads
// and stores are 256-byte rows), a CTA of 4 warps on the four corners of the same 32 cells (the metric terms and
// velocities the corners share come out of L1).  About 70 registers per thread: 6 and more CTAs per SM, where the same
// arithmetic inside the row sweep (140 registers) kept 2.7 warps per scheduler in flight (profiles/r02_eap.md).
template <bool LAST>
__global__ void __launch_bounds__(128)
eap_yield_kernel(const __grid_constant__ LoopArgs a, const __grid_constant__ EapArgs e)
{
    const int q = threadIdx.x >> 5;                        // corner: 0 ne, 1 nw, 2 sw, 3 se
    const int c = 1 + blockIdx.x * 32 + (threadIdx.x & 31), r = 1 + blockIdx.y;
    if (c > a.nxl + 1) return;
    if (!(a.mask[(size_t)r * a.ldm + c] & 1u)) return;
    const size_t idx = (size_t)r * a.ld + c;
    const double u_ij = a.u0[idx], u_wj = a.u0[idx - 1], u_is = a.u0[idx - a.ld], u_ws = a.u0[idx - a.ld - 1];
    const double v_ij = a.v0[idx], v_wj = a.v0[idx - 1], v_is = a.v0[idx - a.ld], v_ws = a.v0[idx - a.ld - 1];
    const double dxt = a.dxt[idx], dyt = a.dyt[idx], cxp = a.cxp[idx], cyp = a.cyp[idx], cxm = a.cxm[idx], cym = a.cym[idx];
    double divu, tension, shear;
    if (q == 0) {
        divu = cyp * u_ij - dyt * u_wj + cxp * v_ij - dxt * v_is;
        tension = -cym * u_ij - dyt * u_wj + cxm * v_ij + dxt * v_is;
        shear = -cym * v_ij - dyt * v_wj - cxm * u_ij - dxt * u_is;
    } else if (q == 1) {
        divu = cym * u_wj + dyt * u_ij + cxp * v_wj - dxt * v_ws;
        tension = -cyp * u_wj + dyt * u_ij + cxm * v_wj + dxt * v_ws;
        shear = -cyp * v_wj + dyt * v_ij - cxm * u_wj - dxt * u_ws;
    } else if (q == 2) {
        divu = cym * u_ws + dyt * u_is + cxm * v_ws + dxt * v_wj;
        tension = -cyp * u_ws + dyt * u_is + cxp * v_ws - dxt * v_wj;
        shear = -cyp * v_ws + dyt * v_is - cxp * u_ws + dxt * u_wj;
    } else {
        divu = cyp * u_is - dyt * u_ws + cxm * v_is + dxt * v_ij;
        tension = -cym * u_is - dyt * u_ws + cxp * v_is - dxt * v_ij;
        shear = -cym * v_is - dyt * v_ws - cxp * u_is + dxt * u_ij;
    }
    // invsin = c1/sin(pi2/c12) * invstressconviso (:1523-1525)
    const double invsin = 1.0 / evp_sin((2.0 * 3.14159265358979323846) / 12.0) * (1.0 / (1.0 + 0.45 * 0.45));
    const EapCorner o = eap_yield<LAST>(e, invsin, divu, tension, shear, e.rot[4 * q][idx], e.rot[4 * q + 1][idx],
                                        e.rot[4 * q + 2][idx], (int)e.rot[4 * q + 3][idx], a.strength[idx]);
    e.yld[3 * q][idx] = o.sp;
    e.yld[3 * q + 1][idx] = o.sm;
    e.yld[3 * q + 2][idx] = o.s12;
    if (LAST) e.alphar[q][idx] = o.alphar;
}

// ---- stage 2: relaxation + combinations + stepu --------------------------------------------------------------------
// One T cell of the rest of stress_eap (:1264-1463): s[12] in/out, the yield stresses from stage 1; str[8] out; on the
// last subcycle the diagnostics (which need the strain rates again: 12 cheap expressions).
struct EapDiag {
    double divu, shear, rdg_conv, e11, e12, e22, s11, s12, s22, ys11, ys12, ys22;
};
template <bool LAST>
__device__ __forceinline__ void stress_eap_cell(
    const Consts &k, const EapArgs &e, const LoopArgs &a, size_t idx, double u_ij, double u_wj, double u_is, double u_ws,
    double v_ij, double v_wj, double v_is, double v_ws, double (&s)[12], double (&str)[8], EapDiag &dg)
{
    double y[12];                                          // ne.sp ne.sm ne.s12 nw.sp ...
#pragma unroll
    for (int t = 0; t < 12; ++t) y[t] = e.yld[t][idx];
    const double dxt = a.dxt[idx], dyt = a.dyt[idx];

    // elastic relaxation (:1303-1331)
    s[0] = (s[0] + y[0] * k.arlx1i) * k.denom1;
    s[1] = (s[1] + y[3] * k.arlx1i) * k.denom1;
    s[2] = (s[2] + y[6] * k.arlx1i) * k.denom1;
    s[3] = (s[3] + y[9] * k.arlx1i) * k.denom1;
    s[4] = (s[4] + y[1] * k.arlx1i) * k.denom1;
    s[5] = (s[5] + y[4] * k.arlx1i) * k.denom1;
    s[6] = (s[6] + y[7] * k.arlx1i) * k.denom1;
    s[7] = (s[7] + y[10] * k.arlx1i) * k.denom1;
    s[8] = (s[8] + y[2] * k.arlx1i) * k.denom1;
    s[9] = (s[9] + y[5] * k.arlx1i) * k.denom1;
    s[10] = (s[10] + y[8] * k.arlx1i) * k.denom1;
    s[11] = (s[11] + y[11] * k.arlx1i) * k.denom1;

    if (LAST) {
        const double cxp = a.cxp[idx], cyp = a.cyp[idx], cxm = a.cxm[idx], cym = a.cym[idx], tarear = a.tarear[idx];
        // strain rates * area (:1194-1232)
        const double divune = cyp * u_ij - dyt * u_wj + cxp * v_ij - dxt * v_is;
        const double divunw = cym * u_wj + dyt * u_ij + cxp * v_wj - dxt * v_ws;
        const double divusw = cym * u_ws + dyt * u_is + cxm * v_ws + dxt * v_wj;
        const double divuse = cyp * u_is - dyt * u_ws + cxm * v_is + dxt * v_ij;
        const double tensionne = -cym * u_ij - dyt * u_wj + cxm * v_ij + dxt * v_is;
        const double tensionnw = -cyp * u_wj + dyt * u_ij + cxm * v_wj + dxt * v_ws;
        const double tensionsw = -cyp * u_ws + dyt * u_is + cxp * v_ws - dxt * v_wj;
        const double tensionse = -cym * u_is - dyt * u_ws + cxp * v_is - dxt * v_ij;
        const double shearne = -cym * v_ij - dyt * v_wj - cxm * u_ij - dxt * u_is;
        const double shearnw = -cyp * v_wj + dyt * v_ij - cxm * u_wj - dxt * u_ws;
        const double shearsw = -cyp * v_ws + dyt * v_is - cxp * u_ws + dxt * u_wj;
        const double shearse = -cym * v_is - dyt * v_ws - cxp * u_is + dxt * u_ij;
        // :1268-1286
        const double ts = tensionne + tensionnw + tensionse + tensionsw;
        const double ss = shearne + shearnw + shearse + shearsw;
        dg.shear = 0.25 * tarear * sqrt(ts * ts + ss * ss);
        dg.divu = 0.25 * (divune + divunw + divuse + divusw) * tarear;
        dg.rdg_conv = -fmin(0.25 * (e.alphar[0][idx] + e.alphar[1][idx] + e.alphar[2][idx] + e.alphar[3][idx]), 0.0) * tarear;
        // :1288-1295 (written on every subcycle by the reference; the last one's values survive the call)
        dg.e11 = 0.5 * 0.25 * (divune + divunw + divuse + divusw + tensionne + tensionnw + tensionse + tensionsw) * tarear;
        dg.e12 = 0.5 * 0.25 * (shearne + shearnw + shearse + shearsw) * tarear;
        dg.e22 = 0.5 * 0.25 * (divune + divunw + divuse + divusw - tensionne - tensionnw - tensionse - tensionsw) * tarear;
        // :1333-1355
        dg.s11 = 0.5 * 0.25 * (s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7]);
        dg.s22 = 0.5 * 0.25 * (s[0] + s[1] + s[2] + s[3] - s[4] - s[5] - s[6] - s[7]);
        dg.s12 = 0.25 * (s[8] + s[9] + s[10] + s[11]);
        dg.ys11 = 0.5 * 0.25 * (y[0] + y[3] + y[6] + y[9] + y[1] + y[4] + y[7] + y[10]);
        dg.ys22 = 0.5 * 0.25 * (y[0] + y[3] + y[6] + y[9] - y[1] - y[4] - y[7] - y[10]);
        dg.ys12 = 0.25 * (y[2] + y[5] + y[8] + y[11]);
    }
    stress_combine(k, s, dxt, dyt, a.dxhy[idx], a.dyhx[idx], str);
}

// relaxation + combinations + stepu + E-W ghost stores + tripole inbox stores: the register sweep of evp_subcycle_kernel
// (evp_kernels.cuh) around stress_eap_cell.  Classic relaxation only (kdyn = 2 with revised_evp is rejected at init).
template <bool LAST>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
eap_subcycle_kernel(const __grid_constant__ LoopArgs a, const __grid_constant__ EapArgs e)
{
    const int lane = threadIdx.x & 31;
    const int strip = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int c0 = 1 + strip * kUColsPerWarp;
    if (c0 > a.nxl) return;                               // warp-uniform
    const int c = c0 + lane;
    const int r0 = 1 + blockIdx.y * a.rows_per_chunk;
    const int rend = min(r0 + a.rows_per_chunk, a.ny + 1);
    const bool tcol = c <= a.nxl + 1;
    const bool ucol = lane < kUColsPerWarp && c <= a.nxl;
    const bool tstore_col = tcol && (lane < kUColsPerWarp || c == a.nxl + 1);
    const Consts &k = a.k;
    const int ld = a.ld;

    size_t idx = (size_t)(r0 - 1) * ld + c;
    double u_s = tcol ? a.u0[idx] : 0.0, v_s = tcol ? a.v0[idx] : 0.0;
    double uw_s = shfl_up1(u_s), vw_s = shfl_up1(v_s);
    if (lane == 0) { uw_s = a.u0[idx - 1]; vw_s = a.v0[idx - 1]; }
    double s1p = 0.0, s5p = 0.0, n2p = 0.0, n7p = 0.0;
    unsigned mprev = 0;

    for (int r = r0; r <= rend; ++r) {
        idx = (size_t)r * ld + c;
        const double u_c = tcol ? a.u0[idx] : 0.0, v_c = tcol ? a.v0[idx] : 0.0;
        double uw_c = shfl_up1(u_c), vw_c = shfl_up1(v_c);
        if (lane == 0) { uw_c = a.u0[idx - 1]; vw_c = a.v0[idx - 1]; }
        const unsigned m = tcol ? a.mask[(size_t)r * a.ldm + c] : 0u;
        const bool tact = (m & 1u) != 0;

        double str[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (tact) {
            double s[12];
#pragma unroll
            for (int q = 0; q < 12; ++q) s[q] = a.s0[q][idx];
            EapDiag dg;
            stress_eap_cell<LAST>(k, e, a, idx, u_c, uw_c, u_s, uw_s, v_c, vw_c, v_s, vw_s, s, str, dg);
            if (tstore_col && (r < rend || r == a.ny + 1)) {
#pragma unroll
                for (int q = 0; q < 12; ++q) a.s1[q][idx] = s[q];
                if (LAST) {
                    a.divu[idx] = dg.divu; a.shear[idx] = dg.shear; a.rdg_conv[idx] = dg.rdg_conv;
                    a.prs_sig[idx] = a.strength[idx];                                      // :1297
                    e.out[0][idx] = dg.e11; e.out[1][idx] = dg.e12; e.out[2][idx] = dg.e22;
                    e.out[3][idx] = dg.s11; e.out[4][idx] = dg.s12; e.out[5][idx] = dg.s22;
                    e.out[6][idx] = dg.ys11; e.out[7][idx] = dg.ys12; e.out[8][idx] = dg.ys22;
                }
            }
        }
        const double e2 = shfl_down1(str[1]), e4 = shfl_down1(str[3]);
        const double e7 = shfl_down1(str[6]), e8 = shfl_down1(str[7]);

        if (r > r0) {                                     // stepu on U row r-1 (ice_dyn_shared.F90:700-746)
            const size_t iu = idx - ld;
            const bool uact = ucol && (mprev & 2u);
            double unew = u_s, vnew = v_s;
            if (uact) {
                const double uold = u_s, vold = v_s;
                const double uo = a.uocn[iu], vo = a.vocn[iu];
                const double du = uo - uold, dv = vo - vold;
                const double vrel = a.aiu[iu] * k.rhow * a.Cw[iu] * sqrt(du * du + dv * dv);
                const double taux = vrel * a.waterx[iu], tauy = vrel * a.watery[iu];
                const double umdt = a.umassdti[iu], fm = a.fm[iu];
                const double cca = (k.brlx + k.revp) * umdt + vrel * k.cosw;
                const double ccb = fm + copysign(1.0, fm) * vrel * k.sinw;
                const double ab2 = cca * cca + ccb * ccb;
                const double uar = a.uarear[iu];
                const double strintx = uar * (s1p + n2p + str[2] + e4);
                const double strinty = uar * (s5p + str[5] + n7p + e8);
                const double cc1 = strintx + a.forcex[iu] + taux + umdt * (k.brlx * uold);
                const double cc2 = strinty + a.forcey[iu] + tauy + umdt * (k.brlx * vold);
                unew = (cca * cc1 + ccb * cc2) / ab2;
                vnew = (cca * cc2 - ccb * cc1) / ab2;
                if (LAST) { a.strintx[iu] = strintx; a.strinty[iu] = strinty; }
            }
            if (ucol) {
                a.u1[iu] = unew; a.v1[iu] = vnew;
                if (c == 1 && a.uW) {
                    const size_t ig = (size_t)(r - 1) * ld + a.nxl_west + 1;
                    a.uW[ig] = unew; a.vW[ig] = vnew;
                }
                if (c == a.nxl && a.uE) {
                    const size_t ig = (size_t)(r - 1) * ld;
                    a.uE[ig] = unew; a.vE[ig] = vnew;
                }
                if (a.tripole && r - 1 >= a.ny - 1) tripole_push(a, c, (r - 1) - (a.ny - 1), unew, vnew);
            }
        }
        s1p = str[0]; s5p = str[4]; n2p = e2; n7p = e7;
        u_s = u_c; v_s = v_c; uw_s = uw_c; vw_s = vw_c; mprev = m;
    }
}

// calc_ffrac (:1795-1864): both components at once (blockno 1 with a11, blockno 2 with a12 share everything up to mresult)
__device__ __forceinline__ void eap_calc_ffrac(double stressp, double stressm, double stress12, double a11, double a12,
                                               double &mresult11, double &mresult12)
{
    const double kfrac = 0.001, threshold = 3.0 * 0.1;
    const double sigma11 = 0.5 * (stressp + stressm), sigma12 = stress12, sigma22 = 0.5 * (stressp - stressm);
    const double gamma = 0.5 * evp_atan2((2.0 * sigma12), (sigma11 - sigma22));
    double Q11, Q12;
    evp_sincos(gamma, &Q12, &Q11);
    const double Q11Q11 = Q11 * Q11, Q11Q12 = Q11 * Q12, Q12Q12 = Q12 * Q12;
    const double sigma_1 = Q11Q11 * sigma11 + 2.0 * Q11Q12 * sigma12 + Q12Q12 * sigma22;
    const double sigma_2 = Q12Q12 * sigma11 - 2.0 * Q11Q12 * sigma12 + Q11Q11 * sigma22;
    bool diffuse;
    if ((sigma_1 >= 0.0) && (sigma_2 >= 0.0)) diffuse = false;                    // pure divergence
    else if ((sigma_1 >= 0.0) && (sigma_2 < 0.0)) diffuse = true;                 // unconfined compression
    else if (sigma_2 == 0.0) diffuse = false;                                     // shear faulting
    else if ((sigma_1 <= 0.0) && (sigma_1 / sigma_2 <= threshold)) diffuse = true;
    else diffuse = false;                                                         // horizontal spalling
    mresult11 = diffuse ? kfrac * (a11 - Q12Q12) : 0.0;
    mresult12 = diffuse ? kfrac * (a12 + Q11Q12) : 0.0;
}

// stepa (:1664-1787), one thread per (T cell of the list, corner) -- mask bit 0; the list includes the north / east ghost
// cells -- followed by the rotation of the new tensor (eap_rotation) for the subcycles to come; the cell averages
// (:1780-1781) through shared memory, summed in the reference's order.
__global__ void __launch_bounds__(128)
eap_stepa_kernel(const __grid_constant__ EapArgs e, int nxl, int ny, int ld, int ldm, const unsigned char *mask, const double *sp0,
                 size_t plane_elems)
{
    __shared__ double sh11[4][32], sh12[4][32];
    const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c = 1 + blockIdx.x * 32 + lane, r = 1 + blockIdx.y;
    const bool act = c <= nxl + 1 && (mask[(size_t)r * ldm + min(c, nxl + 1)] & 1u);
    const size_t idx = (size_t)r * ld + c;
    if (act) {
        const double kth = 0.2 * 0.001;
        const double dteikth = 1.0 / (e.dtei + kth), p5kth = 0.5 * kth;
        // stresses: 12 consecutive planes p1..4 m1..4 12_1..4
        const double stressp = sp0[(size_t)q * plane_elems + idx];
        const double stressm = sp0[(size_t)(4 + q) * plane_elems + idx];
        const double stress12 = sp0[(size_t)(8 + q) * plane_elems + idx];
        const double a11 = e.a11c[q][idx], a12 = e.a12c[q][idx];
        double m11, m12;
        eap_calc_ffrac(stressp, stressm, stress12, a11, a12, m11, m12);
        const double a11n = (a11 * e.dtei + p5kth - m11) * dteikth;
        const double a12n = (a12 * e.dtei - m12) * dteikth;
        e.a11c[q][idx] = a11n;
        e.a12c[q][idx] = a12n;
        sh11[q][lane] = a11n;
        sh12[q][lane] = a12n;
        double gamma, Q11, Q12, kaf;
        eap_rotation(a11n, a12n, gamma, Q11, Q12, kaf);
        e.rot[4 * q][idx] = gamma; e.rot[4 * q + 1][idx] = Q11; e.rot[4 * q + 2][idx] = Q12; e.rot[4 * q + 3][idx] = kaf;
    }
    __syncthreads();
    if (act && q == 0) {
        e.a11[idx] = 0.25 * (sh11[0][lane] + sh11[1][lane] + sh11[2][lane] + sh11[3][lane]);
        e.a12[idx] = 0.25 * (sh12[0][lane] + sh12[1][lane] + sh12[2][lane] + sh12[3][lane]);
    }
}

__global__ void fill_plane_kernel(double *p, double v, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// the structure-tensor reset of the driver (:274-289) on every cell of the block, ghost ring included, then the rotation
// of the tensor on the T cells of the list (mask bit 0: set by prep2_kernel, launched before this kernel)
__global__ void __launch_bounds__(128)
eap_reset_kernel(const __grid_constant__ EapArgs e, int nxb, int nyb, int ld, int ldm, const double *icetmask,
                 const unsigned char *mask)
{
    const int q = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + (threadIdx.x & 31), r = blockIdx.y;
    if (c >= nxb || r >= nyb) return;
    const size_t idx = (size_t)r * ld + c;
    double a11 = e.a11c[q][idx], a12 = e.a12c[q][idx];
    if (icetmask[idx] == 0.0) {
        a11 = 0.5; a12 = 0.0;
        e.a11c[q][idx] = a11; e.a12c[q][idx] = a12;
    }
    if (mask[(size_t)r * ldm + c] & 1u) {
        double gamma, Q11, Q12, kaf;
        eap_rotation(a11, a12, gamma, Q11, Q12, kaf);
        e.rot[4 * q][idx] = gamma; e.rot[4 * q + 1][idx] = Q11; e.rot[4 * q + 2][idx] = Q12; e.rot[4 * q + 3][idx] = kaf;
    }
}

}  // namespace evp
