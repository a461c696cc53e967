// evp_strength.cuh -- ice_strength on the device.
//
// reference: source/ice_mechred.F90:2111-2269 (ice_strength), :936-1285 (ridge_itd), constants
// :65-82; call site source/ice_dyn_evp.F90:291-301.  One thread per T cell, the whole block
// (ghost ring included: the reference's list holds the N/E ghost cells too); prep2_kernel then keeps
// the values on the icetmask == 1 cells only (strength_masked semantics, ice_mechred.F90:2175) and
// the halo update of evp (:311) overwrites the ghosts.
//
// Everything of ridge_itd that the strength needs is a running scalar here: the cumulative
// distribution G(n) is swept once for its normalisation and once more for the participation
// function, so no per-category arrays live in registers beyond what one category needs.
// sqrt is IEEE; exp is evp_exp (evp_exp.h): double-double evaluation rounded once, correctly rounded, the
// same definition the CPU oracle compiles -- so `strength` is bitwise equal to the oracle in every mode,
// the default exponential participation function included (CUDA's own exp is 1 ulp and differs from every
// libm; at ndte = 120 that ulp grows beyond 1e-10 on ill-conditioned states).
#pragma once
#include "evp_kernels.cuh"
#include "evp_exp.h"

namespace evp {

struct StrengthArgs {
    int nxb, nyb, ld;              // block incl. ghosts, plane pitch
    int ncat, kstrength, krdg_partic, krdg_redist;
    double mu_rdg, Cf, Cp;         // Cp = 0.5*gravit*(rhow-rhoi)*rhoi/rhow (ice_mechred.F90:68)
    size_t cat_stride;             // doubles between two categories of aicen / vicen
    const double *aice, *vice, *aice0, *aicen, *vicen;
    double *strength;
};

__global__ void ice_strength_kernel(const __grid_constant__ StrengthArgs a)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c >= a.nxb) return;
    const size_t q = (size_t)r * a.ld + c;
    constexpr double puny = 1.0e-11, Gstar = 0.15, astar = 0.05, maxraft = 1.0, Hstar = 25.0;

    if (a.kstrength != 1) {        // Hibler 79 on the physical cells (:2259-2265), 0 elsewhere
        const bool phys = c >= 1 && c <= a.nxb - 2 && r >= 1 && r <= a.nyb - 2;
        a.strength[q] = phys ? 2.75e4 * a.vice[q] * evp_exp(-20.0 * (1.0 - a.aice[q])) : 0.0;
        return;
    }

    // ---- normalisation of the cumulative distribution: 1 / Gsum(ncat)  (:1030-1058) ----
    const double G0raw = a.aice0[q] > puny ? a.aice0[q] : 0.0;       // Gsum(0) (Gsum(-1) = 0)
    double Gtop = G0raw;
    for (int n = 0; n < a.ncat; ++n) {
        const double an = a.aicen[n * a.cat_stride + q];
        if (an > puny) Gtop = Gtop + an;
    }
    const double work = 1.0 / Gtop;

    const double Gstari = 1.0 / Gstar, astari = 1.0 / astar;
    const double xtmp = 1.0 / (1.0 - evp_exp(-astari));                  // :1108
    // participation of open water, category 0 (:1085-1095 / :1110-1122)
    double Gm1 = 0.0;                       // Gsum(n-1), normalised (Gsum(-1) = 0 is never scaled)
    double Gn = G0raw * work;               // Gsum(n)
    double Em1 = evp_exp(-Gm1 * astari) * xtmp, En = evp_exp(-Gn * astari) * xtmp;
    auto partic = [&](double lo, double hi, double elo, double ehi) -> double {
        if (a.krdg_partic != 0) return elo - ehi;
        if (hi < Gstar) return Gstari * (hi - lo) * (2.0 - (lo + hi) * Gstari);
        if (lo < Gstar) return Gstari * (Gstar - lo) * (2.0 - (lo + Gstar) * Gstari);
        return 0.0;
    };
    double aksum = partic(Gm1, Gn, Em1, En);                         // :1237
    double str = 0.0;
    double Graw = G0raw;                    // un-normalised running sum
    for (int n = 0; n < a.ncat; ++n) {
        const double an = a.aicen[n * a.cat_stride + q];
        const bool present = an > puny;
        if (present) Graw = Graw + an;
        Gm1 = Gn; Em1 = En;
        Gn = Graw * work;
        En = evp_exp(-Gn * astari) * xtmp;
        const double ap = partic(Gm1, Gn, Em1, En);
        double krdg = 1.0, hrmin = 0.0, hrmax = 0.0, hrexp = 0.0, hi = 0.0;
        if (present) {
            hi = a.vicen[n * a.cat_stride + q] / an;
            if (a.krdg_redist == 0) {                                // Hibler 80 (:1153-1170)
                hrmin = fmin(2.0 * hi, hi + maxraft);
                hrmax = 2.0 * sqrt(Hstar * hi);
                hrmax = fmax(hrmax, hrmin + puny);
                krdg = (0.5 * (hrmin + hrmax)) / hi;
            } else {                                                 // exponential (:1203-1224)
                const double hc = fmax(hi, puny);
                hrmin = fmin(2.0 * hc, hc + maxraft);
                hrexp = a.mu_rdg * sqrt(hc);
                krdg = (hrmin + hrexp) / hc;
            }
        }
        aksum = aksum + ap * (1.0 - 1.0 / krdg);                     // :1243-1244
        if (present && ap > 0.0) {                                   // :2204-2236
            double h2rdg;
            if (a.krdg_redist == 0)
                h2rdg = (1.0 / 3.0) * (hrmax * hrmax * hrmax - hrmin * hrmin * hrmin) / (hrmax - hrmin);
            else
                h2rdg = hrmin * hrmin + 2.0 * hrmin * hrexp + 2.0 * hrexp * hrexp;
            const double dh2rdg = -hi * hi + h2rdg / krdg;
            str = str + ap * dh2rdg;
        }
    }
    a.strength[q] = a.Cf * a.Cp * str / aksum;                       // :2249
}

}  // namespace evp
