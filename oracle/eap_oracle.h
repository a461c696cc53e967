/*
 * eap_oracle.h -- CPU restatement of CICE5's elastic-anisotropic-plastic dynamics, `eap(dt)` (kdyn = 2).
 * TEST INFRASTRUCTURE: the same rules as evp_oracle.h (only tests/, smoke() and bench.py's CPU legs load it).
 *
 * Restated from (path:line in the reference checkout):
 *   source/ice_dyn_eap.F90:66-486     eap driver (shares evp_prep1 / evp_prep2 / stepu / evp_finish and the halo
 *                                     pattern with evp; no ice_HaloUpdate_stress)
 *   source/ice_dyn_eap.F90:493-621    init_eap: the lookup tables s11r ... s22s (:557-618)
 *   source/ice_dyn_eap.F90:624-1047   w1, w2, s11kr, s12kr, s22kr, s11ks, s12ks, s22ks
 *   source/ice_dyn_eap.F90:1052-1467  stress_eap
 *   source/ice_dyn_eap.F90:1474-1658  update_stress_rdg
 *   source/ice_dyn_eap.F90:1664-1787  stepa
 *   source/ice_dyn_eap.F90:1795-1864  calc_ffrac
 *
 * PARITY PINNED TO THE REFERENCE TEXT, UP TO THE INTRINSICS atan2 / sin / cos: update_stress_rdg, calc_ffrac, stepa,
 * stress_eap, the six table kernels and w1, w2 are executed mechanically from the checkout
 * (tests/golden/make_eap_vectors.py) with the Fortran intrinsics atan2, sin, cos bound to the routines of
 * cice5_b200/csrc/evp_trig.h (trig_mode = 1 here: what the device uses); this file must reproduce the outputs bit for
 * bit (tests/test_oracle_reference_text.py).  The table generator runs with the C library's sin / cos / tan / exp on
 * both sides (the interpreter's math module calls the same libm), and the tables are an INPUT of the device path.
 */
#ifndef EAP_ORACLE_H
#define EAP_ORACLE_H

#include "evp_oracle.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { EAP_NX_YIELD = 41, EAP_NY_YIELD = 41, EAP_NA_YIELD = 21 };   /* ice_dyn_eap.F90:31-34 */
#define EAP_TABLE_ELEMS (EAP_NX_YIELD * EAP_NY_YIELD * EAP_NA_YIELD)

typedef struct {
    /* lookup tables, Fortran (nx_yield, ny_yield, na_yield), ix fastest (:36-37) */
    const double *s11r, *s12r, *s22r, *s11s, *s12s, *s22s;
    /* inout: structure tensor at the four corners ne, nw, sw, se (:39-41; restart fields :1872-1902) */
    double *a11_c[4], *a12_c[4];
    /* inout: cell averages, written by stepa on the icellt cells (:1780-1781) */
    double *a11, *a12;
    /* out, zeroed at the start of every call (:166-182) */
    double *e11, *e12, *e22, *s11, *s12, *s22, *yieldstress11, *yieldstress12, *yieldstress22;
    int trig_mode;            /* 0: libm atan2 / sin / cos; 1: evp_trig.h (what the device uses) */
} evp_oracle_eap;

/* One call of eap(dt); cfg, g, s as evp_oracle_run (revised_evp must be 0).  Returns 0 on success. */
int eap_oracle_run(const evp_oracle_cfg *cfg, const evp_oracle_grid *g, evp_oracle_state *s, evp_oracle_eap *e,
                   double dt, evp_oracle_stats *stats);

/* init_eap's tables (:557-618); each array EAP_TABLE_ELEMS doubles.  ia_lo..ia_hi, ix_lo.., iy_lo..: 1-based inclusive
 * index ranges to fill (the whole table: 1..21, 1..41, 1..41); elements outside keep their values. */
void eap_oracle_init_tables(double *s11r, double *s12r, double *s22r, double *s11s, double *s12s, double *s22s,
                            int ia_lo, int ia_hi, int ix_lo, int ix_hi, int iy_lo, int iy_hi);

/* the pieces, for the unit comparisons with the reference text */
void eap_oracle_update_stress_rdg(const evp_oracle_eap *tab, int ksub, int ndte, double divu, double tension,
                                  double shear, double a11, double a12, double strength, double *stressp,
                                  double *stressm, double *stress12, double *alphar, double *alphas);
double eap_oracle_calc_ffrac(int trig_mode, int blockno, double stressp, double stressm, double stress12, double a1x);
double eap_oracle_table_kernel(int which, double x, double y, double z, double phi);   /* 0..5: s11kr s12kr s22kr s11ks s12ks s22ks */
double eap_oracle_w(int which, double a);                                             /* 1: w1, 2: w2 */

/* evp_trig.h as plain functions (accuracy tests; the interpreter binds them as the Fortran intrinsics) */
double evp_oracle_sin(double x);
double evp_oracle_cos(double x);
double evp_oracle_atan2(double y, double x);

#ifdef __cplusplus
}
#endif
#endif
