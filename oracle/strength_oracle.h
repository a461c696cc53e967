/* strength_oracle.h -- CPU restatement of ice_strength (source/ice_mechred.F90:2111-2269).
 * TEST INFRASTRUCTURE (see evp_oracle.h); pinned bit for bit to the reference text executed mechanically
 * (tests/test_oracle_reference_text.py), see evp_oracle.h. */
#ifndef STRENGTH_ORACLE_H
#define STRENGTH_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int nx, ny;          /* physical cells; arrays are (ny+2) x (nx+2) planes, i fastest          */
    int ncat;            /* thickness categories (ice_domain_size.F90 ncat; 5 in the BASELINE configs) */
    int kstrength;       /* 1 Rothrock 75 (default), 0 Hibler 79      (ice_mechred.F90:55-56)     */
    int krdg_partic;     /* 0 Thorndike 75, 1 exponential (default)   (:57-58)                    */
    int krdg_redist;     /* 0 Hibler 80, 1 exponential (default)      (:59-60)                    */
    double mu_rdg, Cf;   /* namelist: 3 / 17 (ice_init.F90:276-277)                               */
    double rhow, rhoi, gravit;
    int exp_mode;        /* 0: libm exp (what the reference-text vectors were produced with);
                          * 1: evp_exp (cice5_b200/csrc/evp_exp.h), the correctly rounded exp the device uses */
} strength_oracle_cfg;

/* strength = 0 outside the list; cellmask NULL = every cell is in the list */
int strength_oracle_run(const strength_oracle_cfg *cfg, const double *aice, const double *vice,
                        const double *aice0, const double *aicen, const double *vicen,
                        const int *cellmask, double *strength);

/* y[i] = evp_exp(x[i]) (cice5_b200/csrc/evp_exp.h) */
void strength_oracle_cr_exp(long long n, const double *x, double *y);

#ifdef __cplusplus
}
#endif
#endif
