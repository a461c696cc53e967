/* remap_oracle.h -- CPU restatement of the first stage of CICE5's horizontal transport:
 * departure_points (source/ice_transport_remap.F90:1493-1670) and the ghost-cell update of dpx, dpy that
 * horizontal_remap issues after it (:574-578).  TEST INFRASTRUCTURE (see evp_oracle.h); pinned bit for bit to the
 * reference's own text executed mechanically (tests/golden/make_remap_vectors.py, tests/test_oracle_reference_text.py). */
#ifndef REMAP_ORACLE_H
#define REMAP_ORACLE_H
#include "evp_oracle.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Arrays: (ny+2) x (nx+2) global-with-ghost planes, i fastest, one block (ilo = jlo = 2, nghost = 1).
 * no_halo != 0: leave the ghost cells of dpx, dpy at 0 as departure_points does (what the reference-text vectors hold).
 * l_stop / istop / jstop as the routine leaves them (Fortran 1-based indices of the LAST offending cell in loop order);
 * on l_stop the routine returns before the midpoint correction and before any halo update (the caller aborts). */
int remap_oracle_departure_points(const evp_oracle_cfg *cfg, double dt, int l_dp_midpt, int no_halo,
                                  const double *uvel, const double *vvel, const double *dxu, const double *dyu,
                                  const double *HTN, const double *HTE, double *dpx, double *dpy,
                                  int *l_stop, int *istop, int *jstop);

#ifdef __cplusplus
}
#endif
#endif
