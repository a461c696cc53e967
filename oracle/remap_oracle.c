/*
 * remap_oracle.c -- CPU restatement of departure_points (TEST INFRASTRUCTURE, see remap_oracle.h).
 *
 * Follows, statement by statement: source/ice_transport_remap.F90:1493-1670 (departure_points, nghost = 1) and the
 * two ice_HaloUpdate calls on dpx, dpy (field_loc_NEcorner, field_type_vector) of horizontal_remap (:574-578).
 * Compile with -ffp-contract=off: the products and sums below are the reference's, in its order.
 */
#include <stddef.h>

#include "remap_oracle.h"

static const double c0 = 0.0, p5 = 0.5;

int remap_oracle_departure_points(const evp_oracle_cfg *cfg, double dt, int l_dp_midpt, int no_halo,
                                  const double *uvel, const double *vvel, const double *dxu, const double *dyu,
                                  const double *HTN, const double *HTE, double *dpx, double *dpy,
                                  int *l_stop, int *istop, int *jstop)
{
    if (!cfg || !uvel || !vvel || !dxu || !dyu || !HTN || !HTE || !dpx || !dpy || !l_stop || !istop || !jstop) return 1;
    const int nxb = cfg->nx_global + 2, nyb = cfg->ny_global + 2;
    const int ilo = 2, ihi = nxb - 1, jlo = 2, jhi = nyb - 1;              /* Fortran indices, nghost = 1 */
#define A(f, i, j) f[(size_t)((j) - 1) * nxb + ((i) - 1)]
    for (size_t q = 0; q < (size_t)nxb * nyb; ++q) { dpx[q] = c0; dpy[q] = c0; }        /* :1557-1558 */
    *l_stop = 0; *istop = 0; *jstop = 0;

    for (int j = jlo; j <= jhi; ++j)                                       /* :1560-1575 */
        for (int i = ilo; i <= ihi; ++i) {
            A(dpx, i, j) = -dt * A(uvel, i, j);
            A(dpy, i, j) = -dt * A(vvel, i, j);
            if (A(dpx, i, j) < -A(HTN, i, j) || A(dpx, i, j) > A(HTN, i + 1, j) ||
                A(dpy, i, j) < -A(HTE, i, j) || A(dpy, i, j) > A(HTE, i, j + 1)) {
                *l_stop = 1; *istop = i; *jstop = j;
            }
        }
    if (*l_stop) return 0;                                                 /* :1577-1591: report and return */

    if (l_dp_midpt) {                                                      /* :1593-1668 */
        for (int j = jlo; j <= jhi; ++j)
            for (int i = ilo; i <= ihi; ++i) {
                if (A(uvel, i, j) != c0 || A(vvel, i, j) != c0) {
                    int i2 = i, j2 = j;
                    double mpx, mpy, mpxt = c0, mpyt = c0, ump, vmp;
                    A(dpx, i, j) = A(dpx, i, j) / A(dxu, i, j);
                    A(dpy, i, j) = A(dpy, i, j) / A(dyu, i, j);
                    mpx = p5 * A(dpx, i, j);
                    mpy = p5 * A(dpy, i, j);
                    if (mpx >= c0 && mpy >= c0) { i2 = i + 1; j2 = j + 1; mpxt = mpx - p5; mpyt = mpy - p5; }
                    else if (mpx < c0 && mpy < c0) { i2 = i; j2 = j; mpxt = mpx + p5; mpyt = mpy + p5; }
                    else if (mpx >= c0 && mpy < c0) { i2 = i + 1; j2 = j; mpxt = mpx - p5; mpyt = mpy + p5; }
                    else if (mpx < c0 && mpy >= c0) { i2 = i; j2 = j + 1; mpxt = mpx + p5; mpyt = mpy - p5; }
                    ump = A(uvel, i2 - 1, j2 - 1) * (mpxt - p5) * (mpyt - p5)
                        - A(uvel, i2, j2 - 1) * (mpxt + p5) * (mpyt - p5)
                        + A(uvel, i2, j2) * (mpxt + p5) * (mpyt + p5)
                        - A(uvel, i2 - 1, j2) * (mpxt - p5) * (mpyt + p5);
                    vmp = A(vvel, i2 - 1, j2 - 1) * (mpxt - p5) * (mpyt - p5)
                        - A(vvel, i2, j2 - 1) * (mpxt + p5) * (mpyt - p5)
                        + A(vvel, i2, j2) * (mpxt + p5) * (mpyt + p5)
                        - A(vvel, i2 - 1, j2) * (mpxt - p5) * (mpyt + p5);
                    A(dpx, i, j) = -dt * ump;
                    A(dpy, i, j) = -dt * vmp;
                }
            }
    }
#undef A
    if (!no_halo) {                                                        /* horizontal_remap :574-578 */
        if (evp_oracle_halo_r8(cfg, dpx, 1, 1, 0.0)) return 1;
        if (evp_oracle_halo_r8(cfg, dpy, 1, 1, 0.0)) return 1;
    }
    return 0;
}
