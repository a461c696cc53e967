/* remap_oracle.h -- CPU restatement of the first stages of CICE5's horizontal transport:
 * make_masks / construct_fields / limited_gradient (:867-1484), departure_points (source/ice_transport_remap.F90:1493-1670) and the ghost-cell update of dpx, dpy that
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

/* Second stage: make_masks (:867-1015) + construct_fields (:1024-1340) + limited_gradient (:1348-1484) for the
 * open-water category 0 and the ice categories 1..ncat, as horizontal_remap calls them (:462-511), then -- unless
 * no_halo -- the ghost-cell updates of mc, mx, my, tc, tx, ty (:579-584, :616-621; maskhalo_remap = .false.).
 * mm: ncat+1 planes (category 0 first); tm, tc, tx, ty: planes [n-1][nt-1], n = 1..ncat, nt = 1..ntrace, i.e. the
 * reference's (nx_block, ny_block, ntrace, ncat).  tracer_type / depend / has_dependents: init_transport's arrays
 * (source/ice_transport_driver.F90:82-118), depend 1-based.  mmask_out / tmask_out / icells_out may be NULL. */
int remap_oracle_construct_fields(const evp_oracle_cfg *cfg, int ncat, int ntrace, const int *tracer_type,
                                  const int *depend, const int *has_dependents, int no_halo, const double *hm,
                                  const double *xav, const double *yav, const double *xxav, const double *yyav,
                                  const double *mm, const double *tm, double *mc, double *mx, double *my, double *tc,
                                  double *tx, double *ty, double *mmask_out, double *tmask_out, int *icells_out);

/* Third stage, the reference's compile-time settings l_fixed_area = .false., integral_order = 3
 * (source/ice_transport_driver.F90:31,47).  One block nxb x nyb incl. the ghost ring (ilo = jlo = 2, nghost = 1).
 * Arrays: xp, yp [6 groups][4 vertices 0..3][nyb][nxb]; iflux, jflux (Fortran indices), triarea [6][nyb][nxb].
 *   triangles:           locate_triangles (:1680-3047) for the north (north != 0) or east edges, then, if coords != 0,
 *                        triangle_coordinates (:3078-3187); icells[6] = cells per group; edgearea (may be NULL) as left
 *   transport_integrals: (:3199-3509) from those arrays; tc == NULL: mass only (the open-water call)
 *   update_fields:       (:3517-3729); tm == NULL: mass only
 *   horizontal_remap:    all of :309-850 on a global-with-ghost block (see remap_oracle.c) */
int remap_oracle_triangles(int nxb, int nyb, int north, int coords, const double *dpx, const double *dpy,
                           const double *dxu, const double *dyu, double *xp, double *yp, int *iflux, int *jflux,
                           double *triarea, int *icells, double *edgearea);
int remap_oracle_transport_integrals(int nxb, int nyb, int ntrace, const int *tracer_type, const int *depend,
                                     const double *xp, const double *yp, const int *iflux, const int *jflux,
                                     const double *triarea, const double *mc, const double *mx, const double *my,
                                     double *mflx, const double *tc, const double *tx, const double *ty, double *mtflx);
int remap_oracle_update_fields(int nxb, int nyb, int ntrace, const int *tracer_type, const int *depend,
                               const double *tarear, const double *mflxe, const double *mflxn, double *mm,
                               const double *mtflxe, const double *mtflxn, double *tm, int *l_stop, int *istop, int *jstop);
int remap_oracle_horizontal_remap(const evp_oracle_cfg *cfg, double dt, int ncat, int ntrace, const int *tracer_type,
                                  const int *depend, const int *has_dependents, int no_halo, const double *hm,
                                  const double *xav, const double *yav, const double *xxav, const double *yyav,
                                  const double *dxu, const double *dyu, const double *HTN, const double *HTE,
                                  const double *tarear, const double *uvel, const double *vvel, double *mm, double *tm,
                                  int *stop, int *stop_cat, int *istop, int *jstop);

/* The caller, transport_remap (source/ice_transport_driver.F90:198-627) with l_conservation_check = l_monotonicity_check
 * = .false.: state_to_tracers (:789-900), horizontal_remap, tracers_to_state (:908-1002), bound_state
 * (source/ice_state.F90:173-238).  aicen, vicen, vsnon: ncat planes; trcrn: planes [n-1][it-1] = the reference's
 * (nx_block, ny_block, ntrcr, ncat); ntrace = 2 + ntrcr; nt_qsno 1-based; tracer tables as for construct_fields. */
int remap_oracle_state_to_tracers(int nxb, int nyb, int ncat, int ntrcr, int nt_qsno, int nslyr, double rhos, double Lfresh,
                                  const double *aice0, const double *aicen, const double *trcrn, const double *vicen,
                                  const double *vsnon, double *aim, double *trm);
int remap_oracle_tracers_to_state(int nxb, int nyb, int ncat, int ntrcr, int nt_qsno, int nslyr, double rhos, double Lfresh,
                                  const double *aim, const double *trm, double *aice0, double *aicen, double *trcrn,
                                  double *vicen, double *vsnon);
int remap_oracle_transport_remap(const evp_oracle_cfg *cfg, double dt, int ncat, int ntrcr, int nt_qsno, int nslyr,
                                 double rhos, double Lfresh, const int *tracer_type, const int *depend,
                                 const int *has_dependents, int no_halo, const double *hm, const double *xav,
                                 const double *yav, const double *xxav, const double *yyav, const double *dxu,
                                 const double *dyu, const double *HTN, const double *HTE, const double *tarear,
                                 const double *uvel, const double *vvel, double *aice0, double *aicen, double *trcrn,
                                 double *vicen, double *vsnon, int *stop, int *stop_cat, int *istop, int *jstop);

#ifdef __cplusplus
}
#endif
#endif
