/*
 * evp_host.c -- the C-ABI boundary exercised from plain C, no Python, no torch: what a host model
 * (CICE's Fortran through iso_c_binding) does around `call evp(dt)`.
 *
 *   gcc -O2 -std=c99 -Iinclude examples/evp_host.c -Lcice5_b200/lib -levp_b200_strict \
 *       -Wl,-rpath,$PWD/cice5_b200/lib -lm -o evp_host && ./evp_host
 *
 * A uniform closed basin (a ring of land cells around nx-2 x ny-2 ocean cells), ice at rest
 * under a uniform wind: after one evp(dt) the ice moves down-wind and to the right of it.  Prints
 * a few numbers; exit status = the library's status (5 = no CUDA device: there is no CPU fallback).
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "evp_b200.h"

static double *plane(size_t n, double v)
{
    double *p = (double *)malloc(n * sizeof(double));
    for (size_t q = 0; q < n; ++q) p[q] = v;
    return p;
}

int main(void)
{
    const int nx = 96, ny = 64, nxb = nx + 2, nyb = ny + 2;
    const size_t n = (size_t)nxb * nyb;
    const double dx = 25.0e3, dt = 3600.0;
    evp_config_t cfg;
    evp_handle_t *h = NULL;
    int rc;

    memset(&cfg, 0, sizeof(cfg));
    cfg.nx_global = nx; cfg.ny_global = ny; cfg.nx_block = nxb; cfg.ny_block = nyb;
    cfg.rank = 0; cfg.nranks = 1; cfg.device = 0;
    cfg.ndte = 120; cfg.kdyn = 1; cfg.revised_evp = 0;
    cfg.ew_boundary = EVP_EW_CYCLIC; cfg.ns_boundary = EVP_NS_OPEN;     /* as gx3; the land ring closes the basin */
    cfg.use_cuda_graph = 1;
    cfg.dt_init = dt; cfg.cosw = 1.0; cfg.sinw = 0.0;
    cfg.rhow = 1026.0; cfg.rhoi = 917.0; cfg.rhos = 330.0; cfg.gravit = 9.80616;
    cfg.a_min = 1.0e-3; cfg.m_min = 1.0e-2;
    rc = evp_gpu_init(&cfg, &h);
    if (rc != EVP_OK) {
        fprintf(stderr, "evp_gpu_init: %s\n", evp_gpu_strerror(rc));
        return rc;
    }

    /* uniform Cartesian metrics: HTE = HTN = dx (ice_grid.F90:341-367 with constant lengths) */
    evp_grid_t g;
    memset(&g, 0, sizeof(g));
    double *dxy = plane(n, dx), *zero = plane(n, 0.0), *cp = plane(n, dx), *cm = plane(n, -dx);
    double *area = plane(n, dx * dx), *arear = plane(n, 1.0 / (dx * dx)), *tiny = plane(n, 1.0e-11 * dx * dx);
    double *fcor = plane(n, 1.4e-4);
    int *tmask = (int *)calloc(n, sizeof(int)), *umask = (int *)calloc(n, sizeof(int));
    for (int j = 2; j < nyb - 2; ++j)                  /* one ring of land inside the ghost ring */
        for (int i = 2; i < nxb - 2; ++i) tmask[(size_t)j * nxb + i] = 1;
    for (int j = 1; j < nyb - 1; ++j)                  /* umask = min of the 4 T masks (:1590-1591) */
        for (int i = 1; i < nxb - 1; ++i) {
            const size_t q = (size_t)j * nxb + i;
            umask[q] = tmask[q] && tmask[q + 1] && tmask[q + nxb] && tmask[q + nxb + 1];
        }
    g.dxt = g.dyt = dxy; g.dxhy = g.dyhx = zero;
    g.cxp = g.cyp = cp; g.cxm = g.cym = cm;
    g.tarea = g.uarea = area; g.tarear = g.uarear = arear; g.tinyarea = tiny; g.fcor = fcor;
    g.tmask = tmask; g.umask = umask;
    g.HTE = g.HTN = dxy;                                /* optional: compact metric rows */
    rc = evp_gpu_set_grid(h, &g);
    if (rc != EVP_OK) { fprintf(stderr, "set_grid: %s\n", evp_gpu_strerror(rc)); return rc; }

    /* ice state and forcing */
    evp_in_t in;
    evp_inout_t io;
    evp_out_t out;
    memset(&in, 0, sizeof(in)); memset(&io, 0, sizeof(io)); memset(&out, 0, sizeof(out));
    double *aice = plane(n, 0.0), *vice = plane(n, 0.0), *vsno = plane(n, 0.0), *strength = plane(n, 0.0);
    double *tau = plane(n, 0.0);
    for (size_t q = 0; q < n; ++q)
        if (tmask[q]) {
            aice[q] = 0.9; vice[q] = 1.8; vsno[q] = 0.1;
            strength[q] = 2.75e4 * vice[q] * exp(-20.0 * (1.0 - aice[q]));    /* Hibler, ice_mechred.F90:2263 */
            tau[q] = 0.9 * 0.1;                                                 /* 0.1 N/m2 wind stress in x */
        }
    in.aice = in.aice_init = aice; in.vice = vice; in.vsno = vsno;
    in.strairxT = tau; in.strairyT = zero; in.uocn = in.vocn = zero; in.ss_tltx = in.ss_tlty = zero;
    double *cw = plane(n, 0.00536);
    in.Cdn_ocn = cw; in.strength = strength;
    double *state[14];
    for (int k = 0; k < 14; ++k) state[k] = plane(n, 0.0);
    io.uvel = state[0]; io.vvel = state[1];
    for (int c = 0; c < 4; ++c) { io.stressp[c] = state[2 + c]; io.stressm[c] = state[6 + c]; io.stress12[c] = state[10 + c]; }
    io.iceumask = (int *)calloc(n, sizeof(int));
    io.fm = plane(n, 0.0);
    out.divu = plane(n, 0.0); out.shear = plane(n, 0.0); out.strocnxT = plane(n, 0.0);

    rc = evp_gpu_run(h, dt, &in, &io, &out);            /* == call evp(dt) */
    if (rc != EVP_OK) {
        fprintf(stderr, "evp_gpu_run: %s [%s]\n", evp_gpu_strerror(rc), evp_gpu_last_cuda_error());
        return rc;
    }
    evp_stats_t st;
    evp_gpu_get_stats(h, &st);
    const size_t mid = (size_t)(nyb / 2) * nxb + nxb / 2;
    printf("evp(dt): %lld active T cells, %lld kernel launches, loop %.3f ms\n", st.active_t_cells,
           st.kernel_launches, st.loop_ms);
    printf("centre of the basin: uvel = %.6e  vvel = %.6e m/s  (down-wind, deflected to the right)\n",
           io.uvel[mid], io.vvel[mid]);
    rc = (io.uvel[mid] > 0.0 && io.vvel[mid] < 0.0) ? EVP_OK : 100;
    evp_gpu_finalize(h);
    return rc;
}
