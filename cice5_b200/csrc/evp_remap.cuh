// evp_remap.cuh -- first stage of horizontal transport on the device: departure_points.
//
// reference: source/ice_transport_remap.F90:1493-1670 (departure_points, nghost = 1), called by horizontal_remap
// (:541-550) right after evp in step_dynamics (source/ice_step_mod.F90:1126-1130); the ghost cells of dpx, dpy are then
// filled by two ice_HaloUpdate calls (NE corner, vector; :574-578) -- the generic halo of evp_prep.cuh.  Reads uvel,
// vvel straight from the resident EVP state: the velocities never leave the device between evp and this stage.
#pragma once
#include "evp_kernels.cuh"

namespace evp {

struct DepartArgs {
    int nxb, nyb, ld;              // block incl. ghosts, plane pitch
    double dt;
    const double *u, *v;           // resident velocities (current ping-pong copy)
    const double *dxu, *dyu, *htn, *hte;
    double *dpx, *dpy;
    int *stop;                     // 0: all departure points in bounds; else 1 + (j-1)*nxb + (i-1) of the LAST
                                   // offending cell in the reference's loop order (j outer, i inner; 1-based i, j)
    int midpt;                     // l_dp_midpt
};

// dpx = -dt*uvel, dpy = -dt*vvel on the physical cells, 0 on the ghost ring; bounds check (:1557-1575)
__global__ void departure_first_kernel(const __grid_constant__ DepartArgs a)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c >= a.nxb) return;
    const size_t q = (size_t)r * a.ld + c;
    double px = 0.0, py = 0.0;
    if (c >= 1 && c <= a.nxb - 2 && r >= 1 && r <= a.nyb - 2) {
        px = -a.dt * a.u[q];
        py = -a.dt * a.v[q];
        if (px < -a.htn[q] || px > a.htn[q + 1] || py < -a.hte[q] || py > a.hte[q + a.ld])
            atomicMax(a.stop, 1 + r * a.nxb + c);
    }
    a.dpx[q] = px;
    a.dpy[q] = py;
}

// midpoint correction (:1593-1668): skipped altogether when any cell was out of bounds (the reference returns
// before it) and for cells at rest
__global__ void departure_midpt_kernel(const __grid_constant__ DepartArgs a)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (!a.midpt || c < 1 || c > a.nxb - 2 || r < 1 || r > a.nyb - 2) return;
    if (*(volatile int *)a.stop) return;
    const size_t q = (size_t)r * a.ld + c;
    if (a.u[q] == 0.0 && a.v[q] == 0.0) return;
    const double sx = a.dpx[q] / a.dxu[q], sy = a.dpy[q] / a.dyu[q];        // unit-cell coordinates
    const double mpx = 0.5 * sx, mpy = 0.5 * sy;                             // midpoint of the back trajectory
    // cell (i2, j2) that holds the midpoint, and the midpoint relative to that cell's centre
    const int di = mpx >= 0.0 ? 1 : 0, dj = mpy >= 0.0 ? 1 : 0;
    const double mpxt = mpx >= 0.0 ? mpx - 0.5 : mpx + 0.5;
    const double mpyt = mpy >= 0.0 ? mpy - 0.5 : mpy + 0.5;
    const size_t q2 = q + (size_t)dj * a.ld + di;                            // (i2, j2)
    const double xm = mpxt - 0.5, xp = mpxt + 0.5, ym = mpyt - 0.5, yp = mpyt + 0.5;
    const double ump = a.u[q2 - a.ld - 1] * xm * ym - a.u[q2 - a.ld] * xp * ym + a.u[q2] * xp * yp - a.u[q2 - 1] * xm * yp;
    const double vmp = a.v[q2 - a.ld - 1] * xm * ym - a.v[q2 - a.ld] * xp * ym + a.v[q2] * xp * yp - a.v[q2 - 1] * xm * yp;
    a.dpx[q] = -a.dt * ump;
    a.dpy[q] = -a.dt * vmp;
}

}  // namespace evp
