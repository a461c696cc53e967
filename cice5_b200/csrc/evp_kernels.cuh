// evp_kernels.cuh -- sm_100a FP64 kernels of the EVP subcycle (stress + stepu + halo pushes).
//
// Reference semantics (path:line in the reference checkout):
//   stress     source/ice_dyn_evp.F90:618-847      stepu   source/ice_dyn_shared.F90:700-746
//   halo       serial/ice_boundary.F90:630-896 (NEcorner / vector, tripole u-fold)
//
// Layout: one SoA plane per field, row pitch `ld` doubles, element (c, r) = Fortran (i=c+1, j=r+1).
// Physical cells c = 1..nxl, r = 1..ny; ghost ring c = 0, nxl+1 and r = 0, ny+1.
//
// Fused subcycle kernel ("warp sweep"): a warp owns a strip of 31 U columns (32 T columns, the
// 32nd is the redundant east column) and marches north through a chunk of rows.  Per step it
// computes `stress` for one T row, keeps the previous row's stress combinations in registers,
// takes the east neighbour's via warp shuffles and finishes `stepu` for the U row below.  The
// 8-plane `str` intermediate of the reference never exists in memory.  Stresses and velocities
// are ping-ponged (old -> new planes) so that the redundant ring cells read the same inputs as
// their owners: results are bitwise independent of the tiling and of the GPU decomposition.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace evp {

constexpr int kWarpsPerBlock = 4;
constexpr int kUColsPerWarp = 31;
constexpr int kMaxRanks = 8;

struct Consts {
    double p166, p333, p111, p055, p027, p222;
    double ecci, arlx1i, denom1, brlx, revp;
    double cosw, sinw, rhow;
};

struct LoopArgs {
    int nxl, ny, ld, ldm;          // physical cols / rows, pitches (doubles, mask bytes)
    int rows_per_chunk;
    int nx_global, lo;             // global columns; lo = global index of local column c=0
    int nranks, rank, tripole;
    int start[kMaxRanks + 1];      // x-slabs: rank t owns global columns start[t]+1 .. start[t+1]
    int nxl_west;                  // physical columns of the west neighbour (its east ghost = nxl_west + 1)
    int inbox_ld;                  // pitch of one inbox row (doubles)
    const unsigned char *mask;     // bit0: T cell active (icetmask==1), bit1: U cell active (iceumask)
    const double *u0, *v0;
    double *u1, *v1;
    const double *s0[12];          // stressp_1..4, stressm_1..4, stress12_1..4 (old)
    double *s1[12];                // (new)
    const double *dxt, *dyt, *dxhy, *dyhx, *cxp, *cyp, *cxm, *cym, *tinyarea, *strength, *tarear;
    const double *aiu, *Cw, *uocn, *vocn, *waterx, *watery, *forcex, *forcey, *umassdti, *fm, *uarear;
    double *divu, *shear, *rdg_conv, *rdg_shear, *prs_sig, *strintx, *strinty;
    const double *uvel_init, *vvel_init;   // read by stepu only under revised EVP (revp = 1)
    // E-W ghost pushes: new-velocity planes of the west / east neighbour (self when cyclic on
    // one GPU, peer memory over NVLink otherwise, null on a closed edge)
    double *uW, *vW, *uE, *vE;
    // tripole: top-two-row inboxes of every rank (peer pointers), layout [u|v][row ny-1|ny][c]
    double *inbox[kMaxRanks];
    Consts k;
};

// owner rank t, its local column ct (1-based) and its width wt for global column X (1-based)
__device__ __forceinline__ void slab_owner(const int *start, int nranks, int X, int &t, int &ct, int &wt)
{
    t = 0;
    while (t + 1 < nranks && X > start[t + 1]) ++t;
    ct = X - start[t];
    wt = start[t + 1] - start[t];
}

__device__ __forceinline__ double shfl_up1(double x)
{
    return __shfl_up_sync(0xffffffffu, x, 1);
}
__device__ __forceinline__ double shfl_down1(double x)
{
    return __shfl_down_sync(0xffffffffu, x, 1);
}

// One corner of `stress` (source/ice_dyn_evp.F90:627-721): strain rates * area, Delta, replacement pressure
// over Delta, the three stress updates.  Every statement is the reference's, operand for operand; only the
// ORDER of independent statements differs from the source (corner by corner instead of quantity by quantity),
// which changes no bit and lets the compiler keep one corner's temporaries alive at a time instead of four.
struct Corner {
    double div, tension, shear, Delta, c0raw;      // c0raw = strength / max(Delta, tinyarea)  (:683-686)
};
__device__ __forceinline__ Corner stress_corner(const Consts &k, double div, double tension, double shear,
                                                double tinyarea, double strength, double &sp, double &sm,
                                                double &s12)
{
    Corner c;
    c.div = div; c.tension = tension; c.shear = shear;
    c.Delta = sqrt(div * div + k.ecci * (tension * tension + shear * shear));          // :657-660
    c.c0raw = strength / fmax(c.Delta, tinyarea);                                       // :683-686
    const double c1 = c.c0raw * k.arlx1i;                                               // :690-693
    const double c0 = c1 * k.ecci;                                                      // :694-697
    sp = (sp + c1 * (div - c.Delta)) * k.denom1;                                        // :704-709
    sm = (sm + c0 * tension) * k.denom1;                                                // :710-715
    s12 = (s12 + c0 * shear * 0.5) * k.denom1;                                          // :716-721
    return c;
}

// One T cell of `stress`.  s[12] in/out (p1..4, m1..4, 12_1..4); str[8] out.
template <bool LAST>
__device__ __forceinline__ void stress_cell(
    const Consts &k, double u_ij, double u_wj, double u_is, double u_ws, double v_ij, double v_wj,
    double v_is, double v_ws, double dxt, double dyt, double dxhy, double dyhx, double cxp,
    double cyp, double cxm, double cym, double tinyarea, double strength, double tarear,
    double (&s)[12], double (&str)[8], double &divu, double &shear, double &rdg_conv,
    double &rdg_shear, double &prs_sig)
{
    // strain rates * area (:627-654), corner by corner: ne (1), nw (2), sw (3), se (4)
    double sp1 = s[0], sp2 = s[1], sp3 = s[2], sp4 = s[3];
    double sm1 = s[4], sm2 = s[5], sm3 = s[6], sm4 = s[7];
    double s121 = s[8], s122 = s[9], s123 = s[10], s124 = s[11];
    const Corner ne = stress_corner(k, cyp * u_ij - dyt * u_wj + cxp * v_ij - dxt * v_is,
                                    -cym * u_ij - dyt * u_wj + cxm * v_ij + dxt * v_is,
                                    -cym * v_ij - dyt * v_wj - cxm * u_ij - dxt * u_is, tinyarea, strength, sp1, sm1, s121);
    const Corner nw = stress_corner(k, cym * u_wj + dyt * u_ij + cxp * v_wj - dxt * v_ws,
                                    -cyp * u_wj + dyt * u_ij + cxm * v_wj + dxt * v_ws,
                                    -cyp * v_wj + dyt * v_ij - cxm * u_wj - dxt * u_ws, tinyarea, strength, sp2, sm2, s122);
    const Corner sw = stress_corner(k, cym * u_ws + dyt * u_is + cxm * v_ws + dxt * v_wj,
                                    -cyp * u_ws + dyt * u_is + cxp * v_ws - dxt * v_wj,
                                    -cyp * v_ws + dyt * v_is - cxp * u_ws + dxt * u_wj, tinyarea, strength, sp3, sm3, s123);
    const Corner se = stress_corner(k, cyp * u_is - dyt * u_ws + cxm * v_is + dxt * v_ij,
                                    -cym * u_is - dyt * u_ws + cxp * v_is - dxt * v_ij,
                                    -cym * v_is - dyt * v_ws - cxp * u_is + dxt * u_ij, tinyarea, strength, sp4, sm4, s124);

    if (LAST) {  // ksub == ndte (:665-677)
        divu = 0.25 * (ne.div + nw.div + se.div + sw.div) * tarear;
        const double tmp = 0.25 * (ne.Delta + nw.Delta + se.Delta + sw.Delta) * tarear;
        rdg_conv = -fmin(divu, 0.0);
        rdg_shear = 0.5 * (tmp - fabs(divu));
        const double ts = ne.tension + nw.tension + se.tension + sw.tension;
        const double ss = ne.shear + nw.shear + se.shear + sw.shear;
        shear = 0.25 * tarear * sqrt(ts * ts + ss * ss);
        prs_sig = ne.c0raw * ne.Delta;                                                  // :688
    }
    s[0] = sp1; s[1] = sp2; s[2] = sp3; s[3] = sp4;
    s[4] = sm1; s[5] = sm2; s[6] = sm3; s[7] = sm4;
    s[8] = s121; s[9] = s122; s[10] = s123; s[11] = s124;

    // combinations (:752-795)
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

    // dF/dx (:800-820)
    double strp_tmp = 0.25 * dyt * (k.p333 * ssigpn + k.p166 * ssigps);
    double strm_tmp = 0.25 * dyt * (k.p333 * ssigmn + k.p166 * ssigms);
    str[0] = -strp_tmp - strm_tmp - str12ew + dxhy * (-csigpne + csigmne) + dyhx * csig12ne;
    str[1] = strp_tmp + strm_tmp - str12we + dxhy * (-csigpnw + csigmnw) + dyhx * csig12nw;
    strp_tmp = 0.25 * dyt * (k.p333 * ssigps + k.p166 * ssigpn);
    strm_tmp = 0.25 * dyt * (k.p333 * ssigms + k.p166 * ssigmn);
    str[2] = -strp_tmp - strm_tmp + str12ew + dxhy * (-csigpse + csigmse) + dyhx * csig12se;
    str[3] = strp_tmp + strm_tmp + str12we + dxhy * (-csigpsw + csigmsw) + dyhx * csig12sw;
    // dF/dy (:825-845)
    strp_tmp = 0.25 * dxt * (k.p333 * ssigpe + k.p166 * ssigpw);
    strm_tmp = 0.25 * dxt * (k.p333 * ssigme + k.p166 * ssigmw);
    str[4] = -strp_tmp + strm_tmp - str12ns - dyhx * (csigpne + csigmne) + dxhy * csig12ne;
    str[5] = strp_tmp - strm_tmp - str12sn - dyhx * (csigpse + csigmse) + dxhy * csig12se;
    strp_tmp = 0.25 * dxt * (k.p333 * ssigpw + k.p166 * ssigpe);
    strm_tmp = 0.25 * dxt * (k.p333 * ssigmw + k.p166 * ssigme);
    str[6] = -strp_tmp + strm_tmp + str12ns - dyhx * (csigpnw + csigmnw) + dxhy * csig12nw;
    str[7] = strp_tmp - strm_tmp + str12sn - dyhx * (csigpsw + csigmsw) + dxhy * csig12sw;
}

// Push a freshly computed top-row velocity (U rows ny-1, ny) into the tripole inbox of every
// rank whose local columns (ghosts included) fold onto this global column
// (copy-in :3737-3750, copy-out source nx - ig :3768 with ioffset = 1 :811,874-877).
__device__ __forceinline__ void tripole_push(const LoopArgs &a, double *const *inbox, int c, int rowsel, double u,
                                             double v)
{
    const int ig = a.lo + c;                  // global column, 1-based
    int X = a.nx_global - ig;                 // the column whose fold source is ig
    if (X == 0) X = a.nx_global;
    int t, ct, wt;                            // owner, local column on rank t (1..wt)
    slab_owner(a.start, a.nranks, X, t, ct, wt);
    const size_t urow = (size_t)(0 + rowsel) * a.inbox_ld, vrow = (size_t)(2 + rowsel) * a.inbox_ld;
    inbox[t][urow + ct] = u;
    inbox[t][vrow + ct] = v;
    if (ct == 1) {                            // also the east ghost of rank t-1
        const int t2 = (t + a.nranks - 1) % a.nranks;
        const int w2 = a.start[t2 + 1] - a.start[t2];
        inbox[t2][urow + w2 + 1] = u;
        inbox[t2][vrow + w2 + 1] = v;
    }
    if (ct == wt) {                           // also the west ghost of rank t+1
        const int t2 = (t + 1) % a.nranks;
        inbox[t2][urow] = u;
        inbox[t2][vrow] = v;
    }
}
__device__ __forceinline__ void tripole_push(const LoopArgs &a, int c, int rowsel, double u, double v)
{
    tripole_push(a, a.inbox, c, rowsel, u, v);
}

template <bool LAST, bool REVISED = false>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
evp_subcycle_kernel(const __grid_constant__ LoopArgs a)
{
    const int lane = threadIdx.x & 31;
    const int strip = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int c0 = 1 + strip * kUColsPerWarp;
    if (c0 > a.nxl) return;                               // warp-uniform
    const int c = c0 + lane;                              // T column of this lane (U column if lane < 31)
    const int r0 = 1 + blockIdx.y * a.rows_per_chunk;     // first U row of the chunk
    const int rend = min(r0 + a.rows_per_chunk, a.ny + 1);  // T rows r0..rend, U rows r0..rend-1
    const bool tcol = c <= a.nxl + 1;
    const bool ucol = lane < kUColsPerWarp && c <= a.nxl;
    const bool tstore_col = tcol && (lane < kUColsPerWarp || c == a.nxl + 1);
    const Consts &k = a.k;
    const int ld = a.ld;

    // velocities of the row below the first T row
    size_t idx = (size_t)(r0 - 1) * ld + c;
    double u_s = tcol ? a.u0[idx] : 0.0, v_s = tcol ? a.v0[idx] : 0.0;
    double uw_s = shfl_up1(u_s), vw_s = shfl_up1(v_s);
    if (lane == 0) { uw_s = a.u0[idx - 1]; vw_s = a.v0[idx - 1]; }
    double s1p = 0.0, s5p = 0.0, n2p = 0.0, n7p = 0.0;    // previous T row: own str1,5; east str2,7
    unsigned mprev = 0;

    for (int r = r0; r <= rend; ++r) {
        idx = (size_t)r * ld + c;
        const double u_c = tcol ? a.u0[idx] : 0.0, v_c = tcol ? a.v0[idx] : 0.0;
        double uw_c = shfl_up1(u_c), vw_c = shfl_up1(v_c);
        if (lane == 0) { uw_c = a.u0[idx - 1]; vw_c = a.v0[idx - 1]; }
        const unsigned m = tcol ? a.mask[(size_t)r * a.ldm + c] : 0u;
        const bool tact = (m & 1u) != 0;

        double str[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (__ballot_sync(0xffffffffu, tact)) {
            if (tact) {
                double s[12];
#pragma unroll
                for (int q = 0; q < 12; ++q) s[q] = a.s0[q][idx];
                double divu, shear, rdg_conv, rdg_shear, prs_sig;
                stress_cell<LAST>(k, u_c, uw_c, u_s, uw_s, v_c, vw_c, v_s, vw_s, a.dxt[idx],
                                  a.dyt[idx], a.dxhy[idx], a.dyhx[idx], a.cxp[idx], a.cyp[idx],
                                  a.cxm[idx], a.cym[idx], a.tinyarea[idx], a.strength[idx],
                                  LAST ? a.tarear[idx] : 0.0, s, str, divu, shear, rdg_conv,
                                  rdg_shear, prs_sig);
                if (tstore_col && (r < rend || r == a.ny + 1)) {
#pragma unroll
                    for (int q = 0; q < 12; ++q) a.s1[q][idx] = s[q];
                    if (LAST) {
                        a.divu[idx] = divu; a.shear[idx] = shear; a.rdg_conv[idx] = rdg_conv;
                        a.rdg_shear[idx] = rdg_shear; a.prs_sig[idx] = prs_sig;
                    }
                }
            }
        }
        // east neighbour's combinations (T column c+1)
        const double e2 = shfl_down1(str[1]), e4 = shfl_down1(str[3]);
        const double e7 = shfl_down1(str[6]), e8 = shfl_down1(str[7]);

        if (r > r0) {                                     // stepu on U row r-1 (:700-746)
            const size_t iu = idx - ld;
            const bool uact = ucol && (mprev & 2u);
            double unew = u_s, vnew = v_s;
            if (__ballot_sync(0xffffffffu, uact)) {
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
                    // revp = 0 for classic EVP, so revp*uvel_init vanishes (:731-734)
                    const double cc1 = strintx + a.forcex[iu] + taux +
                                       umdt * (REVISED ? k.brlx * uold + k.revp * a.uvel_init[iu] : k.brlx * uold);
                    const double cc2 = strinty + a.forcey[iu] + tauy +
                                       umdt * (REVISED ? k.brlx * vold + k.revp * a.vvel_init[iu] : k.brlx * vold);
                    unew = (cca * cc1 + ccb * cc2) / ab2;
                    vnew = (cca * cc2 - ccb * cc1) / ab2;
                    if (LAST) { a.strintx[iu] = strintx; a.strinty[iu] = strinty; }
                }
            }
            if (ucol) {
                a.u1[iu] = unew; a.v1[iu] = vnew;
                // width-1 halo, E-W part: my edge columns are the neighbours' ghosts
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

// Tripole u-fold of one column of an NEcorner field (serial/ice_boundary.F90:809-824, 861-890):
// the top physical row r = ny is replaced by the (anti)symmetrised average, the ghost row
// r = ny+1 by the mirrored row ny-1.  own = this column's row-ny value before the fold,
// src_top / src_below = rows ny / ny-1 of the fold source column (global column nx - ig).
// first_pair: first column pair (i, nx - i) that is averaged -- kFirstPair2D in ice_HaloUpdate2DR8
// (mpi/ice_boundary.F90:1562), kFirstPair3D in ice_HaloUpdate3DR8, the routine the velocity pair fld2
// goes through (ice_dyn_evp.F90:314, :397): its loop starts at nxGlobal/2 - lenUsedTripoleBuf + 1
// (mpi/ice_boundary.F90:2817) and ice_HaloCreate leaves lenUsedTripoleBuf = nxGlobal/2 - 1 (:1067-1071),
// so the pair (1, nx-1) next to the pole is mirrored with its sign but not averaged.  Pinned by executing
// the reference text (tests/golden/make_halo_vectors.py, tests/test_oracle_reference_text.py).
constexpr int kFirstPair2D = 1, kFirstPair3D = 2;
__device__ __forceinline__ void fold_necorner_column(double *plane, int ld, int ny, int c, int ig,
                                                     int nx, double own, double src_below,
                                                     double src_top, double sgn, int first_pair)
{
    const int m = (nx - ig == 0) ? nx : nx - ig;
    const int half = nx / 2;
    double x;
    if (ig >= first_pair && ig <= half - 1) {
        x = 0.5 * (own + sgn * src_top);               // xavg; copy-out gives sgn*(sgn*xavg)
        x = sgn * (sgn * x);
    } else if (m >= first_pair && m <= half - 1) {
        x = sgn * (0.5 * (src_top + sgn * own));       // sgn * xavg of the pair (m, ig)
    } else {
        x = sgn * src_top;                             // self-mirrored / unpaired points
    }
    plane[(size_t)ny * ld + c] = x;
    plane[(size_t)(ny + 1) * ld + c] = sgn * src_below;
}

// ---- cross-rank synchronisation (one rank per GPU; flags live in every rank's arena) ------------
// signal: bump my epoch and store it into slot [my rank] of every peer's flag array.
// wait:   spin until every peer's slot in MY flag array has reached my epoch.
// Every rank executes the same sequence of halo operations, so epochs advance in lock step.
struct SyncArgs {
    int rank, nranks;
    unsigned long long *epoch;                    // mine
    unsigned long long *flags[kMaxRanks];         // flag arrays of all ranks (peer pointers)
    int *error;                                   // mine; set on time-out
    unsigned long long timeout_ns;
};

// one warp: everything this rank stored before (made visible to lane 0 by fences / barriers) is
// released to the peers together with the new epoch; lane r stores to peer r, in parallel
__device__ __forceinline__ void peer_signal_warp(const SyncArgs &a, int lane)
{
    unsigned long long e = 0;
    if (lane == 0) {
        e = *(volatile unsigned long long *)a.epoch + 1;
        *(volatile unsigned long long *)a.epoch = e;
        __threadfence_system();
    }
    e = __shfl_sync(0xffffffffu, e, 0);
    if (lane < a.nranks && lane != a.rank) {
        unsigned long long *p = a.flags[lane] + a.rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(e) : "memory");
    }
}

__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// thread r < nranks waits for peer r; callers follow with __syncthreads + __threadfence_system
__device__ __forceinline__ void peer_wait(const SyncArgs &a, int r)
{
    if (r < a.nranks && r != a.rank) {
        const unsigned long long e = *(volatile unsigned long long *)a.epoch;
        const unsigned long long *p = a.flags[a.rank] + r;
        unsigned long long t0, t1, v;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
            if (v >= e) break;
            if (*(volatile int *)a.error) break;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > a.timeout_ns) { *(volatile int *)a.error = 1; break; }
            __nanosleep(100);
        }
    }
}

// Fold of the freshly stepped (u, v): one thread per local column, ghost columns included (their
// row-ny value was pushed into the plane by the E-W neighbour's subcycle kernel).
// inbox rows: [0] u row ny-1, [1] u row ny, [2] v row ny-1, [3] v row ny of the fold source.
// do_wait: first wait until every peer has signalled the end of its subcycle kernel (the barrier
// that otherwise is the separate halo_wait_kernel).
__global__ void fold_uv_kernel(double *u, double *v, int ld, int nxl, int ny, int nx, int lo,
                               const double *inbox, int inbox_ld, const __grid_constant__ SyncArgs sa, int do_wait,
                               unsigned long long *trace)
{
    if (do_wait) {
        if (trace && blockIdx.x == 0 && threadIdx.x == 0) trace[3] = global_ns();
        peer_wait(sa, threadIdx.x);
        __syncthreads();
        __threadfence_system();
        if (trace && blockIdx.x == 0 && threadIdx.x == 0) trace[4] = global_ns();
    }
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c > nxl + 1) return;
    int ig = lo + c;
    if (ig < 1) ig += nx;
    if (ig > nx) ig -= nx;
    fold_necorner_column(u, ld, ny, c, ig, nx, u[(size_t)ny * ld + c], inbox[c], inbox[inbox_ld + c], -1.0,
                         kFirstPair3D);
    fold_necorner_column(v, ld, ny, c, ig, nx, v[(size_t)ny * ld + c], inbox[2 * inbox_ld + c],
                         inbox[3 * inbox_ld + c], -1.0, kFirstPair3D);
}

}  // namespace evp
