/*
 * evp_oracle.h -- CPU restatement of CICE5's EVP momentum solver (TEST INFRASTRUCTURE).
 *
 * This is the parity oracle for the B200 kernels: a plain-C restatement of
 * `subroutine evp(dt)` and its helpers, written from the reference Fortran
 * (citations are path:line relative to the reference checkout):
 *   source/ice_dyn_evp.F90:68-510     evp driver
 *   source/ice_dyn_evp.F90:520-849    stress
 *   source/ice_dyn_shared.F90:185-259 set_evp_parameters
 *   source/ice_dyn_shared.F90:270-365 evp_prep1
 *   source/ice_dyn_shared.F90:377-614 evp_prep2
 *   source/ice_dyn_shared.F90:623-748 stepu
 *   source/ice_dyn_shared.F90:757-844 evp_finish
 *   source/ice_grid.F90:1799-1958     t2ugrid_vector, to_ugrid, u2tgrid_vector, to_tgrid
 *   serial/ice_boundary.F90:630-896   ice_HaloUpdate2DR8 (tripole u-fold included)
 *   serial/ice_boundary.F90:3269-3443 ice_HaloUpdate_stress
 *   mpi/ice_boundary.F90:2625-2632    ghost-ring pre-fill of the MPI build
 *
 * PARITY PINNED TO THE REFERENCE TEXT (through an interpreter).  The reference ships no tests, golden logs or fixtures for this
 * path and cannot be compiled in this image (no Fortran compiler).  What pins this file:
 *   - ARITHMETIC (set_evp_parameters, evp_prep1, to_ugrid, evp_prep2, stress, stepu, evp_finish,
 *     to_tgrid, principal_stress; ice_strength in strength_oracle.c): the reference's own subroutine
 *     texts are executed mechanically from the checkout (tests/golden/fortran_exec.py, a
 *     Fortran-subset interpreter; tests/golden/make_reference_vectors.py) and the outputs committed
 *     as tests/golden/ref_block_vectors.npz; in `no_halo` mode this oracle must reproduce them BIT
 *     FOR BIT (tests/test_oracle_reference_text.py: default build, -Dcoupled, AusCOM with a turning
 *     angle, revised EVP).  An interpreter, not a compiled run: compiler-specific contraction / -ftz
 *     choices of a real build are not represented.
 *   - HALO / TRIPOLE code: the reference's create_blocks, ice_HaloCreate, ice_HaloMsgCreate and
 *     ice_HaloUpdate2DR8 / 3DR8 / 2DI4 / _stress of the MPI build are executed the same way on one task
 *     (tests/golden/make_halo_vectors.py -> ref_halo_vectors.npz) for tripole, open and closed domains
 *     cut into 1x1 ... 4x1 blocks; the halo functions here must reproduce every cell of every block,
 *     ghosts included, bit for bit, for 1 / 2 / 4 slabs.  Two findings: the 3-D routine that carries
 *     the velocity pair does not average the column pair (1, nx-1) of the tripole's top row
 *     (reproduced: halo_r8_ex, FIRST_PAIR_3D); ice_HaloUpdate_stress never writes the ring cells that
 *     have no source, which this oracle sets to the 0 they hold in any run of the reference (stated
 *     in the test).
 *   - THE WHOLE CALL: the driver itself (subroutine evp, source/ice_dyn_evp.F90:68-510) is executed the
 *     same way on tripole and open domains cut into 1x1 ... 4x1 blocks, calling the reference's own
 *     subroutines and halo code (tests/golden/make_driver_vectors.py -> ref_driver_vectors.npz: default
 *     build, -Dcoupled, revised EVP, ice on the tripole's poles); evp_oracle_run with halos on and 1 / 2 / 4
 *     slabs must reproduce all 34 inout / out fields bit for bit on every physical cell, and the
 *     halo-updated ones (uvel, vvel, strength) on every ghost cell.
 *   - Caveat: an interpreter is not a compiler -- contraction / -ftz choices of a real ifort build are
 *     not represented -- and the generic ice_HaloUpdate interface, ice_distributionGet* (one task) and
 *     the global reductions are stood in for.  Further defended by invariants (decomposition
 *     independence, free drift) and an independent numpy restatement (oracle/evp_numpy.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this code.  The product path never does.
 *
 * Array convention: every field is ONE global block with a width-1 ghost ring,
 * Fortran order (i fastest), dimensions (nx_global+2) x (ny_global+2), i.e.
 * element (i,j), 1-based, lives at [(j-1)*(nx_global+2) + (i-1)].  Internally
 * the domain is cut into `nslabs` x-slabs (one reference "block" each) so that
 * the block/halo structure of the reference is exercised; results are merged
 * back into the global-with-ghost arrays.
 */
#ifndef EVP_ORACLE_H
#define EVP_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int nx_global, ny_global; /* physical cells                                  */
    int nslabs;               /* number of x-slabs (blocks); divides nx_global   */
    int ndte;                 /* subcycles (ice_dyn_shared.F90:33)                */
    int ew_cyclic;            /* ew_boundary_type=='cyclic' (else ghosts = fill)  */
    int ns_tripole;           /* ns_boundary_type=='tripole' (else ghosts = fill) */
    int grid_tripole;         /* grid_type=='tripole' (ice_dyn_evp.F90:416)       */
    int tilt_mode;            /* 0: strtlt from geostrophic current (uncoupled, or
                                    AusCOM with .not.use_ocnslope);
                                 1: -gravit*umass*ss_tlt (#ifdef coupled)          */
    int strength_masked;      /* 1: strength kept only on icetmask==1 cells, 0 elsewhere
                                    (kstrength=1 list semantics, ice_mechred.F90:2175);
                                 0: used as given on all physical cells (kstrength=0,
                                    ice_mechred.F90:2259-2263)                         */
    int nthreads;             /* OpenMP threads for the ij loops (0 = default)    */
    double dt_init;           /* dt handed to init_evp -> set_evp_parameters      */
    double cosw, sinw;        /* ocean turning angle                              */
    double rhow, rhoi, rhos, gravit;
    double a_min, m_min;      /* ice_dyn_shared.F90:60-61                         */
    int revised_evp;          /* Bouillon et al. 2013 (ice_dyn_shared.F90:230-233): revp = 1,
                                 arlx1i / brlx from xmin, stresses zeroed every call (:492-504) */
    int no_halo;              /* 1: skip every halo update (nslabs must be 1): the block sequence
                                 prep1 -> to_ugrid -> prep2 -> ndte x (stress, stepu) -> finish -> to_tgrid
                                 on the ghost cells as given -- the mode in which the oracle is compared
                                 with the reference's subroutine texts executed mechanically
                                 (tests/golden/make_reference_vectors.py)                            */
} evp_oracle_cfg;

/* grid arrays, ghosts valid as the reference's init_grid2 leaves them */
typedef struct {
    const double *dxt, *dyt, *dxhy, *dyhx, *cxp, *cyp, *cxm, *cym;
    const double *tarea, *uarea, *tarear, *uarear, *tinyarea, *fcor;
    const int *tmask, *umask; /* 0/1 */
} evp_oracle_grid;

typedef struct {
    /* in (per call) */
    const double *aice, *vice, *vsno, *aice_init;
    const double *strairxT, *strairyT, *uocn, *vocn, *ss_tltx, *ss_tlty, *Cw;
    const double *strength_in; /* host ice_strength result on T cells (>= icetmask cells) */
    /* inout */
    double *uvel, *vvel;
    double *stressp[4], *stressm[4], *stress12[4];
    int *iceumask;
    double *fm;
    /* out */
    double *strairx, *strairy, *strtltx, *strtlty, *strocnx, *strocny, *strintx, *strinty;
    double *strocnxT, *strocnyT, *divu, *shear, *rdg_conv, *rdg_shear, *prs_sig;
    double *uvel_init, *vvel_init, *tmass, *strength;
    int *icetmask;             /* diagnostic: mask after its halo update */
    /* in, optional (NULL: the default build).  The ACCESS build with calc_strair = .false.
     * (source/ice_dyn_evp.F90:220-245): strairx/y are the coupler's U-grid stresses, copied whole,
     * and t2ugrid_vector is not called. */
    const double *strax, *stray;
} evp_oracle_state;

typedef struct {
    double loop_seconds;       /* wall time of the ksub loop only              */
    double total_seconds;      /* wall time of the whole evp call              */
    long long active_t_cells;  /* sum over blocks of icellt                    */
    long long active_u_cells;  /* sum over blocks of icellu                    */
} evp_oracle_stats;

/* One call of evp(dt).  Returns 0 on success. */
int evp_oracle_run(const evp_oracle_cfg *cfg, const evp_oracle_grid *g,
                   evp_oracle_state *s, double dt, evp_oracle_stats *stats);

/* Stand-alone halo updates on a global-with-ghost array (for fixtures & tests).
 * loc: 0 = center, 1 = NEcorner.  kind: 0 = scalar, 1 = vector.               */
int evp_oracle_halo_r8(const evp_oracle_cfg *cfg, double *a, int loc, int kind, double fill);
/* the same through ice_HaloUpdate3DR8 (one level of the velocity pair fld2): the NE-corner fold leaves
 * the column pair (1, nx-1) unaveraged, see halo_r8_ex in evp_oracle.c */
int evp_oracle_halo_r8_fld2(const evp_oracle_cfg *cfg, double *a, int loc, int kind, double fill);
int evp_oracle_halo_i4(const evp_oracle_cfg *cfg, int *a, int loc, int kind, int fill);
/* ice_HaloUpdate_stress(array1, array2, halo, center, scalar): array1's ghosts from array2 */
int evp_oracle_halo_stress(const evp_oracle_cfg *cfg, double *a1, const double *a2);

/* restartfile's treatment of freshly read stresses (source/ice_restart_driver.F90:370-394, 431-466): 12 x
 * ice_HaloUpdate_stress on a tripole grid, then zero south ghost row / west ghost column.  s[12] in place. */
int evp_oracle_restart_fixup(const evp_oracle_cfg *cfg, double *const *s);

/* principal_stress, ice_dyn_shared.F90:853-893 (diagnostic, n elements) */
void evp_oracle_principal_stress(long long n, const double *stressp_1, const double *stressm_1,
                                 const double *stress12_1, const double *prs_sig,
                                 double *sig1, double *sig2);

#ifdef __cplusplus
}
#endif
#endif
