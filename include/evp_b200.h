/*
 * evp_b200.h -- C ABI of the B200-native EVP momentum solver (drop-in for CICE5 `evp(dt)`).
 *
 * The reference has no FFI or plugin API for this path: the boundary is the Fortran module
 * procedure `ice_dyn_evp::evp(dt)` (source/ice_dyn_evp.F90:47,68), called from
 * `step_dynamics` (source/ice_step_mod.F90:1119), which exchanges all of its data through
 * `use`-associated module arrays (source/ice_dyn_evp.F90:70-92).  The entry points below are
 * what a replacement `module ice_dyn_evp` binds with `iso_c_binding`
 * (cice5_b200/fortran/ice_dyn_evp_gpu.F90).
 *
 * Array convention (identical to the reference): every field is the calling task's block,
 * Fortran order, `real(dbl_kind)` / `integer(int_kind)`, dimensions (nx_block, ny_block)
 * with a width-1 ghost ring (nghost = 1, source/ice_blocks.F90:42-47); one block per task
 * (max_blocks = 1).  Logical masks are passed as int32 0/1.  Pointers are borrowed for the
 * duration of the call; the handle owns all device memory.  Every function returns 0 on
 * success and a non-zero status otherwise (text via evp_gpu_strerror; the Fortran shim hands
 * it to abort_ice, mpi/ice_exit.F90).  One handle per task, one task per GPU; calls are not
 * re-entrant (neither is the reference's caller).
 */
#ifndef EVP_B200_H
#define EVP_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct evp_handle evp_handle_t;

enum {
    EVP_OK = 0,
    EVP_ERR_ARG = 1,          /* null pointer / inconsistent sizes                      */
    EVP_ERR_UNSUPPORTED = 2,  /* kdyn not 1 / 2, unsupported boundary type / decomposition */
    EVP_ERR_CUDA = 3,         /* a CUDA runtime call failed (see evp_gpu_last_cuda_error) */
    EVP_ERR_STATE = 4,        /* call order violated (e.g. run before set_grid)         */
    EVP_ERR_NO_DEVICE = 5,    /* no usable CUDA device: there is NO CPU fallback        */
    EVP_ERR_PEER = 6,         /* multi-GPU peer mapping failed                          */
    EVP_ERR_TOO_LARGE = 7     /* block exceeds the static scratch of an allocation-free entry point */
};

enum { EVP_EW_CLOSED = 0, EVP_EW_CYCLIC = 1 };
enum { EVP_NS_OPEN = 0, EVP_NS_CLOSED = 1, EVP_NS_TRIPOLE = 2 };

/* Frozen at init (namelist + CPP switches of the reference build). */
typedef struct {
    int nx_global, ny_global;   /* ice_domain_size.F90:24-25                                  */
    int nx_block, ny_block;     /* block incl. ghosts: nx_global/nranks + 2, ny_global + 2     */
    int rank, nranks;           /* x-slab index / count (slenderX1-style decomposition)        */
    int device;                 /* CUDA device ordinal for this task                           */
    int ndte;                   /* dynamics_nml ndte (ice_dyn_shared.F90:33)                   */
    int kdyn;                   /* 1: EVP; 2: EAP (ice_dyn_eap.F90, classic relaxation only; needs
                                   evp_gpu_set_eap_tables before the first call)                */
    int revised_evp;            /* 0: classic EVP (ice_dyn_shared.F90:239-242); 1: revised EVP
                                   (Bouillon et al. 2013, :230-233,492-504,715,732-734)         */
    int ew_boundary;            /* EVP_EW_*   (ice_domain.F90 ew_boundary_type)                 */
    int ns_boundary;            /* EVP_NS_*   (ns_boundary_type; tripole = u-fold)              */
    int grid_tripole;           /* grid_type == 'tripole' (ice_dyn_evp.F90:416)                 */
    int tilt_mode;              /* 0: strtlt = -+fm*ocn (uncoupled / AusCOM .not.use_ocnslope,
                                      ice_dyn_shared.F90:598-599,605-608);
                                   1: -gravit*umass*ss_tlt (#ifdef coupled, :601-602)           */
    int strength_masked;        /* 1: keep strength only where icetmask==1 (kstrength=1 list
                                      semantics); 0: as given on physical cells (kstrength=0)   */
    int use_cuda_graph;         /* capture the ndte loop in a CUDA graph                        */
    double dt_init;             /* dt given to init_evp -> set_evp_parameters (:185-259)        */
    double cosw, sinw;          /* ocean turning angle (namelist under AusCOM)                  */
    double rhow, rhoi, rhos, gravit;
    double a_min, m_min;        /* ice_dyn_shared.F90:60-61                                     */
    double xmin;                /* revised EVP only: min(dxt, dyt) over ALL tmask cells of the
                                   global domain (ice_dyn_shared.F90:220-223); 0 = take it from this
                                   task's block (only valid when nranks == 1)                    */
    int slab_start[9];          /* first global column (0-based) of every rank's x-slab; all zero =
                                   equal slabs of nx_global/nranks columns.  Unequal slabs are the
                                   load balancing the reference does in ice_distribution
                                   (create_distrb_*, work per block from the land mask)            */
} evp_config_t;

/* Grid arrays (source/ice_grid.F90:48-77,112-115; fcor_blk ice_dyn_shared.F90:83-84),
 * copied to the device once. */
typedef struct {
    const double *dxt, *dyt, *dxhy, *dyhx, *cxp, *cyp, *cxm, *cym;
    const double *tarea, *uarea, *tarear, *uarear, *tinyarea, *fcor;
    const int *tmask, *umask;
    /* optional (NULL = not given): HTE, HTN (ice_grid.F90:51-52).  Where the metric terms above
     * are reproduced bit for bit from them (verified cell by cell at set_grid), the subcycle
     * kernel fetches HTE / HTN instead of the 8 (9) derived planes; results are identical. */
    const double *HTE, *HTN;
} evp_grid_t;

/* Per-call inputs (ice_state.F90:54-56,155; ice_flux.F90:37-48; ice_atmo.F90:40). */
typedef struct {
    const double *aice, *vice, *vsno, *aice_init;
    const double *strairxT, *strairyT, *uocn, *vocn;
    const double *ss_tltx, *ss_tlty;    /* read (and uploaded) only when tilt_mode == 1; may be NULL otherwise */
    const double *Cdn_ocn;
    const double *strength;     /* host ice_strength result on the block's physical T cells */
    /* optional (NULL = the default build).  #ifdef ACCESS with calc_strair = .false.
     * (source/ice_dyn_evp.F90:220-229): the coupler's wind stress, already on the U grid and
     * multiplied by aice; strairx/y become copies of these arrays (ghost cells included) and the
     * T->U averaging of strairxT/yT (:240-241) is not used.  Both or neither; the same choice on
     * every rank. */
    const double *strax, *stray;
} evp_in_t;

/* Optional: the thickness distribution, so that ice_strength (source/ice_mechred.F90:2111-2269, called by
 * evp at source/ice_dyn_evp.F90:291-301) runs on the device instead of on the host.  After
 * evp_gpu_upload_itd, evp_in_t.strength may be NULL.  aicen / vicen: (nx_block, ny_block, ncat). */
typedef struct {
    int ncat;                   /* ice_domain_size.F90 ncat (<= 16)                                   */
    int kstrength;              /* 1 Rothrock 75, 0 Hibler 79           (ice_mechred.F90:55-56)       */
    int krdg_partic;            /* 0 Thorndike 75, 1 exponential        (:57-58)                      */
    int krdg_redist;            /* 0 Hibler 80, 1 exponential           (:59-60)                      */
    double mu_rdg, Cf;          /* namelist (ice_init.F90:276-277: 3, 17)                             */
    const double *aicen, *vicen;
    const double *aice0;        /* open-water fraction (ice_state.F90)                                */
} evp_itd_t;

/* Persistent dynamic state, read and written (ice_state.F90:143-144; ice_flux.F90:95-107). */
typedef struct {
    double *uvel, *vvel;
    double *stressp[4], *stressm[4], *stress12[4];   /* _1.._4 */
    int *iceumask;
    double *fm;
} evp_inout_t;

/* Outputs; any pointer may be NULL to skip its download. */
typedef struct {
    double *strairx, *strairy, *strtltx, *strtlty, *strocnx, *strocny, *strintx, *strinty;
    double *strocnxT, *strocnyT, *divu, *shear, *rdg_conv, *rdg_shear, *prs_sig;
    double *uvel_init, *vvel_init, *tmass, *strength;
    int *icetmask;
} evp_out_t;

typedef struct {
    double prep_ms, loop_ms, finish_ms, total_ms;   /* device times of the last evp_gpu_compute */
    long long active_t_cells, active_u_cells;       /* icellt / icellu of this block            */
    long long kernel_launches;                      /* kernels launched by the last compute     */
    long long loop_kernel_launches;                 /* ... of which inside the ndte loop        */
} evp_stats_t;

int evp_gpu_init(const evp_config_t *cfg, evp_handle_t **h);
int evp_gpu_set_grid(evp_handle_t *h, const evp_grid_t *g);

/* per call, before evp_gpu_run / evp_gpu_upload when evp_in_t.strength is NULL */
int evp_gpu_upload_itd(evp_handle_t *h, const evp_itd_t *itd);

/* == one call of evp(dt): upload + compute + download */
int evp_gpu_run(evp_handle_t *h, double dt, const evp_in_t *in, evp_inout_t *io, evp_out_t *out);

/* The same in three stages (device-resident benchmarking; "next" row: resident state). */
int evp_gpu_upload(evp_handle_t *h, const evp_in_t *in, const evp_inout_t *io);
int evp_gpu_compute(evp_handle_t *h, double dt);
int evp_gpu_download(evp_handle_t *h, evp_inout_t *io, evp_out_t *out);

/* Resident-state protocol of evp_gpu_run.  The dynamic state (uvel, vvel, the 12 stresses,
 * iceumask) lives on the device between calls -- evp is its only writer in the reference's time step
 * (source/ice_step_mod.F90:1119-1141) -- so a call only has to move what changed on the host and what the host
 * reads next:
 *   resident = 1: evp_gpu_run uploads evp_in_t every call, evp_inout_t only on the first call and after
 *                 evp_gpu_invalidate_state (a restart was read, the host edited the state);
 *   download_mask: which groups evp_gpu_run copies back (pointers that are NULL are skipped as before):
 *     EVP_DL_VELOCITY  uvel, vvel                      -> transport_remap (ice_step_mod.F90:1126-1130)
 *     EVP_DL_DEFORM    divu, shear, rdg_conv, rdg_shear -> step_ridge / ridge_ice (:1139), history
 *     EVP_DL_OCN_T     strocnxT, strocnyT              -> coupler (ice_flux.F90:52-53), restart
 *     EVP_DL_STRESS    the 12 stresses                 -> restart files only (ice_restart_driver.F90:122-176)
 *     EVP_DL_HISTORY   everything else (strair*, strtlt*, strocnx/y, strint*, prs_sig, uvel_init, vvel_init,
 *                      tmass, strength, icetmask, fm, iceumask) -> history accumulation / diagnostics
 * evp_gpu_download(h, io, out) still returns every array at any time (history / restart steps).
 * Default: resident = 0, mask = EVP_DL_ALL: every array goes up and comes back, as in round 1. */
enum { EVP_DL_VELOCITY = 1, EVP_DL_STRESS = 2, EVP_DL_DEFORM = 4, EVP_DL_OCN_T = 8, EVP_DL_HISTORY = 16,
       EVP_DL_ALL = 31 };
int evp_gpu_set_resident(evp_handle_t *h, int resident, unsigned download_mask);
int evp_gpu_invalidate_state(evp_handle_t *h);

/* Restart hand-off on the resident state, after the host has read a restart file into
 * its arrays and uploaded them (evp_gpu_upload, or the next evp_gpu_run after evp_gpu_invalidate_state):
 * what restartfile does to the stresses once they are read (source/ice_restart_driver.F90:370-394: the 12
 * ice_HaloUpdate_stress calls of a tripole grid, same pairing as ice_dyn_evp.F90:416-481; :431-466: zero the
 * unused south ghost row and west ghost column of the 12 stresses) -- on the device, in place.  With several
 * ranks every rank calls it (it contains cross-rank halo exchanges, like evp_gpu_compute). */
int evp_gpu_restart_fixup(evp_handle_t *h);
/* the same for the ranks of a one-device group (see evp_gpu_compute_group) */
int evp_gpu_restart_fixup_group(evp_handle_t **handles, int n);

/* Horizontal transport, first stage: departure_points (source/ice_transport_remap.F90:1493-1670)
 * from the device-resident uvel, vvel of the last evp -- what horizontal_remap does first after evp in step_dynamics
 * (source/ice_step_mod.F90:1126-1130, ice_transport_remap.F90:541-550) -- followed by the ghost-cell updates of dpx, dpy
 * (ice_HaloUpdate, NE corner, vector; :574-578).  The velocities do not leave the device in between.
 *   set_transport_grid: dxu, dyu, HTN, HTE (source/ice_grid.F90:48-52) of this task's block, once.
 *   compute: dt = the transport time step; l_dp_midpt = the reference's parameter (ice_transport_driver.F90:51, .true.).
 *            With several ranks every rank calls it (it contains halo exchanges).  The results stay on the device
 *            (work planes that the next evp / principal_stress call reuses).
 *   fetch:   dpx, dpy (whole block incl. ghost cells; NULL = skip) and what the reference reports on failure: l_stop,
 *            and istop, jstop = block indices (1-based) of the last cell, in the reference's loop order, whose departure
 *            point lies more than one cell away.  On l_stop the reference aborts (abort_ice, :558); dpx, dpy then hold the
 *            first-order estimate on the physical cells and their ghost cells are not meaningful.
 *   evp_gpu_departure_points = compute + fetch. */
typedef struct {
    const double *dxu, *dyu, *HTN, *HTE;
} evp_transport_grid_t;
int evp_gpu_set_transport_grid(evp_handle_t *h, const evp_transport_grid_t *g);
int evp_gpu_departure_points_compute(evp_handle_t *h, double dt, int l_dp_midpt);
int evp_gpu_departure_points_group(evp_handle_t **handles, int n, double dt, int l_dp_midpt);
int evp_gpu_departure_points_fetch(evp_handle_t *h, double *dpx, double *dpy, int *l_stop, int *istop, int *jstop);
int evp_gpu_departure_points(evp_handle_t *h, double dt, int l_dp_midpt, double *dpx, double *dpy, int *l_stop, int *istop,
                             int *jstop);

/* Horizontal transport, field reconstruction: make_masks (source/ice_transport_remap.F90:867-1015),
 * construct_fields (:1024-1340) and limited_gradient (:1348-1484) for the open-water category 0 and the ice categories
 * 1..ncat, as horizontal_remap calls them per block (:462-511), followed by its ghost-cell updates of mc, mx, my, tc,
 * tx, ty (:579-584, :616-621; maskhalo_remap = .false., the default of ice_transport_driver.F90).
 *   set_transport_tracers: once (again when the tracer set changes).  tracer_type / depend / has_dependents are
 *            init_transport's arrays (source/ice_transport_driver.F90:82-118; depend 1-based; a type-2 tracer's parent
 *            is an earlier type-1 tracer flagged has_dependents, else EVP_ERR_ARG); hm, xav, yav, xxav, yyav the
 *            block's planes of ice_grid.F90 / ice_transport_remap.F90:300-340.  Allocates the device planes.
 *            ntrace <= EVP_REMAP_MAX_TRACE, at most EVP_REMAP_MAX_PARENTS tracers with dependents (EVP_ERR_TOO_LARGE).
 *   upload:  mm = aim (nx_block, ny_block, 0:ncat), tm = trm (nx_block, ny_block, ntrace, ncat) of this block with
 *            their ghost cells filled, as horizontal_remap receives them.
 *   compute / group: the reconstruction and the ghost-cell updates (every rank of the run calls it, like evp_gpu_run).
 *   fetch:   mc, mx, my (nx_block, ny_block, 0:ncat) and tc, tx, ty (nx_block, ny_block, ntrace, ncat); any may be NULL.
 *   evp_gpu_construct_fields = upload + compute + fetch. */
#define EVP_REMAP_MAX_TRACE 128
#define EVP_REMAP_MAX_PARENTS 16
int evp_gpu_set_transport_tracers(evp_handle_t *h, int ncat, int ntrace, const int *tracer_type, const int *depend,
                                  const int *has_dependents, const double *hm, const double *xav, const double *yav,
                                  const double *xxav, const double *yyav);
int evp_gpu_construct_fields_upload(evp_handle_t *h, const double *mm, const double *tm);
int evp_gpu_construct_fields_compute(evp_handle_t *h);
int evp_gpu_construct_fields_group(evp_handle_t **handles, int n);
int evp_gpu_construct_fields_fetch(evp_handle_t *h, double *mc, double *mx, double *my, double *tc, double *tx, double *ty);
int evp_gpu_construct_fields(evp_handle_t *h, const double *mm, const double *tm, double *mc, double *mx, double *my,
                             double *tc, double *tx, double *ty);
/* All of horizontal_remap (source/ice_transport_remap.F90:309-850) with the reference's compile-time settings l_fixed_area
 * = .false., integral_order = 3, l_dp_midpt = .true. (source/ice_transport_driver.F90:31,47,51): the reconstruction and the
 * departure points above with their ghost-cell updates, then per edge (east, north) locate_triangles (:1680-3047),
 * triangle_coordinates (:3078-3187) and transport_integrals (:3199-3509) of every category, and update_fields (:3517-3729)
 * of open water and of every ice category.  Needs set_grid (tarear), set_transport_grid, set_transport_tracers, the
 * velocities of the last evp resident, and mm / tm uploaded (construct_fields_upload).
 *   compute / group: every rank of the run calls it.
 *   fetch:   mm (nx_block, ny_block, 0:ncat), tm (nx_block, ny_block, ntrace, ncat) after the remap (physical cells; the
 *            caller's ice_HaloUpdate of the state follows as in transport_remap, ice_transport_driver.F90:490-520), and
 *            what the reference would print before abort_ice: *stop = 0 completed, 1 bad departure points (:552-563),
 *            2 negative open-water area (:783-795), 3 negative ice area of category *stop_cat (:797-825); istop, jstop the
 *            block indices it reports.  After a stop the arrays are not meaningful (the reference aborts).
 *   evp_gpu_horizontal_remap = construct_fields_upload + compute + fetch, mm and tm in place. */
int evp_gpu_horizontal_remap_compute(evp_handle_t *h, double dt);
int evp_gpu_horizontal_remap_group(evp_handle_t **handles, int n, double dt);
int evp_gpu_horizontal_remap_fetch(evp_handle_t *h, double *mm, double *tm, int *stop, int *stop_cat, int *istop, int *jstop);
int evp_gpu_horizontal_remap(evp_handle_t *h, double dt, double *mm, double *tm, int *stop, int *stop_cat, int *istop,
                             int *jstop);
/* The caller of horizontal_remap, transport_remap (source/ice_transport_driver.F90:198-627) with its compile-time settings
 * l_conservation_check = l_monotonicity_check = .false. (:249-251): state_to_tracers (:789-900), horizontal_remap,
 * tracers_to_state (:908-1002) and bound_state (source/ice_state.F90:173-238, maskhalo_bound = .false.) -- what
 * step_dynamics calls after evp (source/ice_step_mod.F90:1126-1130).  The tracer table given to set_transport_tracers
 * starts with hice, hsno (ntrace = 2 + ntrcr, init_transport :90-110).
 *   upload:  aice0 (nx_block, ny_block), aicen / vicen / vsnon (nx_block, ny_block, ncat), trcrn (nx_block, ny_block,
 *            ntrcr, ncat) of this block, ghost cells filled.
 *   compute / group: nt_qsno (1-based), nslyr, rhos, Lfresh: ice_state's / ice_constants' values (snow enthalpy is
 *            transported as trcrn + rhos * Lfresh, :880-886).  Every rank of the run calls it.
 *   fetch:   the transported state, ghost cells updated by bound_state; cells whose new area is not positive keep their
 *            old aicen, vicen, vsnon, trcrn as in the reference (:963-975).  stop etc. as evp_gpu_horizontal_remap_fetch.
 *   evp_gpu_transport_remap = upload + compute + fetch, in place. */
int evp_gpu_transport_remap_upload(evp_handle_t *h, const double *aice0, const double *aicen, const double *trcrn,
                                   const double *vicen, const double *vsnon);
int evp_gpu_transport_remap_compute(evp_handle_t *h, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh);
int evp_gpu_transport_remap_group(evp_handle_t **handles, int n, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh);
int evp_gpu_transport_remap_fetch(evp_handle_t *h, double *aice0, double *aicen, double *trcrn, double *vicen, double *vsnon,
                                  int *stop, int *stop_cat, int *istop, int *jstop);
int evp_gpu_transport_remap(evp_handle_t *h, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh, double *aice0,
                            double *aicen, double *trcrn, double *vicen, double *vsnon, int *stop, int *stop_cat, int *istop,
                            int *jstop);
/* Device time (CUDA events, ms) of the last call of each transport stage on this handle: ms[0] departure_points kernels,
 * ms[1] their two ghost updates, ms[2] the reconstruction kernel, ms[3] its ghost updates, ms[4] the edge fluxes and
 * update_fields of the last horizontal_remap, ms[5] that whole horizontal_remap.  What the reference's timer_advect /
 * timer_bound would show for these stages (source/ice_timers.F90); bench.py reads it. */
int evp_gpu_get_transport_timing(evp_handle_t *h, double ms[6]);

/* Several ranks driven by one host thread (handles of ranks 0..n-1 on ONE device, e.g. tests with
 * fewer GPUs than ranks): steps are interleaved on one stream so that no kernel ever waits. */
int evp_gpu_compute_group(evp_handle_t **handles, int n, double dt);

int evp_gpu_get_stats(evp_handle_t *h, evp_stats_t *st);
int evp_gpu_synchronize(evp_handle_t *h);

/* Multi-GPU (one task per GPU): every task exports a blob, the host gathers the blobs of all
 * tasks in rank order (MPI_Allgather in the shim) and hands them back. */
/* principal_stress (source/ice_dyn_shared.F90:853-893; called by the history writer,
 * source/ice_history.F90:1921) from the device-resident stressp_1, stressm_1, stress12_1, prs_sig of
 * the last evp: sig1, sig2 over the whole block, spval_dbl = 1e30 where prs_sig <= puny. */
int evp_gpu_principal_stress(evp_handle_t *h, double *sig1, double *sig2);

/* EAP, the elastic-anisotropic-plastic rheology (kdyn = 2): `eap(dt)`, source/ice_dyn_eap.F90:66-486,
 * called by step_dynamics in place of evp (source/ice_step_mod.F90:1119-1124).  It shares evp_prep1 / evp_prep2 / stepu /
 * evp_finish and the halo pattern with evp, so a handle initialised with kdyn = 2 is driven by the same evp_gpu_run /
 * upload / compute / download calls with the same arrays; stress_eap (:1052-1467, update_stress_rdg :1474-1658) replaces
 * stress, stepa (:1664-1787, calc_ffrac :1795-1864) runs on every 10th subcycle (:408), and there is no
 * ice_HaloUpdate_stress.  What is new at the boundary are the module variables of ice_dyn_eap (:31-57):
 *   tables   s11r ... s22s (nx_yield, ny_yield, na_yield) = (41, 41, 21), Fortran order, as init_eap computed them on the
 *            host (:557-618).  Once, before the first call; they are an input, the device does not recompute them.
 *   state    a11_1..4, a12_1..4 (restart fields, read_restart_eap :1908) and the cell averages a11, a12 (history): device
 *            resident, initialised as init_eap does (a11_k = 0.5, everything else 0, :531-548).  upload after a restart
 *            was read, download when a restart / history file is written; NULL pointers are skipped.
 *   out      e11, e12, e22, s11, s12, s22, yieldstress11 / 12 / 22 (history), zeroed at the start of every call
 *            (:166-182).  rdg_shear stays 0 under EAP (its statement is commented out, :1283-1285), prs_sig = strength.
 * The intrinsics atan2 / sin / cos of update_stress_rdg and calc_ffrac are cice5_b200/csrc/evp_trig.h (error < 1.5 ulp): a
 * Fortran processor is free in their last bit, and the table look-up (`int((x-piq-pi)*invdx)`, :1599-1601) can turn a
 * last-bit difference into the neighbouring table entry.  Every entry point returns EVP_ERR_STATE on a kdyn = 1 handle. */
typedef struct {
    const double *s11r, *s12r, *s22r, *s11s, *s12s, *s22s;
} evp_eap_tables_t;
typedef struct {
    double *a11[4], *a12[4];      /* a11_1..4, a12_1..4: corners ne, nw, sw, se                    */
    double *a11_avg, *a12_avg;    /* a11, a12                                                       */
} evp_eap_state_t;
typedef struct {
    double *e11, *e12, *e22, *s11, *s12, *s22, *yieldstress11, *yieldstress12, *yieldstress22;
} evp_eap_out_t;
int evp_gpu_set_eap_tables(evp_handle_t *h, const evp_eap_tables_t *t);
int evp_gpu_upload_eap(evp_handle_t *h, const evp_eap_state_t *st);
int evp_gpu_download_eap(evp_handle_t *h, evp_eap_state_t *st, evp_eap_out_t *out);

/* No counterpart in the reference.  principal_stress (the same kernel
 * as evp_gpu_principal_stress) for callers that must not allocate -- code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc.  No handle: the four inputs (stressp_1, stressm_1, stress12_1, prs_sig) and
 * the two outputs are HOST arrays of nx_block x ny_block doubles; the device scratch (inputs + outputs, 6 planes) is a
 * file-scope `static __device__` array of EVP_STATIC_MAX_CELLS cells per plane (6 MiB in all) that every GPU which
 * loads the library holds until exit.  The call makes no cudaMalloc / cudaFree / allocating library call; it runs on
 * the default stream of `device` and is complete on return.  nx_block * ny_block > EVP_STATIC_MAX_CELLS ->
 * EVP_ERR_TOO_LARGE.  All calls share the array, which keeps the last call's contents (every element a call reads is
 * written by that call first): calls that overlap on one GPU race -- serialise them. */
#define EVP_STATIC_MAX_CELLS 131072      /* e.g. gx3 102 x 118, gx1 322 x 386, 1 deg tripole 362 x 302 */
int evp_gpu_principal_stress_static(int device, int nx_block, int ny_block, const double *stressp_1,
                                    const double *stressm_1, const double *stress12_1, const double *prs_sig,
                                    double *sig1, double *sig2);

/* Diagnostics (library started with EVP_B200_TRACE=1): per subcycle 8 x u64 device timestamps in
 * ns: [0] subcycle kernel start, [1] its last CTA done, [2] peers signalled, [3] fold wait begin,
 * [4] fold wait end.  No reference counterpart (the reference's timers are ice_timers.F90). */
int evp_gpu_get_trace(evp_handle_t *h, unsigned long long *buf, int n_u64);

int evp_gpu_peer_blob_size(void);
int evp_gpu_get_peer_blob(evp_handle_t *h, void *blob);
int evp_gpu_set_peers(evp_handle_t *h, const void *blobs_all_ranks);

/* Optional, for hosts whose arrays keep their addresses (the reference's module arrays are static:
 * source/ice_state.F90, ice_flux.F90): page-lock every non-NULL array of the three structs once, so that
 * the per-call copies of evp_gpu_run run at the full host-link rate and asynchronously instead of through the
 * driver's staging buffer.  Arrays that are already page-locked are left alone.  The registrations are
 * released by evp_gpu_finalize; the arrays must outlive the handle.  Any argument may be NULL. */
int evp_gpu_pin_host_arrays(evp_handle_t *h, const evp_in_t *in, const evp_inout_t *io, const evp_out_t *out);

int evp_gpu_finalize(evp_handle_t *h);
const char *evp_gpu_strerror(int status);
const char *evp_gpu_last_cuda_error(void);
int evp_gpu_device_count(void);

#ifdef __cplusplus
}
#endif
#endif
