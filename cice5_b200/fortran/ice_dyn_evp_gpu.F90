!=======================================================================
! ice_dyn_evp_gpu.F90 -- replacement for source/ice_dyn_evp.F90 that forwards `call evp(dt)`
! (source/ice_step_mod.F90:1119) to the B200 library through include/evp_b200.h.
!
! Same module name, same public entry, same module arrays as the reference
! (source/ice_dyn_evp.F90:47,68,70-92); build with this file INSTEAD of ice_dyn_evp.F90 and link
! -levp_b200_strict -lcudart.  One MPI task per GPU, one block per task (MXBLCKS=1,
! BLCKX = NXGLOB/ntasks, BLCKY = NYGLOB, distribution_type='cartesian', processor_shape
! 'slenderX1'), as solo_ice_comp:84-86 sets up for 8 tasks.
!
! NOTE: written against the C ABI; NOT compiled in the build image (no Fortran compiler there).
!=======================================================================
      module ice_dyn_evp

      use, intrinsic :: iso_c_binding
      use ice_kinds_mod
      use ice_dyn_shared    ! ndte, kdyn, revised_evp, cosw, sinw, uvel_init, vvel_init, fcor_blk, a_min, m_min
#ifdef AusCOM
      use cpl_parameters, only: use_ocnslope
      use cpl_arrays_setup, only: sicemass
#endif

      implicit none
      private
      public :: evp
      save

      type, bind(C) :: evp_config_t
         integer(c_int) :: nx_global, ny_global, nx_block, ny_block, rank, nranks, device
         integer(c_int) :: ndte, kdyn, revised_evp, ew_boundary, ns_boundary, grid_tripole
         integer(c_int) :: tilt_mode, strength_masked, use_cuda_graph
         real(c_double) :: dt_init, cosw, sinw, rhow, rhoi, rhos, gravit, a_min, m_min
         real(c_double) :: xmin              ! revised EVP: global min(dxt, dyt) over tmask cells
         integer(c_int) :: slab_start(9)     ! all zero: equal x-slabs (CICE blocks have one size)
      end type

      type, bind(C) :: evp_grid_t
         type(c_ptr) :: dxt, dyt, dxhy, dyhx, cxp, cyp, cxm, cym
         type(c_ptr) :: tarea, uarea, tarear, uarear, tinyarea, fcor
         type(c_ptr) :: tmask, umask
         type(c_ptr) :: HTE, HTN             ! optional (c_null_ptr): compact metric rows
      end type

      type, bind(C) :: evp_in_t
         type(c_ptr) :: aice, vice, vsno, aice_init, strairxT, strairyT, uocn, vocn
         type(c_ptr) :: ss_tltx, ss_tlty, Cdn_ocn, strength
         type(c_ptr) :: strax, stray         ! optional (c_null_ptr): ACCESS, .not. calc_strair
      end type

      type, bind(C) :: evp_inout_t
         type(c_ptr) :: uvel, vvel
         type(c_ptr) :: stressp(4), stressm(4), stress12(4)
         type(c_ptr) :: iceumask, fm
      end type

      type, bind(C) :: evp_out_t
         type(c_ptr) :: strairx, strairy, strtltx, strtlty, strocnx, strocny, strintx, strinty
         type(c_ptr) :: strocnxT, strocnyT, divu, shear, rdg_conv, rdg_shear, prs_sig
         type(c_ptr) :: uvel_init, vvel_init, tmass, strength, icetmask
      end type

      interface
         integer(c_int) function evp_gpu_init(cfg, h) bind(C, name='evp_gpu_init')
            import :: c_int, c_ptr, evp_config_t
            type(evp_config_t), intent(in) :: cfg
            type(c_ptr), intent(out) :: h
         end function
         integer(c_int) function evp_gpu_set_grid(h, g) bind(C, name='evp_gpu_set_grid')
            import :: c_int, c_ptr, evp_grid_t
            type(c_ptr), value :: h
            type(evp_grid_t), intent(in) :: g
         end function
         integer(c_int) function evp_gpu_run(h, dt, in, io, out) bind(C, name='evp_gpu_run')
            import :: c_int, c_ptr, c_double, evp_in_t, evp_inout_t, evp_out_t
            type(c_ptr), value :: h
            real(c_double), value :: dt
            type(evp_in_t), intent(in) :: in
            type(evp_inout_t), intent(in) :: io
            type(evp_out_t), intent(in) :: out
         end function
         integer(c_int) function evp_gpu_pin_host_arrays(h, in, io, out) bind(C, name='evp_gpu_pin_host_arrays')
            import :: c_int, c_ptr, evp_in_t, evp_inout_t, evp_out_t
            type(c_ptr), value :: h
            type(evp_in_t), intent(in) :: in
            type(evp_inout_t), intent(in) :: io
            type(evp_out_t), intent(in) :: out
         end function
         integer(c_int) function evp_gpu_peer_blob_size() bind(C, name='evp_gpu_peer_blob_size')
            import :: c_int
         end function
         integer(c_int) function evp_gpu_get_peer_blob(h, blob) bind(C, name='evp_gpu_get_peer_blob')
            import :: c_int, c_ptr, c_char
            type(c_ptr), value :: h
            character(kind=c_char), intent(out) :: blob(*)
         end function
         integer(c_int) function evp_gpu_set_peers(h, blobs) bind(C, name='evp_gpu_set_peers')
            import :: c_int, c_ptr, c_char
            type(c_ptr), value :: h
            character(kind=c_char), intent(in) :: blobs(*)
         end function
         integer(c_int) function evp_gpu_device_count() bind(C, name='evp_gpu_device_count')
            import :: c_int
         end function
         type(c_ptr) function evp_gpu_strerror(rc) bind(C, name='evp_gpu_strerror')
            import :: c_int, c_ptr
            integer(c_int), value :: rc
         end function
      end interface

      type(c_ptr) :: handle = c_null_ptr
      logical :: ready = .false.
      logical :: host_pinned = .false.
      ! logical masks cross the ABI as 0/1 integers (Fortran .true. is compiler dependent)
      integer(c_int), allocatable, target :: tmask_i(:,:), umask_i(:,:), iceumask_i(:,:), icetmask_i(:,:)
      real(c_double), allocatable, target :: tmass_w(:,:)

      contains

!=======================================================================
      subroutine check(rc, where)
      use ice_exit, only: abort_ice
      integer(c_int), intent(in) :: rc
      character(len=*), intent(in) :: where
      character(kind=c_char), pointer :: txt(:)
      character(len=256) :: msg
      integer :: n
      if (rc == 0) return
      call c_f_pointer(evp_gpu_strerror(rc), txt, [256])
      msg = ' '
      do n = 1, 255
         if (txt(n) == c_null_char) exit
         msg(n:n) = txt(n)
      enddo
      call abort_ice('evp (B200): '//where//': '//trim(msg))      ! mpi/ice_exit.F90
      end subroutine check

!=======================================================================
! once, on the first call: after init_grid2 and init_evp have run
      subroutine evp_gpu_setup (dt)
      use ice_blocks, only: nx_block, ny_block
      use ice_communicate, only: my_task, MPI_COMM_ICE, get_num_procs
      use ice_constants, only: rhow, rhoi, rhos, gravit
      use ice_domain, only: nblocks, ew_boundary_type, ns_boundary_type, distrb_info
      use ice_global_reductions, only: global_minval
      use ice_domain_size, only: nx_global, ny_global
      use ice_exit, only: abort_ice
      use ice_grid, only: tmask, umask, dxt, dyt, dxhy, dyhx, cxp, cyp, cxm, cym, &
          tarea, uarea, tarear, uarear, tinyarea, grid_type, HTE, HTN
      include 'mpif.h'
      real (kind=dbl_kind), intent(in) :: dt
      type(evp_config_t) :: cfg
      type(evp_grid_t), target :: g
      integer :: ntasks, nb, ierr, ndev
      character(kind=c_char), allocatable :: mine(:), everyone(:)

      if (nblocks /= 1) call abort_ice('evp (B200): needs exactly one block per task (MXBLCKS=1)')
      ntasks = get_num_procs()
      ndev = evp_gpu_device_count()
      if (ndev < 1) call abort_ice('evp (B200): no CUDA device; there is no CPU fallback')
      cfg%nx_global = nx_global;  cfg%ny_global = ny_global
      cfg%nx_block = nx_block;    cfg%ny_block = ny_block
      cfg%rank = my_task;         cfg%nranks = ntasks
      cfg%device = mod(my_task, ndev)
      cfg%ndte = ndte;            cfg%kdyn = kdyn
      cfg%revised_evp = merge(1, 0, revised_evp)
      cfg%ew_boundary = merge(1, 0, trim(ew_boundary_type) == 'cyclic')
      select case (trim(ns_boundary_type))
         case ('open');    cfg%ns_boundary = 0
         case ('closed');  cfg%ns_boundary = 1
         case ('tripole'); cfg%ns_boundary = 2
         case default;     call abort_ice('evp (B200): unsupported ns_boundary_type')
      end select
      cfg%grid_tripole = merge(1, 0, trim(grid_type) == 'tripole')
#if defined(coupled)
      cfg%tilt_mode = 1
#else
      cfg%tilt_mode = 0
#endif
#ifdef AusCOM
      if (.not. use_ocnslope) cfg%tilt_mode = 0          ! ice_dyn_shared.F90:604-609
#endif
      cfg%strength_masked = 0       ! strength is passed exactly as ice_strength left it
      cfg%use_cuda_graph = 1
      cfg%dt_init = dt              ! == dt_dyn handed to init_evp (CICE_InitMod.F90:104)
      cfg%cosw = cosw;  cfg%sinw = sinw
      cfg%rhow = rhow;  cfg%rhoi = rhoi;  cfg%rhos = rhos;  cfg%gravit = gravit
      cfg%a_min = a_min;  cfg%m_min = m_min
      ! xmin exactly as set_evp_parameters computes it (ice_dyn_shared.F90:220-223)
      cfg%xmin = min(global_minval(dxt, distrb_info, tmask), global_minval(dyt, distrb_info, tmask))
      cfg%slab_start = 0            ! blocks of one size: task r owns columns r*block_size_x+1 ...
      call check(evp_gpu_init(cfg, handle), 'init')

      allocate(tmask_i(nx_block,ny_block), umask_i(nx_block,ny_block), &
               iceumask_i(nx_block,ny_block), icetmask_i(nx_block,ny_block), tmass_w(nx_block,ny_block))
      tmask_i = merge(1, 0, tmask(:,:,1));  umask_i = merge(1, 0, umask(:,:,1))
      g%dxt = c_loc(dxt(1,1,1));   g%dyt = c_loc(dyt(1,1,1))
      g%dxhy = c_loc(dxhy(1,1,1)); g%dyhx = c_loc(dyhx(1,1,1))
      g%cxp = c_loc(cxp(1,1,1));   g%cyp = c_loc(cyp(1,1,1))
      g%cxm = c_loc(cxm(1,1,1));   g%cym = c_loc(cym(1,1,1))
      g%tarea = c_loc(tarea(1,1,1));   g%uarea = c_loc(uarea(1,1,1))
      g%tarear = c_loc(tarear(1,1,1)); g%uarear = c_loc(uarear(1,1,1))
      g%tinyarea = c_loc(tinyarea(1,1,1)); g%fcor = c_loc(fcor_blk(1,1,1))
      g%tmask = c_loc(tmask_i(1,1));   g%umask = c_loc(umask_i(1,1))
      ! HTE, HTN (ice_grid.F90:51-52): where the library verifies, cell by cell, that the metric
      ! arrays above follow from them bit for bit it fetches these two instead of nine planes
      g%HTE = c_loc(HTE(1,1,1));       g%HTN = c_loc(HTN(1,1,1))
      call check(evp_gpu_set_grid(handle, g), 'set_grid')

      if (ntasks > 1) then          ! exchange the peer-memory handles of all tasks of the node
         nb = evp_gpu_peer_blob_size()
         allocate(mine(nb), everyone(nb*ntasks))
         call check(evp_gpu_get_peer_blob(handle, mine), 'get_peer_blob')
         call MPI_ALLGATHER(mine, nb, MPI_CHARACTER, everyone, nb, MPI_CHARACTER, MPI_COMM_ICE, ierr)
         call check(evp_gpu_set_peers(handle, everyone), 'set_peers')
         deallocate(mine, everyone)
      endif
      ready = .true.
      end subroutine evp_gpu_setup

!=======================================================================
! == subroutine evp(dt) of the reference (source/ice_dyn_evp.F90:68)
      subroutine evp (dt)
      use ice_atmo, only: Cdn_ocn
#ifdef ACCESS
      use ice_atmo, only: calc_strair
      use ice_flux, only: strax, stray
#endif
      use ice_blocks, only: block, get_block, nx_block, ny_block
      use ice_domain, only: blocks_ice
      use ice_flux, only: rdg_conv, rdg_shear, prs_sig, strairxT, strairyT, &
          strairx, strairy, uocn, vocn, ss_tltx, ss_tlty, iceumask, fm, &
          strtltx, strtlty, strocnx, strocny, strintx, strinty, strocnxT, strocnyT, &
          stressp_1, stressp_2, stressp_3, stressp_4, stressm_1, stressm_2, stressm_3, stressm_4, &
          stress12_1, stress12_2, stress12_3, stress12_4
      use ice_mechred, only: ice_strength
      use ice_state, only: aice, vice, vsno, uvel, vvel, divu, shear, &
          aice_init, aice0, aicen, vicen, strength
      use ice_timers, only: timer_dynamics, ice_timer_start, ice_timer_stop
      real (kind=dbl_kind), intent(in) :: dt
      type(evp_in_t) :: in
      type(evp_inout_t) :: io
      type(evp_out_t) :: out
      type(block) :: this_block
      integer :: i, j, icellt
      integer, allocatable :: indxti(:), indxtj(:)

      call ice_timer_start(timer_dynamics)
      if (.not. ready) call evp_gpu_setup(dt)

      ! ice_strength stays on the host (SURVEY 8a row a5): evaluate it on every ocean T cell of
      ! the block; the device keeps it where the reference's icellt list would have used it
      this_block = get_block(blocks_ice(1), 1)
      allocate(indxti(nx_block*ny_block), indxtj(nx_block*ny_block))
      icellt = 0
      do j = this_block%jlo, this_block%jhi
      do i = this_block%ilo, this_block%ihi
         icellt = icellt + 1;  indxti(icellt) = i;  indxtj(icellt) = j
      enddo
      enddo
      call ice_strength(nx_block, ny_block, this_block%ilo, this_block%ihi, this_block%jlo, this_block%jhi, &
                        icellt, indxti, indxtj, aice(:,:,1), vice(:,:,1), aice0(:,:,1), &
                        aicen(:,:,:,1), vicen(:,:,:,1), strength(:,:,1))
      deallocate(indxti, indxtj)

      iceumask_i = merge(1, 0, iceumask(:,:,1))
      in%aice = c_loc(aice(1,1,1)); in%vice = c_loc(vice(1,1,1)); in%vsno = c_loc(vsno(1,1,1))
      in%aice_init = c_loc(aice_init(1,1,1))
      in%strairxT = c_loc(strairxT(1,1,1)); in%strairyT = c_loc(strairyT(1,1,1))
      in%uocn = c_loc(uocn(1,1,1)); in%vocn = c_loc(vocn(1,1,1))
      in%ss_tltx = c_loc(ss_tltx(1,1,1)); in%ss_tlty = c_loc(ss_tlty(1,1,1))
      in%Cdn_ocn = c_loc(Cdn_ocn(1,1,1)); in%strength = c_loc(strength(1,1,1))
      in%strax = c_null_ptr;  in%stray = c_null_ptr
#ifdef ACCESS
      ! source/ice_dyn_evp.F90:220-229: the coupler's wind stress is already on the U grid
      if (.not. calc_strair) then
         in%strax = c_loc(strax(1,1,1));  in%stray = c_loc(stray(1,1,1))
      endif
#endif
      io%uvel = c_loc(uvel(1,1,1)); io%vvel = c_loc(vvel(1,1,1))
      io%stressp = [c_loc(stressp_1(1,1,1)), c_loc(stressp_2(1,1,1)), c_loc(stressp_3(1,1,1)), c_loc(stressp_4(1,1,1))]
      io%stressm = [c_loc(stressm_1(1,1,1)), c_loc(stressm_2(1,1,1)), c_loc(stressm_3(1,1,1)), c_loc(stressm_4(1,1,1))]
      io%stress12 = [c_loc(stress12_1(1,1,1)), c_loc(stress12_2(1,1,1)), c_loc(stress12_3(1,1,1)), c_loc(stress12_4(1,1,1))]
      io%iceumask = c_loc(iceumask_i(1,1)); io%fm = c_loc(fm(1,1,1))
      out%strairx = c_loc(strairx(1,1,1)); out%strairy = c_loc(strairy(1,1,1))
      out%strtltx = c_loc(strtltx(1,1,1)); out%strtlty = c_loc(strtlty(1,1,1))
      out%strocnx = c_loc(strocnx(1,1,1)); out%strocny = c_loc(strocny(1,1,1))
      out%strintx = c_loc(strintx(1,1,1)); out%strinty = c_loc(strinty(1,1,1))
      out%strocnxT = c_loc(strocnxT(1,1,1)); out%strocnyT = c_loc(strocnyT(1,1,1))
      out%divu = c_loc(divu(1,1,1)); out%shear = c_loc(shear(1,1,1))
      out%rdg_conv = c_loc(rdg_conv(1,1,1)); out%rdg_shear = c_loc(rdg_shear(1,1,1))
      out%prs_sig = c_loc(prs_sig(1,1,1))
      out%uvel_init = c_loc(uvel_init(1,1,1)); out%vvel_init = c_loc(vvel_init(1,1,1))
      out%tmass = c_loc(tmass_w(1,1)); out%strength = c_loc(strength(1,1,1))
      out%icetmask = c_loc(icetmask_i(1,1))

      ! the module arrays are static: page-lock them once, the per-call copies then run at the full
      ! host-link rate instead of through the driver's staging buffer
      if (.not. host_pinned) then
         call check(evp_gpu_pin_host_arrays(handle, in, io, out), 'pin_host_arrays')
         host_pinned = .true.
      endif

      call check(evp_gpu_run(handle, real(dt, c_double), in, io, out), 'run')

      iceumask(:,:,1) = iceumask_i /= 0
#ifdef AusCOM
      sicemass(:,:,1) = tmass_w          ! ice_dyn_evp.F90:205-207
#endif
      call ice_timer_stop(timer_dynamics)
      end subroutine evp

      end module ice_dyn_evp
