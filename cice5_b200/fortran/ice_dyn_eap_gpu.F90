!=======================================================================
! ice_dyn_eap_gpu.F90 -- the EAP side of the binding (kdyn = 2): what replaces `subroutine eap (dt)` in
! source/ice_dyn_eap.F90 (:66-486) when the build uses ice_dyn_evp_gpu.F90.
!
! A FRAGMENT, not a module: paste it over `subroutine eap` inside the reference's module ice_dyn_eap.  Everything
! else of that module stays the reference's and keeps running on the host: init_eap (the lookup tables s11r ... s22s,
! :557-618, and the initial structure tensor), write_restart_eap / read_restart_eap (:1872-2010), the module arrays
! (:31-57) -- add `, target` to the declarations of s11r ... s22s, a11_1 ... a12_4, a11, a12, e11 ... s22 and
! yieldstress11 / 12 / 22 (they are this module's own arrays; c_loc needs it).  stress_eap, update_stress_rdg, stepa
! and calc_ffrac are no longer called.
!
! A handle whose configuration says kdyn = 2 runs eap inside evp_gpu_run (include/evp_b200.h): evp_prep1 /
! evp_prep2 / stepu / evp_finish and the halo pattern are shared with evp (ice_dyn_eap.F90:195-482), so the call
! itself is the evp shim's -- same arrays, same resident-state protocol.  New here: the tables (once), the structure
! tensor (up after a restart was read, down before one is written) and the EAP history fields (down on request).
!
! NOTE: written against the C ABI; NOT compiled in the build image (no Fortran compiler there).
!=======================================================================
      subroutine eap (dt)
      use, intrinsic :: iso_c_binding
      use ice_dyn_evp, only: evp, evp_gpu_get_handle          ! ice_dyn_evp_gpu.F90
      real (kind=dbl_kind), intent(in) :: dt

      if (.not. eap_ready) then
         eap_handle = evp_gpu_get_handle(dt)                   ! init + set_grid + peers, cfg%kdyn = kdyn = 2
         eap_tab%s11r = c_loc(s11r);  eap_tab%s12r = c_loc(s12r);  eap_tab%s22r = c_loc(s22r)
         eap_tab%s11s = c_loc(s11s);  eap_tab%s12s = c_loc(s12s);  eap_tab%s22s = c_loc(s22s)
         call eap_check(evp_gpu_set_eap_tables(eap_handle, eap_tab), 'set_eap_tables')
         eap_st%a11(1) = c_loc(a11_1);  eap_st%a11(2) = c_loc(a11_2)
         eap_st%a11(3) = c_loc(a11_3);  eap_st%a11(4) = c_loc(a11_4)
         eap_st%a12(1) = c_loc(a12_1);  eap_st%a12(2) = c_loc(a12_2)
         eap_st%a12(3) = c_loc(a12_3);  eap_st%a12(4) = c_loc(a12_4)
         eap_st%a11_avg = c_loc(a11);   eap_st%a12_avg = c_loc(a12)
         eap_out%e11 = c_loc(e11);  eap_out%e12 = c_loc(e12);  eap_out%e22 = c_loc(e22)
         eap_out%s11 = c_loc(s11);  eap_out%s12 = c_loc(s12);  eap_out%s22 = c_loc(s22)
         eap_out%yieldstress11 = c_loc(yieldstress11)
         eap_out%yieldstress12 = c_loc(yieldstress12)
         eap_out%yieldstress22 = c_loc(yieldstress22)
         ! init_eap or read_restart_eap (runtype) has filled a11_1 ... a12_4 by now
         call eap_check(evp_gpu_upload_eap(eap_handle, eap_st), 'upload_eap')
         eap_ready = .true.
      endif
      call evp (dt)       ! the shim's evp: ice_strength on the host, upload, evp_gpu_run (= eap on the device), download
      end subroutine eap

!=======================================================================
! before write_restart_eap (:1872) and before the history writer reads a11, a12, e11 ... yieldstress22
      subroutine eap_gpu_fetch_all
      if (eap_ready) call eap_check(evp_gpu_download_eap(eap_handle, eap_st, eap_out), 'download_eap')
      end subroutine eap_gpu_fetch_all

! after read_restart_eap (:1908) changed a11_1 ... a12_4 on the host
      subroutine eap_gpu_state_changed
      if (eap_ready) call eap_check(evp_gpu_upload_eap(eap_handle, eap_st), 'upload_eap')
      end subroutine eap_gpu_state_changed

      subroutine eap_check(rc, where)
      use ice_exit, only: abort_ice
      integer(c_int), intent(in) :: rc
      character(len=*), intent(in) :: where
      if (rc /= 0) call abort_ice('eap: evp_gpu_'//where//' failed (see evp_gpu_strerror)')
      end subroutine eap_check

!=======================================================================
! ... and in the specification part of module ice_dyn_eap:
!
!     use, intrinsic :: iso_c_binding
!     public :: eap_gpu_fetch_all, eap_gpu_state_changed
!     type, bind(C) :: evp_eap_tables_t
!        type(c_ptr) :: s11r, s12r, s22r, s11s, s12s, s22s
!     end type
!     type, bind(C) :: evp_eap_state_t
!        type(c_ptr) :: a11(4), a12(4), a11_avg, a12_avg
!     end type
!     type, bind(C) :: evp_eap_out_t
!        type(c_ptr) :: e11, e12, e22, s11, s12, s22, yieldstress11, yieldstress12, yieldstress22
!     end type
!     interface
!        integer(c_int) function evp_gpu_set_eap_tables(h, t) bind(C, name='evp_gpu_set_eap_tables')
!           import :: c_int, c_ptr, evp_eap_tables_t
!           type(c_ptr), value :: h
!           type(evp_eap_tables_t), intent(in) :: t
!        end function
!        integer(c_int) function evp_gpu_upload_eap(h, st) bind(C, name='evp_gpu_upload_eap')
!           import :: c_int, c_ptr, evp_eap_state_t
!           type(c_ptr), value :: h
!           type(evp_eap_state_t), intent(in) :: st
!        end function
!        integer(c_int) function evp_gpu_download_eap(h, st, out) bind(C, name='evp_gpu_download_eap')
!           import :: c_int, c_ptr, evp_eap_state_t, evp_eap_out_t
!           type(c_ptr), value :: h
!           type(evp_eap_state_t), intent(in) :: st
!           type(evp_eap_out_t), intent(in) :: out
!        end function
!     end interface
!     type(c_ptr) :: eap_handle = c_null_ptr
!     logical :: eap_ready = .false.
!     type(evp_eap_tables_t) :: eap_tab
!     type(evp_eap_state_t) :: eap_st
!     type(evp_eap_out_t) :: eap_out
!=======================================================================
