! cfd_b200_shim.f90 -- iso_c_binding shim that keeps the reference's subroutine names (calcu, calcv, calcp)
! and forwards them to libcfd_b200.so.  See INTEGRATION.md.  Not compiled in this repository's build image
! (no Fortran compiler there); builds with any Fortran 2003 compiler.
module cfd_b200_c
  use iso_c_binding
  implicit none
  integer(c_int), parameter :: CFD_U=0, CFD_V=1, CFD_P=2, CFD_PP=3, CFD_VIS=4, CFD_DEN=5
  integer(c_int), parameter :: CFD_G_X=0, CFD_G_XU=1, CFD_G_SEW=2, CFD_G_SEWU=3, &
                               CFD_G_Y=4, CFD_G_YV=5, CFD_G_SNS=6, CFD_G_SNSV=7
  type, bind(c) :: cfd_config
    integer(c_int) :: family, ni, nj, iadv, order
    integer(c_int) :: nswp(8)
    real(c_double) :: urf(8)
    integer(c_int) :: ipref, jpref, device, tdma_m, goff, ni_global, halo_w, halo_e
    real(c_double) :: sew_nim1
    integer(c_int) :: solve_flow, solve_t            ! general family: lsolve(1), lsolve(4)
    integer(c_int) :: indcos, istep, jstep           ! TEACH family (control.f90:16,66,67)
    real(c_double) :: viscos, flowin, great          ! TEACH family (control.f90:15,126,197-203)
    integer(c_int) :: joff, nj_global, halo_s, halo_n ! j-slab description (lines cut across contexts)
    real(c_double) :: sns_njm1
    integer(c_int) :: raw_coefficients               ! 1: keep ae..su as stored by the reference (cfd_lin_*)
  end type
  type, bind(c) :: cfd_turbulence                    ! literals of contro (control.f90:105-141)
    real(c_double) :: cmu, cd, c1, c2, cappa, elog, pred, prte, prandl, prandt, pfun, twalln
    real(c_double) :: urfk, urfe, urft, urfvis
    integer(c_int) :: nswpk, nswpd, nswpt, incalk, incald, incalt, inpro
  end type
  interface
    integer(c_int) function cfd_create(ctx, cfg) bind(c, name='cfd_create')
      import; type(c_ptr) :: ctx; type(cfd_config) :: cfg
    end function
    integer(c_int) function cfd_destroy(ctx) bind(c, name='cfd_destroy')
      import; type(c_ptr), value :: ctx
    end function
    integer(c_int) function cfd_set_geometry(ctx, id, a, n) bind(c, name='cfd_set_geometry')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: id, n; real(c_double) :: a(*)
    end function
    integer(c_int) function cfd_upload(ctx, field, a, ld) bind(c, name='cfd_upload')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: field, ld; real(c_double) :: a(ld,*)
    end function
    integer(c_int) function cfd_download(ctx, field, a, ld) bind(c, name='cfd_download')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: field, ld; real(c_double) :: a(ld,*)
    end function
    integer(c_int) function cfd_probe(ctx, field, i, j, val) bind(c, name='cfd_probe')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: field, i, j; real(c_double) :: val
    end function
    integer(c_int) function cfd_calcu(ctx, res) bind(c, name='cfd_calcu')
      import; type(c_ptr), value :: ctx; real(c_double) :: res
    end function
    integer(c_int) function cfd_calcv(ctx, res) bind(c, name='cfd_calcv')
      import; type(c_ptr), value :: ctx; real(c_double) :: res
    end function
    integer(c_int) function cfd_calcp(ctx, res) bind(c, name='cfd_calcp')
      import; type(c_ptr), value :: ctx; real(c_double) :: res
    end function
    integer(c_int) function cfd_lisolv(ctx, istart, jstart, field) bind(c, name='cfd_lisolv')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: istart, jstart, field
    end function
    integer(c_int) function cfd_iterate(ctx, n, res) bind(c, name='cfd_iterate')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: n; real(c_double) :: res(8)
    end function
    ! general family: calct (buoyancy_cav.f90:358-433); TEACH family after cfd_set_turbulence: calct / calcte / calced / props
    integer(c_int) function cfd_calct(ctx, res) bind(c, name='cfd_calct')
      import; type(c_ptr), value :: ctx; real(c_double) :: res
    end function
    integer(c_int) function cfd_set_turbulence(ctx, model) bind(c, name='cfd_set_turbulence')
      import; type(c_ptr), value :: ctx; type(cfd_turbulence) :: model
    end function
    integer(c_int) function cfd_calcte(ctx, res) bind(c, name='cfd_calcte')
      import; type(c_ptr), value :: ctx; real(c_double) :: res
    end function
    integer(c_int) function cfd_calced(ctx, res) bind(c, name='cfd_calced')
      import; type(c_ptr), value :: ctx; real(c_double) :: res
    end function
    integer(c_int) function cfd_props(ctx) bind(c, name='cfd_props')
      import; type(c_ptr), value :: ctx
    end function
    integer(c_int) function cfd_wall_download(ctx, which, a, n) bind(c, name='cfd_wall_download')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: which, n; real(c_double) :: a(*)
    end function
    ! scalar-transport demos: tdma_we / tdma_sn / tdma_solver on uploaded coefficients
    integer(c_int) function cfd_lin_solve(ctx, field, dir, variant, maxit, tol, iters, rsm) bind(c, name='cfd_lin_solve')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: field, dir, variant, maxit
      real(c_double), value :: tol; integer(c_int) :: iters; real(c_double) :: rsm
    end function
  end interface
end module

module calc_mod                      ! replaces cavity_fvm.f90:84-357
  use vars_mod; use cfd_b200_c
  implicit none
  type(c_ptr), save :: ctx = c_null_ptr
contains
  subroutine gpu_init()              ! call once after setgrid and the initial-value assignments (cavity_fvm.f90:538)
    type(cfd_config) :: cfg
    cfg = cfd_config(0, ni, nj, iadv, 0, [nswpu,nswpv,nswpp,0,0,0,0,0], &
                     [urfu,urfv,urfp,0d0,0d0,0d0,0d0,0d0], ipref, jpref, -1, 0, 0, 0, 0, 0, 0d0, &
                     0, 0, 0, 0, 0, 0d0, 0d0, 0d0)
    ierr = cfd_create(ctx, cfg)
    ierr = cfd_set_geometry(ctx, CFD_G_X, x, ni);   ierr = cfd_set_geometry(ctx, CFD_G_XU, xu, ni)
    ierr = cfd_set_geometry(ctx, CFD_G_SEW, sew, ni); ierr = cfd_set_geometry(ctx, CFD_G_SEWU, sewu, ni)
    ierr = cfd_set_geometry(ctx, CFD_G_Y, y, nj);   ierr = cfd_set_geometry(ctx, CFD_G_YV, yv, nj)
    ierr = cfd_set_geometry(ctx, CFD_G_SNS, sns, nj); ierr = cfd_set_geometry(ctx, CFD_G_SNSV, snsv, nj)
    ierr = cfd_upload(ctx, CFD_DEN, den, ni); ierr = cfd_upload(ctx, CFD_VIS, vis, ni)
    ierr = cfd_upload(ctx, CFD_U, u, ni);     ierr = cfd_upload(ctx, CFD_V, v, ni); ierr = cfd_upload(ctx, CFD_P, p, ni)
  end subroutine
  subroutine calcu();  ierr = cfd_calcu(ctx, resoru);  end subroutine
  subroutine calcv();  ierr = cfd_calcv(ctx, resorv);  end subroutine
  subroutine calcp();  ierr = cfd_calcp(ctx, resorm);  end subroutine
  subroutine gpu_fetch()             ! before centred_vel / output (cavity_fvm.f90:567)
    ierr = cfd_download(ctx, CFD_U, u, ni); ierr = cfd_download(ctx, CFD_V, v, ni); ierr = cfd_download(ctx, CFD_P, p, ni)
  end subroutine
end module
