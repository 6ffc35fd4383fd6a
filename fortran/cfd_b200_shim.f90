! cfd_b200_shim.f90 -- iso_c_binding shim that keeps the reference's subroutine names and dummy-argument lists
! and forwards them to libcfd_b200.so (include/cfd_b200.h).  See INTEGRATION.md.
!
!   module cfd_b200_c          constants, bind(c) types and interfaces of the C ABI
!   module calc_mod            calcu / calcv / calcp of lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90:84-357
!   module linsolv_mod         lisolv(istart,jstart,ni,nj,phi) of cavity_fvm.f90:43-82
!   module cfd_b200_general    context + set-up for the `vars` programs (buoyancy_cav.f90, cylin_cond.f90)
!   calcu, calcv, calcp, calct, tdma   external subroutines of buoyancy_cav.f90:122-433, 471-505
!   module calc                calcu .. calced of backward-step-flow-2d/calc.f90:6-580 (module general is default REAL;
!                              the library computes in double precision, the shim converts at the boundary)
!   lisolv(istart,jstart,ni,nj,it,jt,phi)   of backward-step-flow-2d/general.f90:294-327  (build with
!                              -DCFD_B200_TEACH, because it has the cavity routine's name)
!
!   module cfd_b200_heat_c, fdm2d_steady_heat_cond(...), heat2d_explicit_steps / heat2d_implicit_steps   (heat programs)
!   module cavity_mod          sipsol(i1,i2,j1,j2,aw,ae,as,an,ap,su,phi) of the two FD cavity programs
!   tdma_we, tdma_sn, tdma_solver   the line solvers of the scalar-transport demos
!
! Each reference program is built with the part of this file it needs (pre-processor symbols CFD_B200_CAVITY,
! CFD_B200_GENERAL, CFD_B200_TEACH, CFD_B200_HEAT, CFD_B200_FDCAVITY, CFD_B200_SCALAR; gfortran -cpp).  Not compiled in this repository's build image (no Fortran
! compiler there); tests/test_fortran_shim.py checks the types, constants, constructors and interfaces against
! include/cfd_b200.h.  Every component of the bind(c) types is default-initialised and every structure constructor
! uses keywords, so a field added to the C struct cannot silently break a constructor.
module cfd_b200_c
  use iso_c_binding
  implicit none
  ! enum cfd_field
  integer(c_int), parameter :: CFD_U=0, CFD_V=1, CFD_P=2, CFD_PP=3, CFD_VIS=4, CFD_DEN=5, CFD_T=6, CFD_CP=7, &
                               CFD_AE=8, CFD_AW=9, CFD_AN=10, CFD_AS=11, CFD_AP=12, CFD_SU=13, CFD_SP=14, &
                               CFD_DU=15, CFD_DV=16, CFD_GAMT=17, CFD_SUSRC=18, CFD_SPSRC=19, &
                               CFD_TE=20, CFD_ED=21, CFD_GEN=22, CFD_GAMH=23
  ! enum cfd_geom
  integer(c_int), parameter :: CFD_G_X=0, CFD_G_XU=1, CFD_G_SEW=2, CFD_G_SEWU=3, &
                               CFD_G_Y=4, CFD_G_YV=5, CFD_G_SNS=6, CFD_G_SNSV=7, &
                               CFD_G_XDIF=8, CFD_G_XCVI=9, CFD_G_XCVIP=10, CFD_G_FX=11, CFD_G_FXM=12, &
                               CFD_G_YDIF=13, CFD_G_YCVR=14, CFD_G_YCVRS=15, CFD_G_ARX=16, CFD_G_ARXJ=17, &
                               CFD_G_ARXJP=18, CFD_G_R=19, CFD_G_RMN=20, CFD_G_SX=21, CFD_G_SXMN=22, CFD_G_FV=23, &
                               CFD_G_FVP=24, CFD_G_FY=25, CFD_G_FYM=26, &
                               CFD_G_DXEP=27, CFD_G_DXPW=28, CFD_G_DXEPU=29, CFD_G_DXPWU=30, &
                               CFD_G_DYNP=31, CFD_G_DYPS=32, CFD_G_DYNPV=33, CFD_G_DYPSV=34, CFD_G_RV=35, CFD_G_RCV=36
  integer(c_int), parameter :: CFD_FAMILY_CAVITY=0, CFD_FAMILY_GENERAL=1, CFD_FAMILY_TEACH=2
  integer(c_int), parameter :: CFD_ORDER_REDBLACK=0, CFD_ORDER_SEQUENTIAL=1
  integer(c_int), parameter :: CFD_EQ_U=0, CFD_EQ_V=1, CFD_EQ_P=2, CFD_EQ_T=3, CFD_EQ_TE=4, CFD_EQ_ED=5
  integer(c_int), parameter :: CFD_HOOK_NONE=0, CFD_HOOK_BOUSSINESQ=1, CFD_HOOK_FIELD_SOURCES=2
  integer(c_int), parameter :: CFD_WALL_TAUN=0, CFD_WALL_TAUW=1, CFD_WALL_YPLUSN=2, CFD_WALL_XPLUSW=3
  type, bind(c) :: cfd_config
    integer(c_int) :: family = 0
    integer(c_int) :: ni = 0
    integer(c_int) :: nj = 0
    integer(c_int) :: iadv = 3
    integer(c_int) :: order = 0
    integer(c_int) :: nswp(8) = 0
    real(c_double) :: urf(8) = 0.0_c_double
    integer(c_int) :: ipref = 1
    integer(c_int) :: jpref = 1
    integer(c_int) :: device = -1
    integer(c_int) :: tdma_m = 0
    integer(c_int) :: goff = 0
    integer(c_int) :: ni_global = 0
    integer(c_int) :: halo_w = 0
    integer(c_int) :: halo_e = 0
    real(c_double) :: sew_nim1 = 0.0_c_double
    integer(c_int) :: solve_flow = 0
    integer(c_int) :: solve_t = 0
    integer(c_int) :: indcos = 0
    integer(c_int) :: istep = 0
    integer(c_int) :: jstep = 0
    real(c_double) :: viscos = 0.0_c_double
    real(c_double) :: flowin = 0.0_c_double
    real(c_double) :: great = 0.0_c_double
    integer(c_int) :: joff = 0
    integer(c_int) :: nj_global = 0
    integer(c_int) :: halo_s = 0
    integer(c_int) :: halo_n = 0
    real(c_double) :: sns_njm1 = 0.0_c_double
    integer(c_int) :: raw_coefficients = 0
  end type
  type, bind(c) :: cfd_hook_desc
    integer(c_int) :: kind = 0
    real(c_double) :: params(4) = 0.0_c_double
  end type
  type, bind(c) :: cfd_turbulence
    real(c_double) :: cmu = 0.0_c_double
    real(c_double) :: cd = 0.0_c_double
    real(c_double) :: c1 = 0.0_c_double
    real(c_double) :: c2 = 0.0_c_double
    real(c_double) :: cappa = 0.0_c_double
    real(c_double) :: elog = 0.0_c_double
    real(c_double) :: pred = 0.0_c_double
    real(c_double) :: prte = 0.0_c_double
    real(c_double) :: prandl = 0.0_c_double
    real(c_double) :: prandt = 0.0_c_double
    real(c_double) :: pfun = 0.0_c_double
    real(c_double) :: twalln = 0.0_c_double
    real(c_double) :: urfk = 0.0_c_double
    real(c_double) :: urfe = 0.0_c_double
    real(c_double) :: urft = 0.0_c_double
    real(c_double) :: urfvis = 0.0_c_double
    integer(c_int) :: nswpk = 0
    integer(c_int) :: nswpd = 0
    integer(c_int) :: nswpt = 0
    integer(c_int) :: incalk = 0
    integer(c_int) :: incald = 0
    integer(c_int) :: incalt = 0
    integer(c_int) :: inpro = 0
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
    integer(c_int) function cfd_upload_noalloc(ctx, field, a, ld) bind(c, name='cfd_upload_noalloc')
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
    integer(c_int) function cfd_lin_solve(ctx, field, dir, variant, maxit, tol, iters, rsm) bind(c, name='cfd_lin_solve')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: field, dir, variant, maxit; real(c_double), value :: tol
      integer(c_int) :: iters; real(c_double) :: rsm
    end function
    integer(c_int) function cfd_iterate(ctx, n, res) bind(c, name='cfd_iterate')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: n; real(c_double) :: res(8)
    end function
    integer(c_int) function cfd_set_hook(ctx, eq, desc) bind(c, name='cfd_set_hook')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: eq; type(cfd_hook_desc) :: desc
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
    integer(c_int) function cfd_wall_upload(ctx, which, a, n) bind(c, name='cfd_wall_upload')
      import; type(c_ptr), value :: ctx; integer(c_int), value :: which, n; real(c_double) :: a(*)
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

#ifdef CFD_B200_CAVITY
!======================================================================================================================
! lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90
!======================================================================================================================
module calc_mod                      ! replaces cavity_fvm.f90:84-357 (same names, no arguments, results in vars_mod)
  use vars_mod; use cfd_b200_c
  implicit none
  type(c_ptr), save :: ctx = c_null_ptr
  integer(c_int) :: line_order = CFD_ORDER_REDBLACK   ! CFD_ORDER_SEQUENTIAL reproduces the reference's iterates bit for bit
contains
  subroutine gpu_init()              ! call once after setgrid and the initial-value assignments (cavity_fvm.f90:538)
    type(cfd_config) :: cfg
    cfg = cfd_config(family=CFD_FAMILY_CAVITY, ni=ni, nj=nj, iadv=iadv, order=line_order, &
                     nswp=[nswpu,nswpv,nswpp,0,0,0,0,0], urf=[urfu,urfv,urfp,0d0,0d0,0d0,0d0,0d0], &
                     ipref=ipref, jpref=jpref)
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

module linsolv_mod                   ! replaces cavity_fvm.f90:43-82
  use types_mod; use cfd_b200_c
  implicit none
  type(c_ptr), save :: ctx_ls = c_null_ptr
  integer, save :: ls_ni = 0, ls_nj = 0
contains
  ! lisolv(istart,jstart,ni,nj,phi): the reference's dummy list; phi is passed by reference and updated in place.
  ! For callers that assemble ae..su on the HOST (module vars_mod) and want only the line solver on the GPU: the six
  ! coefficient arrays and phi go up, one W->E sweep runs in the reference's line order, phi comes back.  (calc_mod
  ! above never calls this: its sweeps stay on the device between assembly and correction.)
  subroutine lisolv(istart, jstart, ni, nj, phi)
    use vars_mod, only: ap, an, as, ae, aw, su
    integer, intent(in) :: istart, jstart, ni, nj
    real(dp), dimension(ni,nj), intent(inout) :: phi     ! the reference says intent(out) but reads phi (cavity_fvm.f90:63,69)
    type(cfd_config) :: cfg
    integer(c_int) :: rc
    if (.not. c_associated(ctx_ls) .or. ni /= ls_ni .or. nj /= ls_nj) then
      if (c_associated(ctx_ls)) rc = cfd_destroy(ctx_ls)
      cfg = cfd_config(family=CFD_FAMILY_CAVITY, ni=ni, nj=nj, order=CFD_ORDER_SEQUENTIAL, raw_coefficients=1)
      rc = cfd_create(ctx_ls, cfg)
      ls_ni = ni; ls_nj = nj
    end if
    rc = cfd_upload(ctx_ls, CFD_AP, ap, ni); rc = cfd_upload(ctx_ls, CFD_AN, an, ni); rc = cfd_upload(ctx_ls, CFD_AS, as, ni)
    rc = cfd_upload(ctx_ls, CFD_AE, ae, ni); rc = cfd_upload(ctx_ls, CFD_AW, aw, ni); rc = cfd_upload(ctx_ls, CFD_SU, su, ni)
    rc = cfd_upload(ctx_ls, CFD_PP, phi, ni)
    rc = cfd_lisolv(ctx_ls, istart, jstart, CFD_PP)
    rc = cfd_download(ctx_ls, CFD_PP, phi, ni)
  end subroutine
end module
#endif

#ifdef CFD_B200_GENERAL
!======================================================================================================================
! buoyancy-driven-cavity-2d/buoyancy_cav.f90 and cylinder-heat-conduction-2d/cylin_cond.f90 (modules vars, user, serv)
!======================================================================================================================
module cfd_b200_general
  use vars; use cfd_b200_c
  implicit none
  type(c_ptr), save :: ctx = c_null_ptr
  type(c_ptr), save :: ctx_ls = c_null_ptr
  integer(c_int) :: line_order = CFD_ORDER_REDBLACK
contains
  ! call once after grid() and initial() (buoyancy_cav.f90:7-10).  gamsor() is host code the kernels cannot call, so
  ! its effect is handed over as data: `gam` as it stands for the momentum equations (nq = 1, 2) is uploaded as CFD_VIS,
  ! gamt / su_t / sp_t (gamsor's output for nq = 4, per unit volume) as CFD_GAMT / CFD_SUSRC / CFD_SPSRC, and the
  ! Boussinesq force of the v equation as the enumerated hook (rho*g*beta, Tref).
  subroutine gpu_init(gamt, su_t, sp_t, rgbeta, tref)
    real(dp), intent(in) :: gamt(ni,nj), su_t(ni,nj), sp_t(ni,nj), rgbeta, tref
    type(cfd_config) :: cfg
    type(cfd_hook_desc) :: hook
    integer(c_int) :: rc
    cfg = cfd_config(family=CFD_FAMILY_GENERAL, ni=ni, nj=nj, iadv=iadv, order=line_order, &
                     nswp=[nsweep(1),nsweep(2),nsweep(3),nsweep(4),0,0,0,0], &
                     urf=[urf(1),urf(2),urf(3),urf(4),0d0,0d0,0d0,0d0], ipref=ipref, jpref=jpref, &
                     solve_flow=merge(1,0,lsolve(1)), solve_t=merge(1,0,lsolve(4)))
    rc = cfd_create(ctx, cfg)
    rc = cfd_set_geometry(ctx, CFD_G_X, x, ni);       rc = cfd_set_geometry(ctx, CFD_G_XU, xu, ni)
    rc = cfd_set_geometry(ctx, CFD_G_XDIF, xdif, ni); rc = cfd_set_geometry(ctx, CFD_G_SEW, sew, ni)
    rc = cfd_set_geometry(ctx, CFD_G_SEWU, sewu, ni); rc = cfd_set_geometry(ctx, CFD_G_XCVI, xcvi, ni)
    rc = cfd_set_geometry(ctx, CFD_G_XCVIP, xcvip, ni); rc = cfd_set_geometry(ctx, CFD_G_FX, fx, ni)
    rc = cfd_set_geometry(ctx, CFD_G_FXM, fxm, ni)
    rc = cfd_set_geometry(ctx, CFD_G_Y, y, nj);       rc = cfd_set_geometry(ctx, CFD_G_YV, yv, nj)
    rc = cfd_set_geometry(ctx, CFD_G_YDIF, ydif, nj); rc = cfd_set_geometry(ctx, CFD_G_SNS, sns, nj)
    rc = cfd_set_geometry(ctx, CFD_G_SNSV, snsv, nj); rc = cfd_set_geometry(ctx, CFD_G_YCVR, ycvr, nj)
    rc = cfd_set_geometry(ctx, CFD_G_YCVRS, ycvrs, nj); rc = cfd_set_geometry(ctx, CFD_G_ARX, arx, nj)
    rc = cfd_set_geometry(ctx, CFD_G_ARXJ, arxj, nj); rc = cfd_set_geometry(ctx, CFD_G_ARXJP, arxjp, nj)
    rc = cfd_set_geometry(ctx, CFD_G_R, r, nj);       rc = cfd_set_geometry(ctx, CFD_G_RMN, rmn, nj)
    rc = cfd_set_geometry(ctx, CFD_G_SX, sx, nj);     rc = cfd_set_geometry(ctx, CFD_G_SXMN, sxmn, nj)
    rc = cfd_set_geometry(ctx, CFD_G_FV, fv, nj);     rc = cfd_set_geometry(ctx, CFD_G_FVP, fvp, nj)
    rc = cfd_set_geometry(ctx, CFD_G_FY, fy, nj);     rc = cfd_set_geometry(ctx, CFD_G_FYM, fym, nj)
    rc = cfd_upload(ctx, CFD_U, u, ni);   rc = cfd_upload(ctx, CFD_V, v, ni);   rc = cfd_upload(ctx, CFD_P, p, ni)
    rc = cfd_upload(ctx, CFD_T, t, ni);   rc = cfd_upload(ctx, CFD_DEN, den, ni); rc = cfd_upload(ctx, CFD_CP, cp, ni)
    rc = cfd_upload(ctx, CFD_VIS, gam, ni)
    rc = cfd_upload(ctx, CFD_GAMT, gamt, ni); rc = cfd_upload(ctx, CFD_SUSRC, su_t, ni); rc = cfd_upload(ctx, CFD_SPSRC, sp_t, ni)
    hook = cfd_hook_desc(kind=CFD_HOOK_FIELD_SOURCES)
    rc = cfd_set_hook(ctx, CFD_EQ_T, hook)
    if (rgbeta /= 0.0_dp) then
      hook = cfd_hook_desc(kind=CFD_HOOK_BOUSSINESQ, params=[rgbeta, tref, 0d0, 0d0])
      rc = cfd_set_hook(ctx, CFD_EQ_V, hook)
    end if
    ierr = rc
  end subroutine
  subroutine gpu_fetch()             ! before postproc() (buoyancy_cav.f90:23)
    integer(c_int) :: rc
    rc = cfd_download(ctx, CFD_U, u, ni); rc = cfd_download(ctx, CFD_V, v, ni); rc = cfd_download(ctx, CFD_P, p, ni)
    rc = cfd_download(ctx, CFD_T, t, ni)
  end subroutine
end module

subroutine calcu()                   ! buoyancy_cav.f90:122-198
  use cfd_b200_general
  ierr = cfd_calcu(ctx, resor(1))
end subroutine
subroutine calcv()                   ! buoyancy_cav.f90:200-277
  use cfd_b200_general
  ierr = cfd_calcv(ctx, resor(2))
end subroutine
subroutine calcp()                   ! buoyancy_cav.f90:279-356
  use cfd_b200_general
  ierr = cfd_calcp(ctx, resor(3))
end subroutine
subroutine calct()                   ! buoyancy_cav.f90:358-433
  use cfd_b200_general
  ierr = cfd_calct(ctx, resor(4))
end subroutine
! tdma(istart,jstart,ni,nj,phi) (buoyancy_cav.f90:471-505): the line solver alone, on host-assembled coefficients of
! module vars; phi updated in place.
subroutine tdma(istart, jstart, ni, nj, phi)
  use vars, only: dp, ap, an, as, ae, aw, su
  use cfd_b200_general, only: ctx_ls
  use cfd_b200_c
  implicit none
  integer, intent(in) :: istart, jstart, ni, nj
  real(dp), dimension(ni,nj), intent(inout) :: phi
  type(cfd_config) :: cfg
  integer(c_int) :: rc
  integer, save :: ls_ni = 0, ls_nj = 0
  if (.not. c_associated(ctx_ls) .or. ni /= ls_ni .or. nj /= ls_nj) then
    if (c_associated(ctx_ls)) rc = cfd_destroy(ctx_ls)
    ! a cavity-family context is enough: the sweep only needs the six coefficient arrays and phi
    cfg = cfd_config(family=CFD_FAMILY_CAVITY, ni=ni, nj=nj, order=CFD_ORDER_SEQUENTIAL, raw_coefficients=1)
    rc = cfd_create(ctx_ls, cfg)
    ls_ni = ni; ls_nj = nj
  end if
  rc = cfd_upload(ctx_ls, CFD_AP, ap, ni); rc = cfd_upload(ctx_ls, CFD_AN, an, ni); rc = cfd_upload(ctx_ls, CFD_AS, as, ni)
  rc = cfd_upload(ctx_ls, CFD_AE, ae, ni); rc = cfd_upload(ctx_ls, CFD_AW, aw, ni); rc = cfd_upload(ctx_ls, CFD_SU, su, ni)
  rc = cfd_upload(ctx_ls, CFD_PP, phi, ni)
  rc = cfd_lisolv(ctx_ls, istart, jstart, CFD_PP)
  rc = cfd_download(ctx_ls, CFD_PP, phi, ni)
end subroutine
#endif

#ifdef CFD_B200_TEACH
!======================================================================================================================
! backward-step-flow-2d/ (modules general, promod, calc).  Module general is default REAL and dimensioned (it,jt);
! the library is double precision, so fields are converted on the way in and out (dble / real).
!======================================================================================================================
module calc                          ! replaces backward-step-flow-2d/calc.f90:6-580
  use general; use cfd_b200_c
  implicit none
  type(c_ptr), save :: ctx = c_null_ptr
  integer(c_int) :: line_order = CFD_ORDER_REDBLACK
contains
  subroutine up2(field, a)           ! a(it,jt) default real -> device field (ni,nj)
    integer(c_int), intent(in) :: field
    real, intent(in) :: a(:,:)
    real(c_double), allocatable :: d(:,:)
    integer(c_int) :: rc
    allocate(d(ni,nj)); d = dble(a(1:ni,1:nj))
    rc = cfd_upload(ctx, field, d, ni)
  end subroutine
  subroutine down2(field, a)
    integer(c_int), intent(in) :: field
    real, intent(inout) :: a(:,:)
    real(c_double), allocatable :: d(:,:)
    integer(c_int) :: rc
    allocate(d(ni,nj))
    rc = cfd_download(ctx, field, d, ni); a(1:ni,1:nj) = real(d)
  end subroutine
  subroutine geo(id, a, n)
    integer(c_int), intent(in) :: id
    integer, intent(in) :: n
    real, intent(in) :: a(:)
    real(c_double), allocatable :: d(:)
    integer(c_int) :: rc
    allocate(d(n)); d = dble(a(1:n))
    rc = cfd_set_geometry(ctx, id, d, n)
  end subroutine
  ! call once after init, the initial fields and flowin are set (control.f90:204); turbulent = k-epsilon + energy on
  subroutine gpu_init(turbulent)
    logical, intent(in) :: turbulent
    type(cfd_config) :: cfg
    type(cfd_turbulence) :: m
    real(c_double), allocatable :: w(:)
    integer(c_int) :: rc
    cfg = cfd_config(family=CFD_FAMILY_TEACH, ni=ni, nj=nj, iadv=3, order=line_order, &
                     nswp=[nswpu,nswpv,nswpp,0,0,0,0,0], urf=[dble(urfu),dble(urfv),dble(urfp),0d0,0d0,0d0,0d0,0d0], &
                     ipref=ipref, jpref=jpref, indcos=indcos, istep=int(istep), jstep=int(jstep), &
                     viscos=dble(viscos), flowin=dble(flowin), great=dble(great))
    rc = cfd_create(ctx, cfg)
    call geo(CFD_G_X, x, ni);       call geo(CFD_G_XU, xu, ni);     call geo(CFD_G_SEW, sew, ni)
    call geo(CFD_G_SEWU, sewu, ni); call geo(CFD_G_DXEP, dxep, ni); call geo(CFD_G_DXPW, dxpw, ni)
    call geo(CFD_G_DXEPU, dxepu, ni); call geo(CFD_G_DXPWU, dxpwu, ni)
    call geo(CFD_G_Y, y, nj);       call geo(CFD_G_YV, yv, nj);     call geo(CFD_G_R, r, nj)
    call geo(CFD_G_RV, rv, nj);     call geo(CFD_G_RCV, rcv, nj);   call geo(CFD_G_SNS, sns, nj)
    call geo(CFD_G_SNSV, snsv, nj); call geo(CFD_G_DYNP, dynp, nj); call geo(CFD_G_DYPS, dyps, nj)
    call geo(CFD_G_DYNPV, dynpv, nj); call geo(CFD_G_DYPSV, dypsv, nj)
    call up2(CFD_DEN, den); call up2(CFD_VIS, vis)
    call up2(CFD_U, u); call up2(CFD_V, v); call up2(CFD_P, p)
    if (turbulent) then
      m = cfd_turbulence(cmu=dble(cmu), cd=dble(cd), c1=dble(c1), c2=dble(c2), cappa=dble(cappa), elog=dble(elog), &
                         pred=dble(pred), prte=dble(prte), prandl=dble(prandl), prandt=dble(prandt), pfun=dble(pfun), &
                         twalln=dble(twalln), urfk=dble(urfk), urfe=dble(urfe), urft=dble(urft), urfvis=dble(urfvis), &
                         nswpk=nswpk, nswpd=nswpd, nswpt=nswpt, incalk=merge(1,0,incalk), incald=merge(1,0,incald), &
                         incalt=merge(1,0,incalt), inpro=merge(1,0,inpro))
      rc = cfd_set_turbulence(ctx, m)
      call up2(CFD_TE, te); call up2(CFD_ED, ed); call up2(CFD_T, t)
      allocate(w(max(ni,nj)))
      w(1:ni) = dble(yplusn(1:ni)); rc = cfd_wall_upload(ctx, CFD_WALL_YPLUSN, w, ni)
      w(1:nj) = dble(xplusw(1:nj)); rc = cfd_wall_upload(ctx, CFD_WALL_XPLUSW, w, nj)
    end if
  end subroutine
  subroutine calcu()                 ! calc.f90:20-133 (+ modu, promod.f90:26-95)
    real(c_double) :: res; integer(c_int) :: rc
    rc = cfd_calcu(ctx, res); resoru = real(res)
  end subroutine
  subroutine calcv()                 ! calc.f90:137-227 (+ modv, promod.f90:99-133)
    real(c_double) :: res; integer(c_int) :: rc
    rc = cfd_calcv(ctx, res); resorv = real(res)
  end subroutine
  subroutine calcp()                 ! calc.f90:231-305
    real(c_double) :: res; integer(c_int) :: rc
    rc = cfd_calcp(ctx, res); resorm = real(res)
  end subroutine
  subroutine calct()                 ! calc.f90:309-385 (+ modt)
    real(c_double) :: res; integer(c_int) :: rc
    rc = cfd_calct(ctx, res); resort = real(res)
  end subroutine
  subroutine calcte()                ! calc.f90:388-485 (+ modte)
    real(c_double) :: res; integer(c_int) :: rc
    rc = cfd_calcte(ctx, res); resork = real(res)
  end subroutine
  subroutine calced()                ! calc.f90:490-577 (+ moded)
    real(c_double) :: res; integer(c_int) :: rc
    rc = cfd_calced(ctx, res); resore = real(res)
  end subroutine
  subroutine props_gpu()             ! props (general.f90:165-190) on the device fields
    integer(c_int) :: rc
    rc = cfd_props(ctx)
  end subroutine
  function monitor(field, i, j) result(val)   ! u(imon,jmon) etc. in the residual print (control.f90:223-226)
    integer(c_int), intent(in) :: field
    integer, intent(in) :: i, j
    real :: val
    real(c_double) :: d; integer(c_int) :: rc
    rc = cfd_probe(ctx, field, i, j, d); val = real(d)
  end function
  subroutine gpu_fetch()             ! before writeout (control.f90:228-231, 243-296)
    real(c_double), allocatable :: w(:)
    integer(c_int) :: rc
    call down2(CFD_U, u); call down2(CFD_V, v); call down2(CFD_P, p)
    if (incalk .or. incald .or. incalt) then
      call down2(CFD_TE, te); call down2(CFD_ED, ed); call down2(CFD_T, t); call down2(CFD_VIS, vis)
      allocate(w(max(ni,nj)))
      rc = cfd_wall_download(ctx, CFD_WALL_TAUN, w, ni);   taun(1:ni) = real(w(1:ni))
      rc = cfd_wall_download(ctx, CFD_WALL_YPLUSN, w, ni); yplusn(1:ni) = real(w(1:ni))
      rc = cfd_wall_download(ctx, CFD_WALL_TAUW, w, nj);   tauw(1:nj) = real(w(1:nj))
      rc = cfd_wall_download(ctx, CFD_WALL_XPLUSW, w, nj); xplusw(1:nj) = real(w(1:nj))
    end if
  end subroutine
end module

! lisolv(istart,jstart,ni,nj,it,jt,phi) (general.f90:294-327): the line solver alone on host-assembled coefficients of
! module general, leading dimensions (it,jt); phi updated in place.
subroutine lisolv(istart, jstart, ni, nj, it, jt, phi)
  use general, only: ap, an, as, ae, aw, su
  use cfd_b200_c
  implicit none
  integer, intent(in) :: istart, jstart, ni, nj, it, jt
  real, dimension(it,jt), intent(inout) :: phi
  type(c_ptr), save :: ctx_ls = c_null_ptr
  integer, save :: ls_ni = 0, ls_nj = 0
  type(cfd_config) :: cfg
  real(c_double), allocatable :: d(:,:)
  integer(c_int) :: rc
  if (.not. c_associated(ctx_ls) .or. ni /= ls_ni .or. nj /= ls_nj) then
    if (c_associated(ctx_ls)) rc = cfd_destroy(ctx_ls)
    cfg = cfd_config(family=CFD_FAMILY_CAVITY, ni=ni, nj=nj, order=CFD_ORDER_SEQUENTIAL, raw_coefficients=1)
    rc = cfd_create(ctx_ls, cfg)
    ls_ni = ni; ls_nj = nj
  end if
  allocate(d(ni,nj))
  d = dble(ap(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_AP, d, ni)
  d = dble(an(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_AN, d, ni)
  d = dble(as(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_AS, d, ni)
  d = dble(ae(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_AE, d, ni)
  d = dble(aw(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_AW, d, ni)
  d = dble(su(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_SU, d, ni)
  d = dble(phi(1:ni,1:nj)); rc = cfd_upload(ctx_ls, CFD_PP, d, ni)
  rc = cfd_lisolv(ctx_ls, istart, jstart, CFD_PP)
  rc = cfd_download(ctx_ls, CFD_PP, d, ni); phi(1:ni,1:nj) = real(d)
end subroutine
#endif

! ======================================================================================================================
! Heat-conduction programs (finite-difference-method/heat-conduction/): build with -DCFD_B200_HEAT.
!   module cfd_b200_heat_c     bind(c) type + interfaces of the cfd_heat_* part of the C ABI
!   fdm2d_steady_heat_cond(xmin,xmax,ymin,ymax,tleft,tright,ttop,tbottom,t,nx,ny)
!                              steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:68-105 with the same dummy list: the
!                              dense (nx*ny)^2 matrix of interior()/boundary() is never built; LAMBDA and QDOT are still
!                              the program's own functions, evaluated at the points interior() evaluates them (:171-188)
!   heat2d_explicit_steps(...) the update loop of transient-conduction-1d/explicit-time/fdm1d_explicit_heat_cond.f90:
!                              130-139 on a 2-D plate (the 5-point operator of the steady program, constant conductivity)
!   heat2d_implicit_steps(...) the same plate with the implicit 1-D assembly + thomas of .../implicit-time/
!                              fdm1d_explicit_heat_cond.f90:122-184 applied along x, then along y
! ======================================================================================================================
#ifdef CFD_B200_HEAT
module cfd_b200_heat_c
  use iso_c_binding
  implicit none
  type, bind(c) :: cfd_heat_config
    integer(c_int) :: nx = 0
    integer(c_int) :: ny = 0
    real(c_double) :: k = 0.0_c_double
    real(c_double) :: rho = 0.0_c_double
    real(c_double) :: cp = 0.0_c_double
    real(c_double) :: xa = 0.0_c_double
    real(c_double) :: xb = 0.0_c_double
    real(c_double) :: ya = 0.0_c_double
    real(c_double) :: yb = 0.0_c_double
    real(c_double) :: tleft = 0.0_c_double
    real(c_double) :: tright = 0.0_c_double
    real(c_double) :: tbottom = 0.0_c_double
    real(c_double) :: ttop = 0.0_c_double
    real(c_double) :: dt = 0.0_c_double
    real(c_double) :: qthr = 0.0_c_double
    real(c_double) :: qval = 0.0_c_double
    integer(c_int) :: device = -1
    integer(c_int) :: j_goff = 0
    integer(c_int) :: ny_global = 0
    integer(c_int) :: halo_s = 0
    integer(c_int) :: halo_n = 0
  end type
  interface
    integer(c_int) function cfd_heat_create(h, cfg) bind(c, name='cfd_heat_create')
      import; type(c_ptr) :: h; type(cfd_heat_config) :: cfg
    end function
    integer(c_int) function cfd_heat_destroy(h) bind(c, name='cfd_heat_destroy')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function cfd_heat_upload(h, t, ld) bind(c, name='cfd_heat_upload')
      import; type(c_ptr), value :: h; integer(c_int), value :: ld; real(c_double) :: t(ld,*)
    end function
    integer(c_int) function cfd_heat_download(h, t, ld) bind(c, name='cfd_heat_download')
      import; type(c_ptr), value :: h; integer(c_int), value :: ld; real(c_double) :: t(ld,*)
    end function
    integer(c_int) function cfd_heat2d_step(h, nsteps) bind(c, name='cfd_heat2d_step')
      import; type(c_ptr), value :: h; integer(c_int), value :: nsteps
    end function
    integer(c_int) function cfd_heat2d_implicit_step(h, nsteps) bind(c, name='cfd_heat2d_implicit_step')
      import; type(c_ptr), value :: h; integer(c_int), value :: nsteps
    end function
    integer(c_int) function cfd_heat_implicit_halo(fo_y, rows) bind(c, name='cfd_heat_implicit_halo')
      import; real(c_double), value :: fo_y; integer(c_int) :: rows
    end function
    integer(c_int) function cfd_heat_set_operator(h, lamw, lame, lams, lamn, qdot, ld) bind(c, name='cfd_heat_set_operator')
      import; type(c_ptr), value :: h; integer(c_int), value :: ld
      real(c_double) :: lamw(ld,*), lame(ld,*), lams(ld,*), lamn(ld,*), qdot(ld,*)
    end function
    integer(c_int) function cfd_heat2d_apply(h, r, ld) bind(c, name='cfd_heat2d_apply')
      import; type(c_ptr), value :: h; integer(c_int), value :: ld; real(c_double) :: r(ld,*)
    end function
    integer(c_int) function cfd_heat2d_step_var(h, nsteps) bind(c, name='cfd_heat2d_step_var')
      import; type(c_ptr), value :: h; integer(c_int), value :: nsteps
    end function
    integer(c_int) function cfd_heat2d_steady_gs(h, maxit, tol, rf, iters, dx2) bind(c, name='cfd_heat2d_steady_gs')
      import; type(c_ptr), value :: h; integer(c_int), value :: maxit; real(c_double), value :: tol, rf
      integer(c_int) :: iters; real(c_double) :: dx2
    end function
    integer(c_int) function cfd_heat_synchronize(h) bind(c, name='cfd_heat_synchronize')
      import; type(c_ptr), value :: h
    end function
  end interface
end module

! The reference's driver routine, same dummy list (fdm2d_steady_heat_cond.f90:68-105).  t(nx*ny) is numbered
! k = (j-1)*nx + i as in the reference (:70-79), i.e. it is the column-major plate T(nx,ny).
subroutine fdm2d_steady_heat_cond(xmin, xmax, ymin, ymax, tleft, tright, ttop, tbottom, t, nx, ny)
  use cfd_b200_heat_c
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  integer, intent(in) :: nx, ny
  real(dp), intent(in) :: xmin, xmax, ymin, ymax, tleft, tright, ttop, tbottom
  real(dp), intent(out) :: t(nx*ny)
  real(dp), external :: LAMBDA, QDOT
  real(c_double), allocatable :: lw(:,:), le(:,:), ls(:,:), ln(:,:), q(:,:), t2(:,:)
  real(dp) :: x(nx), y(ny), dx, dy
  type(cfd_heat_config) :: cfg
  type(c_ptr) :: h
  integer(c_int) :: rc, iters
  real(c_double) :: dx2
  integer :: i, j
  call xymesh(xmin, xmax, ymin, ymax, x, y, nx, ny)          ! the program's own mesh routine (:107-129)
  dx = x(2) - x(1); dy = y(2) - y(1)
  allocate(lw(nx,ny), le(nx,ny), ls(nx,ny), ln(nx,ny), q(nx,ny), t2(nx,ny))
  lw = 0; le = 0; ls = 0; ln = 0; q = 0; t2 = 0
  do j = 2, ny-1
    do i = 2, nx-1
      lw(i,j) = LAMBDA(x(i)-0.5_dp*dx, y(j)); le(i,j) = LAMBDA(x(i)+0.5_dp*dx, y(j))      ! :171-172
      ls(i,j) = LAMBDA(x(i), y(j)-0.5_dp*dy); ln(i,j) = LAMBDA(x(i), y(j)+0.5_dp*dy)      ! :173-174
      q(i,j) = QDOT(x(i), y(j))                                                            ! :188
    end do
  end do
  cfg = cfd_heat_config(nx=nx, ny=ny, k=1.0_c_double, rho=1.0_c_double, cp=1.0_c_double, xa=xmin, xb=xmax, ya=ymin, yb=ymax, &
                        tleft=tleft, tright=tright, tbottom=tbottom, ttop=ttop, dt=1.0_c_double)
  rc = cfd_heat_create(h, cfg)
  rc = cfd_heat_upload(h, t2, nx)                             ! gauss_seidel starts from t0 = 0 (:95)
  rc = cfd_heat_set_operator(h, lw, le, ls, ln, q, nx)
  rc = cfd_heat2d_steady_gs(h, 500, 1.0e-7_c_double, 0.5_c_double, iters, dx2)           ! maxit, tol, rf of :248-287
  rc = cfd_heat_download(h, t2, nx)
  t = reshape(t2, (/ nx*ny /))
  rc = cfd_heat_destroy(h)
end subroutine

! nsteps explicit / implicit steps of a plate T(nx,ny) held by the caller (uploaded, advanced, downloaded)
subroutine heat2d_explicit_steps(k, rho, cp, xa, xb, ya, yb, tleft, tright, tbottom, ttop, dt, t, nx, ny, nsteps)
  use cfd_b200_heat_c
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  integer, intent(in) :: nx, ny, nsteps
  real(dp), intent(in) :: k, rho, cp, xa, xb, ya, yb, tleft, tright, tbottom, ttop, dt
  real(dp), intent(inout) :: t(nx,ny)
  type(cfd_heat_config) :: cfg
  type(c_ptr) :: h
  integer(c_int) :: rc
  cfg = cfd_heat_config(nx=nx, ny=ny, k=k, rho=rho, cp=cp, xa=xa, xb=xb, ya=ya, yb=yb, tleft=tleft, tright=tright, &
                        tbottom=tbottom, ttop=ttop, dt=dt)
  rc = cfd_heat_create(h, cfg)
  rc = cfd_heat_upload(h, t, nx)
  rc = cfd_heat2d_step(h, nsteps)
  rc = cfd_heat_download(h, t, nx)
  rc = cfd_heat_destroy(h)
end subroutine

subroutine heat2d_implicit_steps(k, rho, cp, xa, xb, ya, yb, tleft, tright, tbottom, ttop, dt, t, nx, ny, nsteps)
  use cfd_b200_heat_c
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  integer, intent(in) :: nx, ny, nsteps
  real(dp), intent(in) :: k, rho, cp, xa, xb, ya, yb, tleft, tright, tbottom, ttop, dt
  real(dp), intent(inout) :: t(nx,ny)
  type(cfd_heat_config) :: cfg
  type(c_ptr) :: h
  integer(c_int) :: rc
  cfg = cfd_heat_config(nx=nx, ny=ny, k=k, rho=rho, cp=cp, xa=xa, xb=xb, ya=ya, yb=yb, tleft=tleft, tright=tright, &
                        tbottom=tbottom, ttop=ttop, dt=dt)
  rc = cfd_heat_create(h, cfg)
  rc = cfd_heat_upload(h, t, nx)
  rc = cfd_heat2d_implicit_step(h, nsteps)
  rc = cfd_heat_download(h, t, nx)
  rc = cfd_heat_destroy(h)
end subroutine
#endif

! ======================================================================================================================
! Finite-difference cavity programs (lid-driven-cavity-2d/fdm-explicit, fdm-implicit): build with -DCFD_B200_FDCAVITY
! (add -DCFD_B200_SIP_EXPLICIT for the explicit program's solver constants).  Replaces `module cavity_mod`:
!   sipsol(i1, i2, j1, j2, aw, ae, as, an, ap, su, phi)      fdm_implicit_cavity.f90:29-92 / fdm_explicit_cavity.f90:24-87,
!                              same dummy list; the arrays arrive already negated (`call sipsol(1, ni, 1, nj+1, -aw, ...)`),
!                              so the library is told to use them as they are (negate = 0)
! ======================================================================================================================
#ifdef CFD_B200_FDCAVITY
module cavity_mod
  use iso_c_binding
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  interface
    integer(c_int) function cfd_sipsol(device, i2, j2, ld, aw, ae, as, an, ap, su, phi, negate, alpha, tol, maxit, iters, rsm) &
        bind(c, name='cfd_sipsol')
      import; integer(c_int), value :: device, i2, j2, ld, negate, maxit; real(c_double), value :: alpha, tol
      real(c_double) :: aw(ld,*), ae(ld,*), as(ld,*), an(ld,*), ap(ld,*), su(ld,*), phi(ld,*)
      integer(c_int) :: iters; real(c_double) :: rsm
    end function
  end interface
contains
  subroutine sipsol(i1, i2, j1, j2, aw, ae, as, an, ap, su, phi)
    integer, intent(in) :: i1, i2, j1, j2
    real(dp), dimension(:,:), intent(in) :: aw, ae, as, an, ap, su
    real(dp), dimension(:,:), intent(out) :: phi            ! in/out in fact: the iteration starts from it (:66-68)
#ifdef CFD_B200_SIP_EXPLICIT
    integer, parameter :: maxit = 600                       ! fdm_explicit_cavity.f90:29-31
    real(dp), parameter :: alpha = 0.88_dp, tol = 1.0e-5_dp
#else
    integer, parameter :: maxit = 3000                      ! fdm_implicit_cavity.f90:34-36
    real(dp), parameter :: alpha = 0.80_dp, tol = 1.0e-6_dp
#endif
    real(c_double), allocatable :: c(:,:,:), x(:,:)
    integer(c_int) :: rc, iters
    real(c_double) :: rsm
    integer :: ld, nc
    ! every caller passes whole arrays with i1 = j1 = 1 (:198, :220, :241 / :206); sections would be copied here anyway
    ld = size(phi, 1); nc = size(phi, 2)
    allocate(c(ld,nc,6), x(ld,nc))
    c(:,:,1) = aw; c(:,:,2) = ae; c(:,:,3) = as; c(:,:,4) = an; c(:,:,5) = ap; c(:,:,6) = su
    x = phi
    rc = cfd_sipsol(-1, i2, j2, ld, c(:,:,1), c(:,:,2), c(:,:,3), c(:,:,4), c(:,:,5), c(:,:,6), x, 0, alpha, tol, maxit, iters, rsm)
    phi = x
  end subroutine
end module
#endif


! ======================================================================================================================
! Scalar-transport demos (stagnation-point-flow-2d/steady and /unsteady, diagonal-convection-2d): build with
! -DCFD_B200_SCALAR.  The iterative line-by-line solvers keep their names and dummy lists:
!   tdma_we / tdma_sn (aw, ae, as, an, ap, su, phi)     fvm2d_stag_point.f90:82-160 (neighbour terms on the right) and, with
!                              -DCFD_B200_SCALAR_LHS, the unsteady program's copies stag_point_unsteady.f90:79-160, which
!                              keep every term on the left: the neighbour arrays are uploaded negated (exact sign flips)
!   tdma_solver(aw, ae, as, an, ap, su, phi, ni, nj)    fvm2d_oblique_conv.f90:284-318 (all terms on the left, its own stop
!                              rule: variant 1 of cfd_lin_solve)
! ======================================================================================================================
#ifdef CFD_B200_SCALAR
module cfd_b200_scalar
  use cfd_b200_c
  implicit none
  integer, parameter, private :: dp = selected_real_kind(15)
  type(c_ptr), private, save :: ctx_lin = c_null_ptr
  integer, private, save :: lin_ni = 0, lin_nj = 0
contains
  ! dir 0: W->E sweeps of S-N lines, 1: S->N sweeps of W-E lines; sgn -1: the caller keeps the neighbour terms on the left
  subroutine lin_solve(aw, ae, as, an, ap, su, phi, dir, variant, maxit, sgn)
    real(dp), dimension(:,:), intent(in) :: aw, ae, as, an, ap, su
    real(dp), dimension(:,:), intent(inout) :: phi
    integer, intent(in) :: dir, variant, maxit
    real(dp), intent(in) :: sgn
    type(cfd_config) :: cfg
    real(c_double), allocatable :: d(:,:)
    integer(c_int) :: rc, iters
    real(c_double) :: rsm
    integer :: ni, nj
    ni = size(phi, 1); nj = size(phi, 2)
    if (.not. c_associated(ctx_lin) .or. ni /= lin_ni .or. nj /= lin_nj) then
      if (c_associated(ctx_lin)) rc = cfd_destroy(ctx_lin)
      cfg = cfd_config(family=CFD_FAMILY_CAVITY, ni=ni, nj=nj, order=CFD_ORDER_SEQUENTIAL, raw_coefficients=1)
      rc = cfd_create(ctx_lin, cfg)
      lin_ni = ni; lin_nj = nj
    end if
    allocate(d(ni,nj))
    d = sgn*ae; rc = cfd_upload(ctx_lin, CFD_AE, d, ni)
    d = sgn*aw; rc = cfd_upload(ctx_lin, CFD_AW, d, ni)
    d = sgn*an; rc = cfd_upload(ctx_lin, CFD_AN, d, ni)
    d = sgn*as; rc = cfd_upload(ctx_lin, CFD_AS, d, ni)
    d = ap;     rc = cfd_upload(ctx_lin, CFD_AP, d, ni)
    d = su;     rc = cfd_upload(ctx_lin, CFD_SU, d, ni)
    d = phi;    rc = cfd_upload(ctx_lin, CFD_PP, d, ni)
    rc = cfd_lin_solve(ctx_lin, CFD_PP, dir, variant, maxit, 1.0e-4_c_double, iters, rsm)
    rc = cfd_download(ctx_lin, CFD_PP, d, ni)
    phi = d
  end subroutine
end module

subroutine tdma_we(aw, ae, as, an, ap, su, phi)
  use cfd_b200_scalar
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  real(dp), dimension(:,:), intent(in) :: aw, ae, as, an, ap, su
  real(dp), dimension(:,:), intent(out) :: phi
#ifdef CFD_B200_SCALAR_LHS
  call lin_solve(aw, ae, as, an, ap, su, phi, 0, 0, 1000, -1.0_dp)
#else
  call lin_solve(aw, ae, as, an, ap, su, phi, 0, 0, 1000, 1.0_dp)
#endif
end subroutine

subroutine tdma_sn(aw, ae, as, an, ap, su, phi)
  use cfd_b200_scalar
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  real(dp), dimension(:,:), intent(in) :: aw, ae, as, an, ap, su
  real(dp), dimension(:,:), intent(out) :: phi
#ifdef CFD_B200_SCALAR_LHS
  call lin_solve(aw, ae, as, an, ap, su, phi, 1, 0, 1000, -1.0_dp)
#else
  call lin_solve(aw, ae, as, an, ap, su, phi, 1, 0, 1000, 1.0_dp)
#endif
end subroutine

subroutine tdma_solver(aw, ae, as, an, ap, su, phi, ni, nj)
  use cfd_b200_scalar
  implicit none
  integer, parameter :: dp = selected_real_kind(15)
  integer, intent(in) :: ni, nj
  real(dp), dimension(1:ni,1:nj), intent(in) :: aw, ae, as, an, ap, su
  real(dp), dimension(1:ni,1:nj), intent(inout) :: phi
  call lin_solve(aw, ae, as, an, ap, su, phi, 0, 1, 100, -1.0_dp)
end subroutine
#endif
