/*
 * cfd_b200.h -- C ABI of libcfd_b200.so: the B200 (sm_100a) implementation of the data-parallel hot path of
 * brli3/CFD (staggered-grid FVM SIMPLE outer iteration + 5-point FD heat stencils).
 *
 * The reference is serial Fortran 90 with no FFI surface: its hot routines take no arguments and talk through
 * module-global allocatable arrays.  Each entry point below therefore names the reference
 * routine (path:line under the reference tree) whose work it replaces; the Fortran iso_c_binding interface a
 * maintainer would add is in INTEGRATION.md.  Plain C types only: pointers, ints, doubles.  Every function
 * returns CFD_OK (0) or a negative cfd_status; no exceptions, no host callbacks, one thread per context.
 *
 * Host arrays are the reference's own: real(dp) A(ld,nj), column-major, 1-based, A(i,j) at a[(i-1)+(j-1)*ld].
 * All (i,j) arguments are 1-based like the Fortran.  Device storage is private to the context (a line-major,
 * chunk-interleaved layout); the only way data crosses is cfd_upload / cfd_download.
 */
#ifndef CFD_B200_H
#define CFD_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cfd_ctx cfd_ctx;

typedef enum cfd_status {
  CFD_OK = 0,
  CFD_ERR_INVALID = -1,      /* bad argument / size / id */
  CFD_ERR_CUDA = -2,         /* a CUDA runtime call failed; cfd_last_error() has the text */
  CFD_ERR_NO_DEVICE = -3,    /* no CUDA device: there is NO CPU fallback */
  CFD_ERR_UNSUPPORTED = -4,  /* valid request outside what this build implements */
  CFD_ERR_NOMEM = -5,
  CFD_ERR_DIVERGED = -6      /* cfd_iterate: a residual sum came back NaN/Inf (the sums are |.|-sums over all cells, so any
                                non-finite field value shows up in them); the fields are left as they are */
} cfd_status;

/* Which reference program family the context mirrors. */
typedef enum cfd_family {
  CFD_FAMILY_CAVITY = 0,   /* lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90 (uniform staggered grid)       */
  CFD_FAMILY_GENERAL = 1,  /* buoyancy-driven-cavity-2d/buoyancy_cav.f90 + cylinder-heat-conduction-2d vars/serv */
  CFD_FAMILY_TEACH = 2     /* backward-step-flow-2d/calc.f90 + promod.f90 + general.f90                       */
} cfd_family;

/* Line ordering of the TDMA sweeps (lisolv, cavity_fvm.f90:62). */
typedef enum cfd_order {
  CFD_ORDER_REDBLACK = 0,  /* two colour passes over the S-N lines (north_star's variant; same fixed point)  */
  CFD_ORDER_SEQUENTIAL = 1 /* the reference's W->E line Gauss-Seidel, bit-comparable debug mode (small grids) */
} cfd_order;

/* 2-D fields, module vars_mod (cavity_fvm.f90:22-41) / vars (cylin_cond.f90:42-71). */
typedef enum cfd_field {
  CFD_U = 0, CFD_V, CFD_P, CFD_PP, CFD_VIS /* = gam */, CFD_DEN, CFD_T, CFD_CP,
  CFD_AE, CFD_AW, CFD_AN, CFD_AS, CFD_AP, CFD_SU, CFD_SP, CFD_DU, CFD_DV,
  /* general family: what gamsor() (cylin_cond.f90:149-169) fills for the scalar equation nq = 4 */
  CFD_GAMT,   /* diffusion coefficient of T (gam when nq == 4), boundary rows included (0 = adiabatic) */
  CFD_SUSRC,  /* su per unit volume */
  CFD_SPSRC,  /* sp per unit volume */
  /* TEACH family with cfd_set_turbulence: te, ed, gen, gamh of module general (general.f90:30-40) */
  CFD_TE, CFD_ED, CFD_GEN, CFD_GAMH,
  CFD_FIELD_COUNT
} cfd_field;

/* 1-D geometry arrays set by setgrid (cavity_fvm.f90:366-414) / grid (buoyancy_cav.f90:28-120). */
typedef enum cfd_geom {
  CFD_G_X = 0, CFD_G_XU, CFD_G_SEW, CFD_G_SEWU,          /* length ni */
  CFD_G_Y, CFD_G_YV, CFD_G_SNS, CFD_G_SNSV,              /* length nj */
  /* general family extras (cylin_cond.f90:57-61) */
  CFD_G_XDIF, CFD_G_XCVI, CFD_G_XCVIP, CFD_G_FX, CFD_G_FXM,                         /* length ni */
  CFD_G_YDIF, CFD_G_YCVR, CFD_G_YCVRS, CFD_G_ARX, CFD_G_ARXJ, CFD_G_ARXJP,
  CFD_G_R, CFD_G_RMN, CFD_G_SX, CFD_G_SXMN, CFD_G_FV, CFD_G_FVP, CFD_G_FY, CFD_G_FYM, /* length nj */
  /* TEACH family extras, computed by init (backward-step-flow-2d/general.f90:69-140) */
  CFD_G_DXEP, CFD_G_DXPW, CFD_G_DXEPU, CFD_G_DXPWU,                                 /* length ni */
  CFD_G_DYNP, CFD_G_DYPS, CFD_G_DYNPV, CFD_G_DYPSV, CFD_G_RV, CFD_G_RCV,            /* length nj */
  CFD_GEOM_COUNT
} cfd_geom;

/* Equation slots for urf[] / nswp[] / residuals: nq of module vars (cylin_cond.f90:47). */
enum { CFD_EQ_U = 0, CFD_EQ_V = 1, CFD_EQ_P = 2, CFD_EQ_T = 3, CFD_EQ_TE = 4, CFD_EQ_ED = 5, CFD_EQ_COUNT = 8 };

/*
 * The literals `program cavity_fvm` sets before its loop (cavity_fvm.f90:519-552), as one struct.
 * Slab fields (goff, ni_global, halo_w, halo_e) describe an i-slab of a larger grid for multi-GPU runs
 * (one context per rank); a single-GPU context leaves them 0.
 */
typedef struct cfd_config {
  int family;              /* cfd_family */
  int ni, nj;              /* local array extents INCLUDING boundary nodes (and halo lines, for a slab) */
  int iadv;                /* 1 CDS, 2 UDS, 3 hybrid (cavity_fvm.f90:30); 4 power law (general family) */
  int order;               /* cfd_order */
  int nswp[CFD_EQ_COUNT];  /* nswpu, nswpv, nswpp, (nsweep(4)) ... */
  double urf[CFD_EQ_COUNT];/* urfu, urfv, urfp, ... */
  int ipref, jpref;        /* 1-based GLOBAL indices of the pressure reference cell (cavity_fvm.f90:549-550) */
  int device;              /* CUDA device ordinal; -1 = the calling thread's current device */
  int tdma_m;              /* elements per thread in the line solver, 0 = choose */
  /* slab description */
  int goff;                /* global i of local line 1, minus 1 */
  int ni_global;           /* 0 = ni */
  int halo_w, halo_e;      /* halo lines at the west / east end of the local arrays (0 at a physical wall) */
  double sew_nim1;         /* value of sew(nim1) of the GLOBAL grid for the stale-scalar quirk of calcv
                              (cavity_fvm.f90:283); 0 = read it from the local geometry */
  /* general family: lsolve(1) and lsolve(4) of module vars (cylin_cond.f90:54, :100-101) */
  int solve_flow, solve_t;
  /* TEACH family (backward-step-flow-2d/control.f90): indcos (1 plane, 2 axisymmetric, :66), istep / jstep
     (:16,:67), laminar viscosity viscos (:126), inlet mass flow flowin (:197-203), great (:15) */
  int indcos, istep, jstep;
  double viscos, flowin, great;
  /* j-slab description (cavity family, red/black order): the S-N lines themselves are cut, each context holds rows
     joff+1 .. joff+nj of every line and solves them with the partition method (cfd_spike_*, below) */
  int joff;                /* global j of local row 1, minus 1 */
  int nj_global;           /* 0 = nj */
  int halo_s, halo_n;      /* halo rows at the south / north end of the local arrays (0 at a physical wall) */
  double sns_njm1;         /* sns(njm1) of the GLOBAL grid for the stale-scalar quirk of calcu (cavity_fvm.f90:186);
                              0 = read it from the local geometry */
  int raw_coefficients;    /* 1: keep the six coefficient arrays as the reference stores them also in a red/black context
                              (needed by cfd_lin_* and by hosts that upload their own coefficients) */
} cfd_config;

/* Device-side form of the `user` module hook gamsor() (cylin_cond.f90:149-169; the buoyancy program's own
   module is missing from the reference): enumerated source terms evaluated inside the kernels. */
typedef enum cfd_hook_kind {
  CFD_HOOK_NONE = 0,
  CFD_HOOK_BOUSSINESQ = 1,  /* eq = CFD_EQ_V: su = params[0]*(ARIMEAN(t(i,j-1),t(i,j),fy(j)) - params[1])
                               per unit volume, params[0] = rho*g*beta, params[1] = Tref */
  CFD_HOOK_FIELD_SOURCES = 2 /* eq = CFD_EQ_T: su, sp per unit volume are the uploaded fields CFD_SUSRC, CFD_SPSRC */
} cfd_hook_kind;
typedef struct cfd_hook_desc {
  int kind;
  double params[4];
} cfd_hook_desc;

/* ---- life cycle ------------------------------------------------------------------------------------- */
/* replaces array_alloc / array_dealloc (cavity_fvm.f90:432-505): fields are zero-initialised on the device */
int cfd_create(cfd_ctx** out, const cfd_config* cfg);
int cfd_destroy(cfd_ctx* ctx);
const char* cfd_strerror(int status);
const char* cfd_last_error(void);
/* Kernels of this context are launched on `cuda_stream` (a cudaStream_t passed as void*); default: stream 0. */
int cfd_set_stream(cfd_ctx* ctx, void* cuda_stream);
int cfd_synchronize(cfd_ctx* ctx);

/* ---- data movement (the only host<->device traffic) ------------------------------------------------- */
/* geometry of setgrid (cavity_fvm.f90:366-414): n doubles, n == ni or nj by id */
int cfd_set_geometry(cfd_ctx* ctx, int geom_id, const double* host, int n);
/* field A(ld,nj) host -> device / device -> host; replaces direct access to the module arrays by the driver
   (initial values cavity_fvm.f90:534-538, output 590-607) */
int cfd_upload(cfd_ctx* ctx, int field, const double* host, int ld);
int cfd_download(cfd_ctx* ctx, int field, double* host, int ld);
/*
 * cfd_upload without any run-time allocation.  (No counterpart in the reference: it exists for hosts of this
 * library that cannot call cudaMalloc after start-up -- a driver whose own runtime has claimed all free device
 * memory, or a solver loop that re-uploads boundary/source data every iteration and cannot afford cudaMalloc's
 * device-wide synchronisation.)  Same arguments and result as cfd_upload, but the column-major image is staged
 * in a 32 MiB buffer that is part of the library image on every GPU that loads it, so it calls no cudaMalloc /
 * cudaFree.  Limit: ni*nj <= 4 194 304 nodes (e.g. 2048 x 2048); larger grids get CFD_ERR_INVALID with a
 * message -- use cfd_upload.  The buffer's device address is looked up (cudaGetSymbolAddress), and with CUDA's
 * lazy module loading its memory is materialised, by cfd_create on the context's device -- the one call that
 * allocates anyway -- so no call of cfd_upload_noalloc itself ever reserves memory (with CUDA_MODULE_LOADING=EAGER
 * it exists from library load).  All contexts on one GPU share the buffer: calls that overlap on the GPU
 * (different streams or host threads) race, so issue them one at a time.
 */
int cfd_upload_noalloc(cfd_ctx* ctx, int field, const double* host, int ld);
/* monitor-point read u(imon,jmon) etc. (cavity_fvm.f90:559-563); LOCAL 1-based indices */
int cfd_probe(cfd_ctx* ctx, int field, int i, int j, double* value);
/* ---- post-processing on the device fields: result files and monitors without whole-field downloads --------------
   cfd_extract_line / cfd_extract_row   A(i, 1:nj) / A(1:ni, j) of a plain field in natural order (n = nj / ni), LOCAL indices
   cfd_centreline          which = 0: uc(index, 1:nj), which = 1: vc(1:ni, index) of centred_vel (cavity_fvm.f90:416-430;
                           postproc buoyancy_cav.f90:441-449 has the same formulas) -- the columns of uvel.txt / vvel.txt
                           (cavity_fvm.f90:600-607)
   cfd_reference_pressure  p(:,:) = p(:,:) - p(ipref,jpref) and the four corner values (cavity_fvm.f90:568-573), in place */
int cfd_extract_line(cfd_ctx* ctx, int field, int i, double* host, int n);
int cfd_extract_row(cfd_ctx* ctx, int field, int j, double* host, int n);
int cfd_centreline(cfd_ctx* ctx, int which, int index, double* host, int n);
int cfd_reference_pressure(cfd_ctx* ctx);
/* raw device view for peer/NCCL halo exchange: base pointer of `field`, doubles per line, number of lines.
   Line i (local, 1-based) is the contiguous range base[(i-1)*line_stride .. +line_stride). */
int cfd_field_device_view(cfd_ctx* ctx, int field, void** base, long long* line_stride, int* nlines);

/* diagnostic: the device line layout chosen for this context (nodes per thread, threads per line, padded
   line length); no counterpart in the reference */
int cfd_layout_info(cfd_ctx* ctx, int* m, int* t, int* ld);
/* diagnostic: number of CUDA kernels this context has launched since cfd_create (bench.py's gpu_launches) */
long long cfd_launch_count(cfd_ctx* ctx);

/* ---- the hot path: one call per reference subroutine -------------------------------------------------- */
/* calcu (cavity_fvm.f90:92-192): assemble + nswpu line sweeps of u; *resoru receives the residual sum */
int cfd_calcu(cfd_ctx* ctx, double* resoru);
/* calcv (cavity_fvm.f90:194-291) */
int cfd_calcv(cfd_ctx* ctx, double* resorv);
/* calcp (cavity_fvm.f90:293-355): p' assembly, pp=0, nswpp sweeps, correction of p,u,v */
int cfd_calcp(cfd_ctx* ctx, double* resorm);
/* calct (buoyancy_cav.f90:358-433): scalar (temperature) equation of the general family */
int cfd_calct(cfd_ctx* ctx, double* resort);
/* install the device-side equivalent of a gamsor() source hook for equation eq (CFD_EQ_*) */
int cfd_set_hook(cfd_ctx* ctx, int eq, const cfd_hook_desc* desc);
/* lisolv(istart,jstart,ni,nj,phi) (cavity_fvm.f90:49-82): ONE sweep over the current coefficients */
int cfd_lisolv(cfd_ctx* ctx, int istart, int jstart, int field);
/* n passes of the loop body cavity_fvm.f90:554-557 without returning to the host; res[0..2] = resoru,resorv,
   resorm of the LAST pass (res[3..7] reserved for T, k, eps ...) */
int cfd_iterate(cfd_ctx* ctx, int n, double res[8]);

/* ---- stage-level entry points (what cfd_calc* are made of; used by the multi-GPU driver, which exchanges
        halo lines between them).  None of them synchronises or touches the host. ------------------------ */
int cfd_assemble_u(cfd_ctx* ctx);                       /* cavity_fvm.f90:104-187 */
int cfd_assemble_v(cfd_ctx* ctx);                       /* cavity_fvm.f90:205-285 */
int cfd_assemble_p(cfd_ctx* ctx);                       /* cavity_fvm.f90:301-342 */
int cfd_assemble_t(cfd_ctx* ctx);                       /* buoyancy_cav.f90:371-428 (general family); TEACH after
                                                           cfd_set_turbulence: calct chapters 1-3, calc.f90:309-385 */
/* half sweep over the lines of one colour (0/1), or a full sequential sweep if colour < 0;
   first != 0 treats phi as zero where the reference has just executed pp(:,:)=0 (cavity_fvm.f90:342) */
int cfd_sweep(cfd_ctx* ctx, int istart, int jstart, int field, int colour, int first);
/* one colour pass in two launches: part 1 = the first / last owned line of an i-slab (if it has this colour and a
   neighbour on that side), part 2 = the other lines of the colour, part 0 = all (= cfd_sweep).  Lets a multi-GPU
   driver send the edge lines while the interior is still being solved. */
int cfd_sweep_part(cfd_ctx* ctx, int istart, int jstart, int field, int colour, int first, int part);
int cfd_correct(cfd_ctx* ctx, const double* ppref_device); /* cavity_fvm.f90:346-354; NULL = pp(ipref,jpref) */
/* the `do n = 1, nswp; call lisolv(...)` loop of calcu / calcv / calcp (cavity_fvm.f90:189-191, 287-289, 343-345):
   nsweeps red/black sweeps of `field` over the current coefficients.  Where the layout allows it (lines of 128..320
   layout threads, whole lines per context) all 2*nsweeps colour passes run in ONE persistent launch ordered by per-line
   completion flags instead of kernel boundaries; an i-slab whose neighbours are attached (cfd_peer_attach) also stores
   its edge lines into the neighbours' halo lines from inside that launch.  Same bits as nsweeps x cfd_sweep(0), cfd_sweep(1). */
int cfd_sweeps(cfd_ctx* ctx, int istart, int jstart, int field, int nsweeps, int first);

/* ---- i-slab neighbours over peer memory (multi-GPU; no counterpart in the reference, which is serial) -------------
   Every slab context exports its plain fields and its flag block (cfd_ipc_export: a 64-byte CUDA IPC handle), the
   neighbouring ranks open them (cfd_ipc_open -> a device pointer valid in this process) and attach them
   (cfd_peer_attach: side 0 = west neighbour, 1 = east neighbour; `what` = a cfd_field or CFD_IPC_FLAGS; the neighbour's
   local extents tell where this context's edge lines land in its arrays).  In one process with several GPUs the
   neighbour's own device pointers (cfd_field_device_view) can be attached directly.  After that
     cfd_halo_push   stores the first `e[k]` / last `w[k]` owned lines of up to 4 fields into the west / east neighbour's
                     halo lines and waits for the lines the neighbours store here: one small kernel, remote STORES and
                     flag polls only -- replaces a grouped NCCL send/recv on the stream
     cfd_sweeps      hands the edge lines over from inside the sweep kernel (above)
   All ranks must issue the same sequence of cfd_sweeps / cfd_halo_push calls (the handshakes count calls).  Waits are
   bounded (about two seconds); cfd_peer_status reports whether one ever timed out.  Loads by bulk copy (TMA) in
   these kernels touch this GPU's memory only. */
#define CFD_IPC_FLAGS (-1)
int cfd_ipc_export(cfd_ctx* ctx, int what, void* handle64);
int cfd_ipc_open(cfd_ctx* ctx, const void* handle64, void** devptr);
int cfd_peer_attach(cfd_ctx* ctx, int side, int what, void* devptr, int peer_ni, int peer_halo_w, int peer_halo_e);
int cfd_peer_detach(cfd_ctx* ctx);
int cfd_halo_push(cfd_ctx* ctx, int nfields, const int* fields, const int* w, const int* e);
int cfd_peer_status(cfd_ctx* ctx, int* aborted);
/* pp(ipref,jpref) to every rank without a collective: the rank that owns the cell attaches every rank's flag block
   (cfd_bcast_attach(ctx, rank, ptr, that rank's ni); ptr = NULL for itself); cfd_scalar_bcast(ctx, owner, field, i, j, &v)
   -- called by ALL ranks, owner != 0 on the owning one -- stores the value into every rank's mailbox (peer stores), and
   returns in v the device address where this rank will find it; the wait for its arrival is queued on the stream.
   Pass v to cfd_correct. */
int cfd_bcast_attach(cfd_ctx* ctx, int rank, void* flag_block, int rank_ni);
int cfd_scalar_bcast(cfd_ctx* ctx, int owner, int field, int i, int j, void** value_dev);
/* *yes = 1 when cfd_sweeps runs as the single flag-ordered launch in this context (with its neighbours attached as they
   are now): an i-slab driver may only drop its per-pass halo exchanges in that case */
int cfd_sweeps_single_launch(cfd_ctx* ctx, int* yes);
/* diagnostic (environment CFD_B200_FLOW_TIMING=1): clock64 sums of the consumer phases of the cluster sweep kernel since
   the last call -- out[0] = lines solved by block 0, out[1..6] = cycles waiting for / eliminating the slices, in the
   backward elimination, the cyclic reduction, the cluster exchange, the stores, the end-of-line barrier */
int cfd_debug_flow_timing(cfd_ctx* ctx, double out[8]);

/* ---- TEACH family: k-epsilon model, energy equation, wall functions -------------------------------
   cfd_set_turbulence allocates te, ed, t, gen, gamh and switches on the wall-function branches of modu / modv
   (promod.f90:36-60, 105-121).  The literals are those `contro` sets (control.f90:105-141).  After it:
     cfd_calcte  = calcte (calc.f90:388-485) with modte (promod.f90:182-256)     residual slot CFD_EQ_TE
     cfd_calced  = calced (calc.f90:490-577) with moded (promod.f90:260-296)     residual slot CFD_EQ_ED
     cfd_calct   = calct  (calc.f90:309-385) with modt  (promod.f90:146-178)     residual slot CFD_EQ_T
     cfd_props   = props  (general.f90:165-190)
   and cfd_iterate runs calcu, calcv, calcp, calcte, calced, calct, props per pass (control.f90:210-217; each
   switched by incalk / incald / incalt / inpro).  The 1-D wall arrays taun, tauw, yplusn, xplusw (general.f90:45-46)
   live on the device; cfd_wall_upload / cfd_wall_download move them (n = ni for taun, yplusn; nj for tauw, xplusw).
   log / pow of the wall laws come from the CUDA math library: results agree with the reference arithmetic to
   rounding (1e-12 class), not bit for bit. */
typedef struct cfd_turbulence {
  double cmu, cd, c1, c2, cappa, elog, pred, prte, prandl, prandt, pfun, twalln;
  double urfk, urfe, urft, urfvis;
  int nswpk, nswpd, nswpt;
  int incalk, incald, incalt, inpro;
} cfd_turbulence;
typedef enum cfd_wall_array { CFD_WALL_TAUN = 0, CFD_WALL_TAUW = 1, CFD_WALL_YPLUSN = 2, CFD_WALL_XPLUSW = 3 } cfd_wall_array;
int cfd_set_turbulence(cfd_ctx* ctx, const cfd_turbulence* model);
int cfd_wall_upload(cfd_ctx* ctx, int which, const double* host, int n);
int cfd_wall_download(cfd_ctx* ctx, int which, double* host, int n);
int cfd_calcte(cfd_ctx* ctx, double* resork);
int cfd_calced(cfd_ctx* ctx, double* resore);
int cfd_props(cfd_ctx* ctx);
int cfd_assemble_te(cfd_ctx* ctx);
int cfd_assemble_ed(cfd_ctx* ctx);

/* ---- generic 5-point systems: the iterative line-by-line TDMA of the scalar-transport demos -------
   The context's coefficient fields (uploaded with cfd_upload, CFD_AE .. CFD_SU, raw storage) define
   ap*phi = ae*phiE + aw*phiW + an*phiN + as*phiS + su  on the interior nodes; boundary values are the Dirichlet
   data in `field`.  cfd_lin_wesweep = wesweep (stagnation-point-flow-2d/steady/fvm2d_stag_point.f90:164-196),
   cfd_lin_snsweep = snsweep (:198-230), both with `tdma` (:232-260, two divisions per row), cfd_lin_residual = the
   sum of :99-106, cfd_lin_solve = the loops tdma_we / tdma_sn (:82-162; variant 1: tdma_solver of
   diagonal-convection-2d/fvm2d_oblique_conv.f90:284-318, whose all-terms-on-the-left convention is obtained by
   uploading -ae, -aw, -an, -as).  Sequential-order contexts follow the reference's line order and arithmetic. */
int cfd_lin_wesweep(cfd_ctx* ctx, int field);
int cfd_lin_snsweep(cfd_ctx* ctx, int field);
int cfd_lin_residual(cfd_ctx* ctx, int field, double* res);
int cfd_lin_solve(cfd_ctx* ctx, int field, int dir, int variant, int maxit, double tol, int* iters, double* rsm);
/* time discretisation of stag_point_unsteady.f90:634-688 on the device: from the steady coefficients kept in four
   plain fields (base_fields = ae, aw, an, as), ct = den/dt*dx*dy (ct_field) and the old solution (phio_field) build
   ae .. su of the step: itsch 1 explicit Euler, 2 implicit Euler, 3 Crank-Nicolson; the neighbour coefficients are
   stored negated, which turns that program's all-terms-on-the-left solver convention (:159-261) into the form
   cfd_lin_* solve.  cfd_field_copy: phio = phi (:586). */
int cfd_lin_time_terms(cfd_ctx* ctx, int itsch, int phio_field, const int base_fields[4], int ct_field);
int cfd_field_copy(cfd_ctx* ctx, int dst_field, int src_field);

/* ---- j-slabs: lisolv's tridiagonal systems (cavity_fvm.f90:62-80) cut across contexts ----------------------
   Partition (Wang / SPIKE) method.  With x_s, x_n the still unknown values just outside a context's piece of a
   line, the piece's solution is  x = y + w*x_s + z*x_n :  y solves the piece with zero end values (that is what
   cfd_sweep does in a j-slab context), w and z are the spike vectors of the matrix alone.  The first and last
   unknown of every piece give a small reduced system per line; its six coefficients per piece are all-gathered
   (by the caller: torch.distributed / NCCL), every context solves the reduced systems redundantly and finishes
   its own pieces.  All buffers are DEVICE pointers; all calls are asynchronous on the context's stream.
     cfd_spike_basis    w, z of every line for the coefficients now in the context (once per assembly)
     cfd_spike_pack     which = 1: (w_first, w_last, z_first, z_last) of all ni local lines -> buf[ni][4]
                        which = 0: (y_first, y_last) of the lines of `colour` in `field`    -> buf[nlines][2]
     cfd_spike_finish   reduced solve + x = y + w*x_s + z*x_n on the lines of `colour`; also stores x_s, x_n in the
                        halo rows, so the field's halos are current afterwards.  wz_all = [world][ni][4],
                        y_all = [world][nlines][2] as gathered in rank order (rank = position along j).
     cfd_rows_pack / cfd_rows_unpack   rows j_first .. j_first+nrows-1 (1-based local) of `field` <-> buf[nrows][ni]
                        (a halo row is strided in the line layout) */
int cfd_spike_basis(cfd_ctx* ctx, int istart, int jstart);
int cfd_spike_pack(cfd_ctx* ctx, int istart, int jstart, int field, int colour, int which, double* buf);
int cfd_spike_finish(cfd_ctx* ctx, int istart, int jstart, int field, int colour, const double* wz_all,
                     const double* y_all, int rank, int world);
int cfd_rows_pack(cfd_ctx* ctx, int field, int j_first, int nrows, double* buf);
int cfd_rows_unpack(cfd_ctx* ctx, int field, int j_first, int nrows, const double* buf);
/* finish the deterministic residual reduction and copy res[0..7] to the host (synchronises) */
int cfd_residuals(cfd_ctx* ctx, double res[8]);
/* device address of the 8 packed residual sums (for an NCCL all-reduce) */
int cfd_residual_device_ptr(cfd_ctx* ctx, void** ptr);

/* ---- explicit / implicit 5-point heat conduction (finite-difference-method/heat-conduction/) ---------- */
typedef struct cfd_heat_config {
  int nx, ny;                    /* nodes including the Dirichlet edges */
  double k, rho, cp;             /* fdm1d_explicit_heat_cond.f90:46 */
  double xa, xb, ya, yb;
  double tleft, tright, tbottom, ttop;
  double dt;
  double qthr, qval;             /* QDOT (fdm1d_explicit_heat_cond.f90:76-88): qval where x < qthr else 0 */
  int device;
  int j_goff, ny_global;         /* j-slab of a larger grid (rows), 0 = whole grid */
  int halo_s, halo_n;
} cfd_heat_config;
typedef struct cfd_heat cfd_heat;
int cfd_heat_create(cfd_heat** out, const cfd_heat_config* cfg);
int cfd_heat_destroy(cfd_heat* h);
int cfd_heat_set_stream(cfd_heat* h, void* cuda_stream);
int cfd_heat_upload(cfd_heat* h, const double* host, int ld);      /* T(ld,ny) column-major */
int cfd_heat_download(cfd_heat* h, double* host, int ld);
/* nsteps passes of the update loop fdm1d_explicit_heat_cond.f90:130-139 with the 5-point operator of
   fdm2d_steady_heat_cond.f90:171-188 (constant conductivity) */
int cfd_heat2d_step(cfd_heat* h, int nsteps);
/* same update evaluated in the restatement's exact operation order with IEEE divides (bit-comparable with
   oracle/heat_oracle.c; slower -- for parity debugging) */
int cfd_heat2d_step_exact(cfd_heat* h, int nsteps);
/* nsteps implicit-Euler steps, locally one-dimensional: the tridiagonal assembly of
   implicit-time/fdm1d_explicit_heat_cond.f90:122-143 along x then y, batched Thomas (thomas, :147-184) */
int cfd_heat2d_implicit_step(cfd_heat* h, int nsteps);
/* same, with IEEE divisions in the reference's operation order (bit-comparable with the oracle; slower) */
int cfd_heat2d_implicit_step_exact(cfd_heat* h, int nsteps);
/* Row slabs (cfd_heat_config.j_goff / ny_global) and implicit steps: the y-lines cross the slabs, but both recurrences of
   `thomas` forget their start at a fixed rate per node, so a slab with `rows` halo rows towards each neighbour -- refreshed
   from the neighbours' owned rows before every step -- reproduces the whole-grid result to the last bit or two.  `rows`
   depends on Fo_y = k dt / (rho cp dy^2) only (e.g. 128 at Fo = 5); cfd_heat2d_implicit_step refuses slabs with less. */
int cfd_heat_implicit_halo(double fo_y, int* rows);
/* ---- variable conductivity: the 5-point operator of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:162-190 ----
   lamw, lame, lams, lamn = LAMBDA at the four face mid-points of every node, (x-dx/2,y), (x+dx/2,y), (x,y-dy/2),
   (x,y+dy/2) (:171-174), as (ld,ny) host arrays; qdot = QDOT at the nodes (:188; NULL = none).  The device keeps
   aw = lamw/dx**2, ae, as = lams/dy**2, an (:176-179); ap = -(aw+ae+as+an) (:180) is formed on the fly.
     cfd_heat2d_apply      r = aw*T(i-1,j) + ae*T(i+1,j) + as*T(i,j-1) + an*T(i,j+1) + ap*T(i,j) + qdot on the interior
                           nodes, 0 on the boundary: the residual of the steady equations, to a host array
     cfd_heat2d_step_var   nsteps of the explicit update (fdm1d_explicit_heat_cond.f90:133-135) with that operator,
                           T += (operator/rho/cp)*dt, Dirichlet edges rewritten every step; 56 B/cell/step
     cfd_heat2d_steady_gs  gauss_seidel (:248-287) on the rows interior() / boundary() define, matrix-free (the reference
                           builds the dense (nx*ny)^2 matrix): x2 = (b - s)*rf/a_kk + (1-rf)*x1, stop when |x1-x2|_2 < tol
                           or after maxit passes (reference: rf 0.5, tol 1e-7, 500).  Nodes are visited in two colours
                           ((i+j) even, then odd) instead of one lexicographic order: same fixed point, different
                           iterates.  Starts from the field in the handle (the reference starts from zero). */
int cfd_heat_set_operator(cfd_heat* h, const double* lamw, const double* lame, const double* lams, const double* lamn,
                          const double* qdot, int ld);
int cfd_heat2d_apply(cfd_heat* h, double* host_r, int ld);
int cfd_heat2d_step_var(cfd_heat* h, int nsteps);
int cfd_heat2d_steady_gs(cfd_heat* h, int maxit, double tol, double rf, int* iters, double* dx2);
int cfd_heat_device_view(cfd_heat* h, void** cur, long long* row_stride, int* nrows);
int cfd_heat_synchronize(cfd_heat* h);

/* ---- explicit finite-difference cavity: stencil passes of lid-driven-cavity-2d/fdm-explicit/
        fdm_explicit_cavity.f90 (cfd_fdcav_solve_pp is the device SIP of SURVEY 8f-1) ------ */
typedef enum cfd_fdc_field {
  CFD_FDC_U = 0, CFD_FDC_V, CFD_FDC_P, CFD_FDC_UO, CFD_FDC_VO, CFD_FDC_PO, CFD_FDC_PP, CFD_FDC_SU,
  CFD_FDC_AE, CFD_FDC_AW, CFD_FDC_AN, CFD_FDC_AS, CFD_FDC_AP,   /* coefficient arrays of the implicit variant */
  CFD_FDC_FIELD_COUNT
} cfd_fdc_field;
typedef struct cfd_fdcav_config {
  int ni, nj;                 /* grid nodes (fdm_explicit_cavity.f90:111-112); arrays are (ni+1, nj+1) */
  double re, dt, utop, urfp;  /* :116-119 */
  int device;
  /* implicit variant (lid-driven-cavity-2d/fdm-implicit/fdm_implicit_cavity.f90:105-125) */
  int steady;                 /* steadysim */
  double urfu, urfv;
} cfd_fdcav_config;
typedef struct cfd_fdcav cfd_fdcav;
int cfd_fdcav_create(cfd_fdcav** out, const cfd_fdcav_config* cfg);
int cfd_fdcav_destroy(cfd_fdcav* h);
/* A(ld, nj+1) column-major, ld >= ni+1; field = cfd_fdc_field */
int cfd_fdcav_upload(cfd_fdcav* h, int field, const double* host, int ld);
int cfd_fdcav_download(cfd_fdcav* h, int field, double* host, int ld);
/* u, v momentum predictors (:170-192), pp = 0 and the p' source su (:194-205) */
int cfd_fdcav_predict(cfd_fdcav* h);
/* correction of p, u, v with the caller-solved pp (:208-222), boundary conditions (:224-237), continuity norm
   `error` (:239-247), uo=u, vo=v, po=p (:252-254) */
int cfd_fdcav_correct(cfd_fdcav* h, double* error);
/* sipsol (fdm_explicit_cavity.f90:24-87: Stone's SIP, maxit 600, alpha 0.88, tol 1e-5) on the p' system set up by
   cfd_fdcav_predict, on the device; *iters receives the iterations used */
int cfd_fdcav_solve_pp(cfd_fdcav* h, int* iters);
/* one pass of the time loop (:168-255): predict, solve pp, correct */
int cfd_fdcav_step(cfd_fdcav* h, double* error, int* sip_iters);
/* implicit-Euler variant, fdm_implicit_cavity.f90: coefficient assembly of u (:178-198), v (:201-220), p' (:222-240)
   [which = 0, 1, 2]; sipsol (:29-92, maxit 3000, alpha 0.80, tol 1e-6) on that system; correction + boundary
   conditions + continuity norm (:243-282); and the whole loop body (:177-288) */
int cfd_fdcav_implicit_assemble(cfd_fdcav* h, int which);
int cfd_fdcav_implicit_solve(cfd_fdcav* h, int which, int* iters);
int cfd_fdcav_implicit_correct(cfd_fdcav* h, double* error);
int cfd_fdcav_implicit_step(cfd_fdcav* h, double* error, int iters[3]);
/* sipsol itself as a drop-in on host arrays: `call sipsol(1, i2, 1, j2, aw, ae, as, an, ap, su, phi)` of
   fdm_implicit_cavity.f90:29-92 (alpha 0.80, tol 1e-6, maxit 3000) and fdm_explicit_cavity.f90:24-87 (0.88, 1e-5, 600).  All
   arrays are (ld, j2) column-major, the unknowns are the interior 2..i2-1 x 2..j2-1, phi is in/out.  negate != 0: the four
   neighbour arrays are the program's own and the solver uses their negatives, as in `call sipsol(.., -aw, -ae, -as, -an, ..)`
   (no negated copies on the host).  Uploads, solves in one kernel, downloads phi; work space is allocated per call
   (device < 0: the current device).  iters / rsm (may be NULL) return the iterations used and the last residual ratio. */
int cfd_sipsol(int device, int i2, int j2, int ld, const double* aw, const double* ae, const double* as, const double* an,
               const double* ap, const double* su, double* phi, int negate, double alpha, double tol, int maxit, int* iters,
               double* rsm);

/* library/build identification */
const char* cfd_version(void);

#ifdef __cplusplus
}
#endif
#endif /* CFD_B200_H */
