// common.cuh -- internal declarations shared by the .cu files of libcfd_b200.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/cfd_b200.h"

namespace cfd {

// ----------------------------------------------------------------------------------------------------
// Device layout of a 2-D field ("line layout").
//
// The reference stores A(i,j) column-major (i fastest) and solves tridiagonal systems along j for one i at a
// time (lisolv, cavity_fvm.f90:62-80).  On the device every S-N line i is one contiguous run of LD = T*M
// doubles, and inside the line the nj nodes are dealt out to T threads in chunks of M consecutive nodes,
// stored chunk-interleaved:   node j0 (0-based) of line i0 lives at   i0*LD + (j0 % M)*T + (j0 / M).
// Thread t of a line owns nodes t*M .. t*M+M-1; its k-th node sits at k*T + t, so a warp reading "its k-th
// node" touches 32 consecutive doubles (fully coalesced) while each thread still holds a contiguous piece of
// the line in registers -- which is what the partitioned Thomas solver needs.
// ----------------------------------------------------------------------------------------------------
struct Layout {
  int ni, nj;   // logical extents (lines, nodes per line)
  int M, T;     // nodes per thread, threads per line (T multiple of 32)
  int LD;       // T*M >= nj
  // The six coefficient arrays of one equation (an, as, ap, ae, aw, su -- in that order) are interleaved so that
  // the line solver can fetch "node k of every thread, all six arrays" of a line as ONE contiguous 6*T block:
  //   coefficient a, line i0, node (k,t)  at  coef + i0*cSL + k*cSK + a*T + t,   cSL = 6*LD, cSK = 6*T.
  // In production (red/black) contexts the equations are stored already divided by ap (an/ap, as/ap, ae/ap,
  // aw/ap, su/ap: nc = 5 arrays, ap itself is not kept): one array less to write per assembly and to read per
  // sweep.  Bit-exact (sequential-order) contexts keep the reference's six raw arrays (nc = 6).
  long long cSL;   // nc*LD
  int cSK;         // nc*T
  int nc;
};

__host__ __device__ __forceinline__ int lpos(int j0, int M, int T) { return (j0 % M) * T + (j0 / M); }

#define CFD_MAX_M 16
// res_dev: the 8 residual sums cfd_residuals / cfd_iterate hand to the host, then TEACH's outlet increment uinc
#define CFD_RES_SLOTS 9

// Everything a cavity-family kernel needs, passed by value.
struct CavParams {
  Layout L;
  int ia, ib;            // 0-based inclusive range of local lines this launch works on
  int jlo, jhi;          // 0-based inclusive node range (j) this launch works on
  int iadv;
  double urf;
  double stale_area;     // sns(njm1) for calcu, sew(nim1) for calcv (stale-scalar quirk, cavity_fvm.f90:186,283)
  const double *u, *v, *p, *den, *vis, *du_in, *dv_in;
  double *ae, *aw, *an, *as, *ap, *su, *dout;
  // geometry: by-i arrays in natural order, by-j arrays in line layout (one "line" of LD doubles each)
  const double *x, *xu, *sew, *sewu;
  const double *yL, *yvL, *snsL, *snsvL;
  double* partial;       // one residual partial per block
  // uniform-property form (production kernels): den and vis hold one value each in every node of this context, so the
  // kernels take them as scalars instead of streaming the two fields (same arithmetic on the same values: same bits)
  int uniform;
  double den0, vis0;
};

// Reciprocals of the 1-D geometry for the production assembly kernels (cavity_fast_kernels.cu); 0 where the
// denominator is 0 (never read there).  by-i arrays natural order, by-j arrays in line layout.
struct FastGeom {
  const double *rdxu;    // 1/(xu(i+1)-xu(i))
  const double *rdx;     // 1/(x(i+1)-x(i))
  const double *rsew;    // 1/sew(i)
  const double *rsewu;   // 1/sewu(i)
  const double *rdyL;    // 1/(y(j+1)-y(j))
  const double *rdyvL;   // 1/(yv(j+1)-yv(j))
  const double *rsnsL;   // 1/sns(j)
  const double *rsnsvL;  // 1/snsv(j)
};

struct SweepParams {
  Layout L;
  int i_first;           // 0-based local index of the first line of this colour
  int i_step;            // 2 for a colour pass
  int nlines;            // number of lines in this launch
  int jlo, jhi;          // 0-based inclusive range of unknowns along the line (jstart-1 .. nj-2)
  int first;             // 1: neighbours' phi taken as zero (pp = 0 folded in)
  int first_colour;      // for first==1: 0 -> both neighbours zero, 1 -> neighbours are freshly written
  const double *ae, *aw, *an, *as, *ap, *su;
  double* phi;
  // partition method across j-slabs (jslab_kernels.cu): the Dirichlet end values phi(jlo-1), phi(jhi+1) may be
  // constants instead of the stored values, the right-hand side may be dropped (spike vectors depend on the
  // matrix only), and the solution may go to another array than phi.
  int basis;                 // 1: rhs = 0
  int end_s_const, end_n_const;
  double end_s, end_n;
  double* out;               // solution array (same layout as phi); phi itself unless a spike vector is built
};

// All colour passes of an equation's sweeps in one launch (tdma_flow_kernels.cu).
struct FlowParams {
  Layout L;
  int npass;                 // colour passes in this launch; pass p has colour (colour0 + p) & 1 and epoch epoch0 + p
  int colour0;
  int i_first[2], nlines[2]; // per colour: first local line (0-based) and number of lines (every second line)
  int jlo, jhi;              // unknowns along a line
  int n_zero_ew;             // pp = 0 folded into the first sweep: the first n_zero_ew passes take both neighbour lines as zero,
  int n_zero_ends;           // the first n_zero_ends passes the Dirichlet end values (full sweeps: 1 and 2)
  const double* coef;        // interleaved coefficient block (array `an`)
  double* phi;
  int* flags;                // per local line: epoch of the last pass that finished it
  int epoch0;
  int wait_lo[2], wait_hi[2];   // per colour: local lines of that colour whose flag a later pass may wait for (owned lines and,
                                // with a neighbouring GPU attached, the halo line that GPU rewrites)
  int edge_line[2];          // first / last owned line if it is swept and a neighbour is attached on that side, else -1
  double* peer_phi[2];       // that line's image in the neighbour's array (its halo line), peer memory
  int* peer_flag[2];         // the neighbour's flag of that halo line
  int* peer_ready[2];        // the neighbour's "ready" slot for this GPU (written once at kernel start)
  const int* ready;          // my ready slots [from west, from east]
  int halo_line[2];          // my west / east halo line that the neighbour rewrites during this launch (-1: none) ...
  int halo_last[2];          // ... and the epoch of the last pass that does: the launch does not end before it has arrived
  int* abort_flag;
  int* arrive;               // lines split over a cluster: per-line count of the blocks that have stored their part
  unsigned long long* dbg;   // optional (CFD_B200_FLOW_TIMING=1): clock64 sums of the consumer phases, per kernel
};

struct HaloPush {
  int nmsg;
  const double* src[8];
  double* dst[8];
  long long count[8];        // doubles
  int xe;                    // exchange epoch
  int* peer_xready[2];       // neighbour's slots for this GPU
  int* peer_xdata[2];
  const int* xready;         // my slots [from west, from east]
  const int* xdata;
  int* abort_flag;
};

// slots behind the ni per-line flags of a context's flag block
#define CFD_FLAG_XDATA 0         /* [2] halo-push data marks from the west / east neighbour */
#define CFD_FLAG_XREADY 2        /* [2] halo-push ready marks */
#define CFD_FLAG_SWEEP_READY 4   /* [2] sweep-kernel ready marks */
#define CFD_FLAG_ABORT 6
#define CFD_FLAG_BCAST 7         /* epoch of the last scalar broadcast that has arrived (cfd_scalar_bcast) */
#define CFD_FLAG_SLOTS 8
#define CFD_MAILBOX_DOUBLES 8    /* behind the flag block: 4 slots of received scalars (by epoch & 3), rest reserved */
#define CFD_BCAST_MAX 64

}  // namespace cfd

// ---- error plumbing -----------------------------------------------------------------------------------
void cfd_set_error(const char* fmt, ...);
#define CFD_CUDA(call)                                                                              \
  do {                                                                                              \
    cudaError_t e__ = (call);                                                                       \
    if (e__ != cudaSuccess) {                                                                       \
      cfd_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));         \
      return CFD_ERR_CUDA;                                                                          \
    }                                                                                               \
  } while (0)
#define CFD_REQUIRE(cond, msg)                                   \
  do {                                                           \
    if (!(cond)) {                                               \
      cfd_set_error("%s:%d: %s", __FILE__, __LINE__, msg);       \
      return CFD_ERR_INVALID;                                    \
    }                                                            \
  } while (0)

// ---- context ------------------------------------------------------------------------------------------
struct cfd_ctx {
  cfd_config cfg;
  cfd::Layout L;
  int device;
  cudaStream_t stream;
  double* field[CFD_FIELD_COUNT];   // line layout, ni*LD doubles each (allocated on demand by family);
                                    // AN,AS,AP,AE,AW,SU point into `coef` (interleaved, see Layout)
  double* coef;                     // 6*ni*LD doubles
  double* geom[CFD_GEOM_COUNT];     // by-i: natural (ni doubles); by-j: line layout (LD doubles)
  double* geomn[CFD_GEOM_COUNT];    // every geometry array in natural order (general-family kernels)
  std::vector<double> geom_host[CFD_GEOM_COUNT];
  int boussinesq;                   // cfd_set_hook(CFD_EQ_V, CFD_HOOK_BOUSSINESQ)
  double rgbeta, tref;
  int has_t_sources;                // cfd_set_hook(CFD_EQ_T, CFD_HOOK_FIELD_SOURCES)
  double* stage;                    // ni*nj staging buffer for upload/download (column-major image)
  double* partial;                  // residual partial sums, CFD_EQ_COUNT * max_blocks
  int max_partials;
  double* res_dev;                  // 8 packed residual sums
  double* seq_scratch;              // sequential-order debug sweep: forward-elimination arrays
  double* gfast[8];                 // FastGeom arrays (device)
  bool turbulent;                    // cfd_set_turbulence was called (TEACH family)
  cfd_turbulence turb;
  double* wall[4];                   // taun, tauw, yplusn, xplusw (max(ni,nj)+2 doubles each)
  double* tmult_n;                   // wall-law multiplier per i (modu) / per j (modv), refreshed before each assembly
  double* tmult_w;
  double* spike_w;                   // j-slab partition method: spike vectors of the current equation (ni*LD each)
  double* spike_z;
  double* spike_small;               // x_s, x_n of the lines of one colour
  double* tfast[16];                 // TEACH-family reciprocal geometry (teach_fast_kernels.cu), built on first use
  bool tfast_ready;
  bool exact_assembly;              // bit-exact assembly kernels (sequential-order mode, or CFD_B200_EXACT_ASSEMBLY=1)
  long long launches;               // kernels launched by this context so far (cfd_launch_count)
  // sweeps ordered by per-line flags, i-slab neighbours over peer memory (tdma_flow_kernels.cu)
  int* flags;                       // ni per-line epochs + CFD_FLAG_SLOTS handshake slots (its own cudaMalloc: exportable)
  int epoch;                        // epoch of the next colour pass (the same on every rank of an SPMD run)
  int xepoch;                       // halo-push counter
  bool peer_attached[2];            // west / east neighbour's memory is mapped
  double* peer_field[2][CFD_FIELD_COUNT];
  int* peer_flags[2];
  int peer_ni[2], peer_hw[2], peer_halo_line[2];
  // one-to-all broadcast of a scalar (pp(ipref,jpref)) over peer memory: the owner holds every rank's flag block
  int bcast_n;                       // ranks attached (owner only), 0 = none
  int* bcast_flags[CFD_BCAST_MAX];   // flag block of rank r (its mailbox lies behind it)
  int bcast_ni[CFD_BCAST_MAX];
  int bepoch;
  std::vector<void*> ipc_opened;    // mappings made by cfd_ipc_open (closed by cfd_destroy)
  // cavity family: value of den / vis if the last upload of the field was uniform over the whole local array
  bool uni_den, uni_vis, allow_uniform;
  double den0, vis0;
  unsigned long long* dbg;          // phase timing of the flag-ordered sweeps (diagnostic, CFD_B200_FLOW_TIMING=1)
  double* post_buf;                 // result vector of the post-processing kernels (post_kernels.cu), sized on first use
  size_t post_n;
  int num_sms;                      // SM count of THIS context's device (persistent grids are sized from it)
  int smem_optin;                   // largest dynamic shared memory a block may opt into on this device
  std::vector<const void*> smem_raised;   // kernels whose dynamic shared-memory limit was raised for this context's device
};

// cudaFuncSetAttribute is per device and costs a driver call: raise a kernel's dynamic shared-memory limit to the
// device's opt-in maximum the first time THIS context launches it (contexts on other devices do their own), never
// on the per-launch path after that.
template <class K>
static inline cudaError_t cfd_raise_smem(cfd_ctx* c, K kernel) {
  const void* f = reinterpret_cast<const void*>(kernel);
  for (const void* g : c->smem_raised)
    if (g == f) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c->smem_optin);
  if (e == cudaSuccess) c->smem_raised.push_back(f);
  return e;
}

// kernels / launchers implemented across the .cu files (all asynchronous on ctx->stream)
int cfd_launch_assemble_u(cfd_ctx* c);
int cfd_launch_assemble_v(cfd_ctx* c);
int cfd_launch_assemble_p(cfd_ctx* c);
int cfd_launch_correct(cfd_ctx* c, const double* ppref_dev);
int cfd_launch_sweep_rb(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first);
int cfd_launch_sweep_rb_part(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first, int part);
int cfd_launch_sweep_seq(cfd_ctx* c, int istart, int jstart, double* phi, int first);
int cfd_launch_spike_basis(cfd_ctx* c, int istart, int jstart);
int cfd_launch_sweep_seq2(cfd_ctx* c, int istart, int jstart, double* phi, int first, int two_div);
void cfd_local_j_range(const cfd_ctx* c, int gj_lo, int gj_hi, int* jlo, int* jhi);
int cfd_launch_to_line_layout(cfd_ctx* c, const double* colmajor, double* dst, int field);
int cfd_launch_from_line_layout(cfd_ctx* c, const double* src, double* colmajor, int field);
static inline bool cfd_is_coef_field(int f) {
  return f == CFD_AE || f == CFD_AW || f == CFD_AN || f == CFD_AS || f == CFD_AP || f == CFD_SU;
}
int cfd_launch_reduce_partials(cfd_ctx* c, int eq, int nblocks);
int cfd_launch_assemble_fast(cfd_ctx* c, int which, cfd::CavParams& P, dim3 grid, dim3 block);
int cfd_update_fast_geometry(cfd_ctx* c);
int cfd_launch_assemble_tma(cfd_ctx* c, int which, cfd::CavParams& P);
int cfd_assemble_tma_split(int T);
int cfd_launch_general_assemble(cfd_ctx* c, int which);
int cfd_launch_general_assemble_fast(cfd_ctx* c, int which, double* partial);
int cfd_launch_teach_assemble(cfd_ctx* c, int which);
int cfd_launch_teach_wall(cfd_ctx* c, int which);
int cfd_launch_teach_scalar(cfd_ctx* c, int eq);
int cfd_launch_teach_props(cfd_ctx* c);
bool cfd_flow_eligible(const cfd_ctx* c);
int cfd_cluster_size(int T);
int cfd_launch_flow_cluster(cfd_ctx* c, cfd::FlowParams& P);
int cfd_launch_flow_from_sweep(cfd_ctx* c, const cfd::SweepParams& Q, int field_unused);
int cfd_launch_sweeps_flow(cfd_ctx* c, int istart, int jstart, int field, int nsweeps, int first);
int cfd_launch_scalar_bcast(cfd_ctx* c, int owner, int field, int i, int j, double** value_dev);
int cfd_launch_halo_push(cfd_ctx* c, int nfields, const int* fields, const int* w, const int* e);
int cfd_launch_teach_assemble_fast(cfd_ctx* c, int which, int ia, int ib, double* partial, const double* uinc);
