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
};

__host__ __device__ __forceinline__ int lpos(int j0, int M, int T) { return (j0 % M) * T + (j0 / M); }

#define CFD_MAX_M 16

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
};

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
  double* field[CFD_FIELD_COUNT];   // line layout, ni*LD doubles each (allocated on demand by family)
  double* geom[CFD_GEOM_COUNT];     // by-i: natural (ni doubles); by-j: line layout (LD doubles)
  std::vector<double> geom_host[CFD_GEOM_COUNT];
  double* stage;                    // ni*nj staging buffer for upload/download (column-major image)
  double* partial;                  // residual partial sums, CFD_EQ_COUNT * max_blocks
  int max_partials;
  double* res_dev;                  // 8 packed residual sums
  double* seq_scratch;              // sequential-order debug sweep: forward-elimination arrays
  int resid_pending[CFD_EQ_COUNT];  // number of partials waiting to be folded per equation
};

// kernels / launchers implemented across the .cu files (all asynchronous on ctx->stream)
int cfd_launch_assemble_u(cfd_ctx* c);
int cfd_launch_assemble_v(cfd_ctx* c);
int cfd_launch_assemble_p(cfd_ctx* c);
int cfd_launch_correct(cfd_ctx* c, const double* ppref_dev);
int cfd_launch_sweep_rb(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first);
int cfd_launch_sweep_seq(cfd_ctx* c, int istart, int jstart, double* phi, int first);
int cfd_launch_to_line_layout(cfd_ctx* c, const double* colmajor, double* dst);
int cfd_launch_from_line_layout(cfd_ctx* c, const double* src, double* colmajor);
int cfd_launch_reduce_partials(cfd_ctx* c, int eq, int nblocks);
