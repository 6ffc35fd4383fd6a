// linsys_kernels.cu -- the iterative line-by-line TDMA of the reference's scalar-transport demos on the
// coefficients held by a context:  ap*phi = ae*phiE + aw*phiW + an*phiN + as*phiS + su.
//   tdma_we / tdma_sn  stagnation-point-flow-2d/steady/fvm2d_stag_point.f90:82-162 (loop: sweep, residual, rsm < tol)
//   wesweep            :164-196  -> the line solver of tdma_kernels.cu (lines along j, swept W -> E)
//   snsweep            :198-230  -> k_lin_sn_seq / k_lin_sn_rb below (lines along i, swept S -> N)
//   tdma               :232-260  convention  a x = b x+ + c x- + d,  two divisions per row
//   tdma_solver        diagonal-convection-2d/fvm2d_oblique_conv.f90:284-346: the same loop with all terms on the
//                      left (as*phiS + ap*phi + an*phiN = su - aw*phiW - ae*phiE); uploading -ae, -aw, -an, -as maps it
//                      onto the form above with identical bits (sign flips are exact).
// Sequential-order contexts reproduce the reference's sweep order and arithmetic; red/black contexts colour the
// lines (rows) by parity.  In the line layout a row is one node of every line: a thread owns a row and marches over
// i, its accesses coalescing across the threads of a warp (neighbouring rows of a chunk-interleaved slice).
// Compiled with -fmad=false.
#include "common.cuh"

namespace cfd {

struct LinParams {
  Layout L;
  const double *ae, *aw, *an, *as, *ap, *su;
  double* phi;
  double *p, *q;        // forward-elimination scratch of the row solver (ni*LD each)
  double* partial;
  int colour;           // row colour of this pass (rows j with (j - 2) % 2 == colour), -1: every row in order
};

// |ae*phiE + aw*phiW + an*phiN + as*phiS + su - ap*phi| summed over the interior (fvm2d_stag_point.f90:99-106)
__global__ void __launch_bounds__(256) k_lin_residual(const LinParams P) {
  __shared__ double wsum[32];
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= ni - 2) {
    const double* ph = P.phi + (long long)i0 * LD;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, q = k * T + t;
      if (j0 < 1 || j0 > nj - 2) continue;
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      const int qn = lpos(j0 + 1, M, T), qs = lpos(j0 - 1, M, T);
      const double r = P.ae[cq] * ph[LD + q] + P.aw[cq] * ph[q - LD] + P.an[cq] * ph[qn] + P.as[cq] * ph[qs] + P.su[cq]
                       - P.ap[cq] * ph[q];
      racc += fabs(r);
    }
  }
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int nw = (blockDim.x * blockDim.y + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) racc += __shfl_down_sync(0xffffffffu, racc, o);
  if (lane == 0) wsum[wid] = racc;
  __syncthreads();
  if (wid == 0) {
    racc = (lane < nw) ? wsum[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) racc += __shfl_down_sync(0xffffffffu, racc, o);
    if (lane == 0) P.partial[blockIdx.y * gridDim.x + blockIdx.x] = racc;
  }
}

// One row solve by one thread: c = aw, a = ap, b = ae, d = su + as*phiS + an*phiN; x(1) = phi(1,j), x(ni) = phi(ni,j).
__device__ __forceinline__ void lin_row_solve(const LinParams& P, int j0) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni;
  const long long LD = P.L.LD, cSL = P.L.cSL;
  const int q = lpos(j0, M, T), qn = lpos(j0 + 1, M, T), qs = lpos(j0 - 1, M, T);
  const long long cq0 = (long long)(j0 % M) * P.L.cSK + (j0 / M);
  double pp = 0.0, qq = P.phi[q];                       // p(1) = 0, q(1) = x(1)
  for (int i0 = 1; i0 <= ni - 2; ++i0) {
    const long long o = (long long)i0 * LD, cq = (long long)i0 * cSL + cq0;
    const double c = P.aw[cq], a = P.ap[cq], b = P.ae[cq];
    const double d = P.su[cq] + P.as[cq] * P.phi[o + qs] + P.an[cq] * P.phi[o + qn];
    const double denom = a - c * pp;
    pp = b / denom;
    qq = (d + c * qq) / denom;
    P.p[o + q] = pp;
    P.q[o + q] = qq;
  }
  double x = P.phi[(long long)(ni - 1) * LD + q];
  for (int i0 = ni - 2; i0 >= 1; --i0) {
    const long long o = (long long)i0 * LD;
    x = P.p[o + q] * x + P.q[o + q];
    P.phi[o + q] = x;
  }
}

// sequential order (the reference's): rows one after the other, each using the row below it already updated
__global__ void k_lin_sn_seq(const LinParams P) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int j0 = 1; j0 <= P.L.nj - 2; ++j0) lin_row_solve(P, j0);
}

// red/black rows: all rows of one colour at once, one thread per row
__global__ void __launch_bounds__(128) k_lin_sn_rb(const LinParams P) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;        // n-th row of this colour
  // thread n -> row: walk rows in line-layout order so that a warp touches consecutive addresses
  const int M = P.L.M, T = P.L.T;
  const int t = n % T, k = n / T;
  if (k >= M) return;
  const int j0 = t * M + k;
  if (j0 < 1 || j0 > P.L.nj - 2) return;
  if (((j0 - 1) & 1) != P.colour) return;
  lin_row_solve(P, j0);
}

}  // namespace cfd

using namespace cfd;

static int lin_fill(cfd_ctx* c, int field, LinParams& P) {
  CFD_REQUIRE(c, "null context");
  CFD_REQUIRE(c->L.nc == 6, "cfd_lin_*: the context must keep raw coefficients (sequential order, or raw_coefficients = 1)");
  if (field < 0 || field >= CFD_FIELD_COUNT || !c->field[field] || cfd_is_coef_field(field)) {
    cfd_set_error("field %d is not a plain field of this context", field);
    return CFD_ERR_INVALID;
  }
  CFD_REQUIRE(c->cfg.goff == 0 && c->cfg.halo_w == 0 && c->cfg.halo_e == 0 && c->cfg.halo_s == 0 && c->cfg.halo_n == 0,
              "cfd_lin_*: single-GPU contexts only");
  P.L = c->L;
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP]; P.su = c->field[CFD_SU];
  P.phi = c->field[field];
  return CFD_OK;
}

extern "C" int cfd_lin_wesweep(cfd_ctx* c, int field) {
  LinParams P{};
  int rc = lin_fill(c, field, P);
  if (rc) return rc;
  CFD_CUDA(cudaSetDevice(c->device));
  if (c->cfg.order == CFD_ORDER_SEQUENTIAL) return cfd_launch_sweep_seq2(c, 2, 2, c->field[field], 0, 1);
  if ((rc = cfd_launch_sweep_rb(c, 2, 2, c->field[field], 0, 0))) return rc;
  return cfd_launch_sweep_rb(c, 2, 2, c->field[field], 1, 0);
}

extern "C" int cfd_lin_snsweep(cfd_ctx* c, int field) {
  LinParams P{};
  int rc = lin_fill(c, field, P);
  if (rc) return rc;
  CFD_CUDA(cudaSetDevice(c->device));
  const size_t fbytes = (size_t)c->L.ni * c->L.LD * sizeof(double);
  if (!c->spike_w) {   // the row solver's p, q share the scratch fields of the partition method
    CFD_CUDA(cudaMalloc(&c->spike_w, fbytes));
    CFD_CUDA(cudaMalloc(&c->spike_z, fbytes));
  }
  P.p = c->spike_w;
  P.q = c->spike_z;
  if (c->cfg.order == CFD_ORDER_SEQUENTIAL) {
    P.colour = -1;
    k_lin_sn_seq<<<1, 32, 0, c->stream>>>(P);
    ++c->launches;
  } else {
    const int n = c->L.M * c->L.T;
    for (int colour = 0; colour < 2; ++colour) {
      P.colour = colour;
      k_lin_sn_rb<<<(n + 127) / 128, 128, 0, c->stream>>>(P);
      ++c->launches;
    }
  }
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

static int lin_residual_async(cfd_ctx* c, LinParams& P) {
  dim3 block(32, 8), grid((c->L.T + 31) / 32, (c->L.ni - 2 + 7) / 8);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  P.partial = c->partial + (size_t)6 * c->max_partials;
  k_lin_residual<<<grid, block, 0, c->stream>>>(P);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return cfd_launch_reduce_partials(c, 6, nblocks);
}

extern "C" int cfd_lin_residual(cfd_ctx* c, int field, double* res) {
  LinParams P{};
  int rc = lin_fill(c, field, P);
  if (rc) return rc;
  CFD_REQUIRE(res, "null argument");
  CFD_CUDA(cudaSetDevice(c->device));
  if ((rc = lin_residual_async(c, P))) return rc;
  CFD_CUDA(cudaMemcpyAsync(res, c->res_dev + 6, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

// dir: 0 = tdma_we, 1 = tdma_sn.  variant: 0 = fvm2d_stag_point.f90:82-162 (stop on rsm < tol, or res1 < 1e-10 at the
// first pass), 1 = tdma_solver of fvm2d_oblique_conv.f90:284-318 (stop on res < 1e-10 or rsm < tol).
extern "C" int cfd_lin_solve(cfd_ctx* c, int field, int dir, int variant, int maxit, double tol, int* iters, double* rsm_out) {
  LinParams P{};
  int rc = lin_fill(c, field, P);
  if (rc) return rc;
  CFD_REQUIRE(dir == 0 || dir == 1, "dir must be 0 (west-east sweeps) or 1 (south-north sweeps)");
  CFD_REQUIRE(maxit >= 1, "maxit must be positive");
  double res = 0.0, res1 = 0.0, rsm = 0.0;
  int it = 0;
  for (it = 1; it <= maxit; ++it) {
    rc = dir == 0 ? cfd_lin_wesweep(c, field) : cfd_lin_snsweep(c, field);
    if (rc) return rc;
    if ((rc = cfd_lin_residual(c, field, &res))) return rc;
    if (it == 1) res1 = res;
    rsm = res / res1;
    if (variant == 1 && res < 1.0e-10) break;
    if (rsm < tol) break;
    if (variant == 0 && it == 1 && res1 < 1.0e-10) break;
  }
  if (it > maxit) it = maxit;
  if (iters) *iters = it;
  if (rsm_out) *rsm_out = rsm;
  return CFD_OK;
}

// ---- time discretisation of the unsteady stagnation-point program (stag_point_unsteady.f90:634-688) ----------------
namespace cfd {
struct LinTimeParams {
  Layout L;
  int itsch;                 // 1 explicit Euler, 2 implicit Euler, 3 Crank-Nicolson
  const double *ae0, *aw0, *an0, *as0, *ct, *phio;      // plain fields: steady coefficients, den/dt*dx*dy, old solution
  double *ae, *aw, *an, *as, *ap, *su;
};
__global__ void __launch_bounds__(256) k_lin_time_terms(const LinTimeParams P) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (t >= T || i0 > ni - 2) return;
  const long long o = (long long)i0 * LD;
  for (int k = 0; k < M; ++k) {
    const int j0 = t * M + k, q = k * T + t;
    if (j0 < 1 || j0 > nj - 2) continue;
    const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
    const int qn = lpos(j0 + 1, M, T), qs = lpos(j0 - 1, M, T);
    double ae = P.ae0[o + q], aw = P.aw0[o + q], an = P.an0[o + q], as = P.as0[o + q];
    const double ct = P.ct[o + q], po = P.phio[o + q];
    double su = 0.0, ap;
    if (P.itsch == 2) {
      su = su + ct * po;
      ap = ct - (ae + aw + an + as);
    } else {
      if (P.itsch == 3) {
        ae = 0.5 * ae; aw = 0.5 * aw; an = 0.5 * an; as = 0.5 * as;
      }
      su = su + ct * po - (ae * P.phio[o + LD + q] + aw * P.phio[o - LD + q] + an * P.phio[o + qn] + as * P.phio[o + qs])
           + (ae + aw + an + as) * po;
      if (P.itsch == 1) {
        ap = ct;
        ae = 0.0; aw = 0.0; an = 0.0; as = 0.0;
      } else {
        ap = ct - (ae + aw + an + as);
      }
    }
    // the unsteady program's solver copies keep every term on the left (stag_point_unsteady.f90:159-261:
    // as*phiS + ap*phi + an*phiN = su - aw*phiW - ae*phiE): stored negated, the system has the form cfd_lin_* solve
    P.ae[cq] = -ae; P.aw[cq] = -aw; P.an[cq] = -an; P.as[cq] = -as; P.ap[cq] = ap; P.su[cq] = su;
  }
}
}  // namespace cfd

extern "C" int cfd_lin_time_terms(cfd_ctx* c, int itsch, int phio_field, const int base_fields[4], int ct_field) {
  CFD_REQUIRE(c && base_fields, "null argument");
  CFD_REQUIRE(c->L.nc == 6, "cfd_lin_*: the context must keep raw coefficients");
  CFD_REQUIRE(itsch >= 1 && itsch <= 3, "itsch must be 1 (explicit Euler), 2 (implicit Euler) or 3 (Crank-Nicolson)");
  const int all[6] = {phio_field, base_fields[0], base_fields[1], base_fields[2], base_fields[3], ct_field};
  for (int f : all) {
    if (f < 0 || f >= CFD_FIELD_COUNT || !c->field[f] || cfd_is_coef_field(f)) {
      cfd_set_error("field %d is not a plain field of this context", f);
      return CFD_ERR_INVALID;
    }
  }
  CFD_CUDA(cudaSetDevice(c->device));
  cfd::LinTimeParams P{};
  P.L = c->L;
  P.itsch = itsch;
  P.phio = c->field[phio_field];
  P.ae0 = c->field[base_fields[0]]; P.aw0 = c->field[base_fields[1]]; P.an0 = c->field[base_fields[2]]; P.as0 = c->field[base_fields[3]];
  P.ct = c->field[ct_field];
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP]; P.su = c->field[CFD_SU];
  dim3 block(32, 8), grid((c->L.T + 31) / 32, (c->L.ni - 2 + 7) / 8);
  cfd::k_lin_time_terms<<<grid, block, 0, c->stream>>>(P);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

// field-to-field copy on the device (phio = phi of the time loop, stag_point_unsteady.f90:586)
extern "C" int cfd_field_copy(cfd_ctx* c, int dst, int src) {
  CFD_REQUIRE(c, "null context");
  for (int f : {dst, src}) {
    if (f < 0 || f >= CFD_FIELD_COUNT || !c->field[f] || cfd_is_coef_field(f)) {
      cfd_set_error("field %d is not a plain field of this context", f);
      return CFD_ERR_INVALID;
    }
  }
  CFD_CUDA(cudaSetDevice(c->device));
  CFD_CUDA(cudaMemcpyAsync(c->field[dst], c->field[src], (size_t)c->L.ni * c->L.LD * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  return CFD_OK;
}
