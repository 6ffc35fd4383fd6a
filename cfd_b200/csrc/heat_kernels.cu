// heat_kernels.cu -- explicit (FTCS) and implicit 5-point heat conduction on a uniform 2-D grid.
//
// Reference: finite-difference-method/heat-conduction/transient-conduction-1d/explicit-time/
// fdm1d_explicit_heat_cond.f90:130-139 (update formula, Dirichlet ends rewritten every step, QDOT :76-88),
// with the 5-point operator of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:171-188 (constant
// conductivity).  The reference has no 2-D transient program; the restatement this must match is
// oracle/heat_oracle.c:orc_heat2d_ftcs.
//
// Storage is the host's own: T(nx,ny) column-major, x fastest, two ping-pong buffers.  One thread owns one
// x-column of a strip of rows and marches in y with a three-row register window, so each T value is read
// from HBM once per step (x-neighbours come from L1, strip halo rows from L2) and written once: 16 B/cell.
#include "common.cuh"

struct cfd_heat {
  cfd_heat_config cfg;
  int device;
  cudaStream_t stream;
  double* buf[2];
  int cur;
  long long steps_done;
};

namespace cfd {

struct HeatParams {
  int nx, ny;            // local extents
  int j_lo, j_hi;        // local rows to update (inclusive)
  int gj_off, NY;        // global row of local row 0; global row count
  double alpha, dx2, dy2, dt, rcp_rho_cp_dummy;
  double cx, cy, qdt;    // fast path: alpha*dt/dx^2, alpha*dt/dy^2, (qval/rho/cp)*dt
  double rho, cp, xa, dx, qthr, qval;
  double tleft, tright, tbottom, ttop;
  const double* src;
  double* dst;
};

#define HEAT_RY 64

// EXACT: the restatement's operation order with IEEE divides (bit-comparable);
// otherwise pre-multiplied coefficients and FMAs (agrees to a few ulp; north_star's 1e-12 budget).
template <bool EXACT>
__global__ void __launch_bounds__(128) k_ftcs(const HeatParams P) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.nx) return;
  int j = P.j_lo + blockIdx.y * HEAT_RY;
  const int jend = min(j + HEAT_RY - 1, P.j_hi);
  const size_t nx = (size_t)P.nx;
  const bool xedge = (i == 0) || (i == P.nx - 1);
  const double xbc = (i == 0) ? P.tleft : P.tright;
  // source term of this column: QDOT(x(i), time), x(i) = x(1) + (i-1)*dx
  const double xi = __dadd_rn(P.xa, __dmul_rn((double)i, P.dx));
  const double q = (xi < P.qthr) ? P.qval : 0.0;
  const double qterm = EXACT ? __ddiv_rn(__ddiv_rn(q, P.rho), P.cp) : ((xi < P.qthr) ? P.qdt : 0.0);
  const double* s = P.src + (size_t)j * nx + i;
  double* d = P.dst + (size_t)j * nx + i;
  double S = (j > 0) ? __ldg(s - nx) : 0.0;
  double C = __ldg(s);
  for (; j <= jend; ++j, s += nx, d += nx) {
    const int gj = P.gj_off + j;
    const double N = (j + 1 < P.ny) ? __ldg(s + nx) : 0.0;
    double out;
    if (gj == 0) {
      out = P.tbottom;
    } else if (gj == P.NY - 1) {
      out = P.ttop;
    } else if (xedge) {
      out = xbc;
    } else {
      const double W = __ldg(s - 1), E = __ldg(s + 1);
      if (EXACT) {
        const double lx = __dadd_rn(__dsub_rn(E, __dmul_rn(2.0, C)), W);
        const double ly = __dadd_rn(__dsub_rn(N, __dmul_rn(2.0, C)), S);
        const double tx = __ddiv_rn(__dmul_rn(P.alpha, lx), P.dx2);
        const double ty = __ddiv_rn(__dmul_rn(P.alpha, ly), P.dy2);
        out = __dadd_rn(C, __dmul_rn(__dadd_rn(__dadd_rn(tx, ty), qterm), P.dt));
      } else {
        const double lx = (E - 2.0 * C) + W;
        const double ly = (N - 2.0 * C) + S;
        out = fma(P.cx, lx, fma(P.cy, ly, C + qterm));
      }
    }
    *d = out;
    S = C;
    C = N;
  }
}

}  // namespace cfd

using namespace cfd;

extern "C" int cfd_heat_create(cfd_heat** out, const cfd_heat_config* cfg) {
  CFD_REQUIRE(out && cfg, "null argument");
  *out = nullptr;
  CFD_REQUIRE(cfg->nx >= 3 && cfg->ny >= 3, "nx, ny must be >= 3");
  CFD_REQUIRE(cfg->halo_s >= 0 && cfg->halo_n >= 0 && cfg->halo_s + cfg->halo_n < cfg->ny, "bad halo rows");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cfd_set_error("no CUDA device visible; libcfd_b200 has no CPU fallback");
    return CFD_ERR_NO_DEVICE;
  }
  int dev = cfg->device;
  if (dev < 0) CFD_CUDA(cudaGetDevice(&dev));
  CFD_REQUIRE(dev < ndev, "device ordinal out of range");
  CFD_CUDA(cudaSetDevice(dev));
  cfd_heat* h = new cfd_heat();
  h->cfg = *cfg;
  h->device = dev;
  h->stream = 0;
  h->cur = 0;
  h->steps_done = 0;
  h->buf[0] = h->buf[1] = nullptr;
  const size_t bytes = (size_t)cfg->nx * cfg->ny * sizeof(double);
  for (int b = 0; b < 2; ++b) {
    if (cudaMalloc(&h->buf[b], bytes) != cudaSuccess) {
      cfd_set_error("cudaMalloc of %zu bytes failed", bytes);
      cfd_heat_destroy(h);
      return CFD_ERR_NOMEM;
    }
    CFD_CUDA(cudaMemset(h->buf[b], 0, bytes));
  }
  *out = h;
  return CFD_OK;
}

extern "C" int cfd_heat_destroy(cfd_heat* h) {
  if (!h) return CFD_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (int b = 0; b < 2; ++b) if (h->buf[b]) cudaFree(h->buf[b]);
  delete h;
  return CFD_OK;
}

extern "C" int cfd_heat_set_stream(cfd_heat* h, void* s) {
  CFD_REQUIRE(h, "null context");
  h->stream = (cudaStream_t)s;
  return CFD_OK;
}

extern "C" int cfd_heat_synchronize(cfd_heat* h) {
  CFD_REQUIRE(h, "null context");
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  return CFD_OK;
}

extern "C" int cfd_heat_upload(cfd_heat* h, const double* host, int ld) {
  CFD_REQUIRE(h && host && ld >= h->cfg.nx, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  const size_t w = (size_t)h->cfg.nx * sizeof(double);
  CFD_CUDA(cudaMemcpy2DAsync(h->buf[h->cur], w, host, (size_t)ld * sizeof(double), w, h->cfg.ny,
                             cudaMemcpyHostToDevice, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  return CFD_OK;
}

extern "C" int cfd_heat_download(cfd_heat* h, double* host, int ld) {
  CFD_REQUIRE(h && host && ld >= h->cfg.nx, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  const size_t w = (size_t)h->cfg.nx * sizeof(double);
  CFD_CUDA(cudaMemcpy2DAsync(host, (size_t)ld * sizeof(double), h->buf[h->cur], w, w, h->cfg.ny,
                             cudaMemcpyDeviceToHost, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  return CFD_OK;
}

extern "C" int cfd_heat_device_view(cfd_heat* h, void** cur, long long* row_stride, int* nrows) {
  CFD_REQUIRE(h, "null context");
  if (cur) *cur = h->buf[h->cur];
  if (row_stride) *row_stride = h->cfg.nx;
  if (nrows) *nrows = h->cfg.ny;
  return CFD_OK;
}

static void fill_heat(const cfd_heat* h, HeatParams& P) {
  const cfd_heat_config& g = h->cfg;
  const int NX = g.nx, NY = g.ny_global ? g.ny_global : g.ny;
  P.nx = g.nx;
  P.ny = g.ny;
  P.j_lo = g.halo_s;
  P.j_hi = g.ny - 1 - g.halo_n;
  P.gj_off = g.j_goff;
  P.NY = NY;
  const double dx = (g.xb - g.xa) / (double)(NX - 1), dy = (g.yb - g.ya) / (double)(NY - 1);
  P.alpha = g.k / g.rho / g.cp;
  P.dx2 = dx * dx;
  P.dy2 = dy * dy;
  P.dt = g.dt;
  P.cx = P.alpha * g.dt / P.dx2;
  P.cy = P.alpha * g.dt / P.dy2;
  P.qdt = g.qval / g.rho / g.cp * g.dt;
  P.rho = g.rho;
  P.cp = g.cp;
  P.xa = g.xa;
  P.dx = dx;
  P.qthr = g.qthr;
  P.qval = g.qval;
  P.tleft = g.tleft;
  P.tright = g.tright;
  P.tbottom = g.tbottom;
  P.ttop = g.ttop;
}

static int heat_steps(cfd_heat* h, int nsteps, bool exact) {
  CFD_REQUIRE(h && nsteps >= 0, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  HeatParams P{};
  fill_heat(h, P);
  const int rows = P.j_hi - P.j_lo + 1;
  dim3 block(128), grid((P.nx + 127) / 128, (rows + HEAT_RY - 1) / HEAT_RY);
  for (int n = 0; n < nsteps; ++n) {
    P.src = h->buf[h->cur];
    P.dst = h->buf[h->cur ^ 1];
    if (exact) {
      k_ftcs<true><<<grid, block, 0, h->stream>>>(P);
    } else {
      k_ftcs<false><<<grid, block, 0, h->stream>>>(P);
    }
    CFD_CUDA(cudaGetLastError());
    h->cur ^= 1;
    ++h->steps_done;
  }
  return CFD_OK;
}

extern "C" int cfd_heat2d_step(cfd_heat* h, int nsteps) { return heat_steps(h, nsteps, false); }
extern "C" int cfd_heat2d_step_exact(cfd_heat* h, int nsteps) { return heat_steps(h, nsteps, true); }

extern "C" int cfd_heat2d_implicit_step(cfd_heat* h, int nsteps) {
  (void)h;
  (void)nsteps;
  cfd_set_error("cfd_heat2d_implicit_step: not built yet");
  return CFD_ERR_UNSUPPORTED;
}
