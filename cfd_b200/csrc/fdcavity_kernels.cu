// fdcavity_kernels.cu -- stencil passes of the explicit finite-difference cavity solver
// (lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90): momentum predictors :170-192, p' source :194-205,
// correction :208-222, mirror-ghost boundary conditions :224-237, continuity norm :239-247, uo=u copies :252-254.
// The p' Poisson solve between predictor and correction (Stone's SIP, :24-87) is NOT here: it is SURVEY 8f-1
// "next"; the caller solves pp (reference sipsol on the host today) between cfd_fdcav_predict and
// cfd_fdcav_correct.  Arrays are the host's own column-major layout with one common leading dimension ni+1.
// Reference operation order, non-contracted IEEE operations: bit-comparable with oracle/fdcavity_oracle.c.
#include "common.cuh"

struct cfd_fdcav {
  cfd_fdcav_config cfg;
  int device;
  cudaStream_t stream;
  double* f[CFD_FDC_FIELD_COUNT];
  double* partial;
  int npartial;
  // SIP solver work space (cfd_fdcav_solve_pp): 6 arrays, band progress counters, block partials, info, rsm
  double* sip_work;
  unsigned long long* sip_prog;
  double* sip_partial;
  int* sip_info;
  double* sip_rsm;
};
size_t cfd_sip_sync_bytes(int ld, int nrows);
int cfd_launch_sip(int device, cudaStream_t stream, int ld, int nrows, int imax, int jmax, const double* c,
                   const double* const* coef, double alpha, double tol, int maxit, const double* su, double* phi, double* work,
                   unsigned long long* prog, double* partial, int* info, double* rsm_out, int max_partials);

namespace cfd {

struct FdcParams {
  int ni, nj, ld;
  double dx, dy, dt, re, utop, urfp;
  double *u, *v, *p, *uo, *vo, *po, *pp, *su;
  double *ae, *aw, *an, *as, *ap;
  double urfu, urfv;
  int steady;
  double* partial;
};
#define B2(a, i, j) ((a)[(size_t)((i) - 1) + (size_t)((j) - 1) * (size_t)P.ld])

__global__ void __launch_bounds__(256) k_fdc_predict(const FdcParams P) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x, j = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int ni = P.ni, nj = P.nj;
  const double dx = P.dx, dy = P.dy;
  const double *uo = P.uo, *vo = P.vo, *po = P.po;
  if (i >= 2 && i <= ni - 1 && j >= 2 && j <= nj) {
    const double a = __ddiv_rn(__dmul_rn(-0.5, __dsub_rn(__dmul_rn(B2(uo, i + 1, j), B2(uo, i + 1, j)), __dmul_rn(B2(uo, i - 1, j), B2(uo, i - 1, j)))), dx);
    const double b = __ddiv_rn(__dmul_rn(0.25, __dsub_rn(__dmul_rn(__dadd_rn(B2(uo, i, j), B2(uo, i, j + 1)), __dadd_rn(B2(vo, i, j), B2(vo, i + 1, j))),
                                                      __dmul_rn(__dadd_rn(B2(uo, i, j), B2(uo, i, j - 1)), __dadd_rn(B2(vo, i, j - 1), B2(vo, i + 1, j - 1))))), dy);
    const double c = __ddiv_rn(__dsub_rn(B2(po, i + 1, j), B2(po, i, j)), dx);
    const double lx = __ddiv_rn(__dadd_rn(__dsub_rn(B2(uo, i - 1, j), __dmul_rn(2.0, B2(uo, i, j))), B2(uo, i + 1, j)), __dmul_rn(dx, dx));
    const double ly = __ddiv_rn(__dadd_rn(__dsub_rn(B2(uo, i, j - 1), __dmul_rn(2.0, B2(uo, i, j))), B2(uo, i, j + 1)), __dmul_rn(dy, dy));
    const double r = __dadd_rn(__dsub_rn(__dsub_rn(a, b), c), __ddiv_rn(__dadd_rn(lx, ly), P.re));
    B2(P.u, i, j) = __dadd_rn(B2(uo, i, j), __dmul_rn(P.dt, r));
  }
  if (i >= 2 && i <= ni && j >= 2 && j <= nj - 1) {
    const double a = __ddiv_rn(__dmul_rn(-0.25, __dsub_rn(__dmul_rn(__dadd_rn(B2(uo, i, j), B2(uo, i, j + 1)), __dadd_rn(B2(vo, i, j), B2(vo, i + 1, j))),
                                                       __dmul_rn(__dadd_rn(B2(uo, i - 1, j), B2(uo, i - 1, j + 1)), __dadd_rn(B2(vo, i, j), B2(vo, i - 1, j))))), dx);
    const double b = __ddiv_rn(__dmul_rn(0.5, __dsub_rn(__dmul_rn(B2(vo, i, j + 1), B2(vo, i, j + 1)), __dmul_rn(B2(vo, i, j - 1), B2(vo, i, j - 1)))), dy);
    const double c = __ddiv_rn(__dsub_rn(B2(po, i, j + 1), B2(po, i, j)), dy);
    const double lx = __ddiv_rn(__dadd_rn(__dsub_rn(B2(vo, i - 1, j), __dmul_rn(2.0, B2(vo, i, j))), B2(vo, i + 1, j)), __dmul_rn(dx, dx));
    const double ly = __ddiv_rn(__dadd_rn(__dsub_rn(B2(vo, i, j - 1), __dmul_rn(2.0, B2(vo, i, j))), B2(vo, i, j + 1)), __dmul_rn(dy, dy));
    const double r = __dadd_rn(__dsub_rn(__dsub_rn(a, b), c), __ddiv_rn(__dadd_rn(lx, ly), P.re));
    B2(P.v, i, j) = __dadd_rn(B2(vo, i, j), __dmul_rn(P.dt, r));
  }
}

// p' source (:203) and pp(:,:) = 0 (:194)
__global__ void __launch_bounds__(256) k_fdc_source(const FdcParams P) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x, j = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  if (i > P.ni + 1 || j > P.nj + 1) return;
  B2(P.pp, i, j) = 0.0;
  if (i >= 2 && i <= P.ni && j >= 2 && j <= P.nj)
    B2(P.su, i, j) = __dsub_rn(__ddiv_rn(-__dsub_rn(B2(P.u, i, j), B2(P.u, i - 1, j)), P.dx),
                               __ddiv_rn(__dsub_rn(B2(P.v, i, j), B2(P.v, i, j - 1)), P.dy));
}

// :208-222
__global__ void __launch_bounds__(256) k_fdc_correct(const FdcParams P) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x, j = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int ni = P.ni, nj = P.nj;
  if (i < 2 || i > ni || j < 2 || j > nj) return;
  const double ppc = B2(P.pp, i, j);
  B2(P.p, i, j) = __dadd_rn(B2(P.p, i, j), __dmul_rn(P.urfp, ppc));
  if (i <= ni - 1) B2(P.u, i, j) = __dsub_rn(B2(P.u, i, j), __dmul_rn(__ddiv_rn(P.dt, P.dx), __dsub_rn(B2(P.pp, i + 1, j), ppc)));
  if (j <= nj - 1) B2(P.v, i, j) = __dsub_rn(B2(P.v, i, j), __dmul_rn(__ddiv_rn(P.dt, P.dy), __dsub_rn(B2(P.pp, i, j + 1), ppc)));
}

// :224-237, one block, the reference's statement order (later statements see the earlier ones' results)
__global__ void __launch_bounds__(1024) k_fdc_bc(const FdcParams P) {
  const int ni = P.ni, nj = P.nj, tid = threadIdx.x, nt = blockDim.x;
  for (int i = 1 + tid; i <= ni; i += nt) B2(P.u, i, 1) = -B2(P.u, i, 2);
  __syncthreads();
  for (int i = 1 + tid; i <= ni; i += nt) B2(P.u, i, nj + 1) = __dsub_rn(__dmul_rn(2.0, P.utop), B2(P.u, i, nj));
  __syncthreads();
  for (int j = 2 + tid; j <= nj; j += nt) { B2(P.u, 1, j) = 0.0; B2(P.u, ni, j) = 0.0; }
  __syncthreads();
  for (int j = 1 + tid; j <= nj; j += nt) B2(P.v, 1, j) = -B2(P.v, 2, j);
  __syncthreads();
  for (int j = 1 + tid; j <= nj; j += nt) B2(P.v, ni + 1, j) = -B2(P.v, ni, j);
  __syncthreads();
  for (int i = 2 + tid; i <= ni; i += nt) { B2(P.v, i, 1) = 0.0; B2(P.v, i, nj) = 0.0; }
  __syncthreads();
  const int m = (ni + 1 < nj + 1) ? ni + 1 : nj + 1;      // `p(1,1:ni+1)`: ni where nj is meant (:234), clamped
  for (int j = 1 + tid; j <= m; j += nt) B2(P.p, 1, j) = B2(P.p, 2, j);
  __syncthreads();
  for (int j = 1 + tid; j <= nj + 1; j += nt) B2(P.p, ni + 1, j) = B2(P.p, ni, j);
  __syncthreads();
  for (int i = 2 + tid; i <= ni; i += nt) B2(P.p, i, nj + 1) = B2(P.p, i, nj);
  __syncthreads();
  for (int i = 2 + tid; i <= ni; i += nt) B2(P.p, i, 1) = B2(P.p, i, 2);
}

// :239-247 squared continuity residual, block partials (fixed-order final sum on one block)
__global__ void __launch_bounds__(256) k_fdc_norm(const FdcParams P) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x, j = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  double v = 0.0;
  if (i >= 2 && i <= P.ni && j >= 2 && j <= P.nj) {
    const double d = __dadd_rn(__ddiv_rn(__dsub_rn(B2(P.u, i, j), B2(P.u, i - 1, j)), P.dx),
                               __ddiv_rn(__dsub_rn(B2(P.v, i, j), B2(P.v, i, j - 1)), P.dy));
    v = __dmul_rn(d, d);
  }
  __shared__ double wsum[32];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if (lane == 0) wsum[wid] = v;
  __syncthreads();
  if (wid == 0) {
    v = (lane < 8) ? wsum[lane] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) P.partial[1 + blockIdx.y * gridDim.x + blockIdx.x] = v;
  }
}
__global__ void k_fdc_norm_final(double* partial, int n) {
  __shared__ double sm[1024];
  double v = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) v += partial[1 + k];
  sm[threadIdx.x] = v;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[0] = sqrt(sm[0]);
}

// ---- implicit-Euler variant: lid-driven-cavity-2d/fdm-implicit/fdm_implicit_cavity.f90 ----------------------------
// which: 0 u-momentum coefficients :178-198, 1 v-momentum :201-220, 2 p' equation :222-240 (pp = 0 included)
__global__ void __launch_bounds__(256) k_fdci_assemble(const FdcParams P, const int which) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x, j = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int ni = P.ni, nj = P.nj;
  if (i > ni + 1 || j > nj + 1) return;
  const double dx = P.dx, dy = P.dy;
  const double rex = __ddiv_rn(__ddiv_rn(1.0, P.re), __dmul_rn(dx, dx)), rey = __ddiv_rn(__ddiv_rn(1.0, P.re), __dmul_rn(dy, dy));
  const double pterm = __dadd_rn(__ddiv_rn(1.0, P.dt), __dmul_rn(2.0, __dadd_rn(rex, rey)));
  const double *u = P.u, *v = P.v, *p = P.p;
  if (which == 0) {
    if (i < 2 || i > ni - 1 || j < 2 || j > nj) return;
    const double vel1 = __dmul_rn(0.5, __dadd_rn(B2(v, i, j), B2(v, i + 1, j)));
    const double vel2 = __dmul_rn(0.5, __dadd_rn(B2(v, i, j - 1), B2(v, i + 1, j - 1)));
    B2(P.ae, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(-0.5, B2(u, i + 1, j)), dx), rex);
    B2(P.aw, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(0.5, B2(u, i - 1, j)), dx), rex);
    B2(P.an, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(-0.5, vel1), dy), rey);
    B2(P.as, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(0.5, vel2), dy), rey);
    double su, ap;
    if (P.steady) {
      su = __ddiv_rn(__dsub_rn(B2(p, i, j), B2(p, i + 1, j)), dx);
      ap = __dmul_rn(2.0, __dadd_rn(rex, rey));
    } else {
      su = __dadd_rn(__ddiv_rn(B2(u, i, j), P.dt), __ddiv_rn(__dsub_rn(B2(p, i, j), B2(p, i + 1, j)), dx));
      ap = pterm;
    }
    su = __dadd_rn(su, __dmul_rn(__ddiv_rn(__dmul_rn(__dsub_rn(1.0, P.urfu), ap), P.urfu), B2(u, i, j)));
    B2(P.su, i, j) = su;
    B2(P.ap, i, j) = __ddiv_rn(ap, P.urfu);
  } else if (which == 1) {
    if (i < 2 || i > ni || j < 2 || j > nj - 1) return;
    const double vel1 = __dmul_rn(0.5, __dadd_rn(B2(u, i, j), B2(u, i, j + 1)));
    const double vel2 = __dmul_rn(0.5, __dadd_rn(B2(u, i - 1, j), B2(u, i - 1, j + 1)));
    B2(P.ae, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(-0.5, vel1), dx), rex);
    B2(P.aw, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(0.5, vel2), dx), rex);
    B2(P.an, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(-0.5, B2(v, i, j + 1)), dy), rey);
    B2(P.as, i, j) = __dadd_rn(__ddiv_rn(__dmul_rn(-0.5, B2(v, i, j - 1)), dy), rey);   // sign as in the source (:208)
    double su, ap;
    if (P.steady) {
      su = __ddiv_rn(__dsub_rn(B2(p, i, j), B2(p, i, j + 1)), dy);
      ap = __dmul_rn(2.0, __dadd_rn(rex, rey));
    } else {
      su = __dadd_rn(__ddiv_rn(B2(v, i, j), P.dt), __ddiv_rn(__dsub_rn(B2(p, i, j), B2(p, i, j + 1)), dy));
      ap = pterm;
    }
    su = __dadd_rn(su, __dmul_rn(__ddiv_rn(__dmul_rn(__dsub_rn(1.0, P.urfv), ap), P.urfv), B2(v, i, j)));
    B2(P.su, i, j) = su;
    B2(P.ap, i, j) = __ddiv_rn(ap, P.urfv);
  } else {
    B2(P.pp, i, j) = 0.0;
    if (i < 2 || i > ni || j < 2 || j > nj) return;
    const double base = P.steady ? __ddiv_rn(1.0, __dmul_rn(2.0, __dadd_rn(rex, rey))) : __ddiv_rn(1.0, pterm);
    const double ax = __ddiv_rn(base, __dmul_rn(dx, dx)), ay = __ddiv_rn(base, __dmul_rn(dy, dy));
    B2(P.ae, i, j) = ax;
    B2(P.aw, i, j) = ax;
    B2(P.an, i, j) = ay;
    B2(P.as, i, j) = ay;
    B2(P.ap, i, j) = __dadd_rn(__dadd_rn(__dadd_rn(ax, ax), ay), ay);
    B2(P.su, i, j) = __dadd_rn(__ddiv_rn(__dsub_rn(B2(u, i - 1, j), B2(u, i, j)), dx), __ddiv_rn(__dsub_rn(B2(v, i, j - 1), B2(v, i, j)), dy));
  }
}

// :243-258
__global__ void __launch_bounds__(256) k_fdci_correct(const FdcParams P) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x, j = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int ni = P.ni, nj = P.nj;
  if (i < 2 || i > ni || j < 2 || j > nj) return;
  const double rex = __ddiv_rn(__ddiv_rn(1.0, P.re), __dmul_rn(P.dx, P.dx)), rey = __ddiv_rn(__ddiv_rn(1.0, P.re), __dmul_rn(P.dy, P.dy));
  const double pterm = __dadd_rn(__ddiv_rn(1.0, P.dt), __dmul_rn(2.0, __dadd_rn(rex, rey)));
  const double ppc = B2(P.pp, i, j);
  B2(P.p, i, j) = __dadd_rn(B2(P.p, i, j), __dmul_rn(P.urfp, ppc));
  if (i <= ni - 1) B2(P.u, i, j) = __dsub_rn(B2(P.u, i, j), __ddiv_rn(__ddiv_rn(__dsub_rn(B2(P.pp, i + 1, j), ppc), pterm), P.dx));
  if (j <= nj - 1) B2(P.v, i, j) = __dsub_rn(B2(P.v, i, j), __ddiv_rn(__ddiv_rn(__dsub_rn(B2(P.pp, i, j + 1), ppc), pterm), P.dy));
}

}  // namespace cfd
using namespace cfd;

static void fdc_fill(const cfd_fdcav* h, FdcParams& P) {
  P.ni = h->cfg.ni; P.nj = h->cfg.nj; P.ld = h->cfg.ni + 1;
  P.dx = 1.0 / (h->cfg.ni - 1); P.dy = 1.0 / (h->cfg.nj - 1);
  P.dt = h->cfg.dt; P.re = h->cfg.re; P.utop = h->cfg.utop; P.urfp = h->cfg.urfp;
  P.u = h->f[CFD_FDC_U]; P.v = h->f[CFD_FDC_V]; P.p = h->f[CFD_FDC_P]; P.uo = h->f[CFD_FDC_UO]; P.vo = h->f[CFD_FDC_VO];
  P.po = h->f[CFD_FDC_PO]; P.pp = h->f[CFD_FDC_PP]; P.su = h->f[CFD_FDC_SU];
  P.ae = h->f[CFD_FDC_AE]; P.aw = h->f[CFD_FDC_AW]; P.an = h->f[CFD_FDC_AN]; P.as = h->f[CFD_FDC_AS]; P.ap = h->f[CFD_FDC_AP];
  P.urfu = h->cfg.urfu; P.urfv = h->cfg.urfv; P.steady = h->cfg.steady;
  P.partial = h->partial;
}
static dim3 fdc_grid(const cfd_fdcav* h) { return dim3((h->cfg.ni + 1 + 31) / 32, (h->cfg.nj + 1 + 7) / 8); }

extern "C" int cfd_fdcav_create(cfd_fdcav** out, const cfd_fdcav_config* cfg) {
  CFD_REQUIRE(out && cfg, "null argument");
  *out = nullptr;
  CFD_REQUIRE(cfg->ni >= 3 && cfg->nj >= 3, "ni, nj must be >= 3");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cfd_set_error("no CUDA device visible; libcfd_b200 has no CPU fallback");
    return CFD_ERR_NO_DEVICE;
  }
  int dev = cfg->device;
  if (dev < 0) CFD_CUDA(cudaGetDevice(&dev));
  CFD_REQUIRE(dev < ndev, "device ordinal out of range");
  CFD_CUDA(cudaSetDevice(dev));
  cfd_fdcav* h = new cfd_fdcav();
  h->cfg = *cfg;
  h->device = dev;
  h->stream = 0;
  h->partial = nullptr;
  h->sip_work = nullptr;
  h->sip_prog = nullptr;
  h->sip_partial = nullptr;
  h->sip_info = nullptr;
  h->sip_rsm = nullptr;
  for (auto& f : h->f) f = nullptr;
  const size_t bytes = (size_t)(cfg->ni + 1) * (cfg->nj + 1) * sizeof(double);
  for (auto& f : h->f) {
    if (cudaMalloc(&f, bytes) != cudaSuccess) {
      cfd_set_error("cudaMalloc of %zu bytes failed", bytes);
      cfd_fdcav_destroy(h);
      return CFD_ERR_NOMEM;
    }
    CFD_CUDA(cudaMemset(f, 0, bytes));
  }
  dim3 g = fdc_grid(h);
  h->npartial = g.x * g.y;
  CFD_CUDA(cudaMalloc(&h->partial, (size_t)(h->npartial + 1) * sizeof(double)));
  *out = h;
  return CFD_OK;
}
extern "C" int cfd_fdcav_destroy(cfd_fdcav* h) {
  if (!h) return CFD_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (auto& f : h->f) if (f) cudaFree(f);
  if (h->partial) cudaFree(h->partial);
  if (h->sip_work) cudaFree(h->sip_work);
  if (h->sip_prog) cudaFree(h->sip_prog);
  if (h->sip_partial) cudaFree(h->sip_partial);
  if (h->sip_info) cudaFree(h->sip_info);
  if (h->sip_rsm) cudaFree(h->sip_rsm);
  delete h;
  return CFD_OK;
}
extern "C" int cfd_fdcav_upload(cfd_fdcav* h, int field, const double* host, int ld) {
  CFD_REQUIRE(h && host && field >= 0 && field < CFD_FDC_FIELD_COUNT && ld >= h->cfg.ni + 1, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  const size_t w = (size_t)(h->cfg.ni + 1) * sizeof(double);
  CFD_CUDA(cudaMemcpy2DAsync(h->f[field], w, host, (size_t)ld * sizeof(double), w, h->cfg.nj + 1, cudaMemcpyHostToDevice, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  return CFD_OK;
}
extern "C" int cfd_fdcav_download(cfd_fdcav* h, int field, double* host, int ld) {
  CFD_REQUIRE(h && host && field >= 0 && field < CFD_FDC_FIELD_COUNT && ld >= h->cfg.ni + 1, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  const size_t w = (size_t)(h->cfg.ni + 1) * sizeof(double);
  CFD_CUDA(cudaMemcpy2DAsync(host, (size_t)ld * sizeof(double), h->f[field], w, w, h->cfg.nj + 1, cudaMemcpyDeviceToHost, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  return CFD_OK;
}
extern "C" int cfd_fdcav_predict(cfd_fdcav* h) {
  CFD_REQUIRE(h, "null context");
  CFD_CUDA(cudaSetDevice(h->device));
  FdcParams P{};
  fdc_fill(h, P);
  dim3 block(32, 8), grid = fdc_grid(h);
  k_fdc_predict<<<grid, block, 0, h->stream>>>(P);
  k_fdc_source<<<grid, block, 0, h->stream>>>(P);
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
extern "C" int cfd_fdcav_correct(cfd_fdcav* h, double* error) {
  CFD_REQUIRE(h, "null context");
  CFD_CUDA(cudaSetDevice(h->device));
  FdcParams P{};
  fdc_fill(h, P);
  dim3 block(32, 8), grid = fdc_grid(h);
  k_fdc_correct<<<grid, block, 0, h->stream>>>(P);
  k_fdc_bc<<<1, 1024, 0, h->stream>>>(P);
  k_fdc_norm<<<grid, block, 0, h->stream>>>(P);
  k_fdc_norm_final<<<1, 1024, 0, h->stream>>>(h->partial, h->npartial);
  CFD_CUDA(cudaGetLastError());
  const size_t bytes = (size_t)(h->cfg.ni + 1) * (h->cfg.nj + 1) * sizeof(double);
  CFD_CUDA(cudaMemcpyAsync(h->f[CFD_FDC_UO], h->f[CFD_FDC_U], bytes, cudaMemcpyDeviceToDevice, h->stream));
  CFD_CUDA(cudaMemcpyAsync(h->f[CFD_FDC_VO], h->f[CFD_FDC_V], bytes, cudaMemcpyDeviceToDevice, h->stream));
  CFD_CUDA(cudaMemcpyAsync(h->f[CFD_FDC_PO], h->f[CFD_FDC_P], bytes, cudaMemcpyDeviceToDevice, h->stream));
  if (error) {
    CFD_CUDA(cudaMemcpyAsync(error, h->partial, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CFD_CUDA(cudaStreamSynchronize(h->stream));
  }
  return CFD_OK;
}

static int sip_alloc(cfd_fdcav* h) {
  const size_t n = (size_t)(h->cfg.ni + 1) * (h->cfg.nj + 1);
  if (!h->sip_work) {
    CFD_CUDA(cudaMalloc(&h->sip_work, 6 * n * sizeof(double)));
    CFD_CUDA(cudaMalloc(&h->sip_prog, cfd_sip_sync_bytes(h->cfg.ni + 1, h->cfg.nj + 1)));
    CFD_CUDA(cudaMalloc(&h->sip_partial, 1024 * sizeof(double)));
    CFD_CUDA(cudaMalloc(&h->sip_info, 2 * sizeof(int)));
    CFD_CUDA(cudaMalloc(&h->sip_rsm, sizeof(double)));
  }
  return CFD_OK;
}
static int sip_finish(cfd_fdcav* h, int* iters) {
  int info[2] = {0, 0};
  CFD_CUDA(cudaMemcpyAsync(info, h->sip_info, sizeof(info), cudaMemcpyDeviceToHost, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  if (info[1]) {
    cfd_set_error("SIP: wavefront wait abandoned (pipeline stalled)");
    return CFD_ERR_CUDA;
  }
  if (iters) *iters = info[0];
  return CFD_OK;
}

// sipsol (fdm_explicit_cavity.f90:24-87) for the p' system of the current step, entirely on the device
extern "C" int cfd_fdcav_solve_pp(cfd_fdcav* h, int* iters) {
  CFD_REQUIRE(h, "null context");
  CFD_CUDA(cudaSetDevice(h->device));
  int rc = sip_alloc(h);
  if (rc) return rc;
  const int ni = h->cfg.ni, nj = h->cfg.nj;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), dt = h->cfg.dt;
  const double ax = dt / (dx * dx), ay = dt / (dy * dy);
  const double c[5] = {-ax, -ax, -ay, -ay, 2 * (dt / (dx * dx) + dt / (dy * dy))};
  rc = cfd_launch_sip(h->device, h->stream, ni + 1, nj + 1, ni - 1, nj - 1, c, nullptr, 0.88, 1.0e-5, 600, h->f[CFD_FDC_SU],
                      h->f[CFD_FDC_PP], h->sip_work, h->sip_prog, h->sip_partial, h->sip_info, h->sip_rsm, 1024);
  return rc ? rc : sip_finish(h, iters);
}

// ---- implicit variant (fdm_implicit_cavity.f90:177-288) -------------------------------------------------------
// which: 0 u, 1 v, 2 p'
extern "C" int cfd_fdcav_implicit_assemble(cfd_fdcav* h, int which) {
  CFD_REQUIRE(h && which >= 0 && which <= 2, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  FdcParams P{};
  fdc_fill(h, P);
  k_fdci_assemble<<<fdc_grid(h), dim3(32, 8), 0, h->stream>>>(P, which);
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
// sipsol (fdm_implicit_cavity.f90:29-92: maxit 3000, alpha 0.80, tol 1e-6) on u (i2 = ni), v (j2 = nj) or pp
extern "C" int cfd_fdcav_implicit_solve(cfd_fdcav* h, int which, int* iters) {
  CFD_REQUIRE(h && which >= 0 && which <= 2, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  int rc = sip_alloc(h);
  if (rc) return rc;
  const int ni = h->cfg.ni, nj = h->cfg.nj;
  const double* coef[5] = {h->f[CFD_FDC_AW], h->f[CFD_FDC_AE], h->f[CFD_FDC_AS], h->f[CFD_FDC_AN], h->f[CFD_FDC_AP]};
  double* phi = which == 0 ? h->f[CFD_FDC_U] : which == 1 ? h->f[CFD_FDC_V] : h->f[CFD_FDC_PP];
  const int imax = which == 0 ? ni - 2 : ni - 1, jmax = which == 1 ? nj - 2 : nj - 1;
  rc = cfd_launch_sip(h->device, h->stream, ni + 1, nj + 1, imax, jmax, nullptr, coef, 0.80, 1.0e-6, 3000, h->f[CFD_FDC_SU], phi,
                      h->sip_work, h->sip_prog, h->sip_partial, h->sip_info, h->sip_rsm, 1024);
  return rc ? rc : sip_finish(h, iters);
}
extern "C" int cfd_fdcav_implicit_correct(cfd_fdcav* h, double* error) {
  CFD_REQUIRE(h, "null context");
  CFD_CUDA(cudaSetDevice(h->device));
  FdcParams P{};
  fdc_fill(h, P);
  dim3 block(32, 8), grid = fdc_grid(h);
  k_fdci_correct<<<grid, block, 0, h->stream>>>(P);
  k_fdc_bc<<<1, 1024, 0, h->stream>>>(P);
  k_fdc_norm<<<grid, block, 0, h->stream>>>(P);
  k_fdc_norm_final<<<1, 1024, 0, h->stream>>>(h->partial, h->npartial);
  CFD_CUDA(cudaGetLastError());
  if (error) {
    CFD_CUDA(cudaMemcpyAsync(error, h->partial, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CFD_CUDA(cudaStreamSynchronize(h->stream));
  }
  return CFD_OK;
}
// one pass of the time loop :177-288
extern "C" int cfd_fdcav_implicit_step(cfd_fdcav* h, double* error, int iters[3]) {
  int rc, it = 0;
  for (int w = 0; w < 3; ++w) {
    if ((rc = cfd_fdcav_implicit_assemble(h, w))) return rc;
    if ((rc = cfd_fdcav_implicit_solve(h, w, &it))) return rc;
    if (iters) iters[w] = it;
  }
  return cfd_fdcav_implicit_correct(h, error);
}

// one pass of the time loop fdm_explicit_cavity.f90:168-255 on the device
extern "C" int cfd_fdcav_step(cfd_fdcav* h, double* error, int* sip_iters) {
  int rc = cfd_fdcav_predict(h);
  if (rc) return rc;
  if ((rc = cfd_fdcav_solve_pp(h, sip_iters))) return rc;
  return cfd_fdcav_correct(h, error);
}
