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
int cfd_launch_sip(int device, cudaStream_t stream, int ni, int nj, double aw, double ae, double as, double an, double ap,
                   const double* su, double* phi, double* work, unsigned long long* prog, double* partial, int* info,
                   double* rsm_out, int max_partials);

namespace cfd {

struct FdcParams {
  int ni, nj, ld;
  double dx, dy, dt, re, utop, urfp;
  double *u, *v, *p, *uo, *vo, *po, *pp, *su;
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

}  // namespace cfd
using namespace cfd;

static void fdc_fill(const cfd_fdcav* h, FdcParams& P) {
  P.ni = h->cfg.ni; P.nj = h->cfg.nj; P.ld = h->cfg.ni + 1;
  P.dx = 1.0 / (h->cfg.ni - 1); P.dy = 1.0 / (h->cfg.nj - 1);
  P.dt = h->cfg.dt; P.re = h->cfg.re; P.utop = h->cfg.utop; P.urfp = h->cfg.urfp;
  P.u = h->f[CFD_FDC_U]; P.v = h->f[CFD_FDC_V]; P.p = h->f[CFD_FDC_P]; P.uo = h->f[CFD_FDC_UO]; P.vo = h->f[CFD_FDC_VO];
  P.po = h->f[CFD_FDC_PO]; P.pp = h->f[CFD_FDC_PP]; P.su = h->f[CFD_FDC_SU];
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

// sipsol (fdm_explicit_cavity.f90:24-87) for the p' system of the current step, entirely on the device
extern "C" int cfd_fdcav_solve_pp(cfd_fdcav* h, int* iters) {
  CFD_REQUIRE(h, "null context");
  CFD_CUDA(cudaSetDevice(h->device));
  const int ni = h->cfg.ni, nj = h->cfg.nj;
  const size_t n = (size_t)(ni + 1) * (nj + 1);
  if (!h->sip_work) {
    CFD_CUDA(cudaMalloc(&h->sip_work, 6 * n * sizeof(double)));
    CFD_CUDA(cudaMalloc(&h->sip_prog, (size_t)((nj + 31) / 32 + 1) * sizeof(unsigned long long)));
    CFD_CUDA(cudaMalloc(&h->sip_partial, 1024 * sizeof(double)));
    CFD_CUDA(cudaMalloc(&h->sip_info, 2 * sizeof(int)));
    CFD_CUDA(cudaMalloc(&h->sip_rsm, sizeof(double)));
  }
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), dt = h->cfg.dt;
  const double ax = dt / (dx * dx), ay = dt / (dy * dy);
  const double ap = 2 * (dt / (dx * dx) + dt / (dy * dy));
  int rc = cfd_launch_sip(h->device, h->stream, ni, nj, -ax, -ax, -ay, -ay, ap, h->f[CFD_FDC_SU], h->f[CFD_FDC_PP],
                          h->sip_work, h->sip_prog, h->sip_partial, h->sip_info, h->sip_rsm, 1024);
  if (rc) return rc;
  int info[2] = {0, 0};
  CFD_CUDA(cudaMemcpyAsync(info, h->sip_info, sizeof(info), cudaMemcpyDeviceToHost, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  if (info[1]) {
    cfd_set_error("cfd_fdcav_solve_pp: wavefront wait abandoned (pipeline stalled)");
    return CFD_ERR_CUDA;
  }
  if (iters) *iters = info[0];
  return CFD_OK;
}

// one pass of the time loop fdm_explicit_cavity.f90:168-255 on the device
extern "C" int cfd_fdcav_step(cfd_fdcav* h, double* error, int* sip_iters) {
  int rc = cfd_fdcav_predict(h);
  if (rc) return rc;
  if ((rc = cfd_fdcav_solve_pp(h, sip_iters))) return rc;
  return cfd_fdcav_correct(h, error);
}
