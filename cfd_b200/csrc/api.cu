// api.cu -- the extern "C" surface of libcfd_b200.so for the SIMPLE families (include/cfd_b200.h):
// context life cycle, host<->device layout conversion, and the composition of the stage kernels into the
// reference's subroutines calcu / calcv / calcp / lisolv and the outer-loop body.
#include <cstdarg>
#include <cmath>
#include "common.cuh"

using namespace cfd;

static thread_local char g_err[1024] = "";
void cfd_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* cfd_last_error(void) { return g_err; }
extern "C" const char* cfd_version(void) { return "cfd_b200 0.1 (sm_100a, fp64)"; }
extern "C" const char* cfd_strerror(int s) {
  switch (s) {
    case CFD_OK: return "ok";
    case CFD_ERR_INVALID: return "invalid argument";
    case CFD_ERR_CUDA: return "CUDA runtime error";
    case CFD_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
    case CFD_ERR_UNSUPPORTED: return "unsupported in this build";
    case CFD_ERR_NOMEM: return "out of device memory";
    case CFD_ERR_DIVERGED: return "non-finite residual (the iteration diverged or a field holds NaN/Inf)";
    default: return "unknown status";
  }
}

// ---- layout conversion kernels -----------------------------------------------------------------------
namespace cfd {
// column-major image A(ni,nj) (i fastest)  ->  line layout.  One 32x32 tile of (i0, t) per block and k.
__global__ void __launch_bounds__(256) k_to_line(const Layout L, const double* __restrict__ src, double* __restrict__ dst,
                                                 long long SL, int SK) {
  __shared__ double tile[32][33];
  const int k = blockIdx.z, ib = blockIdx.x * 32, tb = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int t = tb + r, i0 = ib + threadIdx.x, j0 = t * L.M + k;
    tile[r][threadIdx.x] = (t < L.T && i0 < L.ni && j0 < L.nj) ? src[(size_t)j0 * L.ni + i0] : 0.0;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int i0 = ib + r, t = tb + threadIdx.x, j0 = t * L.M + k;
    if (t < L.T && i0 < L.ni && j0 < L.nj) dst[(size_t)i0 * SL + (size_t)k * SK + t] = tile[threadIdx.x][r];
  }
}
__global__ void __launch_bounds__(256) k_from_line(const Layout L, const double* __restrict__ src, double* __restrict__ dst,
                                                   long long SL, int SK) {
  __shared__ double tile[32][33];
  const int k = blockIdx.z, ib = blockIdx.x * 32, tb = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int i0 = ib + r, t = tb + threadIdx.x, j0 = t * L.M + k;
    tile[r][threadIdx.x] = (t < L.T && i0 < L.ni && j0 < L.nj) ? src[(size_t)i0 * SL + (size_t)k * SK + t] : 0.0;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int t = tb + r, i0 = ib + threadIdx.x, j0 = t * L.M + k;
    if (t < L.T && i0 < L.ni && j0 < L.nj) dst[(size_t)j0 * L.ni + i0] = tile[threadIdx.x][r];
  }
}
}  // namespace cfd

int cfd_launch_to_line_layout(cfd_ctx* c, const double* colmajor, double* dst, int field) {
  dim3 block(32, 8), grid((c->L.ni + 31) / 32, (c->L.T + 31) / 32, c->L.M);
  const bool cf = cfd_is_coef_field(field);
  k_to_line<<<grid, block, 0, c->stream>>>(c->L, colmajor, dst, cf ? c->L.cSL : (long long)c->L.LD, cf ? c->L.cSK : c->L.T);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
int cfd_launch_from_line_layout(cfd_ctx* c, const double* src, double* colmajor, int field) {
  dim3 block(32, 8), grid((c->L.ni + 31) / 32, (c->L.T + 31) / 32, c->L.M);
  const bool cf = cfd_is_coef_field(field);
  k_from_line<<<grid, block, 0, c->stream>>>(c->L, src, colmajor, cf ? c->L.cSL : (long long)c->L.LD, cf ? c->L.cSK : c->L.T);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

// ---- layout choice -------------------------------------------------------------------------------------
// T threads per line (multiple of 32, <= 1024), M nodes per thread (<= CFD_MAX_M), T*M >= nj.  Among the
// (M,T) pairs whose line-solver instantiation keeps its chunk in registers (ptxas: no spills) take the least
// padding, ties to the larger M (less parallel-cyclic-reduction work per node); if none fits, least padding.
// a layout whose lines split evenly over a cluster of blocks of <= 320 threads also fits (tdma_cluster_kernels.cu)
static bool fits_registers(int m, int t) { return t <= 384 || (cfd_cluster_size(t) > 0 && m >= 4) || (t <= 640 && m <= 9) || m <= 5; }

static int choose_layout(int ni, int nj, int want_m, Layout* L) {
  long best = -1;
  int bm = 0, bt = 0;
  for (int pass = 0; pass < 2 && best < 0; ++pass) {
    for (int m = CFD_MAX_M; m >= 1; --m) {
      if (want_m > 0 && m != want_m) continue;
      int t = ((nj + m - 1) / m + 31) / 32 * 32;
      if (t > 1024) {
        // longer than any single-block solver: the line is split over a cluster of 4 or 8 blocks of <= 320 threads
        // (tdma_cluster_kernels.cu), so T must be a multiple of 32 * cluster size
        if (m < 4) continue;
        int tt = 0;
        for (int cs = 4; cs <= 8 && !tt; cs *= 2) {
          const int r = (t + 32 * cs - 1) / (32 * cs) * (32 * cs);
          if (r / cs <= 320) tt = r;
        }
        if (!tt) continue;
        t = tt;
      }
      if (pass == 0 && t <= 1024 && !fits_registers(m, t)) continue;
      long ld = (long)t * m;
      if (best < 0 || ld < best) {
        best = ld;
        bm = m;
        bt = t;
      }
    }
  }
  if (best < 0) return CFD_ERR_UNSUPPORTED;
  L->ni = ni;
  L->nj = nj;
  L->M = bm;
  L->T = bt;
  L->LD = (int)best;
  L->nc = 6;
  L->cSL = 6LL * L->LD;
  L->cSK = 6 * L->T;
  return CFD_OK;
}

static const int kCavityFields[] = {CFD_U, CFD_V, CFD_P, CFD_PP, CFD_VIS, CFD_DEN, CFD_DU, CFD_DV};
// order of the six arrays inside the interleaved coefficient block (Layout::cSL/cSK)
static const int kCoefOrder6[6] = {CFD_AN, CFD_AS, CFD_AP, CFD_AE, CFD_AW, CFD_SU};
static const int kCoefOrder5[5] = {CFD_AN, CFD_AS, CFD_AE, CFD_AW, CFD_SU};

static bool geom_is_j(int id) {
  switch (id) {
    case CFD_G_X: case CFD_G_XU: case CFD_G_SEW: case CFD_G_SEWU:
    case CFD_G_XDIF: case CFD_G_XCVI: case CFD_G_XCVIP: case CFD_G_FX: case CFD_G_FXM:
    case CFD_G_DXEP: case CFD_G_DXPW: case CFD_G_DXEPU: case CFD_G_DXPWU:
      return false;
    default:
      return true;
  }
}

// ---- upload for callers that must not allocate ----------------------------------------------------------
// cfd_upload() stages the host image in a device buffer that it cudaMalloc's on its first call.  Some hosts of
// this library cannot tolerate that: a driver that has already handed all free HBM to its own pool (an
// OpenACC / CUDA-Fortran runtime arena, an MPS client at its memory limit) gets an out-of-memory error from any
// later cudaMalloc, and a time-stepping loop that re-uploads boundary data every iteration (the `boundary()` /
// `gamsor()` hooks of buoyancy_cav.f90) must not hit cudaMalloc's device-wide synchronisation in the middle of
// the loop.  For them the staging area below is part of the library image itself.  With CUDA's lazy module loading
// (the default) the array gets its device memory when this module is first touched on a GPU, so cfd_create --
// which allocates anyway -- looks its address up on the context's device (cfd_noalloc_resolve); from then until
// the process exits nothing is allocated or freed by cfd_upload_noalloc.
#define CFD_NOALLOC_MAX_DOUBLES (4u * 1024u * 1024u) /* 32 MiB: grids up to ni*nj = 4 194 304 nodes (2048 x 2048) */
static __device__ __align__(256) double g_noalloc_stage[CFD_NOALLOC_MAX_DOUBLES];
// Device addresses of g_noalloc_stage, one per device ordinal.  In host code the symbol itself is only a
// stand-in; the address a kernel or a memcpy may use comes from cudaGetSymbolAddress with that device current.
static double* g_noalloc_addr[64] = {nullptr};

// Look the staging area's address up on device `dev` (which must be the current device).  Called by cfd_create.
static int cfd_noalloc_resolve(int dev) {
  if (dev < 0 || dev >= 64 || g_noalloc_addr[dev]) return CFD_OK;   // ordinals >= 64: cfd_upload_noalloc refuses them
  double* p = nullptr;
  CFD_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_stage));
  g_noalloc_addr[dev] = p;
  return CFD_OK;
}

extern "C" int cfd_create(cfd_ctx** out, const cfd_config* cfg) {
  CFD_REQUIRE(out && cfg, "null argument");
  *out = nullptr;
  CFD_REQUIRE(cfg->ni >= 4 && cfg->nj >= 4, "ni, nj must be >= 4 (two boundary nodes + two cells)");
  CFD_REQUIRE(cfg->family == CFD_FAMILY_CAVITY || cfg->family == CFD_FAMILY_GENERAL || cfg->family == CFD_FAMILY_TEACH,
              "family must be CFD_FAMILY_CAVITY, CFD_FAMILY_GENERAL or CFD_FAMILY_TEACH");
  const bool general = cfg->family == CFD_FAMILY_GENERAL, teach = cfg->family == CFD_FAMILY_TEACH;
  CFD_REQUIRE(cfg->iadv >= 1 && cfg->iadv <= (general ? 4 : 3), "iadv out of range");
  if (teach) {
    CFD_REQUIRE(cfg->iadv == 3, "the TEACH family has the hybrid scheme only (calc.f90:69-72): iadv must be 3");
    CFD_REQUIRE(cfg->indcos == 1 || cfg->indcos == 2, "indcos must be 1 (plane) or 2 (axisymmetric)");
    CFD_REQUIRE(cfg->istep >= 2 && cfg->jstep >= 2 && cfg->jstep < cfg->nj - 1, "istep / jstep out of range");
    CFD_REQUIRE(cfg->ni >= 6 && cfg->nj >= 6, "TEACH family needs ni, nj >= 6");
  }
  if (general) {
    // the reference's loops use nj where ni is meant (buoyancy_cav.f90:135,150,151,390): only ni == nj is well defined
    const int NIg = cfg->ni_global ? cfg->ni_global : cfg->ni;
    CFD_REQUIRE(!cfg->solve_flow || NIg == cfg->nj, "the general family's flow equations need ni == nj (reference quirk)");
    CFD_REQUIRE(cfg->nj <= NIg, "the general family's calct reads den(nj,j): nj must not exceed ni");
    CFD_REQUIRE(cfg->ni >= 6 && cfg->nj >= 6, "general family needs ni, nj >= 6");
  }
  CFD_REQUIRE(cfg->order == CFD_ORDER_REDBLACK || cfg->order == CFD_ORDER_SEQUENTIAL, "bad order");
  CFD_REQUIRE(cfg->halo_w >= 0 && cfg->halo_e >= 0 && cfg->halo_w + cfg->halo_e < cfg->ni, "bad halo widths");
  CFD_REQUIRE(cfg->halo_s >= 0 && cfg->halo_n >= 0 && cfg->halo_s + cfg->halo_n < cfg->nj, "bad halo row counts");
  if (cfg->halo_s || cfg->halo_n || cfg->joff) {
    CFD_REQUIRE(cfg->family == CFD_FAMILY_CAVITY && cfg->order == CFD_ORDER_REDBLACK,
                "j-slab contexts: cavity family, red/black order only");
    CFD_REQUIRE(cfg->nj_global >= cfg->nj, "j-slab contexts need nj_global");
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cfd_set_error("no CUDA device visible; libcfd_b200 has no CPU fallback");
    return CFD_ERR_NO_DEVICE;
  }
  int dev = cfg->device;
  if (dev < 0) CFD_CUDA(cudaGetDevice(&dev));
  CFD_REQUIRE(dev < ndev, "device ordinal out of range");
  CFD_CUDA(cudaSetDevice(dev));

  cfd_ctx* c = new cfd_ctx();
  c->cfg = *cfg;
  c->device = dev;
  c->stream = 0;
  for (auto& f : c->field) f = nullptr;
  for (auto& g : c->geom) g = nullptr;
  for (auto& g : c->geomn) g = nullptr;
  c->boussinesq = 0;
  c->rgbeta = c->tref = 0.0;
  c->has_t_sources = 0;
  c->stage = nullptr;
  c->coef = nullptr;
  c->partial = nullptr;
  c->res_dev = nullptr;
  c->seq_scratch = nullptr;
  c->launches = 0;
  c->uni_den = c->uni_vis = false;
  c->den0 = c->vis0 = 0.0;
  {
    const char* e = getenv("CFD_B200_UNIFORM");      // read per context: tests switch it between contexts
    c->allow_uniform = !(e && atoi(e) == 0);
  }
  c->flags = nullptr;
  c->dbg = nullptr;
  c->post_buf = nullptr;
  c->post_n = 0;
  c->epoch = 1;
  c->xepoch = 0;
  c->bcast_n = 0;
  c->bepoch = 0;
  for (auto& bf : c->bcast_flags) bf = nullptr;
  for (int sd = 0; sd < 2; ++sd) {
    c->peer_attached[sd] = false;
    c->peer_flags[sd] = nullptr;
    c->peer_ni[sd] = c->peer_hw[sd] = c->peer_halo_line[sd] = 0;
    for (auto& pf : c->peer_field[sd]) pf = nullptr;
  }
  // from here on every failure path releases what has been allocated so far
#define CREATE_CUDA(call)                                                                       \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess) {                                                                   \
      cfd_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));     \
      cfd_destroy(c);                                                                           \
      return e__ == cudaErrorMemoryAllocation ? CFD_ERR_NOMEM : CFD_ERR_CUDA;                   \
    }                                                                                           \
  } while (0)
  CREATE_CUDA(cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, dev));
  CREATE_CUDA(cudaDeviceGetAttribute(&c->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  for (auto& g : c->gfast) g = nullptr;
  for (auto& g : c->tfast) g = nullptr;
  c->spike_w = c->spike_z = c->spike_small = nullptr;
  c->turbulent = false;
  for (auto& w : c->wall) w = nullptr;
  c->tmult_n = c->tmult_w = nullptr;
  c->tfast_ready = false;
  int want_m = cfg->tdma_m;
  if (want_m == 0) {
    const char* e = getenv("CFD_B200_TDMA_M");
    if (e) want_m = atoi(e);
  }
  if (choose_layout(cfg->ni, cfg->nj, want_m, &c->L) != CFD_OK) {
    delete c;
    cfd_set_error("nj=%d too long for the line layout (max %d)", cfg->nj, 2560 * CFD_MAX_M);
    return CFD_ERR_UNSUPPORTED;
  }
  {
    const char* e = getenv("CFD_B200_EXACT_ASSEMBLY");
    c->exact_assembly = (cfg->order == CFD_ORDER_SEQUENTIAL) || (e && atoi(e) != 0) || cfg->raw_coefficients;
    if ((teach || general) && c->exact_assembly &&
        !(cfg->goff == 0 && cfg->halo_w == 0 && cfg->halo_e == 0 && (cfg->ni_global == 0 || cfg->ni_global == cfg->ni))) {
      delete c;
      cfd_set_error("TEACH / general-family i-slab contexts need the red/black order with the production assembly kernels");
      return CFD_ERR_UNSUPPORTED;
    }
    c->L.nc = c->exact_assembly ? 6 : 5;
    c->L.cSL = (long long)c->L.nc * c->L.LD;
    c->L.cSK = c->L.nc * c->L.T;
  }
  const size_t fbytes = (size_t)c->L.ni * c->L.LD * sizeof(double);
  std::vector<int> flist(kCavityFields, kCavityFields + sizeof(kCavityFields) / sizeof(int));
  if (general) {
    const int extra[] = {CFD_T, CFD_CP, CFD_GAMT, CFD_SUSRC, CFD_SPSRC};
    flist.insert(flist.end(), extra, extra + 5);
  }
  for (int f : flist) {
    if (cudaMalloc(&c->field[f], fbytes) != cudaSuccess) {
      cfd_set_error("cudaMalloc of %zu bytes for field %d failed", fbytes, f);
      cfd_destroy(c);
      return CFD_ERR_NOMEM;
    }
    CREATE_CUDA(cudaMemset(c->field[f], 0, fbytes));
  }
  if (cudaMalloc(&c->coef, c->L.nc * fbytes) != cudaSuccess) {
    cfd_set_error("cudaMalloc of %zu bytes for the coefficient block failed", c->L.nc * fbytes);
    cfd_destroy(c);
    return CFD_ERR_NOMEM;
  }
  CREATE_CUDA(cudaMemset(c->coef, 0, c->L.nc * fbytes));
  for (int a = 0; a < c->L.nc; ++a)
    c->field[c->L.nc == 6 ? kCoefOrder6[a] : kCoefOrder5[a]] = c->coef + (size_t)a * c->L.T;
  for (int g = 0; g < CFD_GEOM_COUNT; ++g) {
    const size_t n = geom_is_j(g) ? (size_t)c->L.LD : (size_t)c->L.ni + 2;
    c->geom_host[g].assign(geom_is_j(g) ? c->L.nj : c->L.ni, 0.0);
    CREATE_CUDA(cudaMalloc(&c->geom[g], n * sizeof(double)));
    CREATE_CUDA(cudaMemset(c->geom[g], 0, n * sizeof(double)));
    const size_t nn = (size_t)(geom_is_j(g) ? c->L.nj : c->L.ni) + 2;
    CREATE_CUDA(cudaMalloc(&c->geomn[g], nn * sizeof(double)));
    CREATE_CUDA(cudaMemset(c->geomn[g], 0, nn * sizeof(double)));
  }
  for (int g = 0; g < 8; ++g) {
    const size_t n = (g >= 4) ? (size_t)c->L.LD : (size_t)c->L.ni + 2;
    CREATE_CUDA(cudaMalloc(&c->gfast[g], n * sizeof(double)));
    CREATE_CUDA(cudaMemset(c->gfast[g], 0, n * sizeof(double)));
  }
  c->max_partials = ((c->L.T + 31) / 32) * ((c->L.ni + 3) / 4);
  CREATE_CUDA(cudaMalloc(&c->partial, (size_t)CFD_EQ_COUNT * c->max_partials * sizeof(double)));
  CREATE_CUDA(cudaMalloc(&c->res_dev, CFD_RES_SLOTS * sizeof(double)));
  CREATE_CUDA(cudaMemset(c->res_dev, 0, CFD_RES_SLOTS * sizeof(double)));
  // per-line epochs, handshake slots, per-line arrival counters (lines split over a cluster of blocks)
  CREATE_CUDA(cudaMalloc(&c->flags, (2 * (size_t)c->L.ni + CFD_FLAG_SLOTS) * sizeof(int) + CFD_MAILBOX_DOUBLES * sizeof(double)));
  CREATE_CUDA(cudaMemset(c->flags, 0, (2 * (size_t)c->L.ni + CFD_FLAG_SLOTS) * sizeof(int) + CFD_MAILBOX_DOUBLES * sizeof(double)));
  {
    // cfd_upload_noalloc's staging area: with lazy module loading the array is only materialised when something in
    // this module is first touched, so touch it HERE, while allocating is still allowed (see cfd_upload_noalloc)
    int rcn = cfd_noalloc_resolve(dev);
    if (rcn) {
      cfd_destroy(c);
      return rcn;
    }
  }
#undef CREATE_CUDA
  *out = c;
  return CFD_OK;
}

extern "C" int cfd_destroy(cfd_ctx* c) {
  if (!c) return CFD_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (int a = 0; a < 6; ++a) c->field[kCoefOrder6[a]] = nullptr;
  if (c->coef) cudaFree(c->coef);
  for (auto& f : c->field) if (f) cudaFree(f);
  for (auto& g : c->geom) if (g) cudaFree(g);
  for (auto& g : c->geomn) if (g) cudaFree(g);
  for (auto& g : c->gfast) if (g) cudaFree(g);
  for (auto& g : c->tfast) if (g) cudaFree(g);
  if (c->spike_w) cudaFree(c->spike_w);
  if (c->spike_z) cudaFree(c->spike_z);
  if (c->spike_small) cudaFree(c->spike_small);
  for (auto& w : c->wall) if (w) cudaFree(w);
  if (c->tmult_n) cudaFree(c->tmult_n);
  if (c->tmult_w) cudaFree(c->tmult_w);
  if (c->stage) cudaFree(c->stage);
  if (c->partial) cudaFree(c->partial);
  if (c->res_dev) cudaFree(c->res_dev);
  for (void* m : c->ipc_opened) cudaIpcCloseMemHandle(m);
  if (c->flags) cudaFree(c->flags);
  if (c->post_buf) cudaFree(c->post_buf);
  if (c->dbg) cudaFree(c->dbg);
  if (c->seq_scratch) cudaFree(c->seq_scratch);
  delete c;
  return CFD_OK;
}

extern "C" int cfd_set_stream(cfd_ctx* c, void* s) {
  CFD_REQUIRE(c, "null context");
  c->stream = (cudaStream_t)s;
  return CFD_OK;
}

extern "C" int cfd_synchronize(cfd_ctx* c) {
  CFD_REQUIRE(c, "null context");
  CFD_CUDA(cudaSetDevice(c->device));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_set_geometry(cfd_ctx* c, int id, const double* host, int n) {
  CFD_REQUIRE(c && host, "null argument");
  CFD_REQUIRE(id >= 0 && id < CFD_GEOM_COUNT, "bad geometry id");
  const bool isj = geom_is_j(id);
  CFD_REQUIRE(n == (isj ? c->L.nj : c->L.ni), "geometry array length must be nj (y-arrays) or ni (x-arrays)");
  CFD_CUDA(cudaSetDevice(c->device));
  c->geom_host[id].assign(host, host + n);
  c->tfast_ready = false;
  CFD_CUDA(cudaMemcpyAsync(c->geomn[id], host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (!isj) {
    CFD_CUDA(cudaMemcpyAsync(c->geom[id], host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CFD_CUDA(cudaStreamSynchronize(c->stream));
  } else {
    std::vector<double> tmp((size_t)c->L.LD, 0.0);
    for (int j0 = 0; j0 < n; ++j0) tmp[lpos(j0, c->L.M, c->L.T)] = host[j0];
    CFD_CUDA(cudaMemcpyAsync(c->geom[id], tmp.data(), tmp.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CFD_CUDA(cudaStreamSynchronize(c->stream));
  }
  if (id <= CFD_G_SNSV) return cfd_update_fast_geometry(c);
  return CFD_OK;
}

// Recompute the reciprocal geometry of the production assembly kernels from the host copies.
int cfd_update_fast_geometry(cfd_ctx* c) {
  const int ni = c->L.ni, nj = c->L.nj;
  auto inv = [](double d) { return d != 0.0 ? 1.0 / d : 0.0; };
  const std::vector<double>&x = c->geom_host[CFD_G_X], &xu = c->geom_host[CFD_G_XU], &sew = c->geom_host[CFD_G_SEW],
                     &sewu = c->geom_host[CFD_G_SEWU], &y = c->geom_host[CFD_G_Y], &yv = c->geom_host[CFD_G_YV],
                     &sns = c->geom_host[CFD_G_SNS], &snsv = c->geom_host[CFD_G_SNSV];
  std::vector<double> bi[4], bj[4];
  for (auto& v : bi) v.assign((size_t)ni + 2, 0.0);
  for (auto& v : bj) v.assign((size_t)c->L.LD, 0.0);
  for (int i = 0; i < ni; ++i) {
    if (i + 1 < ni) {
      bi[0][i] = inv(xu[i + 1] - xu[i]);
      bi[1][i] = inv(x[i + 1] - x[i]);
    }
    bi[2][i] = inv(sew[i]);
    bi[3][i] = inv(sewu[i]);
  }
  for (int j = 0; j < nj; ++j) {
    const int q = lpos(j, c->L.M, c->L.T);
    if (j + 1 < nj) {
      bj[0][q] = inv(y[j + 1] - y[j]);
      bj[1][q] = inv(yv[j + 1] - yv[j]);
    }
    bj[2][q] = inv(sns[j]);
    bj[3][q] = inv(snsv[j]);
  }
  for (int g = 0; g < 4; ++g) {
    CFD_CUDA(cudaMemcpyAsync(c->gfast[g], bi[g].data(), bi[g].size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CFD_CUDA(cudaMemcpyAsync(c->gfast[4 + g], bj[g].data(), bj[g].size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

static int ensure_stage(cfd_ctx* c) {
  if (!c->stage) {
    if (cudaMalloc(&c->stage, (size_t)c->L.ni * c->L.nj * sizeof(double)) != cudaSuccess) {
      cfd_set_error("cudaMalloc of the upload/download staging buffer failed");
      return CFD_ERR_NOMEM;
    }
  }
  return CFD_OK;
}

// cavity family: remember whether an uploaded den / vis image holds a single value (then the production assembly
// kernels read it as a scalar, cavity_tma_kernels.cu).  The reference's driver fills both by whole-array assignment
// (cavity_fvm.f90:536-537), so this is the usual case; any non-uniform upload switches the field path back on.
static void note_uniform(cfd_ctx* c, int field, const double* host, int ld) {
  if (c->cfg.family != CFD_FAMILY_CAVITY || (field != CFD_DEN && field != CFD_VIS)) return;
  const double v0 = host[0];
  bool uni = c->allow_uniform;
  for (int j = 0; j < c->L.nj && uni; ++j) {
    const double* row = host + (size_t)j * ld;
    for (int i = 0; i < c->L.ni; ++i)
      if (row[i] != v0) { uni = false; break; }
  }
  if (field == CFD_DEN) { c->uni_den = uni; c->den0 = v0; }
  else { c->uni_vis = uni; c->vis0 = v0; }
}

static int check_field(const cfd_ctx* c, int field) {
  if (field < 0 || field >= CFD_FIELD_COUNT || !c->field[field]) {
    cfd_set_error("field id %d is not allocated in this context", field);
    return CFD_ERR_INVALID;
  }
  return CFD_OK;
}

extern "C" int cfd_upload(cfd_ctx* c, int field, const double* host, int ld) {
  CFD_REQUIRE(c && host, "null argument");
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(ld >= c->L.ni, "ld < ni");
  CFD_CUDA(cudaSetDevice(c->device));
  if ((rc = ensure_stage(c))) return rc;
  CFD_CUDA(cudaMemcpy2DAsync(c->stage, (size_t)c->L.ni * sizeof(double), host, (size_t)ld * sizeof(double),
                             (size_t)c->L.ni * sizeof(double), c->L.nj, cudaMemcpyHostToDevice, c->stream));
  note_uniform(c, field, host, ld);      // scans the host image while the copy is in flight
  if ((rc = cfd_launch_to_line_layout(c, c->stage, c->field[field], field))) return rc;
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_upload_noalloc(cfd_ctx* c, int field, const double* host, int ld) {
  CFD_REQUIRE(c && host, "null argument");
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(ld >= c->L.ni, "ld < ni");
  const size_t n = (size_t)c->L.ni * (size_t)c->L.nj;
  if (n > (size_t)CFD_NOALLOC_MAX_DOUBLES) {
    cfd_set_error("cfd_upload_noalloc: ni*nj = %zu nodes exceeds the built-in staging area of %u doubles (32 MiB); "
                  "use cfd_upload for larger grids", n, CFD_NOALLOC_MAX_DOUBLES);
    return CFD_ERR_INVALID;
  }
  CFD_REQUIRE(c->device >= 0 && c->device < 64, "device ordinal outside the supported range 0..63");
  CFD_CUDA(cudaSetDevice(c->device));
  double* stage = g_noalloc_addr[c->device];   // looked up by cfd_create on this device
  CFD_REQUIRE(stage, "the built-in staging area was not resolved on this device (cfd_create does it)");
  // the image is overwritten completely (ni*nj doubles), so nothing left by an earlier call is ever read
  CFD_CUDA(cudaMemcpy2DAsync(stage, (size_t)c->L.ni * sizeof(double), host, (size_t)ld * sizeof(double),
                             (size_t)c->L.ni * sizeof(double), c->L.nj, cudaMemcpyHostToDevice, c->stream));
  note_uniform(c, field, host, ld);
  if ((rc = cfd_launch_to_line_layout(c, stage, c->field[field], field))) return rc;
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_download(cfd_ctx* c, int field, double* host, int ld) {
  CFD_REQUIRE(c && host, "null argument");
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(ld >= c->L.ni, "ld < ni");
  CFD_CUDA(cudaSetDevice(c->device));
  if ((rc = ensure_stage(c))) return rc;
  if ((rc = cfd_launch_from_line_layout(c, c->field[field], c->stage, field))) return rc;
  CFD_CUDA(cudaMemcpy2DAsync(host, (size_t)ld * sizeof(double), c->stage, (size_t)c->L.ni * sizeof(double),
                             (size_t)c->L.ni * sizeof(double), c->L.nj, cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_probe(cfd_ctx* c, int field, int i, int j, double* value) {
  CFD_REQUIRE(c && value, "null argument");
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(i >= 1 && i <= c->L.ni && j >= 1 && j <= c->L.nj, "probe index out of range");
  CFD_CUDA(cudaSetDevice(c->device));
  const double* src = c->field[field] + (size_t)(i - 1) * c->L.LD + lpos(j - 1, c->L.M, c->L.T);
  CFD_CUDA(cudaMemcpyAsync(value, src, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_field_device_view(cfd_ctx* c, int field, void** base, long long* line_stride, int* nlines) {
  CFD_REQUIRE(c, "null context");
  int rc = check_field(c, field);
  if (rc) return rc;
  if (base) *base = c->field[field];
  if (line_stride) *line_stride = cfd_is_coef_field(field) ? c->L.cSL : (long long)c->L.LD;
  if (nlines) *nlines = c->L.ni;
  return CFD_OK;
}

extern "C" long long cfd_launch_count(cfd_ctx* c) { return c ? c->launches : -1; }

extern "C" int cfd_layout_info(cfd_ctx* c, int* m, int* t, int* ld) {
  CFD_REQUIRE(c, "null context");
  if (m) *m = c->L.M;
  if (t) *t = c->L.T;
  if (ld) *ld = c->L.LD;
  return CFD_OK;
}

// ---- stage entry points ----------------------------------------------------------------------------------
static inline bool is_general(const cfd_ctx* c) { return c->cfg.family == CFD_FAMILY_GENERAL; }
static inline bool is_teach(const cfd_ctx* c) { return c->cfg.family == CFD_FAMILY_TEACH; }
static int asm_u(cfd_ctx* c) {
  if (is_teach(c) && c->turbulent) {
    int rc = cfd_launch_teach_wall(c, 0);
    if (rc) return rc;
  }
  return is_general(c) ? cfd_launch_general_assemble(c, 0) : is_teach(c) ? cfd_launch_teach_assemble(c, 0) : cfd_launch_assemble_u(c);
}
static int asm_v(cfd_ctx* c) {
  if (is_teach(c) && c->turbulent) {
    int rc = cfd_launch_teach_wall(c, 1);
    if (rc) return rc;
  }
  return is_general(c) ? cfd_launch_general_assemble(c, 1) : is_teach(c) ? cfd_launch_teach_assemble(c, 1) : cfd_launch_assemble_v(c);
}
static int asm_p(cfd_ctx* c) {
  return is_general(c) ? cfd_launch_general_assemble(c, 2) : is_teach(c) ? cfd_launch_teach_assemble(c, 2) : cfd_launch_assemble_p(c);
}
// Every entry point that launches or copies makes the context's device current first: one process may hold
// contexts on several GPUs and call them in any order.
#define CFD_ENTER(c)                         \
  do {                                       \
    CFD_REQUIRE(c, "null context");          \
    CFD_CUDA(cudaSetDevice((c)->device));    \
  } while (0)
extern "C" int cfd_assemble_u(cfd_ctx* c) { CFD_ENTER(c); return asm_u(c); }
extern "C" int cfd_assemble_v(cfd_ctx* c) { CFD_ENTER(c); return asm_v(c); }
extern "C" int cfd_assemble_p(cfd_ctx* c) { CFD_ENTER(c); return asm_p(c); }
extern "C" int cfd_assemble_t(cfd_ctx* c) {
  CFD_ENTER(c);
  CFD_REQUIRE(is_general(c) || (is_teach(c) && c->turbulent),
              "cfd_assemble_t needs a CFD_FAMILY_GENERAL context, or a TEACH context after cfd_set_turbulence");
  return is_general(c) ? cfd_launch_general_assemble(c, 3) : cfd_launch_teach_scalar(c, CFD_EQ_T);
}
extern "C" int cfd_set_hook(cfd_ctx* c, int eq, const cfd_hook_desc* d) {
  CFD_REQUIRE(c && d, "null argument");
  CFD_REQUIRE(is_general(c), "hooks belong to CFD_FAMILY_GENERAL contexts");
  if (eq == CFD_EQ_V && (d->kind == CFD_HOOK_BOUSSINESQ || d->kind == CFD_HOOK_NONE)) {
    c->boussinesq = d->kind == CFD_HOOK_BOUSSINESQ;
    c->rgbeta = d->params[0];
    c->tref = d->params[1];
    return CFD_OK;
  }
  if (eq == CFD_EQ_T && (d->kind == CFD_HOOK_FIELD_SOURCES || d->kind == CFD_HOOK_NONE)) {
    c->has_t_sources = d->kind == CFD_HOOK_FIELD_SOURCES;
    return CFD_OK;
  }
  cfd_set_error("cfd_set_hook: hook kind %d is not defined for equation %d", d->kind, eq);
  return CFD_ERR_UNSUPPORTED;
}
extern "C" int cfd_correct(cfd_ctx* c, const double* ppref) { CFD_ENTER(c); return cfd_launch_correct(c, ppref); }

extern "C" int cfd_sweep(cfd_ctx* c, int istart, int jstart, int field, int colour, int first) {
  CFD_ENTER(c);
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(istart >= 2 && jstart >= 2 && jstart <= c->L.nj - 1, "istart/jstart out of range");
  if (colour < 0) return cfd_launch_sweep_seq(c, istart, jstart, c->field[field], first);
  CFD_REQUIRE(colour <= 1, "colour must be 0, 1 or negative (sequential)");
  return cfd_launch_sweep_rb(c, istart, jstart, c->field[field], colour, first);
}

extern "C" int cfd_sweep_part(cfd_ctx* c, int istart, int jstart, int field, int colour, int first, int part) {
  CFD_ENTER(c);
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(istart >= 2 && jstart >= 2 && jstart <= c->L.nj - 1, "istart/jstart out of range");
  CFD_REQUIRE(colour == 0 || colour == 1, "colour must be 0 or 1");
  CFD_REQUIRE(part >= 0 && part <= 2, "part must be 0 (all lines), 1 (slab-edge lines) or 2 (the others)");
  return cfd_launch_sweep_rb_part(c, istart, jstart, c->field[field], colour, first, part);
}

static int full_sweep(cfd_ctx* c, int istart, int jstart, int field, int first) {
  int rc;
  if (c->cfg.order == CFD_ORDER_SEQUENTIAL) return cfd_launch_sweep_seq(c, istart, jstart, c->field[field], first);
  if ((rc = cfd_launch_sweep_rb(c, istart, jstart, c->field[field], 0, first))) return rc;
  return cfd_launch_sweep_rb(c, istart, jstart, c->field[field], 1, first);
}

// nsw sweeps of one equation: all colour passes in one flag-ordered launch where the layout allows it
// (tdma_flow_kernels.cu), else one launch per colour pass
static int sweeps(cfd_ctx* c, int istart, int jstart, int field, int nsw, int first) {
  if (nsw > 0 && cfd_flow_eligible(c)) return cfd_launch_sweeps_flow(c, istart, jstart, field, nsw, first);
  int rc;
  for (int n = 0; n < nsw; ++n)
    if ((rc = full_sweep(c, istart, jstart, field, first && n == 0))) return rc;
  return CFD_OK;
}

extern "C" int cfd_sweeps(cfd_ctx* c, int istart, int jstart, int field, int nsweeps, int first) {
  CFD_ENTER(c);
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(istart >= 2 && jstart >= 2 && jstart <= c->L.nj - 1 && nsweeps >= 0, "istart/jstart/nsweeps out of range");
  return sweeps(c, istart, jstart, field, nsweeps, first);
}

// ---- i-slab neighbours over peer memory ---------------------------------------------------------------------------
static void* ipc_target(cfd_ctx* c, int what) {
  if (what == CFD_IPC_FLAGS) return c->flags;
  if (what < 0 || what >= CFD_FIELD_COUNT || cfd_is_coef_field(what)) return nullptr;
  return c->field[what];
}
extern "C" int cfd_ipc_export(cfd_ctx* c, int what, void* handle64) {
  CFD_REQUIRE(c && handle64, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  void* p = ipc_target(c, what);
  CFD_REQUIRE(p, "cfd_ipc_export: not a plain field of this context (or CFD_IPC_FLAGS)");
  CFD_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  CFD_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  return CFD_OK;
}
extern "C" int cfd_ipc_open(cfd_ctx* c, const void* handle64, void** devptr) {
  CFD_REQUIRE(c && handle64 && devptr, "null argument");
  CFD_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  CFD_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  c->ipc_opened.push_back(p);
  *devptr = p;
  return CFD_OK;
}
extern "C" int cfd_peer_attach(cfd_ctx* c, int side, int what, void* devptr, int peer_ni, int peer_halo_w, int peer_halo_e) {
  CFD_REQUIRE(c && devptr, "null argument");
  CFD_REQUIRE(side == 0 || side == 1, "side must be 0 (west neighbour) or 1 (east neighbour)");
  CFD_REQUIRE((side == 0 ? c->cfg.halo_w : c->cfg.halo_e) > 0, "this context has no halo (no neighbour) on that side");
  CFD_REQUIRE(what == CFD_IPC_FLAGS || ipc_target(c, what), "not a plain field of this context (or CFD_IPC_FLAGS)");
  CFD_REQUIRE(peer_ni > 0 && peer_halo_w >= 0 && peer_halo_e >= 0, "bad neighbour extents");
  CFD_REQUIRE(side == 0 ? peer_halo_e >= 1 : peer_halo_w >= 1, "the neighbour has no halo line facing this context");
  CFD_CUDA(cudaSetDevice(c->device));
  cudaPointerAttributes at;
  CFD_CUDA(cudaPointerGetAttributes(&at, devptr));
  if (at.type == cudaMemoryTypeDevice && at.device != c->device) {   // same process, other GPU: map it
    int can = 0;
    CFD_CUDA(cudaDeviceCanAccessPeer(&can, c->device, at.device));
    CFD_REQUIRE(can, "the neighbour's GPU is not peer-accessible from this one");
    cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CFD_CUDA(e);
    (void)cudaGetLastError();
  }
  c->peer_ni[side] = peer_ni;
  c->peer_hw[side] = peer_halo_w;
  // where my edge line lands in the neighbour's arrays: its first east halo line / its last west halo line
  c->peer_halo_line[side] = side == 0 ? peer_ni - peer_halo_e : peer_halo_w - 1;
  if (what == CFD_IPC_FLAGS) c->peer_flags[side] = (int*)devptr;
  else c->peer_field[side][what] = (double*)devptr;
  c->peer_attached[side] = c->peer_flags[side] != nullptr;
  return CFD_OK;
}
extern "C" int cfd_bcast_attach(cfd_ctx* c, int rank, void* flag_block, int rank_ni) {
  CFD_REQUIRE(c, "null context");
  CFD_REQUIRE(rank >= 0 && rank < CFD_BCAST_MAX && rank_ni > 0, "cfd_bcast_attach: bad rank / extent");
  c->bcast_flags[rank] = flag_block ? (int*)flag_block : c->flags;     // NULL = this context itself
  c->bcast_ni[rank] = rank_ni;
  if (rank + 1 > c->bcast_n) c->bcast_n = rank + 1;
  return CFD_OK;
}
extern "C" int cfd_scalar_bcast(cfd_ctx* c, int owner, int field, int i, int j, void** value_dev) {
  CFD_ENTER(c);
  CFD_REQUIRE(value_dev, "null argument");
  return cfd_launch_scalar_bcast(c, owner, field, i, j, (double**)value_dev);
}
extern "C" int cfd_peer_detach(cfd_ctx* c) {
  CFD_REQUIRE(c, "null context");
  c->bcast_n = 0;
  for (int sd = 0; sd < 2; ++sd) {
    c->peer_attached[sd] = false;
    c->peer_flags[sd] = nullptr;
    for (auto& pf : c->peer_field[sd]) pf = nullptr;
  }
  return CFD_OK;
}
extern "C" int cfd_halo_push(cfd_ctx* c, int nfields, const int* fields, const int* w, const int* e) {
  CFD_ENTER(c);
  CFD_REQUIRE(fields && w && e && nfields >= 1 && nfields <= 4, "cfd_halo_push: 1..4 fields");
  if (!c->peer_attached[0] && !c->peer_attached[1]) return CFD_OK;
  return cfd_launch_halo_push(c, nfields, fields, w, e);
}
// diagnostic: clock64 sums of the consumer phases of the flag-ordered sweep kernels since the last call
// (out[0] = lines, out[1..6] = cycles in: slices, backward, reduction, exchange, stores, end barrier), block 0 only
extern "C" int cfd_debug_flow_timing(cfd_ctx* c, double out[8]) {
  CFD_REQUIRE(c && out, "null argument");
  for (int k = 0; k < 8; ++k) out[k] = 0.0;
  if (!c->dbg) return CFD_OK;
  CFD_CUDA(cudaSetDevice(c->device));
  unsigned long long h[8];
  CFD_CUDA(cudaMemcpy(h, c->dbg, sizeof(h), cudaMemcpyDeviceToHost));
  CFD_CUDA(cudaMemset(c->dbg, 0, sizeof(h)));
  for (int k = 0; k < 8; ++k) out[k] = (double)h[k];
  return CFD_OK;
}
extern "C" int cfd_sweeps_single_launch(cfd_ctx* c, int* yes) {
  CFD_REQUIRE(c && yes, "null argument");
  *yes = cfd_flow_eligible(c) ? 1 : 0;
  return CFD_OK;
}
extern "C" int cfd_peer_status(cfd_ctx* c, int* aborted) {
  CFD_REQUIRE(c && aborted, "null argument");
  CFD_CUDA(cudaSetDevice(c->device));
  CFD_CUDA(cudaMemcpyAsync(aborted, c->flags + c->L.ni + CFD_FLAG_ABORT, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_lisolv(cfd_ctx* c, int istart, int jstart, int field) {
  CFD_REQUIRE(c, "null context");
  int rc = check_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(istart >= 2 && jstart >= 2 && jstart <= c->L.nj - 1, "istart/jstart out of range");
  CFD_CUDA(cudaSetDevice(c->device));
  if ((rc = full_sweep(c, istart, jstart, field, 0))) return rc;
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

static int read_res(cfd_ctx* c, int eq, double* out) {
  if (!out) return CFD_OK;
  CFD_CUDA(cudaMemcpyAsync(out, c->res_dev + eq, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

static int stage_u(cfd_ctx* c) {
  int rc;
  if ((rc = asm_u(c))) return rc;
  return sweeps(c, 3, 2, CFD_U, c->cfg.nswp[CFD_EQ_U], 0);
}
static int stage_v(cfd_ctx* c) {
  int rc;
  if ((rc = asm_v(c))) return rc;
  return sweeps(c, 2, 3, CFD_V, c->cfg.nswp[CFD_EQ_V], 0);
}
static int stage_p(cfd_ctx* c) {
  int rc;
  if ((rc = asm_p(c))) return rc;
  if (c->cfg.nswp[CFD_EQ_P] == 0) {   // pp(:,:) = 0 with no sweep to fold it into
    CFD_CUDA(cudaMemsetAsync(c->field[CFD_PP], 0, (size_t)c->L.ni * c->L.LD * sizeof(double), c->stream));
  }
  if ((rc = sweeps(c, 2, 2, CFD_PP, c->cfg.nswp[CFD_EQ_P], 1))) return rc;
  return cfd_launch_correct(c, nullptr);
}

static int stage_scalar_teach(cfd_ctx* c, int eq, int field, int nsw) {
  int rc;
  if ((rc = cfd_launch_teach_scalar(c, eq))) return rc;
  return sweeps(c, 2, 2, field, nsw, 0);
}

static int stage_t(cfd_ctx* c) {
  int rc;
  if (is_teach(c)) return stage_scalar_teach(c, CFD_EQ_T, CFD_T, c->turb.nswpt);
  if ((rc = cfd_launch_general_assemble(c, 3))) return rc;
  return sweeps(c, 2, 2, CFD_T, c->cfg.nswp[CFD_EQ_T], 0);
}
extern "C" int cfd_calct(cfd_ctx* c, double* resort) {
  CFD_REQUIRE(c, "null context");
  CFD_REQUIRE(is_general(c) || (is_teach(c) && c->turbulent),
              "cfd_calct needs a CFD_FAMILY_GENERAL context, or a TEACH context after cfd_set_turbulence");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc = stage_t(c);
  return rc ? rc : read_res(c, CFD_EQ_T, resort);
}

// ---- TEACH family: k-epsilon / energy (teach_turb_kernels.cu) ---------------------------------------------------
extern "C" int cfd_set_turbulence(cfd_ctx* c, const cfd_turbulence* m) {
  CFD_REQUIRE(c && m, "null argument");
  CFD_REQUIRE(is_teach(c), "cfd_set_turbulence needs a CFD_FAMILY_TEACH context");
  CFD_REQUIRE(m->urfk > 0 && m->urfe > 0 && m->urft > 0 && m->nswpk >= 0 && m->nswpd >= 0 && m->nswpt >= 0, "bad model values");
  CFD_CUDA(cudaSetDevice(c->device));
  const size_t fbytes = (size_t)c->L.ni * c->L.LD * sizeof(double);
  const int extra[] = {CFD_T, CFD_TE, CFD_ED, CFD_GEN, CFD_GAMH};
  for (int f : extra) {
    if (!c->field[f]) {
      if (cudaMalloc(&c->field[f], fbytes) != cudaSuccess) {
        cfd_set_error("cudaMalloc of %zu bytes for field %d failed", fbytes, f);
        return CFD_ERR_NOMEM;
      }
      CFD_CUDA(cudaMemsetAsync(c->field[f], 0, fbytes, c->stream));
    }
  }
  const size_t n1 = (size_t)(c->L.ni > c->L.nj ? c->L.ni : c->L.nj) + 2;
  for (auto& w : c->wall) {
    if (!w) {
      CFD_CUDA(cudaMalloc(&w, n1 * sizeof(double)));
      CFD_CUDA(cudaMemsetAsync(w, 0, n1 * sizeof(double), c->stream));
    }
  }
  if (!c->tmult_n) {
    CFD_CUDA(cudaMalloc(&c->tmult_n, n1 * sizeof(double)));
    CFD_CUDA(cudaMalloc(&c->tmult_w, n1 * sizeof(double)));
    CFD_CUDA(cudaMemsetAsync(c->tmult_n, 0, n1 * sizeof(double), c->stream));
    CFD_CUDA(cudaMemsetAsync(c->tmult_w, 0, n1 * sizeof(double), c->stream));
  }
  c->turb = *m;
  c->turbulent = true;
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

static int wall_args(cfd_ctx* c, int which, const void* host, int n) {
  CFD_REQUIRE(c && host, "null argument");
  CFD_REQUIRE(c->turbulent, "wall arrays exist after cfd_set_turbulence");
  CFD_REQUIRE(which >= 0 && which < 4, "bad wall array id");
  const int want = (which == CFD_WALL_TAUN || which == CFD_WALL_YPLUSN) ? c->L.ni : c->L.nj;
  CFD_REQUIRE(n == want, "wall array length must be ni (taun, yplusn) or nj (tauw, xplusw)");
  return CFD_OK;
}
extern "C" int cfd_wall_upload(cfd_ctx* c, int which, const double* host, int n) {
  int rc = wall_args(c, which, host, n);
  if (rc) return rc;
  CFD_CUDA(cudaSetDevice(c->device));
  CFD_CUDA(cudaMemcpyAsync(c->wall[which], host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}
extern "C" int cfd_wall_download(cfd_ctx* c, int which, double* host, int n) {
  int rc = wall_args(c, which, host, n);
  if (rc) return rc;
  CFD_CUDA(cudaSetDevice(c->device));
  CFD_CUDA(cudaMemcpyAsync(host, c->wall[which], (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}
extern "C" int cfd_assemble_te(cfd_ctx* c) {
  CFD_REQUIRE(c && is_teach(c) && c->turbulent, "needs a TEACH context after cfd_set_turbulence");
  CFD_CUDA(cudaSetDevice(c->device));
  return cfd_launch_teach_scalar(c, CFD_EQ_TE);
}
extern "C" int cfd_assemble_ed(cfd_ctx* c) {
  CFD_REQUIRE(c && is_teach(c) && c->turbulent, "needs a TEACH context after cfd_set_turbulence");
  CFD_CUDA(cudaSetDevice(c->device));
  return cfd_launch_teach_scalar(c, CFD_EQ_ED);
}
extern "C" int cfd_calcte(cfd_ctx* c, double* resork) {
  CFD_REQUIRE(c && is_teach(c) && c->turbulent, "needs a TEACH context after cfd_set_turbulence");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc = stage_scalar_teach(c, CFD_EQ_TE, CFD_TE, c->turb.nswpk);
  return rc ? rc : read_res(c, CFD_EQ_TE, resork);
}
extern "C" int cfd_calced(cfd_ctx* c, double* resore) {
  CFD_REQUIRE(c && is_teach(c) && c->turbulent, "needs a TEACH context after cfd_set_turbulence");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc = stage_scalar_teach(c, CFD_EQ_ED, CFD_ED, c->turb.nswpd);
  return rc ? rc : read_res(c, CFD_EQ_ED, resore);
}
extern "C" int cfd_props(cfd_ctx* c) {
  CFD_REQUIRE(c && is_teach(c) && c->turbulent, "needs a TEACH context after cfd_set_turbulence");
  CFD_CUDA(cudaSetDevice(c->device));
  return cfd_launch_teach_props(c);
}

extern "C" int cfd_calcu(cfd_ctx* c, double* resoru) {
  CFD_REQUIRE(c, "null context");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc = stage_u(c);
  return rc ? rc : read_res(c, CFD_EQ_U, resoru);
}
extern "C" int cfd_calcv(cfd_ctx* c, double* resorv) {
  CFD_REQUIRE(c, "null context");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc = stage_v(c);
  return rc ? rc : read_res(c, CFD_EQ_V, resorv);
}
extern "C" int cfd_calcp(cfd_ctx* c, double* resorm) {
  CFD_REQUIRE(c, "null context");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc = stage_p(c);
  return rc ? rc : read_res(c, CFD_EQ_P, resorm);
}

extern "C" int cfd_iterate(cfd_ctx* c, int n, double res[8]) {
  CFD_REQUIRE(c && n >= 0, "bad argument");
  CFD_CUDA(cudaSetDevice(c->device));
  int rc;
  const bool flow = !is_general(c) || c->cfg.solve_flow, temp = is_general(c) && c->cfg.solve_t;
  for (int it = 0; it < n; ++it) {
    if (flow) {
      if ((rc = stage_u(c))) return rc;
      if ((rc = stage_v(c))) return rc;
      if ((rc = stage_p(c))) return rc;
    }
    if (temp && (rc = stage_t(c))) return rc;
    if (is_teach(c) && c->turbulent) {   // control.f90:213-217
      if (c->turb.incalk && (rc = stage_scalar_teach(c, CFD_EQ_TE, CFD_TE, c->turb.nswpk))) return rc;
      if (c->turb.incald && (rc = stage_scalar_teach(c, CFD_EQ_ED, CFD_ED, c->turb.nswpd))) return rc;
      if (c->turb.incalt && (rc = stage_scalar_teach(c, CFD_EQ_T, CFD_T, c->turb.nswpt))) return rc;
      if (c->turb.inpro && (rc = cfd_launch_teach_props(c))) return rc;
    }
  }
  if (!res) return CFD_OK;
  if ((rc = cfd_residuals(c, res))) return rc;
  // failure detection: the residual sums are |.|-sums over every cell of their equation
  const bool turb = is_teach(c) && c->turbulent;
  const bool used[6] = {flow, flow, flow, temp || (turb && c->turb.incalt), turb && c->turb.incalk, turb && c->turb.incald};
  for (int k = 0; k < 6; ++k) {
    if (n > 0 && used[k] && !std::isfinite(res[k])) {
      cfd_set_error("cfd_iterate: residual of equation %d is not finite after %d passes", k, n);
      return CFD_ERR_DIVERGED;
    }
  }
  return CFD_OK;
}

extern "C" int cfd_residuals(cfd_ctx* c, double res[8]) {
  CFD_REQUIRE(c && res, "null argument");
  CFD_CUDA(cudaSetDevice(c->device));
  CFD_CUDA(cudaMemcpyAsync(res, c->res_dev, 8 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_residual_device_ptr(cfd_ctx* c, void** ptr) {
  CFD_REQUIRE(c && ptr, "null argument");
  *ptr = c->res_dev;
  return CFD_OK;
}
