// post_kernels.cu -- post-processing on the device fields, so that result files and monitors of large runs do not need
// whole-field downloads (an 8192^2 field is 537 MB; a centre-line profile is 64 KB).
//
//   cfd_extract_line / cfd_extract_row   A(i, 1:nj) / A(1:ni, j) of a field in natural order (the device layout is
//                                        chunk-interleaved)
//   cfd_centreline                       uc(imon, 1:nj) and vc(1:ni, jmon) of centred_vel (cavity_fvm.f90:416-430; the same
//                                        formulas in postproc, buoyancy_cav.f90:441-449) -- what uvel.txt / vvel.txt hold
//                                        (cavity_fvm.f90:600-607)
//   cfd_reference_pressure               p = p - p(ipref,jpref) and the four corner values (cavity_fvm.f90:568-573), in
//                                        place on the device
// Arithmetic is the reference's (one addition and one multiplication per value; no contraction possible).
#include "common.cuh"

namespace cfd {

// mode 0: out[j0] = A(line, j0); mode 1: out[i0] = A(i0, row)
__global__ void k_extract(const Layout L, const double* __restrict__ a, int mode, int index, double* __restrict__ out, int n) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  out[k] = mode == 0 ? a[(size_t)index * L.LD + lpos(k, L.M, L.T)] : a[(size_t)k * L.LD + lpos(index, L.M, L.T)];
}

// which 0: uc(index, :) = u(2,:) | 0.5*(u(i,:)+u(i+1,:)) | u(ni,:)      (rows i = 1 | 2..ni-1 | ni; index is 0-based here)
// which 1: vc(:, index) = v(:,2) | 0.5*(v(:,j)+v(:,j+1)) | v(:,nj)
__global__ void k_centreline(const Layout L, const double* __restrict__ f, int which, int index, double* __restrict__ out, int n) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  auto at = [&](int i0, int j0) { return f[(size_t)i0 * L.LD + lpos(j0, L.M, L.T)]; };
  double r;
  if (which == 0) {
    const int i0 = index, ni = L.ni;
    r = i0 == 0 ? at(1, k) : (i0 == ni - 1 ? at(ni - 1, k) : __dmul_rn(0.5, __dadd_rn(at(i0, k), at(i0 + 1, k))));
  } else {
    const int j0 = index, nj = L.nj;
    r = j0 == 0 ? at(k, 1) : (j0 == nj - 1 ? at(k, nj - 1) : __dmul_rn(0.5, __dadd_rn(at(k, j0), at(k, j0 + 1))));
  }
  out[k] = r;
}

__global__ void k_shift_pressure(const Layout L, double* __restrict__ p, const double* __restrict__ pref) {
  const double r = *pref;
  const size_t n = (size_t)L.ni * L.LD;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (size_t)gridDim.x * blockDim.x) p[q] = p[q] - r;
}
__global__ void k_corner_pressure(const Layout L, double* __restrict__ p) {
  auto at = [&](int i0, int j0) -> double& { return p[(size_t)i0 * L.LD + lpos(j0, L.M, L.T)]; };
  const int ni = L.ni, nj = L.nj;
  at(0, 0) = __dsub_rn(__dadd_rn(at(1, 0), at(0, 1)), at(1, 1));
  at(0, nj - 1) = __dsub_rn(__dadd_rn(at(0, nj - 2), at(1, nj - 1)), at(1, nj - 2));
  at(ni - 1, 0) = __dsub_rn(__dadd_rn(at(ni - 2, 0), at(ni - 1, 1)), at(ni - 2, 1));
  at(ni - 1, nj - 1) = __dsub_rn(__dadd_rn(at(ni - 1, nj - 2), at(ni - 2, nj - 1)), at(ni - 2, nj - 2));
}
__global__ void k_copy_scalar(const double* src, double* dst) { *dst = *src; }

}  // namespace cfd

using namespace cfd;

static int plain_field(const cfd_ctx* c, int field) {
  if (field < 0 || field >= CFD_FIELD_COUNT || !c->field[field] || cfd_is_coef_field(field)) {
    cfd_set_error("field %d is not a plain field of this context", field);
    return CFD_ERR_INVALID;
  }
  return CFD_OK;
}

// the small result vector travels through the partial-sum scratch area (max_partials >= (T/32)*(ni/4) doubles is not
// always >= max(ni, nj)): a dedicated buffer sized on first use
static int post_buffer(cfd_ctx* c, size_t n, double** buf) {
  if (c->post_n < n) {
    if (c->post_buf) CFD_CUDA(cudaFree(c->post_buf));
    c->post_buf = nullptr;
    c->post_n = 0;
    if (cudaMalloc(&c->post_buf, n * sizeof(double)) != cudaSuccess) {
      cfd_set_error("cudaMalloc of %zu bytes failed", n * sizeof(double));
      return CFD_ERR_NOMEM;
    }
    c->post_n = n;
  }
  *buf = c->post_buf;
  return CFD_OK;
}

static int fetch(cfd_ctx* c, double* dev, double* host, int n) {
  CFD_CUDA(cudaGetLastError());
  CFD_CUDA(cudaMemcpyAsync(host, dev, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  return CFD_OK;
}

extern "C" int cfd_extract_line(cfd_ctx* c, int field, int i, double* host, int n) {
  CFD_REQUIRE(c && host, "null argument");
  int rc = plain_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(i >= 1 && i <= c->L.ni && n == c->L.nj, "cfd_extract_line: 1 <= i <= ni, n == nj");
  CFD_CUDA(cudaSetDevice(c->device));
  double* buf;
  if ((rc = post_buffer(c, (size_t)(c->L.ni > c->L.nj ? c->L.ni : c->L.nj), &buf))) return rc;
  k_extract<<<(n + 255) / 256, 256, 0, c->stream>>>(c->L, c->field[field], 0, i - 1, buf, n);
  ++c->launches;
  return fetch(c, buf, host, n);
}

extern "C" int cfd_extract_row(cfd_ctx* c, int field, int j, double* host, int n) {
  CFD_REQUIRE(c && host, "null argument");
  int rc = plain_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(j >= 1 && j <= c->L.nj && n == c->L.ni, "cfd_extract_row: 1 <= j <= nj, n == ni");
  CFD_CUDA(cudaSetDevice(c->device));
  double* buf;
  if ((rc = post_buffer(c, (size_t)(c->L.ni > c->L.nj ? c->L.ni : c->L.nj), &buf))) return rc;
  k_extract<<<(n + 255) / 256, 256, 0, c->stream>>>(c->L, c->field[field], 1, j - 1, buf, n);
  ++c->launches;
  return fetch(c, buf, host, n);
}

extern "C" int cfd_centreline(cfd_ctx* c, int which, int index, double* host, int n) {
  CFD_REQUIRE(c && host, "null argument");
  CFD_REQUIRE(which == 0 || which == 1, "which: 0 = uc(index, 1:nj), 1 = vc(1:ni, index)");
  CFD_REQUIRE(c->cfg.goff == 0 && c->cfg.joff == 0 && c->cfg.halo_w + c->cfg.halo_e + c->cfg.halo_s + c->cfg.halo_n == 0,
              "cfd_centreline works on whole-grid contexts");
  if (which == 0) CFD_REQUIRE(index >= 1 && index <= c->L.ni && n == c->L.nj, "uc line: 1 <= index <= ni, n == nj");
  else CFD_REQUIRE(index >= 1 && index <= c->L.nj && n == c->L.ni, "vc row: 1 <= index <= nj, n == ni");
  CFD_CUDA(cudaSetDevice(c->device));
  double* buf;
  int rc = post_buffer(c, (size_t)(c->L.ni > c->L.nj ? c->L.ni : c->L.nj), &buf);
  if (rc) return rc;
  k_centreline<<<(n + 255) / 256, 256, 0, c->stream>>>(c->L, c->field[which == 0 ? CFD_U : CFD_V], which, index - 1, buf, n);
  ++c->launches;
  return fetch(c, buf, host, n);
}

extern "C" int cfd_reference_pressure(cfd_ctx* c) {
  CFD_REQUIRE(c, "null context");
  CFD_REQUIRE(c->cfg.goff == 0 && c->cfg.joff == 0 && c->cfg.halo_w + c->cfg.halo_e + c->cfg.halo_s + c->cfg.halo_n == 0,
              "cfd_reference_pressure works on whole-grid contexts");
  CFD_CUDA(cudaSetDevice(c->device));
  double* buf;
  int rc = post_buffer(c, (size_t)(c->L.ni > c->L.nj ? c->L.ni : c->L.nj), &buf);
  if (rc) return rc;
  double* p = c->field[CFD_P];
  const double* pref = p + (size_t)(c->cfg.ipref - 1) * c->L.LD + lpos(c->cfg.jpref - 1, c->L.M, c->L.T);
  k_copy_scalar<<<1, 1, 0, c->stream>>>(pref, buf);          // the reference value itself is shifted too: take a copy first
  k_shift_pressure<<<c->num_sms * 4, 256, 0, c->stream>>>(c->L, p, buf);
  k_corner_pressure<<<1, 1, 0, c->stream>>>(c->L, p);
  c->launches += 3;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
