// jslab_kernels.cu -- the pieces of the partition (Wang / SPIKE) method that let the S-N line solves of lisolv
// (cavity_fvm.f90:62-80) span several contexts cut along j (second variant):
//
//   piece solution   x = y + w*x_s + z*x_n      y: cfd_sweep in a j-slab context (zero values at the cut ends),
//                                               w, z: cfd_spike_basis (tdma_kernels.cu, matrix only)
//   interface rows   first / last unknown of every piece -> 6 numbers per piece and line (cfd_spike_pack), all-gathered
//   reduced system   per line, 2*world unknowns (lo_r, hi_r), band width 2:
//                       lo_r - W1_r*hi_{r-1} - Z1_r*lo_{r+1} = Y1_r        hi_r - W2_r*hi_{r-1} - Z2_r*lo_{r+1} = Y2_r
//                    solved redundantly by every context, one thread per line (k_spike_reduced)
//   finish           x = y + w*x_s + z*x_n on the piece, x_s / x_n stored into the halo rows (k_spike_apply)
//
// plus the row gather / scatter that turns a halo ROW (strided in the line layout) into a contiguous message.
#include "common.cuh"

namespace cfd {

constexpr int MAXW = 16;   // contexts along j

struct PieceRange {
  Layout L;
  int i_first, i_step, nlines;
  int jlo, jhi;
};

__global__ void k_spike_pack_wz(const PieceRange R, const double* __restrict__ w, const double* __restrict__ z, double* buf) {
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  if (i0 >= R.L.ni) return;
  const long long o = (long long)i0 * R.L.LD;
  const int qa = lpos(R.jlo, R.L.M, R.L.T), qb = lpos(R.jhi, R.L.M, R.L.T);
  buf[4 * i0 + 0] = w[o + qa];
  buf[4 * i0 + 1] = w[o + qb];
  buf[4 * i0 + 2] = z[o + qa];
  buf[4 * i0 + 3] = z[o + qb];
}

__global__ void k_spike_pack_y(const PieceRange R, const double* __restrict__ phi, double* buf) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= R.nlines) return;
  const long long o = (long long)(R.i_first + l * R.i_step) * R.L.LD;
  buf[2 * l + 0] = phi[o + lpos(R.jlo, R.L.M, R.L.T)];
  buf[2 * l + 1] = phi[o + lpos(R.jhi, R.L.M, R.L.T)];
}

// one thread per line of the colour: dense elimination of the 2*world reduced system (diagonally dominant: no pivoting)
__global__ void k_spike_reduced(const PieceRange R, const double* __restrict__ wz_all, const double* __restrict__ y_all,
                                int rank, int world, double* xsn) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= R.nlines) return;
  const int i0 = R.i_first + l * R.i_step, n = 2 * world, ni = R.L.ni;
  double A[2 * MAXW][2 * MAXW + 1];
  for (int a = 0; a < n; ++a)
    for (int b = 0; b <= n; ++b) A[a][b] = 0.0;
  for (int r = 0; r < world; ++r) {
    const double* wz = wz_all + ((size_t)r * ni + i0) * 4;
    const double* y = y_all + ((size_t)r * R.nlines + l) * 2;
    const int lo = 2 * r, hi = 2 * r + 1;
    A[lo][lo] = 1.0;
    A[hi][hi] = 1.0;
    if (r > 0) {           // x_s of piece r is hi_{r-1}
      A[lo][hi - 2] = -wz[0];
      A[hi][hi - 2] = -wz[1];
    }
    if (r < world - 1) {   // x_n of piece r is lo_{r+1}
      A[lo][lo + 2] = -wz[2];
      A[hi][lo + 2] = -wz[3];
    }
    A[lo][n] = y[0];
    A[hi][n] = y[1];
  }
  for (int k = 0; k < n; ++k) {
    const double inv = 1.0 / A[k][k];
    const int ie = (k + 2 < n - 1) ? k + 2 : n - 1, je = (k + 4 < n - 1) ? k + 4 : n - 1;
    for (int a = k + 1; a <= ie; ++a) {
      const double f = A[a][k] * inv;
      if (f != 0.0) {
        for (int b = k; b <= je; ++b) A[a][b] -= f * A[k][b];
        A[a][n] -= f * A[k][n];
      }
    }
  }
  double x[2 * MAXW];
  for (int k = n - 1; k >= 0; --k) {
    double s = A[k][n];
    const int je = (k + 4 < n - 1) ? k + 4 : n - 1;
    for (int b = k + 1; b <= je; ++b) s -= A[k][b] * x[b];
    x[k] = s / A[k][k];
  }
  xsn[2 * l + 0] = rank > 0 ? x[2 * (rank - 1) + 1] : 0.0;
  xsn[2 * l + 1] = rank < world - 1 ? x[2 * (rank + 1)] : 0.0;
}

__global__ void __launch_bounds__(256) k_spike_apply(const PieceRange R, const double* __restrict__ w, const double* __restrict__ z,
                                                     const double* __restrict__ xsn, int rank, int world, double* phi) {
  const int M = R.L.M, T = R.L.T;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int l = blockIdx.y * blockDim.y + threadIdx.y;
  if (t >= T || l >= R.nlines) return;
  const long long o = (long long)(R.i_first + l * R.i_step) * R.L.LD;
  const double xs = xsn[2 * l], xn = xsn[2 * l + 1];
  for (int k = 0; k < M; ++k) {
    const int j0 = t * M + k, q = k * T + t;
    if (j0 >= R.jlo && j0 <= R.jhi) {
      phi[o + q] = fma(__ldg(w + o + q), xs, fma(__ldg(z + o + q), xn, phi[o + q]));
    } else if (j0 == R.jlo - 1 && rank > 0) {
      phi[o + q] = xs;
    } else if (j0 == R.jhi + 1 && rank < world - 1) {
      phi[o + q] = xn;
    }
  }
}

__global__ void k_rows_pack(Layout L, const double* __restrict__ f, int j0_first, int nrows, double* buf) {
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (i0 >= L.ni || r >= nrows) return;
  buf[(size_t)r * L.ni + i0] = f[(long long)i0 * L.LD + lpos(j0_first + r, L.M, L.T)];
}
__global__ void k_rows_unpack(Layout L, double* f, int j0_first, int nrows, const double* __restrict__ buf) {
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (i0 >= L.ni || r >= nrows) return;
  f[(long long)i0 * L.LD + lpos(j0_first + r, L.M, L.T)] = buf[(size_t)r * L.ni + i0];
}

}  // namespace cfd

using namespace cfd;

// lines of `colour` (or all lines, colour < 0) a sweep starting at istart touches, and the piece's unknown range
static bool piece_range(const cfd_ctx* c, int istart, int jstart, int colour, PieceRange* R) {
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  int glo = c->cfg.goff + c->cfg.halo_w + 1, ghi = c->cfg.goff + c->L.ni - c->cfg.halo_e;
  if (glo < istart) glo = istart;
  if (ghi > NI - 1) ghi = NI - 1;
  R->L = c->L;
  cfd_local_j_range(c, jstart - 1, NJ - 2, &R->jlo, &R->jhi);
  if (colour < 0) {
    R->i_first = glo - 1 - c->cfg.goff;
    R->i_step = 1;
    R->nlines = ghi - glo + 1;
  } else {
    const int g0 = glo + ((((glo - istart) & 1) != colour) ? 1 : 0);
    R->i_first = g0 - 1 - c->cfg.goff;
    R->i_step = 2;
    R->nlines = g0 > ghi ? 0 : (ghi - g0) / 2 + 1;
  }
  return R->nlines > 0 && R->jhi >= R->jlo;
}

static int check_plain_field(const cfd_ctx* c, int field) {
  if (field < 0 || field >= CFD_FIELD_COUNT || !c->field[field] || cfd_is_coef_field(field)) {
    cfd_set_error("field %d is not a plain field of this context", field);
    return CFD_ERR_INVALID;
  }
  return CFD_OK;
}

extern "C" int cfd_spike_basis(cfd_ctx* c, int istart, int jstart) {
  CFD_REQUIRE(c, "null context");
  CFD_CUDA(cudaSetDevice(c->device));
  return cfd_launch_spike_basis(c, istart, jstart);
}

extern "C" int cfd_spike_pack(cfd_ctx* c, int istart, int jstart, int field, int colour, int which, double* buf) {
  CFD_REQUIRE(c && buf, "null argument");
  CFD_CUDA(cudaSetDevice(c->device));
  PieceRange R;
  if (which == 1) {
    CFD_REQUIRE(c->spike_w, "cfd_spike_pack: call cfd_spike_basis first");
    if (!piece_range(c, istart, jstart, -1, &R)) return CFD_OK;
    k_spike_pack_wz<<<(c->L.ni + 127) / 128, 128, 0, c->stream>>>(R, c->spike_w, c->spike_z, buf);
  } else {
    int rc = check_plain_field(c, field);
    if (rc) return rc;
    if (!piece_range(c, istart, jstart, colour, &R)) return CFD_OK;
    k_spike_pack_y<<<(R.nlines + 127) / 128, 128, 0, c->stream>>>(R, c->field[field], buf);
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

extern "C" int cfd_spike_finish(cfd_ctx* c, int istart, int jstart, int field, int colour, const double* wz_all,
                                const double* y_all, int rank, int world) {
  CFD_REQUIRE(c && wz_all && y_all, "null argument");
  CFD_REQUIRE(world >= 1 && world <= MAXW && rank >= 0 && rank < world, "bad rank / world");
  CFD_REQUIRE(c->spike_w, "cfd_spike_finish: call cfd_spike_basis first");
  int rc = check_plain_field(c, field);
  if (rc) return rc;
  CFD_CUDA(cudaSetDevice(c->device));
  PieceRange R;
  if (!piece_range(c, istart, jstart, colour, &R)) return CFD_OK;
  if (!c->spike_small) CFD_CUDA(cudaMalloc(&c->spike_small, (size_t)2 * c->L.ni * sizeof(double)));
  k_spike_reduced<<<(R.nlines + 63) / 64, 64, 0, c->stream>>>(R, wz_all, y_all, rank, world, c->spike_small);
  dim3 block(32, 8), grid((c->L.T + 31) / 32, (R.nlines + 7) / 8);
  k_spike_apply<<<grid, block, 0, c->stream>>>(R, c->spike_w, c->spike_z, c->spike_small, rank, world, c->field[field]);
  c->launches += 2;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

extern "C" int cfd_rows_pack(cfd_ctx* c, int field, int j_first, int nrows, double* buf) {
  CFD_REQUIRE(c && buf, "null argument");
  int rc = check_plain_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(j_first >= 1 && nrows >= 1 && j_first + nrows - 1 <= c->L.nj, "row range outside the local array");
  CFD_CUDA(cudaSetDevice(c->device));
  dim3 grid((c->L.ni + 127) / 128, nrows);
  k_rows_pack<<<grid, 128, 0, c->stream>>>(c->L, c->field[field], j_first - 1, nrows, buf);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

extern "C" int cfd_rows_unpack(cfd_ctx* c, int field, int j_first, int nrows, const double* buf) {
  CFD_REQUIRE(c && buf, "null argument");
  int rc = check_plain_field(c, field);
  if (rc) return rc;
  CFD_REQUIRE(j_first >= 1 && nrows >= 1 && j_first + nrows - 1 <= c->L.nj, "row range outside the local array");
  CFD_CUDA(cudaSetDevice(c->device));
  dim3 grid((c->L.ni + 127) / 128, nrows);
  k_rows_unpack<<<grid, 128, 0, c->stream>>>(c->L, c->field[field], j_first - 1, nrows, buf);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
