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
  double* thomas;      // implicit steps: e_x[nx], temp_x[nx], e_y[ny], temp_y[ny] of the forward sweep (constant)
  double* op[5];       // variable-conductivity operator: aw, ae, as, an (fdm2d_steady_heat_cond.f90:176-179) and QDOT at the nodes
  double* gs_partial;  // Gauss-Seidel: per-block partial sums of (x1-x2)**2, then the total
  int gs_blocks;
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

// ---------------------------------------------------------------------------------------------------
// Implicit Euler, locally one-dimensional: the tridiagonal assembly of implicit-time/fdm1d_explicit_heat_cond.f90
// :122-143 (a = c = -Fo, b = 1 + 2 Fo, identity rows at the Dirichlet ends, rhs = T + dt*QDOT/rho/cp) applied along
// x for every interior row and then along y for every interior column, solved with the reference's `thomas`
// (:147-184: e = c/temp, f = (d - a f_prev)/temp, x = f - e x_next).  The coefficients do not depend on the line,
// so e and temp are computed once on the host (same operation order) and only the f / x recurrences run here,
// one thread per line, in place, with IEEE operations in the reference's order (bit-comparable with
// oracle/heat_oracle.c:orc_heat2d_lod).
// ---------------------------------------------------------------------------------------------------
struct LodParams {
  int nx, ny;
  int wall_s, wall_n;        // the local row 0 / ny-1 is a physical (Dirichlet) boundary row; 0 in a row slab whose first /
                             // last rows are halo rows (there the recurrences just start, see k_lod_y_chunk)
  double fox, foy, dt, rho, cp, xa, dx, qthr, qval, qdt;
  double tleft, tright, tbottom, ttop;
  const double *ex, *tx, *ey, *ty;
  double* t;
};

// Dirichlet edges rewritten at the start of every step (same order as the explicit scheme)
__global__ void k_lod_bc(const LodParams P) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nx = (size_t)P.nx;
  if (n >= (P.wall_s ? 1 : 0) && n <= P.ny - (P.wall_n ? 2 : 1)) {
    P.t[(size_t)n * nx] = P.tleft;
    P.t[(size_t)n * nx + nx - 1] = P.tright;
  }
  if (n < P.nx) {
    if (P.wall_s) P.t[n] = P.tbottom;
    if (P.wall_n) P.t[(size_t)(P.ny - 1) * nx + n] = P.ttop;
  }
}

// x-lines: a warp owns 32 interior rows and walks along x in 32-column tiles, transposing through shared memory
// so that global accesses are 256-byte row segments while each lane runs the recurrence of "its" row.
template <bool EXACT>
__global__ void __launch_bounds__(128) k_lod_x(const LodParams P) {
  __shared__ double tile[4][32][33];
  const int lane = threadIdx.x, w = threadIdx.y;
  const int row0 = 1 + (blockIdx.x * 4 + w) * 32;          // first of this warp's rows (0-based j)
  if (row0 > P.ny - 2) return;
  const int nx = P.nx, ntile = (nx + 31) / 32;
  const size_t snx = (size_t)nx;
  const int myrow = row0 + lane;
  const bool rowok = myrow <= P.ny - 2;
  double (*tl)[33] = tile[w];
  // Each lane keeps the NEXT tile's 32 row segments in registers while "its" row of the current tile is solved,
  // so the global-load latency overlaps the serial recurrence.
  double nxt[32];
  auto fetch = [&](int c0) {
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      const int j = row0 + r;
      nxt[r] = (j <= P.ny - 2 && c0 + lane < nx && c0 >= 0) ? P.t[(size_t)j * snx + c0 + lane] : 0.0;
    }
  };
  // forward sweep: f overwrites T
  double fprev = 0.0;
  fetch(0);
  for (int tI = 0; tI < ntile; ++tI) {
    const int c0 = tI * 32;
#pragma unroll
    for (int r = 0; r < 32; ++r) tl[r][lane] = nxt[r];
    __syncwarp();
    if (tI + 1 < ntile) fetch(c0 + 32);
    if (rowok) {
      for (int k = 0; k < 32 && c0 + k < nx; ++k) {
        const int i = c0 + k;
        double d = tl[lane][k], f;
        if (i == 0) {
          f = d;
        } else if (i == nx - 1) {
          f = d;                      // a = 0, temp = 1
        } else if (EXACT) {
          const double xi = __dadd_rn(P.xa, __dmul_rn((double)i, P.dx));
          const double q = (xi < P.qthr) ? P.qval : 0.0;
          d = __dadd_rn(d, __ddiv_rn(__ddiv_rn(__dmul_rn(P.dt, q), P.rho), P.cp));
          f = __ddiv_rn(__dsub_rn(d, __dmul_rn(-P.fox, fprev)), __ldg(P.tx + i));
        } else {                      // production: reciprocal pivots, FMAs (agrees to a few ulp)
          const double xi = __dadd_rn(P.xa, __dmul_rn((double)i, P.dx));
          d += (xi < P.qthr) ? P.qdt : 0.0;
          f = fma(P.fox, fprev, d) * __ldg(P.tx + i);
        }
        tl[lane][k] = f;
        fprev = f;
      }
    }
    __syncwarp();
    for (int r = 0; r < 32; ++r) {
      const int j = row0 + r;
      if (j <= P.ny - 2 && c0 + lane < nx) P.t[(size_t)j * snx + c0 + lane] = tl[r][lane];
    }
    __syncwarp();
  }
  // back substitution: x overwrites f, tiles right to left
  double xnext = 0.0;
  __threadfence_block();
  fetch((ntile - 1) * 32);
  for (int tI = ntile - 1; tI >= 0; --tI) {
    const int c0 = tI * 32;
#pragma unroll
    for (int r = 0; r < 32; ++r) tl[r][lane] = nxt[r];
    __syncwarp();
    if (tI > 0) fetch(c0 - 32);
    if (rowok) {
      for (int k = 31; k >= 0; --k) {
        const int i = c0 + k;
        if (i >= nx) continue;
        const double f = tl[lane][k];
        const double x = (i == nx - 1) ? f : (EXACT ? __dsub_rn(f, __dmul_rn(__ldg(P.ex + i), xnext)) : fma(-__ldg(P.ex + i), xnext, f));
        tl[lane][k] = x;
        xnext = x;
      }
    }
    __syncwarp();
    for (int r = 0; r < 32; ++r) {
      const int j = row0 + r;
      if (j <= P.ny - 2 && c0 + lane < nx) P.t[(size_t)j * snx + c0 + lane] = tl[r][lane];
    }
    __syncwarp();
  }
}

// y-lines: one thread per interior column, consecutive threads = consecutive x (coalesced at every j).  The
// recurrence is serial in j but the loads are not: they are issued LOD_U rows ahead of the chain.
#define LOD_U 16
template <bool EXACT>
__global__ void __launch_bounds__(128) k_lod_y(const LodParams P) {
  const int i = 1 + blockIdx.x * blockDim.x + threadIdx.x;
  if (i > P.nx - 2) return;
  const size_t nx = (size_t)P.nx;
  const int ny = P.ny;
  double* col = P.t + i;
  double fprev = 0.0;
  for (int j0 = 0; j0 < ny; j0 += LOD_U) {
    double d[LOD_U], tt[LOD_U];
#pragma unroll
    for (int r = 0; r < LOD_U; ++r) {
      const int j = j0 + r;
      d[r] = (j < ny) ? col[(size_t)j * nx] : 0.0;
      tt[r] = (j < ny) ? __ldg(P.ty + j) : 1.0;
    }
#pragma unroll
    for (int r = 0; r < LOD_U; ++r) {
      const int j = j0 + r;
      if (j < ny) {
        double f;
        if (j == 0 || j == ny - 1) {
          f = d[r];                   // identity rows (a = 0, temp = 1)
        } else if (EXACT) {
          f = __ddiv_rn(__dsub_rn(d[r], __dmul_rn(-P.foy, fprev)), tt[r]);
        } else {
          f = fma(P.foy, fprev, d[r]) * tt[r];
        }
        col[(size_t)j * nx] = f;
        fprev = f;
      }
    }
  }
  double xnext = 0.0;
  for (int j0 = ny - 1; j0 >= 0; j0 -= LOD_U) {
    double f[LOD_U], ee[LOD_U];
#pragma unroll
    for (int r = 0; r < LOD_U; ++r) {
      const int j = j0 - r;
      f[r] = (j >= 0) ? col[(size_t)j * nx] : 0.0;
      ee[r] = (j >= 0) ? __ldg(P.ey + j) : 0.0;
    }
#pragma unroll
    for (int r = 0; r < LOD_U; ++r) {
      const int j = j0 - r;
      if (j >= 0) {
        const double x = (j == ny - 1) ? f[r] : (EXACT ? __dsub_rn(f[r], __dmul_rn(ee[r], xnext)) : fma(-ee[r], xnext, f[r]));
        col[(size_t)j * nx] = x;
        xnext = x;
      }
    }
  }
}


// ---------------------------------------------------------------------------------------------------
// Production LOD sweeps with the lines cut into chunks.
//
// One thread per line leaves only nx (or ny) threads on the whole GPU -- a third of the loads needed in flight to
// keep HBM busy.  Both recurrences of `thomas` are contractions: f_i = (d_i + Fo f_{i-1})/temp_i and
// x_i = f_i - e_i x_{i+1} forget their starting value at the rate m = lim e_i = ((1+2Fo) - sqrt(1+4Fo))/(2Fo) < 1 per
// node.  A chunk therefore starts W nodes early from zero, W chosen so that m^W < 2^-70: the neglected coupling is far
// below the last bit of the result, and `nchunk` threads can work on every line.  The passes go out of place
// (forward src -> dst, backward dst -> src), because a chunk's warm-up reads nodes another chunk is producing.
// The exact mode keeps the one-thread-per-line kernels above.
// ---------------------------------------------------------------------------------------------------
struct LodChunk {
  int ch;        // nodes per chunk (multiple of 32)
  int warm;      // warm-up nodes (multiple of 32)
  const double* src;
  double* dst;
};

// y-lines.  DIR 0: forward elimination bottom -> top, DIR 1: back substitution top -> bottom.
template <int DIR>
__global__ void __launch_bounds__(128, 8) k_lod_y_chunk(const LodParams P, const LodChunk C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.nx) return;
  const size_t nx = (size_t)P.nx;
  const int ny = P.ny;
  const int r0 = blockIdx.y * C.ch, r1 = min(ny, r0 + C.ch);       // own rows [r0, r1)
  if (r0 >= ny) return;
  const double* s = C.src + i;
  double* d = C.dst + i;
  const bool ident = (i == 0 || i == P.nx - 1);                     // Dirichlet columns: carried along unchanged
  constexpr int U = 16;
  if (DIR == 0) {
    const int rs = max(0, r0 - C.warm);
    double fprev = 0.0;
    for (int j0 = rs; j0 < r1; j0 += U) {
      double v[U];
#pragma unroll
      for (int r = 0; r < U; ++r) v[r] = (j0 + r < r1) ? s[(size_t)(j0 + r) * nx] : 0.0;
#pragma unroll
      for (int r = 0; r < U; ++r) {
        const int j = j0 + r;
        if (j < r1) {
          const double f = (ident || (j == 0 && P.wall_s) || (j == ny - 1 && P.wall_n)) ? v[r] : fma(P.foy, fprev, v[r]) * __ldg(P.ty + j);
          if (j >= r0) d[(size_t)j * nx] = f;
          fprev = f;
        }
      }
    }
  } else {
    const int re = min(ny - 1, r1 - 1 + C.warm);                    // start above the chunk
    double xnext = 0.0;
    for (int j0 = re; j0 >= r0; j0 -= U) {
      double v[U];
#pragma unroll
      for (int r = 0; r < U; ++r) v[r] = (j0 - r >= r0) ? s[(size_t)(j0 - r) * nx] : 0.0;
#pragma unroll
      for (int r = 0; r < U; ++r) {
        const int j = j0 - r;
        if (j >= r0) {
          const double x = (ident || (j == ny - 1 && P.wall_n)) ? v[r] : fma(-__ldg(P.ey + j), xnext, v[r]);
          if (j < r1) d[(size_t)j * nx] = x;
          xnext = x;
        }
      }
    }
  }
}

// x-lines: a warp owns 32 rows and one chunk of columns, walking 32-column tiles transposed through shared memory.
template <int DIR>
__global__ void __launch_bounds__(128) k_lod_x_chunk(const LodParams P, const LodChunk C) {
  __shared__ double tile[4][32][33];
  const int lane = threadIdx.x, w = threadIdx.y;
  const int row0 = (blockIdx.x * 4 + w) * 32;                       // rows 0 .. ny-1 (boundary rows are copied)
  if (row0 >= P.ny) return;
  const int nx = P.nx;
  const size_t snx = (size_t)nx;
  const int c_lo = blockIdx.y * C.ch, c_hi = min(nx, c_lo + C.ch);  // own columns [c_lo, c_hi)
  if (c_lo >= nx) return;
  const int myrow = row0 + lane;
  const bool rowin = myrow < P.ny;
  const bool solve = rowin && myrow >= (P.wall_s ? 1 : 0) && myrow <= P.ny - (P.wall_n ? 2 : 1);   // boundary row: copy
  double (*tl)[33] = tile[w];
  double nxt[32];
  auto fetch = [&](int c0) {
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      const int j = row0 + r;
      nxt[r] = (j < P.ny && c0 + lane < nx) ? C.src[(size_t)j * snx + c0 + lane] : 0.0;
    }
  };
  auto store = [&](int c0) {
    for (int r = 0; r < 32; ++r) {
      const int j = row0 + r;
      if (j < P.ny && c0 + lane < c_hi) C.dst[(size_t)j * snx + c0 + lane] = tl[r][lane];
    }
  };
  if (DIR == 0) {
    const int cs = max(0, c_lo - C.warm);
    double fprev = 0.0;
    fetch(cs);
    for (int c0 = cs; c0 < c_hi; c0 += 32) {
#pragma unroll
      for (int r = 0; r < 32; ++r) tl[r][lane] = nxt[r];
      __syncwarp();
      if (c0 + 32 < c_hi) fetch(c0 + 32);
      if (solve) {
        for (int k = 0; k < 32 && c0 + k < c_hi; ++k) {
          const int i = c0 + k;
          double dd = tl[lane][k], f;
          if (i == 0 || i == nx - 1) {
            f = dd;
          } else {
            const double xi = __dadd_rn(P.xa, __dmul_rn((double)i, P.dx));
            dd += (xi < P.qthr) ? P.qdt : 0.0;
            f = fma(P.fox, fprev, dd) * __ldg(P.tx + i);
          }
          tl[lane][k] = f;
          fprev = f;
        }
      }
      __syncwarp();
      if (c0 >= c_lo) store(c0);
      __syncwarp();
    }
  } else {
    const int ce = min(((nx + 31) / 32) * 32, c_hi + C.warm);       // tile-aligned start to the right of the chunk
    double xnext = 0.0;
    fetch(ce - 32);
    for (int c0 = ce - 32; c0 >= c_lo; c0 -= 32) {
#pragma unroll
      for (int r = 0; r < 32; ++r) tl[r][lane] = nxt[r];
      __syncwarp();
      if (c0 - 32 >= c_lo) fetch(c0 - 32);
      if (solve) {
        for (int k = 31; k >= 0; --k) {
          const int i = c0 + k;
          if (i >= nx) continue;
          const double f = tl[lane][k];
          const double x = (i == nx - 1) ? f : fma(-__ldg(P.ex + i), xnext, f);
          tl[lane][k] = x;
          xnext = x;
        }
      }
      __syncwarp();
      if (c0 < c_hi) store(c0);
      __syncwarp();
    }
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
  h->thomas = nullptr;
  for (auto& o : h->op) o = nullptr;
  h->gs_partial = nullptr;
  h->gs_blocks = 0;
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
  if (h->thomas) cudaFree(h->thomas);
  for (auto& o : h->op) if (o) cudaFree(o);
  if (h->gs_partial) cudaFree(h->gs_partial);
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

// forward-sweep constants of `thomas` for a line of n nodes with a = c = -fo, b = 1 + 2 fo and identity ends
static void thomas_constants(int n, double fo, std::vector<double>& e, std::vector<double>& temp) {
  e.assign(n, 0.0);
  temp.assign(n, 1.0);
  e[0] = 0.0 / 1.0;
  for (int i = 1; i < n; ++i) {
    const bool last = (i == n - 1);
    const double a = last ? 0.0 : -fo, b = last ? 1.0 : 1.0 + 2.0 * fo, c = last ? 0.0 : -fo;
    temp[i] = b - a * e[i - 1];
    e[i] = c / temp[i];
  }
}

static int lod_steps(cfd_heat* h, int nsteps, bool exact) {
  CFD_REQUIRE(h && nsteps >= 0, "bad argument");
  const cfd_heat_config& g = h->cfg;
  // Row slab of a larger grid: the y-lines cross the slabs.  Both Thomas recurrences forget their start at the rate m per
  // node (see k_lod_y_chunk), so a slab whose halo is `warm` rows deep on each inner side solves its owned rows to the
  // last bit from halo rows the caller refreshed before the step (cfd_heat_implicit_halo says how deep); the exact mode
  // needs whole lines.
  const bool slab = g.ny_global && g.ny_global != g.ny;
  if (slab && exact) {
    cfd_set_error("cfd_heat2d_implicit_step_exact: the y-lines would cross slabs; the exact mode runs on whole grids only");
    return CFD_ERR_UNSUPPORTED;
  }
  CFD_CUDA(cudaSetDevice(h->device));
  const int nx = g.nx, ny = g.ny, nyg = slab ? g.ny_global : g.ny;
  const double dx = (g.xb - g.xa) / (double)(nx - 1), dy = (g.yb - g.ya) / (double)(nyg - 1);
  LodParams P{};
  P.nx = nx;
  P.ny = ny;
  P.wall_s = !slab || g.j_goff == 0;
  P.wall_n = !slab || g.j_goff + ny == nyg;
  P.fox = g.k * g.dt / g.rho / g.cp / (dx * dx);
  P.foy = g.k * g.dt / g.rho / g.cp / (dy * dy);
  P.dt = g.dt; P.rho = g.rho; P.cp = g.cp; P.xa = g.xa; P.dx = dx; P.qthr = g.qthr; P.qval = g.qval;
  P.tleft = g.tleft; P.tright = g.tright; P.tbottom = g.tbottom; P.ttop = g.ttop;
  if (!h->thomas) {
    std::vector<double> ex, tx, ey, ty, all;
    thomas_constants(nx, P.fox, ex, tx);
    thomas_constants(nyg, P.foy, ey, ty);      // by GLOBAL row: the pivots depend on the distance from the wall
    all.insert(all.end(), ex.begin(), ex.end());
    all.insert(all.end(), tx.begin(), tx.end());
    all.insert(all.end(), ey.begin(), ey.end());
    all.insert(all.end(), ty.begin(), ty.end());
    for (double v : tx) all.push_back(1.0 / v);      // reciprocal pivots for the production path
    for (double v : ty) all.push_back(1.0 / v);
    CFD_CUDA(cudaMalloc(&h->thomas, all.size() * sizeof(double)));
    CFD_CUDA(cudaMemcpy(h->thomas, all.data(), all.size() * sizeof(double), cudaMemcpyHostToDevice));
  }
  P.ex = h->thomas;
  P.tx = h->thomas + nx;
  P.ey = h->thomas + 2 * (size_t)nx;
  P.ty = h->thomas + 2 * (size_t)nx + nyg;
  if (!exact) {
    P.tx = h->thomas + 2 * (size_t)nx + 2 * (size_t)nyg;
    P.ty = P.tx + nx;
  }
  const int joff = slab ? g.j_goff : 0;        // the kernels index the by-row constants with the local row
  P.ey += joff;
  P.ty += joff;
  P.qdt = g.dt * g.qval / g.rho / g.cp;
  P.t = h->buf[h->cur];
  const int nmax = nx > ny ? nx : ny;
  // chunked production sweeps: warm-up length from the contraction rate of the recurrences (see k_lod_y_chunk)
  // `slots` = blocks of the sweep kernel resident on the GPU at once: every block lives as long as the kernel, so
  // the grid should fit in one wave (a second, partial wave would double the time)
  int num_sms = 0;   // of this handle's device; one attribute query per call, outside the step loop
  CFD_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, h->device));
  auto plan = [](int n, int nlines, int slots, double fo, LodChunk& c) {
    const double m = fo > 0 ? ((1.0 + 2.0 * fo) - sqrt(1.0 + 4.0 * fo)) / (2.0 * fo) : 0.0;
    int warm = (m > 0 && m < 1) ? (int)ceil(70.0 * log(2.0) / -log(m)) : 1;
    warm = ((warm + 31) / 32) * 32;
    const int line_blocks = (nlines + 127) / 128;
    int nch = slots / line_blocks;
    if (nch > 16) nch = 16;
    if (nch < 1) nch = 1;
    while (nch > 1 && (n / nch < 4 * warm || n / nch < 256)) --nch;
    c.ch = (((n + nch - 1) / nch + 31) / 32) * 32;
    c.warm = warm;
    return nch;
  };
  LodChunk cx{}, cy{};
  const int ncx = plan(nx, ny, num_sms * 4, P.fox, cx), ncy = plan(ny, nx, num_sms * 8, P.foy, cy);
  const bool chunked = !exact && (ncx > 1 || ncy > 1 || slab);      // the chunk kernels are the ones that can start mid-line
  if (slab) {
    const int hs = g.j_goff == 0 ? 0 : cy.warm, hn = (g.j_goff + ny == nyg) ? 0 : cy.warm;
    if (g.halo_s < hs || g.halo_n < hn) {
      cfd_set_error("cfd_heat2d_implicit_step: a row slab needs %d halo rows towards each neighbour at this Fourier number "
                    "(has %d / %d): see cfd_heat_implicit_halo", cy.warm, g.halo_s, g.halo_n);
      return CFD_ERR_INVALID;
    }
  }
  for (int n = 0; n < nsteps; ++n) {
    k_lod_bc<<<(nmax + 127) / 128, 128, 0, h->stream>>>(P);
    if (exact) {
      k_lod_x<true><<<(ny - 2 + 127) / 128, dim3(32, 4), 0, h->stream>>>(P);
      k_lod_y<true><<<(nx - 2 + 127) / 128, 128, 0, h->stream>>>(P);
    } else if (chunked) {
      // forward / backward, out of place: cur -> other -> cur for x, then the same for y; the result is in cur
      double* A = h->buf[h->cur];
      double* B = h->buf[h->cur ^ 1];
      const dim3 gx((ny + 127) / 128, (nx + cx.ch - 1) / cx.ch), gy((nx + 127) / 128, (ny + cy.ch - 1) / cy.ch);
      cx.src = A; cx.dst = B;
      k_lod_x_chunk<0><<<gx, dim3(32, 4), 0, h->stream>>>(P, cx);
      cx.src = B; cx.dst = A;
      k_lod_x_chunk<1><<<gx, dim3(32, 4), 0, h->stream>>>(P, cx);
      cy.src = A; cy.dst = B;
      k_lod_y_chunk<0><<<gy, 128, 0, h->stream>>>(P, cy);
      cy.src = B; cy.dst = A;
      k_lod_y_chunk<1><<<gy, 128, 0, h->stream>>>(P, cy);
    } else {
      k_lod_x<false><<<(ny - 2 + 127) / 128, dim3(32, 4), 0, h->stream>>>(P);
      k_lod_y<false><<<(nx - 2 + 127) / 128, 128, 0, h->stream>>>(P);
    }
    CFD_CUDA(cudaGetLastError());
    ++h->steps_done;
  }
  return CFD_OK;
}

extern "C" int cfd_heat2d_implicit_step(cfd_heat* h, int nsteps) { return lod_steps(h, nsteps, false); }
// halo depth (rows towards each neighbour) a row slab needs for implicit steps at Fourier number fo_y = k dt / (rho cp dy^2):
// the warm-up length of the chunked sweeps, m^W < 2^-70
extern "C" int cfd_heat_implicit_halo(double fo_y, int* rows) {
  CFD_REQUIRE(rows && fo_y > 0, "bad argument");
  const double m = ((1.0 + 2.0 * fo_y) - sqrt(1.0 + 4.0 * fo_y)) / (2.0 * fo_y);
  int warm = (m > 0 && m < 1) ? (int)ceil(70.0 * log(2.0) / -log(m)) : 1;
  *rows = ((warm + 31) / 32) * 32;
  return CFD_OK;
}
extern "C" int cfd_heat2d_implicit_step_exact(cfd_heat* h, int nsteps) { return lod_steps(h, nsteps, true); }


// ---------------------------------------------------------------------------------------------------
// Variable conductivity: the 5-point operator of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:162-190 as it
// stands -- four face coefficients per node, aw = LAMBDA(x-dx/2,y)/dx**2, ae, as = LAMBDA(x,y-dy/2)/dy**2, an
// (:171-179), ap = -(aw+ae+as+an) (:180, formed on the fly), right-hand side -QDOT (:188).  Used three ways:
//   k_heat_apply    the residual of the steady equations
//   k_ftcs_var      the explicit update of fdm1d_explicit_heat_cond.f90:133-135 with this operator: 7 arrays streamed
//                   per step (T, four coefficients, QDOT read; T written) = 56 B/cell against 16 B/cell of the
//                   constant-conductivity kernel; same marching scheme (one thread per x-column, 3-row register window)
//   k_heat_gs       the relaxed Gauss-Seidel of gauss_seidel (:248-287) without the dense (nx*ny)^2 matrix: the five
//                   non-zeros of a row are the coefficient arrays, boundary rows are identity rows (:193-246); nodes in
//                   two colours ((i+j) even, then odd) instead of the single lexicographic order -- same fixed point
// All three use non-contracted IEEE operations in the restatement's order: bit-comparable with oracle/heat_oracle.c.
// ---------------------------------------------------------------------------------------------------
namespace cfd {
struct VarParams {
  int nx, ny;
  const double *aw, *ae, *as, *an, *q;
  double rho, cp, dt, rf;
  double tleft, tright, tbottom, ttop;
  const double* src;
  double* dst;
  double* partial;
};

__device__ __forceinline__ double var_operator(const VarParams& P, size_t k, double W, double E, double S, double N, double C) {
  const double aw = __ldg(P.aw + k), ae = __ldg(P.ae + k), as = __ldg(P.as + k), an = __ldg(P.an + k);
  const double ap = -__dadd_rn(__dadd_rn(__dadd_rn(aw, ae), as), an);
  double r = __dadd_rn(__dmul_rn(aw, W), __dmul_rn(ae, E));
  r = __dadd_rn(r, __dmul_rn(as, S));
  r = __dadd_rn(r, __dmul_rn(an, N));
  r = __dadd_rn(r, __dmul_rn(ap, C));
  return __dadd_rn(r, __ldg(P.q + k));
}

// MODE 0: dst = operator(src) + q on interior nodes, 0 on the boundary; MODE 1: one explicit step src -> dst
template <int MODE>
__global__ void __launch_bounds__(128) k_heat_var(const VarParams P) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.nx) return;
  int j = blockIdx.y * HEAT_RY;
  const int jend = min(j + HEAT_RY - 1, P.ny - 1);
  const size_t nx = (size_t)P.nx;
  const bool xedge = (i == 0) || (i == P.nx - 1);
  const double* s = P.src + (size_t)j * nx + i;
  double* d = P.dst + (size_t)j * nx + i;
  double S = (j > 0) ? __ldg(s - nx) : 0.0;
  double C = __ldg(s);
  for (; j <= jend; ++j, s += nx, d += nx) {
    const double N = (j + 1 < P.ny) ? __ldg(s + nx) : 0.0;
    double out;
    if (j == 0) {
      out = MODE ? P.tbottom : 0.0;
    } else if (j == P.ny - 1) {
      out = MODE ? P.ttop : 0.0;
    } else if (xedge) {
      out = MODE ? (i == 0 ? P.tleft : P.tright) : 0.0;
    } else {
      const double op = var_operator(P, (size_t)j * nx + i, __ldg(s - 1), __ldg(s + 1), S, N, C);
      out = MODE ? __dadd_rn(C, __dmul_rn(__ddiv_rn(__ddiv_rn(op, P.rho), P.cp), P.dt)) : op;
    }
    *d = out;
    S = C;
    C = N;
  }
}

// one colour of a Gauss-Seidel pass, in place; each block leaves its sum of (x1-x2)**2 in partial[block]
// (colour 1 adds to what colour 0 left)
__global__ void __launch_bounds__(256) k_heat_gs(const VarParams P, const int colour) {
  __shared__ double red[256];
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int ii = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = 2 * ii + ((j + colour) & 1);
  double d2 = 0.0;
  if (j < P.ny && i < P.nx) {
    const size_t nx = (size_t)P.nx, k = (size_t)j * nx + i;
    double* x = P.dst;
    const double x1 = x[k];
    double s, b, akk;
    if (i == 0 || i == P.nx - 1 || j == 0 || j == P.ny - 1) {
      s = 0.0;
      akk = 1.0;
      b = (j == 0) ? P.tbottom : (j == P.ny - 1) ? P.ttop : (i == 0) ? P.tleft : P.tright;
    } else {
      const double aw = __ldg(P.aw + k), ae = __ldg(P.ae + k), as = __ldg(P.as + k), an = __ldg(P.an + k);
      s = __dadd_rn(__dmul_rn(as, x[k - nx]), __dmul_rn(aw, x[k - 1]));
      s = __dadd_rn(s, __dmul_rn(ae, x[k + 1]));
      s = __dadd_rn(s, __dmul_rn(an, x[k + nx]));
      akk = -__dadd_rn(__dadd_rn(__dadd_rn(aw, ae), as), an);
      b = -__ldg(P.q + k);
    }
    const double x2 = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn(b, s), P.rf), akk), __dmul_rn(__dsub_rn(1.0, P.rf), x1));
    x[k] = x2;
    const double df = __dsub_rn(x1, x2);
    d2 = __dmul_rn(df, df);
  }
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  red[tid] = d2;
  __syncthreads();
  for (int st = 128; st > 0; st >>= 1) {
    if (tid < st) red[tid] += red[tid + st];
    __syncthreads();
  }
  if (tid == 0) {
    double* p = P.partial + (size_t)blockIdx.y * gridDim.x + blockIdx.x;
    *p = colour ? *p + red[0] : red[0];
  }
}

__global__ void __launch_bounds__(1024) k_heat_gs_norm(const double* partial, int n, double* out) {
  __shared__ double red[1024];
  double a = 0.0;
  for (int k = threadIdx.x; k < n; k += 1024) a += partial[k];
  red[threadIdx.x] = a;
  __syncthreads();
  for (int st = 512; st > 0; st >>= 1) {
    if ((int)threadIdx.x < st) red[threadIdx.x] += red[threadIdx.x + st];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sqrt(red[0]);
}
}  // namespace cfd

static int fill_var(cfd_heat* h, VarParams& P) {
  CFD_REQUIRE(h->op[0], "set the operator first (cfd_heat_set_operator)");
  const cfd_heat_config& g = h->cfg;
  CFD_REQUIRE(g.halo_s == 0 && g.halo_n == 0 && (g.ny_global == 0 || g.ny_global == g.ny),
              "the variable-conductivity path runs on whole grids (no row slabs)");
  P.nx = g.nx; P.ny = g.ny;
  P.aw = h->op[0]; P.ae = h->op[1]; P.as = h->op[2]; P.an = h->op[3]; P.q = h->op[4];
  P.rho = g.rho; P.cp = g.cp; P.dt = g.dt; P.rf = 0.5;
  P.tleft = g.tleft; P.tright = g.tright; P.tbottom = g.tbottom; P.ttop = g.ttop;
  P.partial = nullptr;
  return CFD_OK;
}

extern "C" int cfd_heat_set_operator(cfd_heat* h, const double* lamw, const double* lame, const double* lams,
                                     const double* lamn, const double* qdot, int ld) {
  CFD_REQUIRE(h && lamw && lame && lams && lamn && ld >= h->cfg.nx, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  const cfd_heat_config& g = h->cfg;
  const int nx = g.nx, ny = g.ny;
  const size_t n = (size_t)nx * ny;
  const double dx = (g.xb - g.xa) / (double)(nx - 1), dy = (g.yb - g.ya) / (double)(ny - 1);
  // interior() (:162-190): aw = cdw / dx**2 ... on the host, once (O(n), one division per coefficient)
  std::vector<double> buf(n);
  const double* src[5] = {lamw, lame, lams, lamn, qdot};
  const double div[5] = {dx * dx, dx * dx, dy * dy, dy * dy, 1.0};
  for (int a = 0; a < 5; ++a) {
    if (!h->op[a]) {
      if (cudaMalloc(&h->op[a], n * sizeof(double)) != cudaSuccess) {
        cfd_set_error("cudaMalloc of %zu bytes failed", n * sizeof(double));
        return CFD_ERR_NOMEM;
      }
    }
    for (int j = 0; j < ny; ++j)
      for (int i = 0; i < nx; ++i) {
        const bool interior = i > 0 && i < nx - 1 && j > 0 && j < ny - 1;
        const double v = (src[a] && interior) ? src[a][(size_t)j * ld + i] : 0.0;
        buf[(size_t)j * nx + i] = a < 4 ? v / div[a] : v;
      }
    CFD_CUDA(cudaMemcpyAsync(h->op[a], buf.data(), n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CFD_CUDA(cudaStreamSynchronize(h->stream));
  }
  return CFD_OK;
}

extern "C" int cfd_heat2d_apply(cfd_heat* h, double* host_r, int ld) {
  CFD_REQUIRE(h && host_r && ld >= h->cfg.nx, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  VarParams P{};
  int rc = fill_var(h, P);
  if (rc) return rc;
  P.src = h->buf[h->cur];
  P.dst = h->buf[1 - h->cur];
  dim3 grid((P.nx + 127) / 128, (P.ny + HEAT_RY - 1) / HEAT_RY);
  k_heat_var<0><<<grid, 128, 0, h->stream>>>(P);
  CFD_CUDA(cudaGetLastError());
  const size_t w = (size_t)P.nx * sizeof(double);
  CFD_CUDA(cudaMemcpy2DAsync(host_r, (size_t)ld * sizeof(double), P.dst, w, w, P.ny, cudaMemcpyDeviceToHost, h->stream));
  CFD_CUDA(cudaStreamSynchronize(h->stream));
  return CFD_OK;
}

extern "C" int cfd_heat2d_step_var(cfd_heat* h, int nsteps) {
  CFD_REQUIRE(h && nsteps >= 0, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  VarParams P{};
  int rc = fill_var(h, P);
  if (rc) return rc;
  dim3 grid((P.nx + 127) / 128, (P.ny + HEAT_RY - 1) / HEAT_RY);
  for (int n = 0; n < nsteps; ++n) {
    P.src = h->buf[h->cur];
    P.dst = h->buf[1 - h->cur];
    k_heat_var<1><<<grid, 128, 0, h->stream>>>(P);
    h->cur = 1 - h->cur;
    ++h->steps_done;
  }
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

extern "C" int cfd_heat2d_steady_gs(cfd_heat* h, int maxit, double tol, double rf, int* iters, double* dx2) {
  CFD_REQUIRE(h && maxit >= 1 && rf > 0.0, "bad argument");
  CFD_CUDA(cudaSetDevice(h->device));
  VarParams P{};
  int rc = fill_var(h, P);
  if (rc) return rc;
  P.rf = rf;
  P.dst = h->buf[h->cur];
  dim3 block(32, 8), grid(((P.nx + 1) / 2 + 31) / 32, (P.ny + 7) / 8);
  const int nb = grid.x * grid.y;
  if (!h->gs_partial || h->gs_blocks < nb) {
    if (h->gs_partial) cudaFree(h->gs_partial);
    CFD_CUDA(cudaMalloc(&h->gs_partial, ((size_t)nb + 1) * sizeof(double)));
    h->gs_blocks = nb;
  }
  P.partial = h->gs_partial;
  double norm = 0.0;
  int k = 0;
  for (k = 1; k <= maxit; ++k) {                       // gauss_seidel :263-285
    k_heat_gs<<<grid, block, 0, h->stream>>>(P, 0);
    k_heat_gs<<<grid, block, 0, h->stream>>>(P, 1);
    k_heat_gs_norm<<<1, 1024, 0, h->stream>>>(h->gs_partial, nb, h->gs_partial + nb);
    CFD_CUDA(cudaMemcpyAsync(&norm, h->gs_partial + nb, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CFD_CUDA(cudaStreamSynchronize(h->stream));
    if (norm < tol) break;
  }
  if (k > maxit) k = maxit;
  if (iters) *iters = k;
  if (dx2) *dx2 = norm;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
