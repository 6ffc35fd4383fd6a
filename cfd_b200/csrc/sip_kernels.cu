// sip_kernels.cu -- Stone's strongly implicit procedure (SIP) for the 5-point p' equation of the explicit
// finite-difference cavity solver: sipsol, lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90:24-87
// (incomplete LU factors :44-55, residual + forward substitution :59-67, backward substitution + update :71-76,
// stop rule :78-85; maxit 600, alpha 0.88, tol 1e-5).
//
// Every step of the algorithm has a recurrence in BOTH grid directions (cell (i,j) needs (i-1,j) and (i,j-1)), so
// the parallelism is a wavefront.  One persistent cooperative kernel runs the whole solve: the rows are cut into
// bands of 32 (one warp per band), a warp walks its band chunk by chunk (32 columns) and solves each 32x32 tile
// in shared memory along its 63 anti-diagonals; band b may work on chunk c as soon as band b-1 has published
// chunk c (a release/acquire progress counter per band in global memory), so the bands form a software pipeline
// that keeps nj/32 warps busy.  Tiles are read and written as 256-byte row segments.  The backward substitution
// is the same scheme mirrored.  Between the phases of an iteration the grid synchronises (cooperative groups);
// the residual norm is reduced in a fixed order, so every block takes the same stop decision.
// Operation order and non-contracted IEEE arithmetic as in the reference: the result is bit-identical to
// oracle/fdcavity_oracle.c:orc_fdc_sipsol whenever the iteration counts agree (they differ only if the ratio
// rsm lands within rounding of tol).
#include <cooperative_groups.h>
#include "common.cuh"
namespace cg = cooperative_groups;

namespace cfd {

struct SipParams {
  int imax, jmax, ld;        // arrays are column-major with leading dimension ld; interior (0-based) 1..imax x 1..jmax
  double aw, ae, as, an, ap; // constant coefficients as passed to sipsol (already negated by the caller, :206) ...
  const double *awp, *aep, *asp, *anp, *app;   // ... or coefficient arrays (fdm_implicit_cavity.f90:199): used if non-null,
  double sign;               // the four neighbour arrays multiplied by `sign` (-1: the caller's `-aw, -ae, -as, -an`)
  double alpha, tol;
  int maxit;
  const double* su;
  double* phi;
  double *lw, *ls, *lpr, *un, *ue, *res;   // work arrays, same shape
  unsigned long long* prog;  // one progress counter per band
  double* partial;           // per-block residual partial sums
  int* info;                 // [0] iterations used, [1] abort flag
  double* rsm_out;
};

#ifndef SIP_TIMING
#define SIP_TIMING 0
#endif
#define SIP_W 4              // warps (bands) per block
#define TS 34                // padded tile stride: the anti-diagonal (lane l at row l, column d-l) sits at 33*l + d -> 32 different banks

__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// wait until band `b` has published at least `target`; returns false if the wait was abandoned
__device__ __forceinline__ bool wait_band(const SipParams& P, int b, unsigned long long target) {
  unsigned spins = 0;
  while (ld_acquire(P.prog + b) < target) {
    __nanosleep(64);
    if (++spins > (1u << 24) || *(volatile int*)(P.info + 1)) {
      P.info[1] = 1;
      return false;
    }
  }
  return true;
}

// load / store one 32x32 tile (rows jb..jb+31, columns c0..c0+31, 0-based) as coalesced row segments
__device__ __forceinline__ void tile_load(double* tl, const double* g, int ld, int c0, int jb, int imax, int jmax, int lane) {
  // sixteen row segments in flight per lane before the first one is needed (a rolled loop would wait for every load)
  const int i = c0 + lane;
  const bool iok = i >= 1 && i <= imax;
#pragma unroll
  for (int r0 = 0; r0 < 32; r0 += 16) {
    double v[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int j = jb + r0 + r;
      v[r] = (iok && j >= 1 && j <= jmax) ? __ldcg(g + (size_t)j * ld + i) : 0.0;
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) tl[(r0 + r) * TS + lane] = v[r];
  }
}
__device__ __forceinline__ void tile_store(const double* tl, double* g, int ld, int c0, int jb, int imax, int jmax, int lane) {
  for (int r = 0; r < 32; ++r) {
    const int i = c0 + lane, j = jb + r;
    if (i >= 1 && i <= imax && j >= 1 && j <= jmax) g[(size_t)j * ld + i] = tl[r * TS + lane];
  }
}

// the four operand tiles of a chunk in two round trips: sixteen rows of all four arrays (64 values per lane) are in
// flight together (one block per SM leaves each thread up to 255 registers)
__device__ __forceinline__ void tile_load4(double* A, double* B, double* C, double* D, const double* ga, const double* gb,
                                           const double* gc, const double* gd, int ld, int c0, int jb, int imax, int jmax, int lane) {
  const int i = c0 + lane;
  const bool iok = i >= 1 && i <= imax;
#pragma unroll
  for (int r0 = 0; r0 < 32; r0 += 16) {
    double va[16], vb[16], vc[16], vd[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int j = jb + r0 + r;
      const bool ok = iok && j >= 1 && j <= jmax;
      const size_t k = (size_t)j * ld + i;
      va[r] = ok ? __ldcg(ga + k) : 0.0;
      vb[r] = ok ? __ldcg(gb + k) : 0.0;
      vc[r] = ok ? __ldcg(gc + k) : 0.0;
      vd[r] = ok ? __ldcg(gd + k) : 0.0;
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int o = (r0 + r) * TS + lane;
      A[o] = va[r]; B[o] = vb[r]; C[o] = vc[r]; D[o] = vd[r];
    }
  }
}

// PHASE 0: incomplete LU factors (:44-55); 1: forward substitution (:64-65); 2: backward substitution + update (:71-76)
template <int PHASE>
__device__ void sip_wavefront(const SipParams& P, double* sm, unsigned long long epoch_base, int nbands, int nchunks) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int imax = P.imax, jmax = P.jmax;              // interior 0-based: 1..imax, 1..jmax
  double* t0 = sm + (size_t)w * (4 * 32 * TS + 4 * 32);
  double *A = t0, *B = A + 32 * TS, *C = B + 32 * TS, *D = C + 32 * TS;   // tiles
  double *row0 = D + 32 * TS, *row1 = row0 + 32, *col0 = row1 + 32, *col1 = col0 + 32;
  const bool fwd = PHASE != 2;
  long long tm[6] = {0, 0, 0, 0, 0, 0}, tq = 0;
#define TICK(n) if (SIP_TIMING) { const long long now__ = clock64(); tm[n] += now__ - tq; tq = now__; }
  for (int bb = blockIdx.x * SIP_W + w; bb < nbands; bb += gridDim.x * SIP_W) {
    const int b = fwd ? bb : nbands - 1 - bb;           // band processed (backward: from the top)
    const int jb = 1 + 32 * b;
    col0[lane] = 0.0;                                   // values of the column just outside the first chunk
    col1[lane] = 0.0;
    __syncwarp();
    for (int cc = 0; cc < nchunks; ++cc) {
      const int c = fwd ? cc : nchunks - 1 - cc;
      const int c0 = 1 + 32 * c;
      if (SIP_TIMING) tq = clock64();
      // the tile's own operands do not depend on the neighbouring band: fetch them before waiting for it
      if (PHASE == 1) {
        tile_load4(A, B, C, D, P.res, P.ls, P.lw, P.lpr, P.ld, c0, jb, imax, jmax, lane);
      } else if (PHASE == 2) {
        tile_load4(A, B, C, D, P.res, P.un, P.ue, P.phi, P.ld, c0, jb, imax, jmax, lane);
      }
      TICK(0)
      if (bb > 0) {                                     // the band processed just before this one (other warp/block)
        const int dep = fwd ? b - 1 : b + 1;
        if (!wait_band(P, dep, epoch_base + (unsigned long long)cc + 1)) return;
      }
      TICK(1)
      // row just outside the tile in the sweep direction (south for forward phases, north for backward)
      {
        const int jn = fwd ? jb - 1 : jb + 32, i = c0 + lane;
        const bool ok = jn >= 1 && jn <= jmax && i >= 1 && i <= imax;
        if (PHASE == 0) {
          row0[lane] = ok ? P.un[(size_t)jn * P.ld + i] : 0.0;
          row1[lane] = ok ? P.ue[(size_t)jn * P.ld + i] : 0.0;
        } else {
          row0[lane] = ok ? P.res[(size_t)jn * P.ld + i] : 0.0;
        }
      }
      __syncwarp();
      TICK(2)
      if (PHASE == 0) {
        for (int d = 0; d < 63; ++d) {
          const int r = lane, q = d - lane;               // row / column inside the tile
          const bool act = q >= 0 && q < 32 && (c0 + q) <= imax && (jb + r) <= jmax;
          if (act) {
            const int k = r * TS + q;
            // west / south neighbours of the factors un, ue
            const double un_w = (q == 0) ? col0[r] : A[k - 1], ue_w = (q == 0) ? col1[r] : B[k - 1];
            const double un_s = (r == 0) ? row0[q] : A[k - TS], ue_s = (r == 0) ? row1[q] : B[k - TS];
            const size_t kg = (size_t)(jb + r) * P.ld + c0 + q;
            const double caw = P.awp ? P.sign * P.awp[kg] : P.aw, cas = P.asp ? P.sign * P.asp[kg] : P.as;
            const double can = P.anp ? P.sign * P.anp[kg] : P.an, cae = P.aep ? P.sign * P.aep[kg] : P.ae;
            const double cap = P.app ? P.app[kg] : P.ap;
            const double lw = __ddiv_rn(caw, __dadd_rn(1.0, __dmul_rn(P.alpha, un_w)));
            const double ls = __ddiv_rn(cas, __dadd_rn(1.0, __dmul_rn(P.alpha, ue_s)));
            const double p1 = __dmul_rn(__dmul_rn(P.alpha, lw), un_w);
            const double p2 = __dmul_rn(__dmul_rn(P.alpha, ls), ue_s);
            const double den = __dsub_rn(__dsub_rn(__dadd_rn(__dadd_rn(cap, p1), p2), __dmul_rn(lw, ue_w)), __dmul_rn(ls, un_s));
            const double lpr = __ddiv_rn(1.0, den);
            A[k] = __dmul_rn(__dsub_rn(can, p1), lpr);    // un
            B[k] = __dmul_rn(__dsub_rn(cae, p2), lpr);    // ue
            C[k] = lw;
            D[k] = ls;
            // lpr goes straight to global memory (a fifth tile would not fit the budget)
            P.lpr[kg] = lpr;
          }
          __syncwarp();
        }
      } else {
        // Forward / backward substitution: lane = row, marching along the columns one step behind the lane before it.
        // The two recurrence inputs never touch shared memory: the value of the previous row at this column is what
        // the neighbouring lane produced one step ago (shuffle), the value of the previous column is this lane's own
        // last result (register).  Shared memory only supplies the cell's operands and receives the result.
        double last = 0.0;
        // per-lane constants of the skewed walk: the lane is active for d in [lane, dhi]; its cell at step d sits at
        // kb + d (forward) / kb - d (backward) in the tile, the row-0 / column-0 inputs likewise
        const int r = fwd ? lane : 31 - lane;
        int ncol = imax - c0 + 1;                          // valid columns of this chunk
        if (ncol > 32) ncol = 32;
        const bool rowok = (jb + r) <= jmax;
        // forward: q = d - lane valid for q < ncol; backward: q = 31 - (d - lane) valid for q < ncol <=> d - lane > 31 - ncol
        const int dlo = fwd ? lane : lane + (32 - ncol);
        const int dhi = rowok ? (fwd ? lane + ncol - 1 : lane + 31) : -1;
        const int kb = fwd ? r * TS - lane : r * TS + 31 + lane;
        const double cin = col0[r];
#pragma unroll
        for (int d = 0; d < 63; ++d) {
          const double from_row = __shfl_up_sync(0xffffffffu, last, 1);
          if (d >= dlo && d <= dhi) {
            const int k = fwd ? kb + d : kb - d;
            const int q = fwd ? d - lane : 31 + lane - d;
            const double nb_row = (lane == 0) ? row0[q] : from_row;              // south (forward) / north (backward)
            const double nb_col = (d == lane) ? cin : last;                      // west (forward) / east (backward)
            if (PHASE == 1) {
              last = __dmul_rn(__dsub_rn(__dsub_rn(A[k], __dmul_rn(B[k], nb_row)), __dmul_rn(C[k], nb_col)), D[k]);
              A[k] = last;
            } else {
              last = __dsub_rn(__dsub_rn(A[k], __dmul_rn(B[k], nb_row)), __dmul_rn(C[k], nb_col));
              A[k] = last;
              D[k] = __dadd_rn(D[k], last);
            }
          }
        }
        __syncwarp();
      }
      TICK(3)
      // keep the column next to the following chunk, write the tile back, publish
      {
        const int qe = fwd ? 31 : 0;
        col0[lane] = A[lane * TS + qe];
        if (PHASE == 0) col1[lane] = B[lane * TS + qe];
      }
      if (PHASE == 0) {
        tile_store(A, P.un, P.ld, c0, jb, imax, jmax, lane);
        tile_store(B, P.ue, P.ld, c0, jb, imax, jmax, lane);
        tile_store(C, P.lw, P.ld, c0, jb, imax, jmax, lane);
        tile_store(D, P.ls, P.ld, c0, jb, imax, jmax, lane);
      } else if (PHASE == 1) {
        tile_store(A, P.res, P.ld, c0, jb, imax, jmax, lane);
      } else {
        tile_store(A, P.res, P.ld, c0, jb, imax, jmax, lane);
        tile_store(D, P.phi, P.ld, c0, jb, imax, jmax, lane);
      }
      TICK(4)
      __threadfence();
      __syncwarp();
      if (lane == 0) st_release(P.prog + b, epoch_base + (unsigned long long)cc + 1);
      TICK(5)
    }
    if (SIP_TIMING && lane == 0 && (bb == 1 || bb == nbands / 2) && epoch_base / ((unsigned long long)nchunks + 1) == (PHASE == 1 ? 6 : 7))
      printf("SIP phase %d band %d: loads %lld wait %lld row0 %lld steps %lld stores %lld fence+flag %lld cycles over %d chunks\n",
             PHASE, bb, tm[0], tm[1], tm[2], tm[3], tm[4], tm[5], nchunks);
  }
#undef TICK
}

__global__ void __launch_bounds__(32 * SIP_W) k_sip_solve(const SipParams P) {
  extern __shared__ double sm[];
  cg::grid_group grid = cg::this_grid();
  const int imax = P.imax, jmax = P.jmax;
  const int nbands = (jmax + 31) / 32, nchunks = (imax + 31) / 32;
  const unsigned long long stride = (unsigned long long)nchunks + 1;
  unsigned long long epoch = 1;
  // ---- incomplete factorisation
  sip_wavefront<0>(P, sm, epoch * stride, nbands, nchunks);
  ++epoch;
  grid.sync();
  double res1 = 0.0, rsm = 0.0;
  int iter = 0;
  for (iter = 1; iter <= P.maxit; ++iter) {
    // ---- residual (:61-63) and its 1-norm, fixed-order reduction
    double acc = 0.0;
    // rows are dealt to the warps of the whole grid, a warp reads 256-byte row segments, four per lane in flight
    {
      const int lane = threadIdx.x & 31;
      const int gw = blockIdx.x * SIP_W + (threadIdx.x >> 5), nw = gridDim.x * SIP_W;
      for (int j = 1 + gw; j <= jmax; j += nw) {
        const size_t row = (size_t)j * P.ld;
        for (int i0 = 1; i0 <= imax; i0 += 128) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int i = i0 + 32 * u + lane;
            if (i <= imax) {
              const size_t k = row + i;
              const double caw = P.awp ? P.sign * P.awp[k] : P.aw, cae = P.aep ? P.sign * P.aep[k] : P.ae;
              const double can = P.anp ? P.sign * P.anp[k] : P.an, cas = P.asp ? P.sign * P.asp[k] : P.as;
              const double cap = P.app ? P.app[k] : P.ap;
              double r = __dsub_rn(P.su[k], __dmul_rn(caw, P.phi[k - 1]));
              r = __dsub_rn(r, __dmul_rn(cae, P.phi[k + 1]));
              r = __dsub_rn(r, __dmul_rn(can, P.phi[k + P.ld]));
              r = __dsub_rn(r, __dmul_rn(cas, P.phi[k - P.ld]));
              r = __dsub_rn(r, __dmul_rn(cap, P.phi[k]));
              P.res[k] = r;
              acc += fabs(r);
            }
          }
        }
      }
    }
    {
      __shared__ double wsum[SIP_W];
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
      if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
      __syncthreads();
      if (threadIdx.x == 0) {
        double s = 0.0;
        for (int k = 0; k < SIP_W; ++k) s += wsum[k];
        P.partial[blockIdx.x] = s;
      }
    }
    grid.sync();
    double resl = 0.0;
    for (unsigned k = 0; k < gridDim.x; ++k) resl += P.partial[k];   // same order in every block
    if (iter == 1) res1 = resl;
    rsm = resl / res1;
    // ---- forward and backward substitution
    sip_wavefront<1>(P, sm, epoch * stride, nbands, nchunks);
    ++epoch;
    grid.sync();
    sip_wavefront<2>(P, sm, epoch * stride, nbands, nchunks);
    ++epoch;
    grid.sync();
    if (rsm < P.tol) break;
    if (iter == 1 && res1 < 1.0e-10) break;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    P.info[0] = iter > P.maxit ? P.maxit : iter;
    *P.rsm_out = rsm;
  }
}

}  // namespace cfd
using namespace cfd;

// Launch one SIP solve.  coef != nullptr: five coefficient arrays {aw, ae, as, an, ap} (the four neighbour arrays are
// negated, as the callers pass `-aw, -ae, -as, -an`); otherwise the constants c[5].  Work arrays are the caller's.
int cfd_launch_sip(int device, cudaStream_t stream, int ld, int nrows, int imax, int jmax, const double* c,
                   const double* const* coef, double alpha, double tol, int maxit, const double* su, double* phi, double* work,
                   unsigned long long* prog, double* partial, int* info, double* rsm_out, int max_partials) {
  SipParams P{};
  P.imax = imax; P.jmax = jmax; P.ld = ld;
  if (coef) {
    P.awp = coef[0]; P.aep = coef[1]; P.asp = coef[2]; P.anp = coef[3]; P.app = coef[4];
    P.sign = -1.0;
  } else {
    P.aw = c[0]; P.ae = c[1]; P.as = c[2]; P.an = c[3]; P.ap = c[4];
  }
  P.alpha = alpha; P.tol = tol; P.maxit = maxit;
  P.su = su; P.phi = phi;
  const size_t n = (size_t)ld * nrows;
  P.lw = work; P.ls = work + n; P.lpr = work + 2 * n; P.un = work + 3 * n; P.ue = work + 4 * n; P.res = work + 5 * n;
  P.prog = prog; P.partial = partial; P.info = info; P.rsm_out = rsm_out;
  const int nbands = (jmax + 31) / 32;
  const size_t smem = (size_t)SIP_W * (4 * 32 * TS + 4 * 32) * sizeof(double);
  CFD_CUDA(cudaFuncSetAttribute(k_sip_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0, nsm = 0;
  CFD_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sip_solve, 32 * SIP_W, smem));
  CFD_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device));
  // every SM takes part (the residual phase is a plain streaming pass); bands beyond nbands simply do not exist
  int grid = per_sm * nsm;
  (void)nbands;
  if (grid > max_partials) grid = max_partials;
  if (grid < 1) grid = 1;
  CFD_CUDA(cudaMemsetAsync(work, 0, 6 * n * sizeof(double), stream));
  CFD_CUDA(cudaMemsetAsync(prog, 0, (size_t)(nbands + 1) * sizeof(unsigned long long), stream));
  CFD_CUDA(cudaMemsetAsync(info, 0, 2 * sizeof(int), stream));
  void* args[] = {(void*)&P};
  CFD_CUDA(cudaLaunchCooperativeKernel((void*)k_sip_solve, dim3(grid), dim3(32 * SIP_W), args, smem, stream));
  return CFD_OK;
}
