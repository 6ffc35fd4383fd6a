// tdma_kernels.cu -- the line solver lisolv (cavity_fvm.f90:49-82; general.f90:294-327; tdma buoyancy_cav.f90:471-505).
//
// The reference sweeps W->E, solving one S-N tridiagonal system per line with the Thomas algorithm, each line
// using the freshly updated western neighbour (line Gauss-Seidel).  Two device versions:
//
//  k_tdma_rb<M>  red/black line ordering: all lines of one colour are independent.  One group of T threads
//                per line; thread t holds nodes t*M..t*M+M-1 of the line in registers (coalesced loads thanks
//                to the chunk-interleaved layout), eliminates inside its chunk (Wang's partition method:
//                forward elimination with a spike to the previous chunk, backward elimination onto the
//                chunk's last node), the T chunk-end unknowns form a tridiagonal system solved by parallel
//                cyclic reduction in shared memory, then every thread back-substitutes its chunk.  Each
//                coefficient is read exactly once and phi written once: 8 doubles per cell, the algorithmic
//                figure for a sweep.
//  k_tdma_seq    the reference's sequential line order and operation order, for bit-comparable debugging on
//                small grids: one block, lines one after the other, the recurrence run by a single thread out
//                of shared memory with non-contracted IEEE operations.
#include "common.cuh"

namespace cfd {

// reciprocal for the red/black solver (not bit-matched to the reference by construction)
__device__ __forceinline__ double rcp(double x) { return 1.0 / x; }

template <int M, int BMAX>
__global__ void __launch_bounds__(BMAX, 1) k_tdma_rb(const SweepParams P) {
  extern __shared__ double sh[];
  const int T = P.L.T;
  const long long LD = P.L.LD;
  const int t = threadIdx.x, ly = threadIdx.y, LPB = blockDim.y;
  const int line = blockIdx.x * LPB + ly;
  const bool valid = line < P.nlines;
  const int i0 = P.i_first + (valid ? line : 0) * P.i_step;
  const int nth = LPB * T;
  // shared arrays: two PCR buffers of (l,u,r)
  double* bufA = sh;
  double* bufB = sh + 3 * nth;
  const int sidx = ly * T + t;

  const long long o = (long long)i0 * LD;
  const double* an = P.an + o;
  const double* as = P.as + o;
  const double* ap = P.ap + o;
  const double* ae = P.ae + o;
  const double* aw = P.aw + o;
  const double* su = P.su + o;
  double* phi = P.phi + o;
  const double* phiE = phi + LD;
  const double* phiW = phi - LD;
  const bool zeroE = P.first && (P.first_colour == 0 || P.i_step == 1);
  const bool zeroW = P.first && (P.first_colour == 0) && P.i_step != 1;

  double ca[M], cc[M], cf[M];
  const int j0base = t * M;
  // ---- load + forward elimination inside the chunk (spike cf couples to the previous chunk's last node)
  {
    double a[M], b[M], d[M];
#pragma unroll
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k, q = k * T + t;
      const bool act = valid && j0 >= P.jlo && j0 <= P.jhi;
      if (act) {
        a[k] = __ldg(an + q);
        b[k] = __ldg(as + q);
        d[k] = __ldg(ap + q);
        const double e = zeroE ? 0.0 : phiE[q];
        const double w = zeroW ? 0.0 : phiW[q];
        cc[k] = fma(__ldg(ae + q), e, fma(__ldg(aw + q), w, __ldg(su + q)));
      } else {
        a[k] = 0.0;
        b[k] = 0.0;
        d[k] = 1.0;
        cc[k] = (valid && j0 < P.L.nj && !(P.first)) ? phi[q] : 0.0;
      }
    }
    {
      const double inv = rcp(d[0]);
      ca[0] = a[0] * inv;
      cc[0] = cc[0] * inv;
      cf[0] = b[0] * inv;
    }
#pragma unroll
    for (int k = 1; k < M; ++k) {
      const double inv = rcp(fma(-b[k], ca[k - 1], d[k]));
      ca[k] = a[k] * inv;
      cc[k] = fma(b[k], cc[k - 1], cc[k]) * inv;
      cf[k] = b[k] * cf[k - 1] * inv;
    }
  }
  // ---- backward elimination onto the chunk's last node: x_k = cc_k + cf_k*x_prev + ca_k*x_last  (k < M-1)
#pragma unroll
  for (int k = M - 3; k >= 0; --k) {
    cc[k] = fma(ca[k], cc[k + 1], cc[k]);
    cf[k] = fma(ca[k], cf[k + 1], cf[k]);
    ca[k] = ca[k] * ca[k + 1];
  }
  // ---- reduced system over chunk-last nodes: x_t = l*x_{t-1} + u*x_{t+1} + r
  double l, u, r;
  if (M == 1) {
    l = cf[0];
    u = ca[0];
    r = cc[0];
  } else {
    bufB[sidx] = cc[0];
    bufB[nth + sidx] = cf[0];
    bufB[2 * nth + sidx] = ca[0];
    __syncthreads();
    double ncc = 0.0, ncf = 0.0, ng = 0.0;
    if (t + 1 < T) {
      ncc = bufB[sidx + 1];
      ncf = bufB[nth + sidx + 1];
      ng = bufB[2 * nth + sidx + 1];
    }
    const double inv = rcp(fma(-ca[M - 1], ncf, 1.0));
    l = cf[M - 1] * inv;
    u = ca[M - 1] * ng * inv;
    r = fma(ca[M - 1], ncc, cc[M - 1]) * inv;
  }
  double* cur = bufA;
  double* nxt = bufB;
  cur[sidx] = l;
  cur[nth + sidx] = u;
  cur[2 * nth + sidx] = r;
  __syncthreads();
  // ---- parallel cyclic reduction
  for (int s = 1; s < T; s <<= 1) {
    double lm = 0.0, um = 0.0, rm = 0.0, lp = 0.0, up = 0.0, rp = 0.0;
    if (t - s >= 0) {
      lm = cur[sidx - s];
      um = cur[nth + sidx - s];
      rm = cur[2 * nth + sidx - s];
    }
    if (t + s < T) {
      lp = cur[sidx + s];
      up = cur[nth + sidx + s];
      rp = cur[2 * nth + sidx + s];
    }
    const double inv = rcp(fma(-l, um, fma(-u, lp, 1.0)));
    r = fma(l, rm, fma(u, rp, r)) * inv;
    l = l * lm * inv;
    u = u * up * inv;
    nxt[sidx] = l;
    nxt[nth + sidx] = u;
    nxt[2 * nth + sidx] = r;
    __syncthreads();
    double* tmp = cur;
    cur = nxt;
    nxt = tmp;
  }
  const double xe = r;
  const double xp = (t > 0) ? cur[2 * nth + sidx - 1] : 0.0;
  // ---- back-substitute and store
  if (valid) {
#pragma unroll
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k, q = k * T + t;
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double xk = (k == M - 1) ? xe : fma(cf[k], xp, fma(ca[k], xe, cc[k]));
        phi[q] = xk;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Sequential-order debug sweep (exactly cavity_fvm.f90:62-81).
// ---------------------------------------------------------------------------------------------------
struct SeqParams {
  Layout L;
  int ia, ib;        // 0-based inclusive local line range, swept in increasing order
  int jlo, jhi;
  int first;         // phi(i+1,:) taken as zero (pp(:,:) = 0 folded in)
  const double *ae, *aw, *an, *as, *ap, *su;
  double* phi;
};

__global__ void __launch_bounds__(1024, 1) k_tdma_seq(const SeqParams P) {
  extern __shared__ double sh[];
  const int nj = P.L.nj, M = P.L.M, T = P.L.T;
  const long long LD = P.L.LD;
  double* sa = sh;            // an  -> a'
  double* sb = sa + nj;       // as
  double* sd = sb + nj;       // ap
  double* sc = sd + nj;       // c   -> c'
  double* sx = sc + nj;       // phi of this line
  for (int i0 = P.ia; i0 <= P.ib; ++i0) {
    const long long o = (long long)i0 * LD;
    const double* phiC = P.phi + o;
    for (int j0 = threadIdx.x; j0 < nj; j0 += blockDim.x) {
      const int q = lpos(j0, M, T);
      sx[j0] = phiC[q];
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double e = P.first ? 0.0 : phiC[LD + q];
        const double w = phiC[q - LD];
        sa[j0] = P.an[o + q];
        sb[j0] = P.as[o + q];
        sd[j0] = P.ap[o + q];
        sc[j0] = __dadd_rn(__dadd_rn(__dmul_rn(P.ae[o + q], e), __dmul_rn(P.aw[o + q], w)), P.su[o + q]);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double ap_ = 0.0, cp_ = sx[P.jlo - 1];
      for (int j0 = P.jlo; j0 <= P.jhi; ++j0) {
        const double b = sb[j0];
        const double term = __ddiv_rn(1.0, __dsub_rn(sd[j0], __dmul_rn(b, ap_)));
        ap_ = __dmul_rn(sa[j0], term);
        cp_ = __dmul_rn(__dadd_rn(sc[j0], __dmul_rn(b, cp_)), term);
        sa[j0] = ap_;
        sc[j0] = cp_;
      }
      double xn = sx[P.jhi + 1];
      for (int j0 = P.jhi; j0 >= P.jlo; --j0) {
        xn = __dadd_rn(__dmul_rn(sa[j0], xn), sc[j0]);
        sx[j0] = xn;
      }
    }
    __syncthreads();
    for (int j0 = P.jlo + threadIdx.x; j0 <= P.jhi; j0 += blockDim.x) P.phi[o + lpos(j0, M, T)] = sx[j0];
    __syncthreads();   // make the new line visible to the loads of the next line (same block)
  }
}

// ---- launch tables -------------------------------------------------------------------------------------
template <int M>
static int launch_rb_m(cfd_ctx* c, const SweepParams& P, int nblocks, dim3 block, size_t smem) {
  const int nthreads = block.x * block.y;
  if (nthreads <= 384) {
    k_tdma_rb<M, 384><<<nblocks, block, smem, c->stream>>>(P);
  } else if (nthreads <= 640) {
    k_tdma_rb<M, 640><<<nblocks, block, smem, c->stream>>>(P);
  } else {
    k_tdma_rb<M, 1024><<<nblocks, block, smem, c->stream>>>(P);
  }
  return CFD_OK;
}

}  // namespace cfd

using namespace cfd;

int cfd_launch_sweep_rb(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first) {
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  // owned lines, global 1-based [glo, ghi], restricted to istart..NI-1
  int glo = c->cfg.goff + c->cfg.halo_w + 1, ghi = c->cfg.goff + c->L.ni - c->cfg.halo_e;
  if (glo < istart) glo = istart;
  if (ghi > NI - 1) ghi = NI - 1;
  // first line of this colour: (i - istart) % 2 == colour
  int g0 = glo + ((((glo - istart) & 1) != colour) ? 1 : 0);
  if (g0 > ghi) return CFD_OK;
  SweepParams P{};
  P.L = c->L;
  P.i_first = g0 - 1 - c->cfg.goff;
  P.i_step = 2;
  P.nlines = (ghi - g0) / 2 + 1;
  P.jlo = jstart - 1;
  P.jhi = c->L.nj - 2;
  P.first = first ? 1 : 0;
  P.first_colour = colour;
  P.ae = c->field[CFD_AE];
  P.aw = c->field[CFD_AW];
  P.an = c->field[CFD_AN];
  P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP];
  P.su = c->field[CFD_SU];
  P.phi = phi;
  const int T = c->L.T;
  int lpb = 256 / T;
  if (lpb < 1) lpb = 1;
  dim3 block(T, lpb);
  const int nblocks = (P.nlines + lpb - 1) / lpb;
  const size_t smem = (size_t)6 * lpb * T * sizeof(double);
  switch (c->L.M) {
#define CFD_CASE(m) \
  case m:           \
    launch_rb_m<m>(c, P, nblocks, block, smem); \
    break;
    CFD_CASE(1) CFD_CASE(2) CFD_CASE(3) CFD_CASE(4) CFD_CASE(5) CFD_CASE(6) CFD_CASE(7) CFD_CASE(8)
    CFD_CASE(9) CFD_CASE(10) CFD_CASE(11) CFD_CASE(12) CFD_CASE(13) CFD_CASE(14) CFD_CASE(15) CFD_CASE(16)
#undef CFD_CASE
    default:
      cfd_set_error("unsupported M=%d", c->L.M);
      return CFD_ERR_INVALID;
  }
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

int cfd_launch_sweep_seq(cfd_ctx* c, int istart, int jstart, double* phi, int first) {
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  if (c->cfg.goff != 0 || (c->cfg.ni_global && c->cfg.ni_global != c->L.ni)) {
    cfd_set_error("sequential-order sweeps do not shard: replicas only (single-GPU contexts)");
    return CFD_ERR_UNSUPPORTED;
  }
  SeqParams P{};
  P.L = c->L;
  P.ia = istart - 1;
  P.ib = NI - 2;
  P.jlo = jstart - 1;
  P.jhi = c->L.nj - 2;
  P.first = first ? 1 : 0;
  P.ae = c->field[CFD_AE];
  P.aw = c->field[CFD_AW];
  P.an = c->field[CFD_AN];
  P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP];
  P.su = c->field[CFD_SU];
  P.phi = phi;
  const size_t smem = (size_t)5 * c->L.nj * sizeof(double);
  if (smem > 200 * 1024) {
    cfd_set_error("sequential-order debug sweep supports nj <= 5120 (got %d)", c->L.nj);
    return CFD_ERR_UNSUPPORTED;
  }
  static bool attr_set = false;
  if (!attr_set) {
    CFD_CUDA(cudaFuncSetAttribute(k_tdma_seq, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  k_tdma_seq<<<1, 1024, smem, c->stream>>>(P);
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
