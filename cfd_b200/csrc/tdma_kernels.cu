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
#include "tma.cuh"

namespace cfd {

// Reciprocal for the red/black solver (not bit-matched to the reference by construction): hardware seed
// (MUFU.RCP64H, ~20 good bits) + two Newton steps = 5 FP64-pipe instructions instead of the ~25 of an IEEE
// divide; relative error about 1 ulp.  The pivots of a diagonally dominant system are far from 0/inf/denormal.
__device__ __forceinline__ double rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

// Ask the memory system to pull `bytes` (multiple of 16) starting at `p` (16-byte aligned) into L2.  One
// instruction per array line: the next line of this block streams from HBM while the current one is solved.
__device__ __forceinline__ void l2_prefetch(const void* p, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// Persistent: gridDim.x blocks (about one per SM, register-limited) walk over the lines of one colour.
template <int M, int BMAX>
__global__ void __launch_bounds__(BMAX, 1) k_tdma_rb(const SweepParams P) {
  extern __shared__ double sh[];
  const int T = P.L.T;
  const long long LD = P.L.LD;
  const int t = threadIdx.x, ly = threadIdx.y, LPB = blockDim.y;
  const int nth = LPB * T;
  // shared arrays: two PCR buffers of (l,u,r)
  double* bufA = sh;
  double* bufB = sh + 3 * nth;
  const int sidx = ly * T + t;
  const bool zeroE = P.first && (P.first_colour == 0 || P.i_step == 1);
  const bool zeroW = P.first && (P.first_colour == 0) && P.i_step != 1;
  const int j0base = t * M;
  const unsigned line_bytes = (unsigned)(LD * sizeof(double));

  for (int base = blockIdx.x * LPB; base < P.nlines; base += gridDim.x * LPB) {
    const int line = base + ly;
    const bool valid = line < P.nlines;
    const int i0 = P.i_first + (valid ? line : 0) * P.i_step;
    const long long o = (long long)i0 * LD;
    // ---- L2 prefetch of the line this thread group will solve next
    {
      const int nline = line + gridDim.x * LPB;
      if (nline < P.nlines && t < 8) {
        const long long ni0 = P.i_first + nline * P.i_step;
        const double* src = t < 6 ? P.an + ni0 * P.L.cSL + (long long)t * LD
                                  : (t == 6 ? P.phi + (ni0 + 1) * LD : P.phi + (ni0 - 1) * LD);
        if (!(P.first && t >= 6 && (zeroE || zeroW)) && !(t < 6 && t >= P.L.nc)) l2_prefetch(src, line_bytes);
      }
    }
    const long long oc = (long long)i0 * P.L.cSL;
    const double* an = P.an + oc;
    const double* as = P.as + oc;
    const double* ap = P.ap + oc;
    const double* ae = P.ae + oc;
    const double* aw = P.aw + oc;
    const double* su = P.su + oc;
    const double* phi = P.phi + o;
    double* xout = P.out + o;
    const double* phiE = phi + LD;
    const double* phiW = phi - LD;

    double ca[M], cc[M], cf[M];
    // ---- load + forward elimination inside the chunk (spike cf couples to the previous chunk's last node)
    {
      double a[M], b[M], d[M];
#pragma unroll
      for (int k = 0; k < M; ++k) {
        const int j0 = j0base + k, q = k * T + t;
        const bool act = valid && j0 >= P.jlo && j0 <= P.jhi;
        if (act) {
          const int cq = k * P.L.cSK + t;
          a[k] = __ldg(an + cq);
          b[k] = __ldg(as + cq);
          d[k] = (P.L.nc == 6) ? __ldg(ap + cq) : 1.0;
          const double e = zeroE ? 0.0 : __ldg(phiE + q);
          const double w = zeroW ? 0.0 : __ldg(phiW + q);
          cc[k] = P.basis ? 0.0 : fma(__ldg(ae + cq), e, fma(__ldg(aw + cq), w, __ldg(su + cq)));
        } else {
          a[k] = 0.0;
          b[k] = 0.0;
          d[k] = 1.0;
          cc[k] = (valid && j0 < P.L.nj && !(P.first)) ? __ldg(phi + q) : 0.0;
          if (j0 == P.jlo - 1 && P.end_s_const) cc[k] = P.end_s;
          if (j0 == P.jhi + 1 && P.end_n_const) cc[k] = P.end_n;
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
          xout[q] = xk;
        }
      }
    }
    __syncthreads();   // the shared buffers are reused by the next line
  }
}

// ---------------------------------------------------------------------------------------------------
__host__ __device__ constexpr bool tdma_cf_in_smem(int M, int BMAX) {
  return (BMAX == 672 && M > 8) || (BMAX == 1024 && M > 5);
}

// TMA-staged line solver (production path for long lines).
//
// Same algorithm as k_tdma_rb, but the operands never pass through load registers: an extra producer warp
// streams the line as M "slices" (node k of every thread: one contiguous 6*T block of the interleaved
// coefficients plus the two neighbour-phi slices) into a shared-memory ring with cp.async.bulk (TMA) and
// mbarriers, running ahead of the T consumer threads by up to S slices -- also across line boundaries, so HBM
// keeps streaming while the consumers are in the backward / cyclic-reduction / store phases of the previous
// line.  One persistent block per SM.
// ---------------------------------------------------------------------------------------------------
template <int M, int BMAX, int NC>
__global__ void __launch_bounds__(BMAX, 1) k_tdma_tma(const SweepParams P, const int S) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int T = P.L.T;
  const long long LD = P.L.LD;
  const int tid = threadIdx.x;
  // carve shared memory: [ring S*(nc+2)*T doubles][pcr 6*T doubles][full S][empty S]
  constexpr int nc = NC;                      // coefficient arrays per node (5 normalised / 6 raw)
  const int SW = (nc + 2) * T;                // doubles per ring stage
  // Long lines leave too few registers per thread for the three per-node arrays (ca, cc, cf): the spike array
  // cf then lives in shared memory (conflict-free: node k of thread t at k*T + t) instead of spilling to local.
  constexpr bool CFS = tdma_cf_in_smem(M, BMAX);
  double* ring = reinterpret_cast<double*>(smem_raw);
  double* bufA = ring + (size_t)S * SW;
  double* bufB = bufA + 3 * T;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(bufB + 3 * T);
  double* cfs = reinterpret_cast<double*>(bars + 2 * S);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S);
  const int ncw = T / 32;   // consumer warps
  const bool zeroE = P.first && (P.first_colour == 0 || P.i_step == 1);
  const bool zeroW = P.first && (P.first_colour == 0) && P.i_step != 1;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, ncw);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  if (tid >= T) {
    // ---------------- producer warp: one elected thread issues the bulk copies -------------------------
    if (tid == T) {
      const unsigned slice_bytes = (unsigned)(T * sizeof(double));
      // spike vectors need the matrix only: an, as (and ap) are the first arrays of every slice
      const unsigned nmat = nc == 6 ? 3u : 2u;
      const unsigned tx = P.basis ? slice_bytes * nmat
                                  : slice_bytes * ((unsigned)nc + (zeroE ? 0u : 1u) + (zeroW ? 0u : 1u));
      unsigned rs = 0, rph = 0;   // ring slot and its phase bit, advanced together (no div/mod per slice)
      for (int line = blockIdx.x; line < P.nlines; line += gridDim.x) {
        const long long i0 = P.i_first + (long long)line * P.i_step;
        const double* csrc = P.an + i0 * P.L.cSL;       // an is array 0 of the interleaved block
        const double* esrc = P.phi + (i0 + 1) * LD;
        const double* wsrc = P.phi + (i0 - 1) * LD;
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          mbar_expect_tx(fb, tx);
          if (P.basis) {
            tma_load_1d(dst, csrc + (size_t)k * P.L.cSK, nmat * slice_bytes, fb);
          } else {
            tma_load_1d(dst, csrc + (size_t)k * P.L.cSK, nc * slice_bytes, fb);
            if (!zeroE) tma_load_1d(dst + nc * slice_bytes, esrc + (size_t)k * T, slice_bytes, fb);
            if (!zeroW) tma_load_1d(dst + (nc + 1) * slice_bytes, wsrc + (size_t)k * T, slice_bytes, fb);
          }
        }
      }
    }
    return;
  }

  // ---------------- consumers ---------------------------------------------------------------------------
  const int t = tid, lane = tid & 31;
  const int j0base = t * M;
  const unsigned T8 = (unsigned)T * 8u, stage_bytes = (unsigned)SW * 8u, ring32 = smem_u32(ring) + (unsigned)t * 8u;
  const int q_s = lpos(P.jlo - 1, M, T), q_n = lpos(P.jhi + 1, M, T);
  unsigned rs = 0, rph = 0;   // ring slot and its phase bit, advanced together (no div/mod per slice)
  double bS = 0.0, bN = 0.0;   // Dirichlet end values phi(i, jstart-1), phi(i, nj) of the current line
  if (!P.first && (int)blockIdx.x < P.nlines) {
    const double* ph = P.phi + (P.i_first + (long long)blockIdx.x * P.i_step) * LD;
    bS = __ldg(ph + q_s);
    bN = __ldg(ph + q_n);
  }
  if (P.end_s_const) bS = P.end_s;
  if (P.end_n_const) bN = P.end_n;
  for (int line = blockIdx.x; line < P.nlines; line += gridDim.x) {
    const long long i0 = P.i_first + (long long)line * P.i_step;
    double* phi = P.out + i0 * LD;
    // end values of the NEXT line, fetched a whole line ahead
    double bS_next = 0.0, bN_next = 0.0;
    if (!P.first && line + (int)gridDim.x < P.nlines) {
      const double* ph = P.phi + (i0 + (long long)gridDim.x * P.i_step) * LD;
      bS_next = __ldg(ph + q_s);
      bN_next = __ldg(ph + q_n);
    }
    if (P.end_s_const) bS_next = P.end_s;
    if (P.end_n_const) bN_next = P.end_n;
    double ca[M], cc[M], cf[CFS ? 1 : M];
    double* cft = cfs + t;
#define CF_GET(k) (CFS ? cft[(k) * T] : cf[CFS ? 0 : (k)])
#define CF_SET(k, v)            \
  {                             \
    if (CFS) {                  \
      cft[(k) * T] = (v);       \
    } else {                    \
      cf[CFS ? 0 : (k)] = (v);  \
    }                           \
  }
#pragma unroll
    for (int k = 0; k < M; ++k) {
      const unsigned s = rs, ph = rph;
      if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
      mbar_wait(full0 + 8 * s, ph);
      // slice layout: an, as, [ap], ae, aw, su, phiE, phiW -- T doubles each
      const unsigned sl = ring32 + s * stage_bytes;
      double a = lds_f64(sl), b = lds_f64(sl + T8), d = 1.0;
      unsigned sr = sl + 2 * T8;
      if (nc == 6) {
        d = lds_f64(sr);
        sr += T8;
      }
      double c = 0.0;
      if (!P.basis) {
        c = lds_f64(sr + 2 * T8);
        if (!zeroW) c = fma(lds_f64(sr + T8), lds_f64(sr + 4 * T8), c);
        if (!zeroE) c = fma(lds_f64(sr), lds_f64(sr + 3 * T8), c);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      const int j0 = j0base + k;
      if (!(j0 >= P.jlo && j0 <= P.jhi)) {
        a = 0.0;
        b = 0.0;
        d = 1.0;
        c = (j0 == P.jlo - 1) ? bS : ((j0 == P.jhi + 1) ? bN : 0.0);
      }
      if (k == 0) {
        const double inv = rcp(d);
        ca[0] = a * inv;
        cc[0] = c * inv;
        CF_SET(0, b * inv)
      } else {
        const double inv = rcp(fma(-b, ca[k > 0 ? k - 1 : 0], d));
        ca[k] = a * inv;
        cc[k] = fma(b, cc[k > 0 ? k - 1 : 0], c) * inv;
        CF_SET(k, b * CF_GET(k > 0 ? k - 1 : 0) * inv)
      }
    }
#pragma unroll
    for (int k = M - 3; k >= 0; --k) {
      cc[k] = fma(ca[k], cc[k + 1], cc[k]);
      CF_SET(k, fma(ca[k], CF_GET(k + 1), CF_GET(k)))
      ca[k] = ca[k] * ca[k + 1];
    }
    double l, u, r;
    if (M == 1) {
      l = CF_GET(0);
      u = ca[0];
      r = cc[0];
    } else {
      bufB[t] = cc[0];
      bufB[T + t] = CF_GET(0);
      bufB[2 * T + t] = ca[0];
      consumer_sync(T);
      double ncc = 0.0, ncf = 0.0, ng = 0.0;
      if (t + 1 < T) {
        ncc = bufB[t + 1];
        ncf = bufB[T + t + 1];
        ng = bufB[2 * T + t + 1];
      }
      const double inv = rcp(fma(-ca[M - 1], ncf, 1.0));
      l = CF_GET(M - 1) * inv;
      u = ca[M - 1] * ng * inv;
      r = fma(ca[M - 1], ncc, cc[M - 1]) * inv;
    }
    double* cur = bufA;
    double* nxt = bufB;
    cur[t] = l;
    cur[T + t] = u;
    cur[2 * T + t] = r;
    consumer_sync(T);
    for (int sft = 1; sft < T; sft <<= 1) {
      double lm = 0.0, um = 0.0, rm = 0.0, lp = 0.0, up = 0.0, rp = 0.0;
      if (t - sft >= 0) {
        lm = cur[t - sft];
        um = cur[T + t - sft];
        rm = cur[2 * T + t - sft];
      }
      if (t + sft < T) {
        lp = cur[t + sft];
        up = cur[T + t + sft];
        rp = cur[2 * T + t + sft];
      }
      const double inv = rcp(fma(-l, um, fma(-u, lp, 1.0)));
      r = fma(l, rm, fma(u, rp, r)) * inv;
      l = l * lm * inv;
      u = u * up * inv;
      nxt[t] = l;
      nxt[T + t] = u;
      nxt[2 * T + t] = r;
      consumer_sync(T);
      double* tmp = cur;
      cur = nxt;
      nxt = tmp;
    }
    const double xe = r;
    const double xp = (t > 0) ? cur[2 * T + t - 1] : 0.0;
#pragma unroll
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double xk = (k == M - 1) ? xe : fma(CF_GET(k), xp, fma(ca[k], xe, cc[k]));
        phi[k * T + t] = xk;
      }
    }
    bS = bS_next;
    bN = bN_next;
    consumer_sync(T);   // PCR buffers are reused by the next line
  }
#undef CF_GET
#undef CF_SET
}

// ---------------------------------------------------------------------------------------------------
// Sequential-order debug sweep (exactly cavity_fvm.f90:62-81).
// ---------------------------------------------------------------------------------------------------
struct SeqParams {
  Layout L;
  int ia, ib;        // 0-based inclusive local line range, swept in increasing order
  int jlo, jhi;
  int first;         // phi(i+1,:) taken as zero (pp(:,:) = 0 folded in)
  int two_div;       // 1: the `tdma` of the scalar-transport demos (fvm2d_stag_point.f90:232-260): rhs summed as
                     //    su + aw*phiW + ae*phiE and two divisions per row instead of one reciprocal
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
        const long long cq = (long long)i0 * P.L.cSL + (long long)(j0 % M) * P.L.cSK + (j0 / M);
        sa[j0] = P.an[cq];
        sb[j0] = P.as[cq];
        sd[j0] = (P.L.nc == 6) ? P.ap[cq] : 1.0;
        sc[j0] = P.two_div ? __dadd_rn(__dadd_rn(P.su[cq], __dmul_rn(P.aw[cq], w)), __dmul_rn(P.ae[cq], e))
                           : __dadd_rn(__dadd_rn(__dmul_rn(P.ae[cq], e), __dmul_rn(P.aw[cq], w)), P.su[cq]);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double ap_ = 0.0, cp_ = sx[P.jlo - 1];
      for (int j0 = P.jlo; j0 <= P.jhi; ++j0) {
        const double b = sb[j0];
        if (P.two_div) {
          const double denom = __dsub_rn(sd[j0], __dmul_rn(b, ap_));
          ap_ = __ddiv_rn(sa[j0], denom);
          cp_ = __ddiv_rn(__dadd_rn(sc[j0], __dmul_rn(b, cp_)), denom);
        } else {
          const double term = __ddiv_rn(1.0, __dsub_rn(sd[j0], __dmul_rn(b, ap_)));
          ap_ = __dmul_rn(sa[j0], term);
          cp_ = __dmul_rn(__dadd_rn(sc[j0], __dmul_rn(b, cp_)), term);
        }
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
static int launch_tma_m(cfd_ctx* c, const SweepParams& P, int nblocks, int S, size_t smem) {
  const int nthreads = P.L.T + 32;
  cudaError_t e = cudaSuccess;
#define LAUNCH_TDMA(BM, NCV)                                                                                   \
  {                                                                                                           \
    e = cfd_raise_smem(c, k_tdma_tma<M, BM, NCV>);                                                            \
    if (e == cudaSuccess) k_tdma_tma<M, BM, NCV><<<nblocks, nthreads, smem, c->stream>>>(P, S);                \
  }
  const bool five = P.L.nc == 5;
  if (nthreads <= 352) {
    if (five) LAUNCH_TDMA(352, 5) else LAUNCH_TDMA(352, 6)
  } else if (nthreads <= 672) {
    if (five) LAUNCH_TDMA(672, 5) else LAUNCH_TDMA(672, 6)
  } else {
    if (five) LAUNCH_TDMA(1024, 5) else LAUNCH_TDMA(1024, 6)
  }
#undef LAUNCH_TDMA
  ++c->launches;
  return e == cudaSuccess ? CFD_OK : CFD_ERR_CUDA;
}

template <int M>
static int launch_rb_m(cfd_ctx* c, const SweepParams& P, int nblocks, dim3 block, size_t smem) {
  const int nthreads = block.x * block.y;
  if (nthreads <= 320) {
    k_tdma_rb<M, 320><<<nblocks, block, smem, c->stream>>>(P);
    ++c->launches;
  } else if (nthreads <= 384) {
    k_tdma_rb<M, 384><<<nblocks, block, smem, c->stream>>>(P);
    ++c->launches;
  } else if (nthreads <= 640) {
    k_tdma_rb<M, 640><<<nblocks, block, smem, c->stream>>>(P);
    ++c->launches;
  } else {
    k_tdma_rb<M, 1024><<<nblocks, block, smem, c->stream>>>(P);
    ++c->launches;
  }
  return CFD_OK;
}

}  // namespace cfd

using namespace cfd;

// Local 0-based node range of the GLOBAL 0-based range [gj_lo, gj_hi], clamped to the rows this context owns.
void cfd_local_j_range(const cfd_ctx* c, int gj_lo, int gj_hi, int* jlo, int* jhi) {
  int a = gj_lo - c->cfg.joff, b = gj_hi - c->cfg.joff;
  const int lo_own = c->cfg.halo_s, hi_own = c->L.nj - 1 - c->cfg.halo_n;
  if (a < lo_own) a = lo_own;
  if (b > hi_own) b = hi_own;
  *jlo = a;
  *jhi = b;
}

static int launch_sweep(cfd_ctx* c, cfd::SweepParams& P);
int cfd_launch_sweep_rb_part(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first, int part);

int cfd_launch_sweep_rb(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first) {
  return cfd_launch_sweep_rb_part(c, istart, jstart, phi, colour, first, 0);
}

// part 0: every line of the colour; 1: only the lines a neighbouring i-slab needs (first / last owned line, if they
// have this colour and there is a neighbour on that side); 2: the others.  Parts 1 and 2 let the driver send the edge
// lines while the interior is still being solved.
int cfd_launch_sweep_rb_part(cfd_ctx* c, int istart, int jstart, double* phi, int colour, int first, int part) {
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  // owned lines, global 1-based [glo, ghi], restricted to istart..NI-1
  const int own_lo = c->cfg.goff + c->cfg.halo_w + 1, own_hi = c->cfg.goff + c->L.ni - c->cfg.halo_e;
  int glo = own_lo, ghi = own_hi;
  if (glo < istart) glo = istart;
  if (ghi > NI - 1) ghi = NI - 1;
  // first line of this colour: (i - istart) % 2 == colour
  const int g0 = glo + ((((glo - istart) & 1) != colour) ? 1 : 0);
  if (g0 > ghi) return CFD_OK;
  const int gl = g0 + 2 * ((ghi - g0) / 2);                 // last line of this colour
  const bool west_edge = c->cfg.halo_w > 0 && g0 == own_lo, east_edge = c->cfg.halo_e > 0 && gl == own_hi;
  int seg_first[2], seg_n[2], nseg = 0;
  if (part == 0) {
    seg_first[0] = g0; seg_n[0] = (gl - g0) / 2 + 1; nseg = 1;
  } else if (part == 1) {
    if (west_edge) { seg_first[nseg] = g0; seg_n[nseg] = 1; ++nseg; }
    if (east_edge && !(west_edge && gl == g0)) { seg_first[nseg] = gl; seg_n[nseg] = 1; ++nseg; }
  } else {
    const int a = g0 + (west_edge ? 2 : 0), b = gl - (east_edge ? 2 : 0);
    if (a <= b) { seg_first[0] = a; seg_n[0] = (b - a) / 2 + 1; nseg = 1; }
  }
  for (int sgm = 0; sgm < nseg; ++sgm) {
    SweepParams P{};
    P.L = c->L;
    P.i_first = seg_first[sgm] - 1 - c->cfg.goff;
    P.i_step = 2;
    P.nlines = seg_n[sgm];
    cfd_local_j_range(c, jstart - 1, NJ - 2, &P.jlo, &P.jhi);
    if (P.jhi < P.jlo) return CFD_OK;
    P.first = first ? 1 : 0;
    P.first_colour = colour;
    P.ae = c->field[CFD_AE];
    P.aw = c->field[CFD_AW];
    P.an = c->field[CFD_AN];
    P.as = c->field[CFD_AS];
    P.ap = c->field[CFD_AP];
    P.su = c->field[CFD_SU];
    P.phi = phi;
    P.out = phi;
    // a j-slab solves its piece of every line with zero values at the cut ends (the partition method's particular
    // solution, jslab_kernels.cu); physical ends keep the stored Dirichlet values
    P.end_s_const = c->cfg.halo_s > 0;
    P.end_n_const = c->cfg.halo_n > 0;
    int rc = launch_sweep(c, P);
    if (rc) return rc;
  }
  return CFD_OK;
}

// Spike vectors of the partition method for the current coefficients: for EVERY line i >= istart the solutions of
// the local tridiagonal piece with zero right-hand side and end values (1,0) -> spike_w, (0,1) -> spike_z.
int cfd_launch_spike_basis(cfd_ctx* c, int istart, int jstart) {
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  int glo = c->cfg.goff + c->cfg.halo_w + 1, ghi = c->cfg.goff + c->L.ni - c->cfg.halo_e;
  if (glo < istart) glo = istart;
  if (ghi > NI - 1) ghi = NI - 1;
  if (glo > ghi) return CFD_OK;
  const size_t fbytes = (size_t)c->L.ni * c->L.LD * sizeof(double);
  if (!c->spike_w) {
    CFD_CUDA(cudaMalloc(&c->spike_w, fbytes));
    CFD_CUDA(cudaMalloc(&c->spike_z, fbytes));
    CFD_CUDA(cudaMemsetAsync(c->spike_w, 0, fbytes, c->stream));
    CFD_CUDA(cudaMemsetAsync(c->spike_z, 0, fbytes, c->stream));
  }
  for (int which = 0; which < 2; ++which) {
    SweepParams P{};
    P.L = c->L;
    P.i_first = glo - 1 - c->cfg.goff;
    P.i_step = 1;
    P.nlines = ghi - glo + 1;
    cfd_local_j_range(c, jstart - 1, NJ - 2, &P.jlo, &P.jhi);
    if (P.jhi < P.jlo) return CFD_OK;
    P.first = 0;
    P.first_colour = 0;
    P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS];
    P.ap = c->field[CFD_AP]; P.su = c->field[CFD_SU];
    P.phi = c->field[CFD_PP];          // never read for values that matter (rhs dropped, both ends constant)
    P.out = which == 0 ? c->spike_w : c->spike_z;
    P.basis = 1;
    P.end_s_const = P.end_n_const = 1;
    P.end_s = which == 0 ? 1.0 : 0.0;
    P.end_n = which == 0 ? 0.0 : 1.0;
    int rc = launch_sweep(c, P);
    if (rc) return rc;
  }
  return CFD_OK;
}

static int launch_sweep(cfd_ctx* c, cfd::SweepParams& P) {
  const int T = c->L.T;
  // lines longer than one block's solver takes are split over a cluster of blocks (tdma_cluster_kernels.cu)
  if (T + 32 > 1024 || (T > 640 && cfd_cluster_size(T) > 0 && c->L.M >= 4 && !P.basis && !P.end_s_const && !P.end_n_const))
    return cfd_launch_flow_from_sweep(c, P, 0);
  const int num_sms = c->num_sms;
  static int use_tma = -1;   // environment switch, the same for every device
  if (use_tma < 0) {
    const char* e = getenv("CFD_B200_TDMA_TMA");
    use_tma = e ? atoi(e) : 1;
  }
  if (use_tma && T >= 128 && T + 32 <= 1024) {
    // TMA-staged persistent kernel: ring of S slices (8*T doubles each), S a power of two
    const size_t stage = (size_t)(c->L.nc + 2) * T * sizeof(double), pcr = (size_t)6 * T * sizeof(double);
    const int bmax = (T + 32 <= 352) ? 352 : (T + 32 <= 672 ? 672 : 1024);
    const size_t cfs = tdma_cf_in_smem(c->L.M, bmax) ? (size_t)c->L.M * T * sizeof(double) : 0;
    int S = 16;
    const char* es = getenv("CFD_B200_TDMA_STAGES");
    if (es && atoi(es) > 0) S = atoi(es);
    while (S > 1 && S * stage + pcr + cfs + 2 * S * 8 > 220 * 1024) --S;
    const size_t smem = S * stage + pcr + cfs + 2 * (size_t)S * 8;
    const int nblocks = P.nlines < num_sms ? P.nlines : num_sms;
    int rc = CFD_OK;
    switch (c->L.M) {
#define CFD_CASE(m) \
  case m:           \
    rc = launch_tma_m<m>(c, P, nblocks, S, smem); \
    break;
      CFD_CASE(1) CFD_CASE(2) CFD_CASE(3) CFD_CASE(4) CFD_CASE(5) CFD_CASE(6) CFD_CASE(7) CFD_CASE(8)
      CFD_CASE(9) CFD_CASE(10) CFD_CASE(11) CFD_CASE(12) CFD_CASE(13) CFD_CASE(14) CFD_CASE(15) CFD_CASE(16)
#undef CFD_CASE
      default:
        cfd_set_error("unsupported M=%d", c->L.M);
        return CFD_ERR_INVALID;
    }
    if (rc) {
      cfd_set_error("k_tdma_tma: cudaFuncSetAttribute(%zu bytes of shared memory) failed", smem);
      return rc;
    }
    CFD_CUDA(cudaGetLastError());
    return CFD_OK;
  }
  int lpb = 256 / T;
  if (lpb < 1) lpb = 1;
  dim3 block(T, lpb);
  int nblocks = (P.nlines + lpb - 1) / lpb;
  {
    // persistent grid: one block per SM (the chunk arrays leave room for only one), each walking over lines
    int per_sm = (T * lpb <= 256) ? 2 : 1;
    const char* e = getenv("CFD_B200_TDMA_BLOCKS_PER_SM");
    if (e && atoi(e) > 0) per_sm = atoi(e);
    if (nblocks > num_sms * per_sm) nblocks = num_sms * per_sm;
  }
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

int cfd_launch_sweep_seq2(cfd_ctx* c, int istart, int jstart, double* phi, int first, int two_div);
int cfd_launch_sweep_seq(cfd_ctx* c, int istart, int jstart, double* phi, int first) {
  return cfd_launch_sweep_seq2(c, istart, jstart, phi, first, 0);
}
int cfd_launch_sweep_seq2(cfd_ctx* c, int istart, int jstart, double* phi, int first, int two_div) {
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
  P.two_div = two_div;
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
  CFD_CUDA(cfd_raise_smem(c, k_tdma_seq));
  k_tdma_seq<<<1, 1024, smem, c->stream>>>(P);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
