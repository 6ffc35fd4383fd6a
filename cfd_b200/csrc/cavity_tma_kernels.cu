// cavity_tma_kernels.cu -- TMA-staged production assembly of the cavity family's momentum equations
// (calcu cavity_fvm.f90:104-187, calcv :205-285; same arithmetic as cavity_fast_kernels.cu).
//
// The register-window kernels of cavity_fast_kernels.cu stop at ~65 % of the HBM roofline: every load has to
// land in a register, so only ~70 KB per SM are in flight and the loop is latency-bound.  Here one persistent
// block per SM owns whole S-N lines; a producer warp streams, for every node slice k of the line, the 12-13
// array slices the stencil touches (own line and i-neighbours of u, v, p, den, vis) into a shared-memory ring
// with 1-D bulk copies (TMA) completing on mbarriers, several slices ahead and across line boundaries; the T
// consumer threads read each staged value once, keep the j-1 / j / j+1 windows in registers (the cell of slice
// k-1 is evaluated when slice k arrives) and write the five normalised coefficients and du/dv with coalesced
// stores.  The two chunk-edge neighbours per thread and line come from global memory one line ahead.
// HBM traffic is unchanged (each field line is fetched from DRAM once; the i-neighbour copies are L2 hits).
#include "common.cuh"
#include "tma.cuh"

namespace cfd {

__device__ __forceinline__ double tmax2(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double rcp_tma(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

#define SCHEME_TMA(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS) \
  if (iadv == 1) {                                                       \
    AE = fma(-0.5, ce, de);                                              \
    AW = fma(0.5, cw, dw);                                               \
    AN = fma(-0.5, cn, dn);                                              \
    AS = fma(0.5, cs, ds);                                               \
  } else if (iadv == 2) {                                                \
    AE = de + tmax2(0.0, -ce);                                           \
    AW = dw + tmax2(0.0, cw);                                            \
    AN = dn + tmax2(0.0, -cn);                                           \
    AS = ds + tmax2(0.0, cs);                                            \
  } else {                                                               \
    AN = tmax2(fabs(0.5 * cn), dn) - 0.5 * cn;                           \
    AS = tmax2(fabs(0.5 * cs), ds) + 0.5 * cs;                           \
    AE = tmax2(fabs(0.5 * ce), de) - 0.5 * ce;                           \
    AW = tmax2(fabs(0.5 * cw), dw) + 0.5 * cw;                           \
  }

// block reduction over the consumer threads only (named barrier 1), thread 0 writes partial[blockIdx.x]
__device__ __forceinline__ void consumer_sum_to(double v, double* red, int T, int t, double* out) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((t & 31) == 0) red[t >> 5] = v;
  consumer_sync(T);
  if (t < 32) {
    v = (t < T / 32) ? red[t] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (t == 0) *out = v;
  }
}

// =====================================================================================================
// u-momentum.  Stage layout (T doubles each): 0 uC 1 vW 2 vC 3 dW 4 dC 5 sW 6 sC | 7 uW 8 uE 9 dWW 10 dE 11 pW 12 pC
// =====================================================================================================
#define NU_ARR 13
#define ASM_PRODUCER_WARPS 2   // bulk-copy issue is per-thread serial: 12-13 copies per slice need several issuers
// UNI: den and vis are one value each in this context (CavParams::uniform): their six slices are not staged at all, the
// stage holds the other seven arrays back to back, and every den / vis operand below is the scalar.
__host__ __device__ constexpr bool u_is_prop(int a) { return (a >= 3 && a <= 6) || a == 9 || a == 10; }
__host__ __device__ constexpr int u_slot(int a, bool uni) {
  return !uni ? a : (a <= 2 ? a : (a == 7 ? 3 : (a == 8 ? 4 : (a == 11 ? 5 : (a == 12 ? 6 : -1)))));
}
template <int BMAX, bool UNI>
__global__ void __launch_bounds__(BMAX, 1) k_assemble_u_tma(const CavParams P, const FastGeom G, const int S, const int NSPLIT) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int T = P.L.T, M = P.L.M, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int tid = threadIdx.x;
  // a line of more than 320 layout threads is split over NSPLIT blocks of TT = T / NSPLIT consumers each: block part
  // stages and evaluates chunks toff .. toff+TT-1 of the line (the chunk-edge neighbours come from global memory anyway)
  const int TT = T / NSPLIT, toff = ((int)blockIdx.x % NSPLIT) * TT;
  const int bline = (int)blockIdx.x / NSPLIT, lstride = (int)gridDim.x / NSPLIT;
  constexpr int NA = UNI ? 7 : NU_ARR;
  const int SW = NA * TT;
  const double den0 = UNI ? P.den0 : 0.0, vis0 = UNI ? P.vis0 : 0.0;
  double* ring = reinterpret_cast<double*>(smem_raw);
  double* red = ring + (size_t)S * SW;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(red + 32);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S);
  const int ncw = TT / 32;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, ncw);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const int nlines = P.ib - P.ia + 1;

  if (tid >= TT) {
    if (((tid - TT) & 31) == 0) {   // ---- producers: one elected lane per producer warp, arrays dealt round-robin
      const int pw = (tid - TT) >> 5;
      const unsigned sb = (unsigned)(TT * sizeof(double));
      unsigned rs = 0, rph = 0;   // ring slot and its phase bit, advanced together (no div/mod per slice)
      for (int line = bline; line < nlines; line += lstride) {
        const long long i0 = P.ia + line;
        const double* src[NU_ARR] = {P.u + i0 * LD,         P.v + (i0 - 1) * LD,   P.v + i0 * LD,       P.den + (i0 - 1) * LD,
                                     P.den + i0 * LD,       P.vis + (i0 - 1) * LD, P.vis + i0 * LD,     P.u + (i0 - 1) * LD,
                                     P.u + (i0 + 1) * LD,   P.den + (i0 - 2) * LD, P.den + (i0 + 1) * LD, P.p + (i0 - 1) * LD,
                                     P.p + i0 * LD};
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          if (pw == 0) mbar_expect_tx(fb, NA * sb);
#pragma unroll
          for (int a = 0; a < NU_ARR; ++a) {
            if (UNI && u_is_prop(a)) continue;
            const int sl = u_slot(a, UNI);
            if ((sl % ASM_PRODUCER_WARPS) == pw) tma_load_1d(dst + sl * sb, src[a] + (size_t)k * T + toff, sb, fb);
          }
        }
      }
    }
    return;
  }

  // ---- consumers
  const int t = toff + tid, lane = tid & 31;
  const int j0base = t * M;
  const bool hasprev = j0base >= 1, hasnext = j0base + M < nj;
  const int qprev = (M - 1) * T + t - 1, qnext = t + 1;
  const int iadv = P.iadv;
  const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf, stale = P.stale_area;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;   // ring slot and its phase bit, advanced together (no div/mod per slice)
  // chunk-edge neighbours (node j0base-1) of the first line, fetched ahead
  double em_u = 0, em_dW = 0, em_dC = 0, em_A = 0, em_rdy = 0;
  if (bline < nlines && hasprev) {
    const long long i0 = P.ia + bline;
    em_u = __ldg(P.u + i0 * LD + qprev);
    em_dW = UNI ? den0 : __ldg(P.den + (i0 - 1) * LD + qprev);
    em_dC = UNI ? den0 : __ldg(P.den + i0 * LD + qprev);
    em_A = UNI ? vis0 + vis0 : __ldg(P.vis + (i0 - 1) * LD + qprev) + __ldg(P.vis + i0 * LD + qprev);
    em_rdy = __ldg(G.rdyL + qprev);
  }
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const long long out = i0 * LD, cout = i0 * P.L.cSL;
    // node j0base+M of this line (used after the last slice) and node j0base-1 of the next line
    double ep_u = 0, ep_dW = 0, ep_dC = 0, ep_sW = 0, ep_sC = 0, ep_vW = 0, ep_vC = 0;
    if (hasnext) {
      ep_u = __ldg(P.u + i0 * LD + qnext);
      ep_dW = UNI ? den0 : __ldg(P.den + (i0 - 1) * LD + qnext);
      ep_dC = UNI ? den0 : __ldg(P.den + i0 * LD + qnext);
      ep_sW = UNI ? vis0 : __ldg(P.vis + (i0 - 1) * LD + qnext);
      ep_sC = UNI ? vis0 : __ldg(P.vis + i0 * LD + qnext);
      ep_vW = __ldg(P.v + (i0 - 1) * LD + qnext);
      ep_vC = __ldg(P.v + i0 * LD + qnext);
    }
    double nm_u = 0, nm_dW = 0, nm_dC = 0, nm_A = 0;
    if (line + lstride < nlines && hasprev) {
      const long long n0 = i0 + lstride;
      nm_u = __ldg(P.u + n0 * LD + qprev);
      nm_dW = UNI ? den0 : __ldg(P.den + (n0 - 1) * LD + qprev);
      nm_dC = UNI ? den0 : __ldg(P.den + n0 * LD + qprev);
      nm_A = UNI ? vis0 + vis0 : __ldg(P.vis + (n0 - 1) * LD + qprev) + __ldg(P.vis + n0 * LD + qprev);
    }
    const double sewu_i = __ldg(P.sewu + i0), r_sewu_i = __ldg(G.rsewu + i0);
    const double r_dxe = __ldg(G.rdxu + i0), r_dxw = __ldg(G.rdxu + i0 - 1);
    const double r_sew_i = __ldg(G.rsew + i0), r_sew_im1 = __ldg(G.rsew + i0 - 1);
    // windows
    double u_s = em_u, u_c = 0, u_n = 0, dW_m = em_dW, dW_c = 0, dW_p = 0, dC_m = em_dC, dC_c = 0, dC_p = 0;
    double A_m = em_A, sW_c = 0, sC_c = 0, sW_p = 0, sC_p = 0, vW_c = 0, vC_c = 0, vW_p = 0, vC_p = 0;
    double rdy_m = em_rdy, rdy_c = 0;
    double o_uW = 0, o_uE = 0, o_dWW = 0, o_dE = 0, o_pW = 0, o_pC = 0;      // singles of the previous slice
    double g_sns = 0, g_rsns = 0, g_rdy = 0;                                   // geometry of the previous slice
    for (int k = 0; k <= M; ++k) {
      double n_uW = 0, n_uE = 0, n_dWW = 0, n_dE = 0, n_pW = 0, n_pC = 0, ng_sns = 0, ng_rsns = 0, ng_rdy = 0;
      if (k < M) {
        const unsigned s = rs, ph = rph;
        if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
        mbar_wait(full0 + 8 * s, ph);
        const double* sl = ring + (size_t)s * SW + tid;
        u_n = sl[0];
        vW_p = sl[TT];
        vC_p = sl[2 * TT];
        dW_p = UNI ? den0 : sl[3 * TT];
        dC_p = UNI ? den0 : sl[4 * TT];
        sW_p = UNI ? vis0 : sl[5 * TT];
        sC_p = UNI ? vis0 : sl[6 * TT];
        n_uW = sl[u_slot(7, UNI) * TT];
        n_uE = sl[u_slot(8, UNI) * TT];
        n_dWW = UNI ? den0 : sl[9 * TT];
        n_dE = UNI ? den0 : sl[10 * TT];
        n_pW = sl[u_slot(11, UNI) * TT];
        n_pC = sl[u_slot(12, UNI) * TT];
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * s);
        const int q = k * T + t;
        ng_sns = __ldg(P.snsL + q);
        ng_rsns = __ldg(G.rsnsL + q);
        ng_rdy = __ldg(G.rdyL + q);
      } else {
        u_n = ep_u; vW_p = ep_vW; vC_p = ep_vC; dW_p = ep_dW; dC_p = ep_dC; sW_p = ep_sW; sC_p = ep_sC;
      }
      if (k >= 1) {
        const int kk = k - 1, j0 = j0base + kk;
        rdy_c = g_rdy;
        const double A_c = sW_c + sC_c;
        if (j0 >= P.jlo && j0 <= P.jhi) {
          const int q = kk * T + t, cq = kk * P.L.cSK + t;
          const double sns_j = g_sns, rsns_j = g_rsns;
          const double vol = sewu_i * sns_j;
          // UNI: 0.5*(den0 + den0) is den0 exactly and 0.25*(4 vis0) is vis0 exactly, so the means drop out bit for bit
          const double gee = UNI ? den0 * o_uE : 0.5 * (dC_c + o_dE) * o_uE;
          const double gww = UNI ? den0 * o_uW : 0.5 * (o_dWW + dW_c) * o_uW;
          const double gne = UNI ? den0 * vC_p : 0.5 * (dC_c + dC_p) * vC_p;
          const double gnw = UNI ? den0 * vW_p : 0.5 * (dW_c + dW_p) * vW_p;
          const double gse = UNI ? den0 * vC_c : 0.5 * (dC_m + dC_c) * vC_c;
          const double gsw = UNI ? den0 * vW_c : 0.5 * (dW_m + dW_c) * vW_c;
          const double gp = UNI ? den0 * u_c : 0.5 * (dW_c + dC_c) * u_c;
          const double ce = 0.5 * (gp + gee) * sns_j, cw = 0.5 * (gww + gp) * sns_j;
          const double cn = 0.5 * (gnw + gne) * sewu_i, cs = 0.5 * (gsw + gse) * sewu_i;
          const double smp = ce - cw + cn - cs;
          const double cp = tmax2(smp, 0.0);
          const double vise = UNI ? vis0 : sC_c, visw = UNI ? vis0 : sW_c;
          const double visn = UNI ? vis0 : 0.25 * (A_c + (sW_p + sC_p));
          const double viss = UNI ? vis0 : 0.25 * (A_m + A_c);
          const double de = vise * sns_j * r_dxe;
          const double dw = visw * sns_j * r_dxw;
          const double dn = visn * sewu_i * rdy_c;
          const double ds = viss * sewu_i * rdy_m;
          double AE, AW, AN, AS;
          SCHEME_TMA(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
          double SU = (o_pW - o_pC) * sns_j;
          const double dudxe = (o_uE - u_c) * r_sew_i;
          const double dudxw = (u_c - o_uW) * r_sew_im1;
          const double dvdxn = (vC_p - vW_p) * r_sewu_i;
          const double dvdxs = (vC_c - vW_c) * r_sewu_i;
          SU = fma((vise * dudxe - visw * dudxw) * r_sewu_i + (visn * dvdxn - viss * dvdxs) * rsns_j, vol, SU);
          SU = fma(cp, u_c, SU);
          double AP = AE + AW + AN + AS + cp;
          const double resor = AP * u_c - AE * o_uE - AW * o_uW - AN * u_n - AS * u_s - SU;
          racc += fabs(resor);
          AP = AP * r_urf;
          SU = fma(omu * AP, u_c, SU);
          const double rAP = rcp_tma(AP);
          P.ae[cout + cq] = AE * rAP;
          P.aw[cout + cq] = AW * rAP;
          P.an[cout + cq] = AN * rAP;
          P.as[cout + cq] = AS * rAP;
          P.su[cout + cq] = SU * rAP;
          P.dout[out + q] = stale * rAP;
        }
        // shift the windows one node north
        u_s = u_c; dW_m = dW_c; dC_m = dC_c; A_m = A_c; rdy_m = rdy_c;
      }
      u_c = u_n; dW_c = dW_p; dC_c = dC_p; sW_c = sW_p; sC_c = sC_p; vW_c = vW_p; vC_c = vC_p;
      o_uW = n_uW; o_uE = n_uE; o_dWW = n_dWW; o_dE = n_dE; o_pW = n_pW; o_pC = n_pC;
      g_sns = ng_sns; g_rsns = ng_rsns; g_rdy = ng_rdy;
    }
    em_u = nm_u; em_dW = nm_dW; em_dC = nm_dC; em_A = nm_A;
  }
  consumer_sum_to(racc, red, TT, tid, P.partial + blockIdx.x);
}

// =====================================================================================================
// v-momentum.  Stage layout: 0 vC 1 dC | 2 uC 3 uE 4 dW 5 dE 6 sW 7 sC 8 sE 9 pC | 10 vW 11 vE
//   vC and dC need the node ahead (j+1): evaluated with one slice of lag; everything else looks back only.
// =====================================================================================================
#define NV_ARR 12
__host__ __device__ constexpr bool v_is_prop(int a) { return a == 1 || (a >= 4 && a <= 8); }
__host__ __device__ constexpr int v_slot(int a, bool uni) {
  return !uni ? a : (a == 0 ? 0 : (a == 2 ? 1 : (a == 3 ? 2 : (a == 9 ? 3 : (a == 10 ? 4 : (a == 11 ? 5 : -1))))));
}
template <int BMAX, bool UNI>
__global__ void __launch_bounds__(BMAX, 1) k_assemble_v_tma(const CavParams P, const FastGeom G, const int S, const int NSPLIT) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int T = P.L.T, M = P.L.M, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int tid = threadIdx.x;
  // a line of more than 320 layout threads is split over NSPLIT blocks of TT = T / NSPLIT consumers each: block part
  // stages and evaluates chunks toff .. toff+TT-1 of the line (the chunk-edge neighbours come from global memory anyway)
  const int TT = T / NSPLIT, toff = ((int)blockIdx.x % NSPLIT) * TT;
  const int bline = (int)blockIdx.x / NSPLIT, lstride = (int)gridDim.x / NSPLIT;
  constexpr int NA = UNI ? 6 : NV_ARR;
  const int SW = NA * TT;
  const double den0 = UNI ? P.den0 : 0.0, vis0 = UNI ? P.vis0 : 0.0;
  double* ring = reinterpret_cast<double*>(smem_raw);
  double* red = ring + (size_t)S * SW;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(red + 32);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S);
  const int ncw = TT / 32;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, ncw);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const int nlines = P.ib - P.ia + 1;

  if (tid >= TT) {
    if (((tid - TT) & 31) == 0) {
      const int pw = (tid - TT) >> 5;
      const unsigned sb = (unsigned)(TT * sizeof(double));
      unsigned rs = 0, rph = 0;   // ring slot and its phase bit, advanced together (no div/mod per slice)
      for (int line = bline; line < nlines; line += lstride) {
        const long long i0 = P.ia + line;
        const double* src[NV_ARR] = {P.v + i0 * LD,         P.den + i0 * LD,       P.u + i0 * LD,         P.u + (i0 + 1) * LD,
                                     P.den + (i0 - 1) * LD, P.den + (i0 + 1) * LD, P.vis + (i0 - 1) * LD, P.vis + i0 * LD,
                                     P.vis + (i0 + 1) * LD, P.p + i0 * LD,         P.v + (i0 - 1) * LD,   P.v + (i0 + 1) * LD};
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          if (pw == 0) mbar_expect_tx(fb, NA * sb);
#pragma unroll
          for (int a = 0; a < NV_ARR; ++a) {
            if (UNI && v_is_prop(a)) continue;
            const int sl = v_slot(a, UNI);
            if ((sl % ASM_PRODUCER_WARPS) == pw) tma_load_1d(dst + sl * sb, src[a] + (size_t)k * T + toff, sb, fb);
          }
        }
      }
    }
    return;
  }

  const int t = toff + tid, lane = tid & 31;
  const int j0base = t * M;
  const bool hasprev = j0base >= 1, hasnext = j0base + M < nj;
  const int qprev = (M - 1) * T + t - 1, qnext = t + 1;
  const int qprev2 = (j0base >= 2) ? lpos(j0base - 2, M, T) : 0;
  const int iadv = P.iadv;
  const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf, stale = P.stale_area;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;   // ring slot and its phase bit, advanced together (no div/mod per slice)
  // node j0base-1 (and j0base-2 of den) of a line: 12 values
  struct Edge {
    double v, dC, dCmm, uC, uE, dW, dE, sW, sC, sE, p;
  };
  auto load_edge = [&](long long i0) {
    Edge e{};
    if (hasprev) {
      e.v = __ldg(P.v + i0 * LD + qprev);
      e.dC = UNI ? den0 : __ldg(P.den + i0 * LD + qprev);
      e.dCmm = (j0base >= 2) ? (UNI ? den0 : __ldg(P.den + i0 * LD + qprev2)) : 0.0;
      e.uC = __ldg(P.u + i0 * LD + qprev);
      e.uE = __ldg(P.u + (i0 + 1) * LD + qprev);
      e.dW = UNI ? den0 : __ldg(P.den + (i0 - 1) * LD + qprev);
      e.dE = UNI ? den0 : __ldg(P.den + (i0 + 1) * LD + qprev);
      e.sW = UNI ? vis0 : __ldg(P.vis + (i0 - 1) * LD + qprev);
      e.sC = UNI ? vis0 : __ldg(P.vis + i0 * LD + qprev);
      e.sE = UNI ? vis0 : __ldg(P.vis + (i0 + 1) * LD + qprev);
      e.p = __ldg(P.p + i0 * LD + qprev);
    }
    return e;
  };
  Edge em{};
  if (bline < nlines) em = load_edge(P.ia + bline);
  const double g_rdyv_prev0 = hasprev ? __ldg(G.rdyvL + qprev) : 0.0, g_rsns_prev0 = hasprev ? __ldg(G.rsnsL + qprev) : 0.0;
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const long long out = i0 * LD, cout = i0 * P.L.cSL;
    double ep_v = 0, ep_dC = 0;
    if (hasnext) {
      ep_v = __ldg(P.v + i0 * LD + qnext);
      ep_dC = UNI ? den0 : __ldg(P.den + i0 * LD + qnext);
    }
    Edge nm{};
    if (line + lstride < nlines) nm = load_edge(i0 + lstride);
    const double sew_i = __ldg(P.sew + i0), r_sew_i = __ldg(G.rsew + i0);
    const double r_dxe = __ldg(G.rdx + i0), r_dxw = __ldg(G.rdx + i0 - 1);
    // state "at node kk-1" (m) and "at node kk" (c); vC / dC also one node ahead (n / p)
    double v_s = em.v, v_c = 0, v_n = 0, dC_mm = em.dCmm, dC_m = em.dC, dC_c = 0, dC_p = 0;
    double uC_m = em.uC, uE_m = em.uE, dW_m = em.dW, dE_m = em.dE, sC_m = em.sC;
    double SW_m = em.sW + em.sC, SE_m = em.sC + em.sE, p_m = em.p;
    double rdyv_m = g_rdyv_prev0, rsns_m = g_rsns_prev0;
    // values of slice kk (held until slice kk+1 has arrived)
    double c_uC = 0, c_uE = 0, c_dW = 0, c_dE = 0, c_sW = 0, c_sC = 0, c_sE = 0, c_p = 0, c_vW = 0, c_vE = 0;
    double c_snsv = 0, c_rsnsv = 0, c_rdyv = 0, c_rsns = 0;
    for (int k = 0; k <= M; ++k) {
      double n_uC = 0, n_uE = 0, n_dW = 0, n_dE = 0, n_sW = 0, n_sC = 0, n_sE = 0, n_p = 0, n_vW = 0, n_vE = 0;
      double n_snsv = 0, n_rsnsv = 0, n_rdyv = 0, n_rsns = 0;
      if (k < M) {
        const unsigned s = rs, ph = rph;
        if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
        mbar_wait(full0 + 8 * s, ph);
        const double* sl = ring + (size_t)s * SW + tid;
        v_n = sl[0];
        dC_p = UNI ? den0 : sl[TT];
        n_uC = sl[v_slot(2, UNI) * TT];
        n_uE = sl[v_slot(3, UNI) * TT];
        n_dW = UNI ? den0 : sl[4 * TT];
        n_dE = UNI ? den0 : sl[5 * TT];
        n_sW = UNI ? vis0 : sl[6 * TT];
        n_sC = UNI ? vis0 : sl[7 * TT];
        n_sE = UNI ? vis0 : sl[8 * TT];
        n_p = sl[v_slot(9, UNI) * TT];
        n_vW = sl[v_slot(10, UNI) * TT];
        n_vE = sl[v_slot(11, UNI) * TT];
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * s);
        const int q = k * T + t;
        n_snsv = __ldg(P.snsvL + q);
        n_rsnsv = __ldg(G.rsnsvL + q);
        n_rdyv = __ldg(G.rdyvL + q);
        n_rsns = __ldg(G.rsnsL + q);
      } else {
        v_n = ep_v;
        dC_p = ep_dC;
      }
      if (k >= 1) {
        const int kk = k - 1, j0 = j0base + kk;
        const double SW_c = c_sW + c_sC, SE_c = c_sC + c_sE;
        if (j0 >= P.jlo && j0 <= P.jhi) {
          const int q = kk * T + t, cq = kk * P.L.cSK + t;
          const double snsv_j = c_snsv, rsnsv_j = c_rsnsv;
          const double vol = sew_i * snsv_j;
          const double gne = UNI ? den0 * c_uE : 0.5 * (dC_c + c_dE) * c_uE;
          const double gnw = UNI ? den0 * c_uC : 0.5 * (c_dW + dC_c) * c_uC;
          const double gnn = UNI ? den0 * v_n : 0.5 * (dC_c + dC_p) * v_n;
          const double gse = UNI ? den0 * uE_m : 0.5 * (dC_m + dE_m) * uE_m;
          const double gsw = UNI ? den0 * uC_m : 0.5 * (dW_m + dC_m) * uC_m;
          const double gss = UNI ? den0 * v_s : 0.5 * (dC_mm + dC_m) * v_s;
          const double gp = UNI ? den0 * v_c : 0.5 * (dC_m + dC_c) * v_c;
          const double ce = 0.5 * (gne + gse) * snsv_j, cw = 0.5 * (gnw + gsw) * snsv_j;
          const double cn = 0.5 * (gp + gnn) * sew_i, cs = 0.5 * (gss + gp) * sew_i;
          const double smp = ce - cw + cn - cs;
          const double cp = tmax2(smp, 0.0);
          const double vise = UNI ? vis0 : 0.25 * (SE_m + SE_c);
          const double visw = UNI ? vis0 : 0.25 * (SW_m + SW_c);
          const double visn = UNI ? vis0 : c_sC, viss = UNI ? vis0 : sC_m;
          const double de = vise * snsv_j * r_dxe;
          const double dw = visw * snsv_j * r_dxw;
          const double dn = visn * sew_i * c_rdyv;
          const double ds = viss * sew_i * rdyv_m;
          double AE, AW, AN, AS;
          SCHEME_TMA(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
          double SU = (p_m - c_p) * sew_i;
          const double dudye = (c_uE - uE_m) * rsnsv_j;
          const double dudyw = (c_uC - uC_m) * rsnsv_j;
          const double dvdyn = (v_n - v_c) * c_rsns;
          const double dvdys = (v_c - v_s) * rsns_m;
          SU = fma((vise * dudye - visw * dudyw) * r_sew_i + (visn * dvdyn - viss * dvdys) * rsnsv_j, vol, SU);
          SU = fma(cp, v_c, SU);
          double AP = AE + AW + AN + AS + cp;
          const double resor = AP * v_c - AE * c_vE - AW * c_vW - AN * v_n - AS * v_s - SU;
          racc += fabs(resor);
          AP = AP * r_urf;
          SU = fma(omu * AP, v_c, SU);
          const double rAP = rcp_tma(AP);
          P.ae[cout + cq] = AE * rAP;
          P.aw[cout + cq] = AW * rAP;
          P.an[cout + cq] = AN * rAP;
          P.as[cout + cq] = AS * rAP;
          P.su[cout + cq] = SU * rAP;
          P.dout[out + q] = stale * rAP;
        }
        // node kk becomes "previous"
        v_s = v_c; dC_mm = dC_m; dC_m = dC_c;
        uC_m = c_uC; uE_m = c_uE; dW_m = c_dW; dE_m = c_dE; sC_m = c_sC; SW_m = SW_c; SE_m = SE_c; p_m = c_p;
        rdyv_m = c_rdyv; rsns_m = c_rsns;
      }
      v_c = v_n; dC_c = dC_p;
      c_uC = n_uC; c_uE = n_uE; c_dW = n_dW; c_dE = n_dE; c_sW = n_sW; c_sC = n_sC; c_sE = n_sE; c_p = n_p;
      c_vW = n_vW; c_vE = n_vE; c_snsv = n_snsv; c_rsnsv = n_rsnsv; c_rdyv = n_rdyv; c_rsns = n_rsns;
    }
    em = nm;
  }
  consumer_sum_to(racc, red, TT, tid, P.partial + blockIdx.x);
}

}  // namespace cfd

using namespace cfd;

// blocks per line for the TMA-staged assembly: the smallest split whose consumer count fits the 384-thread build
// (168 registers, no spills); 0 = the layout does not split that way
int cfd_assemble_tma_split(int T) {
  for (int n = 1; n <= 8; ++n)
    if (T % (32 * n) == 0 && T / n >= 128 && T / n + 32 * ASM_PRODUCER_WARPS <= 384) return n;
  return 0;
}

// returns the number of partial sums written (grid size), or a negative status
int cfd_launch_assemble_tma(cfd_ctx* c, int which, CavParams& P) {
  FastGeom G;
  G.rdxu = c->gfast[0]; G.rdx = c->gfast[1]; G.rsew = c->gfast[2]; G.rsewu = c->gfast[3];
  G.rdyL = c->gfast[4]; G.rdyvL = c->gfast[5]; G.rsnsL = c->gfast[6]; G.rsnsvL = c->gfast[7];
  const int num_sms = c->num_sms;
  const bool uni = P.uniform != 0;
  const int T = c->L.T, narr = uni ? (which == 0 ? 7 : 6) : (which == 0 ? NU_ARR : NV_ARR);
  const int nsplit = cfd_assemble_tma_split(T);
  if (!nsplit) {
    cfd_set_error("k_assemble_*_tma: lines of %d layout threads do not split into blocks of 128..320 consumers", T);
    return CFD_ERR_INVALID;
  }
  const int TT = T / nsplit;
  const size_t stage = (size_t)narr * TT * sizeof(double);
  int S = uni ? 10 : 8;      // the uniform-property stage is about half as large: a deeper ring fits
  const char* es = getenv("CFD_B200_ASM_STAGES");
  if (es && atoi(es) > 0) S = atoi(es);
  while (S > 2 && S * stage + 32 * 8 + 2 * S * 8 > 200 * 1024) --S;
  const size_t smem = S * stage + 32 * 8 + 2 * (size_t)S * 8;
  const int nlines = P.ib - P.ia + 1;
  const int nblocks = nlines * nsplit < num_sms ? nlines * nsplit : num_sms / nsplit * nsplit;
  cudaError_t e;
#define LAUNCH_ASM(KERNEL)                                                                         \
  e = cfd_raise_smem(c, KERNEL);                                                                   \
  if (e == cudaSuccess) KERNEL<<<nblocks, TT + 32 * ASM_PRODUCER_WARPS, smem, c->stream>>>(P, G, S, nsplit);
  if (which == 0) {
    if (uni) { LAUNCH_ASM((k_assemble_u_tma<384, true>)) } else { LAUNCH_ASM((k_assemble_u_tma<384, false>)) }
  } else {
    if (uni) { LAUNCH_ASM((k_assemble_v_tma<384, true>)) } else { LAUNCH_ASM((k_assemble_v_tma<384, false>)) }
  }
#undef LAUNCH_ASM
  if (e != cudaSuccess) {
    cfd_set_error("k_assemble_*_tma: %s", cudaGetErrorString(e));
    return CFD_ERR_CUDA;
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return nblocks;
}
