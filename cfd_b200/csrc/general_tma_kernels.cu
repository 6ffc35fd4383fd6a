// general_tma_kernels.cu -- TMA-staged production assembly of the general family's equations
// (calcu buoyancy-driven-cavity-2d/buoyancy_cav.f90:122-198, calcv :200-277, calct :358-428; same arithmetic and
// the same reference quirks as general_fast_kernels.cu, which remains the path for short lines and j-slabs).
//
// The register-window kernels issue the loads of one node row per loop trip and wait for them: ~70 KB per SM in
// flight, 40 % of the HBM roofline at 8192^2.  Here, as in cavity_tma_kernels.cu, one persistent block per SM owns
// whole S-N lines; producer warps stream, per node slice k of the line, every array slice the stencil touches into
// a shared-memory ring with 1-D bulk copies (TMA) completing on mbarriers, slices ahead and across line ends; the
// T consumer threads read each staged value once, keep the j windows in registers (the cell of slice k-1 is
// evaluated when slice k has arrived, because its north face needs row k) and store the five normalised
// coefficients and du/dv coalesced.  The chunk-edge neighbours (two per thread and line) come from global memory,
// fetched one line ahead.  The 1-D j geometry (line layout, same for every line) is read through L1.
#include "common.cuh"
#include "general_fast.cuh"
#include "tma.cuh"

namespace cfd {

#define GT_PRODUCER_WARPS 2   // bulk-copy issue is per-thread serial: 13 copies per slice need several issuers

// block reduction over the consumer threads only (named barrier 1), thread 0 writes *out
__device__ __forceinline__ void gt_consumer_sum(double v, double* red, int T, int t, double* out) {
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

// Ring set-up shared by the kernels: S stages of NARR*TT doubles, 32 doubles of reduction scratch, 2*S mbarriers.
// A block owns TT = T/NSPLIT consecutive positions (toff ..) of every slice of its lines: long lines (T > 320) are
// split over NSPLIT blocks so that the consumer count stays within the 384-thread build (168 registers, no spills;
// the 768-thread build spilt and ran at half the speed of the register-window kernel).  The grid is a multiple of
// NSPLIT: block b works on part b % NSPLIT of lines b / NSPLIT, b / NSPLIT + gridDim.x / NSPLIT, ...
#define GT_RING_SETUP(NARR)                                                          \
  extern __shared__ __align__(128) unsigned char smem_raw[];                         \
  const int T = P.L.T, M = P.L.M, nj = P.L.nj;                                       \
  const long long LD = P.L.LD;                                                       \
  const int tid = threadIdx.x;                                                       \
  const int TT = T / NSPLIT, part = blockIdx.x % NSPLIT, toff = part * TT;           \
  const int bline = blockIdx.x / NSPLIT, lstride = gridDim.x / NSPLIT;               \
  const int SW = (NARR) * TT;                                                        \
  double* ring = reinterpret_cast<double*>(smem_raw);                                \
  double* red = ring + (size_t)S * SW;                                               \
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(red + 32);        \
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S);                \
  if (tid == 0) {                                                                    \
    for (int s = 0; s < S; ++s) {                                                    \
      mbar_init(full0 + 8 * s, 1);                                                   \
      mbar_init(empty0 + 8 * s, TT / 32);                                            \
    }                                                                                \
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");               \
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");                     \
  }                                                                                  \
  __syncthreads();                                                                   \
  const int nlines = P.ib - P.ia + 1;

// Producer loop: the variadic part declares `const double* src[NARR]` for line i0.
#define GT_PRODUCE(NARR, ...)                                                                         \
  if (tid >= TT) {                                                                                     \
    if (((tid - TT) & 31) == 0) {                                                                      \
      const int pw = (tid - TT) >> 5;                                                                  \
      const unsigned sb = (unsigned)(TT * sizeof(double));                                             \
      unsigned rs = 0, rph = 0;                                                                        \
      for (int line = bline; line < nlines; line += lstride) {                                         \
        const long long i0 = P.ia + line;                                                              \
        __VA_ARGS__                                                                                    \
        for (int k = 0; k < M; ++k) {                                                                  \
          const unsigned s = rs, ph = rph;                                                             \
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }                                              \
          mbar_wait(empty0 + 8 * s, ph ^ 1u);                                                          \
          const unsigned fb = full0 + 8 * s;                                                           \
          const unsigned dst = smem_u32(ring + (size_t)s * SW);                                        \
          if (pw == 0) mbar_expect_tx(fb, (NARR) * sb);                                                \
          _Pragma("unroll") for (int a = 0; a < (NARR); ++a)                                           \
            if ((a % GT_PRODUCER_WARPS) == pw) tma_load_1d(dst + a * sb, src[a] + (size_t)k * T + toff, sb, fb); \
        }                                                                                              \
      }                                                                                                \
    }                                                                                                  \
    return;                                                                                            \
  }

#define GT_STORE(AE, AW, AN, AS, SU, AP)   \
  const double rAP = gf_rcp(AP);           \
  P.ae[cout + cq] = (AE) * rAP;            \
  P.aw[cout + cq] = (AW) * rAP;            \
  P.an[cout + cq] = (AN) * rAP;            \
  P.as[cout + cq] = (AS) * rAP;            \
  P.su[cout + cq] = (SU) * rAP;

// =====================================================================================================
// u-momentum (buoyancy_cav.f90:135-194).
// Stage layout (T doubles each): 0 uC 1 vW 2 vC 3 dW 4 dC 5 gW 6 gC | 7 uW 8 uE 9 dWW 10 dE 11 pW 12 pC
//   arrays 0-6 are needed one row ahead (north face), 7-12 only at the cell itself (held for one slice).
// =====================================================================================================
#define GT_NU 13
__global__ void __launch_bounds__(384, 1) k_gen_assemble_u_tma(const GenFastParams P, const int S, const int NSPLIT) {
  GT_RING_SETUP(GT_NU)
  GT_PRODUCE(GT_NU,
             const double* src[GT_NU] = {P.u + i0 * LD,         P.v + (i0 - 1) * LD,   P.v + i0 * LD,
                                         P.den + (i0 - 1) * LD, P.den + i0 * LD,       P.gam + (i0 - 1) * LD,
                                         P.gam + i0 * LD,       P.u + (i0 - 1) * LD,   P.u + (i0 + 1) * LD,
                                         P.den + (i0 - 2) * LD, P.den + (i0 + 1) * LD, P.p + (i0 - 1) * LD,
                                         P.p + i0 * LD};)
  // ---- consumers
  const int t = toff + tid, lane = tid & 31;     // t: position inside a slice; tid: position inside the staged part
  const int j0base = t * M;
  const bool hasprev = j0base >= 1, hasnext = j0base + M < nj;
  const int qprev = (M - 1) * T + t - 1, qnext = t + 1;
  const int iadv = P.iadv, ni = P.L.ni;
  const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;
  // row j0base-1 of the first line (the row below the chunk), fetched ahead: u, den W/C, gam W/C
  double em_u = 0, em_dW = 0, em_dC = 0, em_gW = 0, em_gC = 0;
  const double em_r = hasprev ? __ldg(P.r + qprev) : 0.0;
  if (bline < nlines && hasprev) {
    const long long i0 = P.ia + bline;
    em_u = __ldg(P.u + i0 * LD + qprev);
    em_dW = __ldg(P.den + (i0 - 1) * LD + qprev);
    em_dC = __ldg(P.den + i0 * LD + qprev);
    em_gW = __ldg(P.gam + (i0 - 1) * LD + qprev);
    em_gC = __ldg(P.gam + i0 * LD + qprev);
  }
  // 1-D j geometry of the row that arrives next, loaded one slice ahead (it does not depend on the line; with the
  // ring taking the shared memory these loads miss L1, and waiting for them in the slice they are used in was the
  // largest stall of the first version of this kernel)
  double x_fy = __ldg(P.fy + t), x_rmn = __ldg(P.rmn + t), x_rydif = __ldg(P.r_ydif + t), x_r = __ldg(P.r + t);
  double x_arx = __ldg(P.arx + t), x_rsx = __ldg(P.r_sx + t), x_ycvr = __ldg(P.ycvr + t);
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const int i = (int)i0 + 1 + P.goff;                               // global line number (boundary tests only)
    const long long out = i0 * LD, cout = i0 * P.L.cSL;
    // row j0base+M of this line (used after the last slice) and row j0base-1 of the block's next line
    double ep_u = 0, ep_dW = 0, ep_dC = 0, ep_gW = 0, ep_gC = 0, ep_vW = 0, ep_vC = 0;
    if (hasnext) {
      ep_u = __ldg(P.u + i0 * LD + qnext);
      ep_dW = __ldg(P.den + (i0 - 1) * LD + qnext);
      ep_dC = __ldg(P.den + i0 * LD + qnext);
      ep_gW = __ldg(P.gam + (i0 - 1) * LD + qnext);
      ep_gC = __ldg(P.gam + i0 * LD + qnext);
      ep_vW = __ldg(P.v + (i0 - 1) * LD + qnext);
      ep_vC = __ldg(P.v + i0 * LD + qnext);
    }
    double nm_u = 0, nm_dW = 0, nm_dC = 0, nm_gW = 0, nm_gC = 0;
    if (line + lstride < nlines && hasprev) {
      const long long n0 = i0 + lstride;
      nm_u = __ldg(P.u + n0 * LD + qprev);
      nm_dW = __ldg(P.den + (n0 - 1) * LD + qprev);
      nm_dC = __ldg(P.den + n0 * LD + qprev);
      nm_gW = __ldg(P.gam + (n0 - 1) * LD + qprev);
      nm_gC = __ldg(P.gam + n0 * LD + qprev);
    }
    const double xcvi = __ldg(P.xcvi + i0), xcvip = __ldg(P.xcvip + i0 - 1);
    const double fx_w = __ldg(P.fx + i0 - 1), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rsew_w = __ldg(P.r_sew + i0 - 1), rsew_c = __ldg(P.r_sew + i0);
    const double sewu_i = __ldg(P.sewu + i0), rxdif_i = __ldg(P.r_xdif + i0);
    const bool west_b = (i == 3), east_b = (i == nj - 1);          // quirk: `do i = 3, njm1`, `if (i == njm1)` (:135,:150)
    // windows: _c = row of the pending cell (slice k-1), _p = the row that has just arrived (slice k)
    double u_s = 0, u_c = em_u, u_n = 0;
    double dW_c = em_dW, dC_c = em_dC, gW_c = em_gW, gC_c = em_gC, r_c = em_r;
    double dW_p = 0, dC_p = 0, gW_p = 0, gC_p = 0, vW_p = 0, vC_p = 0, r_p = 0;
    double o_uW = 0, o_uE = 0, o_dWW = 0, o_dE = 0, o_pW = 0, o_pC = 0;     // singles of the pending cell
    double g_arx = 0, g_rsx = 0, g_ycvr = 0;                                // its geometry
    double aS = 0.0;                                                        // acof + max(0, flow) of its south face
    for (int k = 0; k <= M; ++k) {
      double n_uW = 0, n_uE = 0, n_dWW = 0, n_dE = 0, n_pW = 0, n_pC = 0;
      const double h_fy = x_fy, h_rmn = x_rmn, h_rydif = x_rydif, ng_arx = x_arx, ng_rsx = x_rsx, ng_ycvr = x_ycvr;
      r_p = x_r;
      {   // geometry of the next row to arrive: slice k+1, the row above the chunk, or slice 0 of the next line
        const int qx = (k + 1 < M) ? (k + 1) * T + t : (k + 1 == M ? qnext : t);
        x_fy = __ldg(P.fy + qx); x_rmn = __ldg(P.rmn + qx); x_rydif = __ldg(P.r_ydif + qx); x_r = __ldg(P.r + qx);
        x_arx = __ldg(P.arx + qx); x_rsx = __ldg(P.r_sx + qx); x_ycvr = __ldg(P.ycvr + qx);
      }
      if (k < M) {
        const unsigned s = rs, ph = rph;
        if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
        mbar_wait(full0 + 8 * s, ph);
        const double* sl = ring + (size_t)s * SW + tid;
        u_n = sl[0];
        vW_p = sl[TT];
        vC_p = sl[2 * TT];
        dW_p = sl[3 * TT];
        dC_p = sl[4 * TT];
        gW_p = sl[5 * TT];
        gC_p = sl[6 * TT];
        n_uW = sl[7 * TT];
        n_uE = sl[8 * TT];
        n_dWW = sl[9 * TT];
        n_dE = sl[10 * TT];
        n_pW = sl[11 * TT];
        n_pC = sl[12 * TT];
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * s);
      } else {
        u_n = ep_u; vW_p = ep_vW; vC_p = ep_vC; dW_p = ep_dW; dC_p = ep_dC; gW_p = ep_gW; gC_p = ep_gC;
      }
      // face between the pending row (jh-1) and the arrived row jh
      const int jh = j0base + k;
      double flowN = 0.0, acofN = 0.0;
      if (jh >= 1 && jh < nj)
        gf_u_face_y(dC_c, dC_p, dW_c, dW_p, gC_c, gC_p, gW_c, gW_p, vC_p, vW_p, xcvi, xcvip, h_fy, h_rmn, h_rydif, r_c,
                    r_p, jh == 1, jh == nj - 1, iadv, flowN, acofN);
      if (k >= 1) {
        const int kk = k - 1, j0 = jh - 1;
        if (j0 >= P.jlo && j0 <= P.jhi) {
          const int q = kk * T + t, cq = kk * P.L.cSK + t;
          const double arx = g_arx, rsx = g_rsx;
          const double F_c = u_c * gf_arim(dW_c, dC_c, fx_c);
          double flow_w, diff_w, flow_e, diff_e;
          if (west_b) {
            flow_w = arx * o_uW * o_dWW;                          // u(2,j)*den(1,j), gam(1,j) (:135-142)
            diff_w = arx * __ldg(P.gam + q) * (rsew_w * rsx);
          } else {
            flow_w = arx * 0.5 * (o_uW * gf_arim(o_dWW, dW_c, fx_w) + F_c);
            diff_w = arx * gW_c * (rsew_w * rsx);
          }
          if (east_b) {
            flow_e = arx * o_uE * o_dE;                           // u(ni,j)*den(ni,j)
            diff_e = arx * __ldg(P.gam + (long long)(ni - 1) * LD + q) * (rsew_c * rsx);
          } else {
            flow_e = arx * 0.5 * (F_c + o_uE * gf_arim(dC_c, o_dE, fx_e));
            diff_e = arx * gC_c * (rsew_c * rsx);
          }
          const double AW = gf_dapec(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
          const double AE = (gf_dapec(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
          const double AS = aS;
          const double AN = (acofN + gf_max(0.0, flowN)) - flowN;
          const double vol = g_ycvr * sewu_i;
          const double DU = vol * rxdif_i * rsx;
          double SU = DU * (o_pW - o_pC);
          double AP = AE + AW + AN + AS;
          const double reseq = AP * u_c - AE * o_uE - AW * o_uW - AN * u_n - AS * u_s - SU;
          racc += fabs(reseq);
          AP = AP * r_urf;
          SU = fma(omu * AP, u_c, SU);
          GT_STORE(AE, AW, AN, AS, SU, AP)
          P.dout[out + q] = DU * rAP;
        }
      }
      // the arrived row becomes the pending one
      aS = acofN + gf_max(0.0, flowN);
      u_s = u_c; u_c = u_n;
      dW_c = dW_p; dC_c = dC_p; gW_c = gW_p; gC_c = gC_p; r_c = r_p;
      o_uW = n_uW; o_uE = n_uE; o_dWW = n_dWW; o_dE = n_dE; o_pW = n_pW; o_pC = n_pC;
      g_arx = ng_arx; g_rsx = ng_rsx; g_ycvr = ng_ycvr;
    }
    em_u = nm_u; em_dW = nm_dW; em_dC = nm_dC; em_gW = nm_gW; em_gC = nm_gC;
  }
  gt_consumer_sum(racc, red, TT, tid, P.partial + blockIdx.x);
}

// =====================================================================================================
// v-momentum (buoyancy_cav.f90:212-273).
// Stage layout: 0 vC 1 dC 2 uC 3 uE 4 dW 5 dE 6 gW 7 gC 8 gE 9 pC 10 vW 11 vE | 12 vN 13 dN | 14 tC (Boussinesq)
//   vN, dN are vC, dC one row north (slice k+1): with them the cell of slice k is finished when slice k arrives and
//   nothing but the south-looking windows is carried.  In the last slice the row north belongs to the next chunk
//   (position t+1 of slice 0, not 16-byte aligned for a bulk copy): those two values come from global memory.
//   r(j)/sns(j) enters as one precombined array (GenFastParams::r_rsns); the wall rows reload what they need.
// =====================================================================================================
#define GT_NV 15
struct GtVEdge { double fw, fe, hw, he, dC, gC, v, p, t, G; };

__global__ void __launch_bounds__(384, 1) k_gen_assemble_v_tma(const GenFastParams P, const int S, const int NSPLIT) {
  GT_RING_SETUP(GT_NV)
  const bool has_t = P.t != nullptr;
  if (tid >= TT) {   // ---- producers
    if (((tid - TT) & 31) == 0) {
      const int pw = (tid - TT) >> 5;
      const unsigned sb = (unsigned)(TT * sizeof(double));
      unsigned rs = 0, rph = 0;
      for (int line = bline; line < nlines; line += lstride) {
        const long long i0 = P.ia + line;
        const double* src[GT_NV] = {P.v + i0 * LD,         P.den + i0 * LD,       P.u + i0 * LD,         P.u + (i0 + 1) * LD,
                                    P.den + (i0 - 1) * LD, P.den + (i0 + 1) * LD, P.gam + (i0 - 1) * LD, P.gam + i0 * LD,
                                    P.gam + (i0 + 1) * LD, P.p + i0 * LD,         P.v + (i0 - 1) * LD,   P.v + (i0 + 1) * LD,
                                    P.v + i0 * LD + T,     P.den + i0 * LD + T,   has_t ? P.t + i0 * LD : nullptr};
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          const bool north = k + 1 < M;
          if (pw == 0) mbar_expect_tx(fb, (12 + (north ? 2 : 0) + (has_t ? 1 : 0)) * sb);
#pragma unroll
          for (int a = 0; a < GT_NV; ++a) {
            if ((a % GT_PRODUCER_WARPS) != pw) continue;
            if ((a == 12 || a == 13) && !north) continue;
            if (a == 14 && !has_t) continue;
            tma_load_1d(dst + a * sb, src[a] + (size_t)k * T + toff, sb, fb);
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
  const int qprev2 = j0base >= 2 ? lpos(j0base - 2, M, T) : 0;
  const int iadv = P.iadv;
  const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;
  // 1-D j geometry of the rows below the chunk (the same for every line)
  const double e_arxjp = hasprev ? __ldg(P.arxjp + qprev) : 0.0, e_rrsns = hasprev ? __ldg(P.r_rsns + qprev) : 0.0;
  const double e_fvp = hasprev ? __ldg(P.fvp + qprev) : 0.0;
  const double rmn0 = __ldg(P.rmn + t), fy0 = __ldg(P.fy + t);
  // row j0base-1 of line i0: x-face ingredients, windows and the mass flux through its south scalar plane
  auto load_edge = [&](long long i0) {
    GtVEdge e{};
    if (!hasprev) return e;
    const int i = (int)i0 + 1 + P.goff;
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    const double fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double dW_q = __ldg(P.den + (i0 - 1) * LD + qprev), dE_q = __ldg(P.den + (i0 + 1) * LD + qprev);
    const double uC_q = __ldg(P.u + i0 * LD + qprev), uE_q = __ldg(P.u + (i0 + 1) * LD + qprev);
    const double gW_q = __ldg(P.gam + (i0 - 1) * LD + qprev), gE_q = __ldg(P.gam + (i0 + 1) * LD + qprev);
    e.dC = __ldg(P.den + i0 * LD + qprev);
    const double dCmm = j0base >= 2 ? __ldg(P.den + i0 * LD + qprev2) : 0.0;
    e.gC = __ldg(P.gam + i0 * LD + qprev);
    e.v = __ldg(P.v + i0 * LD + qprev);
    e.p = __ldg(P.p + i0 * LD + qprev);
    e.t = has_t ? __ldg(P.t + i0 * LD + qprev) : 0.0;
    e.fw = west_b ? uC_q * dW_q : uC_q * gf_arim(dW_q, e.dC, fx_c);
    e.fe = east_b ? uE_q * dE_q : uE_q * gf_arim(e.dC, dE_q, fx_e);
    e.hw = west_b ? gW_q : gf_harm(gW_q, e.gC, fx_c);
    e.he = east_b ? gE_q : gf_harm(e.gC, gE_q, fx_e);
    e.G = e.v * __ldg(P.rmn + qprev) * __ldg(P.sew + i0) * gf_arim(dCmm, e.dC, __ldg(P.fy + qprev));
    return e;
  };
  GtVEdge em{};
  if (bline < nlines) em = load_edge(P.ia + bline);
  // geometry one slice ahead: x_ = row of the next slice, y_ = the row north of it
  double x_rrsns = __ldg(P.r_rsns + t), x_fvp = __ldg(P.fvp + t);
  double x_arxjp = __ldg(P.arxjp + t), x_arxj = __ldg(P.arxj + t), x_rsxmn = __ldg(P.r_sxmn + t);
  double x_ycvrs = __ldg(P.ycvrs + t), x_rydif = __ldg(P.r_ydif + t);
  const int q1 = M > 1 ? T + t : t + 1;
  double y_rmn = __ldg(P.rmn + q1), y_fy = __ldg(P.fy + q1);
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const int i = (int)i0 + 1 + P.goff;
    const long long out = i0 * LD, cout = i0 * P.L.cSL;
    double ep_v = 0, ep_d = 0;
    if (hasnext) {
      ep_v = __ldg(P.v + i0 * LD + qnext);
      ep_d = __ldg(P.den + i0 * LD + qnext);
    }
    GtVEdge nm{};
    if (line + lstride < nlines) nm = load_edge(i0 + lstride);
    const double fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rxdif_c = __ldg(P.r_xdif + i0), rxdif_e = __ldg(P.r_xdif + i0 + 1);
    const double sew_i = __ldg(P.sew + i0);
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    double fw_m = em.fw, fe_m = em.fe, hw_m = em.hw, he_m = em.he;
    double dC_m = em.dC, gC_m = em.gC, v_s = em.v, p_m = em.p, t_m = em.t;
    double arxjp_m = e_arxjp, rrsns_m = e_rrsns, fvp_m = e_fvp;
    double G_m = em.G, G_c = 0.0, rmn_c = rmn0, fy_c = fy0;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k, j = j0 + 1;
      const double rrsns_c = x_rrsns, fvp_c = x_fvp, arxjp_c = x_arxjp, arxj = x_arxj;
      const double rsxmn = x_rsxmn, ycvrs = x_ycvrs, rydif = x_rydif, rmn_p = y_rmn, fy_p = y_fy;
      {
        const int kn = (k + 1 < M) ? k + 1 : 0;
        const int qx = kn * T + t, qy = (kn + 1 < M) ? (kn + 1) * T + t : t + 1;
        x_rrsns = __ldg(P.r_rsns + qx); x_fvp = __ldg(P.fvp + qx);
        x_arxjp = __ldg(P.arxjp + qx); x_arxj = __ldg(P.arxj + qx); x_rsxmn = __ldg(P.r_sxmn + qx);
        x_ycvrs = __ldg(P.ycvrs + qx); x_rydif = __ldg(P.r_ydif + qx);
        y_rmn = __ldg(P.rmn + qy); y_fy = __ldg(P.fy + qy);
      }
      const unsigned s = rs, ph = rph;
      if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
      mbar_wait(full0 + 8 * s, ph);
      const double* sl = ring + (size_t)s * SW + tid;
      const double v_c = sl[0], dC_c = sl[TT], uC_c = sl[2 * TT], uE_c = sl[3 * TT], dW_c = sl[4 * TT], dE_c = sl[5 * TT];
      const double gW_c = sl[6 * TT], gC_c = sl[7 * TT], gE_c = sl[8 * TT], p_c = sl[9 * TT];
      const double v_w = sl[10 * TT], v_e = sl[11 * TT];
      double v_n = ep_v, dC_p = ep_d;
      if (k + 1 < M) {
        v_n = sl[12 * TT];
        dC_p = sl[13 * TT];
      }
      const double t_c = has_t ? sl[14 * TT] : 0.0;
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      // mass flux through the scalar-node planes: G(j) = v(i,j)*rmn(j)*sew*ARIMEAN(den(j-1),den(j),fy(j))
      if (k == 0) G_c = v_c * rmn_c * sew_i * gf_arim(dC_m, dC_c, fy_c);
      const double G_p = v_n * rmn_p * sew_i * gf_arim(dC_c, dC_p, fy_p);
      const double fw_c = west_b ? uC_c * dW_c : uC_c * gf_arim(dW_c, dC_c, fx_c);
      const double fe_c = east_b ? uE_c * dE_c : uE_c * gf_arim(dC_c, dE_c, fx_e);
      const double hw_c = west_b ? gW_c : gf_harm(gW_c, gC_c, fx_c);
      const double he_c = east_b ? gE_c : gf_harm(gC_c, gE_c, fx_e);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int q = k * T + t, cq = k * P.L.cSK + t;
        // x-faces (:212-233): two half faces, rows j and j-1
        const double flow_w = arxj * fw_c + arxjp_m * fw_m;
        const double diff_w = (hw_c * arxj + hw_m * arxjp_m) * (rxdif_c * rsxmn);
        const double flow_e = arxj * fe_c + arxjp_m * fe_m;
        const double diff_e = (he_c * arxj + he_m * arxjp_m) * (rxdif_e * rsxmn);
        // y-faces: north face sits on scalar node j, south face on node j-1
        double flow_n, diff_n, flow_s, diff_s;
        if (j == nj - 1) {                                      // (:246-249) wall row nj
          const int qn = (k + 1 < M) ? (k + 1) * T + t : t + 1;
          const double area = __ldg(P.r + qn) * sew_i;
          flow_n = area * v_n * dC_p;
          diff_n = area * __ldg(P.gam + i0 * LD + qn) * __ldg(P.r_sns + q);
        } else {
          flow_n = gf_arim(G_c, G_p, fvp_c);
          diff_n = rrsns_c * sew_i * gC_c;
        }
        if (j == 3) {                                           // (:212-217) wall row 1; quirk: *sns(2)
          const int q0 = lpos(0, M, T);
          const double area = __ldg(P.r + q0) * sew_i;
          flow_s = area * v_s * __ldg(P.den + i0 * LD + q0);
          diff_s = area * __ldg(P.gam + i0 * LD + q0) * __ldg(P.sns + lpos(1, M, T));
        } else {
          flow_s = gf_arim(G_m, G_c, fvp_m);
          diff_s = rrsns_m * sew_i * gC_m;
        }
        const double AW = gf_dapec(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
        const double AE = (gf_dapec(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
        const double AS = gf_dapec(flow_s, diff_s, iadv) + gf_max(0.0, flow_s);
        const double AN = (gf_dapec(flow_n, diff_n, iadv) + gf_max(0.0, flow_n)) - flow_n;
        const double vol = ycvrs * sew_i;
        const double src = P.boussinesq ? P.rgbeta * (gf_arim(t_m, t_c, fy_c) - P.tref) : 0.0;
        const double DV = vol * rydif;
        double SU = fma(src, vol, DV * (p_m - p_c));
        double AP = AE + AW + AN + AS;
        const double reseq = AP * v_c - AE * v_e - AW * v_w - AN * v_n - AS * v_s - SU;
        racc += fabs(reseq);
        AP = AP * r_urf;
        SU = fma(omu * AP, v_c, SU);
        GT_STORE(AE, AW, AN, AS, SU, AP)
        P.dout[out + q] = DV * rAP;
      }
      fw_m = fw_c; fe_m = fe_c; hw_m = hw_c; he_m = he_c;
      dC_m = dC_c;
      gC_m = gC_c;
      v_s = v_c;
      p_m = p_c; t_m = t_c;
      arxjp_m = arxjp_c;
      rrsns_m = rrsns_c; fvp_m = fvp_c;
      G_m = G_c; G_c = G_p;
      rmn_c = rmn_p; fy_c = fy_p;
    }
    em = nm;
  }
  gt_consumer_sum(racc, red, TT, tid, P.partial + blockIdx.x);
}

}  // namespace cfd

using namespace cfd;

// Parts per line: the smallest split whose consumer count fits the 384-thread build; 0 = not eligible.
static int gt_nsplit(int T) {
  for (int n = 1; n <= 4; ++n)
    if (T % (32 * n) == 0 && T / n >= 128 && T / n + 32 * GT_PRODUCER_WARPS <= 384) return n;
  return 0;
}

// True when the TMA-staged kernels can take this context (long lines in one piece; i-slabs are fine).
bool cfd_general_tma_eligible(const cfd_ctx* c, int which) {
  static int use_tma = -1;
  if (use_tma < 0) {
    const char* e = getenv("CFD_B200_GEN_TMA");
    use_tma = e ? atoi(e) : 1;
  }
  return use_tma && which <= 1 && gt_nsplit(c->L.T) > 0;
}

// which: 0 u, 1 v.  Returns the number of residual partials (grid size), or a negative status.
int cfd_launch_general_assemble_tma(cfd_ctx* c, int which, const GenFastParams& P) {
  static int num_sms = 0;
  if (!num_sms) CFD_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, c->device));
  const int T = c->L.T, narr = which == 0 ? GT_NU : GT_NV, nsplit = gt_nsplit(T), TT = T / nsplit;
  const size_t stage = (size_t)narr * TT * sizeof(double);
  int S = 8;
  const char* es = getenv("CFD_B200_ASM_STAGES");
  if (es && atoi(es) > 0) S = atoi(es);
  while (S > 2 && S * stage + 32 * 8 + 2 * S * 8 > 200 * 1024) --S;
  const size_t smem = S * stage + 32 * 8 + 2 * (size_t)S * 8;
  const int nlines = P.ib - P.ia + 1;
  int nblocks = nlines * nsplit < num_sms ? nlines * nsplit : num_sms / nsplit * nsplit;
  const int nthreads = TT + 32 * GT_PRODUCER_WARPS;
  void (*kernel)(const GenFastParams, const int, const int) = which == 0 ? k_gen_assemble_u_tma : k_gen_assemble_v_tma;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e == cudaSuccess) kernel<<<nblocks, nthreads, smem, c->stream>>>(P, S, nsplit);
  if (e != cudaSuccess) {
    cfd_set_error("k_gen_assemble_*_tma: %s", cudaGetErrorString(e));
    return CFD_ERR_CUDA;
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return nblocks;
}
