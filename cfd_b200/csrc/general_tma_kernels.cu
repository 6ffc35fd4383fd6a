// general_tma_kernels.cu -- TMA-staged production assembly of the general family's equations
// (calcu buoyancy-driven-cavity-2d/buoyancy_cav.f90:122-198, calcv :200-277, calcp :279-343, calct :358-428; same
// arithmetic and the same reference quirks as general_fast_kernels.cu, which remains the path for short lines and
// j-slabs).
//
// The register-window kernels issue the loads of one node row per loop trip and wait for them: ~70 KB per SM in
// flight, 40 % of the HBM roofline at 8192^2.  Here, as in cavity_tma_kernels.cu, persistent blocks own whole S-N
// lines (or an aligned part of every slice of them, see GT_RING_SETUP); producer warps stream, per node slice k of
// the line, every array slice the stencil touches into a shared-memory ring with 1-D bulk copies (TMA) completing
// on mbarriers, slices ahead and across line ends; the consumer threads read each staged value once and store the
// five normalised coefficients and du/dv coalesced.  The north face of a cell needs the row above it: the kernels
// stage that row a second time beside the row itself ("north copies", L2 hits), finish the cell of
// slice k when slice k arrives and carry only the south-looking windows (so does the u kernel, with seven copies).  The
// chunk-edge neighbours (two rows per thread and line) come from global memory, fetched one line ahead.  The 1-D
// j geometry (line layout, same for every line) is read through L2 one slice ahead.
#include "common.cuh"
#include "general_fast.cuh"
#include "tma.cuh"

namespace cfd {

#define GT_PRODUCER_WARPS 2   // bulk-copy issue is per-thread serial: 10-21 copies per slice need several issuers

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

// Producer loop.  NORTH_MASK: arrays that are another array one row north (their src pointer is already advanced by
// T; not copied in the last slice, whose row north is not 16-byte aligned in the layout -- the consumers take it
// from global memory).  ABSENT_MASK: optional arrays this launch does not have.  The variadic part declares
// `const double* src[NARR]` for line i0.  The wait on a free slot backs off with nanosleep: a spinning producer
// warp took a fifth of the block's issue slots.
#define GT_PRODUCE(NARR, NORTH_MASK, ABSENT_MASK, ...)                                                 \
  if (tid >= TT) {                                                                                     \
    if (((tid - TT) & 31) == 0) {                                                                      \
      const int pw = (tid - TT) >> 5;                                                                  \
      const unsigned sb = (unsigned)(TT * sizeof(double));                                             \
      const unsigned absent = (ABSENT_MASK);                                                           \
      const int n_all = (NARR) - __popc(absent), n_last = n_all - __popc((NORTH_MASK) & ~absent);      \
      unsigned rs = 0, rph = 0;                                                                        \
      for (int line = bline; line < nlines; line += lstride) {                                         \
        const long long i0 = P.ia + line;                                                              \
        __VA_ARGS__                                                                                    \
        for (int k = 0; k < M; ++k) {                                                                  \
          const unsigned s = rs, ph = rph;                                                             \
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }                                              \
          mbar_wait_backoff(empty0 + 8 * s, ph ^ 1u);                                                  \
          const unsigned fb = full0 + 8 * s;                                                           \
          const unsigned dst = smem_u32(ring + (size_t)s * SW);                                        \
          const bool last = k + 1 >= M;                                                                \
          const unsigned skip = absent | (last ? (unsigned)(NORTH_MASK) : 0u);                         \
          if (pw == 0) mbar_expect_tx(fb, (last ? n_last : n_all) * sb);                               \
          _Pragma("unroll") for (int a = 0; a < (NARR); ++a)                                           \
            if ((a % GT_PRODUCER_WARPS) == pw && !((skip >> a) & 1u))                                  \
              tma_load_1d(dst + a * sb, src[a] + (size_t)k * T + toff, sb, fb);                        \
        }                                                                                              \
      }                                                                                                \
    }                                                                                                  \
    return;                                                                                            \
  }

// The five arrays of a production (nc = 5) coefficient block sit T doubles apart in the order an, as, ae, aw, su
// (api.cu: kCoefOrder5), so one base pointer serves the five stores.
#define GT_STORE(AE, AW, AN, AS, SU, AP)   \
  const double rAP = gf_rcp(AP);           \
  double* cb__ = P.an + cout + cq;         \
  cb__[0] = (AN) * rAP;                    \
  cb__[T] = (AS) * rAP;                    \
  cb__[2 * T] = (AE) * rAP;                \
  cb__[3 * T] = (AW) * rAP;                \
  cb__[4 * T] = (SU) * rAP;

// =====================================================================================================
// u-momentum (buoyancy_cav.f90:135-194).
// Stage layout (TT doubles each): 0 uC 1 dW 2 dC 3 gW 4 gC | 5 uW 6 uE 7 dWW 8 dE 9 pW 10 pC |
//   north copies 11 uN 12 vWN 13 vCN 14 dWN 15 dCN 16 gWN 17 gCN
//   Only the south face (aS) and u(j-1) are carried from cell to cell.  The two faces at the chunk ends (south of
//   the first cell, north of the last) are evaluated from global memory one line ahead.
//   (A variant that staged each row once and evaluated the cell of slice k-1 when slice k had arrived was 7 %
//   faster but not reproducible from launch to launch in the one warp of a line that holds padding nodes; it was
//   dropped.)
// =====================================================================================================
#define GT_NU 18
#define GT_NU_NORTH 0x3F800u
struct GtUEdge { double aS, us, flowN, acofN, uN; };

template <int IADV>   // scheme number known at compile time (3 = hybrid, the default of every shipped case), 0 = read it from P
__global__ void __launch_bounds__(384, 1) k_gen_assemble_u_tma(const GenFastParams P, const int S, const int NSPLIT) {
  GT_RING_SETUP(GT_NU)
  GT_PRODUCE(GT_NU, GT_NU_NORTH, 0u,
             const double* src[GT_NU] = {P.u + i0 * LD,           P.den + (i0 - 1) * LD,     P.den + i0 * LD,
                                         P.gam + (i0 - 1) * LD,   P.gam + i0 * LD,           P.u + (i0 - 1) * LD,
                                         P.u + (i0 + 1) * LD,     P.den + (i0 - 2) * LD,     P.den + (i0 + 1) * LD,
                                         P.p + (i0 - 1) * LD,     P.p + i0 * LD,             P.u + i0 * LD + T,
                                         P.v + (i0 - 1) * LD + T, P.v + i0 * LD + T,         P.den + (i0 - 1) * LD + T,
                                         P.den + i0 * LD + T,     P.gam + (i0 - 1) * LD + T, P.gam + i0 * LD + T};)
  // ---- consumers
  const int t = toff + tid, lane = tid & 31;
  const int j0base = t * M;
  const bool hasprev = j0base >= 1 && j0base < nj, hasnext = j0base + M < nj;
  const int qprev = (M - 1) * T + t - 1, qnext = t + 1, qlast = (M - 1) * T + t;
  const int iadv = IADV ? IADV : P.iadv, ni = P.L.ni;
  const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;
  // faces at the two ends of this thread's chunk of line i0, and the u values beyond them
  auto load_edge = [&](long long i0) {
    GtUEdge e{};
    const double xcvi = __ldg(P.xcvi + i0), xcvip = __ldg(P.xcvip + i0 - 1);
    const double *dW = P.den + (i0 - 1) * LD, *dC = P.den + i0 * LD, *gW = P.gam + (i0 - 1) * LD, *gC = P.gam + i0 * LD;
    const double *vW = P.v + (i0 - 1) * LD, *vC = P.v + i0 * LD;
    if (hasprev) {
      double flow, acof;
      gf_u_face_y<true>(__ldg(dC + qprev), __ldg(dC + t), __ldg(dW + qprev), __ldg(dW + t), __ldg(gC + qprev), __ldg(gC + t),
                  __ldg(gW + qprev), __ldg(gW + t), __ldg(vC + t), __ldg(vW + t), xcvi, xcvip, __ldg(P.fy + t),
                  __ldg(P.rmn + t), __ldg(P.r_ydif + t), __ldg(P.r + qprev), __ldg(P.r + t), j0base == 1,
                  j0base == nj - 1, iadv, flow, acof);
      e.aS = acof + gf_max(0.0, flow);
      e.us = __ldg(P.u + i0 * LD + qprev);
    }
    if (hasnext) {
      gf_u_face_y<true>(__ldg(dC + qlast), __ldg(dC + qnext), __ldg(dW + qlast), __ldg(dW + qnext), __ldg(gC + qlast),
                  __ldg(gC + qnext), __ldg(gW + qlast), __ldg(gW + qnext), __ldg(vC + qnext), __ldg(vW + qnext), xcvi, xcvip,
                  __ldg(P.fy + qnext), __ldg(P.rmn + qnext), __ldg(P.r_ydif + qnext), __ldg(P.r + qlast), __ldg(P.r + qnext),
                  j0base + M == 1, j0base + M == nj - 1, iadv, e.flowN, e.acofN);
      e.uN = __ldg(P.u + i0 * LD + qnext);
    }
    return e;
  };
  GtUEdge em{};
  if (bline < nlines) em = load_edge(P.ia + bline);
  // 1-D j geometry one slice ahead (it does not depend on the line; with the ring taking the shared memory these
  // loads miss L1): x_ = the row of the next slice, y_ = the row north of it
  double x_arx = __ldg(P.arx + t), x_rsx = __ldg(P.r_sx + t), x_ycvr = __ldg(P.ycvr + t), x_r = __ldg(P.r + t);
  const int q1 = M > 1 ? T + t : t + 1;
  double y_fy = __ldg(P.fy + q1), y_rmn = __ldg(P.rmn + q1), y_rydif = __ldg(P.r_ydif + q1), y_r = __ldg(P.r + q1);
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const int i = (int)i0 + 1 + P.goff;                               // global line number (boundary tests only)
    const long long out = i0 * LD, cout = i0 * P.L.cSL;
    GtUEdge nm{};
    if (line + lstride < nlines) nm = load_edge(i0 + lstride);
    const double xcvi = __ldg(P.xcvi + i0), xcvip = __ldg(P.xcvip + i0 - 1);
    const double fx_w = __ldg(P.fx + i0 - 1), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rsew_w = __ldg(P.r_sew + i0 - 1), rsew_c = __ldg(P.r_sew + i0);
    const double sewu_i = __ldg(P.sewu + i0), rxdif_i = __ldg(P.r_xdif + i0);
    const bool west_b = (i == 3), east_b = (i == nj - 1);          // quirk: `do i = 3, njm1`, `if (i == njm1)` (:135,:150)
    double aS = em.aS, u_s = em.us;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const double arx = x_arx, rsx = x_rsx, ycvr = x_ycvr, r_c = x_r;
      const double fy_p = y_fy, rmn_p = y_rmn, rydif_p = y_rydif, r_p = y_r;
      {
        const int kn = (k + 1 < M) ? k + 1 : 0;
        const int qx = kn * T + t, qy = (kn + 1 < M) ? (kn + 1) * T + t : t + 1;
        x_arx = __ldg(P.arx + qx); x_rsx = __ldg(P.r_sx + qx); x_ycvr = __ldg(P.ycvr + qx); x_r = __ldg(P.r + qx);
        y_fy = __ldg(P.fy + qy); y_rmn = __ldg(P.rmn + qy); y_rydif = __ldg(P.r_ydif + qy); y_r = __ldg(P.r + qy);
      }
      const unsigned s = rs, ph = rph;
      if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
      mbar_wait(full0 + 8 * s, ph);
      const double* sl = ring + (size_t)s * SW + tid;
      const double u_c = sl[0], dW_c = sl[TT], dC_c = sl[2 * TT], gW_c = sl[3 * TT], gC_c = sl[4 * TT];
      const double u_w = sl[5 * TT], u_e = sl[6 * TT], den_ww = sl[7 * TT], den_e = sl[8 * TT];
      const double p_w = sl[9 * TT], p_c = sl[10 * TT];
      const bool north = k + 1 < M;          // the north copies of the last slice are not staged (stale values, unused)
      const double s_uN = sl[11 * TT], vW_p = sl[12 * TT], vC_p = sl[13 * TT], dW_p = sl[14 * TT], dC_p = sl[15 * TT];
      const double gW_p = sl[16 * TT], gC_p = sl[17 * TT];
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      double u_n = em.uN, flowN = em.flowN, acofN = em.acofN;
      if (north) {
        u_n = s_uN;
        flowN = 0.0;
        acofN = 0.0;
        if (j0 + 1 < nj)
          gf_u_face_y<true>(dC_c, dC_p, dW_c, dW_p, gC_c, gC_p, gW_c, gW_p, vC_p, vW_p, xcvi, xcvip, fy_p, rmn_p, rydif_p, r_c,
                      r_p, j0 == 0, j0 + 1 == nj - 1, iadv, flowN, acofN);
      }
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int q = k * T + t, cq = k * P.L.cSK + t;
        // x-faces: mass flux at the u nodes i-1, i, i+1
        const double F_c = u_c * gf_arim(dW_c, dC_c, fx_c);
        double flow_w, diff_w, flow_e, diff_e;
        if (west_b) {
          flow_w = arx * u_w * den_ww;                          // u(2,j)*den(1,j), gam(1,j) (:135-142)
          diff_w = arx * __ldg(P.gam + q) * (rsew_w * rsx);
        } else {
          flow_w = arx * 0.5 * (u_w * gf_arim(den_ww, dW_c, fx_w) + F_c);
          diff_w = arx * gW_c * (rsew_w * rsx);
        }
        if (east_b) {
          flow_e = arx * u_e * den_e;                           // u(ni,j)*den(ni,j)
          diff_e = arx * __ldg(P.gam + (long long)(ni - 1) * LD + q) * (rsew_c * rsx);
        } else {
          flow_e = arx * 0.5 * (F_c + u_e * gf_arim(dC_c, den_e, fx_e));
          diff_e = arx * gC_c * (rsew_c * rsx);
        }
        const double AW = gf_dapec_sel(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
        const double AE = (gf_dapec_sel(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
        const double AS = aS;
        const double AN = (acofN + gf_max(0.0, flowN)) - flowN;
        const double vol = ycvr * sewu_i;
        const double DU = vol * rxdif_i * rsx;
        double SU = DU * (p_w - p_c);
        double AP = AE + AW + AN + AS;
        const double reseq = AP * u_c - AE * u_e - AW * u_w - AN * u_n - AS * u_s - SU;
        racc += fabs(reseq);
        AP = AP * r_urf;
        SU = fma(omu * AP, u_c, SU);
        GT_STORE(AE, AW, AN, AS, SU, AP)
        P.dout[out + q] = DU * rAP;
      }
      aS = acofN + gf_max(0.0, flowN);
      u_s = u_c;
    }
    em = nm;
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

template <int IADV>   // scheme number known at compile time (3 = hybrid, the default of every shipped case), 0 = read it from P
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
          mbar_wait_backoff(empty0 + 8 * s, ph ^ 1u);
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
  const int iadv = IADV ? IADV : P.iadv;
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

// =====================================================================================================
// Energy equation (buoyancy_cav.f90:373-428).
// Stage layout: 0 dC 1 cC 2 gC 3 tC | 4 uC 5 uE 6 dW 7 cW 8 gW 9 gE 10 dE 11 cE 12 tW 13 tE |
//   north copies 14 dN 15 cN 16 gN 17 tN 18 vN | optional 19 su_src 20 sp_src
// =====================================================================================================
#define GT_NT 21
#define GT_NT_NORTH 0x7C000u
struct GtTEdge { double aS, ts, flowN, acofN, tN; };

// y-face between rows lo and hi of the scalar column: D = den*cp, g = gamt; v and the geometry belong to row hi
__device__ __forceinline__ void gt_t_face(double D_lo, double D_hi, double g_lo, double g_hi, double v_hi, double sew_i,
                                          double r_lo, double rmn_hi, double fy_hi, double rydif_hi, bool south_b,
                                          bool north_b, int iadv, double& flow, double& acof) {
  double diff;
  if (south_b) {            // lo is the wall row 1 (:373-378)
    const double areas = r_lo * sew_i;
    flow = D_lo * v_hi * areas;
    diff = g_lo * areas * rydif_hi;
  } else {
    const double arean = rmn_hi * sew_i;
    if (north_b) {          // hi is the wall row nj
      flow = D_hi * v_hi * arean;
      diff = g_hi * arean * rydif_hi;
    } else {
      flow = gf_arim(D_lo, D_hi, fy_hi) * v_hi * arean;
      diff = gf_harm(g_lo, g_hi, fy_hi) * arean * rydif_hi;
    }
  }
  acof = gf_dapec_sel(flow, diff, iadv);
}

template <int IADV>   // scheme number known at compile time (3 = hybrid, the default of every shipped case), 0 = read it from P
__global__ void __launch_bounds__(384, 1) k_gen_assemble_t_tma(const GenFastParams P, const int S, const int NSPLIT) {
  GT_RING_SETUP(GT_NT)
  const bool has_src = P.su_src != nullptr;
  GT_PRODUCE(GT_NT, GT_NT_NORTH, has_src ? 0u : 0x180000u,
             const double* src[GT_NT] = {P.den + i0 * LD,       P.cp + i0 * LD,         P.gam + i0 * LD,
                                         P.t + i0 * LD,         P.u + i0 * LD,          P.u + (i0 + 1) * LD,
                                         P.den + (i0 - 1) * LD, P.cp + (i0 - 1) * LD,   P.gam + (i0 - 1) * LD,
                                         P.gam + (i0 + 1) * LD, P.den + (i0 + 1) * LD,  P.cp + (i0 + 1) * LD,
                                         P.t + (i0 - 1) * LD,   P.t + (i0 + 1) * LD,    P.den + i0 * LD + T,
                                         P.cp + i0 * LD + T,    P.gam + i0 * LD + T,    P.t + i0 * LD + T,
                                         P.v + i0 * LD + T,     has_src ? P.su_src + i0 * LD : nullptr,
                                         has_src ? P.sp_src + i0 * LD : nullptr};)
  // ---- consumers
  const int t = toff + tid, lane = tid & 31;
  const int j0base = t * M;
  const bool hasprev = j0base >= 1 && j0base < nj, hasnext = j0base + M < nj;
  const int qprev = (M - 1) * T + t - 1, qnext = t + 1, qlast = (M - 1) * T + t;
  const int iadv = IADV ? IADV : P.iadv;
  const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;
  auto load_edge = [&](long long i0) {
    GtTEdge e{};
    const double sew_i = __ldg(P.sew + i0);
    const double *dC = P.den + i0 * LD, *cC = P.cp + i0 * LD, *gC = P.gam + i0 * LD, *vC = P.v + i0 * LD;
    if (hasprev) {
      double flow, acof;
      gt_t_face(__ldg(dC + qprev) * __ldg(cC + qprev), __ldg(dC + t) * __ldg(cC + t), __ldg(gC + qprev), __ldg(gC + t),
                __ldg(vC + t), sew_i, __ldg(P.r + qprev), __ldg(P.rmn + t), __ldg(P.fy + t), __ldg(P.r_ydif + t),
                j0base == 1, j0base == nj - 1, iadv, flow, acof);
      e.aS = acof + gf_max(0.0, flow);
      e.ts = __ldg(P.t + i0 * LD + qprev);
    }
    if (hasnext) {
      gt_t_face(__ldg(dC + qlast) * __ldg(cC + qlast), __ldg(dC + qnext) * __ldg(cC + qnext), __ldg(gC + qlast),
                __ldg(gC + qnext), __ldg(vC + qnext), sew_i, __ldg(P.r + qlast), __ldg(P.rmn + qnext), __ldg(P.fy + qnext),
                __ldg(P.r_ydif + qnext), j0base + M == 1, j0base + M == nj - 1, iadv, e.flowN, e.acofN);
      e.tN = __ldg(P.t + i0 * LD + qnext);
    }
    return e;
  };
  GtTEdge em{};
  if (bline < nlines) em = load_edge(P.ia + bline);
  double x_arx = __ldg(P.arx + t), x_rsx = __ldg(P.r_sx + t), x_ycvr = __ldg(P.ycvr + t);
  const int q1 = M > 1 ? T + t : t + 1;
  double y_fy = __ldg(P.fy + q1), y_rmn = __ldg(P.rmn + q1), y_rydif = __ldg(P.r_ydif + q1);
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const int i = (int)i0 + 1 + P.goff;
    const long long cout = i0 * P.L.cSL;
    GtTEdge nm{};
    if (line + lstride < nlines) nm = load_edge(i0 + lstride);
    const double sew_i = __ldg(P.sew + i0), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rxdif_c = __ldg(P.r_xdif + i0), rxdif_e = __ldg(P.r_xdif + i0 + 1);
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    double aS = em.aS, t_s = em.ts;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const double arx = x_arx, rsx = x_rsx, ycvr = x_ycvr, fy_p = y_fy, rmn_p = y_rmn, rydif_p = y_rydif;
      {
        const int kn = (k + 1 < M) ? k + 1 : 0;
        const int qx = kn * T + t, qy = (kn + 1 < M) ? (kn + 1) * T + t : t + 1;
        x_arx = __ldg(P.arx + qx); x_rsx = __ldg(P.r_sx + qx); x_ycvr = __ldg(P.ycvr + qx);
        y_fy = __ldg(P.fy + qy); y_rmn = __ldg(P.rmn + qy); y_rydif = __ldg(P.r_ydif + qy);
      }
      const unsigned s = rs, ph = rph;
      if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
      mbar_wait(full0 + 8 * s, ph);
      const double* sl = ring + (size_t)s * SW + tid;
      const double D_c = sl[0] * sl[TT], g_c = sl[2 * TT], t_c = sl[3 * TT], u_c = sl[4 * TT], u_e = sl[5 * TT];
      const double D_w = sl[6 * TT] * sl[7 * TT], g_w = sl[8 * TT], g_e = sl[9 * TT];
      const double D_e = sl[10 * TT] * sl[11 * TT], t_w = sl[12 * TT], t_e = sl[13 * TT];
      const double su_s = has_src ? sl[19 * TT] : 0.0, sp_s = has_src ? sl[20 * TT] : 0.0;
      double t_n = em.tN, flowN = em.flowN, acofN = em.acofN;
      const int q = k * T + t;
      if (k + 1 < M) {
        const double D_p = sl[14 * TT] * sl[15 * TT], g_p = sl[16 * TT], v_p = sl[18 * TT];
        t_n = sl[17 * TT];
        flowN = 0.0;
        acofN = 0.0;
        if (j0 + 1 < nj)
          gt_t_face(D_c, D_p, g_c, g_p, v_p, sew_i, j0 == 0 ? __ldg(P.r + q) : 0.0, rmn_p, fy_p, rydif_p, j0 == 0,
                    j0 + 1 == nj - 1, iadv, flowN, acofN);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int cq = k * P.L.cSK + t;
        double flow_w, diff_w, flow_e, diff_e;
        if (west_b) {
          flow_w = D_w * u_c * arx;
          diff_w = g_w * arx * (rxdif_c * rsx);
        } else {
          flow_w = gf_arim(D_w, D_c, fx_c) * u_c * arx;
          diff_w = gf_harm(g_w, g_c, fx_c) * arx * (rxdif_c * rsx);
        }
        if (east_b) {                                            // quirk: den(nj,j) (:390)
          const long long lq = (long long)(nj - 1 - P.goff) * LD + q;     // global line nj, held by the last slab
          const double dq = __ldg(P.den + lq) * __ldg(P.cp + lq);
          flow_e = dq * u_e * arx;
          diff_e = g_e * arx * (rxdif_e * rsx);
        } else {
          flow_e = gf_arim(D_c, D_e, fx_e) * u_e * arx;
          diff_e = gf_harm(g_c, g_e, fx_e) * arx * (rxdif_e * rsx);
        }
        const double AW = gf_dapec_sel(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
        const double AE = (gf_dapec_sel(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
        const double AS = aS;
        const double AN = (acofN + gf_max(0.0, flowN)) - flowN;
        const double vol = ycvr * sew_i;
        const double SP = sp_s * vol;
        double SU = su_s * vol;
        double AP = AE + AW + AN + AS - SP;
        const double reseq = AP * t_c - AE * t_e - AW * t_w - AN * t_n - AS * t_s - SU;
        racc += fabs(reseq);
        AP = AP * r_urf;
        SU = fma(omu * AP, t_c, SU);
        GT_STORE(AE, AW, AN, AS, SU, AP)
      }
      aS = acofN + gf_max(0.0, flowN);
      t_s = t_c;
    }
    em = nm;
  }
  gt_consumer_sum(racc, red, TT, tid, P.partial + blockIdx.x);
}

// =====================================================================================================
// Pressure correction (buoyancy_cav.f90:294-340).
// Stage layout: 0 dC | 1 uC 2 uE 3 dW 4 dE 5 duC 6 duE | north copies 7 dN 8 vN 9 dvN
// =====================================================================================================
#define GT_NP 10
#define GT_NP_NORTH 0x380u
struct GtPEdge { double fS, aS, fN, aN; };

// mass flux through and conductance of the y-face between rows lo and hi (v, dv and the geometry belong to row hi)
__device__ __forceinline__ void gt_p_face(double d_lo, double d_hi, double v_hi, double dv_hi, double arean, double fy_hi,
                                          bool south_b, bool north_b, double& f, double& a) {
  if (north_b) {              // den(i,nj)*v(i,nj)
    f = d_hi * v_hi * arean;
    a = 0.0;
  } else if (south_b) {       // (:301-306) den(i,1)*v(i,2)
    f = d_lo * v_hi * arean;
    a = 0.0;
  } else {
    const double denn = gf_arim(d_lo, d_hi, fy_hi);
    f = denn * v_hi * arean;
    a = denn * arean * dv_hi;
  }
}

__global__ void __launch_bounds__(384, 1) k_gen_assemble_p_tma(const GenFastParams P, const int S, const int NSPLIT) {
  GT_RING_SETUP(GT_NP)
  GT_PRODUCE(GT_NP, GT_NP_NORTH, 0u,
             const double* src[GT_NP] = {P.den + i0 * LD,       P.u + i0 * LD,         P.u + (i0 + 1) * LD,
                                         P.den + (i0 - 1) * LD, P.den + (i0 + 1) * LD, P.du_in + i0 * LD,
                                         P.du_in + (i0 + 1) * LD, P.den + i0 * LD + T, P.v + i0 * LD + T,
                                         P.dv_in + i0 * LD + T};)
  // ---- consumers
  const int t = toff + tid, lane = tid & 31;
  const int j0base = t * M;
  const bool hasprev = j0base >= 1 && j0base < nj, hasnext = j0base + M < nj;
  const int qprev = (M - 1) * T + t - 1, qnext = t + 1, qlast = (M - 1) * T + t;
  double racc = 0.0;
  unsigned rs = 0, rph = 0;
  auto load_edge = [&](long long i0) {
    GtPEdge e{};
    const double sew_i = __ldg(P.sew + i0);
    const double *dC = P.den + i0 * LD, *vC = P.v + i0 * LD, *dvC = P.dv_in + i0 * LD;
    if (hasprev)
      gt_p_face(__ldg(dC + qprev), __ldg(dC + t), __ldg(vC + t), __ldg(dvC + t), __ldg(P.rmn + t) * sew_i, __ldg(P.fy + t),
                j0base == 1, j0base == nj - 1, e.fS, e.aS);
    if (hasnext)
      gt_p_face(__ldg(dC + qlast), __ldg(dC + qnext), __ldg(vC + qnext), __ldg(dvC + qnext), __ldg(P.rmn + qnext) * sew_i,
                __ldg(P.fy + qnext), j0base + M == 1, j0base + M == nj - 1, e.fN, e.aN);
    return e;
  };
  GtPEdge em{};
  if (bline < nlines) em = load_edge(P.ia + bline);
  double x_arx = __ldg(P.arx + t);
  const int q1 = M > 1 ? T + t : t + 1;
  double y_fy = __ldg(P.fy + q1), y_rmn = __ldg(P.rmn + q1);
  for (int line = bline; line < nlines; line += lstride) {
    const long long i0 = P.ia + line;
    const int i = (int)i0 + 1 + P.goff;
    const long long cout = i0 * P.L.cSL;
    GtPEdge nm{};
    if (line + lstride < nlines) nm = load_edge(i0 + lstride);
    const double sew_i = __ldg(P.sew + i0), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    double fS = em.fS, aS = em.aS;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const double arx = x_arx, fy_p = y_fy, rmn_p = y_rmn;
      {
        const int kn = (k + 1 < M) ? k + 1 : 0;
        const int qx = kn * T + t, qy = (kn + 1 < M) ? (kn + 1) * T + t : t + 1;
        x_arx = __ldg(P.arx + qx);
        y_fy = __ldg(P.fy + qy); y_rmn = __ldg(P.rmn + qy);
      }
      const unsigned s = rs, ph = rph;
      if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
      mbar_wait(full0 + 8 * s, ph);
      const double* sl = ring + (size_t)s * SW + tid;
      const double d_c = sl[0], u_c = sl[TT], u_e = sl[2 * TT], den_w = sl[3 * TT], den_e = sl[4 * TT];
      const double du_c = sl[5 * TT], du_e = sl[6 * TT];
      double fN = em.fN, aN = em.aN;
      if (k + 1 < M) {
        const double d_p = sl[7 * TT], v_p = sl[8 * TT], dv_p = sl[9 * TT];
        fN = 0.0;
        aN = 0.0;
        if (j0 + 1 < nj) gt_p_face(d_c, d_p, v_p, dv_p, rmn_p * sew_i, fy_p, j0 == 0, j0 + 1 == nj - 1, fN, aN);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int cq = k * P.L.cSK + t;
        double SU = fS, AW, AE;
        if (west_b) {
          SU = SU + den_w * u_c * arx;
          AW = 0.0;
        } else {
          SU = SU + d_c * u_c * arx;                             // quirk: den(i+1,j) of the upstream cell's formula
          AW = gf_arim(den_w, d_c, fx_c) * arx * du_c;
        }
        SU = SU - den_e * u_e * arx;
        AE = east_b ? 0.0 : gf_arim(d_c, den_e, fx_e) * arx * du_e;
        SU = SU - fN;
        racc += fabs(SU);
        const double APs = AE + AW + aN + aS;
        const double rr = APs != 0.0 ? 1.0 / APs : 0.0;
        double* cb = P.an + cout + cq;
        cb[0] = aN * rr;
        cb[T] = aS * rr;
        cb[2 * T] = AE * rr;
        cb[3 * T] = AW * rr;
        cb[4 * T] = SU * rr;
      }
      fS = fN;
      aS = aN;
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
  return use_tma && which <= 3 && gt_nsplit(c->L.T) > 0;
}

// which: 0 u, 1 v, 2 p', 3 T.  Returns the number of residual partials (grid size), or a negative status.
int cfd_launch_general_assemble_tma(cfd_ctx* c, int which, const GenFastParams& P) {
  const int num_sms = c->num_sms;
  const int T = c->L.T, narr = which == 0 ? GT_NU : which == 1 ? GT_NV : which == 2 ? GT_NP : GT_NT, nsplit = gt_nsplit(T), TT = T / nsplit;
  const size_t stage = (size_t)narr * TT * sizeof(double);
  int S = 8;
  const char* es = getenv("CFD_B200_ASM_STAGES");
  if (es && atoi(es) > 0) S = atoi(es);
  while (S > 2 && S * stage + 32 * 8 + 2 * S * 8 > 200 * 1024) --S;
  const size_t smem = S * stage + 32 * 8 + 2 * (size_t)S * 8;
  const int nlines = P.ib - P.ia + 1;
  int nblocks = nlines * nsplit < num_sms ? nlines * nsplit : num_sms / nsplit * nsplit;
  const int nthreads = TT + 32 * GT_PRODUCER_WARPS;
  const bool hybrid = P.iadv == 3;
  void (*kernel)(const GenFastParams, const int, const int) =
      which == 0   ? (hybrid ? k_gen_assemble_u_tma<3> : k_gen_assemble_u_tma<0>)
      : which == 1 ? (hybrid ? k_gen_assemble_v_tma<3> : k_gen_assemble_v_tma<0>)
      : which == 2 ? k_gen_assemble_p_tma
                   : (hybrid ? k_gen_assemble_t_tma<3> : k_gen_assemble_t_tma<0>);
  cudaError_t e = cfd_raise_smem(c, kernel);
  if (e == cudaSuccess) kernel<<<nblocks, nthreads, smem, c->stream>>>(P, S, nsplit);
  if (e != cudaSuccess) {
    cfd_set_error("k_gen_assemble_*_tma: %s", cudaGetErrorString(e));
    return CFD_ERR_CUDA;
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return nblocks;
}
