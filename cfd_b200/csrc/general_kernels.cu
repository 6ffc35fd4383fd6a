// general_kernels.cu -- coefficient assembly of the reference's "general" SIMPLE family (Cartesian / cylindrical /
// polar, DAPEC scheme selector, coupled temperature equation):
//   calcu buoyancy-driven-cavity-2d/buoyancy_cav.f90:122-198, calcv :200-277, calcp :279-343, calct :358-428,
//   DAPEC / ARIMEAN / HARMEAN cylinder-heat-conduction-2d/cylin_cond.f90:245-285.
//
// The reference walks the grid face by face and scatters (aw(i+1,j) and as(i,j+1) are written while visiting
// cell (i,j); calcp also scatters mass fluxes into su(i+1,j), su(i,j+1)).  Here every cell evaluates its own
// four faces with the formula the reference used when it produced that face -- the same expression on the
// same operands gives the same bits, so the results are identical without any scatter -- and accumulates the
// p' source in the reference's order ((src + south) + west) - east - north.  Compiled with -fmad=false; all
// arithmetic follows the reference's operation order (bit-comparable with oracle/general_oracle.c).
//
// Thread mapping as in cavity_kernels.cu (one thread = M consecutive nodes of one line).  Neighbours are addressed
// through LinePos (adds / selects relative to the thread's chunk, linepos.cuh); the production path with register
// windows is general_fast_kernels.cu.
#include "common.cuh"
#include "linepos.cuh"

namespace cfd {

struct GenParams {
  Layout L;
  int ia, ib, jlo, jhi;     // 0-based inclusive cell ranges
  int iadv;
  double urf;
  int boussinesq;           // v equation: su = rgbeta*(ARIMEAN(t(i,j-1),t(i,j),fy(j)) - tref)
  double rgbeta, tref;
  int scale_den_by_cp;      // calct: den(:,:) = den(:,:)*cp(:,:) (buoyancy_cav.f90:372)
  const double *u, *v, *p, *t, *den, *gam, *cp, *du_in, *dv_in, *su_src, *sp_src;
  double *ae, *aw, *an, *as, *ap, *su, *dout;
  // by-i geometry (natural order)
  const double *sew, *sewu, *xdif, *xcvi, *xcvip, *fx;
  // by-j geometry (line layout, read through LinePos)
  const double *sns, *ydif, *ycvr, *ycvrs, *arx, *arxj, *arxjp, *r, *rmn, *sx, *sxmn, *fvp, *fy;
  double* partial;
};

__device__ __forceinline__ double gmax(double a, double b) { return a > b ? a : b; }

// cylin_cond.f90:245-269
__device__ __forceinline__ double DAPEC(double flow, double diff, int iadv) {
  if (flow == 0.0) return diff;
  switch (iadv) {
    case 1: return diff - 0.5 * fabs(flow);
    case 2: return diff;
    case 3: return gmax(0.0, diff - 0.5 * fabs(flow));
    case 4: {
      const double q = 1.0 - 0.1 * fabs(flow / diff);
      const double q2 = q * q;
      return gmax(0.0, q2 * q2 * q) * diff;
    }
    default: return diff;
  }
}
__device__ __forceinline__ double ARIMEAN(double a, double b, double w) { return a * (1 - w) + b * w; }
__device__ __forceinline__ double HARMEAN(double a, double b, double w) { return a * b / (a * (1.0 - w) + b * w); }

// field access by 0-based (i0, j0)
struct Fld {
  const double* p;
  long long LD;
  LinePos lp;
  __device__ __forceinline__ double operator()(int i0, int j0) const { return __ldg(p + (long long)i0 * LD + lp(j0)); }
};

__device__ __forceinline__ void block_sum_gen(double v, double* partial) {
  __shared__ double wsum[32];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int lane = tid & 31, wid = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if (lane == 0) wsum[wid] = v;
  __syncthreads();
  if (wid == 0) {
    v = (lane < nw) ? wsum[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) partial[blockIdx.y * gridDim.x + blockIdx.x] = v;
  }
}

// 1-based helpers on 1-D arrays: by-i arrays in natural order, by-j arrays in line layout (needs `lp` in scope)
#define G1(a, i) __ldg((a) + (i) - 1)
#define GJ(a, j) __ldg((a) + lp((j) - 1))

// ---- u equation: buoyancy_cav.f90:135-194 ------------------------------------------------------------------
// x-face produced while the reference visits u-cell (ii, j): returns flow, acof of the face EAST of cell ii
__device__ __forceinline__ void u_face_e(const GenParams& P, const Fld& U, const Fld& D, const Fld& Gm, int ii, int j,
                                         int ni, int nim1_loop, double& flow, double& acof) {
  const LinePos& lp = U.lp;
  double diff;
  if (ii == nim1_loop) {   // east boundary (`i == njm1`, :150)
    flow = GJ(P.arx, j) * U(ni - 1, j - 1) * D(ni - 1, j - 1);
    diff = GJ(P.arx, j) * Gm(ni - 1, j - 1) / (G1(P.sew, ni - 1) * GJ(P.sx, j));
  } else {
    const double fl = U(ii - 1, j - 1) * ARIMEAN(D(ii - 2, j - 1), D(ii - 1, j - 1), G1(P.fx, ii));
    const double flp = U(ii, j - 1) * ARIMEAN(D(ii - 1, j - 1), D(ii, j - 1), G1(P.fx, ii + 1));
    flow = GJ(P.arx, j) * 0.5 * (fl + flp);
    diff = GJ(P.arx, j) * Gm(ii - 1, j - 1) / (G1(P.sew, ii) * GJ(P.sx, j));
  }
  acof = DAPEC(flow, diff, P.iadv);
}
// y-face NORTH of u-cell (i, jj)
__device__ __forceinline__ void u_face_n(const GenParams& P, const Fld& V, const Fld& D, const Fld& Gm, int i, int jj,
                                         int nj, double& flow, double& acof) {
  const LinePos& lp = V.lp;
  double fl, flm, diff;
  if (jj == nj - 1) {
    fl = G1(P.xcvi, i) * V(i - 1, nj - 1) * D(i - 1, nj - 1);
    flm = G1(P.xcvip, i - 1) * V(i - 2, nj - 1) * D(i - 2, nj - 1);
    diff = GJ(P.r, nj) * (G1(P.xcvi, i) * Gm(i - 1, nj - 1) + G1(P.xcvip, i - 1) * Gm(i - 2, nj - 1)) / GJ(P.ydif, nj);
  } else {
    const double fyp = GJ(P.fy, jj + 1);
    fl = G1(P.xcvi, i) * V(i - 1, jj) * ARIMEAN(D(i - 1, jj - 1), D(i - 1, jj), fyp);
    flm = G1(P.xcvip, i - 1) * V(i - 2, jj) * ARIMEAN(D(i - 2, jj - 1), D(i - 2, jj), fyp);
    const double vis_ne = HARMEAN(Gm(i - 1, jj - 1), Gm(i - 1, jj), fyp);
    const double vis_nw = HARMEAN(Gm(i - 2, jj - 1), Gm(i - 2, jj), fyp);
    diff = GJ(P.rmn, jj + 1) * (vis_ne * G1(P.xcvi, i) + vis_nw * G1(P.xcvip, i - 1)) / GJ(P.ydif, jj + 1);
  }
  flow = GJ(P.rmn, jj + 1) * (fl + flm);
  acof = DAPEC(flow, diff, P.iadv);
}

__global__ void __launch_bounds__(256) k_gen_assemble_u(const GenParams P) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const Fld U{P.u, LD, lp}, V{P.v, LD, lp}, D{P.den, LD, lp}, Gm{P.gam, LD, lp}, Pr{P.p, LD, lp};
    const int i = i0 + 1;
    const int iloop_last = nj - 1;     // quirk: the reference loops `do i = 3, njm1` (:135,:149); ni == nj is enforced
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      double flow, acof, AW, AE, AS, AN;
      // west face
      if (i == 3) {
        flow = GJ(P.arx, j) * U(1, j0) * D(0, j0);
        const double diff = GJ(P.arx, j) * Gm(0, j0) / (G1(P.sew, 2) * GJ(P.sx, j));
        acof = DAPEC(flow, diff, P.iadv);
      } else {
        u_face_e(P, U, D, Gm, i - 1, j, ni, iloop_last, flow, acof);
      }
      AW = acof + gmax(0.0, flow);
      // east face
      u_face_e(P, U, D, Gm, i, j, ni, iloop_last, flow, acof);
      AE = (acof + gmax(0.0, flow)) - flow;
      // south face
      if (j == 2) {
        const double fl = G1(P.xcvi, i) * V(i0, 1) * D(i0, 0);
        const double flm = G1(P.xcvip, i - 1) * V(i0 - 1, 1) * D(i0 - 1, 0);
        flow = GJ(P.r, 1) * (fl + flm);
        const double diff = GJ(P.r, 1) * (G1(P.xcvi, i) * Gm(i0, 0) + G1(P.xcvip, i - 1) * Gm(i0 - 1, 0)) / GJ(P.ydif, 2);
        acof = DAPEC(flow, diff, P.iadv);
      } else {
        u_face_n(P, V, D, Gm, i, j - 1, nj, flow, acof);
      }
      AS = acof + gmax(0.0, flow);
      // north face
      u_face_n(P, V, D, Gm, i, j, nj, flow, acof);
      AN = (acof + gmax(0.0, flow)) - flow;
      const double vol = GJ(P.ycvr, j) * G1(P.sewu, i);
      double SU = 0.0 * vol, SP = 0.0 * vol;
      double DU = vol / (G1(P.xdif, i) * GJ(P.sx, j));
      SU = SU + DU * (Pr(i0 - 1, j0) - Pr(i0, j0));
      double AP = AE + AW + AN + AS - SP;
      const double reseq = AP * U(i0, j0) - AE * U(i0 + 1, j0) - AW * U(i0 - 1, j0) - AN * U(i0, j0 + 1)
                           - AS * U(i0, j0 - 1) - SU;
      racc += fabs(reseq);
      AP = AP / P.urf;
      SU = SU + (1 - P.urf) * AP * U(i0, j0);
      DU = DU / AP;
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
      P.dout[(long long)i0 * LD + k * T + t] = DU;
    }
  }
  block_sum_gen(racc, P.partial);
}

// ---- v equation: buoyancy_cav.f90:212-273 ------------------------------------------------------------------
__device__ __forceinline__ void v_face_e(const GenParams& P, const Fld& U, const Fld& D, const Fld& Gm, int ii, int j,
                                         int ni, double& flow, double& acof) {
  const LinePos& lp = U.lp;
  double fl, flm, diff;
  if (ii == ni - 1) {
    fl = GJ(P.arxj, j) * U(ni - 1, j - 1) * D(ni - 1, j - 1);
    flm = GJ(P.arxjp, j - 1) * U(ni - 1, j - 2) * D(ni - 1, j - 2);
    diff = (GJ(P.arxj, j) * Gm(ni - 1, j - 1) + GJ(P.arxjp, j - 1) * Gm(ni - 1, j - 2)) / (G1(P.xdif, ni) * GJ(P.sxmn, j));
  } else {
    const double fxp = G1(P.fx, ii + 1);
    fl = GJ(P.arxj, j) * U(ii, j - 1) * ARIMEAN(D(ii - 1, j - 1), D(ii, j - 1), fxp);
    flm = GJ(P.arxjp, j - 1) * U(ii, j - 2) * ARIMEAN(D(ii - 1, j - 2), D(ii, j - 2), fxp);
    const double vis_ne = HARMEAN(Gm(ii - 1, j - 1), Gm(ii, j - 1), fxp);
    const double vis_se = HARMEAN(Gm(ii - 1, j - 2), Gm(ii, j - 2), fxp);
    diff = (vis_ne * GJ(P.arxj, j) + vis_se * GJ(P.arxjp, j - 1)) / (G1(P.xdif, ii + 1) * GJ(P.sxmn, j));
  }
  flow = fl + flm;
  acof = DAPEC(flow, diff, P.iadv);
}
__device__ __forceinline__ void v_face_n(const GenParams& P, const Fld& V, const Fld& D, const Fld& Gm, int i, int jj,
                                         int nj, double& flow, double& acof) {
  const LinePos& lp = V.lp;
  double diff;
  if (jj == nj - 1) {
    const double area = GJ(P.r, nj) * G1(P.sew, i);
    flow = area * V(i - 1, nj - 1) * D(i - 1, nj - 1);
    diff = area * Gm(i - 1, nj - 1) / GJ(P.sns, nj - 1);
  } else {
    const double area = GJ(P.r, jj) * G1(P.sew, i);
    const double fl = V(i - 1, jj - 1) * GJ(P.rmn, jj) * G1(P.sew, i) * ARIMEAN(D(i - 1, jj - 2), D(i - 1, jj - 1), GJ(P.fy, jj));
    const double flp = V(i - 1, jj) * GJ(P.rmn, jj + 1) * G1(P.sew, i) * ARIMEAN(D(i - 1, jj - 1), D(i - 1, jj), GJ(P.fy, jj + 1));
    flow = ARIMEAN(fl, flp, GJ(P.fvp, jj));
    diff = area * Gm(i - 1, jj - 1) / GJ(P.sns, jj);
  }
  acof = DAPEC(flow, diff, P.iadv);
}

__global__ void __launch_bounds__(256) k_gen_assemble_v(const GenParams P) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const Fld U{P.u, LD, lp}, V{P.v, LD, lp}, D{P.den, LD, lp}, Gm{P.gam, LD, lp}, Pr{P.p, LD, lp},
        Tt{P.t, LD, lp};
    const int i = i0 + 1;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      double flow, acof, AW, AE, AS, AN;
      if (i == 2) {
        const double fl = GJ(P.arxj, j) * U(1, j0) * D(0, j0);
        const double flm = GJ(P.arxjp, j - 1) * U(1, j0 - 1) * D(0, j0 - 1);
        flow = fl + flm;
        const double diff = (GJ(P.arxj, j) * Gm(0, j0) + GJ(P.arxjp, j - 1) * Gm(0, j0 - 1)) / (G1(P.xdif, 2) * GJ(P.sxmn, j));
        acof = DAPEC(flow, diff, P.iadv);
      } else {
        v_face_e(P, U, D, Gm, i - 1, j, ni, flow, acof);
      }
      AW = acof + gmax(0.0, flow);
      v_face_e(P, U, D, Gm, i, j, ni, flow, acof);
      AE = (acof + gmax(0.0, flow)) - flow;
      if (j == 3) {
        const double area = GJ(P.r, 1) * G1(P.sew, i);
        flow = area * V(i0, 1) * D(i0, 0);
        const double diff = area * Gm(i0, 0) * GJ(P.sns, 2);     // quirk: multiply (buoyancy_cav.f90:216)
        acof = DAPEC(flow, diff, P.iadv);
      } else {
        v_face_n(P, V, D, Gm, i, j - 1, nj, flow, acof);
      }
      AS = acof + gmax(0.0, flow);
      v_face_n(P, V, D, Gm, i, j, nj, flow, acof);
      AN = (acof + gmax(0.0, flow)) - flow;
      const double vol = GJ(P.ycvrs, j) * G1(P.sew, i);
      double src = 0.0;
      if (P.boussinesq) src = P.rgbeta * (ARIMEAN(Tt(i0, j0 - 1), Tt(i0, j0), GJ(P.fy, j)) - P.tref);
      double SU = src * vol, SP = 0.0 * vol;
      double DV = vol / GJ(P.ydif, j);
      SU = SU + DV * (Pr(i0, j0 - 1) - Pr(i0, j0));
      double AP = AE + AW + AN + AS - SP;
      const double reseq = AP * V(i0, j0) - AE * V(i0 + 1, j0) - AW * V(i0 - 1, j0) - AN * V(i0, j0 + 1)
                           - AS * V(i0, j0 - 1) - SU;
      racc += fabs(reseq);
      AP = AP / P.urf;
      SU = SU + (1.0 - P.urf) * AP * V(i0, j0);
      DV = DV / AP;
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
      P.dout[(long long)i0 * LD + k * T + t] = DV;
    }
  }
  block_sum_gen(racc, P.partial);
}

// ---- p' equation: buoyancy_cav.f90:294-340 -----------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gen_assemble_p(const GenParams P) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const Fld U{P.u, LD, lp}, V{P.v, LD, lp}, D{P.den, LD, lp}, DU{P.du_in, LD, lp}, DV{P.dv_in, LD, lp};
    const int i = i0 + 1, nim1 = ni - 1, njm1 = nj - 1;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      const double areaew = GJ(P.arx, j);
      double SU = 0.0, AE, AW, AN, AS;
      // inflow through the south face (scattered by cell (i,j-1), or the boundary loop :301-306)
      if (j == 2) {
        SU = SU + D(i0, 0) * V(i0, 1) * (GJ(P.rmn, 2) * G1(P.sew, i));
        AS = 0.0;
      } else {
        const double denn = ARIMEAN(D(i0, j0 - 1), D(i0, j0), GJ(P.fy, j));
        const double arean = GJ(P.rmn, j) * G1(P.sew, i);
        SU = SU + denn * V(i0, j0) * arean;
        AS = denn * arean * DV(i0, j0);
      }
      // inflow through the west face
      if (i == 2) {
        SU = SU + D(0, j0) * U(1, j0) * areaew;
        AW = 0.0;
      } else {
        SU = SU + D(i0, j0) * U(i0, j0) * areaew;          // quirk: den(i+1,j) of the upstream cell's formula
        AW = ARIMEAN(D(i0 - 1, j0), D(i0, j0), G1(P.fx, i)) * areaew * DU(i0, j0);
      }
      // east
      SU = SU - D(i0 + 1, j0) * U(i0 + 1, j0) * areaew;
      AE = (i == nim1) ? 0.0 : ARIMEAN(D(i0, j0), D(i0 + 1, j0), G1(P.fx, i + 1)) * areaew * DU(i0 + 1, j0);
      // north
      const double arean = GJ(P.rmn, j + 1) * G1(P.sew, i);
      if (j == njm1) {
        SU = SU - D(i0, nj - 1) * V(i0, nj - 1) * arean;
        AN = 0.0;
      } else {
        const double denn = ARIMEAN(D(i0, j0), D(i0, j0 + 1), GJ(P.fy, j + 1));
        SU = SU - denn * V(i0, j0 + 1) * arean;
        AN = denn * arean * DV(i0, j0 + 1);
      }
      const double AP = AE + AW + AN + AS - 0.0;
      racc += fabs(SU);
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
    }
  }
  block_sum_gen(racc, P.partial);
}

// ---- scalar (temperature) equation: buoyancy_cav.f90:373-428 ---------------------------------------------------
struct DenCp {
  Fld D, C;
  int scale;
  __device__ __forceinline__ double operator()(int i0, int j0) const { return scale ? D(i0, j0) * C(i0, j0) : D(i0, j0); }
};
__device__ __forceinline__ void t_face_e(const GenParams& P, const Fld& U, const DenCp& D, const Fld& Gm, int ii, int j,
                                         int ni, int nj, double& flow, double& acof) {
  const LinePos& lp = U.lp;
  const double areaew = GJ(P.arx, j);
  double diff;
  if (ii == ni - 1) {
    const double dene = D(nj - 1, j - 1);                     // quirk: den(nj,j) (buoyancy_cav.f90:390)
    flow = dene * U(ni - 1, j - 1) * areaew;
    diff = Gm(ni - 1, j - 1) * areaew / (G1(P.xdif, ni) * GJ(P.sx, j));
  } else {
    const double dene = ARIMEAN(D(ii - 1, j - 1), D(ii, j - 1), G1(P.fx, ii + 1));
    flow = dene * U(ii, j - 1) * areaew;
    const double game = HARMEAN(Gm(ii - 1, j - 1), Gm(ii, j - 1), G1(P.fx, ii + 1));
    diff = game * areaew / (G1(P.xdif, ii + 1) * GJ(P.sx, j));
  }
  acof = DAPEC(flow, diff, P.iadv);
}
__device__ __forceinline__ void t_face_n(const GenParams& P, const Fld& V, const DenCp& D, const Fld& Gm, int i, int jj,
                                         int nj, double& flow, double& acof) {
  const LinePos& lp = V.lp;
  const double arean = GJ(P.rmn, jj + 1) * G1(P.sew, i);
  double diff;
  if (jj == nj - 1) {
    flow = D(i - 1, nj - 1) * V(i - 1, nj - 1) * arean;
    diff = Gm(i - 1, nj - 1) * arean / GJ(P.ydif, nj);
  } else {
    const double denn = ARIMEAN(D(i - 1, jj - 1), D(i - 1, jj), GJ(P.fy, jj + 1));
    flow = denn * V(i - 1, jj) * arean;
    const double gamn = HARMEAN(Gm(i - 1, jj - 1), Gm(i - 1, jj), GJ(P.fy, jj + 1));
    diff = gamn * arean / GJ(P.ydif, jj + 1);
  }
  acof = DAPEC(flow, diff, P.iadv);
}

__global__ void __launch_bounds__(256) k_gen_assemble_t(const GenParams P) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const Fld U{P.u, LD, lp}, V{P.v, LD, lp}, Gm{P.gam, LD, lp}, Tt{P.t, LD, lp};
    const DenCp D{Fld{P.den, LD, lp}, Fld{P.cp, LD, lp}, P.scale_den_by_cp};
    const Fld Ss{P.su_src, LD, lp}, Sp{P.sp_src, LD, lp};
    const int i = i0 + 1;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      double flow, acof, AW, AE, AS, AN;
      if (i == 2) {
        flow = D(0, j0) * U(1, j0) * GJ(P.arx, j);
        const double diff = Gm(0, j0) * GJ(P.arx, j) / (G1(P.xdif, 2) * GJ(P.sx, j));
        acof = DAPEC(flow, diff, P.iadv);
      } else {
        t_face_e(P, U, D, Gm, i - 1, j, ni, nj, flow, acof);
      }
      AW = acof + gmax(0.0, flow);
      t_face_e(P, U, D, Gm, i, j, ni, nj, flow, acof);
      AE = (acof + gmax(0.0, flow)) - flow;
      if (j == 2) {
        const double areas = GJ(P.r, 1) * G1(P.sew, i);
        flow = D(i0, 0) * V(i0, 1) * areas;
        const double diff = Gm(i0, 0) * areas / GJ(P.ydif, 2);
        acof = DAPEC(flow, diff, P.iadv);
      } else {
        t_face_n(P, V, D, Gm, i, j - 1, nj, flow, acof);
      }
      AS = acof + gmax(0.0, flow);
      t_face_n(P, V, D, Gm, i, j, nj, flow, acof);
      AN = (acof + gmax(0.0, flow)) - flow;
      const double vol = GJ(P.ycvr, j) * G1(P.sew, i);
      const double SP = (P.sp_src ? Sp(i0, j0) : 0.0) * vol;
      double SU = (P.su_src ? Ss(i0, j0) : 0.0) * vol;
      double AP = AE + AW + AN + AS - SP;
      const double reseq = AP * Tt(i0, j0) - AE * Tt(i0 + 1, j0) - AW * Tt(i0 - 1, j0) - AN * Tt(i0, j0 + 1)
                           - AS * Tt(i0, j0 - 1) - SU;
      racc += fabs(reseq);
      AP = AP / P.urf;
      SU = SU + (1.0 - P.urf) * AP * Tt(i0, j0);
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
    }
  }
  block_sum_gen(racc, P.partial);
}

}  // namespace cfd

using namespace cfd;

// which: 0 u, 1 v, 2 p', 3 T
int cfd_launch_general_assemble(cfd_ctx* c, int which) {
  GenParams P{};
  P.L = c->L;
  P.iadv = c->cfg.iadv;
  P.u = c->field[CFD_U]; P.v = c->field[CFD_V]; P.p = c->field[CFD_P]; P.t = c->field[CFD_T];
  P.den = c->field[CFD_DEN]; P.cp = c->field[CFD_CP];
  P.du_in = c->field[CFD_DU]; P.dv_in = c->field[CFD_DV];
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP]; P.su = c->field[CFD_SU];
  P.sew = c->geomn[CFD_G_SEW]; P.sewu = c->geomn[CFD_G_SEWU]; P.xdif = c->geomn[CFD_G_XDIF];
  P.xcvi = c->geomn[CFD_G_XCVI]; P.xcvip = c->geomn[CFD_G_XCVIP]; P.fx = c->geomn[CFD_G_FX];
  P.sns = c->geom[CFD_G_SNS]; P.ydif = c->geom[CFD_G_YDIF]; P.ycvr = c->geom[CFD_G_YCVR];
  P.ycvrs = c->geom[CFD_G_YCVRS]; P.arx = c->geom[CFD_G_ARX]; P.arxj = c->geom[CFD_G_ARXJ];
  P.arxjp = c->geom[CFD_G_ARXJP]; P.r = c->geom[CFD_G_R]; P.rmn = c->geom[CFD_G_RMN]; P.sx = c->geom[CFD_G_SX];
  P.sxmn = c->geom[CFD_G_SXMN]; P.fvp = c->geom[CFD_G_FVP]; P.fy = c->geom[CFD_G_FY];
  const int ni = c->L.ni, nj = c->L.nj;
  int eq = which;
  P.jhi = nj - 2;
  if (which == 0) {
    P.ia = 2; P.ib = ni - 2; P.jlo = 1;
    P.gam = c->field[CFD_VIS]; P.dout = c->field[CFD_DU];
  } else if (which == 1) {
    P.ia = 1; P.ib = ni - 2; P.jlo = 2;
    P.gam = c->field[CFD_VIS]; P.dout = c->field[CFD_DV];
    P.boussinesq = c->boussinesq; P.rgbeta = c->rgbeta; P.tref = c->tref;
  } else if (which == 2) {
    P.ia = 1; P.ib = ni - 2; P.jlo = 1;
  } else {
    P.ia = 1; P.ib = ni - 2; P.jlo = 1;
    P.gam = c->field[CFD_GAMT];
    P.su_src = c->has_t_sources ? c->field[CFD_SUSRC] : nullptr;
    P.sp_src = c->has_t_sources ? c->field[CFD_SPSRC] : nullptr;
    P.scale_den_by_cp = 1;
  }
  P.urf = c->cfg.urf[eq];
  const int nlines = P.ib - P.ia + 1;
  const int BXg = 32, BYg = 8;
  dim3 block(BXg, BYg), grid((c->L.T + BXg - 1) / BXg, (nlines + BYg - 1) / BYg);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  P.partial = c->partial + (size_t)eq * c->max_partials;
  if (!c->exact_assembly) {   // production (red/black) contexts: general_fast_kernels.cu, equations stored divided by ap
    const int nb = cfd_launch_general_assemble_fast(c, which, P.partial);
    if (nb < 0) return nb;
    return cfd_launch_reduce_partials(c, eq, nb);
  }
  switch (which) {
    case 0: k_gen_assemble_u<<<grid, block, 0, c->stream>>>(P); break;
    case 1: k_gen_assemble_v<<<grid, block, 0, c->stream>>>(P); break;
    case 2: k_gen_assemble_p<<<grid, block, 0, c->stream>>>(P); break;
    default: k_gen_assemble_t<<<grid, block, 0, c->stream>>>(P); break;
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return cfd_launch_reduce_partials(c, eq, nblocks);
}
