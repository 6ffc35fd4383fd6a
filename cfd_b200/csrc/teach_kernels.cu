// teach_kernels.cu -- TEACH-T sudden-expansion ("backward step") family, laminar subset:
//   calcu backward-step-flow-2d/calc.f90:20-124, calcv :137-220, calcp :231-280 (non-uniform, axisymmetric
//   geometry: radii r, rv, rcv in the areas), with the device side of the boundary hooks modu / modv
//   (promod.f90:26-95, 99-133: wall shear in sp, zeroed wall links, symmetry row, outlet mass balance).
//
// The reference runs "coefficients (chapter 1) -> call modX (chapter 2) -> ap, residual, relaxation (chapter 3)".
// Here chapters 1-3 are one kernel per equation: the hook's coefficient edits are applied per cell before the
// cell's own chapter-3 work, and the two FIELD edits of modu -- u(i,1) = u(i,2) and the outlet column
// u(ni,j) = u(nim1,j) + uinc -- are used *virtually* in the residual (chapter 3 reads the edited values, chapter 1
// the old ones) and written to memory by k_teach_modu_fields right after, before the line sweeps.  uinc needs
// two sums over the outlet column; one thread adds them in the reference's order (k_teach_outlet), so the
// sequential-order mode stays bit-comparable with oracle/teach_oracle.c (*_f64).  Compiled with -fmad=false.
// FP64 throughout (the reference is FP32 as written).  k-epsilon / energy: not in this subset.
#include "common.cuh"
#include "linepos.cuh"

namespace cfd {

struct TeachParams {
  Layout L;
  int ia, ib, jlo, jhi;
  int indcos, istep, jstep;
  double urf, viscos, great, flowin;
  const double *u, *v, *p, *den, *vis, *du_in, *dv_in;
  double *ae, *aw, *an, *as, *ap, *su, *dout;
  double* ufield;           // u, writable (k_teach_modu_fields)
  double* uinc;             // device scalar
  const double *tmult_n, *tmult_w;   // wall-law multipliers per i / per j (turbulent contexts), else null
  int nsym;                 // local lines 1..nsym get u(i,1) = u(i,2) (global i <= ni-1)
  int has_outlet;           // this context holds the outlet column (global i = ni) as its last local line
  const double *x, *y, *xu, *yv, *r, *rv, *rcv, *sew, *sns, *sewu, *snsv;
  const double *dxep, *dxpw, *dynp, *dyps, *dxepu, *dxpwu, *dynpv, *dypsv;
  double* partial;
};

__device__ __forceinline__ double tmax(double a, double b) { return a > b ? a : b; }

struct TF {
  const double* p;
  long long LD;
  LinePos lp;
  __device__ __forceinline__ double operator()(int i, int j) const {   // 1-based (i,j)
    return __ldg(p + (long long)(i - 1) * LD + lp(j - 1));
  }
};
// 1-based 1-D arrays: by-i in natural order, by-j in line layout (needs `lp` in scope)
#define T1(a, i) __ldg((a) + (i) - 1)
#define TJ(a, j) __ldg((a) + lp((j) - 1))

__device__ __forceinline__ void block_sum_teach(double v, double* partial) {
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

// promod.f90:79-91: outlet mass balance.  The two sums run in the reference's order (thread 0 adds them one by
// one), but the operands are gathered by the whole block into shared memory first, tile by tile, so the serial
// part is a chain of additions on shared memory instead of a chain of dependent global loads.
__global__ void __launch_bounds__(1024) k_teach_outlet(const TeachParams P) {
  __shared__ double s_ard[1024], s_u[1024];
  const int nj = P.L.nj, nim1 = P.L.ni - 1, M = P.L.M, T = P.L.T;
  const long long LD = P.L.LD;
  const double* dA = P.den + (long long)(nim1 - 1) * LD;   // den(nim1, :)
  const double* dB = dA - LD;                              // den(nim1-1, :)
  const double* uA = P.u + (long long)(nim1 - 1) * LD;     // u(nim1, :)
  double ardent = 0.0, flow = 0.0;
  for (int base = 2; base <= nj - 1; base += 1024) {
    const int j = base + threadIdx.x;
    if (j <= nj - 1) {
      const int q = lpos(j - 1, M, T);
      s_ard[threadIdx.x] = 0.5 * (__ldg(dA + q) + __ldg(dB + q)) * __ldg(P.r + q) * __ldg(P.sns + q);
      s_u[threadIdx.x] = __ldg(uA + q);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      const int n = (nj - base < 1024) ? nj - base : 1024;
      for (int k = 0; k < n; ++k) {
        ardent = ardent + s_ard[k];
        flow = flow + s_ard[k] * s_u[k];
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *P.uinc = (P.flowin - flow) / ardent;
}

// promod.f90:72-77, 92-94: the field edits of modu, after the assembly kernel has used the old values
__global__ void k_teach_modu_fields(const TeachParams P) {
  const int ni = P.L.ni, nj = P.L.nj, M = P.L.M, T = P.L.T;
  const long long LD = P.L.LD;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  auto at = [&](int i, int j) -> double& { return P.ufield[(long long)(i - 1) * LD + ((j - 1) % M) * T + ((j - 1) / M)]; };
  if (n >= 1 && n <= P.nsym) at(n, 1) = at(n, 2);                 // symmetry axis, global i = 1..ni-1
  if (P.has_outlet && n >= 2 && n <= nj - 1) {                    // outlet column, j = 2..nj-1
    const double uinc = *P.uinc;
    if (n == 2) at(ni, 1) = at(ni, 2);                            // i = ni of the symmetry loop sees the OLD u(ni,2)
    at(ni, n) = at(ni - 1, n) + uinc;
  }
}

// ---- u: calc.f90:28-124 + modu coefficient edits ------------------------------------------------------------
__global__ void __launch_bounds__(256) k_teach_assemble_u(const TeachParams P) {
  const int M = P.L.M, T = P.L.T, ni = P.L.ni, nj = P.L.nj, nim1 = ni - 1, njm1 = nj - 1;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const TF U{P.u, LD, lp}, V{P.v, LD, lp}, D{P.den, LD, lp}, S{P.vis, LD, lp}, Pr{P.p, LD, lp};
    const int i = i0 + 1;
    const double uinc = *P.uinc;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      const double arean = TJ(P.rv, j + 1) * T1(P.sewu, i);
      const double areas = TJ(P.rv, j) * T1(P.sewu, i);
      const double areaew = TJ(P.r, j) * TJ(P.sns, j);
      double vol = TJ(P.r, j) * T1(P.sewu, i) * TJ(P.sns, j);
      const double gn = 0.5 * (D(i, j + 1) + D(i, j)) * V(i, j + 1);
      const double gnw = 0.5 * (D(i - 1, j) + D(i - 1, j + 1)) * V(i - 1, j + 1);
      const double gs = 0.5 * (D(i, j - 1) + D(i, j)) * V(i, j);
      const double gsw = 0.5 * (D(i - 1, j) + D(i - 1, j - 1)) * V(i - 1, j);
      const double ge = 0.5 * (D(i + 1, j) + D(i, j)) * U(i + 1, j);
      const double gp = 0.5 * (D(i, j) + D(i - 1, j)) * U(i, j);
      const double gw = 0.5 * (D(i - 1, j) + D(i - 2, j)) * U(i - 1, j);
      const double cn = 0.5 * (gn + gnw) * arean;
      const double cs = 0.5 * (gs + gsw) * areas;
      const double ce = 0.5 * (ge + gp) * areaew;
      const double cw = 0.5 * (gp + gw) * areaew;
      const double visn = 0.25 * (S(i, j) + S(i, j + 1) + S(i - 1, j) + S(i - 1, j + 1));
      const double viss = 0.25 * (S(i, j) + S(i, j - 1) + S(i - 1, j) + S(i - 1, j - 1));
      const double dn = visn * arean / TJ(P.dynp, j);
      const double ds = viss * areas / TJ(P.dyps, j);
      const double de = S(i, j) * areaew / T1(P.dxepu, i);
      const double dw = S(i - 1, j) * areaew / T1(P.dxpwu, i);
      const double smp = cn - cs + ce - cw;
      const double cp = tmax(0.0, smp);
      const double cpo = cp;
      double AN = tmax(fabs(0.5 * cn), dn) - 0.5 * cn;
      double AS = tmax(fabs(0.5 * cs), ds) + 0.5 * cs;
      double AE = tmax(fabs(0.5 * ce), de) - 0.5 * ce;
      double AW = tmax(fabs(0.5 * cw), dw) + 0.5 * cw;
      double DU = areaew;
      double SU = cpo * U(i, j) + DU * (Pr(i - 1, j) - Pr(i, j));
      double SP = -cp;
      const double dudxp = (U(i + 1, j) - U(i, j)) / T1(P.sew, i);
      const double dudxm = (U(i, j) - U(i - 1, j)) / T1(P.sew, i - 1);
      SU = (S(i, j) * dudxp - S(i - 1, j) * dudxm) / T1(P.sewu, i) * vol + SU;
      const double gamp = 0.25 * (S(i, j) + S(i - 1, j) + S(i, j + 1) + S(i - 1, j + 1));
      const double dvdxp = TJ(P.rv, j + 1) * (V(i, j + 1) - V(i - 1, j + 1)) / T1(P.dxep, i);
      const double gamm = 0.25 * (S(i, j) + S(i - 1, j) + S(i, j - 1) + S(i - 1, j - 1));
      const double dvdxm = TJ(P.rv, j) * (V(i, j) - V(i - 1, j)) / T1(P.dxep, i);
      SU = SU + (gamp * dvdxp - gamm * dvdxm) / TJ(P.sns, j) / TJ(P.r, j) * vol;
      // ---- modu (promod.f90:36-77), laminar wall law: yplus = 11 <= 11.63 -> tmult = viscos/yp
      if (j == njm1) {
        const double yp = TJ(P.yv, nj) - TJ(P.y, njm1);
        const double tmult = P.tmult_n ? __ldg(P.tmult_n + i - 1) : P.viscos / yp;
        SP = SP - tmult * T1(P.sewu, i) * TJ(P.rv, nj);
        AN = 0.0;
      }
      if (i == 3 && j >= P.jstep + 1) AW = 0.0;
      if (j == 2) AS = 0.0;
      // ---- chapter 3 with the edited field values: u(i,1) = u(i,2), u(ni,j) = u(nim1,j) + uinc
      const double u_s = (j == 2) ? U(i, 2) : U(i, j - 1);
      const double u_e = (i == nim1) ? U(nim1, j) + uinc : U(i + 1, j);
      double AP = AN + AS + AE + AW - SP;
      DU = DU / AP;
      double resor = AN * U(i, j + 1) + AS * u_s + AE * u_e + AW * U(i - 1, j) - AP * U(i, j) + SU;
      vol = TJ(P.r, j) * T1(P.sew, i) * TJ(P.sns, j);
      const double sorvol = P.great * vol;
      if (-SP > 0.5 * sorvol) resor = resor / sorvol;
      racc += fabs(resor);
      AP = AP / P.urf;
      SU = SU + (1. - P.urf) * AP * U(i, j);
      DU = DU * P.urf;
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
      P.dout[(long long)i0 * LD + k * T + t] = DU;
    }
  }
  block_sum_teach(racc, P.partial);
}

// ---- v: calc.f90:145-220 + modv coefficient edits -------------------------------------------------------------
__global__ void __launch_bounds__(256) k_teach_assemble_v(const TeachParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj, njm1 = nj - 1;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const TF U{P.u, LD, lp}, V{P.v, LD, lp}, D{P.den, LD, lp}, S{P.vis, LD, lp}, Pr{P.p, LD, lp};
    const int i = i0 + 1;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      const double arean = TJ(P.rcv, j + 1) * T1(P.sew, i);
      const double areas = TJ(P.rcv, j) * T1(P.sew, i);
      const double areaew = TJ(P.rv, j) * TJ(P.snsv, j);
      double vol = TJ(P.rv, j) * T1(P.sew, i) * TJ(P.snsv, j);
      const double gn = 0.5 * (D(i, j + 1) + D(i, j)) * V(i, j + 1);
      const double gp = 0.5 * (D(i, j) + D(i, j - 1)) * V(i, j);
      const double gs = 0.5 * (D(i, j - 1) + D(i, j - 2)) * V(i, j - 1);
      const double ge = 0.5 * (D(i + 1, j) + D(i, j)) * U(i + 1, j);
      const double gse = 0.5 * (D(i, j - 1) + D(i + 1, j - 1)) * U(i + 1, j - 1);
      const double gw = 0.5 * (D(i, j) + D(i - 1, j)) * U(i, j);
      const double gsw = 0.5 * (D(i, j - 1) + D(i - 1, j - 1)) * U(i, j - 1);
      const double cn = 0.5 * (gn + gp) * arean;
      const double cs = 0.5 * (gp + gs) * areas;
      const double ce = 0.5 * (ge + gse) * areaew;
      const double cw = 0.5 * (gw + gsw) * areaew;
      const double vise = 0.25 * (S(i, j) + S(i + 1, j) + S(i, j - 1) + S(i + 1, j - 1));
      const double visw = 0.25 * (S(i, j) + S(i - 1, j) + S(i, j - 1) + S(i - 1, j - 1));
      const double dn = S(i, j) * arean / TJ(P.dynpv, j);
      const double ds = S(i, j - 1) * areas / TJ(P.dypsv, j);
      const double de = vise * areaew / T1(P.dxep, i);
      const double dw = visw * areaew / T1(P.dxpw, i);
      const double smp = cn - cs + ce - cw;
      const double cp = tmax(0.0, smp);
      const double cpo = cp;
      double AN = tmax(fabs(0.5 * cn), dn) - 0.5 * cn;
      double AS = tmax(fabs(0.5 * cs), ds) + 0.5 * cs;
      double AE = tmax(fabs(0.5 * ce), de) - 0.5 * ce;
      double AW = tmax(fabs(0.5 * cw), dw) + 0.5 * cw;
      double DV = 0.5 * (arean + areas);
      double SU = cpo * V(i, j) + DV * (Pr(i, j - 1) - Pr(i, j));
      double SP = -cp;
      if (P.indcos == 2) SP = SP - S(i, j) * vol / (TJ(P.rv, j) * TJ(P.rv, j));   // applied twice in the reference
      if (P.indcos == 2) SP = SP - S(i, j) * vol / (TJ(P.rv, j) * TJ(P.rv, j));   // (calc.f90:183-184)
      const double dudyp = (U(i + 1, j) - U(i + 1, j - 1)) / TJ(P.dyps, j);
      const double gamp = 0.25 * (S(i, j) + S(i + 1, j) + S(i, j - 1) + S(i + 1, j - 1));
      const double gamm = 0.25 * (S(i, j) + S(i - 1, j) + S(i, j - 1) + S(i - 1, j - 1));
      const double dudym = (U(i, j) - U(i, j - 1)) / TJ(P.dyps, j);
      SU = SU + (gamp * dudyp - gamm * dudym) / T1(P.sew, i) * vol;
      const double dvdyp = (V(i, j + 1) - V(i, j)) / TJ(P.sns, j);
      const double rgamp = S(i, j) * TJ(P.r, j);
      const double dvdym = (V(i, j) - V(i, j - 1)) / TJ(P.sns, j - 1);
      const double rgamm = S(i, j - 1) * TJ(P.r, j - 1);
      SU = SU + (rgamp * dvdyp - rgamm * dvdym) / (TJ(P.r, j) * TJ(P.sns, j)) * vol;
      // ---- modv (promod.f90:105-128)
      if (i == P.istep && j >= P.jstep + 1) {
        const double xp = T1(P.x, P.istep) - T1(P.xu, P.istep);
        const double tmult = P.tmult_w ? __ldg(P.tmult_w + j - 1) : P.viscos / xp;
        SP = SP - tmult * TJ(P.snsv, j) * TJ(P.rv, j);
        AW = 0.0;
      }
      if (j == njm1 && i >= P.istep) AN = 0.0;
      if (j == 3) AS = 0.0;
      // ---- chapter 3
      double AP = AN + AS + AE + AW - SP;
      DV = DV / AP;
      double resor = AN * V(i, j + 1) + AS * V(i, j - 1) + AE * V(i + 1, j) + AW * V(i - 1, j) - AP * V(i, j) + SU;
      vol = TJ(P.r, j) * T1(P.sew, i) * TJ(P.sns, j);
      const double sorvol = P.great * vol;
      if (-SP > 0.5 * sorvol) resor = resor / sorvol;
      racc += fabs(resor);
      AP = AP / P.urf;
      SU = SU + (1. - P.urf) * AP * V(i, j);
      DV = DV * P.urf;
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
      P.dout[(long long)i0 * LD + k * T + t] = DV;
    }
  }
  block_sum_teach(racc, P.partial);
}

// ---- p': calc.f90:241-280 -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_teach_assemble_p(const TeachParams P) {
  const int M = P.L.M, T = P.L.T;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib) {
    const LinePos lp(M, T, t);
    const TF U{P.u, LD, lp}, V{P.v, LD, lp}, D{P.den, LD, lp}, DU{P.du_in, LD, lp}, DV{P.dv_in, LD, lp};
    const int i = i0 + 1;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < P.jlo || j0 > P.jhi) continue;
      const double arean = TJ(P.rv, j + 1) * T1(P.sew, i);
      const double areas = TJ(P.rv, j) * T1(P.sew, i);
      const double areaew = TJ(P.r, j) * TJ(P.sns, j);
      const double denn = 0.5 * (D(i, j) + D(i, j + 1));
      const double dens = 0.5 * (D(i, j) + D(i, j - 1));
      const double dene = 0.5 * (D(i, j) + D(i + 1, j));
      const double denw = 0.5 * (D(i, j) + D(i - 1, j));
      const double AN = denn * arean * DV(i, j + 1);
      const double AS = dens * areas * DV(i, j);
      const double AE = dene * areaew * DU(i + 1, j);
      const double AW = denw * areaew * DU(i, j);
      const double cn = denn * V(i, j + 1) * arean;
      const double cs = dens * V(i, j) * areas;
      const double ce = dene * U(i + 1, j) * areaew;
      const double cw = denw * U(i, j) * areaew;
      const double smp = cn - cs + ce - cw;
      racc += fabs(smp);
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AN + AS + AE + AW - 0.0; P.su[cq] = -smp;
    }
  }
  block_sum_teach(racc, P.partial);
}

}  // namespace cfd

using namespace cfd;

// which: 0 u, 1 v, 2 p'
int cfd_launch_teach_assemble(cfd_ctx* c, int which) {
  TeachParams P{};
  P.L = c->L;
  P.indcos = c->cfg.indcos;
  P.istep = c->cfg.istep;
  P.jstep = c->cfg.jstep;
  P.viscos = c->cfg.viscos;
  P.great = c->cfg.great;
  P.flowin = c->cfg.flowin;
  P.u = c->field[CFD_U]; P.v = c->field[CFD_V]; P.p = c->field[CFD_P]; P.den = c->field[CFD_DEN]; P.vis = c->field[CFD_VIS];
  P.du_in = c->field[CFD_DU]; P.dv_in = c->field[CFD_DV];
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP]; P.su = c->field[CFD_SU];
  P.ufield = c->field[CFD_U];
  P.uinc = c->res_dev + 8;          // device scalar behind the 8 residual slots the host sees (CFD_RES_SLOTS)
  P.tmult_n = c->turbulent ? c->tmult_n : nullptr;
  P.tmult_w = c->turbulent ? c->tmult_w : nullptr;
  P.x = c->geomn[CFD_G_X]; P.y = c->geom[CFD_G_Y]; P.xu = c->geomn[CFD_G_XU]; P.yv = c->geom[CFD_G_YV];
  P.r = c->geom[CFD_G_R]; P.rv = c->geom[CFD_G_RV]; P.rcv = c->geom[CFD_G_RCV];
  P.sew = c->geomn[CFD_G_SEW]; P.sns = c->geom[CFD_G_SNS]; P.sewu = c->geomn[CFD_G_SEWU]; P.snsv = c->geom[CFD_G_SNSV];
  P.dxep = c->geomn[CFD_G_DXEP]; P.dxpw = c->geomn[CFD_G_DXPW]; P.dynp = c->geom[CFD_G_DYNP]; P.dyps = c->geom[CFD_G_DYPS];
  P.dxepu = c->geomn[CFD_G_DXEPU]; P.dxpwu = c->geomn[CFD_G_DXPWU]; P.dynpv = c->geom[CFD_G_DYNPV];
  P.dypsv = c->geom[CFD_G_DYPSV];
  const int ni = c->L.ni, nj = c->L.nj;
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : ni, goff = c->cfg.goff;
  P.jhi = nj - 2;
  // global 1-based line ranges 3..NI-1 (u) / 2..NI-1 (v, p'), clamped to the lines this context owns
  const int lo_own = c->cfg.halo_w, hi_own = ni - 1 - c->cfg.halo_e;
  P.ia = (which == 0 ? 2 : 1) - goff;
  P.ib = NI - 2 - goff;
  if (P.ia < lo_own) P.ia = lo_own;
  if (P.ib > hi_own) P.ib = hi_own;
  P.has_outlet = (goff + ni == NI);
  P.nsym = (NI - 1 - goff < ni) ? NI - 1 - goff : ni;
  if (which == 0) {
    P.jlo = 1; P.dout = c->field[CFD_DU];
  } else if (which == 1) {
    P.jlo = 2; P.dout = c->field[CFD_DV];
  } else {
    P.jlo = 1;
  }
  if (P.ib < P.ia) {   // a slab that owns none of this equation's lines still publishes a zero residual
    CFD_CUDA(cudaMemsetAsync(c->res_dev + which, 0, sizeof(double), c->stream));
    return CFD_OK;
  }
  P.urf = c->cfg.urf[which];
  const int nlines = P.ib - P.ia + 1;
  dim3 block(32, 8), grid((c->L.T + 31) / 32, (nlines + 7) / 8);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  P.partial = c->partial + (size_t)which * c->max_partials;
  if (!c->exact_assembly) {   // production (red/black) contexts: teach_fast_kernels.cu, equations stored divided by ap
    if (which == 0 && P.has_outlet) {
      k_teach_outlet<<<1, 1024, 0, c->stream>>>(P);
      ++c->launches;
    }
    const int nb = cfd_launch_teach_assemble_fast(c, which, P.ia, P.ib, P.partial, P.uinc);
    if (nb < 0) return nb;
    if (which == 0) {
      const int n = (ni > nj ? ni : nj) + 1;
      k_teach_modu_fields<<<(n + 127) / 128, 128, 0, c->stream>>>(P);
      ++c->launches;
    }
    CFD_CUDA(cudaGetLastError());
    return cfd_launch_reduce_partials(c, which, nb);
  }
  if (which == 0) {
    k_teach_outlet<<<1, 1024, 0, c->stream>>>(P);
    k_teach_assemble_u<<<grid, block, 0, c->stream>>>(P);
    const int n = (ni > nj ? ni : nj) + 1;
    k_teach_modu_fields<<<(n + 127) / 128, 128, 0, c->stream>>>(P);
    c->launches += 3;
  } else if (which == 1) {
    k_teach_assemble_v<<<grid, block, 0, c->stream>>>(P);
    ++c->launches;
  } else {
    k_teach_assemble_p<<<grid, block, 0, c->stream>>>(P);
    ++c->launches;
  }
  CFD_CUDA(cudaGetLastError());
  return cfd_launch_reduce_partials(c, which, nblocks);
}
