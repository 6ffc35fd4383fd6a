// teach_turb_kernels.cu -- k-epsilon model, energy equation and wall functions of the TEACH sudden-expansion
// program:
//   calcte calc.f90:388-485 + modte promod.f90:182-256,  calced :490-577 + moded :260-296,
//   calct  :309-385 + modt :146-178,  props general.f90:165-190,
//   the wall-function branches of modu / modv (promod.f90:36-60, 105-121) as a per-line / per-row pre-pass.
//
// As in teach_kernels.cu the reference's chapters 1-3 are one kernel per equation: every cell builds its
// convection / diffusion coefficients, applies the hook edits that concern it (top wall row njm1, step wall column
// istep, symmetry row 2) in the order the reference's loops would have applied them, then does the chapter-3 work.
// Per-cell side effects of the hooks stay with the cell's thread: gen(i,j) (read later by calced), yplusn(i) by the
// top-row cell, xplusw(j) by the step-column cell, te(i,1) = te(i,2) by the row-2 cell (its residual uses the edited
// value).  Compiled with -fmad=false and the reference's operation order; log / pow / sqrt come from the CUDA math
// library, so agreement with oracle/teach_turb.inc is to rounding, not bitwise.  Works for both coefficient
// storages (nc = 6 raw, nc = 5 divided by ap).
#include <cmath>
#include "common.cuh"

namespace cfd {

struct TurbParams {
  Layout L;
  int indcos, istep, jstep;
  int goff, ni_global;       // global 1-based i of local line i0 is i0 + 1 + goff (i-slab contexts)
  int il_lo, il_hi;          // owned interior lines (local, 0-based, inclusive)
  int halo_w, halo_e;
  double viscos, great, urf, rurf;
  double cdterm, cmu75;      // cmu^0.25, cmu^0.75 (the reference evaluates them per cell; constants of the model)
  cfd_turbulence m;
  const double *u, *v, *den;
  double *vis, *te, *ed, *t, *gen, *gamh;
  double *ae, *aw, *an, *as, *ap, *su;
  double *taun, *tauw, *yplusn, *xplusw, *tmult_n, *tmult_w;
  const double *x, *xu, *sew, *sewu, *dxep, *dxpw;                      // by-i, natural order
  const double *y, *yv, *r, *rv, *sns, *snsv, *dynp, *dyps;             // by-j, line layout
  const double *yn, *yvn;                                               // by-j, natural order (single entries)
  double* partial;
};

__device__ __forceinline__ double tq_max(double a, double b) { return a > b ? a : b; }

#define Q1(a, i) __ldg((a) + (i) - 1)

// field access relative to the current cell: at<DI, DJ>() = f(i+DI, j+DJ), DI, DJ in {-1, 0, 1}
struct RelQ {
  const double* c;      // start of line i
  long long LD;
  int qm, qc, qp;       // positions of rows j-1, j, j+1 inside a line
  template <int DI, int DJ>
  __device__ __forceinline__ double at() const {
    return __ldg(c + DI * LD + (DJ < 0 ? qm : (DJ > 0 ? qp : qc)));
  }
};

__device__ __forceinline__ void tq_block_sum(double v, double* partial) {
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

// ---- wall-function pre-pass of modu (which = 0, threads over i) / modv (which = 1, threads over j) ----------------
__global__ void k_teach_wall(const TurbParams P, int which) {
  const int ni = P.L.ni, nj = P.L.nj, njm1 = nj - 1, M = P.L.M, T = P.L.T;
  const long long LD = P.L.LD;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  auto F = [&](const double* f, int i, int j) { return f[(long long)(i - 1) * LD + lpos(j - 1, M, T)]; };
  const double cdterm = P.cdterm;
  if (which == 0) {
    // local 1-based line i; taun(i) needs the lines i-1, i, so it exists for every owned line and for the east halo
    // line next to the slab (calcte reads taun(i+1)); global tests through ig
    const int i = n, ig = n + P.goff, NI = P.ni_global;
    if (ig < 3 || ig > NI - 1 || i < 2 || i - 1 < P.il_lo || i - 1 > P.il_hi + (P.halo_e ? 1 : 0)) return;
    const int j = njm1;
    const double yp = P.yvn[nj - 1] - P.yn[njm1 - 1];
    const double sqrtk = sqrt(0.5 * (F(P.te, i, j) + F(P.te, i - 1, j)));
    const double denu = 0.5 * (F(P.den, i, j) + F(P.den, i - 1, j));
    const double yplusa = 0.5 * (P.yplusn[i - 1] + P.yplusn[i - 2]);
    double tmult;
    if (yplusa <= 11.63) {
      tmult = P.viscos / yp;
    } else {
      tmult = denu * cdterm * sqrtk * P.m.cappa / log(P.m.elog * yplusa);
    }
    const double tau = -tmult * F(P.u, i, j);
    P.tmult_n[i - 1] = tmult;
    P.taun[i - 1] = tau;
    if (ig == 3 && i >= 2) P.taun[i - 2] = tau;          // taun(2) = taun(3)
    if (ig == NI - 1 && i < ni) P.taun[i] = tau;         // taun(ni) = taun(nim1)
  } else {
    const int j = n, i = P.istep - P.goff, jstp1 = P.jstep + 1;      // the slab that owns line istep
    if (i - 1 < P.il_lo || i - 1 > P.il_hi) return;
    if (j < jstp1 || j > njm1) return;
    const double xp = P.x[i - 1] - P.xu[i - 1];                         // by-i geometry is the slab's slice
    const double sqrtk = sqrt(0.5 * (F(P.te, i, j) + F(P.te, i, j - 1)));
    const double denv = 0.5 * (F(P.den, i, j) + F(P.den, i, j - 1));
    const double xplusa = 0.5 * (P.xplusw[j - 1] + P.xplusw[j - 2]);
    double tmult;
    if (xplusa <= 11.63) {
      tmult = P.viscos / xp;
    } else {
      tmult = denv * cdterm * sqrtk * P.m.cappa / log(P.m.elog * xplusa);
    }
    const double tau = -tmult * F(P.v, i, j);
    P.tmult_w[j - 1] = tmult;
    P.tauw[j - 1] = tau;
    if (j == jstp1) P.tauw[P.jstep - 1] = tau;   // tauw(jstep) = tauw(jstp1)
    if (j == njm1) P.tauw[nj - 1] = tau;          // tauw(nj) = tauw(njm1)
  }
}

// ---- props: general.f90:165-190 -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_teach_props(const TurbParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.il_lo + blockIdx.y * blockDim.y + threadIdx.y;
  if (t >= T || i0 > P.il_hi) return;
  const long long o = (long long)i0 * LD;
  for (int k = 0; k < M; ++k) {
    const int j0 = t * M + k, q = k * T + t;
    if (j0 < 1 || j0 > nj - 2) continue;
    const double visold = P.vis[o + q];
    if (j0 == 1) P.vis[o + lpos(0, M, T)] = visold;           // vis(i,1) = vis(i,2), the value before the update
    double vnew;
    const double ed = P.ed[o + q], te = P.te[o + q];
    if (ed == 0.) {
      vnew = P.viscos;
    } else {
      vnew = __ldg(P.den + o + q) * (te * te) * P.m.cmu / ed + P.viscos;
    }
    vnew = P.m.urfvis * vnew + (1. - P.m.urfvis) * visold;
    P.vis[o + q] = vnew;
    P.gamh[o + q] = P.viscos / P.m.prandl + (vnew - P.viscos) / P.m.prandt;
  }
}

// ---- the three scalar equations --------------------------------------------------------------------------------
// EQ: CFD_EQ_TE, CFD_EQ_ED, CFD_EQ_T
// FAST (production, red/black contexts): every IEEE division (about 25 instructions each, some 25 per cell) becomes a
// multiplication by a reciprocal from the hardware seed + two Newton steps (5 instructions, ~1 ulp), and the five
// normalising divisions by ap share one reciprocal.  Same formulas, results equal to rounding (tested at 1e-9 over 20
// passes against the oracle); the sequential-order contexts keep the exact form.
__device__ __forceinline__ double tq_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}
#define TQ_DIV(a, b) (FAST ? (a) * tq_rcp(b) : (a) / (b))
template <int EQ, bool FAST>
__global__ void __launch_bounds__(256) k_teach_scalar(const TurbParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj, njm1 = nj - 1;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.il_lo + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.il_hi) {
    // i: local 1-based line (geometry, wall arrays); ig: global (the tests against istep and the inlet plane)
    const int i = i0 + 1, ig = i + P.goff, jstp1 = P.jstep + 1, istep = P.istep;
    double* phi_w = EQ == CFD_EQ_TE ? P.te : (EQ == CFD_EQ_ED ? P.ed : P.t);
    const long long oline = (long long)i0 * LD;
    const double pr = EQ == CFD_EQ_TE ? P.m.prte : P.m.pred;
    const double rpr = 1.0 / pr;
    const double cdterm = P.cdterm, cmu75 = P.cmu75;
    for (int k = 0; k < M; ++k) {
      const int j0 = t * M + k, j = j0 + 1;
      if (j0 < 1 || j0 > nj - 2) continue;
      // positions of rows j-1, j, j+1 in the line; every access below is base + {-LD, 0, +LD} + one of them
      const int qc = k * T + t, qp = (k + 1 < M) ? qc + T : t + 1, qm = (k > 0) ? qc - T : (M - 1) * T + t - 1;
      const RelQ U{P.u + oline, LD, qm, qc, qp}, V{P.v + oline, LD, qm, qc, qp}, D{P.den + oline, LD, qm, qc, qp};
      const RelQ S{P.vis + oline, LD, qm, qc, qp}, TE{P.te + oline, LD, qm, qc, qp}, PHI{phi_w + oline, LD, qm, qc, qp};
      const RelQ GM{(EQ == CFD_EQ_T ? P.gamh : P.vis) + oline, LD, qm, qc, qp};
      // ---- chapter 1 (calc.f90:321-352 / 402-425 / 503-526)
      const double arean = __ldg(P.rv + qp) * Q1(P.sew, i);
      const double areas = __ldg(P.rv + qc) * Q1(P.sew, i);
      const double areaew = __ldg(P.r + qc) * __ldg(P.sns + qc);
      double vol = __ldg(P.r + qc) * __ldg(P.sns + qc) * Q1(P.sew, i);
      const double gn = 0.5 * (D.at<0, 0>() + D.at<0, 1>()) * V.at<0, 1>();
      const double gs = 0.5 * (D.at<0, 0>() + D.at<0, -1>()) * V.at<0, 0>();
      const double ge = 0.5 * (D.at<0, 0>() + D.at<1, 0>()) * U.at<1, 0>();
      const double gw = 0.5 * (D.at<0, 0>() + D.at<-1, 0>()) * U.at<0, 0>();
      const double cn = gn * arean, cs = gs * areas, ce = ge * areaew, cw = gw * areaew;
      double gamn = 0.5 * (GM.at<0, 0>() + GM.at<0, 1>());
      double gams = 0.5 * (GM.at<0, 0>() + GM.at<0, -1>());
      double game = 0.5 * (GM.at<0, 0>() + GM.at<1, 0>());
      double gamw = 0.5 * (GM.at<0, 0>() + GM.at<-1, 0>());
      if (EQ != CFD_EQ_T) {
        if (FAST) {
          gamn = gamn * rpr; gams = gams * rpr; game = game * rpr; gamw = gamw * rpr;
        } else {
          gamn = gamn / pr; gams = gams / pr; game = game / pr; gamw = gamw / pr;
        }
      }
      const double dn = TQ_DIV(gamn * arean, __ldg(P.dynp + qc));
      const double ds = TQ_DIV(gams * areas, __ldg(P.dyps + qc));
      const double de = TQ_DIV(game * areaew, Q1(P.dxep, i));
      const double dw = TQ_DIV(gamw * areaew, Q1(P.dxpw, i));
      const double smp = cn - cs + ce - cw;
      const double cp = tq_max(0.0, smp);
      const double cpo = cp;
      double AN = tq_max(fabs(0.5 * cn), dn) - 0.5 * cn;
      double AS = tq_max(fabs(0.5 * cs), ds) + 0.5 * cs;
      double AE = tq_max(fabs(0.5 * ce), de) - 0.5 * ce;
      double AW = tq_max(fabs(0.5 * cw), dw) + 0.5 * cw;
      const double phi_c = PHI.at<0, 0>();
      double SU = cpo * phi_c;
      double SP = -cp;
      double phi_s = PHI.at<0, -1>();
      const long long oq = (long long)i0 * LD + k * T + t;
      if (EQ == CFD_EQ_TE) {
        // velocity gradients at the cell centre (calc.f90:430-436; the same expressions in promod.f90:199-200, 227-228)
        const double rsew_f = FAST ? tq_rcp(Q1(P.sew, i)) : 0.0, rsns_f = FAST ? tq_rcp(__ldg(P.sns + qc)) : 0.0;
        const double dudx = FAST ? (U.at<1, 0>() - U.at<0, 0>()) * rsew_f : (U.at<1, 0>() - U.at<0, 0>()) / Q1(P.sew, i);
        const double dvdy = FAST ? (V.at<0, 1>() - V.at<0, 0>()) * rsns_f : (V.at<0, 1>() - V.at<0, 0>()) / __ldg(P.sns + qc);
        const double dudy_n = (U.at<0, 0>() + U.at<1, 0>() + U.at<0, 1>() + U.at<1, 1>()) / 4.
                              - (U.at<0, 0>() + U.at<1, 0>() + U.at<0, -1>() + U.at<1, -1>()) / 4.;
        const double dvdx_n = (V.at<0, 0>() + V.at<0, 1>() + V.at<1, 0>() + V.at<1, 1>()) / 4.
                              - (V.at<0, 0>() + V.at<0, 1>() + V.at<-1, 0>() + V.at<-1, 1>()) / 4.;
        const double dudy = FAST ? dudy_n * rsns_f : dudy_n / __ldg(P.sns + qc);
        const double dvdx = FAST ? dvdx_n * rsew_f : dvdx_n / Q1(P.sew, i);
        const double vis_c = S.at<0, 0>(), den_c = D.at<0, 0>();
        double gen = (2. * (dudx * dudx + dvdy * dvdy) + (dudy + dvdx) * (dudy + dvdx)) * vis_c;
        const double v_c0 = V.at<0, 0>(), v_n0 = V.at<0, 1>();
        double vdr = TQ_DIV(v_c0, __ldg(P.rv + qc));
        if (j == 2) vdr = 0.0;
        if (P.indcos == 2) {
          const double h = vdr + TQ_DIV(v_n0, __ldg(P.rv + qp));
          gen = gen + vis_c * 0.5 * (h * h);
        }
        const double sukd = SU;
        SU = SU + gen * vol;
        const double spkd = SP;
        SP = SP - TQ_DIV(P.m.cd * P.m.cmu * (den_c * den_c) * phi_c * vol, vis_c);
        // ---- modte (promod.f90:182-256): top wall, then step wall, then symmetry axis
        if (j == njm1) {
          const double yp = P.yvn[nj - 1] - P.yn[njm1 - 1];
          const double sqrtk = sqrt(phi_c);
          const double gencou = 0.5 * (fabs(P.taun[i] * U.at<1, 0>()) + fabs(P.taun[i - 1] * U.at<0, 0>())) / yp;
          const double yplus = den_c * sqrtk * cdterm * yp / P.viscos;
          P.yplusn[i - 1] = yplus;
          // i-slabs: the wall law of the next iteration reads yplusn of the halo line beside the slab; its owner
          // computes it from these same operands, so the neighbour restates it here instead of exchanging it
          if (P.halo_w && i0 == P.il_lo) P.yplusn[i - 2] = D.at<-1, 0>() * sqrt(PHI.at<-1, 0>()) * cdterm * yp / P.viscos;
          if (P.halo_e && i0 == P.il_hi) P.yplusn[i] = D.at<1, 0>() * sqrt(PHI.at<1, 0>()) * cdterm * yp / P.viscos;
          const double genres = gen - vis_c * (dudy * dudy);
          gen = genres + gencou;
          double diterm;
          if (yplus <= 11.63) {
            diterm = den_c * cmu75 * sqrtk * yplus / yp;
          } else {
            diterm = den_c * cmu75 * sqrtk * log(P.m.elog * yplus) / (P.m.cappa * yp);
          }
          SU = gen * vol + sukd;
          SP = -diterm * vol + spkd;
          AN = 0.0;
        }
        if (ig == istep && j >= jstp1) {
          const double xp = P.x[i - 1] - P.xu[i - 1];
          const double sqrtk = sqrt(phi_c);
          const double xplus = den_c * sqrtk * cdterm * xp / P.viscos;
          P.xplusw[j - 1] = xplus;
          const double gencou = 0.5 * (fabs(P.tauw[j] * V.at<0, 1>()) + fabs(P.tauw[j - 1] * V.at<0, 0>())) / xp;
          const double genres = gen - vis_c * (dvdx * dvdx);
          gen = genres + gencou;
          double diterm;
          if (xplus <= 11.63) {
            diterm = den_c * cmu75 * sqrtk * xplus / xp;
          } else {
            diterm = den_c * cmu75 * sqrtk * log(P.m.elog * xplus) / (P.m.cappa * xp);
          }
          SU = sukd + gen * vol;
          SP = spkd - diterm * vol;
          AW = 0.0;
        }
        if (j == 2) {
          P.te[oline + lpos(0, M, T)] = phi_c;      // te(i,1) = te(i,2)
          phi_s = phi_c;
          gen = gen - vis_c * (dudy * dudy);
          SU = sukd + gen * vol;
          AS = 0.0;
        }
        P.gen[oq] = gen;
      } else if (EQ == CFD_EQ_ED) {
        const double vis_c = S.at<0, 0>(), den_c = D.at<0, 0>(), te_c = TE.at<0, 0>();
        SU = SU + TQ_DIV(P.m.c1 * P.m.cmu * __ldg(P.gen + oq) * vol * den_c * te_c, vis_c);
        SP = SP - TQ_DIV(P.m.c2 * den_c * phi_c * vol, te_c);
        // ---- moded (promod.f90:260-296): fixed values through the great-source trick
        if (j == njm1) {
          const double yp = P.yvn[nj - 1] - P.yn[njm1 - 1];
          const double term = cmu75 / (P.m.cappa * yp);
          SU = P.great * term * pow(te_c, 1.5);
          SP = -P.great;
        }
        if (ig == istep && j >= jstp1 && j <= nj - 2) {
          const double xp = P.x[i - 1] - P.xu[i - 1];
          const double term = cmu75 / (P.m.cappa * xp);
          SU = P.great * term * pow(te_c, 1.5);
          SP = -P.great;
        }
        if (j == 2) AS = 0.0;
      } else {
        // ---- modt (promod.f90:146-178)
        if (j == njm1) {
          const double dyn = P.yvn[nj - 1] - P.yn[njm1 - 1];
          AN = 0.0;
          const double area = __ldg(P.rv + qp) * Q1(P.sew, i);
          const double yplus = P.yplusn[i - 1];
          double gt;
          if (yplus <= 11.63) {
            gt = P.viscos / (P.m.prandl * dyn);
          } else {
            const double uplus = log(P.m.elog * yplus) / P.m.cappa;
            gt = D.at<0, 0>() * cdterm * sqrt(TE.at<0, 0>()) / (P.m.prandt * (uplus + P.m.pfun));
          }
          const double term = gt * area;
          SU = SU + term * P.m.twalln;
          SP = SP - term;
        }
        if (ig == 2 && j >= jstp1) AW = 0.0;
        if (j == 2) AS = 0.0;
      }
      // ---- chapter 3
      double AP = AN + AS + AE + AW - SP;
      double resor = AN * PHI.at<0, 1>() + AS * phi_s + AE * PHI.at<1, 0>() + AW * PHI.at<-1, 0>() - AP * phi_c + SU;
      vol = __ldg(P.r + qc) * Q1(P.sew, i) * __ldg(P.sns + qc);
      const double sorvol = P.great * vol;
      if (-SP > 0.5 * sorvol) resor = resor / sorvol;
      racc += fabs(resor);
      AP = FAST ? AP * P.rurf : AP / P.urf;
      SU = SU + (1. - P.urf) * AP * phi_c;
      const long long cq = (long long)i0 * P.L.cSL + (long long)k * P.L.cSK + t;
      if (P.L.nc == 6) {
        P.ae[cq] = AE; P.aw[cq] = AW; P.an[cq] = AN; P.as[cq] = AS; P.ap[cq] = AP; P.su[cq] = SU;
      } else {
        if (FAST) {
          const double rAP = tq_rcp(AP);
          P.ae[cq] = AE * rAP; P.aw[cq] = AW * rAP; P.an[cq] = AN * rAP; P.as[cq] = AS * rAP; P.su[cq] = SU * rAP;
        } else {
          P.ae[cq] = AE / AP; P.aw[cq] = AW / AP; P.an[cq] = AN / AP; P.as[cq] = AS / AP; P.su[cq] = SU / AP;
        }
      }
    }
  }
  tq_block_sum(racc, P.partial);
}

}  // namespace cfd

using namespace cfd;

static void fill_turb(const cfd_ctx* c, TurbParams& P) {
  P.L = c->L;
  P.indcos = c->cfg.indcos; P.istep = c->cfg.istep; P.jstep = c->cfg.jstep;
  P.viscos = c->cfg.viscos; P.great = c->cfg.great;
  P.goff = c->cfg.goff;
  P.ni_global = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  P.halo_w = c->cfg.halo_w; P.halo_e = c->cfg.halo_e;
  // owned lines without the two global boundary lines
  P.il_lo = c->cfg.halo_w > 0 ? c->cfg.halo_w : (c->cfg.goff == 0 ? 1 : 0);
  P.il_hi = c->L.ni - 1 - c->cfg.halo_e - ((c->cfg.goff + c->L.ni == P.ni_global) ? 1 : 0);
  P.m = c->turb;
  P.cdterm = std::pow(c->turb.cmu, 0.25);
  P.cmu75 = std::pow(c->turb.cmu, .75);
  P.u = c->field[CFD_U]; P.v = c->field[CFD_V]; P.den = c->field[CFD_DEN];
  P.vis = c->field[CFD_VIS]; P.te = c->field[CFD_TE]; P.ed = c->field[CFD_ED]; P.t = c->field[CFD_T];
  P.gen = c->field[CFD_GEN]; P.gamh = c->field[CFD_GAMH];
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP]; P.su = c->field[CFD_SU];
  P.taun = c->wall[CFD_WALL_TAUN]; P.tauw = c->wall[CFD_WALL_TAUW];
  P.yplusn = c->wall[CFD_WALL_YPLUSN]; P.xplusw = c->wall[CFD_WALL_XPLUSW];
  P.tmult_n = c->tmult_n; P.tmult_w = c->tmult_w;
  P.x = c->geomn[CFD_G_X]; P.xu = c->geomn[CFD_G_XU]; P.sew = c->geomn[CFD_G_SEW]; P.sewu = c->geomn[CFD_G_SEWU];
  P.dxep = c->geomn[CFD_G_DXEP]; P.dxpw = c->geomn[CFD_G_DXPW];
  P.y = c->geom[CFD_G_Y]; P.yv = c->geom[CFD_G_YV]; P.r = c->geom[CFD_G_R]; P.rv = c->geom[CFD_G_RV];
  P.sns = c->geom[CFD_G_SNS]; P.snsv = c->geom[CFD_G_SNSV]; P.dynp = c->geom[CFD_G_DYNP]; P.dyps = c->geom[CFD_G_DYPS];
  P.yn = c->geomn[CFD_G_Y]; P.yvn = c->geomn[CFD_G_YV];
}

// which: 0 = modu's wall law (before the u assembly), 1 = modv's
int cfd_launch_teach_wall(cfd_ctx* c, int which) {
  TurbParams P{};
  fill_turb(c, P);
  const int n = (which == 0 ? c->L.ni : c->L.nj) + 1;
  k_teach_wall<<<(n + 127) / 128, 128, 0, c->stream>>>(P, which);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

int cfd_launch_teach_props(cfd_ctx* c) {
  TurbParams P{};
  fill_turb(c, P);
  dim3 block(32, 8), grid((c->L.T + 31) / 32, (P.il_hi - P.il_lo + 1 + 7) / 8);
  k_teach_props<<<grid, block, 0, c->stream>>>(P);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

// eq: CFD_EQ_TE, CFD_EQ_ED, CFD_EQ_T -- assembly (chapters 1-3) and the residual sum
int cfd_launch_teach_scalar(cfd_ctx* c, int eq) {
  TurbParams P{};
  fill_turb(c, P);
  P.urf = eq == CFD_EQ_TE ? c->turb.urfk : (eq == CFD_EQ_ED ? c->turb.urfe : c->turb.urft);
  P.rurf = 1.0 / P.urf;
  const bool fast = !c->exact_assembly;     // red/black contexts: reciprocal form of the same formulas
  dim3 block(32, 8), grid((c->L.T + 31) / 32, (P.il_hi - P.il_lo + 1 + 7) / 8);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  P.partial = c->partial + (size_t)eq * c->max_partials;
  if (eq == CFD_EQ_TE) {
    if (fast) k_teach_scalar<CFD_EQ_TE, true><<<grid, block, 0, c->stream>>>(P);
    else k_teach_scalar<CFD_EQ_TE, false><<<grid, block, 0, c->stream>>>(P);
  } else if (eq == CFD_EQ_ED) {
    if (fast) k_teach_scalar<CFD_EQ_ED, true><<<grid, block, 0, c->stream>>>(P);
    else k_teach_scalar<CFD_EQ_ED, false><<<grid, block, 0, c->stream>>>(P);
  } else {
    if (fast) k_teach_scalar<CFD_EQ_T, true><<<grid, block, 0, c->stream>>>(P);
    else k_teach_scalar<CFD_EQ_T, false><<<grid, block, 0, c->stream>>>(P);
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return cfd_launch_reduce_partials(c, eq, nblocks);
}
