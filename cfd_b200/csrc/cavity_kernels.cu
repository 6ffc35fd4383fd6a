// cavity_kernels.cu -- coefficient assembly and pressure/velocity correction of the cavity family
// (lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90: calcu 92-187, calcv 194-285, calcp 293-354).
//
// One thread owns M consecutive nodes of one S-N line and marches along them with three-wide register
// windows (j-1, j, j+1) for every array whose j-neighbours the stencil needs; i-neighbours are the same
// positions of the adjacent lines, fetched through L1 (the 8 lines of a block share them).  All loads and
// stores of a warp are 32 consecutive doubles.  Pass 1 (coefficients) and pass 2 (ap, residual, under-
// relaxation, du) of the reference are fused: 5 field reads + 7 writes per cell, the algorithmic 12 doubles.
// Arithmetic keeps the reference's operation order; the file is compiled with -fmad=false so results are
// bit-identical to the CPU restatement (oracle/cavity_oracle.c) apart from the order of the residual sum.
#include "common.cuh"

namespace cfd {

__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }

// Three-wide sliding window along a line held by one thread.
struct Win3 {
  double m, c, p;
};

struct LineCursor {
  int t, M, T, nj, j0base;
  __device__ __forceinline__ int pos(int k) const { return k * T + t; }
  __device__ __forceinline__ int pos_next(int k) const { return (k + 1 < M) ? (k + 1) * T + t : t + 1; }
  // window start: m = node j0base-1 (if it exists), c = node j0base
  __device__ __forceinline__ void init(Win3& w, const double* line) const {
    w.m = (j0base >= 1) ? __ldg(line + (M - 1) * T + t - 1) : 0.0;   // node t*M-1 = last node of thread t-1
    w.c = __ldg(line + t);
    w.p = 0.0;
  }
  // fetch node j0+1 (if it exists) into w.p
  __device__ __forceinline__ void fetch(Win3& w, const double* line, int k) const {
    w.p = (j0base + k + 1 < nj) ? __ldg(line + pos_next(k)) : 0.0;
  }
  __device__ __forceinline__ static void shift(Win3& w) {
    w.m = w.c;
    w.c = w.p;
  }
};

// Deterministic block reduction of one double per thread; thread 0 writes partial[block].
__device__ __forceinline__ void block_sum_to(double v, double* partial) {
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

// hybrid / upwind / central coefficient pair selection, cavity_fvm.f90:141-156 (same order of operations)
#define CFD_SCHEME(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)  \
  if (iadv == 1) {                                                        \
    AE = de - 0.5 * ce;                                                   \
    AW = dw + 0.5 * cw;                                                   \
    AN = dn - 0.5 * cn;                                                   \
    AS = ds + 0.5 * cs;                                                   \
  } else if (iadv == 2) {                                                 \
    AE = de + dmax(0.0, -ce);                                             \
    AW = dw + dmax(0.0, cw);                                              \
    AN = dn + dmax(0.0, -cn);                                             \
    AS = ds + dmax(0.0, cs);                                              \
  } else {                                                                \
    AN = dmax(fabs(0.5 * cn), dn) - 0.5 * cn;                             \
    AS = dmax(fabs(0.5 * cs), ds) + 0.5 * cs;                             \
    AE = dmax(fabs(0.5 * ce), de) - 0.5 * ce;                             \
    AW = dmax(fabs(0.5 * cw), dw) + 0.5 * cw;                             \
  }

// ---------------------------------------------------------------------------------------------------
// u-momentum: cavity_fvm.f90:104-187
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_assemble_u(const CavParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const LineCursor lc{t, M, T, nj, t * M};
    const double* uW = P.u + (long long)(i0 - 1) * LD;
    const double* uC = uW + LD;
    const double* uE = uC + LD;
    const double* vW = P.v + (long long)(i0 - 1) * LD;
    const double* vC = vW + LD;
    const double* dWW = P.den + (long long)(i0 - 2) * LD;
    const double* dW = dWW + LD;
    const double* dC = dW + LD;
    const double* dE = dC + LD;
    const double* sW = P.vis + (long long)(i0 - 1) * LD;
    const double* sC = sW + LD;
    const double* pW = P.p + (long long)(i0 - 1) * LD;
    const double* pC = pW + LD;
    const long long out = (long long)i0 * LD;
    const long long cout = (long long)i0 * P.L.cSL;
    // i-geometry (1-based i = i0+1): sewu(i), xu(i+1)-xu(i), xu(i)-xu(i-1), sew(i), sew(i-1)
    const double sewu_i = __ldg(P.sewu + i0);
    const double dxe = __ldg(P.xu + i0 + 1) - __ldg(P.xu + i0);
    const double dxw = __ldg(P.xu + i0) - __ldg(P.xu + i0 - 1);
    const double sew_i = __ldg(P.sew + i0), sew_im1 = __ldg(P.sew + i0 - 1);
    const int iadv = P.iadv;
    const double urfu = P.urf, stale = P.stale_area;

    Win3 wu, wdW, wdC, wsW, wsC, wy, wvW, wvC;
    lc.init(wu, uC);
    lc.init(wdW, dW);
    lc.init(wdC, dC);
    lc.init(wsW, sW);
    lc.init(wsC, sC);
    lc.init(wy, P.yL);
    lc.init(wvW, vW);
    lc.init(wvC, vC);
    for (int k = 0; k < M; ++k) {
      const int j0 = lc.j0base + k;
      lc.fetch(wu, uC, k);
      lc.fetch(wdW, dW, k);
      lc.fetch(wdC, dC, k);
      lc.fetch(wsW, sW, k);
      lc.fetch(wsC, sC, k);
      lc.fetch(wy, P.yL, k);
      lc.fetch(wvW, vW, k);
      lc.fetch(wvC, vC, k);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int q = lc.pos(k);
        const int cq = k * P.L.cSK + t;
        const double u_w = __ldg(uW + q), u_e = __ldg(uE + q);
        const double den_ww = __ldg(dWW + q), den_e = __ldg(dE + q);
        const double p_w = __ldg(pW + q), p_c = __ldg(pC + q);
        const double sns_j = __ldg(P.snsL + q);
        const double arean = sewu_i, areas = sewu_i, areaw = sns_j, areae = sns_j;
        const double vol = sewu_i * sns_j;
        // mass fluxes around the u-cell (112-118)
        const double gee = 0.5 * (wdC.c + den_e) * u_e;
        const double gww = 0.5 * (den_ww + wdW.c) * u_w;
        const double gne = 0.5 * (wdC.c + wdC.p) * wvC.p;
        const double gnw = 0.5 * (wdW.c + wdW.p) * wvW.p;
        const double gse = 0.5 * (wdC.m + wdC.c) * wvC.c;
        const double gsw = 0.5 * (wdW.m + wdW.c) * wvW.c;
        const double gp = 0.5 * (wdW.c + wdC.c) * wu.c;
        const double ge = 0.5 * (gp + gee);
        const double gw = 0.5 * (gww + gp);
        const double gn = 0.5 * (gnw + gne);
        const double gs = 0.5 * (gsw + gse);
        const double ce = ge * areae, cw = gw * areaw, cn = gn * arean, cs = gs * areas;
        const double smp = ce - cw + cn - cs;
        const double cp = dmax(smp, 0.0);
        const double vise = wsC.c, visw = wsW.c;
        const double visn = 0.25 * (wsW.c + wsW.p + wsC.p + wsC.c);
        const double viss = 0.25 * (wsW.m + wsW.c + wsC.c + wsC.m);
        const double de = vise * areae / dxe;
        const double dw = visw * areaw / dxw;
        const double dn = visn * arean / (wy.p - wy.c);
        const double ds = viss * areas / (wy.c - wy.m);
        double AE, AW, AN, AS;
        CFD_SCHEME(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
        double SU = (p_w - p_c) * 0.5 * (areae + areaw);
        const double dudxe = (u_e - wu.c) / sew_i;
        const double dudxw = (wu.c - u_w) / sew_im1;
        const double dvdxn = (wvC.p - wvW.p) / sewu_i;
        const double dvdxs = (wvC.c - wvW.c) / sewu_i;
        SU = SU + ((vise * dudxe - visw * dudxw) / sewu_i + (visn * dvdxn - viss * dvdxs) / sns_j) * vol;
        SU = SU + cp * wu.c;
        const double SP = -cp;
        // second pass of the reference (175-187), fused
        double AP = AE + AW + AN + AS - SP;
        const double resor = AP * wu.c - AE * u_e - AW * u_w - AN * wu.p - AS * wu.m - SU;
        racc += fabs(resor);
        AP = AP / urfu;
        SU = SU + (1 - urfu) * AP * wu.c;
        P.ae[cout + cq] = AE;
        P.aw[cout + cq] = AW;
        P.an[cout + cq] = AN;
        P.as[cout + cq] = AS;
        P.ap[cout + cq] = AP;
        P.su[cout + cq] = SU;
        P.dout[out + q] = 0.5 * (stale + stale) / AP;
      }
      LineCursor::shift(wu);
      LineCursor::shift(wdW);
      LineCursor::shift(wdC);
      LineCursor::shift(wsW);
      LineCursor::shift(wsC);
      LineCursor::shift(wy);
      LineCursor::shift(wvW);
      LineCursor::shift(wvC);
    }
  }
  block_sum_to(racc, P.partial);
}

// ---------------------------------------------------------------------------------------------------
// v-momentum: cavity_fvm.f90:205-285
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_assemble_v(const CavParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const LineCursor lc{t, M, T, nj, t * M};
    const double* uC = P.u + (long long)i0 * LD;
    const double* uE = uC + LD;
    const double* vW = P.v + (long long)(i0 - 1) * LD;
    const double* vC = vW + LD;
    const double* vE = vC + LD;
    const double* dW = P.den + (long long)(i0 - 1) * LD;
    const double* dC = dW + LD;
    const double* dE = dC + LD;
    const double* sW = P.vis + (long long)(i0 - 1) * LD;
    const double* sC = sW + LD;
    const double* sE = sC + LD;
    const double* pC = P.p + (long long)i0 * LD;
    const long long out = (long long)i0 * LD;
    const long long cout = (long long)i0 * P.L.cSL;
    // i-geometry (i = i0+1): sew(i), x(i+1)-x(i), x(i)-x(i-1)
    const double sew_i = __ldg(P.sew + i0);
    const double dxe = __ldg(P.x + i0 + 1) - __ldg(P.x + i0);
    const double dxw = __ldg(P.x + i0) - __ldg(P.x + i0 - 1);
    const int iadv = P.iadv;
    const double urfv = P.urf, stale = P.stale_area;

    Win3 wuC, wuE, wv, wdW, wdC, wdE, wsW, wsC, wsE, wp, wyv, wsns;
    lc.init(wuC, uC);
    lc.init(wuE, uE);
    lc.init(wv, vC);
    lc.init(wdW, dW);
    lc.init(wdC, dC);
    lc.init(wdE, dE);
    lc.init(wsW, sW);
    lc.init(wsC, sC);
    lc.init(wsE, sE);
    lc.init(wp, pC);
    lc.init(wyv, P.yvL);
    lc.init(wsns, P.snsL);
    // den(i, j-2): one more value behind the window
    double den_mm = (lc.j0base >= 2) ? __ldg(dC + lpos(lc.j0base - 2, M, T)) : 0.0;
    for (int k = 0; k < M; ++k) {
      const int j0 = lc.j0base + k;
      lc.fetch(wuC, uC, k);
      lc.fetch(wuE, uE, k);
      lc.fetch(wv, vC, k);
      lc.fetch(wdW, dW, k);
      lc.fetch(wdC, dC, k);
      lc.fetch(wdE, dE, k);
      lc.fetch(wsW, sW, k);
      lc.fetch(wsC, sC, k);
      lc.fetch(wsE, sE, k);
      lc.fetch(wp, pC, k);
      lc.fetch(wyv, P.yvL, k);
      lc.fetch(wsns, P.snsL, k);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int q = lc.pos(k);
        const int cq = k * P.L.cSK + t;
        const double v_w = __ldg(vW + q), v_e = __ldg(vE + q);
        const double snsv_j = __ldg(P.snsvL + q);
        const double areae = snsv_j, areaw = snsv_j, arean = sew_i, areas = sew_i;
        const double vol = sew_i * snsv_j;
        const double gne = 0.5 * (wdC.c + wdE.c) * wuE.c;
        const double gnw = 0.5 * (wdW.c + wdC.c) * wuC.c;
        const double gnn = 0.5 * (wdC.c + wdC.p) * wv.p;
        const double gse = 0.5 * (wdC.m + wdE.m) * wuE.m;
        const double gsw = 0.5 * (wdW.m + wdC.m) * wuC.m;
        const double gss = 0.5 * (den_mm + wdC.m) * wv.m;
        const double gp = 0.5 * (wdC.m + wdC.c) * wv.c;
        const double ge = 0.5 * (gne + gse);
        const double gw = 0.5 * (gnw + gsw);
        const double gn = 0.5 * (gp + gnn);
        const double gs = 0.5 * (gss + gp);
        const double ce = ge * areae, cw = gw * areaw, cn = gn * arean, cs = gs * areas;
        const double smp = ce - cw + cn - cs;
        const double cp = dmax(smp, 0.0);
        const double vise = 0.25 * (wsC.m + wsC.c + wsE.c + wsE.m);
        const double visw = 0.25 * (wsW.m + wsW.c + wsC.c + wsC.m);
        const double visn = wsC.c, viss = wsC.m;
        const double de = vise * areae / dxe;
        const double dw = visw * areaw / dxw;
        const double dn = visn * arean / (wyv.p - wyv.c);
        const double ds = viss * areas / (wyv.c - wyv.m);
        double AE, AW, AN, AS;
        CFD_SCHEME(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
        double SU = (wp.m - wp.c) * 0.5 * (arean + areas);
        const double dudye = (wuE.c - wuE.m) / snsv_j;
        const double dudyw = (wuC.c - wuC.m) / snsv_j;
        const double dvdyn = (wv.p - wv.c) / wsns.c;
        const double dvdys = (wv.c - wv.m) / wsns.m;
        SU = SU + (vise * dudye - visw * dudyw) / sew_i * vol + (visn * dvdyn - viss * dvdys) / snsv_j * vol;
        const double SP = -cp;
        SU = SU + cp * wv.c;
        double AP = AE + AW + AN + AS - SP;
        const double resor = AP * wv.c - AE * v_e - AW * v_w - AN * wv.p - AS * wv.m - SU;
        racc += fabs(resor);
        AP = AP / urfv;
        SU = SU + (1.0 - urfv) * AP * wv.c;
        P.ae[cout + cq] = AE;
        P.aw[cout + cq] = AW;
        P.an[cout + cq] = AN;
        P.as[cout + cq] = AS;
        P.ap[cout + cq] = AP;
        P.su[cout + cq] = SU;
        P.dout[out + q] = 0.5 * (stale + stale) / AP;
      }
      den_mm = wdC.m;
      LineCursor::shift(wuC);
      LineCursor::shift(wuE);
      LineCursor::shift(wv);
      LineCursor::shift(wdW);
      LineCursor::shift(wdC);
      LineCursor::shift(wdE);
      LineCursor::shift(wsW);
      LineCursor::shift(wsC);
      LineCursor::shift(wsE);
      LineCursor::shift(wp);
      LineCursor::shift(wyv);
      LineCursor::shift(wsns);
    }
  }
  block_sum_to(racc, P.partial);
}

// ---------------------------------------------------------------------------------------------------
// p' equation: cavity_fvm.f90:301-340 (pp = 0 of line 342 is folded into the first sweep)
// ---------------------------------------------------------------------------------------------------
template <bool UNI>
__global__ void __launch_bounds__(256) k_assemble_p(const CavParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const LineCursor lc{t, M, T, nj, t * M};
    const double* uC = P.u + (long long)i0 * LD;
    const double* uE = uC + LD;
    const double* vC = P.v + (long long)i0 * LD;
    const double* dW = P.den + (long long)(i0 - 1) * LD;
    const double* dC = dW + LD;
    const double* dE = dC + LD;
    const double* duC = P.du_in + (long long)i0 * LD;
    const double* duE = duC + LD;
    const double* dvC = P.dv_in + (long long)i0 * LD;
        const long long cout = (long long)i0 * P.L.cSL;
    const double sew_i = __ldg(P.sew + i0);
    Win3 wd, wv, wdv;
    if (UNI) wd.m = wd.c = wd.p = P.den0; else lc.init(wd, dC);
    lc.init(wv, vC);
    lc.init(wdv, dvC);
    for (int k = 0; k < M; ++k) {
      const int j0 = lc.j0base + k;
      if (!UNI) lc.fetch(wd, dC, k);
      lc.fetch(wv, vC, k);
      lc.fetch(wdv, dvC, k);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const int q = lc.pos(k);
        const int cq = k * P.L.cSK + t;
        const double den_w = UNI ? P.den0 : __ldg(dW + q), den_e = UNI ? P.den0 : __ldg(dE + q);
        const double u_c = __ldg(uC + q), u_e = __ldg(uE + q);
        const double du_c = __ldg(duC + q), du_e = __ldg(duE + q);
        const double sns_j = __ldg(P.snsL + q);
        const double areae = sns_j, areaw = sns_j, arean = sew_i, areas = sew_i;
        const double dene = 0.5 * (wd.c + den_e);
        const double denw = 0.5 * (den_w + wd.c);
        const double denn = 0.5 * (wd.c + wd.p);
        const double dens = 0.5 * (wd.m + wd.c);
        const double AE = dene * areae * du_e;
        const double AW = denw * areaw * du_c;
        const double AN = denn * arean * wdv.p;
        const double AS = dens * areas * wdv.c;
        const double ge = dene * u_e, gw = denw * u_c, gn = denn * wv.p, gs = dens * wv.c;
        const double ce = ge * areae, cw = gw * areaw, cn = gn * arean, cs = gs * areas;
        const double smp = ce - cw + cn - cs;
        racc += fabs(smp);
        if (P.L.nc == 6) {
          P.ae[cout + cq] = AE;
          P.aw[cout + cq] = AW;
          P.an[cout + cq] = AN;
          P.as[cout + cq] = AS;
          P.ap[cout + cq] = AE + AW + AN + AS - 0.0;
          P.su[cout + cq] = -smp;
        } else {   // production contexts store the equation divided by ap
          const double APs = AE + AW + AN + AS;
          const double r = APs != 0.0 ? 1.0 / APs : 0.0;
          P.ae[cout + cq] = AE * r;
          P.aw[cout + cq] = AW * r;
          P.an[cout + cq] = AN * r;
          P.as[cout + cq] = AS * r;
          P.su[cout + cq] = -smp * r;
        }
      }
      if (!UNI) LineCursor::shift(wd);
      LineCursor::shift(wv);
      LineCursor::shift(wdv);
    }
  }
  block_sum_to(racc, P.partial);
}

// ---------------------------------------------------------------------------------------------------
// correction of p, u, v: cavity_fvm.f90:346-354
// ---------------------------------------------------------------------------------------------------
struct CorrParams {
  Layout L;
  int ia, ib, jlo, jhi;
  int i_skip;            // local 0-based line where u is NOT corrected (global i == 2), or -1
  int j_skip;            // local 0-based node where v is NOT corrected (global j == 2)
  double urfp;
  const double *pp, *du, *dv;
  const double* ppref;   // device address of pp(ipref,jpref)
  double *p, *u, *v;
};

__global__ void __launch_bounds__(256) k_correct(const CorrParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  if (!(t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo)) return;
  const double ppref = __ldg(P.ppref);
  const double* ppW = P.pp + (long long)(i0 - 1) * LD;
  const double* ppC = ppW + LD;
  const long long o = (long long)i0 * LD;
  const int j0base = t * M;
  double pp_m = (j0base >= 1) ? __ldg(ppC + (M - 1) * T + t - 1) : 0.0;
  for (int k = 0; k < M; ++k) {
    const int j0 = j0base + k, q = k * T + t;
    if (j0 >= nj) break;
    const double pp_c = __ldg(ppC + q);
    if (j0 >= P.jlo && j0 <= P.jhi) {
      P.p[o + q] = P.p[o + q] + P.urfp * (pp_c - ppref);
      if (i0 != P.i_skip) P.u[o + q] = P.u[o + q] + __ldg(P.du + o + q) * (__ldg(ppW + q) - pp_c);
      if (j0 != P.j_skip) P.v[o + q] = P.v[o + q] + __ldg(P.dv + o + q) * (pp_m - pp_c);
    }
    pp_m = pp_c;
  }
}

// Sum the per-block partials of one equation in a fixed order -> res[eq].
__global__ void k_reduce_partials(const double* partial, int n, double* out) {
  __shared__ double sm[1024];
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v += partial[i];
  sm[threadIdx.x] = v;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sm[0];
}

}  // namespace cfd

using namespace cfd;

// ---- host-side launchers ------------------------------------------------------------------------------
// Clamp a GLOBAL 1-based inclusive line range to the lines this context owns; returns 0-based local range.
static void owned_range(const cfd_ctx* c, int gi_lo, int gi_hi, int* ia, int* ib) {
  const int lo_own = c->cfg.halo_w, hi_own = c->L.ni - 1 - c->cfg.halo_e;   // local 0-based
  int a = gi_lo - 1 - c->cfg.goff, b = gi_hi - 1 - c->cfg.goff;
  if (a < lo_own) a = lo_own;
  if (b > hi_own) b = hi_own;
  *ia = a;
  *ib = b;
}

static void fill_common(const cfd_ctx* c, CavParams& P) {
  P.L = c->L;
  P.iadv = c->cfg.iadv;
  P.u = c->field[CFD_U];
  P.v = c->field[CFD_V];
  P.p = c->field[CFD_P];
  P.den = c->field[CFD_DEN];
  P.vis = c->field[CFD_VIS];
  // production kernels only: both property fields hold one value each (cfd_upload saw it)
  P.uniform = (!c->exact_assembly && c->uni_den && c->uni_vis) ? 1 : 0;
  P.den0 = c->den0;
  P.vis0 = c->vis0;
  P.du_in = c->field[CFD_DU];
  P.dv_in = c->field[CFD_DV];
  P.ae = c->field[CFD_AE];
  P.aw = c->field[CFD_AW];
  P.an = c->field[CFD_AN];
  P.as = c->field[CFD_AS];
  P.ap = c->field[CFD_AP];
  P.su = c->field[CFD_SU];
  P.x = c->geom[CFD_G_X];
  P.xu = c->geom[CFD_G_XU];
  P.sew = c->geom[CFD_G_SEW];
  P.sewu = c->geom[CFD_G_SEWU];
  P.yL = c->geom[CFD_G_Y];
  P.yvL = c->geom[CFD_G_YV];
  P.snsL = c->geom[CFD_G_SNS];
  P.snsvL = c->geom[CFD_G_SNSV];
}

static const int BX = 32, BY = 8;

// which: 0 = u, 1 = v, 2 = p'
static int launch_assemble(cfd_ctx* c, int which, CavParams& P, int eq) {
  const int nlines = P.ib - P.ia + 1;
  if (nlines <= 0 || P.jhi < P.jlo) {   // a slab that owns none of this equation's cells still publishes a zero residual
    CFD_CUDA(cudaMemsetAsync(c->res_dev + eq, 0, sizeof(double), c->stream));
    return CFD_OK;
  }
  const bool fast = !c->exact_assembly && which < 2;
  static int use_tma = -1;
  if (use_tma < 0) {
    const char* e = getenv("CFD_B200_ASM_TMA");
    use_tma = e ? atoi(e) : 1;
  }
  const bool jslab = c->cfg.halo_s > 0 || c->cfg.halo_n > 0;
  // TMA-staged persistent kernels (the 384-thread build: 168 registers, no spills): lines of 128..320 layout threads in
  // one block, longer lines split over 2 / 4 / 8 blocks of that size (a 768-thread build spilt and measured 17.1 ms per
  // iteration at 8192^2 against 11.2 ms with the register-window kernels, which remain the fallback).
  if (fast && use_tma && !jslab && cfd_assemble_tma_split(c->L.T) > 0) {
    P.partial = c->partial + (size_t)eq * c->max_partials;
    const int nb = cfd_launch_assemble_tma(c, which, P);
    if (nb < 0) return nb;
    k_reduce_partials<<<1, 1024, 0, c->stream>>>(P.partial, nb, c->res_dev + eq);
    ++c->launches;
    CFD_CUDA(cudaGetLastError());
    return CFD_OK;
  }
  const int by = fast ? 4 : BY;
  dim3 block(BX, by), grid((c->L.T + BX - 1) / BX, (nlines + by - 1) / by);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  P.partial = c->partial + (size_t)eq * c->max_partials;
  if (fast) {
    int rc = cfd_launch_assemble_fast(c, which, P, grid, block);
    if (rc) return rc;
  } else {
    if (which == 0) {
      k_assemble_u<<<grid, block, 0, c->stream>>>(P);
    } else if (which == 1) {
      k_assemble_v<<<grid, block, 0, c->stream>>>(P);
    } else {
      if (P.uniform) k_assemble_p<true><<<grid, block, 0, c->stream>>>(P);
      else k_assemble_p<false><<<grid, block, 0, c->stream>>>(P);
    }
    ++c->launches;
    CFD_CUDA(cudaGetLastError());
  }
  k_reduce_partials<<<1, 1024, 0, c->stream>>>(P.partial, nblocks, c->res_dev + eq);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

int cfd_launch_assemble_u(cfd_ctx* c) {
  CavParams P{};
  fill_common(c, P);
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  owned_range(c, 3, NI - 1, &P.ia, &P.ib);
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  cfd_local_j_range(c, 1, NJ - 2, &P.jlo, &P.jhi);
  P.urf = c->cfg.urf[CFD_EQ_U];
  P.stale_area = c->cfg.sns_njm1 != 0.0 ? c->cfg.sns_njm1 : c->geom_host[CFD_G_SNS][c->L.nj - 2];   // sns(njm1)
  P.dout = c->field[CFD_DU];
  return launch_assemble(c, 0, P, CFD_EQ_U);
}

int cfd_launch_assemble_v(cfd_ctx* c) {
  CavParams P{};
  fill_common(c, P);
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  owned_range(c, 2, NI - 1, &P.ia, &P.ib);
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  cfd_local_j_range(c, 2, NJ - 2, &P.jlo, &P.jhi);
  P.urf = c->cfg.urf[CFD_EQ_V];
  if (c->cfg.sew_nim1 != 0.0) {
    P.stale_area = c->cfg.sew_nim1;
  } else {
    P.stale_area = c->geom_host[CFD_G_SEW][NI - 2 - c->cfg.goff];   // sew(nim1)
  }
  P.dout = c->field[CFD_DV];
  return launch_assemble(c, 1, P, CFD_EQ_V);
}

int cfd_launch_assemble_p(cfd_ctx* c) {
  CavParams P{};
  fill_common(c, P);
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  owned_range(c, 2, NI - 1, &P.ia, &P.ib);
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  cfd_local_j_range(c, 1, NJ - 2, &P.jlo, &P.jhi);
  P.dout = nullptr;
  return launch_assemble(c, 2, P, CFD_EQ_P);
}

int cfd_launch_correct(cfd_ctx* c, const double* ppref_dev) {
  CorrParams P{};
  P.L = c->L;
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  owned_range(c, 2, NI - 1, &P.ia, &P.ib);
  if (P.ib < P.ia) return CFD_OK;
  const int NJ = c->cfg.nj_global ? c->cfg.nj_global : c->L.nj;
  cfd_local_j_range(c, 1, NJ - 2, &P.jlo, &P.jhi);
  if (P.jhi < P.jlo) return CFD_OK;
  P.j_skip = 2 - 1 - c->cfg.joff;
  P.i_skip = 2 - 1 - c->cfg.goff;
  P.urfp = c->cfg.urf[CFD_EQ_P];
  P.pp = c->field[CFD_PP];
  P.du = c->field[CFD_DU];
  P.dv = c->field[CFD_DV];
  P.p = c->field[CFD_P];
  P.u = c->field[CFD_U];
  P.v = c->field[CFD_V];
  if (ppref_dev) {
    P.ppref = ppref_dev;
  } else {
    const int il = c->cfg.ipref - 1 - c->cfg.goff, jl = c->cfg.jpref - 1 - c->cfg.joff;
    if (il < 0 || il >= c->L.ni || jl < 0 || jl >= c->L.nj) {
      cfd_set_error("cfd_correct: pressure reference cell (%d,%d) is not in this context", c->cfg.ipref, c->cfg.jpref);
      return CFD_ERR_INVALID;
    }
    P.ppref = c->field[CFD_PP] + (long long)il * c->L.LD + lpos(jl, c->L.M, c->L.T);
  }
  const int nlines = P.ib - P.ia + 1;
  dim3 block(BX, BY), grid((c->L.T + BX - 1) / BX, (nlines + BY - 1) / BY);
  k_correct<<<grid, block, 0, c->stream>>>(P);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

int cfd_launch_reduce_partials(cfd_ctx* c, int eq, int nblocks) {
  k_reduce_partials<<<1, 1024, 0, c->stream>>>(c->partial + (size_t)eq * c->max_partials, nblocks, c->res_dev + eq);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
