// teach_fast_kernels.cu -- production (red/black mode) coefficient assembly of the TEACH sudden-expansion family.
//
// Same equations as teach_kernels.cu (calcu backward-step-flow-2d/calc.f90:20-124, calcv :137-220, calcp :231-280,
// hook edits of modu / modv promod.f90:26-133), arranged like cavity_fast_kernels.cu: every thread walks its M
// consecutive nodes of one line with the j-neighbours carried in registers, the IEEE divisions of the reference's
// formulas become multiplications by reciprocals of the 1-D geometry (built once on the host per geometry), products
// may contract into FMAs, and the equations are stored divided by ap (nc = 5).  vol/(sewu) = r*sns and similar
// identities are applied analytically, so results differ from the bit-exact kernels in the last ulps; converged
// fields agree to ~1e-12 (contract 1e-8).  The sequential-order mode keeps using teach_kernels.cu.
//
// Global-i tests (i == 3, i == nim1, i == istep) use i0 + goff, so the kernels also serve i-slab contexts.
#include "common.cuh"

namespace cfd {

struct TeachFastParams {
  Layout L;
  int ia, ib, jlo, jhi;      // 0-based inclusive local line / node ranges
  int goff, ni_global;       // global 1-based i of local line i0 is i0 + 1 + goff
  int indcos, istep, jstep;
  double urf, great;
  double tmult_n;            // viscos / (yv(nj) - y(njm1))       (modu, laminar wall law)
  double tmult_w;            // viscos / (x(istep) - xu(istep))   (modv)
  const double *u, *v, *p, *den, *vis, *du_in, *dv_in;
  double *ae, *aw, *an, *as, *su, *dout;
  const double* uinc;
  const double *tmult_n_arr, *tmult_w_arr;   // turbulent contexts: wall-law multipliers per i / per j, else null
  // by-i (natural order, local index)
  const double *sew, *sewu, *r_sew, *r_dxep, *r_dxpw, *r_dxepu, *r_dxpwu;
  // by-j (line layout)
  const double *rv, *rcv, *r, *aew, *aewv, *r_dynp, *r_dyps, *r_dynpv, *r_dypsv, *snsv_rv, *r_sns, *c3;
  double* partial;
};

__device__ __forceinline__ double tf_max(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double tf_rcp(double x) {   // hardware seed + two Newton steps (about 1 ulp)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

__device__ __forceinline__ void tf_block_sum(double v, double* partial) {
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

#define TPOS(k) ((k) * T + t)
#define TPOS_NEXT(k) (((k) + 1 < M) ? ((k) + 1) * T + t : t + 1)
#define TPOS_PREV0 ((M - 1) * T + t - 1)

#define TEACH_HYBRID(ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS) \
  AN = tf_max(fabs(0.5 * cn), dn) - 0.5 * cn;                        \
  AS = tf_max(fabs(0.5 * cs), ds) + 0.5 * cs;                        \
  AE = tf_max(fabs(0.5 * ce), de) - 0.5 * ce;                        \
  AW = tf_max(fabs(0.5 * cw), dw) + 0.5 * cw;

// ---- u ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 4) k_teach_assemble_u_fast(const TeachFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M;
    const int ig = i0 + 1 + P.goff;                                   // global 1-based i
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
    const double sewu_i = __ldg(P.sewu + i0), sew_i = __ldg(P.sew + i0);
    const double r_dxe = __ldg(P.r_dxepu + i0), r_dxw = __ldg(P.r_dxpwu + i0);
    const double r_sew_i = __ldg(P.r_sew + i0), r_sew_im1 = __ldg(P.r_sew + i0 - 1);
    const double r_dxep_i = __ldg(P.r_dxep + i0);
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
    const double uinc = (ig == P.ni_global - 1) ? __ldg(P.uinc) : 0.0;
    const bool outlet = ig == P.ni_global - 1, first_u = ig == 3;
    const bool hasprev = j0base >= 1;
    const int qp0 = TPOS_PREV0;
    double u_s = hasprev ? __ldg(uC + qp0) : 0.0, u_c = __ldg(uC + t), u_n;
    double dW_m = hasprev ? __ldg(dW + qp0) : 0.0, dW_c = __ldg(dW + t), dW_p;
    double dC_m = hasprev ? __ldg(dC + qp0) : 0.0, dC_c = __ldg(dC + t), dC_p;
    double A_m = hasprev ? __ldg(sW + qp0) + __ldg(sC + qp0) : 0.0;
    double sW_c = __ldg(sW + t), sC_c = __ldg(sC + t), sW_p, sC_p;
    double vW_c = __ldg(vW + t), vC_c = __ldg(vC + t), vW_p, vC_p;
    double rv_c = __ldg(P.rv + t), rv_p;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const bool okn = j0 + 1 < nj;
      const int qn = TPOS_NEXT(k), q = TPOS(k);
      const int cq = k * P.L.cSK + t;
      u_n = okn ? __ldg(uC + qn) : 0.0;
      dW_p = okn ? __ldg(dW + qn) : 0.0;
      dC_p = okn ? __ldg(dC + qn) : 0.0;
      sW_p = okn ? __ldg(sW + qn) : 0.0;
      sC_p = okn ? __ldg(sC + qn) : 0.0;
      vW_p = okn ? __ldg(vW + qn) : 0.0;
      vC_p = okn ? __ldg(vC + qn) : 0.0;
      rv_p = okn ? __ldg(P.rv + qn) : 0.0;
      const double A_c = sW_c + sC_c;
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double u_w = __ldg(uW + q), u_e = __ldg(uE + q);
        const double den_ww = __ldg(dWW + q), den_e = __ldg(dE + q);
        const double p_w = __ldg(pW + q), p_c = __ldg(pC + q);
        const double aew = __ldg(P.aew + q), rdn = __ldg(P.r_dynp + q), rds = __ldg(P.r_dyps + q);
        const double arean = rv_p * sewu_i, areas = rv_c * sewu_i;
        const double gn = 0.5 * (dC_p + dC_c) * vC_p;
        const double gnw = 0.5 * (dW_c + dW_p) * vW_p;
        const double gs = 0.5 * (dC_m + dC_c) * vC_c;
        const double gsw = 0.5 * (dW_c + dW_m) * vW_c;
        const double ge = 0.5 * (den_e + dC_c) * u_e;
        const double gp = 0.5 * (dC_c + dW_c) * u_c;
        const double gw = 0.5 * (dW_c + den_ww) * u_w;
        const double cn = 0.5 * (gn + gnw) * arean, cs = 0.5 * (gs + gsw) * areas;
        const double ce = 0.5 * (ge + gp) * aew, cw = 0.5 * (gp + gw) * aew;
        const double visn = 0.25 * (A_c + (sW_p + sC_p));
        const double viss = 0.25 * (A_c + A_m);
        const double dn = visn * arean * rdn, ds = viss * areas * rds;
        const double de = sC_c * aew * r_dxe, dw = sW_c * aew * r_dxw;
        const double smp = cn - cs + ce - cw;
        const double cp = tf_max(0.0, smp);
        double AN, AS, AE, AW;
        TEACH_HYBRID(ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
        double SU = fma(cp, u_c, aew * (p_w - p_c));
        double SP = -cp;
        const double dudxp = (u_e - u_c) * r_sew_i, dudxm = (u_c - u_w) * r_sew_im1;
        SU = fma(sC_c * dudxp - sW_c * dudxm, aew, SU);                       // (...)/sewu*vol, vol = r*sewu*sns
        const double dvdxp = rv_p * (vC_p - vW_p) * r_dxep_i, dvdxm = rv_c * (vC_c - vW_c) * r_dxep_i;
        SU = fma(visn * dvdxp - viss * dvdxm, sewu_i, SU);                    // (...)/sns/r*vol
        if (j0 == nj - 2) {                                                   // modu: top wall (promod.f90:36-60)
          SP = SP - (P.tmult_n_arr ? __ldg(P.tmult_n_arr + i0) : P.tmult_n) * sewu_i * rv_p;
          AN = 0.0;
        }
        if (first_u && j0 >= P.jstep) AW = 0.0;                               // step wall (:67-69)
        if (j0 == 1) AS = 0.0;                                                // symmetry axis (:72-77)
        const double u_s_eff = (j0 == 1) ? u_c : u_s;                         // u(i,1) = u(i,2)
        const double u_e_eff = outlet ? u_c + uinc : u_e;                     // u(ni,j) = u(nim1,j) + uinc
        double AP = AN + AS + AE + AW - SP;
        double resor = AN * u_n + AS * u_s_eff + AE * u_e_eff + AW * u_w - AP * u_c + SU;
        const double sorvol = P.great * aew * sew_i;
        if (-SP > 0.5 * sorvol) resor = resor / sorvol;
        racc += fabs(resor);
        AP = AP * r_urf;
        SU = fma(omu * AP, u_c, SU);
        const double rAP = tf_rcp(AP);
        P.ae[cout + cq] = AE * rAP;
        P.aw[cout + cq] = AW * rAP;
        P.an[cout + cq] = AN * rAP;
        P.as[cout + cq] = AS * rAP;
        P.su[cout + cq] = SU * rAP;
        P.dout[out + q] = aew * rAP;                                          // du = areaew/ap*urf with ap before /urf
      }
      u_s = u_c; u_c = u_n;
      dW_m = dW_c; dW_c = dW_p;
      dC_m = dC_c; dC_c = dC_p;
      A_m = A_c; sW_c = sW_p; sC_c = sC_p;
      vW_c = vW_p; vC_c = vC_p;
      rv_c = rv_p;
    }
  }
  tf_block_sum(racc, P.partial);
}

// ---- v ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 4) k_teach_assemble_v_fast(const TeachFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M;
    const int ig = i0 + 1 + P.goff;
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
    const double sew_i = __ldg(P.sew + i0);
    const double r_dxe = __ldg(P.r_dxep + i0), r_dxw = __ldg(P.r_dxpw + i0);
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
    const bool at_step = ig == P.istep, past_step = ig >= P.istep, axi = P.indcos == 2;
    const bool hasprev = j0base >= 1;
    const int qp = TPOS_PREV0;
    double uC_m = hasprev ? __ldg(uC + qp) : 0.0, uE_m = hasprev ? __ldg(uE + qp) : 0.0;
    double dW_m = hasprev ? __ldg(dW + qp) : 0.0, dE_m = hasprev ? __ldg(dE + qp) : 0.0;
    double sC_m = hasprev ? __ldg(sC + qp) : 0.0;
    double SW_m = hasprev ? __ldg(sW + qp) + sC_m : 0.0, SE_m = hasprev ? sC_m + __ldg(sE + qp) : 0.0;
    double p_m = hasprev ? __ldg(pC + qp) : 0.0;
    double rsns_m = hasprev ? __ldg(P.r_sns + qp) : 0.0, r_m = hasprev ? __ldg(P.r + qp) : 0.0;
    double v_s = hasprev ? __ldg(vC + qp) : 0.0, v_c = __ldg(vC + t), v_n;
    double dC_mm = (j0base >= 2) ? __ldg(dC + lpos(j0base - 2, M, T)) : 0.0;
    double dC_m = hasprev ? __ldg(dC + qp) : 0.0, dC_c = __ldg(dC + t), dC_p;
    double rcv_c = __ldg(P.rcv + t), rcv_p;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const bool okn = j0 + 1 < nj;
      const int qn = TPOS_NEXT(k), q = TPOS(k);
      const int cq = k * P.L.cSK + t;
      v_n = okn ? __ldg(vC + qn) : 0.0;
      dC_p = okn ? __ldg(dC + qn) : 0.0;
      rcv_p = okn ? __ldg(P.rcv + qn) : 0.0;
      const double uC_c = __ldg(uC + q), uE_c = __ldg(uE + q);
      const double dW_c = __ldg(dW + q), dE_c = __ldg(dE + q);
      const double sC_c = __ldg(sC + q);
      const double SW_c = __ldg(sW + q) + sC_c, SE_c = sC_c + __ldg(sE + q);
      const double p_c = __ldg(pC + q);
      const double rsns_c = __ldg(P.r_sns + q), r_c = __ldg(P.r + q);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double v_w = __ldg(vW + q), v_e = __ldg(vE + q);
        const double aewv = __ldg(P.aewv + q), rdn = __ldg(P.r_dynpv + q), rds = __ldg(P.r_dypsv + q);
        const double rdyps = __ldg(P.r_dyps + q), c3 = __ldg(P.c3 + q), aew = __ldg(P.aew + q);
        const double arean = rcv_p * sew_i, areas = rcv_c * sew_i;
        const double gn = 0.5 * (dC_p + dC_c) * v_n;
        const double gp = 0.5 * (dC_c + dC_m) * v_c;
        const double gs = 0.5 * (dC_m + dC_mm) * v_s;
        const double ge = 0.5 * (dE_c + dC_c) * uE_c;
        const double gse = 0.5 * (dC_m + dE_m) * uE_m;
        const double gw = 0.5 * (dC_c + dW_c) * uC_c;
        const double gsw = 0.5 * (dC_m + dW_m) * uC_m;
        const double cn = 0.5 * (gn + gp) * arean, cs = 0.5 * (gp + gs) * areas;
        const double ce = 0.5 * (ge + gse) * aewv, cw = 0.5 * (gw + gsw) * aewv;
        const double vise = 0.25 * (SE_c + SE_m), visw = 0.25 * (SW_c + SW_m);
        const double dn = sC_c * arean * rdn, ds = sC_m * areas * rds;
        const double de = vise * aewv * r_dxe, dw = visw * aewv * r_dxw;
        const double smp = cn - cs + ce - cw;
        const double cp = tf_max(0.0, smp);
        double AN, AS, AE, AW;
        TEACH_HYBRID(ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
        double SU = fma(cp, v_c, 0.5 * (arean + areas) * (p_m - p_c));
        double SP = -cp;
        if (axi) SP = fma(-2.0 * sC_c * sew_i, __ldg(P.snsv_rv + q), SP);      // applied twice (calc.f90:183-184)
        const double dudyp = (uE_c - uE_m) * rdyps, dudym = (uC_c - uC_m) * rdyps;
        SU = fma(vise * dudyp - visw * dudym, aewv, SU);                       // (...)/sew*vol, vol = rv*sew*snsv
        const double dvdyp = (v_n - v_c) * rsns_c, dvdym = (v_c - v_s) * rsns_m;
        SU = fma((sC_c * r_c) * dvdyp - (sC_m * r_m) * dvdym, c3 * sew_i, SU);  // (...)/(r*sns)*vol
        if (at_step && j0 >= P.jstep) {                                        // modv: side wall (promod.f90:105-121)
          SP = SP - (P.tmult_w_arr ? __ldg(P.tmult_w_arr + j0) : P.tmult_w) * aewv;
          AW = 0.0;
        }
        if (j0 == nj - 2 && past_step) AN = 0.0;
        if (j0 == 2) AS = 0.0;
        double AP = AN + AS + AE + AW - SP;
        double resor = AN * v_n + AS * v_s + AE * v_e + AW * v_w - AP * v_c + SU;
        const double sorvol = P.great * aew * sew_i;
        if (-SP > 0.5 * sorvol) resor = resor / sorvol;
        racc += fabs(resor);
        AP = AP * r_urf;
        SU = fma(omu * AP, v_c, SU);
        const double rAP = tf_rcp(AP);
        P.ae[cout + cq] = AE * rAP;
        P.aw[cout + cq] = AW * rAP;
        P.an[cout + cq] = AN * rAP;
        P.as[cout + cq] = AS * rAP;
        P.su[cout + cq] = SU * rAP;
        P.dout[out + q] = 0.5 * (arean + areas) * rAP;
      }
      uC_m = uC_c; uE_m = uE_c;
      dW_m = dW_c; dE_m = dE_c;
      sC_m = sC_c; SW_m = SW_c; SE_m = SE_c;
      p_m = p_c; rsns_m = rsns_c; r_m = r_c;
      v_s = v_c; v_c = v_n;
      dC_mm = dC_m; dC_m = dC_c; dC_c = dC_p;
      rcv_c = rcv_p;
    }
  }
  tf_block_sum(racc, P.partial);
}

// ---- p' -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 4) k_teach_assemble_p_fast(const TeachFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M;
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
    const bool hasprev = j0base >= 1;
    double d_m = hasprev ? __ldg(dC + TPOS_PREV0) : 0.0, d_c = __ldg(dC + t), d_p;
    double v_c = __ldg(vC + t), v_p, dv_c = __ldg(dvC + t), dv_p, rv_c = __ldg(P.rv + t), rv_p;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const bool okn = j0 + 1 < nj;
      const int qn = TPOS_NEXT(k), q = TPOS(k);
      const int cq = k * P.L.cSK + t;
      d_p = okn ? __ldg(dC + qn) : 0.0;
      v_p = okn ? __ldg(vC + qn) : 0.0;
      dv_p = okn ? __ldg(dvC + qn) : 0.0;
      rv_p = okn ? __ldg(P.rv + qn) : 0.0;
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double aew = __ldg(P.aew + q);
        const double arean = rv_p * sew_i, areas = rv_c * sew_i;
        const double denn = 0.5 * (d_c + d_p), dens = 0.5 * (d_c + d_m);
        const double dene = 0.5 * (d_c + __ldg(dE + q)), denw = 0.5 * (d_c + __ldg(dW + q));
        const double u_c = __ldg(uC + q), u_e = __ldg(uE + q);
        const double AN = denn * arean * dv_p, AS = dens * areas * dv_c;
        const double AE = dene * aew * __ldg(duE + q), AW = denw * aew * __ldg(duC + q);
        const double smp = denn * v_p * arean - dens * v_c * areas + dene * u_e * aew - denw * u_c * aew;
        racc += fabs(smp);
        const double APs = AN + AS + AE + AW;
        const double rr = APs != 0.0 ? 1.0 / APs : 0.0;
        P.ae[cout + cq] = AE * rr;
        P.aw[cout + cq] = AW * rr;
        P.an[cout + cq] = AN * rr;
        P.as[cout + cq] = AS * rr;
        P.su[cout + cq] = -smp * rr;
      }
      d_m = d_c; d_c = d_p;
      v_c = v_p; dv_c = dv_p; rv_c = rv_p;
    }
  }
  tf_block_sum(racc, P.partial);
}

}  // namespace cfd

using namespace cfd;

// Reciprocal / product geometry of the fast kernels, rebuilt from the host copies whenever the geometry changed.
// tfast[0..5] by-i (ni+2 doubles), tfast[6..16] by-j (LD doubles, line layout).
static int ensure_teach_fast(cfd_ctx* c) {
  if (c->tfast_ready) return CFD_OK;
  const int ni = c->L.ni, nj = c->L.nj;
  auto inv = [](double d) { return d != 0.0 ? 1.0 / d : 0.0; };
  auto& G = c->geom_host;
  std::vector<double> bi[5], bj[11];
  for (auto& v : bi) v.assign((size_t)ni + 2, 0.0);
  for (auto& v : bj) v.assign((size_t)c->L.LD, 0.0);
  for (int i = 0; i < ni; ++i) {
    bi[0][i] = inv(G[CFD_G_SEW][i]);
    bi[1][i] = inv(G[CFD_G_DXEP][i]);
    bi[2][i] = inv(G[CFD_G_DXPW][i]);
    bi[3][i] = inv(G[CFD_G_DXEPU][i]);
    bi[4][i] = inv(G[CFD_G_DXPWU][i]);
  }
  for (int j = 0; j < nj; ++j) {
    const int q = lpos(j, c->L.M, c->L.T);
    const double r = G[CFD_G_R][j], rv = G[CFD_G_RV][j], sns = G[CFD_G_SNS][j], snsv = G[CFD_G_SNSV][j];
    bj[0][q] = r * sns;                      // aew
    bj[1][q] = rv * snsv;                    // aewv
    bj[2][q] = inv(G[CFD_G_DYNP][j]);
    bj[3][q] = inv(G[CFD_G_DYPS][j]);
    bj[4][q] = inv(G[CFD_G_DYNPV][j]);
    bj[5][q] = inv(G[CFD_G_DYPSV][j]);
    bj[6][q] = snsv * inv(rv);               // vol/(rv*rv)/sew
    bj[7][q] = inv(sns);
    bj[8][q] = rv * snsv * inv(r * sns);     // vol/(r*sns)/sew
    bj[9][q] = 0.0;
    bj[10][q] = 0.0;
  }
  CFD_CUDA(cudaSetDevice(c->device));
  for (int g = 0; g < 5; ++g) {
    if (!c->tfast[g]) CFD_CUDA(cudaMalloc(&c->tfast[g], bi[g].size() * sizeof(double)));
    CFD_CUDA(cudaMemcpyAsync(c->tfast[g], bi[g].data(), bi[g].size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  for (int g = 0; g < 9; ++g) {
    if (!c->tfast[6 + g]) CFD_CUDA(cudaMalloc(&c->tfast[6 + g], bj[g].size() * sizeof(double)));
    CFD_CUDA(cudaMemcpyAsync(c->tfast[6 + g], bj[g].data(), bj[g].size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  c->tfast_ready = true;
  return CFD_OK;
}

// which: 0 u, 1 v, 2 p'; ia..ib local 0-based lines.  Returns the number of residual partials, or a negative status.
int cfd_launch_teach_assemble_fast(cfd_ctx* c, int which, int ia, int ib, double* partial, const double* uinc) {
  int rc = ensure_teach_fast(c);
  if (rc) return rc;
  TeachFastParams P{};
  P.L = c->L;
  P.ia = ia; P.ib = ib;
  P.jhi = c->L.nj - 2;
  P.jlo = (which == 1) ? 2 : 1;
  P.goff = c->cfg.goff;
  P.ni_global = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  P.indcos = c->cfg.indcos; P.istep = c->cfg.istep; P.jstep = c->cfg.jstep;
  P.urf = c->cfg.urf[which];
  P.great = c->cfg.great;
  auto& G = c->geom_host;
  const int nj = c->L.nj;
  P.tmult_n = c->cfg.viscos / (G[CFD_G_YV][nj - 1] - G[CFD_G_Y][nj - 2]);
  {
    const int il = c->cfg.istep - 1 - c->cfg.goff;     // local index of istep, if this context holds it
    P.tmult_w = (il >= 0 && il < c->L.ni) ? c->cfg.viscos / (G[CFD_G_X][il] - G[CFD_G_XU][il]) : 0.0;
  }
  P.u = c->field[CFD_U]; P.v = c->field[CFD_V]; P.p = c->field[CFD_P]; P.den = c->field[CFD_DEN]; P.vis = c->field[CFD_VIS];
  P.du_in = c->field[CFD_DU]; P.dv_in = c->field[CFD_DV];
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS]; P.su = c->field[CFD_SU];
  P.dout = which == 0 ? c->field[CFD_DU] : c->field[CFD_DV];
  P.uinc = uinc;
  P.tmult_n_arr = c->turbulent ? c->tmult_n : nullptr;
  P.tmult_w_arr = c->turbulent ? c->tmult_w : nullptr;
  P.sew = c->geom[CFD_G_SEW]; P.sewu = c->geom[CFD_G_SEWU];
  P.r_sew = c->tfast[0]; P.r_dxep = c->tfast[1]; P.r_dxpw = c->tfast[2]; P.r_dxepu = c->tfast[3]; P.r_dxpwu = c->tfast[4];
  P.rv = c->geom[CFD_G_RV]; P.rcv = c->geom[CFD_G_RCV]; P.r = c->geom[CFD_G_R];
  P.aew = c->tfast[6]; P.aewv = c->tfast[7]; P.r_dynp = c->tfast[8]; P.r_dyps = c->tfast[9]; P.r_dynpv = c->tfast[10];
  P.r_dypsv = c->tfast[11]; P.snsv_rv = c->tfast[12]; P.r_sns = c->tfast[13]; P.c3 = c->tfast[14];
  P.partial = partial;
  const int nlines = ib - ia + 1;
  dim3 block(32, 4), grid((c->L.T + 31) / 32, (nlines + 3) / 4);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  if (which == 0) {
    k_teach_assemble_u_fast<<<grid, block, 0, c->stream>>>(P);
  } else if (which == 1) {
    k_teach_assemble_v_fast<<<grid, block, 0, c->stream>>>(P);
  } else {
    k_teach_assemble_p_fast<<<grid, block, 0, c->stream>>>(P);
  }
  ++c->launches;
  if (cudaGetLastError() != cudaSuccess) {
    cfd_set_error("TEACH fast assembly launch failed");
    return CFD_ERR_CUDA;
  }
  return nblocks;
}
