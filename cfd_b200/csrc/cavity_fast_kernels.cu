// cavity_fast_kernels.cu -- production (red/black mode) coefficient assembly of the cavity family.
//
// Same equations as cavity_kernels.cu (calcu cavity_fvm.f90:104-187, calcv 205-285) with the arithmetic made
// cheap enough to stay HBM-bound on B200: the ~12 IEEE divisions per cell of the reference's formulas become
// multiplications by reciprocals of the 1-D geometry (computed once on the host when the geometry is set),
// products may contract into FMAs, and the four-point viscosity averages share pair sums between j and j+1.
// Results differ from the bit-exact kernels in the last couple of ulps per coefficient; converged fields agree
// with the reference to ~1e-13 (the contract is 1e-8).  The sequential-order debug mode keeps using the exact
// kernels.  Thread mapping and windows as in cavity_kernels.cu, but arrays that only look back (j-1, j) or only
// ahead (j, j+1) keep two-entry windows to hold more blocks per SM.
#include "common.cuh"

namespace cfd {

__device__ __forceinline__ double fmax2(double a, double b) { return a > b ? a : b; }
// hardware reciprocal seed + two Newton steps (about 1 ulp); ap is a sum of positive coefficients
__device__ __forceinline__ double rcp_fast(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

__device__ __forceinline__ void block_sum_to_fast(double v, double* partial) {
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

#define CFD_SCHEME_FAST(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS) \
  if (iadv == 1) {                                                            \
    AE = fma(-0.5, ce, de);                                                   \
    AW = fma(0.5, cw, dw);                                                    \
    AN = fma(-0.5, cn, dn);                                                   \
    AS = fma(0.5, cs, ds);                                                    \
  } else if (iadv == 2) {                                                     \
    AE = de + fmax2(0.0, -ce);                                                \
    AW = dw + fmax2(0.0, cw);                                                 \
    AN = dn + fmax2(0.0, -cn);                                                \
    AS = ds + fmax2(0.0, cs);                                                 \
  } else {                                                                    \
    AN = fmax2(fabs(0.5 * cn), dn) - 0.5 * cn;                                \
    AS = fmax2(fabs(0.5 * cs), ds) + 0.5 * cs;                                \
    AE = fmax2(fabs(0.5 * ce), de) - 0.5 * ce;                                \
    AW = fmax2(fabs(0.5 * cw), dw) + 0.5 * cw;                                \
  }

// node positions of a thread's chunk
#define POS(k) ((k) * T + t)
#define POS_NEXT(k) (((k) + 1 < M) ? ((k) + 1) * T + t : t + 1)
#define POS_PREV0 ((M - 1) * T + t - 1)

__global__ void __launch_bounds__(128, 4) k_assemble_u_fast(const CavParams P, const FastGeom G) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M;
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
    const double sewu_i = __ldg(P.sewu + i0), r_sewu_i = __ldg(G.rsewu + i0);
    const double r_dxe = __ldg(G.rdxu + i0), r_dxw = __ldg(G.rdxu + i0 - 1);
    const double r_sew_i = __ldg(G.rsew + i0), r_sew_im1 = __ldg(G.rsew + i0 - 1);
    const int iadv = P.iadv;
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf, stale = P.stale_area;
    const bool hasprev = j0base >= 1;
    // windows: u (s,c,n); den W/C (m,c,p); vis pair sums A = visW + visC (m,c,p); v W/C (c,p); 1/dy (m,c)
    double u_s = hasprev ? __ldg(uC + POS_PREV0) : 0.0, u_c = __ldg(uC + t), u_n;
    double dW_m = hasprev ? __ldg(dW + POS_PREV0) : 0.0, dW_c = __ldg(dW + t), dW_p;
    double dC_m = hasprev ? __ldg(dC + POS_PREV0) : 0.0, dC_c = __ldg(dC + t), dC_p;
    double A_m = hasprev ? __ldg(sW + POS_PREV0) + __ldg(sC + POS_PREV0) : 0.0;
    double sW_c = __ldg(sW + t), sC_c = __ldg(sC + t), sW_p, sC_p;
    double vW_c = __ldg(vW + t), vC_c = __ldg(vC + t), vW_p, vC_p;
    double rdy_m = hasprev ? __ldg(G.rdyL + POS_PREV0) : 0.0;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const bool okn = j0 + 1 < nj;
      const int qn = POS_NEXT(k), q = POS(k);
      const int cq = k * P.L.cSK + t;
      u_n = okn ? __ldg(uC + qn) : 0.0;
      dW_p = okn ? __ldg(dW + qn) : 0.0;
      dC_p = okn ? __ldg(dC + qn) : 0.0;
      sW_p = okn ? __ldg(sW + qn) : 0.0;
      sC_p = okn ? __ldg(sC + qn) : 0.0;
      vW_p = okn ? __ldg(vW + qn) : 0.0;
      vC_p = okn ? __ldg(vC + qn) : 0.0;
      const double rdy_c = __ldg(G.rdyL + q);
      const double A_c = sW_c + sC_c;
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double u_w = __ldg(uW + q), u_e = __ldg(uE + q);
        const double den_ww = __ldg(dWW + q), den_e = __ldg(dE + q);
        const double p_w = __ldg(pW + q), p_c = __ldg(pC + q);
        const double sns_j = __ldg(P.snsL + q), rsns_j = __ldg(G.rsnsL + q);
        const double vol = sewu_i * sns_j;
        const double gee = 0.5 * (dC_c + den_e) * u_e;
        const double gww = 0.5 * (den_ww + dW_c) * u_w;
        const double gne = 0.5 * (dC_c + dC_p) * vC_p;
        const double gnw = 0.5 * (dW_c + dW_p) * vW_p;
        const double gse = 0.5 * (dC_m + dC_c) * vC_c;
        const double gsw = 0.5 * (dW_m + dW_c) * vW_c;
        const double gp = 0.5 * (dW_c + dC_c) * u_c;
        const double ce = 0.5 * (gp + gee) * sns_j, cw = 0.5 * (gww + gp) * sns_j;
        const double cn = 0.5 * (gnw + gne) * sewu_i, cs = 0.5 * (gsw + gse) * sewu_i;
        const double smp = ce - cw + cn - cs;
        const double cp = fmax2(smp, 0.0);
        const double vise = sC_c, visw = sW_c;
        const double visn = 0.25 * (A_c + (sW_p + sC_p));
        const double viss = 0.25 * (A_m + A_c);
        const double de = vise * sns_j * r_dxe;
        const double dw = visw * sns_j * r_dxw;
        const double dn = visn * sewu_i * rdy_c;
        const double ds = viss * sewu_i * rdy_m;
        double AE, AW, AN, AS;
        CFD_SCHEME_FAST(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
        double SU = (p_w - p_c) * sns_j;
        const double dudxe = (u_e - u_c) * r_sew_i;
        const double dudxw = (u_c - u_w) * r_sew_im1;
        const double dvdxn = (vC_p - vW_p) * r_sewu_i;
        const double dvdxs = (vC_c - vW_c) * r_sewu_i;
        SU = fma((vise * dudxe - visw * dudxw) * r_sewu_i + (visn * dvdxn - viss * dvdxs) * rsns_j, vol, SU);
        SU = fma(cp, u_c, SU);
        double AP = AE + AW + AN + AS + cp;
        const double resor = AP * u_c - AE * u_e - AW * u_w - AN * u_n - AS * u_s - SU;
        racc += fabs(resor);
        AP = AP * r_urf;
        SU = fma(omu * AP, u_c, SU);
        const double rAP = rcp_fast(AP);
        P.ae[cout + cq] = AE * rAP;
        P.aw[cout + cq] = AW * rAP;
        P.an[cout + cq] = AN * rAP;
        P.as[cout + cq] = AS * rAP;
        P.su[cout + cq] = SU * rAP;
        P.dout[out + q] = stale * rAP;
      }
      u_s = u_c; u_c = u_n;
      dW_m = dW_c; dW_c = dW_p;
      dC_m = dC_c; dC_c = dC_p;
      A_m = A_c; sW_c = sW_p; sC_c = sC_p;
      vW_c = vW_p; vC_c = vC_p;
      rdy_m = rdy_c;
    }
  }
  block_sum_to_fast(racc, P.partial);
}

__global__ void __launch_bounds__(128, 4) k_assemble_v_fast(const CavParams P, const FastGeom G) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M;
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
    const double sew_i = __ldg(P.sew + i0), r_sew_i = __ldg(G.rsew + i0);
    const double r_dxe = __ldg(G.rdx + i0), r_dxw = __ldg(G.rdx + i0 - 1);
    const int iadv = P.iadv;
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf, stale = P.stale_area;
    const bool hasprev = j0base >= 1;
    const int qp = POS_PREV0;
    // back windows (m = j-1): u C/E, den W/E, vis sums SW = visW+visC, SE = visC+visE, visC, p, 1/dyv, 1/sns
    double uC_m = hasprev ? __ldg(uC + qp) : 0.0, uE_m = hasprev ? __ldg(uE + qp) : 0.0;
    double dW_m = hasprev ? __ldg(dW + qp) : 0.0, dE_m = hasprev ? __ldg(dE + qp) : 0.0;
    double sC_m = hasprev ? __ldg(sC + qp) : 0.0;
    double SW_m = hasprev ? __ldg(sW + qp) + sC_m : 0.0, SE_m = hasprev ? sC_m + __ldg(sE + qp) : 0.0;
    double p_m = hasprev ? __ldg(pC + qp) : 0.0;
    double rdyv_m = hasprev ? __ldg(G.rdyvL + qp) : 0.0, rsns_m = hasprev ? __ldg(G.rsnsL + qp) : 0.0;
    // v (s,c,n) and den C (mm,m,c,p)
    double v_s = hasprev ? __ldg(vC + qp) : 0.0, v_c = __ldg(vC + t), v_n;
    double dC_mm = (j0base >= 2) ? __ldg(dC + lpos(j0base - 2, M, T)) : 0.0;
    double dC_m = hasprev ? __ldg(dC + qp) : 0.0, dC_c = __ldg(dC + t), dC_p;
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const bool okn = j0 + 1 < nj;
      const int qn = POS_NEXT(k), q = POS(k);
      const int cq = k * P.L.cSK + t;
      v_n = okn ? __ldg(vC + qn) : 0.0;
      dC_p = okn ? __ldg(dC + qn) : 0.0;
      const double uC_c = __ldg(uC + q), uE_c = __ldg(uE + q);
      const double dW_c = __ldg(dW + q), dE_c = __ldg(dE + q);
      const double sC_c = __ldg(sC + q);
      const double SW_c = __ldg(sW + q) + sC_c, SE_c = sC_c + __ldg(sE + q);
      const double p_c = __ldg(pC + q);
      const double rdyv_c = __ldg(G.rdyvL + q), rsns_c = __ldg(G.rsnsL + q);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double v_w = __ldg(vW + q), v_e = __ldg(vE + q);
        const double snsv_j = __ldg(P.snsvL + q), rsnsv_j = __ldg(G.rsnsvL + q);
        const double vol = sew_i * snsv_j;
        const double gne = 0.5 * (dC_c + dE_c) * uE_c;
        const double gnw = 0.5 * (dW_c + dC_c) * uC_c;
        const double gnn = 0.5 * (dC_c + dC_p) * v_n;
        const double gse = 0.5 * (dC_m + dE_m) * uE_m;
        const double gsw = 0.5 * (dW_m + dC_m) * uC_m;
        const double gss = 0.5 * (dC_mm + dC_m) * v_s;
        const double gp = 0.5 * (dC_m + dC_c) * v_c;
        const double ce = 0.5 * (gne + gse) * snsv_j, cw = 0.5 * (gnw + gsw) * snsv_j;
        const double cn = 0.5 * (gp + gnn) * sew_i, cs = 0.5 * (gss + gp) * sew_i;
        const double smp = ce - cw + cn - cs;
        const double cp = fmax2(smp, 0.0);
        const double vise = 0.25 * (SE_m + SE_c);
        const double visw = 0.25 * (SW_m + SW_c);
        const double visn = sC_c, viss = sC_m;
        const double de = vise * snsv_j * r_dxe;
        const double dw = visw * snsv_j * r_dxw;
        const double dn = visn * sew_i * rdyv_c;
        const double ds = viss * sew_i * rdyv_m;
        double AE, AW, AN, AS;
        CFD_SCHEME_FAST(iadv, ce, cw, cn, cs, de, dw, dn, ds, AE, AW, AN, AS)
        double SU = (p_m - p_c) * sew_i;
        const double dudye = (uE_c - uE_m) * rsnsv_j;
        const double dudyw = (uC_c - uC_m) * rsnsv_j;
        const double dvdyn = (v_n - v_c) * rsns_c;
        const double dvdys = (v_c - v_s) * rsns_m;
        SU = fma((vise * dudye - visw * dudyw) * r_sew_i + (visn * dvdyn - viss * dvdys) * rsnsv_j, vol, SU);
        SU = fma(cp, v_c, SU);
        double AP = AE + AW + AN + AS + cp;
        const double resor = AP * v_c - AE * v_e - AW * v_w - AN * v_n - AS * v_s - SU;
        racc += fabs(resor);
        AP = AP * r_urf;
        SU = fma(omu * AP, v_c, SU);
        const double rAP = rcp_fast(AP);
        P.ae[cout + cq] = AE * rAP;
        P.aw[cout + cq] = AW * rAP;
        P.an[cout + cq] = AN * rAP;
        P.as[cout + cq] = AS * rAP;
        P.su[cout + cq] = SU * rAP;
        P.dout[out + q] = stale * rAP;
      }
      uC_m = uC_c; uE_m = uE_c;
      dW_m = dW_c; dE_m = dE_c;
      sC_m = sC_c; SW_m = SW_c; SE_m = SE_c;
      p_m = p_c; rdyv_m = rdyv_c; rsns_m = rsns_c;
      v_s = v_c; v_c = v_n;
      dC_mm = dC_m; dC_m = dC_c; dC_c = dC_p;
    }
  }
  block_sum_to_fast(racc, P.partial);
}

}  // namespace cfd

using namespace cfd;

int cfd_launch_assemble_fast(cfd_ctx* c, int which, CavParams& P, dim3 grid, dim3 block) {
  FastGeom G;
  G.rdxu = c->gfast[0];
  G.rdx = c->gfast[1];
  G.rsew = c->gfast[2];
  G.rsewu = c->gfast[3];
  G.rdyL = c->gfast[4];
  G.rdyvL = c->gfast[5];
  G.rsnsL = c->gfast[6];
  G.rsnsvL = c->gfast[7];
  if (which == 0) {
    k_assemble_u_fast<<<grid, block, 0, c->stream>>>(P, G);
  } else {
    k_assemble_v_fast<<<grid, block, 0, c->stream>>>(P, G);
  }
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}
