// general_fast.cuh -- parameter block and arithmetic helpers shared by the production assembly kernels of the
// general family (general_fast_kernels.cu: register windows; general_tma_kernels.cu: TMA-staged lines).
#pragma once
#include "common.cuh"

namespace cfd {

struct GenFastParams {
  Layout L;
  int ia, ib, jlo, jhi;
  int goff, NI;              // i-slabs: global 1-based i of local line i0 is i0 + 1 + goff; NI = global ni
  int iadv;
  double urf;
  int boussinesq;
  double rgbeta, tref;
  const double *u, *v, *p, *t, *den, *gam, *cp, *du_in, *dv_in, *su_src, *sp_src;
  double *ae, *aw, *an, *as, *su, *dout;
  // by-i (natural order)
  const double *sew, *sewu, *xcvi, *xcvip, *fx, *r_sew, *r_xdif;
  // by-j (line layout)
  const double *arx, *arxj, *arxjp, *r, *rmn, *fy, *fvp, *ycvr, *ycvrs, *sns, *r_sx, *r_ydif, *r_sxmn, *r_sns, *r_rsns;   // r_rsns = r/sns
  double* partial;
};

__device__ __forceinline__ double gf_max(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double gf_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}
__device__ __forceinline__ double gf_arim(double a, double b, double w) { return fma(b - a, w, a); }
// HARMEAN (cylin_cond.f90:268-285).  Equal arguments (uniform properties, the usual case) return the argument: the
// quotient a*a/a is the same number to rounding and the reciprocal with its Newton steps is a tenth of a kernel.
__device__ __forceinline__ double gf_harm(double a, double b, double w) {
  if (a == b) return a;
  return a * b * gf_rcp(fma(b - a, w, a));
}
// DAPEC (cylin_cond.f90:245-266)
__device__ __forceinline__ double gf_dapec(double flow, double diff, int iadv) {
  if (flow == 0.0) return diff;
  switch (iadv) {
    case 1: return fma(-0.5, fabs(flow), diff);
    case 2: return diff;
    case 3: return gf_max(0.0, fma(-0.5, fabs(flow), diff));
    case 4: {
      const double q = 1.0 - 0.1 * fabs(flow / diff);
      const double q2 = q * q;
      return gf_max(0.0, q2 * q2 * q) * diff;
    }
    default: return diff;
  }
}
// The same without data-dependent branches: central / upwind / hybrid share one FMA and are told apart by selects on
// the (launch-uniform) scheme number; only the power law keeps its own branch.  Same values; measured faster in
// the TMA-staged u and T kernels (-3 %) and slower in the v kernel (+9 %, register pressure), which keeps gf_dapec.
__device__ __forceinline__ double gf_dapec_sel(double flow, double diff, int iadv) {
  if (iadv == 4) {
    if (flow == 0.0) return diff;
    const double q = 1.0 - 0.1 * fabs(flow / diff);
    const double q2 = q * q;
    return gf_max(0.0, q2 * q2 * q) * diff;
  }
  const double c = fma(-0.5, fabs(flow), diff);
  const double r = iadv == 1 ? c : (iadv == 3 ? gf_max(0.0, c) : diff);
  return flow == 0.0 ? diff : r;
}

// u-momentum (buoyancy_cav.f90:135-194):
// y-face between rows lo (j) and hi (j+1) of the u-cell column i: flow and diffusion conductance
template <bool SEL = false>
__device__ __forceinline__ void gf_u_face_y(double Dc_lo, double Dc_hi, double Dw_lo, double Dw_hi, double Gc_lo, double Gc_hi,
                                            double Gw_lo, double Gw_hi, double Vc, double Vw, double xcvi, double xcvip,
                                            double fy_h, double rmn_h, double rydif_h, double r_lo, double r_hi,
                                            bool south_b, bool north_b, int iadv, double& flow, double& acof) {
  double fl, flm, diff;
  if (north_b) {          // hi is the wall row nj (:163-166)
    fl = xcvi * Vc * Dc_hi;
    flm = xcvip * Vw * Dw_hi;
    diff = r_hi * (xcvi * Gc_hi + xcvip * Gw_hi) * rydif_h;
    flow = rmn_h * (fl + flm);
  } else if (south_b) {   // lo is the wall row 1 (:135-142)
    fl = xcvi * Vc * Dc_lo;
    flm = xcvip * Vw * Dw_lo;
    diff = r_lo * (xcvi * Gc_lo + xcvip * Gw_lo) * rydif_h;
    flow = r_lo * (fl + flm);
  } else {
    fl = xcvi * Vc * gf_arim(Dc_lo, Dc_hi, fy_h);
    flm = xcvip * Vw * gf_arim(Dw_lo, Dw_hi, fy_h);
    diff = rmn_h * (gf_harm(Gc_lo, Gc_hi, fy_h) * xcvi + gf_harm(Gw_lo, Gw_hi, fy_h) * xcvip) * rydif_h;
    flow = rmn_h * (fl + flm);
  }
  acof = SEL ? gf_dapec_sel(flow, diff, iadv) : gf_dapec(flow, diff, iadv);
}

}  // namespace cfd

// general_tma_kernels.cu
bool cfd_general_tma_eligible(const cfd_ctx* c, int which);
int cfd_launch_general_assemble_tma(cfd_ctx* c, int which, const cfd::GenFastParams& P);
