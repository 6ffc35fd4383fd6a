// general_fast_kernels.cu -- production (red/black mode) coefficient assembly of the "general" SIMPLE family
// (calcu buoyancy-driven-cavity-2d/buoyancy_cav.f90:122-198, calcv :200-277, calcp :279-343, calct :358-428,
// DAPEC / ARIMEAN / HARMEAN cylinder-heat-conduction-2d/cylin_cond.f90:245-285).
//
// Same equations and the same reference quirks as general_kernels.cu (loop bound `njm1` used for i, `den(nj,j)` at
// the east boundary of calct, `*sns(2)` at the south boundary of calcv, den(i+1,j) in the west mass flux of calcp),
// arranged for bandwidth: one thread walks its M consecutive nodes of a line with the j-neighbours in registers,
// every y-face is evaluated once and handed from cell j to cell j+1, divisions by geometry are multiplications by
// reciprocals built once per geometry, the harmonic mean uses the hardware reciprocal + two Newton steps, products
// contract into FMAs, and the equations are stored divided by ap (nc = 5).  Differences from the bit-exact kernels
// are a few ulps per coefficient; the sequential-order mode keeps using general_kernels.cu.
#include "common.cuh"
#include "general_fast.cuh"

namespace cfd {

__device__ __forceinline__ void gf_block_sum(double v, double* partial) {
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

#define GPOS(k) ((k) * T + t)
#define GPOS_NEXT(k) (((k) + 1 < M) ? ((k) + 1) * T + t : t + 1)
#define GPOS_PREV0 ((M - 1) * T + t - 1)

// normalise by ap (after under-relaxation) and store; returns nothing
#define GEN_STORE(AE, AW, AN, AS, SU, AP)        \
  {                                              \
    const double rAP__ = gf_rcp(AP);             \
    P.ae[cout + cq] = (AE) * rAP__;              \
    P.aw[cout + cq] = (AW) * rAP__;              \
    P.an[cout + cq] = (AN) * rAP__;              \
    P.as[cout + cq] = (AS) * rAP__;              \
    P.su[cout + cq] = (SU) * rAP__;              \
    rAPout = rAP__;                              \
  }

// ---- u: buoyancy_cav.f90:135-194 (y-face helper gf_u_face_y in general_fast.cuh) ---------------------------
__global__ void __launch_bounds__(128, 4) k_gen_assemble_u_fast(const GenFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj, ni = P.L.ni;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M, i = i0 + 1 + P.goff, iadv = P.iadv;   // i: global line number (boundary tests only)
    const double* uW = P.u + (long long)(i0 - 1) * LD;
    const double* uC = uW + LD;
    const double* uE = uC + LD;
    const double* vW = P.v + (long long)(i0 - 1) * LD;
    const double* vC = vW + LD;
    const double* dWW = P.den + (long long)(i0 - 2) * LD;
    const double* dW = dWW + LD;
    const double* dC = dW + LD;
    const double* dE = dC + LD;
    const double* gW = P.gam + (long long)(i0 - 1) * LD;
    const double* gC = gW + LD;
    const double* pW = P.p + (long long)(i0 - 1) * LD;
    const double* pC = pW + LD;
    const long long out = (long long)i0 * LD, cout = (long long)i0 * P.L.cSL;
    const double xcvi = __ldg(P.xcvi + i0), xcvip = __ldg(P.xcvip + i0 - 1);
    const double fx_w = __ldg(P.fx + i0 - 1), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rsew_w = __ldg(P.r_sew + i0 - 1), rsew_c = __ldg(P.r_sew + i0);
    const double sewu_i = __ldg(P.sewu + i0), rxdif_i = __ldg(P.r_xdif + i0);
    const bool west_b = (i == 3), east_b = (i == nj - 1);       // quirk: `do i = 3, njm1`, `if (i == njm1)` (:135,:150)
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
    const bool hasprev = j0base >= 1;
    const int qp0 = GPOS_PREV0;
    double u_s = hasprev ? __ldg(uC + qp0) : 0.0, u_c = __ldg(uC + t), u_n;
    double dW_m = hasprev ? __ldg(dW + qp0) : 0.0, dW_c = __ldg(dW + t), dW_p;
    double dC_m = hasprev ? __ldg(dC + qp0) : 0.0, dC_c = __ldg(dC + t), dC_p;
    double gW_m = hasprev ? __ldg(gW + qp0) : 0.0, gW_c = __ldg(gW + t), gW_p;
    double gC_m = hasprev ? __ldg(gC + qp0) : 0.0, gC_c = __ldg(gC + t), gC_p;
    double vW_c = __ldg(vW + t), vC_c = __ldg(vC + t), vW_p, vC_p;
    double r_m = hasprev ? __ldg(P.r + qp0) : 0.0, r_c = __ldg(P.r + t), r_p;
    // south face of the chunk's first cell (the north face of row j0base-1)
    double aS = 0.0;   // acof + max(0, flow) of the face below the current cell
    if (hasprev) {
      double flow, acof;
      gf_u_face_y(dC_m, dC_c, dW_m, dW_c, gC_m, gC_c, gW_m, gW_c, vC_c, vW_c, xcvi, xcvip, __ldg(P.fy + t),
                  __ldg(P.rmn + t), __ldg(P.r_ydif + t), r_m, r_c, j0base == 1, false, iadv, flow, acof);
      aS = acof + gf_max(0.0, flow);
    }
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      const bool okn = j0 + 1 < nj;
      const int qn = GPOS_NEXT(k), q = GPOS(k);
      const int cq = k * P.L.cSK + t;
      u_n = okn ? __ldg(uC + qn) : 0.0;
      dW_p = okn ? __ldg(dW + qn) : 0.0;
      dC_p = okn ? __ldg(dC + qn) : 0.0;
      gW_p = okn ? __ldg(gW + qn) : 0.0;
      gC_p = okn ? __ldg(gC + qn) : 0.0;
      vW_p = okn ? __ldg(vW + qn) : 0.0;
      vC_p = okn ? __ldg(vC + qn) : 0.0;
      r_p = okn ? __ldg(P.r + qn) : 0.0;
      double flowN = 0.0, acofN = 0.0;
      if (okn)
        gf_u_face_y(dC_c, dC_p, dW_c, dW_p, gC_c, gC_p, gW_c, gW_p, vC_p, vW_p, xcvi, xcvip, __ldg(P.fy + qn),
                    __ldg(P.rmn + qn), __ldg(P.r_ydif + qn), r_c, r_p, j0 == 0, j0 + 1 == nj - 1, iadv, flowN, acofN);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double u_w = __ldg(uW + q), u_e = __ldg(uE + q);
        const double den_ww = __ldg(dWW + q), den_e = __ldg(dE + q);
        const double arx = __ldg(P.arx + q), rsx = __ldg(P.r_sx + q);
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
        const double AW = gf_dapec(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
        const double AE = (gf_dapec(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
        const double AS = aS;
        const double AN = (acofN + gf_max(0.0, flowN)) - flowN;
        const double vol = __ldg(P.ycvr + q) * sewu_i;
        double DU = vol * rxdif_i * rsx;
        double SU = DU * (__ldg(pW + q) - __ldg(pC + q));
        double AP = AE + AW + AN + AS;
        const double reseq = AP * u_c - AE * u_e - AW * u_w - AN * u_n - AS * u_s - SU;
        racc += fabs(reseq);
        AP = AP * r_urf;
        SU = fma(omu * AP, u_c, SU);
        double rAPout;
        GEN_STORE(AE, AW, AN, AS, SU, AP)
        P.dout[out + q] = DU * rAPout;
      }
      aS = acofN + gf_max(0.0, flowN);
      u_s = u_c; u_c = u_n;
      dW_m = dW_c; dW_c = dW_p;
      dC_m = dC_c; dC_c = dC_p;
      gW_m = gW_c; gW_c = gW_p;
      gC_m = gC_c; gC_c = gC_p;
      vW_c = vW_p; vC_c = vC_p;
      r_m = r_c; r_c = r_p;
    }
  }
  gf_block_sum(racc, P.partial);
}

// ---- v: buoyancy_cav.f90:212-273 ------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 4) k_gen_assemble_v_fast(const GenFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M, i = i0 + 1 + P.goff, iadv = P.iadv;   // i: global line number (boundary tests only)
    const double* uC = P.u + (long long)i0 * LD;        // u(i,.)   : west face flux
    const double* uE = uC + LD;                         // u(i+1,.) : east face flux
    const double* vW = P.v + (long long)(i0 - 1) * LD;
    const double* vC = vW + LD;
    const double* vE = vC + LD;
    const double* dW = P.den + (long long)(i0 - 1) * LD;
    const double* dC = dW + LD;
    const double* dE = dC + LD;
    const double* gW = P.gam + (long long)(i0 - 1) * LD;
    const double* gC = gW + LD;
    const double* gE = gC + LD;
    const double* pC = P.p + (long long)i0 * LD;
    const double* tC = P.t ? P.t + (long long)i0 * LD : nullptr;
    const long long out = (long long)i0 * LD, cout = (long long)i0 * P.L.cSL;
    const double fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rxdif_c = __ldg(P.r_xdif + i0), rxdif_e = __ldg(P.r_xdif + i0 + 1);
    const double sew_i = __ldg(P.sew + i0);
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
    const bool hasprev = j0base >= 1;
    const int qp0 = GPOS_PREV0;
    // per-row x-face ingredients of row j-1 (m) : mass flux and conductance pieces on the west and east faces
    double fw_m = 0.0, fe_m = 0.0, hw_m = 0.0, he_m = 0.0;
    double dC_mm = (j0base >= 2) ? __ldg(dC + lpos(j0base - 2, M, T)) : 0.0;
    double dC_m = hasprev ? __ldg(dC + qp0) : 0.0, dC_c = __ldg(dC + t), dC_p;
    double gC_m = hasprev ? __ldg(gC + qp0) : 0.0;
    double v_s = hasprev ? __ldg(vC + qp0) : 0.0, v_c = __ldg(vC + t), v_n;
    double p_m = hasprev ? __ldg(pC + qp0) : 0.0, t_m = (hasprev && tC) ? __ldg(tC + qp0) : 0.0;
    double arxjp_m = hasprev ? __ldg(P.arxjp + qp0) : 0.0;
    double r_m = hasprev ? __ldg(P.r + qp0) : 0.0, r_mm = (j0base >= 2) ? __ldg(P.r + lpos(j0base - 2, M, T)) : 0.0;
    double rsns_m = hasprev ? __ldg(P.r_sns + qp0) : 0.0, sns_m = hasprev ? __ldg(P.sns + qp0) : 0.0;
    double fvp_m = hasprev ? __ldg(P.fvp + qp0) : 0.0;
    double rmn_c = __ldg(P.rmn + t), fy_c = __ldg(P.fy + t);
    if (hasprev) {
      const double dW_q = __ldg(dW + qp0), dE_q = __ldg(dE + qp0), uC_q = __ldg(uC + qp0), uE_q = __ldg(uE + qp0);
      const double gW_q = __ldg(gW + qp0), gE_q = __ldg(gE + qp0);
      fw_m = west_b ? uC_q * dW_q : uC_q * gf_arim(dW_q, dC_m, fx_c);
      fe_m = east_b ? uE_q * dE_q : uE_q * gf_arim(dC_m, dE_q, fx_e);
      hw_m = west_b ? gW_q : gf_harm(gW_q, gC_m, fx_c);
      he_m = east_b ? gE_q : gf_harm(gC_m, gE_q, fx_e);
    }
    // mass flux through the scalar-node plane j (v-cell face): G(j) = v(i,j)*rmn(j)*sew*ARIMEAN(den(j-1),den(j),fy(j))
    double G_c = v_c * rmn_c * sew_i * gf_arim(dC_m, dC_c, fy_c);
    double G_m = 0.0;
    if (hasprev) {
      const double rmn_q = __ldg(P.rmn + qp0), fy_q = __ldg(P.fy + qp0);
      G_m = v_s * rmn_q * sew_i * gf_arim(dC_mm, dC_m, fy_q);
    }
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k, j = j0 + 1;
      const bool okn = j0 + 1 < nj;
      const int qn = GPOS_NEXT(k), q = GPOS(k);
      const int cq = k * P.L.cSK + t;
      v_n = okn ? __ldg(vC + qn) : 0.0;
      dC_p = okn ? __ldg(dC + qn) : 0.0;
      const double rmn_p = okn ? __ldg(P.rmn + qn) : 0.0, fy_p = okn ? __ldg(P.fy + qn) : 0.0;
      const double G_p = v_n * rmn_p * sew_i * gf_arim(dC_c, dC_p, fy_p);
      const double dW_c = __ldg(dW + q), dE_c = __ldg(dE + q), uC_c = __ldg(uC + q), uE_c = __ldg(uE + q);
      const double gW_c = __ldg(gW + q), gC_c = __ldg(gC + q), gE_c = __ldg(gE + q);
      const double p_c = __ldg(pC + q), t_c = tC ? __ldg(tC + q) : 0.0;
      const double r_c = __ldg(P.r + q), rsns_c = __ldg(P.r_sns + q), sns_c = __ldg(P.sns + q), fvp_c = __ldg(P.fvp + q);
      const double arxjp_c = __ldg(P.arxjp + q);
      const double fw_c = west_b ? uC_c * dW_c : uC_c * gf_arim(dW_c, dC_c, fx_c);
      const double fe_c = east_b ? uE_c * dE_c : uE_c * gf_arim(dC_c, dE_c, fx_e);
      const double hw_c = west_b ? gW_c : gf_harm(gW_c, gC_c, fx_c);
      const double he_c = east_b ? gE_c : gf_harm(gC_c, gE_c, fx_e);
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double v_w = __ldg(vW + q), v_e = __ldg(vE + q);
        const double arxj = __ldg(P.arxj + q), rsxmn = __ldg(P.r_sxmn + q);
        // x-faces (:212-233): two half faces, rows j and j-1
        const double flow_w = arxj * fw_c + arxjp_m * fw_m;
        const double diff_w = (hw_c * arxj + hw_m * arxjp_m) * (rxdif_c * rsxmn);
        const double flow_e = arxj * fe_c + arxjp_m * fe_m;
        const double diff_e = (he_c * arxj + he_m * arxjp_m) * (rxdif_e * rsxmn);
        // y-faces: north face sits on scalar node j, south face on node j-1
        double flow_n, diff_n, flow_s, diff_s;
        if (j == nj - 1) {                                      // (:246-249) wall row nj
          const double area = __ldg(P.r + qn) * sew_i;
          flow_n = area * v_n * dC_p;
          diff_n = area * __ldg(gC + qn) * rsns_c;
        } else {
          flow_n = gf_arim(G_c, G_p, fvp_c);
          diff_n = r_c * sew_i * gC_c * rsns_c;
        }
        if (j == 3) {                                           // (:212-217) wall row 1; quirk: *sns(2)
          const double area = r_mm * sew_i;
          flow_s = area * v_s * dC_mm;
          diff_s = area * __ldg(gC + lpos(0, M, T)) * sns_m;
        } else {
          flow_s = gf_arim(G_m, G_c, fvp_m);
          diff_s = r_m * sew_i * gC_m * rsns_m;
        }
        const double AW = gf_dapec(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
        const double AE = (gf_dapec(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
        const double AS = gf_dapec(flow_s, diff_s, iadv) + gf_max(0.0, flow_s);
        const double AN = (gf_dapec(flow_n, diff_n, iadv) + gf_max(0.0, flow_n)) - flow_n;
        const double vol = __ldg(P.ycvrs + q) * sew_i;
        const double src = P.boussinesq ? P.rgbeta * (gf_arim(t_m, t_c, fy_c) - P.tref) : 0.0;
        double DV = vol * __ldg(P.r_ydif + q);
        double SU = fma(src, vol, DV * (p_m - p_c));
        double AP = AE + AW + AN + AS;
        const double reseq = AP * v_c - AE * v_e - AW * v_w - AN * v_n - AS * v_s - SU;
        racc += fabs(reseq);
        AP = AP * r_urf;
        SU = fma(omu * AP, v_c, SU);
        double rAPout;
        GEN_STORE(AE, AW, AN, AS, SU, AP)
        P.dout[out + q] = DV * rAPout;
      }
      fw_m = fw_c; fe_m = fe_c; hw_m = hw_c; he_m = he_c;
      dC_mm = dC_m; dC_m = dC_c; dC_c = dC_p;
      gC_m = gC_c;
      v_s = v_c; v_c = v_n;
      p_m = p_c; t_m = t_c;
      arxjp_m = arxjp_c;
      r_mm = r_m; r_m = r_c;
      rsns_m = rsns_c; sns_m = sns_c; fvp_m = fvp_c;
      G_m = G_c; G_c = G_p;
      rmn_c = rmn_p; fy_c = fy_p;
    }
  }
  gf_block_sum(racc, P.partial);
}

// ---- p': buoyancy_cav.f90:294-340 ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 4) k_gen_assemble_p_fast(const GenFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M, i = i0 + 1 + P.goff;
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
    const double sew_i = __ldg(P.sew + i0), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    const bool hasprev = j0base >= 1;
    double d_m = hasprev ? __ldg(dC + GPOS_PREV0) : 0.0, d_c = __ldg(dC + t), d_p;
    // south face of the chunk's first cell: mass flux in and conductance
    double fS = 0.0, aS = 0.0;
    if (hasprev) {
      const double v_c = __ldg(vC + t), arean = __ldg(P.rmn + t) * sew_i;
      if (j0base == 1) {
        fS = d_m * v_c * arean;                                  // (:301-306) den(i,1)*v(i,2)
        aS = 0.0;
      } else {
        const double denn = gf_arim(d_m, d_c, __ldg(P.fy + t));
        fS = denn * v_c * arean;
        aS = denn * arean * __ldg(dvC + t);
      }
    }
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k, j = j0 + 1;
      const bool okn = j0 + 1 < nj;
      const int qn = GPOS_NEXT(k), q = GPOS(k);
      const int cq = k * P.L.cSK + t;
      d_p = okn ? __ldg(dC + qn) : 0.0;
      double fN = 0.0, aN = 0.0;
      if (okn) {
        const double v_p = __ldg(vC + qn), arean = __ldg(P.rmn + qn) * sew_i;
        if (j == nj - 1) {
          fN = d_p * v_p * arean;                                // den(i,nj)*v(i,nj)
          aN = 0.0;
        } else if (j0 == 0) {
          fN = d_c * v_p * arean;
          aN = 0.0;
        } else {
          const double denn = gf_arim(d_c, d_p, __ldg(P.fy + qn));
          fN = denn * v_p * arean;
          aN = denn * arean * __ldg(dvC + qn);
        }
      }
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double arx = __ldg(P.arx + q);
        const double u_c = __ldg(uC + q), u_e = __ldg(uE + q), den_w = __ldg(dW + q), den_e = __ldg(dE + q);
        double SU = fS, AW, AE;
        if (west_b) {
          SU = SU + den_w * u_c * arx;
          AW = 0.0;
        } else {
          SU = SU + d_c * u_c * arx;                             // quirk: den(i+1,j) of the upstream cell's formula
          AW = gf_arim(den_w, d_c, fx_c) * arx * __ldg(duC + q);
        }
        SU = SU - den_e * u_e * arx;
        AE = east_b ? 0.0 : gf_arim(d_c, den_e, fx_e) * arx * __ldg(duE + q);
        SU = SU - fN;
        racc += fabs(SU);
        const double APs = AE + AW + aN + aS;
        const double rr = APs != 0.0 ? 1.0 / APs : 0.0;
        P.ae[cout + cq] = AE * rr;
        P.aw[cout + cq] = AW * rr;
        P.an[cout + cq] = aN * rr;
        P.as[cout + cq] = aS * rr;
        P.su[cout + cq] = SU * rr;
      }
      fS = fN; aS = aN;
      d_m = d_c; d_c = d_p;
    }
  }
  gf_block_sum(racc, P.partial);
}

// ---- T: buoyancy_cav.f90:373-428 ------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 4) k_gen_assemble_t_fast(const GenFastParams P) {
  const int M = P.L.M, T = P.L.T, nj = P.L.nj;
  const long long LD = P.L.LD;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i0 = P.ia + blockIdx.y * blockDim.y + threadIdx.y;
  double racc = 0.0;
  if (t < T && i0 <= P.ib && t * M <= P.jhi && t * M + M - 1 >= P.jlo) {
    const int j0base = t * M, i = i0 + 1 + P.goff, iadv = P.iadv;   // i: global line number (boundary tests only)
    const double* uC = P.u + (long long)i0 * LD;
    const double* uE = uC + LD;
    const double* vC = P.v + (long long)i0 * LD;
    const double* dW = P.den + (long long)(i0 - 1) * LD;
    const double* dC = dW + LD;
    const double* dE = dC + LD;
    const double* cW = P.cp + (long long)(i0 - 1) * LD;
    const double* cC = cW + LD;
    const double* cE = cC + LD;
    const double* gW = P.gam + (long long)(i0 - 1) * LD;
    const double* gC = gW + LD;
    const double* gE = gC + LD;
    const double* tW = P.t + (long long)(i0 - 1) * LD;
    const double* tC = tW + LD;
    const double* tE = tC + LD;
    const long long cout = (long long)i0 * P.L.cSL;
    const double sew_i = __ldg(P.sew + i0), fx_c = __ldg(P.fx + i0), fx_e = __ldg(P.fx + i0 + 1);
    const double rxdif_c = __ldg(P.r_xdif + i0), rxdif_e = __ldg(P.r_xdif + i0 + 1);
    const bool west_b = (i == 2), east_b = (i == P.NI - 1);
    const double r_urf = 1.0 / P.urf, omu = 1.0 - P.urf;
    const bool hasprev = j0base >= 1;
    const int qp0 = GPOS_PREV0;
    double D_m = hasprev ? __ldg(dC + qp0) * __ldg(cC + qp0) : 0.0, D_c = __ldg(dC + t) * __ldg(cC + t), D_p;
    double g_m = hasprev ? __ldg(gC + qp0) : 0.0, g_c = __ldg(gC + t), g_p;
    double t_s = hasprev ? __ldg(tC + qp0) : 0.0, t_c = __ldg(tC + t), t_n;
    double aS = 0.0;
    if (hasprev) {
      const double v_c = __ldg(vC + t);
      double flow, diff;
      if (j0base == 1) {                                         // (:373-378) wall row 1
        const double areas = __ldg(P.r + qp0) * sew_i;
        flow = D_m * v_c * areas;
        diff = g_m * areas * __ldg(P.r_ydif + t);
      } else {
        const double arean = __ldg(P.rmn + t) * sew_i, fy = __ldg(P.fy + t);
        flow = gf_arim(D_m, D_c, fy) * v_c * arean;
        diff = gf_harm(g_m, g_c, fy) * arean * __ldg(P.r_ydif + t);
      }
      aS = gf_dapec(flow, diff, iadv) + gf_max(0.0, flow);
    }
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k, j = j0 + 1;
      const bool okn = j0 + 1 < nj;
      const int qn = GPOS_NEXT(k), q = GPOS(k);
      const int cq = k * P.L.cSK + t;
      D_p = okn ? __ldg(dC + qn) * __ldg(cC + qn) : 0.0;
      g_p = okn ? __ldg(gC + qn) : 0.0;
      t_n = okn ? __ldg(tC + qn) : 0.0;
      double flowN = 0.0, acofN = 0.0;
      if (okn) {
        const double v_p = __ldg(vC + qn), rydif = __ldg(P.r_ydif + qn);
        double diff;
        if (j0 == 0) {                                           // face between wall row 1 and row 2
          const double areas = __ldg(P.r + q) * sew_i;
          flowN = D_c * v_p * areas;
          diff = g_c * areas * rydif;
        } else {
          const double arean = __ldg(P.rmn + qn) * sew_i;
          if (j == nj - 1) {
            flowN = D_p * v_p * arean;
            diff = g_p * arean * rydif;
          } else {
            const double fy = __ldg(P.fy + qn);
            flowN = gf_arim(D_c, D_p, fy) * v_p * arean;
            diff = gf_harm(g_c, g_p, fy) * arean * rydif;
          }
        }
        acofN = gf_dapec(flowN, diff, iadv);
      }
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double arx = __ldg(P.arx + q), rsx = __ldg(P.r_sx + q);
        const double u_c = __ldg(uC + q), u_e = __ldg(uE + q);
        const double D_w = __ldg(dW + q) * __ldg(cW + q), g_w = __ldg(gW + q), g_e = __ldg(gE + q);
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
          const double D_e = __ldg(dE + q) * __ldg(cE + q);
          flow_e = gf_arim(D_c, D_e, fx_e) * u_e * arx;
          diff_e = gf_harm(g_c, g_e, fx_e) * arx * (rxdif_e * rsx);
        }
        const double AW = gf_dapec(flow_w, diff_w, iadv) + gf_max(0.0, flow_w);
        const double AE = (gf_dapec(flow_e, diff_e, iadv) + gf_max(0.0, flow_e)) - flow_e;
        const double AS = aS;
        const double AN = (acofN + gf_max(0.0, flowN)) - flowN;
        const double vol = __ldg(P.ycvr + q) * sew_i;
        const double SP = P.sp_src ? __ldg(P.sp_src + (long long)i0 * LD + q) * vol : 0.0;
        double SU = P.su_src ? __ldg(P.su_src + (long long)i0 * LD + q) * vol : 0.0;
        double AP = AE + AW + AN + AS - SP;
        const double reseq = AP * t_c - AE * __ldg(tE + q) - AW * __ldg(tW + q) - AN * t_n - AS * t_s - SU;
        racc += fabs(reseq);
        AP = AP * r_urf;
        SU = fma(omu * AP, t_c, SU);
        [[maybe_unused]] double rAPout;
        GEN_STORE(AE, AW, AN, AS, SU, AP)
      }
      aS = acofN + gf_max(0.0, flowN);
      D_m = D_c; D_c = D_p;
      g_m = g_c; g_c = g_p;
      t_s = t_c; t_c = t_n;
    }
  }
  gf_block_sum(racc, P.partial);
}

}  // namespace cfd

using namespace cfd;

// Reciprocal geometry, rebuilt from the host copies whenever the geometry changed (shares cfd_ctx::tfast with the
// TEACH family -- a context belongs to one family): [0] 1/sew(i), [1] 1/xdif(i); [6..10] 1/sx, 1/ydif, 1/sxmn, 1/sns, r/sns (j).
static int ensure_gen_fast(cfd_ctx* c) {
  if (c->tfast_ready) return CFD_OK;
  const int ni = c->L.ni, nj = c->L.nj;
  auto inv = [](double d) { return d != 0.0 ? 1.0 / d : 0.0; };
  auto& G = c->geom_host;
  std::vector<double> bi[2], bj[5];
  for (auto& v : bi) v.assign((size_t)ni + 2, 0.0);
  for (auto& v : bj) v.assign((size_t)c->L.LD, 0.0);
  for (int i = 0; i < ni; ++i) {
    bi[0][i] = inv(G[CFD_G_SEW][i]);
    bi[1][i] = inv(G[CFD_G_XDIF][i]);
  }
  for (int j = 0; j < nj; ++j) {
    const int q = lpos(j, c->L.M, c->L.T);
    bj[0][q] = inv(G[CFD_G_SX][j]);
    bj[1][q] = inv(G[CFD_G_YDIF][j]);
    bj[2][q] = inv(G[CFD_G_SXMN][j]);
    bj[3][q] = inv(G[CFD_G_SNS][j]);
    bj[4][q] = G[CFD_G_R][j] * bj[3][q];
  }
  CFD_CUDA(cudaSetDevice(c->device));
  for (int g = 0; g < 2; ++g) {
    if (!c->tfast[g]) CFD_CUDA(cudaMalloc(&c->tfast[g], bi[g].size() * sizeof(double)));
    CFD_CUDA(cudaMemcpyAsync(c->tfast[g], bi[g].data(), bi[g].size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  for (int g = 0; g < 5; ++g) {
    if (!c->tfast[6 + g]) CFD_CUDA(cudaMalloc(&c->tfast[6 + g], bj[g].size() * sizeof(double)));
    CFD_CUDA(cudaMemcpyAsync(c->tfast[6 + g], bj[g].data(), bj[g].size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  CFD_CUDA(cudaStreamSynchronize(c->stream));
  c->tfast_ready = true;
  return CFD_OK;
}

// which: 0 u, 1 v, 2 p', 3 T.  Returns the number of residual partials, or a negative status.
int cfd_launch_general_assemble_fast(cfd_ctx* c, int which, double* partial) {
  int rc = ensure_gen_fast(c);
  if (rc) return rc;
  GenFastParams P{};
  P.L = c->L;
  P.iadv = c->cfg.iadv;
  P.urf = c->cfg.urf[which];
  P.u = c->field[CFD_U]; P.v = c->field[CFD_V]; P.p = c->field[CFD_P]; P.t = c->field[CFD_T];
  P.den = c->field[CFD_DEN]; P.cp = c->field[CFD_CP];
  P.du_in = c->field[CFD_DU]; P.dv_in = c->field[CFD_DV];
  P.ae = c->field[CFD_AE]; P.aw = c->field[CFD_AW]; P.an = c->field[CFD_AN]; P.as = c->field[CFD_AS]; P.su = c->field[CFD_SU];
  P.sew = c->geom[CFD_G_SEW]; P.sewu = c->geom[CFD_G_SEWU]; P.xcvi = c->geom[CFD_G_XCVI]; P.xcvip = c->geom[CFD_G_XCVIP];
  P.fx = c->geom[CFD_G_FX]; P.r_sew = c->tfast[0]; P.r_xdif = c->tfast[1];
  P.arx = c->geom[CFD_G_ARX]; P.arxj = c->geom[CFD_G_ARXJ]; P.arxjp = c->geom[CFD_G_ARXJP]; P.r = c->geom[CFD_G_R];
  P.rmn = c->geom[CFD_G_RMN]; P.fy = c->geom[CFD_G_FY]; P.fvp = c->geom[CFD_G_FVP]; P.ycvr = c->geom[CFD_G_YCVR];
  P.ycvrs = c->geom[CFD_G_YCVRS]; P.sns = c->geom[CFD_G_SNS];
  P.r_sx = c->tfast[6]; P.r_ydif = c->tfast[7]; P.r_sxmn = c->tfast[8]; P.r_sns = c->tfast[9]; P.r_rsns = c->tfast[10];
  const int ni = c->L.ni, nj = c->L.nj;
  P.goff = c->cfg.goff;
  P.NI = c->cfg.ni_global ? c->cfg.ni_global : ni;
  P.jhi = nj - 2;
  // global line ranges 3..NI-1 (u) / 2..NI-1 (v, p', T), clamped to the lines this context owns
  P.ia = (which == 0 ? 2 : 1) - P.goff;
  P.ib = P.NI - 2 - P.goff;
  if (P.ia < c->cfg.halo_w) P.ia = c->cfg.halo_w;
  if (P.ib > ni - 1 - c->cfg.halo_e) P.ib = ni - 1 - c->cfg.halo_e;
  if (which == 0) {
    P.jlo = 1; P.gam = c->field[CFD_VIS]; P.dout = c->field[CFD_DU];
  } else if (which == 1) {
    P.jlo = 2; P.gam = c->field[CFD_VIS]; P.dout = c->field[CFD_DV];
    P.boussinesq = c->boussinesq; P.rgbeta = c->rgbeta; P.tref = c->tref;
    if (!P.boussinesq) P.t = nullptr;
  } else if (which == 2) {
    P.jlo = 1;
  } else {
    P.jlo = 1; P.gam = c->field[CFD_GAMT];
    P.su_src = c->has_t_sources ? c->field[CFD_SUSRC] : nullptr;
    P.sp_src = c->has_t_sources ? c->field[CFD_SPSRC] : nullptr;
  }
  P.partial = partial;
  const int nlines = P.ib - P.ia + 1;
  if (nlines <= 0) return 0;          // a slab that owns none of this equation's lines
  if (cfd_general_tma_eligible(c, which)) return cfd_launch_general_assemble_tma(c, which, P);
  dim3 block(32, 4), grid((c->L.T + 31) / 32, (nlines + 3) / 4);
  const int nblocks = grid.x * grid.y;
  if (nblocks > c->max_partials) {
    cfd_set_error("residual partial buffer too small (%d > %d)", nblocks, c->max_partials);
    return CFD_ERR_INVALID;
  }
  switch (which) {
    case 0: k_gen_assemble_u_fast<<<grid, block, 0, c->stream>>>(P); break;
    case 1: k_gen_assemble_v_fast<<<grid, block, 0, c->stream>>>(P); break;
    case 2: k_gen_assemble_p_fast<<<grid, block, 0, c->stream>>>(P); break;
    default: k_gen_assemble_t_fast<<<grid, block, 0, c->stream>>>(P); break;
  }
  ++c->launches;
  if (cudaGetLastError() != cudaSuccess) {
    cfd_set_error("general-family fast assembly launch failed");
    return CFD_ERR_CUDA;
  }
  return nblocks;
}
