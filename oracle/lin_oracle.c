/*
 * oracle/lin_oracle.c -- CPU restatement of the reference's scalar-transport demos on a 5-point stencil.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/cavity_oracle.c header for the rule).
 *
 * Sources restated (paths under /root/reference/):
 *   orc_lin_grid        <- grid / interpolation factors   stagnation-point-flow-2d/steady/fvm2d_stag_point.f90:405-447
 *                          (identical lines in diagonal-convection-2d/fvm2d_oblique_conv.f90:118-160)
 *   orc_stag_setup      <- velocities, boundary phi, coefficients, boundary edits   fvm2d_stag_point.f90:448-505
 *   orc_stag_tdma       <- tdma      :232-260   (a x = b x+ + c x- + d, two divisions per row)
 *   orc_stag_wesweep    <- wesweep   :164-196   (x(ni) = phi(i,ni): ni where nj is meant -- kept; callers use ni == nj)
 *   orc_stag_snsweep    <- snsweep   :198-230
 *   orc_stag_solve      <- tdma_we / tdma_sn   :82-162
 *   orc_stag_finish     <- outlet / symmetry copies and the wall flux   :518-527
 *   orc_stag_time_terms <- time discretisation of stag_point_unsteady.f90:634-688 (explicit / implicit Euler / Crank-Nicolson)
 *   orc_oblique_setup   <- fvm2d_oblique_conv.f90:161-232
 *   orc_oblique_tdma / _wesweep / _solve <- :348-387, :320-346, :284-318  (a x- + b x + c x+ = d, no end values)
 * `order` = 0 follows the reference's line order; 1 colours the lines (rows) by parity like the red/black device mode.
 *
 * PARITY PIN: the reference ships no outputs for these programs -> "parity unpinned"; the profile along the
 * diagonal and the wall flux they print are checked for plausibility only (tests/test_oracle_lin.py).
 */
#include <math.h>
#include <stdlib.h>

#define A2(a, i, j) ((a)[(size_t)((i) - 1) + (size_t)((j) - 1) * (size_t)ni])
#define A1(a, i) ((a)[(i) - 1])

void orc_lin_grid(int nicv, int njcv, double xmin, double xmax, double ymin, double ymax, double *x, double *y, double *xc,
                  double *yc, double *fx, double *fy) {
  const int ni = nicv + 2, nj = njcv + 2, nim1 = ni - 1, njm1 = nj - 1;
  double dx = (xmax - xmin) / (double)nicv, dy = (ymax - ymin) / (double)njcv;   /* expfx = expfy = 1 */
  int i, j;
  A1(x, 1) = xmin;
  for (i = 2; i <= nim1; ++i) A1(x, i) = A1(x, i - 1) + dx;
  A1(x, ni) = A1(x, nim1);
  A1(xc, 1) = A1(x, 1);
  for (i = 2; i <= nim1; ++i) A1(xc, i) = 0.5 * (A1(x, i) + A1(x, i - 1));
  A1(xc, ni) = A1(x, nim1);
  A1(y, 1) = ymin;
  for (j = 2; j <= njm1; ++j) A1(y, j) = A1(y, j - 1) + dy;
  A1(y, nj) = A1(y, njm1);
  A1(yc, 1) = A1(y, 1);
  for (j = 2; j <= njm1; ++j) A1(yc, j) = 0.5 * (A1(y, j) + A1(y, j - 1));
  A1(yc, nj) = A1(y, njm1);
  A1(fx, 1) = 0.0;
  for (i = 2; i <= nim1; ++i) A1(fx, i) = (A1(x, i) - A1(xc, i)) / (A1(xc, i + 1) - A1(xc, i));
  A1(fx, ni) = 0.0;
  A1(fy, 1) = 0.0;
  for (j = 2; j <= njm1; ++j) A1(fy, j) = (A1(y, j) - A1(yc, j)) / (A1(yc, j + 1) - A1(yc, j));
  A1(fy, nj) = 0.0;
}

static double dmin(double a, double b) { return a < b ? a : b; }

/* convection / diffusion coefficients of the stagnation-point flow u = x, v = -y (fvm2d_stag_point.f90:448-490) */
static void stag_coeffs(int ni, int nj, int isch, double den, double gam, const double *x, const double *y, const double *xc,
                        const double *yc, const double *fx, const double *fy, const double *u, const double *v, double *ae,
                        double *aw, double *an, double *as, double *su) {
  const int nim1 = ni - 1, njm1 = nj - 1;
  int i, j;
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i) {
      const double de = -gam * (A1(y, j) - A1(y, j - 1)) / (A1(xc, i + 1) - A1(xc, i));
      const double dw = -gam * (A1(y, j) - A1(y, j - 1)) / (A1(xc, i) - A1(xc, i - 1));
      const double dn = -gam * (A1(x, i) - A1(x, i - 1)) / (A1(yc, j + 1) - A1(yc, j));
      const double ds = -gam * (A1(x, i) - A1(x, i - 1)) / (A1(yc, j) - A1(yc, j - 1));
      const double ue = A2(u, i, j), uw = A2(u, i - 1, j), vn = A2(v, i, j), vs = A2(v, i, j - 1);
      const double ge = den * ue * (A1(y, j) - A1(y, j - 1));
      const double gw = -den * uw * (A1(y, j) - A1(y, j - 1));
      const double gn = den * vn * (A1(x, i) - A1(x, i - 1));
      const double gs = -den * vs * (A1(x, i) - A1(x, i - 1));
      double ce, cw, cn, cs;
      if (isch == 1) {
        ce = dmin(ge, 0.0); cw = dmin(gw, 0.0); cn = dmin(gn, 0.0); cs = dmin(gs, 0.0);
      } else {
        ce = ge * A1(fx, i); cw = gw * (1.0 - A1(fx, i - 1)); cn = gn * A1(fy, j); cs = gs * (1.0 - A1(fy, j - 1));
      }
      A2(ae, i, j) = ce + de;
      A2(aw, i, j) = cw + dw;
      A2(an, i, j) = cn + dn;
      A2(as, i, j) = cs + ds;
      A2(su, i, j) = 0.0;
    }
}

void orc_stag_setup(int nicv, int njcv, int isch, double den, double gam, const double *x, const double *y, const double *xc,
                    const double *yc, const double *fx, const double *fy, double *u, double *v, double *phi, double *ae,
                    double *aw, double *an, double *as, double *ap, double *su) {
  const int ni = nicv + 2, nj = njcv + 2, nim1 = ni - 1, njm1 = nj - 1;
  const double ymin = A1(y, 1), ymax = A1(y, njm1);
  int i, j;
  for (j = 1; j <= nj; ++j)
    for (i = 1; i <= ni; ++i) {
      A2(u, i, j) = A1(x, i);
      A2(v, i, j) = -A1(y, j);
      A2(phi, i, j) = 0.0;
      A2(ae, i, j) = 0; A2(aw, i, j) = 0; A2(an, i, j) = 0; A2(as, i, j) = 0; A2(ap, i, j) = 0; A2(su, i, j) = 0;
    }
  for (j = 1; j <= njm1; ++j) A2(phi, 1, j) = 1.0 - (A1(yc, j) - ymin) / (ymax - ymin);
  stag_coeffs(ni, nj, isch, den, gam, x, y, xc, yc, fx, fy, u, v, ae, aw, an, as, su);
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i) A2(ap, i, j) = (A2(ae, i, j) + A2(aw, i, j) + A2(an, i, j) + A2(as, i, j));
  for (j = 2; j <= njm1; ++j) {   /* east: outflow, zero gradient */
    A2(ap, nim1, j) = A2(ap, nim1, j) - A2(ae, nim1, j);
    A2(ae, nim1, j) = 0.0;
  }
  for (i = 2; i <= nim1; ++i) {   /* south: symmetry */
    A2(ap, i, 2) = A2(ap, i, 2) - A2(as, i, 2);
    A2(as, i, 2) = 0.0;
  }
}

void orc_stag_tdma(int n, const double *a, const double *b, const double *c, const double *d, double *x, double *p, double *q) {
  int i;
  A1(p, 1) = 0.0;
  A1(q, 1) = A1(x, 1);
  for (i = 2; i <= n - 1; ++i) {
    A1(p, i) = A1(b, i) / (A1(a, i) - A1(c, i) * A1(p, i - 1));
    A1(q, i) = (A1(d, i) + A1(c, i) * A1(q, i - 1)) / (A1(a, i) - A1(c, i) * A1(p, i - 1));
  }
  for (i = n - 1; i >= 2; --i) A1(x, i) = A1(p, i) * A1(x, i + 1) + A1(q, i);
}

static void we_line(int ni, int nj, int i, const double *aw, const double *ae, const double *as, const double *an, const double *ap,
                    const double *su, double *phi, double *w) {
  double *a = w, *b = w + nj, *c = w + 2 * nj, *d = w + 3 * nj, *x = w + 4 * nj, *p = w + 5 * nj, *q = w + 6 * nj;
  int j;
  for (j = 2; j <= nj - 1; ++j) {
    A1(c, j) = A2(as, i, j);
    A1(a, j) = A2(ap, i, j);
    A1(b, j) = A2(an, i, j);
    A1(d, j) = A2(su, i, j) + A2(aw, i, j) * A2(phi, i - 1, j) + A2(ae, i, j) * A2(phi, i + 1, j);
  }
  A1(x, 1) = A2(phi, i, 1);
  A1(x, nj) = A2(phi, i, nj);      /* the source writes x(ni) = phi(i,ni); ni == nj in every shipped case */
  orc_stag_tdma(nj, a, b, c, d, x, p, q);
  for (j = 2; j <= nj - 1; ++j) A2(phi, i, j) = A1(x, j);
}
void orc_stag_wesweep(int ni, int nj, int order, const double *aw, const double *ae, const double *as, const double *an,
                      const double *ap, const double *su, double *phi) {
  double *w = (double *)malloc(sizeof(double) * 7 * (size_t)(ni > nj ? ni : nj));
  int i, colour;
  if (order == 0) {
    for (i = 2; i <= ni - 1; ++i) we_line(ni, nj, i, aw, ae, as, an, ap, su, phi, w);
  } else {
    for (colour = 0; colour < 2; ++colour)
      for (i = 2 + colour; i <= ni - 1; i += 2) we_line(ni, nj, i, aw, ae, as, an, ap, su, phi, w);
  }
  free(w);
}
static void sn_line(int ni, int nj, int j, const double *aw, const double *ae, const double *as, const double *an, const double *ap,
                    const double *su, double *phi, double *w) {
  double *a = w, *b = w + ni, *c = w + 2 * ni, *d = w + 3 * ni, *x = w + 4 * ni, *p = w + 5 * ni, *q = w + 6 * ni;
  int i;
  (void)nj;
  for (i = 2; i <= ni - 1; ++i) {
    A1(c, i) = A2(aw, i, j);
    A1(a, i) = A2(ap, i, j);
    A1(b, i) = A2(ae, i, j);
    A1(d, i) = A2(su, i, j) + A2(as, i, j) * A2(phi, i, j - 1) + A2(an, i, j) * A2(phi, i, j + 1);
  }
  A1(x, 1) = A2(phi, 1, j);
  A1(x, ni) = A2(phi, ni, j);
  orc_stag_tdma(ni, a, b, c, d, x, p, q);
  for (i = 2; i <= ni - 1; ++i) A2(phi, i, j) = A1(x, i);
}
void orc_stag_snsweep(int ni, int nj, int order, const double *aw, const double *ae, const double *as, const double *an,
                      const double *ap, const double *su, double *phi) {
  double *w = (double *)malloc(sizeof(double) * 7 * (size_t)(ni > nj ? ni : nj));
  int j, colour;
  if (order == 0) {
    for (j = 2; j <= nj - 1; ++j) sn_line(ni, nj, j, aw, ae, as, an, ap, su, phi, w);
  } else {
    for (colour = 0; colour < 2; ++colour)
      for (j = 2 + colour; j <= nj - 1; j += 2) sn_line(ni, nj, j, aw, ae, as, an, ap, su, phi, w);
  }
  free(w);
}
double orc_stag_residual(int ni, int nj, const double *aw, const double *ae, const double *as, const double *an, const double *ap,
                         const double *su, const double *phi) {
  double res = 0.0;
  int i, j;
  for (j = 2; j <= nj - 1; ++j)
    for (i = 2; i <= ni - 1; ++i)
      res = res + fabs(A2(ae, i, j) * A2(phi, i + 1, j) + A2(aw, i, j) * A2(phi, i - 1, j) + A2(an, i, j) * A2(phi, i, j + 1)
                       + A2(as, i, j) * A2(phi, i, j - 1) + A2(su, i, j) - A2(ap, i, j) * A2(phi, i, j));
  return res;
}
/* dir 0 = tdma_we, 1 = tdma_sn; returns the iteration count, *rsm the last normalised residual */
int orc_stag_solve(int ni, int nj, int order, int dir, int maxit, double tol, const double *aw, const double *ae, const double *as,
                   const double *an, const double *ap, const double *su, double *phi, double *rsm_out) {
  double res, res1 = 0.0, rsm = 0.0;
  int iter;
  for (iter = 1; iter <= maxit; ++iter) {
    if (dir == 0) orc_stag_wesweep(ni, nj, order, aw, ae, as, an, ap, su, phi);
    else orc_stag_snsweep(ni, nj, order, aw, ae, as, an, ap, su, phi);
    res = orc_stag_residual(ni, nj, aw, ae, as, an, ap, su, phi);
    if (iter == 1) res1 = res;
    rsm = res / res1;
    if (rsm < tol) break;
    if (iter == 1 && res1 < 1.0e-10) break;
  }
  if (iter > maxit) iter = maxit;
  *rsm_out = rsm;
  return iter;
}
double orc_stag_finish(int ni, int nj, double gam, const double *y, const double *xc, double *phi) {
  const int nim1 = ni - 1, njm1 = nj - 1;
  double fwall = 0.0;
  int i, j;
  for (i = 2; i <= nim1; ++i) A2(phi, i, 1) = A2(phi, i, 2);
  for (j = 1; j <= nj; ++j) A2(phi, ni, j) = A2(phi, nim1, j);
  for (j = 2; j <= njm1; ++j)
    fwall = fwall + gam * (A1(y, j) - A1(y, j - 1)) * (A2(phi, 2, j) - A2(phi, 1, j)) / (A1(xc, 2) - A1(xc, 1));
  return fwall;
}

/* stag_point_unsteady.f90:583-688: one time step's coefficients from phio (itsch 1 explicit, 2 implicit Euler, 3 CN) */
void orc_stag_time_terms(int nicv, int njcv, int isch, int itsch, double den, double gam, double dt, const double *x,
                         const double *y, const double *xc, const double *yc, const double *fx, const double *fy, const double *u,
                         const double *v, const double *phio, double *ae, double *aw, double *an, double *as, double *ap, double *su) {
  const int ni = nicv + 2, nj = njcv + 2, nim1 = ni - 1, njm1 = nj - 1;
  int i, j;
  for (j = 1; j <= nj; ++j)
    for (i = 1; i <= ni; ++i) A2(ap, i, j) = 0.;
  stag_coeffs(ni, nj, isch, den, gam, x, y, xc, yc, fx, fy, u, v, ae, aw, an, as, su);
  for (j = 2; j <= njm1; ++j) A2(ae, nim1, j) = 0.0;
  for (i = 2; i <= nim1; ++i) A2(as, i, 2) = 0.0;
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i) {
      const double dx = A1(x, i) - A1(x, i - 1);
      const double ct = den / dt * dx * (A1(y, j) - A1(y, j - 1));
      if (itsch == 1) {
        A2(su, i, j) = A2(su, i, j) + ct * A2(phio, i, j)
                       - (A2(ae, i, j) * A2(phio, i + 1, j) + A2(aw, i, j) * A2(phio, i - 1, j) + A2(an, i, j) * A2(phio, i, j + 1)
                          + A2(as, i, j) * A2(phio, i, j - 1))
                       + (A2(ae, i, j) + A2(aw, i, j) + A2(an, i, j) + A2(as, i, j)) * A2(phio, i, j);
        A2(ap, i, j) = ct;
        A2(ae, i, j) = 0.0; A2(aw, i, j) = 0.0; A2(an, i, j) = 0.0; A2(as, i, j) = 0.0;
      } else if (itsch == 2) {
        A2(su, i, j) = A2(su, i, j) + ct * A2(phio, i, j);
        A2(ap, i, j) = ct - (A2(ae, i, j) + A2(aw, i, j) + A2(an, i, j) + A2(as, i, j));
      } else {
        A2(ae, i, j) = 0.5 * A2(ae, i, j);
        A2(aw, i, j) = 0.5 * A2(aw, i, j);
        A2(an, i, j) = 0.5 * A2(an, i, j);
        A2(as, i, j) = 0.5 * A2(as, i, j);
        A2(su, i, j) = A2(su, i, j) + ct * A2(phio, i, j)
                       - (A2(ae, i, j) * A2(phio, i + 1, j) + A2(aw, i, j) * A2(phio, i - 1, j) + A2(an, i, j) * A2(phio, i, j + 1)
                          + A2(as, i, j) * A2(phio, i, j - 1))
                       + (A2(ae, i, j) + A2(aw, i, j) + A2(an, i, j) + A2(as, i, j)) * A2(phio, i, j);
        A2(ap, i, j) = ct - (A2(ae, i, j) + A2(aw, i, j) + A2(an, i, j) + A2(as, i, j));
      }
    }
}

/* ---- oblique convection of a step profile: fvm2d_oblique_conv.f90:161-232 ---------------------------------------- */
void orc_oblique_setup(int nicv, int njcv, double vel, int east_north_bc, const double *x, const double *y, double *u, double *v,
                       double *phi, double *ae, double *aw, double *an, double *as, double *ap, double *su) {
  const int ni = nicv + 2, nj = njcv + 2, nim1 = ni - 1, njm1 = nj - 1;
  int i, j;
  for (j = 1; j <= nj; ++j)
    for (i = 1; i <= ni; ++i) {
      A2(u, i, j) = vel; A2(v, i, j) = vel; A2(phi, i, j) = 0.0;
      A2(ae, i, j) = 0; A2(aw, i, j) = 0; A2(an, i, j) = 0; A2(as, i, j) = 0; A2(ap, i, j) = 0; A2(su, i, j) = 0;
    }
  for (j = 2; j <= njm1; ++j) A2(phi, 1, j) = 1.0;
  for (i = 2; i <= nim1; ++i) A2(phi, i, 1) = 0.0;
  A2(phi, 1, 1) = 0.5;
  if (east_north_bc == 2) {
    for (i = 1; i <= nim1; ++i) A2(phi, i, nj) = 1.0;
    for (j = 1; j <= njm1; ++j) A2(phi, ni, j) = 0.0;
    A2(phi, ni, nj) = 0.5;
  }
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i) {
      const double ue = A2(u, i, j), uw = A2(u, i - 1, j), vn = A2(v, i, j), vs = A2(v, i, j - 1);
      const double ge = ue * (A1(y, j) - A1(y, j - 1)), gw = -uw * (A1(y, j) - A1(y, j - 1));
      const double gn = vn * (A1(x, i) - A1(x, i - 1)), gs = -vs * (A1(x, i) - A1(x, i - 1));
      A2(ae, i, j) = dmin(ge, 0.0);
      A2(aw, i, j) = dmin(gw, 0.0);
      A2(an, i, j) = dmin(gn, 0.0);
      A2(as, i, j) = dmin(gs, 0.0);
      A2(ap, i, j) = -(A2(ae, i, j) + A2(aw, i, j) + A2(an, i, j) + A2(as, i, j));
      A2(su, i, j) = 0.0;
    }
  i = 2;
  for (j = 2; j <= njm1; ++j) {
    A2(su, i, j) = A2(su, i, j) - A2(aw, i, j) * A2(phi, 1, j);
    A2(aw, i, j) = 0.0;
  }
  i = nim1;
  for (j = 2; j <= njm1; ++j) {
    if (east_north_bc == 1) A2(ap, i, j) = A2(ae, i, j) + A2(ap, i, j);
    else A2(su, i, j) = A2(su, i, j) - A2(ae, i, j) * A2(phi, ni, j);
    A2(ae, i, j) = 0.0;
  }
  j = njm1;
  for (i = 2; i <= nim1; ++i) {
    if (east_north_bc == 1) A2(ap, i, j) = A2(an, i, j) + A2(ap, i, j);
    else A2(su, i, j) = A2(su, i, j) - A2(an, i, j) * A2(phi, i, nj);
    A2(an, i, j) = 0.0;
  }
  j = 2;
  for (i = 2; i <= nim1; ++i) {
    A2(su, i, j) = A2(su, i, j) - A2(as, i, j) * A2(phi, i, 1);
    A2(as, i, j) = 0.0;
  }
}
/* :348-387 */
static void oblique_tdma(int n, const double *a, const double *b, const double *c, const double *d, double *x, double *p, double *q) {
  int i;
  double denom;
  for (i = 1; i <= n; ++i) { A1(p, i) = 0.0; A1(q, i) = 0.0; }
  A1(p, 2) = A1(c, 2) / A1(b, 2);
  A1(q, 2) = A1(d, 2) / A1(b, 2);
  for (i = 3; i <= n - 1; ++i) {
    denom = A1(b, i) - A1(a, i) * A1(p, i - 1);
    A1(p, i) = A1(c, i) / denom;
    A1(q, i) = (A1(d, i) - A1(a, i) * A1(q, i - 1)) / denom;
  }
  A1(x, n - 1) = A1(q, n - 1);
  for (i = n - 2; i >= 2; --i) A1(x, i) = A1(q, i) - A1(p, i) * A1(x, i + 1);
}
static void oblique_line(int ni, int nj, int i, const double *aw, const double *ae, const double *as, const double *an,
                         const double *ap, const double *su, double *phi, double *w) {
  double *a = w, *b = w + nj, *c = w + 2 * nj, *d = w + 3 * nj, *x = w + 4 * nj, *p = w + 5 * nj, *q = w + 6 * nj;
  int j;
  for (j = 2; j <= nj - 1; ++j) {
    A1(a, j) = A2(as, i, j);
    A1(b, j) = A2(ap, i, j);
    A1(c, j) = A2(an, i, j);
    A1(d, j) = A2(su, i, j) - A2(aw, i, j) * A2(phi, i - 1, j) - A2(ae, i, j) * A2(phi, i + 1, j);
  }
  oblique_tdma(nj, a, b, c, d, x, p, q);
  for (j = 2; j <= nj - 1; ++j) A2(phi, i, j) = A1(x, j);
}
/* :284-346; returns the iteration count */
int orc_oblique_solve(int ni, int nj, int order, int maxit, double tol, const double *aw, const double *ae, const double *as,
                      const double *an, const double *ap, const double *su, double *phi, double *rsm_out) {
  double *w = (double *)malloc(sizeof(double) * 7 * (size_t)(ni > nj ? ni : nj));
  double res, res1 = 0.0, rsm = 0.0;
  int iter, i, j, colour;
  for (iter = 1; iter <= maxit; ++iter) {
    if (order == 0) {
      for (i = 2; i <= ni - 1; ++i) oblique_line(ni, nj, i, aw, ae, as, an, ap, su, phi, w);
    } else {
      for (colour = 0; colour < 2; ++colour)
        for (i = 2 + colour; i <= ni - 1; i += 2) oblique_line(ni, nj, i, aw, ae, as, an, ap, su, phi, w);
    }
    res = 0.0;
    for (j = 2; j <= nj - 1; ++j)
      for (i = 2; i <= ni - 1; ++i)
        res = res + fabs(A2(ae, i, j) * A2(phi, i + 1, j) + A2(aw, i, j) * A2(phi, i - 1, j) + A2(an, i, j) * A2(phi, i, j + 1)
                         + A2(as, i, j) * A2(phi, i, j - 1) + A2(ap, i, j) * A2(phi, i, j) - A2(su, i, j));
    if (iter == 1) res1 = res;
    rsm = res / res1;
    if (res < 1.0e-10) break;
    if (rsm < tol) break;
  }
  if (iter > maxit) iter = maxit;
  free(w);
  *rsm_out = rsm;
  return iter;
}

/* ---- the unsteady program's own solver copies: every term on the left, Dirichlet end values kept ------------------
   tdma stag_point_unsteady.f90:225-261, wesweep :159-190, snsweep :192-223, tdma_we / tdma_sn :79-157 */
static void ustag_tdma(int n, const double *a, const double *b, const double *c, const double *d, double *x, double *p, double *q) {
  int i;
  A1(p, 1) = 0.0;
  A1(q, 1) = A1(x, 1);
  for (i = 2; i <= n - 1; ++i) {
    A1(p, i) = A1(c, i) / (A1(b, i) - A1(a, i) * A1(p, i - 1));
    A1(q, i) = (A1(d, i) - A1(a, i) * A1(q, i - 1)) / (A1(b, i) - A1(a, i) * A1(p, i - 1));
  }
  for (i = n - 1; i >= 2; --i) A1(x, i) = A1(q, i) - A1(p, i) * A1(x, i + 1);
}
static void ustag_we_line(int ni, int nj, int i, const double *aw, const double *ae, const double *as, const double *an,
                          const double *ap, const double *su, double *phi, double *w) {
  double *a = w, *b = w + nj, *c = w + 2 * nj, *d = w + 3 * nj, *x = w + 4 * nj, *p = w + 5 * nj, *q = w + 6 * nj;
  int j;
  for (j = 2; j <= nj - 1; ++j) {
    A1(a, j) = A2(as, i, j);
    A1(b, j) = A2(ap, i, j);
    A1(c, j) = A2(an, i, j);
    A1(d, j) = A2(su, i, j) - A2(aw, i, j) * A2(phi, i - 1, j) - A2(ae, i, j) * A2(phi, i + 1, j);
  }
  A1(x, 1) = A2(phi, i, 1);
  A1(x, nj) = A2(phi, i, nj);
  ustag_tdma(nj, a, b, c, d, x, p, q);
  for (j = 2; j <= nj - 1; ++j) A2(phi, i, j) = A1(x, j);
}
static void ustag_sn_line(int ni, int j, const double *aw, const double *ae, const double *as, const double *an, const double *ap,
                          const double *su, double *phi, double *w) {
  double *a = w, *b = w + ni, *c = w + 2 * ni, *d = w + 3 * ni, *x = w + 4 * ni, *p = w + 5 * ni, *q = w + 6 * ni;
  int i;
  for (i = 2; i <= ni - 1; ++i) {
    A1(a, i) = A2(aw, i, j);
    A1(b, i) = A2(ap, i, j);
    A1(c, i) = A2(ae, i, j);
    A1(d, i) = A2(su, i, j) - A2(as, i, j) * A2(phi, i, j - 1) - A2(an, i, j) * A2(phi, i, j + 1);
  }
  A1(x, 1) = A2(phi, 1, j);
  A1(x, ni) = A2(phi, ni, j);
  ustag_tdma(ni, a, b, c, d, x, p, q);
  for (i = 2; i <= ni - 1; ++i) A2(phi, i, j) = A1(x, i);
}
int orc_ustag_solve(int ni, int nj, int order, int dir, int maxit, double tol, const double *aw, const double *ae, const double *as,
                    const double *an, const double *ap, const double *su, double *phi, double *rsm_out) {
  double *w = (double *)malloc(sizeof(double) * 7 * (size_t)(ni > nj ? ni : nj));
  double res, res1 = 0.0, rsm = 0.0;
  int iter, i, j, colour;
  for (iter = 1; iter <= maxit; ++iter) {
    if (dir == 0) {
      if (order == 0) {
        for (i = 2; i <= ni - 1; ++i) ustag_we_line(ni, nj, i, aw, ae, as, an, ap, su, phi, w);
      } else {
        for (colour = 0; colour < 2; ++colour)
          for (i = 2 + colour; i <= ni - 1; i += 2) ustag_we_line(ni, nj, i, aw, ae, as, an, ap, su, phi, w);
      }
    } else {
      if (order == 0) {
        for (j = 2; j <= nj - 1; ++j) ustag_sn_line(ni, j, aw, ae, as, an, ap, su, phi, w);
      } else {
        for (colour = 0; colour < 2; ++colour)
          for (j = 2 + colour; j <= nj - 1; j += 2) ustag_sn_line(ni, j, aw, ae, as, an, ap, su, phi, w);
      }
    }
    res = 0.0;
    for (j = 2; j <= nj - 1; ++j)
      for (i = 2; i <= ni - 1; ++i)
        res = res + fabs(A2(ae, i, j) * A2(phi, i + 1, j) + A2(aw, i, j) * A2(phi, i - 1, j) + A2(an, i, j) * A2(phi, i, j + 1)
                         + A2(as, i, j) * A2(phi, i, j - 1) + A2(ap, i, j) * A2(phi, i, j) - A2(su, i, j));
    if (iter == 1) res1 = res;
    rsm = res / res1;
    if (rsm < tol) break;
    if (iter == 1 && res1 < 1.0e-10) break;
  }
  if (iter > maxit) iter = maxit;
  free(w);
  *rsm_out = rsm;
  return iter;
}
