/*
 * oracle/fdcavity_oracle.c -- CPU restatement of the reference's explicit finite-difference cavity solver
 * (primitive variables, staggered grid, explicit Euler predictor + p' Poisson equation solved by Stone's SIP).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/cavity_oracle.c header for the rule).
 *
 * Source: /root/reference/lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90
 *   orc_fdc_sipsol   <- sipsol            :24-87   (maxit 600, alpha 0.88, tol 1e-5)
 *   orc_fdc_init     <- initial condition :160-166
 *   orc_fdc_predict  <- u/v momentum predictor, p' coefficients and source, pp = 0   :168-205
 *   orc_fdc_correct  <- correction of p,u,v, mirror-ghost boundary conditions, continuity norm, uo=u... :207-254
 *   orc_fdc_step     <- one pass of the time loop :168-255
 *
 * All 2-D arrays are stored with the common leading dimension ld = ni+1 and nj+1 columns (the reference's
 * u(1:ni,1:nj+1), v(1:ni+1,1:nj), p(1:ni+1,1:nj+1) are the leading parts).  A(i,j) at a[(i-1)+(j-1)*ld].
 * Quirk kept: the west-boundary pressure copy runs over 1:ni+1 although the extent is nj+1 (:234); it is
 * clamped to the array here (identical whenever ni == nj, which every shipped case has).
 * PARITY PIN: none exists (no reference outputs); consistency checks only -- "parity unpinned".
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct orc_fdc {
  int ni, nj;
  double re, dt, utop, urfp;
  double *u, *v, *p, *uo, *vo, *po, *pp, *su;
  int sip_iters;      /* iterations used by the last sipsol call */
} orc_fdc;

#define A2(a, i, j) ((a)[(size_t)((i) - 1) + (size_t)((j) - 1) * (size_t)ld])

/* :24-87.  Coefficient arrays are passed as scalars here because the caller's are constant over the grid
   (aw = ae = dt/dx**2, an = as = dt/dy**2, negated by the call :206, ap = 2*(dt/dx**2+dt/dy**2)); interior
   (i1+1..i2-1, j1+1..j2-1) = (2..ni, 2..nj). */
int orc_fdc_sipsol(int ni, int nj, double aw, double ae, double as, double an, double ap, const double *su, double *phi) {
  const int ld = ni + 1, i1 = 1, i2 = ni + 1, j1 = 1, j2 = nj + 1, maxit = 600;
  const double alpha = 0.88, tol = 1.0e-5;
  size_t n = (size_t)ld * (size_t)(nj + 1);
  double *lw = calloc(n, 8), *ls = calloc(n, 8), *lpr = calloc(n, 8), *un = calloc(n, 8), *ue = calloc(n, 8), *res = calloc(n, 8);
  double p1, p2, rsm, resl, res1 = 0.0;
  int i, j, iter, used = 0;
  for (j = j1 + 1; j <= j2 - 1; ++j)
    for (i = i1 + 1; i <= i2 - 1; ++i) {
      A2(lw, i, j) = aw / (1.0 + alpha * A2(un, i - 1, j));
      A2(ls, i, j) = as / (1.0 + alpha * A2(ue, i, j - 1));
      p1 = alpha * A2(lw, i, j) * A2(un, i - 1, j);
      p2 = alpha * A2(ls, i, j) * A2(ue, i, j - 1);
      A2(lpr, i, j) = 1.0 / (ap + p1 + p2 - A2(lw, i, j) * A2(ue, i - 1, j) - A2(ls, i, j) * A2(un, i, j - 1));
      A2(un, i, j) = (an - p1) * A2(lpr, i, j);
      A2(ue, i, j) = (ae - p2) * A2(lpr, i, j);
    }
  for (iter = 1; iter <= maxit; ++iter) {
    used = iter;
    resl = 0.0;
    for (j = j1 + 1; j <= j2 - 1; ++j)
      for (i = i1 + 1; i <= i2 - 1; ++i) {
        A2(res, i, j) = A2(su, i, j) - aw * A2(phi, i - 1, j) - ae * A2(phi, i + 1, j) - an * A2(phi, i, j + 1)
                        - as * A2(phi, i, j - 1) - ap * A2(phi, i, j);
        resl = resl + fabs(A2(res, i, j));
        A2(res, i, j) = (A2(res, i, j) - A2(ls, i, j) * A2(res, i, j - 1) - A2(lw, i, j) * A2(res, i - 1, j)) * A2(lpr, i, j);
      }
    if (iter == 1) res1 = resl;
    rsm = resl / res1;
    for (j = j2 - 1; j >= j1 + 1; --j)
      for (i = i2 - 1; i >= i1 + 1; --i) {
        A2(res, i, j) = A2(res, i, j) - A2(un, i, j) * A2(res, i, j + 1) - A2(ue, i, j) * A2(res, i + 1, j);
        A2(phi, i, j) = A2(phi, i, j) + A2(res, i, j);
      }
    if (rsm < tol) break;
    else if (iter == 1 && res1 < 1.0e-10) break;
  }
  free(lw); free(ls); free(lpr); free(un); free(ue); free(res);
  return used;
}

/* :160-166 */
void orc_fdc_init(orc_fdc *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  int i, j;
  size_t n = (size_t)ld * (size_t)(nj + 1);
  memset(s->u, 0, n * 8); memset(s->v, 0, n * 8); memset(s->p, 0, n * 8); memset(s->pp, 0, n * 8);
  memset(s->uo, 0, n * 8); memset(s->vo, 0, n * 8); memset(s->po, 0, n * 8); memset(s->su, 0, n * 8);
  for (i = 1; i <= ni; ++i) { A2(s->uo, i, nj) = 1.0; A2(s->uo, i, nj + 1) = 1.0; }
  for (j = 1; j <= nj + 1; ++j) for (i = 1; i <= ni + 1; ++i) A2(s->po, i, j) = 1.0;
}

/* :168-205 */
void orc_fdc_predict(orc_fdc *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), dt = s->dt, re = s->re;
  double *u = s->u, *v = s->v, *uo = s->uo, *vo = s->vo, *po = s->po, *su = s->su;
  int i, j;
  for (j = 2; j <= nj; ++j)
    for (i = 2; i <= ni - 1; ++i) {
      A2(u, i, j) = -0.5 * (A2(uo, i + 1, j) * A2(uo, i + 1, j) - A2(uo, i - 1, j) * A2(uo, i - 1, j)) / dx
                    - 0.25 * ((A2(uo, i, j) + A2(uo, i, j + 1)) * (A2(vo, i, j) + A2(vo, i + 1, j))
                              - (A2(uo, i, j) + A2(uo, i, j - 1)) * (A2(vo, i, j - 1) + A2(vo, i + 1, j - 1))) / dy
                    - (A2(po, i + 1, j) - A2(po, i, j)) / dx
                    + ((A2(uo, i - 1, j) - 2.0 * A2(uo, i, j) + A2(uo, i + 1, j)) / (dx * dx)
                       + (A2(uo, i, j - 1) - 2.0 * A2(uo, i, j) + A2(uo, i, j + 1)) / (dy * dy)) / re;
      A2(u, i, j) = A2(uo, i, j) + dt * A2(u, i, j);
    }
  for (j = 2; j <= nj - 1; ++j)
    for (i = 2; i <= ni; ++i) {
      A2(v, i, j) = -0.25 * ((A2(uo, i, j) + A2(uo, i, j + 1)) * (A2(vo, i, j) + A2(vo, i + 1, j))
                             - (A2(uo, i - 1, j) + A2(uo, i - 1, j + 1)) * (A2(vo, i, j) + A2(vo, i - 1, j))) / dx
                    - 0.5 * (A2(vo, i, j + 1) * A2(vo, i, j + 1) - A2(vo, i, j - 1) * A2(vo, i, j - 1)) / dy
                    - (A2(po, i, j + 1) - A2(po, i, j)) / dy
                    + ((A2(vo, i - 1, j) - 2.0 * A2(vo, i, j) + A2(vo, i + 1, j)) / (dx * dx)
                       + (A2(vo, i, j - 1) - 2.0 * A2(vo, i, j) + A2(vo, i, j + 1)) / (dy * dy)) / re;
      A2(v, i, j) = A2(vo, i, j) + dt * A2(v, i, j);
    }
  memset(s->pp, 0, (size_t)ld * (size_t)(nj + 1) * 8);
  for (j = 2; j <= nj; ++j)
    for (i = 2; i <= ni; ++i)
      A2(su, i, j) = -(A2(u, i, j) - A2(u, i - 1, j)) / dx - (A2(v, i, j) - A2(v, i, j - 1)) / dy;
}

void orc_fdc_solve_pp(orc_fdc *s) {
  const int ni = s->ni, nj = s->nj;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), dt = s->dt;
  const double ax = dt / (dx * dx), ay = dt / (dy * dy);
  s->sip_iters = orc_fdc_sipsol(ni, nj, -ax, -ax, -ay, -ay, 2 * (dt / (dx * dx) + dt / (dy * dy)), s->su, s->pp);
}

/* :207-254 ; returns the continuity norm `error` (:239-247) */
double orc_fdc_correct(orc_fdc *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), dt = s->dt;
  double *u = s->u, *v = s->v, *p = s->p, *pp = s->pp;
  double error = 0.0, resnorm;
  int i, j, m;
  size_t n = (size_t)ld * (size_t)(nj + 1);
  for (j = 2; j <= nj; ++j) for (i = 2; i <= ni; ++i) A2(p, i, j) = A2(p, i, j) + s->urfp * A2(pp, i, j);
  for (j = 2; j <= nj; ++j) for (i = 2; i <= ni - 1; ++i) A2(u, i, j) = A2(u, i, j) - dt / dx * (A2(pp, i + 1, j) - A2(pp, i, j));
  for (j = 2; j <= nj - 1; ++j) for (i = 2; i <= ni; ++i) A2(v, i, j) = A2(v, i, j) - dt / dy * (A2(pp, i, j + 1) - A2(pp, i, j));
  for (i = 1; i <= ni; ++i) A2(u, i, 1) = -A2(u, i, 2);
  for (i = 1; i <= ni; ++i) A2(u, i, nj + 1) = 2.0 * s->utop - A2(u, i, nj);
  for (j = 2; j <= nj; ++j) A2(u, 1, j) = 0.0;
  for (j = 2; j <= nj; ++j) A2(u, ni, j) = 0.0;
  for (j = 1; j <= nj; ++j) A2(v, 1, j) = -A2(v, 2, j);
  for (j = 1; j <= nj; ++j) A2(v, ni + 1, j) = -A2(v, ni, j);
  for (i = 2; i <= ni; ++i) A2(v, i, 1) = 0.0;
  for (i = 2; i <= ni; ++i) A2(v, i, nj) = 0.0;
  m = (ni + 1 < nj + 1) ? ni + 1 : nj + 1;           /* p(1,1:ni+1) = p(2,1:ni+1) : ni used for nj (:234) */
  for (j = 1; j <= m; ++j) A2(p, 1, j) = A2(p, 2, j);
  for (j = 1; j <= nj + 1; ++j) A2(p, ni + 1, j) = A2(p, ni, j);
  for (i = 2; i <= ni; ++i) A2(p, i, nj + 1) = A2(p, i, nj);
  for (i = 2; i <= ni; ++i) A2(p, i, 1) = A2(p, i, 2);
  for (j = 2; j <= nj; ++j)
    for (i = 2; i <= ni; ++i) {
      double d = (A2(u, i, j) - A2(u, i - 1, j)) / dx + (A2(v, i, j) - A2(v, i, j - 1)) / dy;
      resnorm = d * d;
      error = error + resnorm;
    }
  error = sqrt(error);
  memcpy(s->uo, u, n * 8);
  memcpy(s->vo, v, n * 8);
  memcpy(s->po, p, n * 8);
  return error;
}

double orc_fdc_step(orc_fdc *s) {
  orc_fdc_predict(s);
  orc_fdc_solve_pp(s);
  return orc_fdc_correct(s);
}

/* ==============================================================================================================
 * Implicit-Euler FD cavity: /root/reference/lid-driven-cavity-2d/fdm-implicit/fdm_implicit_cavity.f90
 *   orc_sipsol_arr   <- sipsol :29-92 (array coefficients, maxit 3000, alpha 0.80, tol 1e-6); phi is in/out
 *   orc_fdci_*       <- time loop :177-288 (linearised u, v coefficient assembly, p' equation, correction, BCs, norm)
 * All arrays (ni+1, nj+1), ld = ni+1.  The caller's negated sections (-aw, -ae, -as, -an) are formed here.
 * ============================================================================================================== */
int orc_sipsol_arr(int ld, int nrow, int i2, int j2, const double *aw, const double *ae, const double *as,
                   const double *an, const double *ap, const double *su, double *phi, int maxit, double alpha, double tol) {
  const int i1 = 1, j1 = 1;
  size_t n = (size_t)ld * (size_t)nrow;
  double *lw = calloc(n, 8), *ls = calloc(n, 8), *lpr = calloc(n, 8), *un = calloc(n, 8), *ue = calloc(n, 8), *res = calloc(n, 8);
  double p1, p2, rsm, resl, res1 = 0.0;
  int i, j, iter, used = 0;
  for (j = j1 + 1; j <= j2 - 1; ++j)
    for (i = i1 + 1; i <= i2 - 1; ++i) {
      A2(lw, i, j) = -A2(aw, i, j) / (1.0 + alpha * A2(un, i - 1, j));
      A2(ls, i, j) = -A2(as, i, j) / (1.0 + alpha * A2(ue, i, j - 1));
      p1 = alpha * A2(lw, i, j) * A2(un, i - 1, j);
      p2 = alpha * A2(ls, i, j) * A2(ue, i, j - 1);
      A2(lpr, i, j) = 1.0 / (A2(ap, i, j) + p1 + p2 - A2(lw, i, j) * A2(ue, i - 1, j) - A2(ls, i, j) * A2(un, i, j - 1));
      A2(un, i, j) = (-A2(an, i, j) - p1) * A2(lpr, i, j);
      A2(ue, i, j) = (-A2(ae, i, j) - p2) * A2(lpr, i, j);
    }
  for (iter = 1; iter <= maxit; ++iter) {
    used = iter;
    resl = 0.0;
    for (j = j1 + 1; j <= j2 - 1; ++j)
      for (i = i1 + 1; i <= i2 - 1; ++i) {
        A2(res, i, j) = A2(su, i, j) - (-A2(aw, i, j)) * A2(phi, i - 1, j) - (-A2(ae, i, j)) * A2(phi, i + 1, j)
                        - (-A2(an, i, j)) * A2(phi, i, j + 1) - (-A2(as, i, j)) * A2(phi, i, j - 1) - A2(ap, i, j) * A2(phi, i, j);
        resl = resl + fabs(A2(res, i, j));
        A2(res, i, j) = (A2(res, i, j) - A2(ls, i, j) * A2(res, i, j - 1) - A2(lw, i, j) * A2(res, i - 1, j)) * A2(lpr, i, j);
      }
    if (iter == 1) res1 = resl;
    rsm = resl / res1;
    for (j = j2 - 1; j >= j1 + 1; --j)
      for (i = i2 - 1; i >= i1 + 1; --i) {
        A2(res, i, j) = A2(res, i, j) - A2(un, i, j) * A2(res, i, j + 1) - A2(ue, i, j) * A2(res, i + 1, j);
        A2(phi, i, j) = A2(phi, i, j) + A2(res, i, j);
      }
    if (rsm < tol) break;
    else if (iter == 1 && res1 < 1.0e-10) break;
  }
  free(lw); free(ls); free(lpr); free(un); free(ue); free(res);
  return used;
}

typedef struct orc_fdci {
  int ni, nj, steady;
  double re, dt, utop, urfu, urfv, urfp;
  double *u, *v, *p, *pp, *ae, *aw, *an, *as, *ap, *su;
  int sip_iters[3];
} orc_fdci;

/* :160-164 */
void orc_fdci_init(orc_fdci *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  int i;
  size_t n = (size_t)ld * (size_t)(nj + 1);
  double *all[10] = {s->u, s->v, s->p, s->pp, s->ae, s->aw, s->an, s->as, s->ap, s->su};
  for (i = 0; i < 10; ++i) memset(all[i], 0, n * 8);
  for (i = 1; i <= ni; ++i) { A2(s->u, i, nj) = 1.0; A2(s->u, i, nj + 1) = 1.0; }
}

/* :178-198 */
void orc_fdci_assemble_u(orc_fdci *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), re = s->re, dt = s->dt, urfu = s->urfu;
  const double rex = 1.0 / re / (dx * dx), rey = 1.0 / re / (dy * dy), pterm = 1.0 / dt + 2.0 * (rex + rey);
  double *u = s->u, *v = s->v, *p = s->p, vel1, vel2;
  int i, j;
  for (j = 2; j <= nj; ++j)
    for (i = 2; i <= ni - 1; ++i) {
      vel1 = 0.5 * (A2(v, i, j) + A2(v, i + 1, j));
      vel2 = 0.5 * (A2(v, i, j - 1) + A2(v, i + 1, j - 1));
      A2(s->ae, i, j) = -0.5 * A2(u, i + 1, j) / dx + rex;
      A2(s->aw, i, j) = 0.5 * A2(u, i - 1, j) / dx + rex;
      A2(s->an, i, j) = -0.5 * vel1 / dy + rey;
      A2(s->as, i, j) = 0.5 * vel2 / dy + rey;
      if (s->steady) {
        A2(s->su, i, j) = (A2(p, i, j) - A2(p, i + 1, j)) / dx;
        A2(s->ap, i, j) = 2.0 * (rex + rey);
      } else {
        A2(s->su, i, j) = A2(u, i, j) / dt + (A2(p, i, j) - A2(p, i + 1, j)) / dx;
        A2(s->ap, i, j) = pterm;
      }
      A2(s->su, i, j) = A2(s->su, i, j) + (1 - urfu) * A2(s->ap, i, j) / urfu * A2(u, i, j);
      A2(s->ap, i, j) = A2(s->ap, i, j) / urfu;
    }
}
/* :201-220 */
void orc_fdci_assemble_v(orc_fdci *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), re = s->re, dt = s->dt, urfv = s->urfv;
  const double rex = 1.0 / re / (dx * dx), rey = 1.0 / re / (dy * dy), pterm = 1.0 / dt + 2.0 * (rex + rey);
  double *u = s->u, *v = s->v, *p = s->p, vel1, vel2;
  int i, j;
  for (j = 2; j <= nj - 1; ++j)
    for (i = 2; i <= ni; ++i) {
      vel1 = 0.5 * (A2(u, i, j) + A2(u, i, j + 1));
      vel2 = 0.5 * (A2(u, i - 1, j) + A2(u, i - 1, j + 1));
      A2(s->ae, i, j) = -0.5 * vel1 / dx + rex;
      A2(s->aw, i, j) = 0.5 * vel2 / dx + rex;
      A2(s->an, i, j) = -0.5 * A2(v, i, j + 1) / dy + rey;
      A2(s->as, i, j) = -0.5 * A2(v, i, j - 1) / dy + rey;      /* sign as in the source (:208), SURVEY 8a a12 */
      if (s->steady) {
        A2(s->su, i, j) = (A2(p, i, j) - A2(p, i, j + 1)) / dy;
        A2(s->ap, i, j) = 2.0 * (rex + rey);
      } else {
        A2(s->su, i, j) = A2(v, i, j) / dt + (A2(p, i, j) - A2(p, i, j + 1)) / dy;
        A2(s->ap, i, j) = pterm;
      }
      A2(s->su, i, j) = A2(s->su, i, j) + (1 - urfv) * A2(s->ap, i, j) / urfv * A2(v, i, j);
      A2(s->ap, i, j) = A2(s->ap, i, j) / urfv;
    }
}
/* :223-240 */
void orc_fdci_assemble_p(orc_fdci *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), re = s->re, dt = s->dt;
  const double rex = 1.0 / re / (dx * dx), rey = 1.0 / re / (dy * dy), pterm = 1.0 / dt + 2.0 * (rex + rey);
  int i, j;
  memset(s->pp, 0, (size_t)ld * (size_t)(nj + 1) * 8);
  for (j = 2; j <= nj; ++j)
    for (i = 2; i <= ni; ++i) {
      if (s->steady) {
        A2(s->ae, i, j) = 1.0 / (2.0 * (rex + rey)) / (dx * dx);
        A2(s->aw, i, j) = 1.0 / (2.0 * (rex + rey)) / (dx * dx);
        A2(s->an, i, j) = 1.0 / (2.0 * (rex + rey)) / (dy * dy);
        A2(s->as, i, j) = 1.0 / (2.0 * (rex + rey)) / (dy * dy);
      } else {
        A2(s->ae, i, j) = 1.0 / pterm / (dx * dx);
        A2(s->aw, i, j) = 1.0 / pterm / (dx * dx);
        A2(s->an, i, j) = 1.0 / pterm / (dy * dy);
        A2(s->as, i, j) = 1.0 / pterm / (dy * dy);
      }
      A2(s->ap, i, j) = A2(s->ae, i, j) + A2(s->aw, i, j) + A2(s->an, i, j) + A2(s->as, i, j);
      A2(s->su, i, j) = (A2(s->u, i - 1, j) - A2(s->u, i, j)) / dx + (A2(s->v, i, j - 1) - A2(s->v, i, j)) / dy;
    }
}
int orc_fdci_solve(orc_fdci *s, int which) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  double *phi = which == 0 ? s->u : which == 1 ? s->v : s->pp;
  const int i2 = which == 0 ? ni : ni + 1, j2 = which == 1 ? nj : nj + 1;
  int n = orc_sipsol_arr(ld, nj + 1, i2, j2, s->aw, s->ae, s->as, s->an, s->ap, s->su, phi, 3000, 0.80, 1.0e-6);
  s->sip_iters[which] = n;
  return n;
}
/* :243-282 ; returns `error` */
double orc_fdci_correct(orc_fdci *s) {
  const int ni = s->ni, nj = s->nj, ld = ni + 1;
  const double dx = 1.0 / (ni - 1), dy = 1.0 / (nj - 1), re = s->re, dt = s->dt;
  const double rex = 1.0 / re / (dx * dx), rey = 1.0 / re / (dy * dy), pterm = 1.0 / dt + 2.0 * (rex + rey);
  double *u = s->u, *v = s->v, *p = s->p, *pp = s->pp, error = 0.0;
  int i, j, m;
  for (j = 2; j <= nj; ++j) for (i = 2; i <= ni; ++i) A2(p, i, j) = A2(p, i, j) + s->urfp * A2(pp, i, j);
  for (j = 2; j <= nj; ++j) for (i = 2; i <= ni - 1; ++i) A2(u, i, j) = A2(u, i, j) - (A2(pp, i + 1, j) - A2(pp, i, j)) / pterm / dx;
  for (j = 2; j <= nj - 1; ++j) for (i = 2; i <= ni; ++i) A2(v, i, j) = A2(v, i, j) - (A2(pp, i, j + 1) - A2(pp, i, j)) / pterm / dy;
  for (i = 1; i <= ni; ++i) A2(u, i, 1) = -A2(u, i, 2);
  for (i = 1; i <= ni; ++i) A2(u, i, nj + 1) = 2.0 * s->utop - A2(u, i, nj);
  for (j = 2; j <= nj; ++j) A2(u, 1, j) = 0.0;
  for (j = 2; j <= nj; ++j) A2(u, ni, j) = 0.0;
  for (j = 1; j <= nj; ++j) A2(v, 1, j) = -A2(v, 2, j);
  for (j = 1; j <= nj; ++j) A2(v, ni + 1, j) = -A2(v, ni, j);
  for (i = 2; i <= ni; ++i) A2(v, i, 1) = 0.0;
  for (i = 2; i <= ni; ++i) A2(v, i, nj) = 0.0;
  m = (ni + 1 < nj + 1) ? ni + 1 : nj + 1;
  for (j = 1; j <= m; ++j) A2(p, 1, j) = A2(p, 2, j);
  for (j = 1; j <= nj + 1; ++j) A2(p, ni + 1, j) = A2(p, ni, j);
  for (i = 2; i <= ni; ++i) A2(p, i, nj + 1) = A2(p, i, nj);
  for (i = 2; i <= ni; ++i) A2(p, i, 1) = A2(p, i, 2);
  for (j = 2; j <= nj; ++j)
    for (i = 2; i <= ni; ++i) {
      double d = (A2(u, i, j) - A2(u, i - 1, j)) / dx + (A2(v, i, j) - A2(v, i, j - 1)) / dy;
      error = error + d * d;
    }
  return sqrt(error);
}
double orc_fdci_step(orc_fdci *s) {
  orc_fdci_assemble_u(s); orc_fdci_solve(s, 0);
  orc_fdci_assemble_v(s); orc_fdci_solve(s, 1);
  orc_fdci_assemble_p(s); orc_fdci_solve(s, 2);
  return orc_fdci_correct(s);
}
