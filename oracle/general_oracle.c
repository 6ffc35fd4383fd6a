/*
 * oracle/general_oracle.c -- CPU restatement of the reference's "general" SIMPLE family (Cartesian / axisymmetric
 * cylindrical / polar, DAPEC scheme selector, coupled temperature equation).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/cavity_oracle.c header for the rule).
 *
 * Sources restated (paths under /root/reference/):
 *   orc_gen_grid    <- grid      buoyancy-driven-cavity-2d/buoyancy_cav.f90:28-120 (== cylin_cond.f90:315-407)
 *   DAPEC/ARIMEAN/HARMEAN <- module serv, cylinder-heat-conduction-2d/cylin_cond.f90:245-285
 *   orc_gen_calcu   <- calcu     buoyancy_cav.f90:122-198
 *   orc_gen_calcv   <- calcv     buoyancy_cav.f90:200-277
 *   orc_gen_calcp   <- calcp     buoyancy_cav.f90:279-356
 *   orc_gen_calct   <- calct     buoyancy_cav.f90:358-433
 *   orc_gen_tdma    <- tdma      buoyancy_cav.f90:471-505 (order 0 reference W->E lines, 1 red/black lines)
 *   user module "cylinder" <- module user, cylin_cond.f90:73-169 (control/initial/boundary/gamsor as shipped)
 *   user module "buoyancy": the reference's buoyancy_cav.f90 `use`s a module `user` that is NOT in the
 *                   repository.  The one here is authored by this build after the cylinder
 *                   template and SURVEY 8a's notes: Cartesian unit square, T=1 west / T=0 east, adiabatic N/S
 *                   (gam = 0 on those boundary rows for nq=4), Boussinesq source for the v equation
 *                   su = rho*g*beta*(T_v - Tref) with T_v = ARIMEAN(t(i,j-1), t(i,j), fy(j)).  Its parity is
 *                   therefore defined by this file, not by a reference output.
 *   orc_gen_run     <- program main, cylin_cond.f90:289-312 / buoyancy_cav.f90:2-26
 *
 * Quirks kept on purpose: `do i = 3, njm1` / `i == njm1` in calcu use nj for ni;
 * `dene = den(nj,j)` in calct; `diff = area*gam(i,1)*sns(2)` (multiply) in calcv; ce in calcp uses den(i+1,j)
 * and ignores dene; su accumulation order in calcp; den*cp ... den/cp round trip in calct.
 *
 * PARITY PIN: the known-answer table for the cylinder case in SURVEY.md 8(c)-3 (tests/golden/cylinder_kat.json).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct orc_general {
  int ni, nj, icrd, iadv, ipref, jpref, order;
  int user;                      /* 1 cylinder (as shipped), 2 buoyancy (authored) */
  int lsolve[4];                 /* u(=v,p), -, -, T : lsolve(1) and lsolve(4) of the reference */
  int nsweep[4];
  double urf[4], resor[4];
  double xstart, xend, ystart, yend, rstart;
  /* user-module parameters */
  int ibnd, is1, is2, js1, js2;
  double t0, gam0, gam1, htc, tf, qin;            /* cylinder */
  double visc, cond, rgbeta, tref;                /* buoyancy: mu, k, rho*g*beta, reference temperature */
  int llast;
  /* 1-D geometry */
  double *x, *xu, *xdif, *sew, *sewu, *xcvi, *xcvip, *fx, *fxm;
  double *y, *yv, *ydif, *sns, *snsv, *ycvr, *ycvrs, *arx, *arxj, *arxjp, *r, *rmn, *sx, *sxmn, *fv, *fvp, *fy, *fym;
  /* 2-D fields */
  double *u, *v, *p, *pp, *du, *dv, *t, *den, *gam, *cp, *ae, *aw, *an, *as, *ap, *su, *sp;
} orc_general;

#define IX(i, j) ((size_t)((i) - 1) + (size_t)((j) - 1) * (size_t)ni)
#define X1(a, i) ((a)[(i) - 1])
static inline double dmax(double a, double b) { return a > b ? a : b; }

/* cylin_cond.f90:245-269 */
static double DAPEC(double flow, double diff, int iadv) {
  double d;
  if (flow == 0.0) return diff;
  switch (iadv) {
    case 1: return diff - 0.5 * fabs(flow);
    case 2: return diff;
    case 3: return dmax(0.0, diff - 0.5 * fabs(flow));
    case 4: {
      double q = 1.0 - 0.1 * fabs(flow / diff);
      double q2 = q * q;
      d = dmax(0.0, q2 * q2 * q);             /* x**5 as gfortran's powi expansion: (x*x)*(x*x)*x */
      return d * diff;
    }
    default: return diff;
  }
}
/* cylin_cond.f90:271-285 */
static double ARIMEAN(double phi1, double phi2, double wf2) { return phi1 * (1 - wf2) + phi2 * wf2; }
static double HARMEAN(double phi1, double phi2, double wf2) { return phi1 * phi2 / (phi1 * (1.0 - wf2) + phi2 * wf2); }

/* buoyancy_cav.f90:28-120 ; all arrays must come zero-filled (array_init, :515-560) */
void orc_gen_grid(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, nim2 = ni - 2, njm1 = nj - 1, njm2 = nj - 2, icrd = s->icrd;
  double *x = s->x, *xu = s->xu, *xdif = s->xdif, *sew = s->sew, *sewu = s->sewu, *xcvi = s->xcvi, *xcvip = s->xcvip;
  double *fx = s->fx, *fxm = s->fxm, *y = s->y, *yv = s->yv, *ydif = s->ydif, *sns = s->sns, *snsv = s->snsv;
  double *ycvr = s->ycvr, *ycvrs = s->ycvrs, *arx = s->arx, *arxj = s->arxj, *arxjp = s->arxjp, *r = s->r, *rmn = s->rmn;
  double *sx = s->sx, *sxmn = s->sxmn, *fv = s->fv, *fvp = s->fvp, *fy = s->fy, *fym = s->fym;
  int i, j;
  double dx, dy;
  dx = (s->xend - s->xstart) / (ni - 2);
  X1(xu, 2) = s->xstart;
  for (i = 3; i <= ni; ++i) X1(xu, i) = X1(xu, i - 1) + dx;
  for (i = 2; i <= nim1; ++i) X1(sew, i) = X1(xu, i + 1) - X1(xu, i);
  X1(x, 1) = X1(xu, 2);
  for (i = 2; i <= nim1; ++i) X1(x, i) = 0.5 * (X1(xu, i) + X1(xu, i + 1));
  X1(x, ni) = X1(xu, ni);
  for (i = 2; i <= ni; ++i) X1(xdif, i) = X1(x, i) - X1(x, i - 1);
  for (i = 3; i <= nim1; ++i) X1(sewu, i) = X1(xdif, i);
  X1(sewu, 3) = X1(sewu, 3) + X1(xdif, 2);
  X1(sewu, nim1) = X1(sewu, nim1) + X1(xdif, ni);
  for (i = 3; i <= nim2; ++i) X1(xcvi, i) = 0.5 * X1(sew, i);
  for (i = 3; i <= nim2; ++i) X1(xcvip, i) = X1(xcvi, i);
  X1(xcvip, 2) = X1(sew, 2);
  X1(xcvi, nim1) = X1(sew, nim1);
  dy = (s->yend - s->ystart) / (nj - 2);
  X1(yv, 2) = s->ystart;
  for (j = 3; j <= nj; ++j) X1(yv, j) = X1(yv, j - 1) + dy;
  for (j = 2; j <= njm1; ++j) X1(sns, j) = X1(yv, j + 1) - X1(yv, j);
  X1(y, 1) = X1(yv, 2);
  for (j = 2; j <= njm1; ++j) X1(y, j) = 0.5 * (X1(yv, j) + X1(yv, j + 1));
  X1(y, nj) = X1(yv, nj);
  for (j = 2; j <= nj; ++j) X1(ydif, j) = X1(y, j) - X1(y, j - 1);
  for (j = 3; j <= njm1; ++j) X1(snsv, j) = X1(ydif, j);
  X1(snsv, 3) = X1(snsv, 3) + X1(ydif, 2);
  X1(snsv, njm1) = X1(snsv, njm1) + X1(ydif, nj);
  X1(r, 1) = s->rstart;
  if (icrd == 1) {
    for (j = 1; j <= nj; ++j) { X1(rmn, j) = 1.0; X1(r, j) = 1.0; }
  } else {
    for (j = 2; j <= nj; ++j) X1(r, j) = X1(r, j - 1) + X1(ydif, j);
    X1(rmn, 2) = X1(r, 1);
    for (j = 3; j <= nj; ++j) X1(rmn, j) = X1(rmn, j - 1) + X1(sns, j - 1);
  }
  for (j = 1; j <= nj; ++j) { X1(sx, j) = 1.0; X1(sxmn, j) = 1.0; }
  if (icrd == 3) {
    for (j = 1; j <= nj; ++j) X1(sx, j) = X1(r, j);
    for (j = 2; j <= nj; ++j) X1(sxmn, j) = X1(rmn, j);
  }
  for (j = 2; j <= njm1; ++j) X1(ycvr, j) = X1(r, j) * X1(sns, j);
  for (j = 2; j <= njm1; ++j) X1(arx, j) = X1(ycvr, j);
  if (icrd == 3) for (j = 1; j <= nj; ++j) X1(arx, j) = X1(sns, j);
  for (j = 4; j <= njm2; ++j) X1(ycvrs, j) = 0.5 * (X1(r, j) + X1(r, j - 1)) * X1(ydif, j);
  X1(ycvrs, 3) = 0.5 * (X1(r, 3) + X1(r, 1)) * X1(snsv, 3);
  X1(ycvrs, njm1) = 0.5 * (X1(r, nj) + X1(r, njm2)) * X1(snsv, njm1);
  if (icrd == 2) {
    for (j = 3; j <= njm2; ++j) X1(arxj, j) = 0.25 * (1.0 + X1(rmn, j) / X1(r, j)) * X1(arx, j);
    for (j = 3; j <= njm2; ++j) X1(arxjp, j) = X1(arx, j) - X1(arxj, j);
  } else {
    for (j = 3; j <= njm2; ++j) X1(arxj, j) = 0.5 * X1(arx, j);
    for (j = 3; j <= njm2; ++j) X1(arxjp, j) = X1(arxj, j);
  }
  X1(arxjp, 2) = X1(arx, 2);
  X1(arxj, njm1) = X1(arx, njm1);
  for (j = 3; j <= njm2; ++j) X1(fv, j) = X1(arxjp, j) / X1(arx, j);
  for (j = 3; j <= njm2; ++j) X1(fvp, j) = 1.0 - X1(fv, j);
  for (i = 2; i <= ni; ++i) X1(fx, i) = 0.5 * X1(sew, i - 1) / X1(xdif, i);
  for (i = 2; i <= ni; ++i) X1(fxm, i) = 1.0 - X1(fx, i);
  for (j = 2; j <= nj; ++j) X1(fy, j) = 0.5 * X1(sns, j - 1) / X1(ydif, j);
  for (j = 2; j <= nj; ++j) X1(fym, j) = 1.0 - X1(fy, j);
}

/* buoyancy_cav.f90:507-513 */
static void clrsor(orc_general *s) {
  size_t n = (size_t)s->ni * (size_t)s->nj;
  memset(s->sp, 0, n * sizeof(double));
  memset(s->su, 0, n * sizeof(double));
}

/* ---- user modules ------------------------------------------------------------------------------------ */
/* cylin_cond.f90:149-169 (nq is not consulted by the shipped routine: only T is solved) */
static void gamsor_cylinder(orc_general *s, int nq) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  double *gam = s->gam, *su = s->su, *sp = s->sp;
  int i, j;
  (void)nq;
  for (j = 1; j <= nj; ++j) for (i = 1; i <= ni; ++i) gam[IX(i, j)] = s->gam0;
  for (j = s->js1; j <= s->js2; ++j)
    for (i = s->is1; i <= s->is2; ++i) {
      gam[IX(i, j)] = s->gam1;
      su[IX(i, j)] = 200.0;
      sp[IX(i, j)] = -0.1;
    }
  if (s->ibnd == 1) return;
  for (i = 2; i <= nim1; ++i) gam[IX(i, nj)] = 0.0;
  for (i = 2; i <= nim1; ++i) {
    su[IX(i, 2)] = su[IX(i, 2)] + X1(s->r, 1) / X1(s->arx, 2) * s->tf / (X1(s->ydif, 2) / gam[IX(i, 2)] + 1.0 / s->htc);
    sp[IX(i, 2)] = sp[IX(i, 2)] - X1(s->r, 1) / X1(s->arx, 2) / (X1(s->ydif, 2) / gam[IX(i, 2)] + 1.0 / s->htc);
    gam[IX(i, 1)] = 0.0;
  }
  for (j = 2; j <= njm1; ++j) su[IX(nim1, j)] = su[IX(nim1, j)] + s->qin / X1(s->sew, nim1);
  for (j = 2; j <= njm1; ++j) gam[IX(ni, j)] = 0.0;
}

/* authored (see header): Boussinesq cavity */
static void gamsor_buoyancy(orc_general *s, int nq) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  double *gam = s->gam, *su = s->su;
  int i, j;
  if (nq == 1 || nq == 2) {
    for (j = 1; j <= nj; ++j) for (i = 1; i <= ni; ++i) gam[IX(i, j)] = s->visc;
    if (nq == 2)
      for (j = 3; j <= njm1; ++j)
        for (i = 2; i <= nim1; ++i)
          su[IX(i, j)] = s->rgbeta * (ARIMEAN(s->t[IX(i, j - 1)], s->t[IX(i, j)], X1(s->fy, j)) - s->tref);
  } else if (nq == 4) {
    for (j = 1; j <= nj; ++j) for (i = 1; i <= ni; ++i) gam[IX(i, j)] = s->cond;
    for (i = 2; i <= nim1; ++i) {   /* adiabatic north and south walls */
      gam[IX(i, nj)] = 0.0;
      gam[IX(i, 1)] = 0.0;
    }
  }
}

static void gamsor(orc_general *s, int nq) {
  if (s->user == 1) gamsor_cylinder(s, nq);
  else if (s->user == 2) gamsor_buoyancy(s, nq);
}

/* cylin_cond.f90:140-147 ; buoyancy (authored): adiabatic rows copy the adjacent interior value */
void orc_gen_boundary(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  double *t = s->t;
  int i, j;
  if (s->user == 1) {
    if (s->ibnd == 2 && !s->llast) return;
    for (i = 2; i <= nim1; ++i) t[IX(i, nj)] = t[IX(i, njm1)];
    for (i = 2; i <= nim1; ++i)
      t[IX(i, 1)] = (s->gam0 * t[IX(i, 2)] + s->htc * X1(s->ydif, 2) * s->tf) / (s->gam0 + s->htc * X1(s->ydif, 2));
    for (j = 2; j <= njm1; ++j) t[IX(ni, j)] = t[IX(nim1, j)] + s->qin * X1(s->xdif, ni) / s->gam0;
  } else if (s->user == 2) {
    if (!s->llast) return;
    for (i = 2; i <= nim1; ++i) { t[IX(i, nj)] = t[IX(i, njm1)]; t[IX(i, 1)] = t[IX(i, 2)]; }
  }
}

/* cylin_cond.f90:117-134 ; buoyancy (authored) */
void orc_gen_initial(orc_general *s) {
  const int ni = s->ni, nj = s->nj;
  int i, j;
  size_t n = (size_t)ni * (size_t)nj, k;
  for (k = 0; k < n; ++k) { s->den[k] = 1.0; s->cp[k] = 1.0; }     /* den_ini, cp_ini (grid(), :118-119) */
  if (s->user == 1) {
    for (j = 1; j <= nj; ++j) for (i = 2; i <= ni; ++i) s->t[IX(i, j)] = s->t0;
    for (j = 1; j <= nj; ++j) s->t[IX(1, j)] = s->t0 * (1.0 + X1(s->y, j));
  } else if (s->user == 2) {
    for (j = 1; j <= nj; ++j) {
      for (i = 2; i <= ni - 1; ++i) s->t[IX(i, j)] = s->tref;
      s->t[IX(1, j)] = 1.0;
      s->t[IX(ni, j)] = 0.0;
    }
  }
}

/* buoyancy_cav.f90:471-505 */
static void tdma_line(int i, int jstart, int ni, int nj, double *phi, const orc_general *s, double *a, double *c) {
  int j, jj, njm1 = nj - 1, jsrm1 = jstart - 1;
  double b, d, term;
  X1(a, jsrm1) = 0.0;
  X1(c, jsrm1) = phi[IX(i, jsrm1)];
  for (j = jstart; j <= njm1; ++j) {
    X1(a, j) = s->an[IX(i, j)];
    b = s->as[IX(i, j)];
    X1(c, j) = s->ae[IX(i, j)] * phi[IX(i + 1, j)] + s->aw[IX(i, j)] * phi[IX(i - 1, j)] + s->su[IX(i, j)];
    d = s->ap[IX(i, j)];
    term = 1.0 / (d - b * X1(a, j - 1));
    X1(a, j) = X1(a, j) * term;
    X1(c, j) = (X1(c, j) + b * X1(c, j - 1)) * term;
  }
  for (jj = jstart; jj <= njm1; ++jj) {
    j = nj + jsrm1 - jj;
    phi[IX(i, j)] = X1(a, j) * phi[IX(i, j + 1)] + X1(c, j);
  }
}
void orc_gen_tdma(const orc_general *s, int istart, int jstart, double *phi) {
  int ni = s->ni, nj = s->nj, nim1 = ni - 1, i, colour;
  double *a = (double *)malloc(sizeof(double) * (size_t)(nj + 1)), *c = (double *)malloc(sizeof(double) * (size_t)(nj + 1));
  if (s->order == 0) {
    for (i = istart; i <= nim1; ++i) tdma_line(i, jstart, ni, nj, phi, s, a, c);
  } else {
    for (colour = 0; colour < 2; ++colour)
      for (i = istart + colour; i <= nim1; i += 2) tdma_line(i, jstart, ni, nj, phi, s, a, c);
  }
  free(a);
  free(c);
}

/* buoyancy_cav.f90:122-198 (assembly only; the sweeps follow in orc_gen_calcu) */
void orc_gen_assemble_u(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1, iadv = s->iadv;
  double *u = s->u, *v = s->v, *p = s->p, *den = s->den, *gam = s->gam, *du = s->du;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp;
  const double *xcvi = s->xcvi, *xcvip = s->xcvip, *r = s->r, *ydif = s->ydif, *arx = s->arx, *sew = s->sew, *sx = s->sx;
  const double *fx = s->fx, *fy = s->fy, *rmn = s->rmn, *ycvr = s->ycvr, *sewu = s->sewu, *xdif = s->xdif;
  int i, j;
  double vis_ne, vis_nw, vol, fl, flm, flp, flow, diff, acof, reseq;
  clrsor(s);
  gamsor(s, 1);
  for (i = 3; i <= njm1; ++i) {
    fl = X1(xcvi, i) * v[IX(i, 2)] * den[IX(i, 1)];
    flm = X1(xcvip, i - 1) * v[IX(i - 1, 2)] * den[IX(i - 1, 1)];
    flow = X1(r, 1) * (fl + flm);
    diff = X1(r, 1) * (X1(xcvi, i) * gam[IX(i, 1)] + X1(xcvip, i - 1) * gam[IX(i - 1, 1)]) / X1(ydif, 2);
    acof = DAPEC(flow, diff, iadv);
    as[IX(i, 2)] = acof + dmax(0.0, flow);
  }
  s->resor[0] = 0.0;
  for (j = 2; j <= njm1; ++j) {
    flow = X1(arx, j) * u[IX(2, j)] * den[IX(1, j)];
    diff = X1(arx, j) * gam[IX(1, j)] / (X1(sew, 2) * X1(sx, j));
    acof = DAPEC(flow, diff, iadv);
    aw[IX(3, j)] = acof + dmax(0.0, flow);
    for (i = 3; i <= njm1; ++i) {
      if (i == njm1) {
        flow = X1(arx, j) * u[IX(ni, j)] * den[IX(ni, j)];
        diff = X1(arx, j) * gam[IX(ni, j)] / (X1(sew, nim1) * X1(sx, j));
      } else {
        fl = u[IX(i, j)] * ARIMEAN(den[IX(i - 1, j)], den[IX(i, j)], X1(fx, i));
        flp = u[IX(i + 1, j)] * ARIMEAN(den[IX(i, j)], den[IX(i + 1, j)], X1(fx, i + 1));
        flow = X1(arx, j) * 0.5 * (fl + flp);
        diff = X1(arx, j) * gam[IX(i, j)] / (X1(sew, i) * X1(sx, j));
      }
      acof = DAPEC(flow, diff, iadv);
      aw[IX(i + 1, j)] = acof + dmax(0.0, flow);
      ae[IX(i, j)] = aw[IX(i + 1, j)] - flow;
      if (j == njm1) {
        fl = X1(xcvi, i) * v[IX(i, nj)] * den[IX(i, nj)];
        flm = X1(xcvip, i - 1) * v[IX(i - 1, nj)] * den[IX(i - 1, nj)];
        diff = X1(r, nj) * (X1(xcvi, i) * gam[IX(i, nj)] + X1(xcvip, i - 1) * gam[IX(i - 1, nj)]) / X1(ydif, nj);
      } else {
        fl = X1(xcvi, i) * v[IX(i, j + 1)] * ARIMEAN(den[IX(i, j)], den[IX(i, j + 1)], X1(fy, j + 1));
        flm = X1(xcvip, i - 1) * v[IX(i - 1, j + 1)] * ARIMEAN(den[IX(i - 1, j)], den[IX(i - 1, j + 1)], X1(fy, j + 1));
        vis_ne = HARMEAN(gam[IX(i, j)], gam[IX(i, j + 1)], X1(fy, j + 1));
        vis_nw = HARMEAN(gam[IX(i - 1, j)], gam[IX(i - 1, j + 1)], X1(fy, j + 1));
        diff = X1(rmn, j + 1) * (vis_ne * X1(xcvi, i) + vis_nw * X1(xcvip, i - 1)) / X1(ydif, j + 1);
      }
      flow = X1(rmn, j + 1) * (fl + flm);
      acof = DAPEC(flow, diff, iadv);
      as[IX(i, j + 1)] = acof + dmax(0.0, flow);
      an[IX(i, j)] = as[IX(i, j + 1)] - flow;
      vol = X1(ycvr, j) * X1(sewu, i);
      su[IX(i, j)] = su[IX(i, j)] * vol;
      sp[IX(i, j)] = sp[IX(i, j)] * vol;
      du[IX(i, j)] = vol / (X1(xdif, i) * X1(sx, j));
      su[IX(i, j)] = su[IX(i, j)] + du[IX(i, j)] * (p[IX(i - 1, j)] - p[IX(i, j)]);
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
      reseq = ap[IX(i, j)] * u[IX(i, j)] - ae[IX(i, j)] * u[IX(i + 1, j)] - aw[IX(i, j)] * u[IX(i - 1, j)]
              - an[IX(i, j)] * u[IX(i, j + 1)] - as[IX(i, j)] * u[IX(i, j - 1)] - su[IX(i, j)];
      s->resor[0] = s->resor[0] + fabs(reseq);
      ap[IX(i, j)] = ap[IX(i, j)] / s->urf[0];
      su[IX(i, j)] = su[IX(i, j)] + (1 - s->urf[0]) * ap[IX(i, j)] * u[IX(i, j)];
      du[IX(i, j)] = du[IX(i, j)] / ap[IX(i, j)];
    }
  }
}
void orc_gen_calcu(orc_general *s) {
  int n;
  orc_gen_assemble_u(s);
  for (n = 1; n <= s->nsweep[0]; ++n) orc_gen_tdma(s, 3, 2, s->u);
}

/* buoyancy_cav.f90:200-277 */
void orc_gen_assemble_v(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1, iadv = s->iadv;
  double *u = s->u, *v = s->v, *p = s->p, *den = s->den, *gam = s->gam, *dv = s->dv;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp;
  const double *r = s->r, *sew = s->sew, *sns = s->sns, *arxj = s->arxj, *arxjp = s->arxjp, *xdif = s->xdif, *sxmn = s->sxmn;
  const double *fx = s->fx, *fy = s->fy, *rmn = s->rmn, *fvp = s->fvp, *ycvrs = s->ycvrs, *ydif = s->ydif;
  int i, j;
  double vis_ne, vis_se, area, vol, fl, flm, flp, flow, diff, acof, reseq;
  clrsor(s);
  gamsor(s, 2);
  for (i = 2; i <= nim1; ++i) {
    area = X1(r, 1) * X1(sew, i);
    flow = area * v[IX(i, 2)] * den[IX(i, 1)];
    diff = area * gam[IX(i, 1)] * X1(sns, 2);
    acof = DAPEC(flow, diff, iadv);
    as[IX(i, 3)] = acof + dmax(0.0, flow);
  }
  s->resor[1] = 0.0;
  for (j = 3; j <= njm1; ++j) {
    fl = X1(arxj, j) * u[IX(2, j)] * den[IX(1, j)];
    flm = X1(arxjp, j - 1) * u[IX(2, j - 1)] * den[IX(1, j - 1)];
    flow = fl + flm;
    diff = (X1(arxj, j) * gam[IX(1, j)] + X1(arxjp, j - 1) * gam[IX(1, j - 1)]) / (X1(xdif, 2) * X1(sxmn, j));
    acof = DAPEC(flow, diff, iadv);
    aw[IX(2, j)] = acof + dmax(0.0, flow);
    for (i = 2; i <= nim1; ++i) {
      if (i == nim1) {
        fl = X1(arxj, j) * u[IX(ni, j)] * den[IX(ni, j)];
        flm = X1(arxjp, j - 1) * u[IX(ni, j - 1)] * den[IX(ni, j - 1)];
        diff = (X1(arxj, j) * gam[IX(ni, j)] + X1(arxjp, j - 1) * gam[IX(ni, j - 1)]) / (X1(xdif, ni) * X1(sxmn, j));
      } else {
        fl = X1(arxj, j) * u[IX(i + 1, j)] * ARIMEAN(den[IX(i, j)], den[IX(i + 1, j)], X1(fx, i + 1));
        flm = X1(arxjp, j - 1) * u[IX(i + 1, j - 1)] * ARIMEAN(den[IX(i, j - 1)], den[IX(i + 1, j - 1)], X1(fx, i + 1));
        vis_ne = HARMEAN(gam[IX(i, j)], gam[IX(i + 1, j)], X1(fx, i + 1));
        vis_se = HARMEAN(gam[IX(i, j - 1)], gam[IX(i + 1, j - 1)], X1(fx, i + 1));
        diff = (vis_ne * X1(arxj, j) + vis_se * X1(arxjp, j - 1)) / (X1(xdif, i + 1) * X1(sxmn, j));
      }
      flow = fl + flm;
      acof = DAPEC(flow, diff, iadv);
      aw[IX(i + 1, j)] = acof + dmax(0.0, flow);
      ae[IX(i, j)] = aw[IX(i + 1, j)] - flow;
      if (j == njm1) {
        area = X1(r, nj) * X1(sew, i);
        flow = area * v[IX(i, nj)] * den[IX(i, nj)];
        diff = area * gam[IX(i, nj)] / X1(sns, njm1);
      } else {
        area = X1(r, j) * X1(sew, i);
        fl = v[IX(i, j)] * X1(rmn, j) * X1(sew, i) * ARIMEAN(den[IX(i, j - 1)], den[IX(i, j)], X1(fy, j));
        flp = v[IX(i, j + 1)] * X1(rmn, j + 1) * X1(sew, i) * ARIMEAN(den[IX(i, j)], den[IX(i, j + 1)], X1(fy, j + 1));
        flow = ARIMEAN(fl, flp, X1(fvp, j));
        diff = area * gam[IX(i, j)] / X1(sns, j);
      }
      acof = DAPEC(flow, diff, iadv);
      as[IX(i, j + 1)] = acof + dmax(0.0, flow);
      an[IX(i, j)] = as[IX(i, j + 1)] - flow;
      vol = X1(ycvrs, j) * X1(sew, i);
      su[IX(i, j)] = su[IX(i, j)] * vol;
      sp[IX(i, j)] = sp[IX(i, j)] * vol;
      dv[IX(i, j)] = vol / X1(ydif, j);
      su[IX(i, j)] = su[IX(i, j)] + dv[IX(i, j)] * (p[IX(i, j - 1)] - p[IX(i, j)]);
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
      reseq = ap[IX(i, j)] * v[IX(i, j)] - ae[IX(i, j)] * v[IX(i + 1, j)] - aw[IX(i, j)] * v[IX(i - 1, j)]
              - an[IX(i, j)] * v[IX(i, j + 1)] - as[IX(i, j)] * v[IX(i, j - 1)] - su[IX(i, j)];
      s->resor[1] = s->resor[1] + fabs(reseq);
      ap[IX(i, j)] = ap[IX(i, j)] / s->urf[1];
      su[IX(i, j)] = su[IX(i, j)] + (1.0 - s->urf[1]) * ap[IX(i, j)] * v[IX(i, j)];
      dv[IX(i, j)] = dv[IX(i, j)] / ap[IX(i, j)];
    }
  }
}
void orc_gen_calcv(orc_general *s) {
  int n;
  orc_gen_assemble_v(s);
  for (n = 1; n <= s->nsweep[1]; ++n) orc_gen_tdma(s, 2, 3, s->v);
}

/* buoyancy_cav.f90:279-343 */
void orc_gen_assemble_p(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  double *u = s->u, *v = s->v, *den = s->den, *du = s->du, *dv = s->dv, *pp = s->pp;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp;
  const double *ycvr = s->ycvr, *sew = s->sew, *rmn = s->rmn, *arx = s->arx, *fx = s->fx, *fy = s->fy;
  int i, j;
  double areaew, arean, areas, vol, dene, denn, ce, cw, cn, cs;
  clrsor(s);
  gamsor(s, 3);
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i) {
      vol = X1(ycvr, j) * X1(sew, i);
      sp[IX(i, j)] = sp[IX(i, j)] * vol;
      su[IX(i, j)] = su[IX(i, j)] * vol;
    }
  for (i = 2; i <= nim1; ++i) {
    areas = X1(rmn, 2) * X1(sew, i);
    cs = den[IX(i, 1)] * v[IX(i, 2)] * areas;
    su[IX(i, 2)] = su[IX(i, 2)] + cs;
    as[IX(i, 2)] = 0.0;
  }
  s->resor[2] = 0.0;
  for (j = 2; j <= njm1; ++j) {
    areaew = X1(arx, j);
    cw = den[IX(1, j)] * u[IX(2, j)] * areaew;
    su[IX(2, j)] = su[IX(2, j)] + cw;
    aw[IX(2, j)] = 0.0;
    for (i = 2; i <= nim1; ++i) {
      if (i == nim1) {
        ce = den[IX(ni, j)] * u[IX(ni, j)] * areaew;
        su[IX(i, j)] = su[IX(i, j)] - ce;
        ae[IX(i, j)] = 0.0;
      } else {
        dene = ARIMEAN(den[IX(i, j)], den[IX(i + 1, j)], X1(fx, i + 1));
        ce = den[IX(i + 1, j)] * u[IX(i + 1, j)] * areaew;
        su[IX(i, j)] = su[IX(i, j)] - ce;
        ae[IX(i, j)] = dene * areaew * du[IX(i + 1, j)];
        su[IX(i + 1, j)] = su[IX(i + 1, j)] + ce;
        aw[IX(i + 1, j)] = ae[IX(i, j)];
      }
      arean = X1(rmn, j + 1) * X1(sew, i);
      if (j == njm1) {
        cn = den[IX(i, nj)] * v[IX(i, nj)] * arean;
        su[IX(i, j)] = su[IX(i, j)] - cn;
        an[IX(i, j)] = 0.0;
      } else {
        denn = ARIMEAN(den[IX(i, j)], den[IX(i, j + 1)], X1(fy, j + 1));
        cn = denn * v[IX(i, j + 1)] * arean;
        su[IX(i, j)] = su[IX(i, j)] - cn;
        an[IX(i, j)] = denn * arean * dv[IX(i, j + 1)];
        su[IX(i, j + 1)] = su[IX(i, j + 1)] + cn;
        as[IX(i, j + 1)] = an[IX(i, j)];
      }
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
      s->resor[2] = s->resor[2] + fabs(su[IX(i, j)]);
      pp[IX(i, j)] = 0.0;
    }
  }
}
/* buoyancy_cav.f90:347-355 */
void orc_gen_correct(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  double *u = s->u, *v = s->v, *p = s->p, *pp = s->pp, *du = s->du, *dv = s->dv;
  int i, j;
  double ppref = pp[IX(s->ipref, s->jpref)];
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i) {
      p[IX(i, j)] = p[IX(i, j)] + s->urf[2] * (pp[IX(i, j)] - ppref);
      if (i != 2) u[IX(i, j)] = u[IX(i, j)] + du[IX(i, j)] * (pp[IX(i - 1, j)] - pp[IX(i, j)]);
      if (j != 2) v[IX(i, j)] = v[IX(i, j)] + dv[IX(i, j)] * (pp[IX(i, j - 1)] - pp[IX(i, j)]);
    }
}
void orc_gen_calcp(orc_general *s) {
  int n;
  orc_gen_assemble_p(s);
  for (n = 1; n <= s->nsweep[2]; ++n) orc_gen_tdma(s, 2, 2, s->pp);
  orc_gen_correct(s);
}

/* buoyancy_cav.f90:358-428 */
void orc_gen_assemble_t(orc_general *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1, iadv = s->iadv;
  double *u = s->u, *v = s->v, *t = s->t, *den = s->den, *gam = s->gam, *cp = s->cp;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp;
  const double *r = s->r, *sew = s->sew, *ydif = s->ydif, *arx = s->arx, *xdif = s->xdif, *sx = s->sx;
  const double *fx = s->fx, *fy = s->fy, *rmn = s->rmn, *ycvr = s->ycvr;
  int i, j;
  size_t k, n = (size_t)ni * (size_t)nj;
  double areaew, arean, areas, vol, dene, denw, denn, game, gamn, flow, diff, acof, reseq;
  clrsor(s);
  gamsor(s, 4);
  if (s->lsolve[3]) for (k = 0; k < n; ++k) den[k] = den[k] * cp[k];
  for (i = 2; i <= nim1; ++i) {
    areas = X1(r, 1) * X1(sew, i);
    flow = den[IX(i, 1)] * v[IX(i, 2)] * areas;
    diff = gam[IX(i, 1)] * areas / X1(ydif, 2);
    acof = DAPEC(flow, diff, iadv);
    as[IX(i, 2)] = acof + dmax(0.0, flow);
  }
  s->resor[3] = 0.0;
  for (j = 2; j <= njm1; ++j) {
    areaew = X1(arx, j);
    denw = den[IX(1, j)];
    flow = denw * u[IX(2, j)] * areaew;
    diff = gam[IX(1, j)] * areaew / (X1(xdif, 2) * X1(sx, j));
    acof = DAPEC(flow, diff, iadv);
    aw[IX(2, j)] = acof + dmax(0.0, flow);
    for (i = 2; i <= nim1; ++i) {
      if (i == nim1) {
        dene = den[IX(nj, j)];                      /* quirk: nj for ni (buoyancy_cav.f90:390) */
        flow = dene * u[IX(ni, j)] * areaew;
        diff = gam[IX(ni, j)] * areaew / (X1(xdif, ni) * X1(sx, j));
      } else {
        dene = ARIMEAN(den[IX(i, j)], den[IX(i + 1, j)], X1(fx, i + 1));
        flow = dene * u[IX(i + 1, j)] * areaew;
        game = HARMEAN(gam[IX(i, j)], gam[IX(i + 1, j)], X1(fx, i + 1));
        diff = game * areaew / (X1(xdif, i + 1) * X1(sx, j));
      }
      acof = DAPEC(flow, diff, iadv);
      aw[IX(i + 1, j)] = acof + dmax(0.0, flow);
      ae[IX(i, j)] = aw[IX(i + 1, j)] - flow;
      arean = X1(rmn, j + 1) * X1(sew, i);
      if (j == njm1) {
        denn = den[IX(i, nj)];
        flow = denn * v[IX(i, nj)] * arean;
        diff = gam[IX(i, nj)] * arean / X1(ydif, nj);
      } else {
        denn = ARIMEAN(den[IX(i, j)], den[IX(i, j + 1)], X1(fy, j + 1));
        flow = denn * v[IX(i, j + 1)] * arean;
        gamn = HARMEAN(gam[IX(i, j)], gam[IX(i, j + 1)], X1(fy, j + 1));
        diff = gamn * arean / X1(ydif, j + 1);
      }
      acof = DAPEC(flow, diff, iadv);
      as[IX(i, j + 1)] = acof + dmax(0.0, flow);
      an[IX(i, j)] = as[IX(i, j + 1)] - flow;
      vol = X1(ycvr, j) * X1(sew, i);
      sp[IX(i, j)] = sp[IX(i, j)] * vol;
      su[IX(i, j)] = su[IX(i, j)] * vol;
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
      reseq = ap[IX(i, j)] * t[IX(i, j)] - ae[IX(i, j)] * t[IX(i + 1, j)] - aw[IX(i, j)] * t[IX(i - 1, j)]
              - an[IX(i, j)] * t[IX(i, j + 1)] - as[IX(i, j)] * t[IX(i, j - 1)] - su[IX(i, j)];
      s->resor[3] = s->resor[3] + fabs(reseq);
      ap[IX(i, j)] = ap[IX(i, j)] / s->urf[3];
      su[IX(i, j)] = su[IX(i, j)] + (1.0 - s->urf[3]) * ap[IX(i, j)] * t[IX(i, j)];
    }
  }
}
void orc_gen_calct(orc_general *s) {
  int n;
  size_t k, nn = (size_t)s->ni * (size_t)s->nj;
  orc_gen_assemble_t(s);
  for (n = 1; n <= s->nsweep[3]; ++n) orc_gen_tdma(s, 2, 2, s->t);
  if (s->lsolve[3]) for (k = 0; k < nn; ++k) s->den[k] = s->den[k] / s->cp[k];
}

/* program main (cylin_cond.f90:289-312 / buoyancy_cav.f90:2-26) incl. the stop rule of output() (:171-184).
   hist (optional) receives 5 doubles per iteration: resor(1..4), t(imon,jmon).  Returns iterations done
   (calc calls).  out_before_exit: 1 = buoyancy ordering (output() before the llast test). */
int orc_gen_run(orc_general *s, int max_iter, double tol, int imon, int jmon, double *hist) {
  const int ni = s->ni;
  int iter, done = 0;
  s->llast = 0;
  for (iter = 1; iter <= max_iter; ++iter) {
    double m;
    int k;
    orc_gen_boundary(s);
    if (s->lsolve[0]) {
      orc_gen_calcu(s);
      orc_gen_calcv(s);
      orc_gen_calcp(s);
    }
    if (s->lsolve[3]) orc_gen_calct(s);
    done = iter;
    if (hist) {
      double *h = hist + 5 * (size_t)(iter - 1);
      for (k = 0; k < 4; ++k) h[k] = s->resor[k];
      h[4] = s->t[IX(imon, jmon)];
    }
    if (s->llast) break;
    m = s->resor[0];
    for (k = 1; k < 4; ++k) m = dmax(m, s->resor[k]);
    if (m < tol || iter == max_iter - 1) s->llast = 1;
  }
  return done;
}
