/*
 * oracle/cavity_oracle.c -- CPU restatement of the reference's staggered-grid SIMPLE cavity solver.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (cfd_b200/, include/) may link, import or call
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Follows /root/reference/lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90 statement by statement:
 *   orc_setgrid   <- setgrid      cavity_fvm.f90:366-414
 *   orc_lisolv    <- lisolv       cavity_fvm.f90:49-82      (order=0: the reference W->E line Gauss-Seidel;
 *                                                            order=1: red/black line colouring, the variant
 *                                                            BASELINE.json:north_star names -- same fixed point)
 *   orc_calcu     <- calcu        cavity_fvm.f90:92-192
 *   orc_calcv     <- calcv        cavity_fvm.f90:194-291
 *   orc_calcp     <- calcp        cavity_fvm.f90:293-355
 *   orc_cavity_init / orc_cavity_run <- program cavity_fvm  cavity_fvm.f90:509-566 (literals + outer loop)
 *   orc_centred_vel <- centred_vel cavity_fvm.f90:416-430 ; orc_corner_p <- 568-573
 *
 * Arrays are Fortran column-major, 1-based: A(i,j) lives at a[(i-1) + (j-1)*ni].
 * Build with -ffp-contract=off for the bit reference (no FMA contraction; gfortran -O3 for generic x86-64
 * does not contract either), and -O3 -march=native for the timed CPU baseline.
 *
 * PARITY PIN: the reference ships no tests, fixtures or golden outputs and no Fortran compiler exists in the
 * build container, so this restatement is pinned only against the known-answer table in SURVEY.md 8(c)
 * (values from an independent transliteration) -- see tests/test_oracle_cavity.py.  By the letter of the
 * contract that is "parity unpinned by the reference's own vectors"; DESIGN.md says so too.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct orc_cavity {
  int ni, nj;
  int iadv;                 /* 1 CDS, 2 UDS, 3 hybrid            cavity_fvm.f90:30 */
  int nswpu, nswpv, nswpp;
  int ipref, jpref;
  int order;                /* 0 sequential lines (reference), 1 red/black lines */
  double urfu, urfv, urfp;
  double resoru, resorv, resorm;
  double *x, *y, *xu, *yv, *sew, *sns, *sewu, *snsv;      /* 1-D, length ni or nj */
  double *u, *v, *p, *pp, *vis, *den;                    /* (ni,nj) */
  double *ae, *aw, *an, *as, *ap, *su, *sp, *du, *dv;    /* (ni,nj) */
} orc_cavity;

#define IX(i, j) ((size_t)((i) - 1) + (size_t)((j) - 1) * (size_t)ni)
#define X1(a, i) ((a)[(i) - 1])

static inline double dmax(double a, double b) { return a > b ? a : b; }

/* cavity_fvm.f90:366-414 */
void orc_setgrid(int nicv, int njcv, double xstart, double xend, double ystart, double yend,
                 double *x, double *y, double *xu, double *yv,
                 double *sew, double *sns, double *sewu, double *snsv) {
  int ni = nicv + 2, nj = njcv + 2, i, j;
  double dx, dy;
  X1(xu, 1) = xstart;
  X1(xu, 2) = xstart;
  dx = (xend - xstart) / (nicv);
  for (i = 3; i <= ni; ++i) X1(xu, i) = X1(xu, i - 1) + dx;
  X1(x, 1) = X1(xu, 2);
  for (i = 2; i <= ni - 1; ++i) {
    X1(x, i) = 0.5 * (X1(xu, i) + X1(xu, i + 1));
    X1(sew, i) = X1(xu, i + 1) - X1(xu, i);
  }
  X1(x, ni) = X1(xu, ni);
  for (i = 3; i <= ni - 1; ++i) X1(sewu, i) = X1(x, i) - X1(x, i - 1);
  X1(yv, 1) = ystart;
  X1(yv, 2) = ystart;
  dy = (yend - ystart) / njcv;
  for (j = 3; j <= nj; ++j) X1(yv, j) = X1(yv, j - 1) + dy;
  X1(y, 1) = X1(yv, 2);
  for (j = 2; j <= nj - 1; ++j) {
    X1(y, j) = 0.5 * (X1(yv, j) + X1(yv, j + 1));
    X1(sns, j) = X1(yv, j + 1) - X1(yv, j);
  }
  X1(y, nj) = X1(yv, nj);
  for (j = 3; j <= nj - 1; ++j) X1(snsv, j) = X1(y, j) - X1(y, j - 1);
}

/* One line of cavity_fvm.f90:63-80 (scratch a,c indexed 1..nj like the Fortran automatic arrays). */
static void lisolv_line(int i, int jstart, int ni, int nj, double *phi,
                        const double *ap, const double *an, const double *as,
                        const double *ae, const double *aw, const double *su,
                        double *a, double *c) {
  int j, jj, njm1 = nj - 1, jsrm1 = jstart - 1;
  double b, d, term;
  X1(a, jsrm1) = 0.0;
  X1(c, jsrm1) = phi[IX(i, jsrm1)];
  for (j = jstart; j <= njm1; ++j) {
    X1(a, j) = an[IX(i, j)];
    b = as[IX(i, j)];
    X1(c, j) = ae[IX(i, j)] * phi[IX(i + 1, j)] + aw[IX(i, j)] * phi[IX(i - 1, j)] + su[IX(i, j)];
    d = ap[IX(i, j)];
    term = 1.0 / (d - b * X1(a, j - 1));
    X1(a, j) = X1(a, j) * term;
    X1(c, j) = (X1(c, j) + b * X1(c, j - 1)) * term;
  }
  for (jj = jstart; jj <= njm1; ++jj) {
    j = nj + jsrm1 - jj;
    phi[IX(i, j)] = X1(a, j) * phi[IX(i, j + 1)] + X1(c, j);
  }
}

/* cavity_fvm.f90:49-82.  order==1: two colour passes, colour 0 = lines with (i-istart) even. */
void orc_lisolv(const orc_cavity *s, int istart, int jstart, double *phi) {
  int ni = s->ni, nj = s->nj, nim1 = ni - 1, i, colour;
  double *a = (double *)malloc(sizeof(double) * (size_t)(nj + 1));
  double *c = (double *)malloc(sizeof(double) * (size_t)(nj + 1));
  if (s->order == 0) {
    for (i = istart; i <= nim1; ++i)
      lisolv_line(i, jstart, ni, nj, phi, s->ap, s->an, s->as, s->ae, s->aw, s->su, a, c);
  } else {
    for (colour = 0; colour < 2; ++colour)
      for (i = istart + colour; i <= nim1; i += 2)
        lisolv_line(i, jstart, ni, nj, phi, s->ap, s->an, s->as, s->ae, s->aw, s->su, a, c);
  }
  free(a);
  free(c);
}

/* cavity_fvm.f90:104-187 (coefficient assembly, residual, under-relaxation, du) -- no solve. */
void orc_assemble_u(orc_cavity *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1, iadv = s->iadv;
  const double *den = s->den, *vis = s->vis, *u = s->u, *v = s->v, *p = s->p;
  const double *xu = s->xu, *y = s->y, *sew = s->sew, *sns = s->sns, *sewu = s->sewu;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp, *du = s->du;
  const double urfu = s->urfu;
  int i, j;
  double arean = 0.0, areas = 0.0, areaw = 0.0, areae = 0.0, vol;
  double gee, gww, gne, gnw, gse, gsw, gp, ge, gw, gn, gs;
  double ce, cw, cn, cs, cp, smp;
  double vise, visw, visn, viss, de, dw, dn, ds;
  double dudxe, dudxw, dvdxn, dvdxs, resor;
  for (j = 2; j <= njm1; ++j) {
    for (i = 3; i <= nim1; ++i) {
      arean = X1(sewu, i);
      areas = X1(sewu, i);
      areaw = X1(sns, j);
      areae = X1(sns, j);
      vol = X1(sewu, i) * X1(sns, j);
      gee = 0.5 * (den[IX(i, j)] + den[IX(i + 1, j)]) * u[IX(i + 1, j)];
      gww = 0.5 * (den[IX(i - 2, j)] + den[IX(i - 1, j)]) * u[IX(i - 1, j)];
      gne = 0.5 * (den[IX(i, j)] + den[IX(i, j + 1)]) * v[IX(i, j + 1)];
      gnw = 0.5 * (den[IX(i - 1, j)] + den[IX(i - 1, j + 1)]) * v[IX(i - 1, j + 1)];
      gse = 0.5 * (den[IX(i, j - 1)] + den[IX(i, j)]) * v[IX(i, j)];
      gsw = 0.5 * (den[IX(i - 1, j - 1)] + den[IX(i - 1, j)]) * v[IX(i - 1, j)];
      gp = 0.5 * (den[IX(i - 1, j)] + den[IX(i, j)]) * u[IX(i, j)];
      ge = 0.5 * (gp + gee);
      gw = 0.5 * (gww + gp);
      gn = 0.5 * (gnw + gne);
      gs = 0.5 * (gsw + gse);
      ce = ge * areae;
      cw = gw * areaw;
      cn = gn * arean;
      cs = gs * areas;
      smp = ce - cw + cn - cs;
      cp = dmax(smp, 0.0);
      vise = vis[IX(i, j)];
      visw = vis[IX(i - 1, j)];
      visn = 0.25 * (vis[IX(i - 1, j)] + vis[IX(i - 1, j + 1)] + vis[IX(i, j + 1)] + vis[IX(i, j)]);
      viss = 0.25 * (vis[IX(i - 1, j - 1)] + vis[IX(i - 1, j)] + vis[IX(i, j)] + vis[IX(i, j - 1)]);
      de = vise * areae / (X1(xu, i + 1) - X1(xu, i));
      dw = visw * areaw / (X1(xu, i) - X1(xu, i - 1));
      dn = visn * arean / (X1(y, j + 1) - X1(y, j));
      ds = viss * areas / (X1(y, j) - X1(y, j - 1));
      if (iadv == 1) {
        ae[IX(i, j)] = de - 0.5 * ce;
        aw[IX(i, j)] = dw + 0.5 * cw;
        an[IX(i, j)] = dn - 0.5 * cn;
        as[IX(i, j)] = ds + 0.5 * cs;
      } else if (iadv == 2) {
        ae[IX(i, j)] = de + dmax(0.0, -ce);
        aw[IX(i, j)] = dw + dmax(0.0, cw);
        an[IX(i, j)] = dn + dmax(0.0, -cn);
        as[IX(i, j)] = ds + dmax(0.0, cs);
      } else {
        an[IX(i, j)] = dmax(fabs(0.5 * cn), dn) - 0.5 * cn;
        as[IX(i, j)] = dmax(fabs(0.5 * cs), ds) + 0.5 * cs;
        ae[IX(i, j)] = dmax(fabs(0.5 * ce), de) - 0.5 * ce;
        aw[IX(i, j)] = dmax(fabs(0.5 * cw), dw) + 0.5 * cw;
      }
      su[IX(i, j)] = (p[IX(i - 1, j)] - p[IX(i, j)]) * 0.5 * (areae + areaw);
      dudxe = (u[IX(i + 1, j)] - u[IX(i, j)]) / X1(sew, i);
      dudxw = (u[IX(i, j)] - u[IX(i - 1, j)]) / X1(sew, i - 1);
      dvdxn = (v[IX(i, j + 1)] - v[IX(i - 1, j + 1)]) / X1(sewu, i);
      dvdxs = (v[IX(i, j)] - v[IX(i - 1, j)]) / X1(sewu, i);
      su[IX(i, j)] = su[IX(i, j)] + ((vise * dudxe - visw * dudxw) / X1(sewu, i)
                                     + (visn * dvdxn - viss * dvdxs) / X1(sns, j)) * vol;
      su[IX(i, j)] = su[IX(i, j)] + cp * u[IX(i, j)];
      sp[IX(i, j)] = -cp;
    }
  }
  s->resoru = 0.0;
  for (j = 2; j <= njm1; ++j) {
    for (i = 3; i <= nim1; ++i) {
      /* quirk: areae/areaw are NOT reassigned here (cavity_fvm.f90:186) -- they keep the value left by
         the last pass of the loop above, i.e. sns(njm1), for every cell. */
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
      resor = ap[IX(i, j)] * u[IX(i, j)] - ae[IX(i, j)] * u[IX(i + 1, j)] - aw[IX(i, j)] * u[IX(i - 1, j)]
              - an[IX(i, j)] * u[IX(i, j + 1)] - as[IX(i, j)] * u[IX(i, j - 1)] - su[IX(i, j)];
      s->resoru = s->resoru + fabs(resor);
      ap[IX(i, j)] = ap[IX(i, j)] / urfu;
      su[IX(i, j)] = su[IX(i, j)] + (1 - urfu) * ap[IX(i, j)] * u[IX(i, j)];
      du[IX(i, j)] = 0.5 * (areae + areaw) / ap[IX(i, j)];
    }
  }
}

void orc_calcu(orc_cavity *s) {
  int n;
  orc_assemble_u(s);
  for (n = 1; n <= s->nswpu; ++n) orc_lisolv(s, 3, 2, s->u);   /* cavity_fvm.f90:189-191 */
}

/* cavity_fvm.f90:205-285 */
void orc_assemble_v(orc_cavity *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1, iadv = s->iadv;
  const double *den = s->den, *vis = s->vis, *u = s->u, *v = s->v, *p = s->p;
  const double *x = s->x, *yv = s->yv, *sew = s->sew, *sns = s->sns, *snsv = s->snsv;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp, *dv = s->dv;
  const double urfv = s->urfv;
  int i, j;
  double arean = 0.0, areas = 0.0, areaw = 0.0, areae = 0.0, vol;
  double gne, gnw, gnn, gse, gsw, gss, gp, ge, gw, gn, gs;
  double ce, cw, cn, cs, cp, smp;
  double vise, visw, visn, viss, de, dw, dn, ds;
  double dudye, dudyw, dvdyn, dvdys, resor;
  for (j = 3; j <= njm1; ++j) {
    for (i = 2; i <= nim1; ++i) {
      areae = X1(snsv, j);
      areaw = X1(snsv, j);
      arean = X1(sew, i);
      areas = X1(sew, i);
      vol = X1(sew, i) * X1(snsv, j);
      gne = 0.5 * (den[IX(i, j)] + den[IX(i + 1, j)]) * u[IX(i + 1, j)];
      gnw = 0.5 * (den[IX(i - 1, j)] + den[IX(i, j)]) * u[IX(i, j)];
      gnn = 0.5 * (den[IX(i, j)] + den[IX(i, j + 1)]) * v[IX(i, j + 1)];
      gse = 0.5 * (den[IX(i, j - 1)] + den[IX(i + 1, j - 1)]) * u[IX(i + 1, j - 1)];
      gsw = 0.5 * (den[IX(i - 1, j - 1)] + den[IX(i, j - 1)]) * u[IX(i, j - 1)];
      gss = 0.5 * (den[IX(i, j - 2)] + den[IX(i, j - 1)]) * v[IX(i, j - 1)];
      gp = 0.5 * (den[IX(i, j - 1)] + den[IX(i, j)]) * v[IX(i, j)];
      ge = 0.5 * (gne + gse);
      gw = 0.5 * (gnw + gsw);
      gn = 0.5 * (gp + gnn);
      gs = 0.5 * (gss + gp);
      ce = ge * areae;
      cw = gw * areaw;
      cn = gn * arean;
      cs = gs * areas;
      smp = ce - cw + cn - cs;
      cp = dmax(smp, 0.0);
      vise = 0.25 * (vis[IX(i, j - 1)] + vis[IX(i, j)] + vis[IX(i + 1, j)] + vis[IX(i + 1, j - 1)]);
      visw = 0.25 * (vis[IX(i - 1, j - 1)] + vis[IX(i - 1, j)] + vis[IX(i, j)] + vis[IX(i, j - 1)]);
      visn = vis[IX(i, j)];
      viss = vis[IX(i, j - 1)];
      de = vise * areae / (X1(x, i + 1) - X1(x, i));
      dw = visw * areaw / (X1(x, i) - X1(x, i - 1));
      dn = visn * arean / (X1(yv, j + 1) - X1(yv, j));
      ds = viss * areas / (X1(yv, j) - X1(yv, j - 1));
      if (iadv == 1) {
        ae[IX(i, j)] = de - 0.5 * ce;
        aw[IX(i, j)] = dw + 0.5 * cw;
        an[IX(i, j)] = dn - 0.5 * cn;
        as[IX(i, j)] = ds + 0.5 * cs;
      } else if (iadv == 2) {
        ae[IX(i, j)] = de + dmax(0.0, -ce);
        aw[IX(i, j)] = dw + dmax(0.0, cw);
        an[IX(i, j)] = dn + dmax(0.0, -cn);
        as[IX(i, j)] = ds + dmax(0.0, cs);
      } else {
        an[IX(i, j)] = dmax(fabs(0.5 * cn), dn) - 0.5 * cn;
        as[IX(i, j)] = dmax(fabs(0.5 * cs), ds) + 0.5 * cs;
        ae[IX(i, j)] = dmax(fabs(0.5 * ce), de) - 0.5 * ce;
        aw[IX(i, j)] = dmax(fabs(0.5 * cw), dw) + 0.5 * cw;
      }
      su[IX(i, j)] = (p[IX(i, j - 1)] - p[IX(i, j)]) * 0.5 * (arean + areas);
      dudye = (u[IX(i + 1, j)] - u[IX(i + 1, j - 1)]) / X1(snsv, j);
      dudyw = (u[IX(i, j)] - u[IX(i, j - 1)]) / X1(snsv, j);
      dvdyn = (v[IX(i, j + 1)] - v[IX(i, j)]) / X1(sns, j);
      dvdys = (v[IX(i, j)] - v[IX(i, j - 1)]) / X1(sns, j - 1);
      su[IX(i, j)] = su[IX(i, j)] + (vise * dudye - visw * dudyw) / X1(sew, i) * vol
                     + (visn * dvdyn - viss * dvdys) / X1(snsv, j) * vol;
      sp[IX(i, j)] = -cp;
      su[IX(i, j)] = su[IX(i, j)] + cp * v[IX(i, j)];
    }
  }
  s->resorv = 0.0;
  for (j = 3; j <= njm1; ++j) {
    for (i = 2; i <= nim1; ++i) {
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
      resor = ap[IX(i, j)] * v[IX(i, j)] - ae[IX(i, j)] * v[IX(i + 1, j)] - aw[IX(i, j)] * v[IX(i - 1, j)]
              - an[IX(i, j)] * v[IX(i, j + 1)] - as[IX(i, j)] * v[IX(i, j - 1)] - su[IX(i, j)];
      s->resorv = s->resorv + fabs(resor);
      ap[IX(i, j)] = ap[IX(i, j)] / urfv;
      su[IX(i, j)] = su[IX(i, j)] + (1.0 - urfv) * ap[IX(i, j)] * v[IX(i, j)];
      /* quirk: arean/areas are stale from the loop above = sew(nim1) for every cell (cavity_fvm.f90:283) */
      dv[IX(i, j)] = 0.5 * (arean + areas) / ap[IX(i, j)];
    }
  }
}

void orc_calcv(orc_cavity *s) {
  int n;
  orc_assemble_v(s);
  for (n = 1; n <= s->nswpv; ++n) orc_lisolv(s, 2, 3, s->v);   /* cavity_fvm.f90:287-289 */
}

/* cavity_fvm.f90:301-342 : p' coefficients, mass residual, ap, pp = 0 */
void orc_assemble_p(orc_cavity *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  const double *den = s->den, *u = s->u, *v = s->v, *du = s->du, *dv = s->dv;
  const double *sew = s->sew, *sns = s->sns;
  double *ae = s->ae, *aw = s->aw, *an = s->an, *as = s->as, *ap = s->ap, *su = s->su, *sp = s->sp;
  int i, j;
  double arean, areas, areaw, areae;
  double ge, gw, gn, gs, dene, denw, denn, dens;
  double ce, cw, cn, cs, smp;
  s->resorm = 0.0;
  for (j = 2; j <= njm1; ++j) {
    for (i = 2; i <= nim1; ++i) {
      areae = X1(sns, j);
      areaw = X1(sns, j);
      arean = X1(sew, i);
      areas = X1(sew, i);
      dene = 0.5 * (den[IX(i, j)] + den[IX(i + 1, j)]);
      denw = 0.5 * (den[IX(i - 1, j)] + den[IX(i, j)]);
      denn = 0.5 * (den[IX(i, j)] + den[IX(i, j + 1)]);
      dens = 0.5 * (den[IX(i, j - 1)] + den[IX(i, j)]);
      ae[IX(i, j)] = dene * areae * du[IX(i + 1, j)];
      aw[IX(i, j)] = denw * areaw * du[IX(i, j)];
      an[IX(i, j)] = denn * arean * dv[IX(i, j + 1)];
      as[IX(i, j)] = dens * areas * dv[IX(i, j)];
      ge = dene * u[IX(i + 1, j)];
      gw = denw * u[IX(i, j)];
      gn = denn * v[IX(i, j + 1)];
      gs = dens * v[IX(i, j)];
      ce = ge * areae;
      cw = gw * areaw;
      cn = gn * arean;
      cs = gs * areas;
      smp = ce - cw + cn - cs;
      su[IX(i, j)] = -smp;
      sp[IX(i, j)] = 0.0;
      s->resorm = s->resorm + fabs(smp);
    }
  }
  for (j = 2; j <= njm1; ++j)
    for (i = 2; i <= nim1; ++i)
      ap[IX(i, j)] = ae[IX(i, j)] + aw[IX(i, j)] + an[IX(i, j)] + as[IX(i, j)] - sp[IX(i, j)];
  memset(s->pp, 0, sizeof(double) * (size_t)ni * (size_t)nj);
}

/* cavity_fvm.f90:346-354 */
void orc_correct(orc_cavity *s) {
  const int ni = s->ni, nj = s->nj, nim1 = ni - 1, njm1 = nj - 1;
  double *u = s->u, *v = s->v, *p = s->p;
  const double *pp = s->pp, *du = s->du, *dv = s->dv;
  const double urfp = s->urfp;
  int i, j;
  double ppref = pp[IX(s->ipref, s->jpref)];
  for (j = 2; j <= njm1; ++j) {
    for (i = 2; i <= nim1; ++i) {
      p[IX(i, j)] = p[IX(i, j)] + urfp * (pp[IX(i, j)] - ppref);
      if (i != 2) u[IX(i, j)] = u[IX(i, j)] + du[IX(i, j)] * (pp[IX(i - 1, j)] - pp[IX(i, j)]);
      if (j != 2) v[IX(i, j)] = v[IX(i, j)] + dv[IX(i, j)] * (pp[IX(i, j - 1)] - pp[IX(i, j)]);
    }
  }
}

void orc_calcp(orc_cavity *s) {
  int n;
  orc_assemble_p(s);
  for (n = 1; n <= s->nswpp; ++n) orc_lisolv(s, 2, 2, s->pp);  /* cavity_fvm.f90:343-345 */
  orc_correct(s);
}

/* Outer loop of program cavity_fvm (cavity_fvm.f90:553-566).  Runs at most maxit iterations, stops when
   max(resoru,resorv,resorm) < tol (reference literal: 1e-10).  Returns the number of iterations done.
   If hist != NULL it receives 5 doubles per iteration: resoru,resorv,resorm,u(imon,jmon),v(imon,jmon). */
int orc_cavity_run(orc_cavity *s, int maxit, double tol, int imon, int jmon, double *hist) {
  const int ni = s->ni;
  int iter;
  for (iter = 1; iter <= maxit; ++iter) {
    double m;
    orc_calcu(s);
    orc_calcv(s);
    orc_calcp(s);
    if (hist) {
      double *h = hist + 5 * (size_t)(iter - 1);
      h[0] = s->resoru; h[1] = s->resorv; h[2] = s->resorm;
      h[3] = s->u[IX(imon, jmon)]; h[4] = s->v[IX(imon, jmon)];
    }
    m = dmax(dmax(s->resoru, s->resorv), s->resorm);
    if (m < tol) return iter;
  }
  return maxit;
}

/* cavity_fvm.f90:416-430 */
void orc_centred_vel(int ni, int nj, const double *u, const double *v, double *uc, double *vc) {
  int i, j, nim1 = ni - 1, njm1 = nj - 1;
  for (j = 1; j <= nj; ++j) {
    uc[IX(1, j)] = u[IX(2, j)];
    uc[IX(ni, j)] = u[IX(ni, j)];
    for (i = 2; i <= nim1; ++i) uc[IX(i, j)] = 0.5 * (u[IX(i, j)] + u[IX(i + 1, j)]);
  }
  for (i = 1; i <= ni; ++i) {
    vc[IX(i, 1)] = v[IX(i, 2)];
    vc[IX(i, nj)] = v[IX(i, nj)];
    for (j = 2; j <= njm1; ++j) vc[IX(i, j)] = 0.5 * (v[IX(i, j)] + v[IX(i, j + 1)]);
  }
}

/* cavity_fvm.f90:568-573 */
void orc_corner_p(int ni, int nj, int ipref, int jpref, double *p) {
  int nim1 = ni - 1, njm1 = nj - 1;
  size_t k, n = (size_t)ni * (size_t)nj;
  double pref = p[IX(ipref, jpref)];
  for (k = 0; k < n; ++k) p[k] = p[k] - pref;
  p[IX(1, 1)] = p[IX(2, 1)] + p[IX(1, 2)] - p[IX(2, 2)];
  p[IX(1, nj)] = p[IX(1, njm1)] + p[IX(2, nj)] - p[IX(2, njm1)];
  p[IX(ni, 1)] = p[IX(nim1, 1)] + p[IX(ni, 2)] - p[IX(nim1, 2)];
  p[IX(ni, nj)] = p[IX(ni, njm1)] + p[IX(nim1, nj)] - p[IX(nim1, njm1)];
}
