/*
 * oracle/heat_oracle.c -- CPU restatement of the reference's finite-difference heat-conduction stencils.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/cavity_oracle.c header for the rule).
 *
 * Reference (paths under /root/reference/finite-difference-method/heat-conduction/):
 *   orc_qdot_step          <- QDOT        transient-conduction-1d/explicit-time/fdm1d_explicit_heat_cond.f90:76-88
 *   orc_heat1d_explicit    <- fd1d_explicit_heat_cond   same file :90-141 (update :130-139)
 *   orc_thomas             <- thomas      transient-conduction-1d/implicit-time/fdm1d_explicit_heat_cond.f90:147-184
 *   orc_heat1d_implicit    <- fd1d_implicit_heat_cond   same file :100-146 (assembly :122-143)
 *   orc_heat2d_ftcs        <- the 1-D update formula (:133-135) applied with the 5-point operator of
 *                             steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:171-188 (constant lambda).
 *                             The reference has NO 2-D transient program; this is the
 *                             extrapolation BASELINE.json config 2 asks for, so its parity is defined by this
 *                             file, not by a reference output.
 *   orc_heat2d_coeffs      <- interior()  steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:162-190
 *   orc_heat2d_adi         <- the implicit 1-D assembly (:122-143) applied line by line in x then y
 *                             (Peaceman-Rachford-free "locally one-dimensional" splitting), solved with orc_thomas.
 *
 * 2-D arrays are Fortran column-major, T(i,j) at t[(i-1)+(j-1)*nx], i along x.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define T2(a, i, j) ((a)[(size_t)((i) - 1) + (size_t)((j) - 1) * (size_t)nx])
#define X1(a, i) ((a)[(i) - 1])

/* fdm1d_explicit_heat_cond.f90:76-88.  The Fortran compares the double x with the default-real literal 0.1,
   i.e. with (double)0.1f; qthr carries that value so the restatement is exact. */
double orc_qdot_step(double x, double qthr, double qval) { return (x < qthr) ? qval : 0.0; }

/* fdm1d_explicit_heat_cond.f90:120-140.  t0 is an explicit array (the reference passes a scalar to an array
   dummy -- undefined behaviour).  Returns the number of steps taken. */
int orc_heat1d_explicit(double k, double rho, double cp, double xa, double xb, double ta, double tb,
                        const double *t0, double endtime, double dt, double qthr, double qval,
                        double *x, double *t, int nx, int maxsteps) {
  double *tcopy = (double *)malloc(sizeof(double) * (size_t)nx);
  double dx, time;
  int i, n = 0;
  memcpy(t, t0, sizeof(double) * (size_t)nx);
  memcpy(tcopy, t0, sizeof(double) * (size_t)nx);
  dx = (xb - xa) / (double)(nx - 1);
  X1(x, 1) = xa;
  for (i = 2; i <= nx; ++i) X1(x, i) = X1(x, 1) + (i - 1) * dx;
  time = 0.0;
  while (time <= endtime && n < maxsteps) {
    X1(t, 1) = ta;
    X1(t, nx) = tb;
    for (i = 2; i <= nx - 1; ++i)
      X1(t, i) = X1(tcopy, i) + (k / rho / cp * (X1(tcopy, i + 1) - 2 * X1(tcopy, i) + X1(tcopy, i - 1)) /
                                 (dx * dx) + orc_qdot_step(X1(x, i), qthr, qval) / rho / cp) * dt;
    memcpy(tcopy, t, sizeof(double) * (size_t)nx);
    time = time + dt;
    ++n;
  }
  free(tcopy);
  return n;
}

/* implicit-time/fdm1d_explicit_heat_cond.f90:147-184 : a x(i-1) + b x(i) + c x(i+1) = d, two divides per row */
void orc_thomas(const double *a, const double *b, const double *c, const double *d, double *x, int n) {
  double *e = (double *)malloc(sizeof(double) * (size_t)n);
  double *f = (double *)malloc(sizeof(double) * (size_t)n);
  double temp;
  int i;
  X1(e, 1) = X1(c, 1) / X1(b, 1);
  X1(f, 1) = X1(d, 1) / X1(b, 1);
  for (i = 2; i <= n; ++i) {
    temp = X1(b, i) - X1(a, i) * X1(e, i - 1);
    X1(e, i) = X1(c, i) / temp;
    X1(f, i) = (X1(d, i) - X1(a, i) * X1(f, i - 1)) / temp;
  }
  X1(x, n) = X1(f, n);
  for (i = n - 1; i >= 1; --i) X1(x, i) = X1(f, i) - X1(e, i) * X1(x, i + 1);
  free(e);
  free(f);
}

/* implicit-time/fdm1d_explicit_heat_cond.f90:100-146 */
int orc_heat1d_implicit(double k, double rho, double cp, double xa, double xb, double ta, double tb,
                        const double *t0, double endtime, double dt, double qthr, double qval,
                        double *x, double *t, int nx, int maxsteps) {
  size_t nb = sizeof(double) * (size_t)nx;
  double *tcopy = (double *)malloc(nb), *tri_a = (double *)calloc((size_t)nx, sizeof(double));
  double *tri_b = (double *)calloc((size_t)nx, sizeof(double)), *tri_c = (double *)calloc((size_t)nx, sizeof(double));
  double *rhs = (double *)calloc((size_t)nx, sizeof(double));
  double dx, fo, time;
  int i, n = 0;
  memcpy(t, t0, nb);
  memcpy(tcopy, t0, nb);
  dx = (xb - xa) / (double)(nx - 1);
  X1(x, 1) = xa;
  for (i = 2; i <= nx; ++i) X1(x, i) = X1(x, 1) + (i - 1) * dx;
  fo = k * dt / rho / cp / (dx * dx);
  time = 0.0;
  while (time <= endtime && n < maxsteps) {
    X1(t, 1) = ta;
    X1(tri_a, 1) = 0.0; X1(tri_b, 1) = 1.0; X1(tri_c, 1) = 0.0;
    X1(rhs, 1) = X1(t, 1);
    for (i = 2; i <= nx - 1; ++i) {
      X1(tri_a, i) = -fo;
      X1(tri_b, i) = 1.0 + 2.0 * fo;
      X1(tri_c, i) = -fo;
      X1(rhs, i) = X1(tcopy, i) + dt * orc_qdot_step(X1(x, i), qthr, qval) / rho / cp;
    }
    X1(t, nx) = tb;
    X1(tri_a, nx) = 0.0; X1(tri_b, nx) = 1.0; X1(tri_c, nx) = 0.0;
    X1(rhs, nx) = X1(t, nx);
    orc_thomas(tri_a, tri_b, tri_c, rhs, t, nx);
    memcpy(tcopy, t, nb);
    time = time + dt;
    ++n;
  }
  free(tcopy); free(tri_a); free(tri_b); free(tri_c); free(rhs);
  return n;
}

/*
 * 2-D FTCS: nsteps steps of
 *   t(i,j) = tc(i,j) + ( k/rho/cp*(tc(i+1,j)-2*tc(i,j)+tc(i-1,j))/dx**2
 *                      + k/rho/cp*(tc(i,j+1)-2*tc(i,j)+tc(i,j-1))/dy**2
 *                      + QDOT(x(i),time)/rho/cp ) * dt
 * with Dirichlet edges rewritten every step (tleft at i=1, tright at i=nx, tbottom at j=1, ttop at j=ny;
 * corners take the bottom/top value, matching the order of boundary() in fdm2d_steady_heat_cond.f90:212-246
 * where the bottom/top rows are written last).  tc holds the state on entry and on exit.
 */
void orc_heat2d_ftcs(double k, double rho, double cp, double xa, double xb, double ya, double yb,
                     double tleft, double tright, double tbottom, double ttop,
                     double dt, double qthr, double qval, double *tc, double *t, int nx, int ny, int nsteps) {
  double dx = (xb - xa) / (double)(nx - 1), dy = (yb - ya) / (double)(ny - 1);
  double *x = (double *)malloc(sizeof(double) * (size_t)nx);
  int i, j, n;
  X1(x, 1) = xa;
  for (i = 2; i <= nx; ++i) X1(x, i) = X1(x, 1) + (i - 1) * dx;
  for (n = 0; n < nsteps; ++n) {
    for (j = 2; j <= ny - 1; ++j) {
      T2(t, 1, j) = tleft;
      T2(t, nx, j) = tright;
    }
    for (i = 1; i <= nx; ++i) {
      T2(t, i, 1) = tbottom;
      T2(t, i, ny) = ttop;
    }
    for (j = 2; j <= ny - 1; ++j)
      for (i = 2; i <= nx - 1; ++i)
        T2(t, i, j) = T2(tc, i, j) +
                      (k / rho / cp * (T2(tc, i + 1, j) - 2 * T2(tc, i, j) + T2(tc, i - 1, j)) / (dx * dx) +
                       k / rho / cp * (T2(tc, i, j + 1) - 2 * T2(tc, i, j) + T2(tc, i, j - 1)) / (dy * dy) +
                       orc_qdot_step(X1(x, i), qthr, qval) / rho / cp) * dt;
    memcpy(tc, t, sizeof(double) * (size_t)nx * (size_t)ny);
  }
  free(x);
}

/* fdm2d_steady_heat_cond.f90:162-190 with a caller-supplied conductivity sampled at the four face mid-points
   (lamw..lamn are (nx,ny) arrays holding LAMBDA at (x-dx/2,y), (x+dx/2,y), (x,y-dy/2), (x,y+dy/2)). */
void orc_heat2d_coeffs(const double *lamw, const double *lame, const double *lams, const double *lamn,
                       double dx, double dy, double *aw, double *ae, double *as, double *an, double *ap,
                       int nx, int ny) {
  int i, j;
  for (i = 2; i <= nx - 1; ++i)
    for (j = 2; j <= ny - 1; ++j) {
      T2(aw, i, j) = T2(lamw, i, j) / (dx * dx);
      T2(ae, i, j) = T2(lame, i, j) / (dx * dx);
      T2(as, i, j) = T2(lams, i, j) / (dy * dy);
      T2(an, i, j) = T2(lamn, i, j) / (dy * dy);
      T2(ap, i, j) = -(T2(aw, i, j) + T2(ae, i, j) + T2(as, i, j) + T2(an, i, j));
    }
}

/*
 * Implicit Euler, locally one-dimensional in 2-D: each step solves, for every interior row j, the 1-D implicit
 * system of implicit-time/fdm1d_explicit_heat_cond.f90:122-143 along x (source added in this half), then for
 * every interior column i the same system along y (no source), with identity rows at the Dirichlet ends.
 */
void orc_heat2d_lod(double k, double rho, double cp, double xa, double xb, double ya, double yb,
                    double tleft, double tright, double tbottom, double ttop,
                    double dt, double qthr, double qval, double *tc, int nx, int ny, int nsteps) {
  double dx = (xb - xa) / (double)(nx - 1), dy = (yb - ya) / (double)(ny - 1);
  double fox = k * dt / rho / cp / (dx * dx), foy = k * dt / rho / cp / (dy * dy);
  int nmax = nx > ny ? nx : ny, i, j, n;
  size_t nb = sizeof(double) * (size_t)nmax;
  double *a = (double *)malloc(nb), *b = (double *)malloc(nb), *c = (double *)malloc(nb);
  double *r = (double *)malloc(nb), *s = (double *)malloc(nb), *x = (double *)malloc(sizeof(double) * (size_t)nx);
  X1(x, 1) = xa;
  for (i = 2; i <= nx; ++i) X1(x, i) = X1(x, 1) + (i - 1) * dx;
  for (n = 0; n < nsteps; ++n) {
    for (j = 2; j <= ny - 1; ++j) {
      T2(tc, 1, j) = tleft;
      T2(tc, nx, j) = tright;
    }
    for (i = 1; i <= nx; ++i) {
      T2(tc, i, 1) = tbottom;
      T2(tc, i, ny) = ttop;
    }
    for (j = 2; j <= ny - 1; ++j) {            /* x-lines */
      X1(a, 1) = 0.0; X1(b, 1) = 1.0; X1(c, 1) = 0.0; X1(r, 1) = T2(tc, 1, j);
      for (i = 2; i <= nx - 1; ++i) {
        X1(a, i) = -fox; X1(b, i) = 1.0 + 2.0 * fox; X1(c, i) = -fox;
        X1(r, i) = T2(tc, i, j) + dt * orc_qdot_step(X1(x, i), qthr, qval) / rho / cp;
      }
      X1(a, nx) = 0.0; X1(b, nx) = 1.0; X1(c, nx) = 0.0; X1(r, nx) = T2(tc, nx, j);
      orc_thomas(a, b, c, r, s, nx);
      for (i = 1; i <= nx; ++i) T2(tc, i, j) = X1(s, i);
    }
    for (i = 2; i <= nx - 1; ++i) {            /* y-lines */
      X1(a, 1) = 0.0; X1(b, 1) = 1.0; X1(c, 1) = 0.0; X1(r, 1) = T2(tc, i, 1);
      for (j = 2; j <= ny - 1; ++j) {
        X1(a, j) = -foy; X1(b, j) = 1.0 + 2.0 * foy; X1(c, j) = -foy;
        X1(r, j) = T2(tc, i, j);
      }
      X1(a, ny) = 0.0; X1(b, ny) = 1.0; X1(c, ny) = 0.0; X1(r, ny) = T2(tc, i, ny);
      orc_thomas(a, b, c, r, s, ny);
      for (j = 1; j <= ny; ++j) T2(tc, i, j) = X1(s, j);
    }
  }
  free(a); free(b); free(c); free(r); free(s); free(x);
}

/* ---- variable conductivity: the operator of fdm2d_steady_heat_cond.f90:162-190 as it stands ------------------------
   Coefficient arrays aw, ae, as, an come from orc_heat2d_coeffs (LAMBDA at the four face mid-points / dx**2, dy**2);
   ap = -(aw + ae + as + an) (:180); rhs = -QDOT (:188).  q holds QDOT at the nodes. */

/* residual of the steady equations on the interior nodes: aw*T(i-1,j) + ae*T(i+1,j) + as*T(i,j-1) + an*T(i,j+1) + ap*T + q
   (the matrix row :182-186 applied to T minus the right-hand side); 0 on the boundary nodes */
void orc_heat2d_apply(const double *aw, const double *ae, const double *as, const double *an, const double *q,
                      const double *t, double *r, int nx, int ny) {
  int i, j;
  memset(r, 0, sizeof(double) * (size_t)nx * (size_t)ny);
  for (j = 2; j <= ny - 1; ++j)
    for (i = 2; i <= nx - 1; ++i) {
      double ap = -(T2(aw, i, j) + T2(ae, i, j) + T2(as, i, j) + T2(an, i, j));
      T2(r, i, j) = T2(aw, i, j) * T2(t, i - 1, j) + T2(ae, i, j) * T2(t, i + 1, j) + T2(as, i, j) * T2(t, i, j - 1) +
                    T2(an, i, j) * T2(t, i, j + 1) + ap * T2(t, i, j) + T2(q, i, j);
    }
}

/* the explicit update of fdm1d_explicit_heat_cond.f90:133-135 with that operator in place of k*(T(i+1)-2T(i)+T(i-1))/dx**2:
   T = Tc + ((aw*Tw + ae*Te + as*Ts + an*Tn + ap*Tc + q)/rho/cp)*dt, Dirichlet edges rewritten every step (:131-132) */
void orc_heat2d_ftcs_var(const double *aw, const double *ae, const double *as, const double *an, const double *q,
                         double rho, double cp, double dt, double tleft, double tright, double tbottom, double ttop,
                         double *tc, int nx, int ny, int nsteps) {
  double *t = (double *)malloc(sizeof(double) * (size_t)nx * (size_t)ny);
  int i, j, n;
  for (n = 0; n < nsteps; ++n) {
    for (j = 2; j <= ny - 1; ++j) {
      T2(t, 1, j) = tleft;
      T2(t, nx, j) = tright;
    }
    for (i = 1; i <= nx; ++i) {
      T2(t, i, 1) = tbottom;
      T2(t, i, ny) = ttop;
    }
    for (j = 2; j <= ny - 1; ++j)
      for (i = 2; i <= nx - 1; ++i) {
        double ap = -(T2(aw, i, j) + T2(ae, i, j) + T2(as, i, j) + T2(an, i, j));
        double op = T2(aw, i, j) * T2(tc, i - 1, j) + T2(ae, i, j) * T2(tc, i + 1, j) + T2(as, i, j) * T2(tc, i, j - 1) +
                    T2(an, i, j) * T2(tc, i, j + 1) + ap * T2(tc, i, j) + T2(q, i, j);
        T2(t, i, j) = T2(tc, i, j) + (op / rho / cp) * dt;
      }
    memcpy(tc, t, sizeof(double) * (size_t)nx * (size_t)ny);
  }
  free(t);
}

/* gauss_seidel (fdm2d_steady_heat_cond.f90:248-287) on the matrix that interior() (:162-190) and boundary() (:193-246)
   build, without forming it: row k = (j-1)*nx + i has a(k,k-nx) = as, a(k,k-1) = aw, a(k,k) = ap, a(k,k+1) = ae,
   a(k,k+nx) = an and rhs -QDOT on interior nodes, a(k,k) = 1 and rhs = wall temperature on boundary nodes (bottom / top
   rows include the corners).  The dense inner loop adds a(i,j)*x(j) in increasing j, zeros included: s = as*xS + aw*xW
   + ae*xE + an*xN in that order, with the already updated x2 for j < i and x1 for j > i.  x2 = (b - s)*rf/a(k,k) +
   (1-rf)*x1; stop when sqrt(sum (x1-x2)**2) < tol or after imax passes.  order 0: the reference's single lexicographic
   pass; order 1: two colours, (i+j) even first (what the CUDA path does; same fixed point).  x holds the start vector
   (the reference starts from 0) and receives the result.  Returns the number of passes; *dx2_out = the last norm. */
int orc_heat2d_gs(const double *aw, const double *ae, const double *as, const double *an, const double *q,
                  double tleft, double tright, double tbottom, double ttop, double rf, double tol, int imax, int order,
                  double *x, int nx, int ny, double *dx2_out) {
  size_t n = (size_t)nx * (size_t)ny;
  double *x1 = (double *)malloc(sizeof(double) * n), *x2 = (double *)malloc(sizeof(double) * n);
  double dx2 = 0.0;
  int i, j, k, col, done = 0;
  memcpy(x1, x, sizeof(double) * n);
  memcpy(x2, x, sizeof(double) * n);
  for (k = 1; k <= imax; ++k) {
    for (col = 0; col < (order ? 2 : 1); ++col)
      for (j = 1; j <= ny; ++j)
        for (i = 1; i <= nx; ++i) {
          double s, b, akk;
          if (order && ((i + j) & 1) != col) continue;
          if (i == 1 || i == nx || j == 1 || j == ny) {
            s = 0.0;
            akk = 1.0;
            b = (j == 1) ? tbottom : (j == ny) ? ttop : (i == 1) ? tleft : tright;
          } else {
            /* in both orders the neighbours already visited in this pass hold x2, the others x1 */
            double xs = order ? (col ? T2(x2, i, j - 1) : T2(x1, i, j - 1)) : T2(x2, i, j - 1);
            double xw = order ? (col ? T2(x2, i - 1, j) : T2(x1, i - 1, j)) : T2(x2, i - 1, j);
            double xe = order ? (col ? T2(x2, i + 1, j) : T2(x1, i + 1, j)) : T2(x1, i + 1, j);
            double xn = order ? (col ? T2(x2, i, j + 1) : T2(x1, i, j + 1)) : T2(x1, i, j + 1);
            s = T2(as, i, j) * xs + T2(aw, i, j) * xw + T2(ae, i, j) * xe + T2(an, i, j) * xn;
            akk = -(T2(aw, i, j) + T2(ae, i, j) + T2(as, i, j) + T2(an, i, j));
            b = -T2(q, i, j);
          }
          T2(x2, i, j) = (b - s) * rf / akk + (1 - rf) * T2(x1, i, j);
        }
    dx2 = 0.0;
    for (i = 0; i < (int)n; ++i) dx2 = dx2 + (x1[i] - x2[i]) * (x1[i] - x2[i]);
    dx2 = sqrt(dx2);
    done = k;
    if (dx2 < tol) break;
    memcpy(x1, x2, sizeof(double) * n);
  }
  memcpy(x, x2, sizeof(double) * n);
  if (dx2_out) *dx2_out = dx2;
  free(x1);
  free(x2);
  return done;
}
