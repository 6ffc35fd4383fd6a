/* examples/cavity_driver.c -- `program cavity_fvm` (lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90:507-609) call for
 * call over the C ABI: what the Fortran driver does once its calcu / calcv / calcp forward to libcfd_b200.so
 * (INTEGRATION.md).  The host keeps the reference's array shapes (column-major, (ni,nj)); only the three calc
 * routines and the solver run on the GPU.
 *
 *   gcc -O2 -Iinclude examples/cavity_driver.c -Lcfd_b200/lib -lcfd_b200 -Wl,-rpath,$PWD/cfd_b200/lib -lm -o cavity_driver
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "cfd_b200.h"

#define CHECK(call)                                                                      \
  do {                                                                                   \
    int rc__ = (call);                                                                   \
    if (rc__ != CFD_OK) {                                                                \
      fprintf(stderr, "%s -> %s: %s\n", #call, cfd_strerror(rc__), cfd_last_error());   \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

/* setgrid (cavity_fvm.f90:366-414): uniform, cell-centred, backward-staggered */
static void setgrid(int nicv, int njcv, double xs, double xe, double ys, double ye, double *x, double *xu, double *sew,
                    double *sewu, double *y, double *yv, double *sns, double *snsv) {
  const int ni = nicv + 2, nj = njcv + 2;
  const double dx = (xe - xs) / (double)nicv, dy = (ye - ys) / (double)njcv;
  int i, j;
  for (i = 0; i < ni; ++i) x[i] = xu[i] = sew[i] = sewu[i] = 0.0;
  for (j = 0; j < nj; ++j) y[j] = yv[j] = sns[j] = snsv[j] = 0.0;
  xu[0] = xu[1] = xs;
  for (i = 2; i < ni; ++i) xu[i] = xu[i - 1] + dx;
  x[0] = xu[1];
  for (i = 1; i < ni - 1; ++i) { x[i] = 0.5 * (xu[i] + xu[i + 1]); sew[i] = xu[i + 1] - xu[i]; }
  x[ni - 1] = xu[ni - 1];
  for (i = 2; i < ni - 1; ++i) sewu[i] = x[i] - x[i - 1];
  yv[0] = yv[1] = ys;
  for (j = 2; j < nj; ++j) yv[j] = yv[j - 1] + dy;
  y[0] = yv[1];
  for (j = 1; j < nj - 1; ++j) { y[j] = 0.5 * (yv[j] + yv[j + 1]); sns[j] = yv[j + 1] - yv[j]; }
  y[nj - 1] = yv[nj - 1];
  for (j = 2; j < nj - 1; ++j) snsv[j] = y[j] - y[j - 1];
}

int main(int argc, char **argv) {
  const int nicv = argc > 1 ? atoi(argv[1]) : 10, njcv = argc > 2 ? atoi(argv[2]) : 10;   /* :519-520 */
  const int ni = nicv + 2, nj = njcv + 2;
  const double reynolds = 100.0, rho = 1000.0, utop = 1.0, tol = 1.0e-10;                  /* :521-527 */
  const int maxit = 90000;
  cfd_config cfg = {0};
  cfd_ctx *ctx = NULL;
  double *g[8], *den, *vis, *u, resoru = 0, resorv = 0, resorm = 0, umon;
  int k, i, j, iter;

  cfg.family = CFD_FAMILY_CAVITY;
  cfg.ni = ni; cfg.nj = nj; cfg.iadv = 3;                        /* hybrid (:528) */
  cfg.order = CFD_ORDER_SEQUENTIAL;                              /* the reference's sweep order; CFD_ORDER_REDBLACK = production */
  cfg.nswp[0] = cfg.nswp[1] = cfg.nswp[2] = 3;                   /* :545-547 */
  cfg.urf[0] = cfg.urf[1] = 0.5; cfg.urf[2] = 0.8;               /* :541-543 */
  cfg.ipref = ni / 2; cfg.jpref = nj / 2;                        /* :549-550 */
  cfg.device = -1;
  CHECK(cfd_create(&ctx, &cfg));                                 /* array_alloc (:530) */

  for (k = 0; k < 4; ++k) g[k] = (double *)malloc(sizeof(double) * (size_t)ni);
  for (k = 4; k < 8; ++k) g[k] = (double *)malloc(sizeof(double) * (size_t)nj);
  setgrid(nicv, njcv, 0.0, 1.0, 0.0, 1.0, g[0], g[1], g[2], g[3], g[4], g[5], g[6], g[7]);   /* :532 */
  for (k = 0; k < 8; ++k) CHECK(cfd_set_geometry(ctx, CFD_G_X + k, g[k], k < 4 ? ni : nj));

  den = (double *)malloc(sizeof(double) * (size_t)ni * nj);
  vis = (double *)malloc(sizeof(double) * (size_t)ni * nj);
  u = (double *)calloc((size_t)ni * nj, sizeof(double));
  for (j = 0; j < nj; ++j)
    for (i = 0; i < ni; ++i) {
      den[i + (size_t)j * ni] = rho;                             /* :534 */
      vis[i + (size_t)j * ni] = rho * utop * 1.0 / reynolds;     /* :535 */
    }
  for (i = 1; i < ni; ++i) u[i + (size_t)(nj - 1) * ni] = utop;  /* lid: u(2:ni,nj) = utop (:538) */
  CHECK(cfd_upload(ctx, CFD_DEN, den, ni));
  CHECK(cfd_upload(ctx, CFD_VIS, vis, ni));
  CHECK(cfd_upload(ctx, CFD_U, u, ni));

  for (iter = 1; iter <= maxit; ++iter) {                        /* :553-566 */
    CHECK(cfd_calcu(ctx, &resoru));
    CHECK(cfd_calcv(ctx, &resorv));
    CHECK(cfd_calcp(ctx, &resorm));
    if (iter % 500 == 0 || iter == 1) {
      CHECK(cfd_probe(ctx, CFD_U, cfg.ipref, cfg.jpref, &umon));
      printf("%6d %12.4e %12.4e %12.4e  u(mon) = %14.7e\n", iter, resoru, resorv, resorm, umon);
    }
    if (fmax(resoru, fmax(resorv, resorm)) < tol) break;
  }
  CHECK(cfd_download(ctx, CFD_U, u, ni));                        /* results back for centred_vel / Tecplot (:567-607) */
  printf("converged after %d outer iterations; u(%d,%d) = %.15e\n", iter, ni / 2, nj / 2,
         u[(ni / 2 - 1) + (size_t)(nj / 2 - 1) * ni]);
  CHECK(cfd_destroy(ctx));
  for (k = 0; k < 8; ++k) free(g[k]);
  free(den); free(vis); free(u);
  return 0;
}
