/*
 * oracle_promol.c -- restatement of grid1 radial tables (read_critic, interp) and environ%promolecular.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Citations: /root/reference/src/grid1mod@proc.f90, /root/reference/src/environmod@proc.f90:1202-1327.
 */
#include "oracle.h"
#include "oracle_util.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

orc_grid1 *orc_grid1_from_table(int z, int qat, int ngrid, double a, double b, const double *r, const double *f,
                                const double *fp, const double *fpp) {
  orc_grid1 *g = calloc(1, sizeof(orc_grid1));
  g->isinit = 1; g->z = z; g->qat = qat; g->ngrid = ngrid; g->a = a; g->b = b;
  size_t nb = sizeof(double) * ngrid;
  g->r = malloc(nb); g->f = malloc(nb); g->fp = malloc(nb); g->fpp = malloc(nb);
  memcpy(g->r, r, nb); memcpy(g->f, f, nb); memcpy(g->fp, fp, nb); memcpy(g->fpp, fpp, nb);
  g->rmax = r[ngrid - 1];
  g->rmax2 = g->rmax * g->rmax;
  return g;
}

void orc_grid1_free(orc_grid1 *g) {
  if (!g) return;
  free(g->r); free(g->f); free(g->fp); free(g->fpp);
  free(g);
}

/* 6-point derivation formulas grid1mod@proc.f90:25-38 */
static const int noef[3][6] = {{0, 1, 2, 3, 4, 5}, {-2, -1, 0, 1, 2, 3}, {-5, -4, -3, -2, -1, 0}};
static const double coef1[3][6] = {{-274, 600, -600, 400, -150, 24}, {6, -60, -40, 120, -30, 4}, {-24, 150, -400, 600, -600, 274}};
static const double coef2[3][6] = {{225, -770, 1070, -780, 305, -50}, {-5, 80, -150, 80, -5, 0}, {-50, 305, -780, 1070, -770, 225}};
static const double core_cutdens = 1e-8; /* :41 */

/* read_critic :204-332 with n = z - q electrons (read_db :58-81) */
orc_grid1 *orc_grid1_read_critic(const char *file, int z, int q) {
  FILE *fp = fopen(file, "r");
  if (!fp) return NULL;
  int nn;
  if (fscanf(fp, "%d", &nn) != 1) { fclose(fp); return NULL; }
  char lab[16];
  int *occ = calloc(nn, sizeof(int));
  double *wfcin = malloc(sizeof(double) * nn);
  for (int i = 0; i < nn; i++) if (fscanf(fp, "%15s", lab) != 1) { fclose(fp); return NULL; }
  for (int i = 0; i < nn; i++) if (fscanf(fp, "%d", &occ[i]) != 1) { fclose(fp); return NULL; }
  double xmin, zz, dx;
  int ngrid;
  if (fscanf(fp, "%lf %lf %lf %d", &xmin, &zz, &dx, &ngrid) != 4) { fclose(fp); return NULL; }
  int n = z - q, tot = 0;
  for (int i = 0; i < nn; i++) tot += occ[i];
  if (tot != n) {                                               /* :240-251 */
    int ns = 0;
    for (int i = 0; i < nn; i++) {
      if (ns + occ[i] > n) {
        occ[i] = n - ns;
        for (int j = i + 1; j < nn; j++) occ[j] = 0;
        break;
      } else
        ns += occ[i];
    }
  }
  int nalloc = ngrid;
  double *r = malloc(sizeof(double) * nalloc);
  double *rr0 = calloc(nalloc, sizeof(double)), *rr1 = calloc(nalloc, sizeof(double)), *rr2 = calloc(nalloc, sizeof(double));
  const double pi = 3.14159265358979323846264338328;
  for (int i = 0; i < nalloc; i++) {                            /* :253-265 */
    double rv;
    if (fscanf(fp, "%lf", &rv) != 1) { ngrid = i; break; }
    for (int j = 0; j < nn; j++) if (fscanf(fp, "%lf", &wfcin[j]) != 1) wfcin[j] = 0;
    r[i] = rv;
    double s = 0;
    for (int j = 0; j < nn; j++) s += occ[j] * (wfcin[j] * wfcin[j]);
    rr0[i] = s;
    if (rr0[i] / (4.0 * pi * rv * rv) < core_cutdens && i > 0) { ngrid = i + 1; break; }
  }
  fclose(fp);
  orc_grid1 *g = calloc(1, sizeof(orc_grid1));
  g->isinit = 1; g->z = z; g->qat = q;
  g->a = exp(xmin) / zz; g->b = dx; g->ngrid = ngrid;
  g->rmax = r[ngrid - 1]; g->rmax2 = g->rmax * g->rmax;
  g->r = malloc(sizeof(double) * ngrid); g->f = malloc(sizeof(double) * ngrid);
  g->fp = malloc(sizeof(double) * ngrid); g->fpp = malloc(sizeof(double) * ngrid);
  const double fac1 = 1.0 / 120.0, fac2 = 2.0 / 120.0;
  for (int i = 1; i <= ngrid; i++) {                            /* :276-306, 1-based i */
    int ic = (i <= 2) ? 0 : ((i >= ngrid - 2) ? 2 : 1);
    for (int j = 0; j < 6; j++) {
      rr1[i - 1] += coef1[ic][j] * rr0[i - 1 + noef[ic][j]];
      rr2[i - 1] += coef2[ic][j] * rr0[i - 1 + noef[ic][j]];
    }
    rr1[i - 1] *= fac1;
    rr2[i - 1] *= fac2;
    double rv = r[i - 1];
    double r1 = 1.0 / rv, r2 = r1 * r1, r3 = r2 * r1, r4 = r3 * r1;
    double delta = 1.0 / g->b, delta2 = delta * delta;
    g->r[i - 1] = rv;
    g->f[i - 1] = rr0[i - 1] * r2;
    g->fp[i - 1] = (rr1[i - 1] * delta - 2.0 * rr0[i - 1]) * r3;
    g->fpp[i - 1] = (rr2[i - 1] * delta2 - 5.0 * rr1[i - 1] * delta + 6.0 * rr0[i - 1]) * r4;
  }
  for (int i = 0; i < ngrid; i++) {                             /* :307-309 */
    g->f[i] = g->f[i] / (4.0 * pi);
    g->fp[i] = g->fp[i] / (4.0 * pi);
    g->fpp[i] = g->fpp[i] / (4.0 * pi);
  }
  free(occ); free(wfcin); free(r); free(rr0); free(rr1); free(rr2);
  return g;
}

/* interp :84-135 */
void orc_grid1_interp(const orc_grid1 *g, double r0, double *f, double *fp, double *fpp) {
  *f = 0; *fp = 0; *fpp = 0;
  if (!g || !g->isinit) return;
  if (r0 >= g->rmax) return;
  if (g->z - g->qat <= 0) return;
  int ir;
  double r;
  if (r0 <= g->r[0]) { ir = 1; r = g->r[0]; }
  else { ir = 1 + (int)floor(log(r0 / g->a) / g->b); r = r0; }
  int base = ir;
  if (base < 2) base = 2;
  if (base > g->ngrid - 2) base = g->ngrid - 2;
  base -= 2; /* ii = base + i, i = 1..4 (1-based) */
  double rr[4], dr1[4], x1dr12[4][4];
  memset(x1dr12, 0, sizeof(x1dr12));
  for (int i = 0; i < 4; i++) {
    int ii = base + i; /* 0-based */
    rr[i] = g->r[ii];
    dr1[i] = r - rr[i];
    for (int j = 0; j < i; j++) {
      x1dr12[i][j] = 1.0 / (rr[i] - rr[j]);
      x1dr12[j][i] = -x1dr12[i][j];
    }
  }
  for (int i = 0; i < 4; i++) {
    int ii = base + i;
    double prod = 1.0;
    for (int j = 0; j < 4; j++) {
      if (i == j) continue;
      prod = prod * dr1[j] * x1dr12[i][j];
    }
    *f += g->f[ii] * prod;
    *fp += g->fp[ii] * prod;
    *fpp += g->fpp[ii] * prod;
  }
}

/* promolecular :1202-1327 (no zpsp/core variant; x0 Cartesian) */
void orc_promolecular(const orc_environ *e, orc_grid1 *const *agrid, const double *x0, int nder, const unsigned char *fragmask,
                      double *f, double *fp, double *fpp) {
  *f = 0;
  for (int i = 0; i < 3; i++) fp[i] = 0;
  for (int i = 0; i < 9; i++) fpp[i] = 0;
  double xc[3] = {x0[0], x0[1], x0[2]};
  orc_center_cart(e, xc, NULL);                                 /* :1242-1243 */
  double *rcutmax = calloc(e->nspc, sizeof(double));            /* :1246-1263 */
  for (int i = 0; i < e->nspc; i++) {
    int iz = e->zspc[i];
    if (iz == 0 || iz > 118) continue;
    const orc_grid1 *g = agrid[iz];
    if (g && g->isinit) rcutmax[i] = fmin(orc_cutrad(iz), g->rmax);
    else rcutmax[i] = orc_cutrad(iz);
  }
  int *nid = malloc(sizeof(int) * e->n);
  double *dist = malloc(sizeof(double) * e->n);
  int nat, lvec[3];
  orc_list_near_atoms(e, xc, 1, rcutmax, &nat, nid, dist, lvec); /* :1268 */
  free(rcutmax);
  double rl[3] = {lvec[0], lvec[1], lvec[2]}, rlvec[3];
  orc_matvec(e->m_x2c, rl, rlvec);                              /* :1270-1271 */
  for (int ii = 0; ii < nat; ii++) {                            /* :1284-1323 */
    int i = nid[ii];
    if (fragmask && !fragmask[e->cidx[i] - 1]) continue;
    int iz = e->zspc[e->is[i] - 1];
    if (iz == 0 || iz > 118) continue;
    double r = dist[ii];
    if (r > orc_cutrad(iz)) continue;
    const orc_grid1 *g = agrid[iz];
    if (!g || !g->isinit) continue;
    r = fmax(fmax(r, g->r[0]), 1e-14);
    double rho, rhop, rhopp;
    orc_grid1_interp(g, r, &rho, &rhop, &rhopp);
    rho = fmax(rho, 0.0);
    *f += rho;
    if (nder < 1) continue;
    double xx[3];
    for (int d = 0; d < 3; d++) xx[d] = xc[d] - (e->r[3 * i + d] + rlvec[d]);
    double rinv1 = 1.0 / r;
    double rinv1rp = rinv1 * rhop;
    for (int d = 0; d < 3; d++) fp[d] += xx[d] * rinv1rp;
    if (nder < 2) continue;
    double rfac = (rhopp - rinv1rp) * rinv1 * rinv1;
    for (int j = 0; j < 3; j++) {
      fpp[j + 3 * j] += rinv1rp + rfac * xx[j] * xx[j];
      for (int k = 0; k < j; k++) {
        double radd = rfac * xx[j] * xx[k];
        fpp[j + 3 * k] += radd;
        fpp[k + 3 * j] += radd;
      }
    }
  }
  free(nid);
  free(dist);
}

/* The atoms of promolecular's sum loop (:1284-1300) that reach interp: inside list_near_atoms' species cutoff
 * min(cutrad, rmax), in the fragment, with an initialised table.  x0 Cartesian; folded like :1242-1243. */
static long orc_promol_nat(const orc_environ *e, orc_grid1 *const *agrid, const double *x0, const unsigned char *fragmask) {
  double xc[3] = {x0[0], x0[1], x0[2]};
  orc_center_cart(e, xc, NULL);
  double *rcutmax = calloc(e->nspc, sizeof(double));
  for (int i = 0; i < e->nspc; i++) {
    int iz = e->zspc[i];
    if (iz == 0 || iz > 118) continue;
    const orc_grid1 *g = agrid[iz];
    if (g && g->isinit) rcutmax[i] = fmin(orc_cutrad(iz), g->rmax);
    else rcutmax[i] = orc_cutrad(iz);
  }
  int *nid = malloc(sizeof(int) * e->n);
  double *dist = malloc(sizeof(double) * e->n);
  int nat, lvec[3];
  orc_list_near_atoms(e, xc, 1, rcutmax, &nat, nid, dist, lvec);
  free(rcutmax);
  long cnt = 0;
  for (int ii = 0; ii < nat; ii++) {
    int i = nid[ii];
    if (fragmask && !fragmask[e->cidx[i] - 1]) continue;
    int iz = e->zspc[e->is[i] - 1];
    if (iz == 0 || iz > 118) continue;
    if (dist[ii] > orc_cutrad(iz)) continue;
    const orc_grid1 *g = agrid[iz];
    if (!g || !g->isinit) continue;
    cnt++;
  }
  free(nid);
  free(dist);
  return cnt;
}

void orc_promol_nat_points(const orc_environ *e, orc_grid1 *const *agrid, long npts, const double *pts,
                           const unsigned char *fragmask, long *nat, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (long p = 0; p < npts; p++) nat[p] = orc_promol_nat(e, agrid, pts + 3 * p, fragmask);
}
