/*
 * oracle_field.c -- restatement of field%grd / grd0, eig (trace/tredig/tqlgrm) and the CUBE / NCIPLOT grid loops.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Citations: /root/reference/src/fieldmod@proc.f90, src/rhoplot@proc.f90, src/nci@proc.f90, src/oldlibs/{trace,tredig,tqlgrm}.f
 */
#include "oracle.h"
#include "oracle_util.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static const double flooreps = 1e-4; /* fieldmod@proc.f90:30 */

/* wrap + transform of grd :698-721.  Returns wc (via cell matrices) and wcr (via reduced-cell matrices). */
static void grd_points(const orc_environ *e, const double *v, int per, double *wc, double *wcr) {
  double wx[3], wxr[3];
  orc_matvec(e->m_c2x, v, wx);
  orc_matvec(e->m_c2xr, v, wxr);
  if (per) {
    for (int i = 0; i < 3; i++) {
      if (wx[i] < -flooreps || wx[i] > 1.0 + flooreps) wx[i] = wx[i] - floor(wx[i]);
      if (wxr[i] < -flooreps || wxr[i] > 1.0 + flooreps) wxr[i] = wxr[i] - floor(wxr[i]);
    }
  }
  orc_matvec(e->m_x2c, wx, wc);
  orc_matvec(e->m_xr2c, wxr, wcr);
}

/* grd :613-842, branches type_wfn (:759-794), type_promol / type_promol_frag (:802-808), tail :834-840 */
void orc_grd(const orc_field *f, const double *v, int nder, int periodic, double *res) {
  memset(res, 0, sizeof(double) * ORC_RES_LEN);
  double wc[3], wcr[3];
  grd_points(f->env, v, periodic, wc, wcr);
  double *gf = res + 1, *hf = res + 4;
  if (f->type == ORC_TYPE_WFN) {
    double o[91];
    orc_rho2(f->wfn, f->env, wcr, nder, o);
    res[0] = o[0];
    res[16] = o[3];
    res[28] = o[1]; res[29] = o[2];
    for (int k = 0; k < 3; k++) gf[k] = o[6 + k];
    for (int k = 0; k < 9; k++) hf[k] = o[24 + k];
    res[17] = o[78];
    res[18] = o[81];
    for (int k = 0; k < 9; k++) res[19 + k] = o[82 + k];
  } else {
    orc_promolecular(f->env, f->agrid, wcr, nder, f->fragmask, &res[0], gf, hf);
    res[16] = res[0];
  }
  if (f->cgrid) {                                               /* :827-832 core augmentation, evaluated at wc */
    double rho, grad[3], h[9];
    orc_promolecular(f->env, f->cgrid, wc, nder, NULL, &rho, grad, h);
    res[0] = res[0] + rho;
    for (int k = 0; k < 3; k++) gf[k] = gf[k] + grad[k];
    for (int k = 0; k < 9; k++) hf[k] = hf[k] + h[k];
  }
  int nid = orc_identify_atom(f->env, wc, 1e-5);                /* :836 */
  res[15] = (nid > 0);
  if (nid > 0) gf[0] = gf[1] = gf[2] = 0.0;                      /* :838 */
  res[13] = orc_norm2(gf, 3);                                   /* :839 */
  res[14] = hf[0] + hf[4] + hf[8];                              /* :840 */
}

/* grd0 :848-928 */
double orc_grd0(const orc_field *f, const double *v, int periodic) {
  double wc[3], wcr[3];
  grd_points(f->env, v, periodic, wc, wcr);
  double rho, g[3], h[9];
  if (f->type == ORC_TYPE_WFN) {
    double o[91];
    orc_rho2(f->wfn, f->env, wcr, 0, o);
    rho = o[0];
  } else
    orc_promolecular(f->env, f->agrid, wcr, 0, f->fragmask, &rho, g, h);
  if (f->cgrid) {                                               /* :922-925 */
    double rhoaux;
    orc_promolecular(f->env, f->cgrid, wc, 0, NULL, &rhoaux, g, h);
    rho = rho + rhoaux;
  }
  return rho;
}

/* ---- eig: Householder tridiagonalisation (tredig.f) + QL with implicit shifts (tqlgrm.f) ----
   z is n x n column-major: Z(i,j) = z[(i-1) + n*(j-1)].  Written with 1-based macros to follow the F77 flow. */
#define Z(i, j) z[((i)-1) + (size_t)n * ((j)-1)]
#define D(i) d[(i)-1]
#define E(i) e[(i)-1]
static void tredig(int n, double *d, double *e, double *z) {
  if (n > 1) {
    for (int ii = 2; ii <= n; ii++) {
      int i = n + 2 - ii, l = i - 1;
      double h = 0, scale = 0;
      int simple = (l < 2);
      if (!simple) {
        for (int k = 1; k <= l; k++) scale += fabs(Z(i, k));
        if (scale == 0) simple = 1;
      }
      if (simple) {
        E(i) = Z(i, l);
      } else {
        double rscale = 1.0 / scale;
        for (int k = 1; k <= l; k++) {
          Z(i, k) = Z(i, k) * rscale;
          h += Z(i, k) * Z(i, k);
        }
        double f = Z(i, l);
        double g = -copysign(sqrt(h), f);
        E(i) = scale * g;
        h = h - f * g;
        Z(i, l) = f - g;
        f = 0;
        double rh = 1.0 / h, rhscale = rh * rscale;
        for (int j = 1; j <= l; j++) {
          Z(j, i) = Z(i, j) * rhscale;
          g = 0;
          for (int k = 1; k <= j; k++) g += Z(j, k) * Z(i, k);
          for (int k = j + 1; k <= l; k++) g += Z(k, j) * Z(i, k);
          E(j) = g * rh;
          f += E(j) * Z(i, j);
        }
        double hh = f / (h + h);
        for (int j = 1; j <= l; j++) {
          f = Z(i, j);
          g = E(j) - hh * f;
          E(j) = g;
          for (int k = 1; k <= j; k++) Z(j, k) = Z(j, k) - f * E(k) - g * Z(i, k);
        }
        for (int k = 1; k <= l; k++) Z(i, k) = scale * Z(i, k);
      }
      D(i) = h;
    }
  }
  D(1) = 0;
  E(1) = 0;
  for (int i = 1; i <= n; i++) {
    int l = i - 1;
    if (D(i) != 0) {
      for (int j = 1; j <= l; j++) {
        double g = 0;
        for (int k = 1; k <= l; k++) g += Z(i, k) * Z(k, j);
        for (int k = 1; k <= l; k++) Z(k, j) = Z(k, j) - g * Z(k, i);
      }
    }
    D(i) = Z(i, i);
    Z(i, i) = 1.0;
    for (int j = 1; j <= l; j++) { Z(j, i) = 0; Z(i, j) = 0; }
  }
}

static int tqlgrm(int n, double *d, double *e, double *z) {
  const double amach = (double)16.0e-13f; /* PARAMETER (AMACH = 16.0E-13): a default-real literal, tqlgrm.f:10 */
  if (n == 1) return 0;
  for (int i = 2; i <= n; i++) E(i - 1) = E(i);
  double f = 0, b = 0;
  E(n) = 0;
  for (int l = 1; l <= n; l++) {
    int j = 0;
    double h = amach * (fabs(D(l)) + fabs(E(l)));
    if (b < h) b = h;
    int m;
    for (m = l; m <= n; m++)
      if (fabs(E(m)) <= b) break;
    if (m != l) {
      do {
        if (j == 30) return l;
        j++;
        int l1 = l + 1;
        double g = D(l);
        double p = (D(l1) - g) / (2 * E(l));
        double r = (fabs(p * amach) > 1.0) ? p : sqrt(p * p + 1);
        D(l) = E(l) / (p + copysign(r, p));
        h = g - D(l);
        for (int i = l1; i <= n; i++) D(i) = D(i) - h;
        f = f + h;
        p = D(m);
        double c = 1.0, s = 0.0;
        for (int ii = 1; ii <= m - l; ii++) {
          int i = m - ii;
          g = c * E(i);
          h = c * p;
          if (fabs(p) >= fabs(E(i))) {
            c = E(i) / p;
            r = sqrt(c * c + 1);
            E(i + 1) = s * p * r;
            s = c / r;
            c = 1.0 / r;
          } else {
            c = p / E(i);
            r = sqrt(c * c + 1);
            E(i + 1) = s * E(i) * r;
            s = 1.0 / r;
            c = c * s;
          }
          p = c * D(i) - s * g;
          D(i + 1) = h + s * (c * g + s * D(i));
          for (int k = 1; k <= n; k++) {
            h = Z(k, i + 1);
            Z(k, i + 1) = s * Z(k, i) + c * h;
            Z(k, i) = c * Z(k, i) - s * h;
          }
        }
        E(l) = s * p;
        D(l) = c * p;
      } while (fabs(E(l)) > b);
    }
    D(l) = D(l) + f;
  }
  for (int ii = 2; ii <= n; ii++) {
    int i = ii - 1, k = i;
    double p = D(i);
    for (int j = ii; j <= n; j++) {
      if (D(j) >= p) continue;
      k = j;
      p = D(j);
    }
    if (k == i) continue;
    D(k) = D(i);
    D(i) = p;
    for (int j = 1; j <= n; j++) {
      p = Z(j, i);
      Z(j, i) = Z(j, k);
      Z(j, k) = p;
    }
  }
  return 0;
}
#undef Z
#undef D
#undef E

int orc_eig(double *h, double *eval, int n) { /* tools_math@proc.f90:822-841, trace.f */
  double work[16];
  double *w = (n <= 16) ? work : malloc(sizeof(double) * n);
  for (int i = 0; i < n; i++) w[i] = 0;
  tredig(n, eval, w, h);
  int ierr = tqlgrm(n, eval, w, h);
  if (w != work) free(w);
  return ierr;
}

static double res_prop(const double *res, int nti) { /* rhoplot@proc.f90:596-621 */
  switch (nti) {
    case 0: return res[0];
    case 1: return res[1];
    case 2: return res[2];
    case 3: return res[3];
    case 4: return res[13];
    case 5: return res[4 + 0];      /* hf(1,1) */
    case 6: return res[4 + 0 + 3];  /* hf(1,2) */
    case 7: return res[4 + 0 + 6];  /* hf(1,3) */
    case 8: return res[4 + 1 + 3];  /* hf(2,2) */
    case 9: return res[4 + 1 + 6];  /* hf(2,3) */
    case 10: return res[4 + 2 + 6]; /* hf(3,3) */
    default: return res[14];
  }
}

/* rhoplot_cube :587-631 */
void orc_cube(const orc_field *f, const double *x0, const double *xd, const int *nn, int periodic, unsigned outmask,
              double *const *out, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic)
  for (int iz = 0; iz < nn[2]; iz++)
    for (int iy = 0; iy < nn[1]; iy++)
      for (int ix = 0; ix < nn[0]; ix++) {
        double xp[3], res[ORC_RES_LEN];
        for (int d = 0; d < 3; d++)
          xp[d] = x0[d] + (double)ix * xd[d] + (double)iy * xd[3 + d] + (double)iz * xd[6 + d]; /* :591 */
        orc_grd(f, xp, 2, periodic, res);                       /* :595 always nder = 2 */
        size_t idx = (size_t)ix + (size_t)nn[0] * ((size_t)iy + (size_t)nn[1] * iz);
        int slot = 0;
        for (int nti = 0; nti < 12; nti++)
          if (outmask & (1u << nti)) out[slot++][idx] = res_prop(res, nti);
      }
}

/* nciplot :499-561 (non-grid reference field, no fragments) */
void orc_nci(const orc_field *fref, const orc_field *fpromol, const double *x0, const double *xmat, const int *nstep,
             double *crho, double *cgrad, double *rhoat, int nthreads) {
  const double pi = 3.14159265358979323846264338328;
  const double fthirds = 4.0 / 3.0;
  const double cnst = 2.0 * pow(3.0 * pi * pi, 1.0 / 3.0);       /* :87 */
  const double vsmall = 1e-80;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic)
  for (int i = 0; i < nstep[0]; i++)
    for (int j = 0; j < nstep[1]; j++)
      for (int k = 0; k < nstep[2]; k++) {
        double x[3], res[ORC_RES_LEN], ehess[3], hf[9];
        for (int d = 0; d < 3; d++) x[d] = x0[d] + i * xmat[d] + j * xmat[3 + d] + k * xmat[6 + d]; /* :504 */
        orc_grd(fref, x, 2, 1, res);                            /* :523 periodic defaults to .true. */
        memcpy(hf, res + 4, sizeof(hf));
        orc_eig(hf, ehess, 3);                                  /* :524 */
        double dimgrad = res[13] / (cnst * pow(fmax(res[0], vsmall), fthirds)); /* :525 */
        size_t idx = (size_t)k + (size_t)nstep[2] * ((size_t)j + (size_t)nstep[1] * i);
        cgrad[idx] = dimgrad;                                   /* :553 */
        crho[idx] = copysign(fabs(res[0]), ehess[1]) * 100.0;   /* sign(res%f,ehess(2))*100 :554 */
        if (rhoat) rhoat[idx] = orc_grd0(fpromol, x, 1);        /* :530 */
      }
}

/* nciplot fragment densities :534-550: plain sums over the fragment's own atoms (no periodic images, the point is
   not folded), cutoff g%rmax only.  xyz/z: the atoms of all fragments back to back (nat[ifr] each), Cartesian.
   rhofrag: [nfrag][n1*n2*n3], each (k,j,i) with k fastest. */
void orc_nci_fragments(int nfrag, const int *nat, const double *xyz, const int *z, orc_grid1 *const *agrid,
                       const double *x0, const double *xmat, const int *nstep, double *rhofrag, int nthreads) {
  const size_t npt = (size_t)nstep[0] * nstep[1] * nstep[2];
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic)
  for (int i = 0; i < nstep[0]; i++)
    for (int j = 0; j < nstep[1]; j++)
      for (int k = 0; k < nstep[2]; k++) {
        double x[3];
        for (int d = 0; d < 3; d++) x[d] = x0[d] + i * xmat[d] + j * xmat[3 + d] + k * xmat[6 + d]; /* :504 */
        size_t idx = (size_t)k + (size_t)nstep[2] * ((size_t)j + (size_t)nstep[1] * i);
        size_t a0 = 0;
        for (int ifr = 0; ifr < nfrag; ifr++) {
          double sum = 0.0;                                       /* :535 */
          for (int ia = 0; ia < nat[ifr]; ia++) {
            const double *r = xyz + 3 * (a0 + ia);
            double xd[3] = {x[0] - r[0], x[1] - r[1], x[2] - r[2]};
            double dist = orc_norm2(xd, 3);                       /* :539 */
            int iz = z[a0 + ia];
            if (iz < 1 || iz > 118 || !agrid[iz] || !agrid[iz]->isinit) continue; /* :541 */
            if (dist > agrid[iz]->rmax) continue;                 /* :542 */
            dist = fmax(dist, agrid[iz]->r[0]);                   /* :543 */
            dist = fmax(dist, 1e-14);                             /* :544 */
            double rrho, rrho1, rrho2;
            orc_grid1_interp(agrid[iz], dist, &rrho, &rrho1, &rrho2);
            rrho = fmax(rrho, 0.0);                               /* :546 */
            sum += rrho;
          }
          rhofrag[(size_t)ifr * npt + idx] = sum;
          a0 += nat[ifr];
        }
      }
}

/* nciplot cutoff pass :569-604.  nfrag > 0: rhofrag [nfrag][npts] and rhoat are required (:579-582). */
long orc_nci_cutoffs_frag(const int *nstep, double *crho, double *cgrad, const double *rhoat, int nfrag,
                          const double *rhofrag, double rhoparam, double rhoparam2, double rhocut, double dimcut,
                          double rhoplot, const double *srhorange, int onlyneg, double rho_void, double omega,
                          double *sums, double *dat, long maxdat) {
  double sumq = 0, sumqp = 0, sumv = 0;
  long ndat = 0;
  size_t ntot = (size_t)nstep[0] * nstep[1] * nstep[2];
  for (size_t idx = 0; idx < ntot; idx++) { /* (i,j,k) order with k fastest == linear order of the arrays */
    int inter = 1;
    if (nfrag > 0) {                                              /* :579-582 */
      double s = 0.0;
      for (int ifr = 0; ifr < nfrag; ifr++) s += rhofrag[(size_t)ifr * ntot + idx];
      inter = s >= rhoparam2 * rhoat[idx];
      for (int ifr = 0; ifr < nfrag; ifr++) inter = inter && (rhofrag[(size_t)ifr * ntot + idx] <= s * rhoparam);
    }
    if (rho_void > 0) {
      inter = inter && (rhoat[idx] < rho_void);
      if (rhoat[idx] < rho_void) {
        sumq += fabs(crho[idx]) / 100.0;
        sumqp += rhoat[idx];
        sumv += 1;
      }
    }
    if (fabs(crho[idx]) < rhocut * 100.0 && cgrad[idx] < dimcut && inter) {
      if (dat && ndat < maxdat) { dat[2 * ndat] = crho[idx] / 100.0; dat[2 * ndat + 1] = cgrad[idx]; }
      ndat++;
    }
    if (fabs(crho[idx]) > rhoplot * 100.0 || !inter) cgrad[idx] = 100.0;
    if (crho[idx] < srhorange[0] * 100.0 || crho[idx] > srhorange[1] * 100.0 || !inter) cgrad[idx] = 100.0;
    if (crho[idx] > 0 && onlyneg) cgrad[idx] = 100.0;
  }
  double nt = (double)nstep[0] * nstep[1] * nstep[2];
  sums[0] = sumq * omega / nt;
  sums[1] = sumqp * omega / nt;
  sums[2] = sumv * omega / nt;
  return ndat;
}

long orc_nci_cutoffs(const int *nstep, double *crho, double *cgrad, const double *rhoat, double rhocut, double dimcut,
                     double rhoplot, const double *srhorange, int onlyneg, double rho_void, double omega, double *sums,
                     double *dat, long maxdat) {
  return orc_nci_cutoffs_frag(nstep, crho, cgrad, rhoat, 0, NULL, 0.95, 0.75, rhocut, dimcut, rhoplot, srhorange, onlyneg,
                              rho_void, omega, sums, dat, maxdat);
}

/* promolecular_grid crystalmod@proc.f90:5118-5158 */
void orc_promolecular_grid(const orc_environ *e, orc_grid1 *const *agrid, const int *n, const unsigned char *fragmask,
                           double *out, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic)
  for (int k = 0; k < n[2]; k++)
    for (int j = 0; j < n[1]; j++)
      for (int i = 0; i < n[0]; i++) {
        /* x = (i-1)*xdelta with xdelta = 1/n (:5137-5148); icrd_crys -> reduced -> fold -> Cartesian (y2z_center) */
        double x[3] = {i * (1.0 / n[0]), j * (1.0 / n[1]), k * (1.0 / n[2])}, xr[3], xc[3], rho, g[3], h[9];
        orc_matvec(e->m_x2xr, x, xr);
        if (!e->ismolecule)
          for (int d = 0; d < 3; d++) xr[d] = xr[d] - floor(xr[d]);
        orc_matvec(e->m_xr2c, xr, xc);
        orc_promolecular(e, agrid, xc, 0, fragmask, &rho, g, h);   /* fr: :5149 */
        out[(size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * k)] = rho;
      }
}

void orc_mo_points(const orc_field *f, long npts, const double *pts, int imo, int periodic, double *phi, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (long p = 0; p < npts; p++) {
    double wc[3], wcr[3];
    grd_points(f->env, pts + 3 * p, periodic, wc, wcr);
    phi[p] = orc_calculate_mo(f->wfn, f->env, wcr, imo);
  }
}

void orc_grd_points(const orc_field *f, long npts, const double *pts, int nder, int periodic, double *res, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (long p = 0; p < npts; p++) orc_grd(f, pts + 3 * p, nder, periodic, res + (size_t)p * ORC_RES_LEN);
}

void orc_nact_points(const orc_wfn *f, const orc_environ *env, long npts, const double *pts, long *nact, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (long p = 0; p < npts; p++) nact[p] = orc_nact(f, env, pts + 3 * p);
}
