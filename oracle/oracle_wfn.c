/*
 * oracle_wfn.c -- restatement of molwfn%rho2 / calculate_mo_gto / calculate_edf / complete_struct.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Citations: /root/reference/src/wfn_private@proc.F90.
 */
#include "oracle.h"
#include "oracle_util.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* primitive type -> (lx,ly,lz), :2695-2705 */
static const int li[56][3] = {
    {0, 0, 0},
    {1, 0, 0}, {0, 1, 0}, {0, 0, 1},
    {2, 0, 0}, {0, 2, 0}, {0, 0, 2}, {1, 1, 0}, {1, 0, 1}, {0, 1, 1},
    {3, 0, 0}, {0, 3, 0}, {0, 0, 3}, {2, 1, 0}, {2, 0, 1}, {0, 2, 1}, {1, 2, 0}, {1, 0, 2}, {0, 1, 2}, {1, 1, 1},
    {4, 0, 0}, {0, 4, 0}, {0, 0, 4}, {3, 1, 0}, {3, 0, 1}, {1, 3, 0}, {0, 3, 1}, {1, 0, 3},
    {0, 1, 3}, {2, 2, 0}, {2, 0, 2}, {0, 2, 2}, {2, 1, 1}, {1, 2, 1}, {1, 1, 2},
    {0, 0, 5}, {0, 1, 4}, {0, 2, 3}, {0, 3, 2}, {0, 4, 1}, {0, 5, 0}, {1, 0, 4}, {1, 1, 3},
    {1, 2, 2}, {1, 3, 1}, {1, 4, 0}, {2, 0, 3}, {2, 1, 2}, {2, 2, 1}, {2, 3, 0}, {3, 0, 2},
    {3, 1, 1}, {3, 2, 0}, {4, 0, 1}, {4, 1, 0}, {5, 0, 0}};
/* :3084-3086 */
static const int limax[56] = {0, 1, 1, 1, 2, 2, 2, 1, 1, 1, 3, 3, 3, 2, 2, 2, 2, 2, 2, 1, 4, 4, 4, 3, 3, 3, 3, 3,
                              3, 2, 2, 2, 2, 2, 2, 5, 4, 3, 3, 4, 5, 4, 3, 2, 3, 4, 3, 2, 2, 3, 3, 3, 3, 4, 4, 5};
static const double rprim_thres = 1e-12; /* :145 */

typedef struct { int key; int idx; } ikey;
static int cmp_ikey(const void *a, const void *b) {
  const ikey *x = a, *y = b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->idx - y->idx; /* stable (reference qcksort is not): only the summation order within an atom differs */
}

static void order_by_center(int n, const int *icenter, int natom, int *icord, int *iprilo, int *iprihi) {
  ikey *k = malloc(sizeof(ikey) * (n > 0 ? n : 1));
  for (int i = 0; i < n; i++) { k[i].key = icenter[i]; k[i].idx = i + 1; }
  qsort(k, n, sizeof(ikey), cmp_ikey);
  for (int i = 0; i < natom; i++) { iprilo[i] = 0; iprihi[i] = -1; }
  for (int i = 0; i < n; i++) {
    icord[i] = k[i].idx;
    int iat = icenter[icord[i] - 1];
    if (iprilo[iat - 1] == 0) iprilo[iat - 1] = i + 1;
    iprihi[iat - 1] = i + 1;
  }
  free(k);
}

static void *dup_mem(const void *p, size_t n) {
  if (!p || n == 0) return NULL;
  void *q = malloc(n);
  memcpy(q, p, n);
  return q;
}

/* holds the raw arrays + complete_struct :3077-3184 (molecular environment: cidx(i) = i) */
orc_wfn *orc_wfn_create(int natom, int npri, const int *icenter, const int *itype, const double *e, int nmoocc,
                        int nmoall, const double *cmo, const double *occ, int wfntyp, int nalpha, int nedf,
                        const int *icenter_edf, const int *itype_edf, const double *e_edf, const double *c_edf) {
  orc_wfn *f = calloc(1, sizeof(orc_wfn));
  f->natom = natom; f->npri = npri; f->nmoocc = nmoocc; f->nmoall = nmoall; f->wfntyp = wfntyp; f->nalpha = nalpha;
  f->nedf = nedf;
  f->icenter = dup_mem(icenter, sizeof(int) * npri);
  f->itype = dup_mem(itype, sizeof(int) * npri);
  f->e = dup_mem(e, sizeof(double) * npri);
  f->occ = dup_mem(occ, sizeof(double) * nmoocc);
  f->cmo = dup_mem(cmo, sizeof(double) * (size_t)nmoall * npri);
  if (nedf > 0) {
    f->icenter_edf = dup_mem(icenter_edf, sizeof(int) * nedf);
    f->itype_edf = dup_mem(itype_edf, sizeof(int) * nedf);
    f->e_edf = dup_mem(e_edf, sizeof(double) * nedf);
    f->c_edf = dup_mem(c_edf, sizeof(double) * nedf);
  }
  f->spcutoff = calloc(natom, sizeof(double));
  f->lmax = calloc(natom, sizeof(int));
  f->dran = malloc(sizeof(double) * npri);
  for (int i = 0; i < npri; i++) {                             /* :3101-3110 */
    f->dran[i] = sqrt(-log(rprim_thres) / e[i]);
    int c = icenter[i] - 1;
    f->spcutoff[c] = fmax(f->spcutoff[c], f->dran[i]);
    if (limax[itype[i] - 1] > f->lmax[c]) f->lmax[c] = limax[itype[i] - 1];
  }
  f->icord = malloc(sizeof(int) * npri);
  f->iprilo = malloc(sizeof(int) * natom);
  f->iprihi = malloc(sizeof(int) * natom);
  order_by_center(npri, icenter, natom, f->icord, f->iprilo, f->iprihi); /* :3112-3130 */
  if (nedf > 0) {                                              /* :3133-3161 */
    f->dran_edf = malloc(sizeof(double) * nedf);
    for (int i = 0; i < nedf; i++) {
      f->dran_edf[i] = sqrt(-log(rprim_thres) / e_edf[i]);
      int c = icenter_edf[i] - 1;
      f->spcutoff[c] = fmax(f->spcutoff[c], f->dran_edf[i]);
    }
    f->icord_edf = malloc(sizeof(int) * nedf);
    f->iprilo_edf = malloc(sizeof(int) * natom);
    f->iprihi_edf = malloc(sizeof(int) * natom);
    order_by_center(nedf, icenter_edf, natom, f->icord_edf, f->iprilo_edf, f->iprihi_edf);
  }
  f->globalcutoff = 0;
  for (int i = 0; i < natom; i++) f->globalcutoff = fmax(f->globalcutoff, f->spcutoff[i]); /* :3164 */
  return f;
}

void orc_wfn_free(orc_wfn *f) {
  if (!f) return;
  free(f->icenter); free(f->itype); free(f->e); free(f->occ); free(f->cmo);
  free(f->icenter_edf); free(f->itype_edf); free(f->e_edf); free(f->c_edf);
  free(f->dran); free(f->spcutoff); free(f->dran_edf); free(f->lmax); free(f->icord); free(f->iprilo);
  free(f->iprihi); free(f->icord_edf); free(f->iprilo_edf); free(f->iprihi_edf);
  free(f);
}

void orc_wfn_tables(const orc_wfn *f, double *dran, double *spcutoff, int *lmax, int *icord, int *iprilo, int *iprihi) {
  memcpy(dran, f->dran, sizeof(double) * f->npri);
  memcpy(spcutoff, f->spcutoff, sizeof(double) * f->natom);
  memcpy(lmax, f->lmax, sizeof(int) * f->natom);
  memcpy(icord, f->icord, sizeof(int) * f->npri);
  memcpy(iprilo, f->iprilo, sizeof(int) * f->natom);
  memcpy(iprihi, f->iprihi, sizeof(int) * f->natom);
}

/* powers table xl(ix,0:2,0:lmax) :2717-2742.  Entries above lmax are left untouched by the reference; a primitive of
   the atom never reads them because lmax is the atom's maximum l. */
static void powers(const double *xx, int lmax, double xl[3][3][6]) {
  for (int ix = 0; ix < 3; ix++) {
    xl[ix][0][0] = 1.0; xl[ix][1][0] = 0.0; xl[ix][2][0] = 0.0;
    if (lmax < 1) continue;
    xl[ix][0][1] = xx[ix]; xl[ix][1][1] = 1.0; xl[ix][2][1] = 0.0;
    if (lmax < 2) continue;
    xl[ix][0][2] = xl[ix][0][1] * xx[ix]; xl[ix][1][2] = 2.0 * xl[ix][0][1]; xl[ix][2][2] = 2.0;
    if (lmax < 3) continue;
    xl[ix][0][3] = xl[ix][0][2] * xx[ix]; xl[ix][1][3] = 3.0 * xl[ix][0][2]; xl[ix][2][3] = 6.0 * xl[ix][0][1];
    if (lmax < 4) continue;
    xl[ix][0][4] = xl[ix][0][3] * xx[ix]; xl[ix][1][4] = 4.0 * xl[ix][0][3]; xl[ix][2][4] = 12.0 * xl[ix][0][2];
    if (lmax < 5) continue;
    xl[ix][0][5] = xl[ix][0][4] * xx[ix]; xl[ix][1][5] = 5.0 * xl[ix][0][4]; xl[ix][2][5] = 20.0 * xl[ix][0][3];
  }
}

/* x**n with integer n as gfortran evaluates it (libgcc __powidf2, square-and-multiply) */
static inline double ipow(double x, int n) { return __builtin_powi(x, n); }

/* chi(1:10) for one primitive :2755-2768 (shared by calculate_mo_gto and calculate_edf :2858-2869) */
static void prim_chi(const double *xx, double xl[3][3][6], const int *l, double al, double ex, int nder, double *chi) {
  double x0 = xl[0][0][l[0]], y0 = xl[1][0][l[1]], z0 = xl[2][0][l[2]];
  chi[0] = x0 * y0 * z0 * ex;
  if (nder > 0) {
    double dx = xl[0][1][l[0]] - 2 * al * ipow(xx[0], l[0] + 1);
    double dy = xl[1][1][l[1]] - 2 * al * ipow(xx[1], l[1] + 1);
    double dz = xl[2][1][l[2]] - 2 * al * ipow(xx[2], l[2] + 1);
    chi[1] = dx * y0 * z0 * ex;
    chi[2] = dy * x0 * z0 * ex;
    chi[3] = dz * x0 * y0 * ex;
    if (nder > 1) {
      chi[4] = (xl[0][2][l[0]] - 2 * al * (2 * l[0] + 1) * x0 + 4 * al * al * ipow(xx[0], l[0] + 2)) * y0 * z0 * ex;
      chi[5] = (xl[1][2][l[1]] - 2 * al * (2 * l[1] + 1) * y0 + 4 * al * al * ipow(xx[1], l[1] + 2)) * z0 * x0 * ex;
      chi[6] = (xl[2][2][l[2]] - 2 * al * (2 * l[2] + 1) * z0 + 4 * al * al * ipow(xx[2], l[2] + 2)) * x0 * y0 * ex;
      chi[7] = dx * dy * z0 * ex;
      chi[8] = dx * dz * y0 * ex;
      chi[9] = dz * dy * x0 * ex;
    }
  }
}

static const int imax_nder[3] = {1, 4, 10};

/* calculate_mo_gto :2673-2781. phi is (nmoocc x imax) column-major. */
static void calculate_mo_gto(const orc_wfn *f, const orc_environ *env, const double *xpos, double *phi, int nder,
                             int nenv, const int *eid, const double *dist) {
  int imax = imax_nder[nder];
  int nmo = f->nmoocc;
  memset(phi, 0, sizeof(double) * (size_t)nmo * imax);
  double xl[3][3][6];
  for (int i = 0; i < nenv; i++) {
    int iat = env->cidx[eid[i]]; /* 1-based */
    double d2 = dist[i] * dist[i];
    /* xx = xpos - f%env%at(iat)%r : note it is the cell atom iat, not the environment atom eid(i) (:2715) */
    const double *ra = env->r + 3 * (iat - 1);
    double xx[3] = {xpos[0] - ra[0], xpos[1] - ra[1], xpos[2] - ra[2]};
    powers(xx, f->lmax[iat - 1], xl);
    for (int j = f->iprilo[iat - 1]; j <= f->iprihi[iat - 1]; j++) {
      int ipri = f->icord[j - 1] - 1;
      if (dist[i] > f->dran[ipri]) continue;                    /* :2747 */
      double al = f->e[ipri];
      double ex = exp(-al * d2);
      double chi[10];
      prim_chi(xx, xl, li[f->itype[ipri] - 1], al, ex, nder, chi);
      const double *c = f->cmo + (size_t)ipri * f->nmoall;
      for (int ix = 0; ix < imax; ix++) {
        double ch = chi[ix];
        double *ph = phi + (size_t)ix * nmo;
        for (int imo = 0; imo < nmo; imo++) ph[imo] += c[imo] * ch; /* :2771-2776 */
      }
    }
  }
}

/* calculate_edf :2783-2873 */
static void calculate_edf(const orc_wfn *f, const orc_environ *env, const double *xpos, double *rhoc, double *gradc,
                          double *hc, int nder, int nenv, const int *eid, const double *dist) {
  *rhoc = 0;
  for (int i = 0; i < 3; i++) gradc[i] = 0;
  for (int i = 0; i < 6; i++) hc[i] = 0;
  double xl[3][3][6];
  for (int i = 0; i < nenv; i++) {
    int iat = env->cidx[eid[i]];
    double d2 = dist[i] * dist[i];
    const double *ra = env->r + 3 * (iat - 1);
    double xx[3] = {xpos[0] - ra[0], xpos[1] - ra[1], xpos[2] - ra[2]};
    /* the EDF routine also builds xl only up to the valence lmax of the atom (:2825); EDF types are s in practice */
    int lm = f->lmax[iat - 1];
    for (int j = f->iprilo_edf[iat - 1]; j <= f->iprihi_edf[iat - 1]; j++) {
      int t = f->itype_edf[f->icord_edf[j - 1] - 1] - 1;
      if (limax[t] > lm) lm = limax[t]; /* avoid reading uninitialised table entries (undefined in the reference) */
    }
    powers(xx, lm, xl);
    for (int j = f->iprilo_edf[iat - 1]; j <= f->iprihi_edf[iat - 1]; j++) {
      int ipri = f->icord_edf[j - 1] - 1;
      if (dist[i] > f->dran_edf[ipri]) continue;
      double al = f->e_edf[ipri];
      double ex = exp(-al * d2);
      double exc = f->c_edf[ipri] * ex;
      const int *l = li[f->itype_edf[ipri] - 1];
      double x0 = xl[0][0][l[0]], y0 = xl[1][0][l[1]], z0 = xl[2][0][l[2]];
      *rhoc += f->c_edf[ipri] * x0 * y0 * z0 * ex;
      if (nder < 1) continue;
      double dx = xl[0][1][l[0]] - 2 * al * ipow(xx[0], l[0] + 1);
      double dy = xl[1][1][l[1]] - 2 * al * ipow(xx[1], l[1] + 1);
      double dz = xl[2][1][l[2]] - 2 * al * ipow(xx[2], l[2] + 1);
      gradc[0] += dx * y0 * z0 * exc;
      gradc[1] += dy * x0 * z0 * exc;
      gradc[2] += dz * x0 * y0 * exc;
      if (nder < 2) continue;
      hc[0] += (xl[0][2][l[0]] - 2 * al * (2 * l[0] + 1) * x0 + 4 * al * al * ipow(xx[0], l[0] + 2)) * y0 * z0 * exc;
      hc[1] += (xl[1][2][l[1]] - 2 * al * (2 * l[1] + 1) * y0 + 4 * al * al * ipow(xx[1], l[1] + 2)) * z0 * x0 * exc;
      hc[2] += (xl[2][2][l[2]] - 2 * al * (2 * l[2] + 1) * z0 + 4 * al * al * ipow(xx[2], l[2] + 2)) * x0 * y0 * exc;
      hc[3] += dx * dy * z0 * exc;
      hc[4] += dx * dz * y0 * exc;
      hc[5] += dz * dy * x0 * exc;
    }
  }
}

/* calculate_mo :2447-2553 for a numeric selector: phi_imo(xpos) through calculate_mo_gto(imo,imo,imo,nder=0); imo is
   1-based and may address a virtual orbital (<= nmoall) */
double orc_calculate_mo(const orc_wfn *f, const orc_environ *env, const double *xpos, int imo) {
  double phi = 0.0;
  if (imo < 1 || imo > f->nmoall) return 0.0 / 0.0;             /* :2531-2532 is a fatal error in the reference */
  int *eid = malloc(sizeof(int) * env->n);
  double *dist = malloc(sizeof(double) * env->n);
  int nenv, lvec[3];
  int ierr = orc_list_near_atoms(env, xpos, 0, f->spcutoff, &nenv, eid, dist, lvec); /* :2535 */
  if (!ierr) {
    double xl[3][3][6];
    for (int i = 0; i < nenv; i++) {
      int iat = env->cidx[eid[i]];
      double d2 = dist[i] * dist[i];
      const double *ra = env->r + 3 * (iat - 1);
      double xx[3] = {xpos[0] - ra[0], xpos[1] - ra[1], xpos[2] - ra[2]};
      powers(xx, f->lmax[iat - 1], xl);
      for (int j = f->iprilo[iat - 1]; j <= f->iprihi[iat - 1]; j++) {
        int ipri = f->icord[j - 1] - 1;
        if (dist[i] > f->dran[ipri]) continue;                  /* :2747 */
        double al = f->e[ipri], chi[10];
        prim_chi(xx, xl, li[f->itype[ipri] - 1], al, exp(-al * d2), 0, chi);
        phi += f->cmo[(size_t)ipri * f->nmoall + (imo - 1)] * chi[0];
      }
    }
  }
  free(eid);
  free(dist);
  return phi;
}

long orc_nact(const orc_wfn *f, const orc_environ *env, const double *xpos) {
  int *eid = malloc(sizeof(int) * env->n);
  double *dist = malloc(sizeof(double) * env->n);
  int nenv, lvec[3];
  long n = 0;
  int ierr = orc_list_near_atoms(env, xpos, 0, f->spcutoff, &nenv, eid, dist, lvec);
  if (!ierr) {
    for (int i = 0; i < nenv; i++) {
      int iat = env->cidx[eid[i]];
      for (int j = f->iprilo[iat - 1]; j <= f->iprihi[iat - 1]; j++)
        if (!(dist[i] > f->dran[f->icord[j - 1] - 1])) n++;
    }
  }
  free(eid);
  free(dist);
  return n;
}

/* rho2 :2032-2228 */
void orc_rho2(const orc_wfn *f, const orc_environ *env, const double *xpos, int nder, double *out) {
  double *rho = out, *rhoval = out + 3, *grad = out + 6, *gradval = out + 15, *h = out + 24, *hval = out + 51;
  double *gkin = out + 78, *vir = out + 81, *stress = out + 82;
  memset(out, 0, sizeof(double) * 91);
  double hh[3][6], hhval[3][6];
  memset(hh, 0, sizeof(hh));
  memset(hhval, 0, sizeof(hhval));

  int *eid = malloc(sizeof(int) * env->n);
  double *dist = malloc(sizeof(double) * env->n);
  int nenv, lvec[3];
  int ierr = orc_list_near_atoms(env, xpos, 0, f->spcutoff, &nenv, eid, dist, lvec); /* :2070 */
  if (ierr > 0) { free(eid); free(dist); return; }             /* :2071 */

  int imax = imax_nder[nder], nmo = f->nmoocc;
  double *phi = malloc(sizeof(double) * (size_t)nmo * imax);
  calculate_mo_gto(f, env, xpos, phi, nder, nenv, eid, dist);
#define PHI(imo, ix) phi[(size_t)((ix)-1) * nmo + (imo)]
  int uhf = (f->wfntyp == 1);
  for (int imo = 0; imo < nmo; imo++) {                         /* :2091-2147 */
    double aocc = f->occ[imo];
    int sp = uhf ? ((imo + 1 <= f->nalpha) ? 1 : 2) : 0;
    double rhosum = aocc * PHI(imo, 1) * PHI(imo, 1);
    rho[0] += rhosum;
    if (uhf) rho[sp] += rhosum;
    if (nder > 0) {
      double gkinsum = aocc * (PHI(imo, 2) * PHI(imo, 2) + PHI(imo, 3) * PHI(imo, 3) + PHI(imo, 4) * PHI(imo, 4));
      for (int k = 0; k < 3; k++) {
        double g = 2 * aocc * PHI(imo, 1) * PHI(imo, 2 + k);
        grad[k] += g;
        if (uhf) grad[3 * sp + k] += g;
      }
      gkin[0] += gkinsum;
      if (uhf) gkin[sp] += gkinsum;
    }
    if (nder > 1) {
      double hs[6];
      for (int k = 0; k < 3; k++)
        hs[k] = 2 * aocc * (PHI(imo, 1) * PHI(imo, 5 + k) + PHI(imo, 2 + k) * PHI(imo, 2 + k));
      hs[3] = 2 * aocc * (PHI(imo, 1) * PHI(imo, 8) + PHI(imo, 2) * PHI(imo, 3));
      hs[4] = 2 * aocc * (PHI(imo, 1) * PHI(imo, 9) + PHI(imo, 2) * PHI(imo, 4));
      hs[5] = 2 * aocc * (PHI(imo, 1) * PHI(imo, 10) + PHI(imo, 3) * PHI(imo, 4));
      for (int k = 0; k < 6; k++) {
        hh[0][k] += hs[k];
        if (uhf) hh[sp][k] += hs[k];
      }
      /* stress(i,j) at [i + 3*j] */
      stress[0 + 3 * 0] += aocc * (PHI(imo, 1) * PHI(imo, 5) - PHI(imo, 2) * PHI(imo, 2));
      stress[0 + 3 * 1] += aocc * (PHI(imo, 1) * PHI(imo, 8) - PHI(imo, 2) * PHI(imo, 3));
      stress[0 + 3 * 2] += aocc * (PHI(imo, 1) * PHI(imo, 9) - PHI(imo, 2) * PHI(imo, 4));
      stress[1 + 3 * 1] += aocc * (PHI(imo, 1) * PHI(imo, 6) - PHI(imo, 3) * PHI(imo, 3));
      stress[1 + 3 * 2] += aocc * (PHI(imo, 1) * PHI(imo, 10) - PHI(imo, 3) * PHI(imo, 4));
      stress[2 + 3 * 2] += aocc * (PHI(imo, 1) * PHI(imo, 7) - PHI(imo, 4) * PHI(imo, 4));
    }
  }
#undef PHI
  free(phi);

  double rhoc = 0, gradc[3] = {0, 0, 0}, hhc[6] = {0, 0, 0, 0, 0, 0};
  if (f->nedf > 0) calculate_edf(f, env, xpos, &rhoc, gradc, hhc, nder, nenv, eid, dist); /* :2151 */
  free(eid);
  free(dist);

  if (!uhf) {                                                   /* :2159-2183 */
    rhoval[0] = rho[0]; rhoval[1] = 0.5 * rhoval[0]; rhoval[2] = rhoval[1];
    rho[0] = rho[0] + rhoc; rho[1] = 0.5 * rho[0]; rho[2] = rho[1];
    if (nder > 0) {
      for (int k = 0; k < 3; k++) {
        gradval[k] = grad[k]; gradval[3 + k] = 0.5 * gradval[k]; gradval[6 + k] = gradval[3 + k];
        grad[k] = grad[k] + gradc[k]; grad[3 + k] = 0.5 * grad[k]; grad[6 + k] = grad[3 + k];
      }
      gkin[1] = gkin[0] / 2.0; gkin[2] = gkin[1];
    }
    if (nder > 1)
      for (int k = 0; k < 6; k++) {
        hhval[0][k] = hh[0][k]; hhval[1][k] = 0.5 * hhval[0][k]; hhval[2][k] = hhval[1][k];
        hh[0][k] = hh[0][k] + hhc[k]; hh[1][k] = 0.5 * hh[0][k]; hh[2][k] = hh[1][k];
      }
  } else {                                                      /* :2184-2200 */
    for (int k = 0; k < 3; k++) rhoval[k] = rho[k];
    rho[0] += rhoc; rho[1] += 0.5 * rhoc; rho[2] += 0.5 * rhoc;
    if (nder > 0)
      for (int s = 0; s < 3; s++)
        for (int k = 0; k < 3; k++) {
          gradval[3 * s + k] = grad[3 * s + k];
          grad[3 * s + k] += (s == 0 ? 1.0 : 0.5) * gradc[k];
        }
    if (nder > 1)
      for (int s = 0; s < 3; s++)
        for (int k = 0; k < 6; k++) {
          hhval[s][k] = hh[s][k];
          hh[s][k] += (s == 0 ? 1.0 : 0.5) * hhc[k];
        }
  }
  /* :2202-2226 */
  stress[1 + 3 * 0] = stress[0 + 3 * 1];
  stress[2 + 3 * 0] = stress[0 + 3 * 2];
  stress[2 + 3 * 1] = stress[1 + 3 * 2];
  for (int k = 0; k < 9; k++) stress[k] *= 0.5;
  *vir = stress[0] + stress[4] + stress[8];
  for (int j = 0; j < 3; j++) {
    gkin[j] = 0.5 * gkin[j];
    double *hj = h + 9 * j, *hvj = hval + 9 * j;
    for (int i = 0; i < 3; i++) { hj[i + 3 * i] = hh[j][i]; hvj[i + 3 * i] = hhval[j][i]; }
    hj[0 + 3 * 1] = hj[1 + 3 * 0] = hh[j][3];
    hj[0 + 3 * 2] = hj[2 + 3 * 0] = hh[j][4];
    hj[1 + 3 * 2] = hj[2 + 3 * 1] = hh[j][5];
    hvj[0 + 3 * 1] = hvj[1 + 3 * 0] = hhval[j][3];
    hvj[0 + 3 * 2] = hvj[2 + 3 * 0] = hhval[j][4];
    hvj[1 + 3 * 2] = hvj[2 + 3 * 1] = hhval[j][5];
  }
}
