/*
 * oracle_env.c -- restatement of critic2's `environ` (atom environment + region search).
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Citations: /root/reference/src/environmod@proc.f90.
 */
#include "oracle.h"
#include "oracle_util.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* cutrad: src/global.f90:51-93 (cutoff radius for 1d-12 atomic densities, bohr) */
static const double cutrad_tab[123] = {
    2.149886192475e+01, 1.169139170668e+01, 3.430831385801e+01, 2.502075396007e+01, 2.801001395722e+01,
    2.167675592180e+01, 1.749805708313e+01, 1.465173060207e+01, 1.263885024136e+01, 1.110521599057e+01,
    3.523728402162e+01, 2.763528367271e+01, 3.395549316507e+01, 2.847261278601e+01, 2.487715217494e+01,
    2.222930087269e+01, 2.022231415676e+01, 1.857150175607e+01, 3.884428729523e+01, 3.144587224767e+01,
    2.970981796151e+01, 2.864438811442e+01, 2.784088946336e+01, 2.925194799711e+01, 2.660566532177e+01,
    2.609916866690e+01, 2.558901439004e+01, 2.517359152887e+01, 2.691554955610e+01, 2.435659411320e+01,
    3.467478603212e+01, 2.914443825602e+01, 2.572575006996e+01, 2.323452863278e+01, 2.134146595122e+01,
    1.981582897591e+01, 3.976877622180e+01, 3.266858263171e+01, 3.027851405458e+01, 2.899491720657e+01,
    2.967865003580e+01, 2.914637014504e+01, 2.697201600611e+01, 2.844039136970e+01, 2.814409350112e+01,
    1.659926809140e+01, 2.771163603049e+01, 2.519886588880e+01, 3.538116802480e+01, 3.026454800251e+01,
    2.695514982633e+01, 2.460024202780e+01, 2.277601677390e+01, 2.130658017554e+01, 4.137458546886e+01,
    3.442036204804e+01, 3.242561614450e+01, 3.212250868201e+01, 3.306457792690e+01, 3.284026775197e+01,
    3.262654222620e+01, 3.242292112974e+01, 3.222895504883e+01, 3.128431696800e+01, 3.180465714311e+01,
    3.157435555252e+01, 3.135291924508e+01, 3.120231512615e+01, 3.099709515455e+01, 3.079969409503e+01,
    3.160515129459e+01, 2.709458469010e+01, 2.614193052742e+01, 2.548104664032e+01, 2.489113924347e+01,
    2.441668377017e+01, 2.405143298004e+01, 2.466268008529e+01, 2.439924398342e+01, 2.305709117567e+01,
    3.643576493190e+01, 3.110226831614e+01, 2.780342993946e+01, 2.541102668192e+01, 2.360240806573e+01,
    2.210165966674e+01, 4.053200388132e+01, 3.407838067822e+01, 3.585071927373e+01, 3.175945034367e+01,
    3.478340806986e+01, 3.489038964505e+01, 3.514212336660e+01, 3.120895952111e+01, 37., 37., 37., 37., 37.,
    37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37., 37.,
    0., 0., 0., 0., 0.};

double orc_cutrad(int z) { return (z >= 1 && z <= 123) ? cutrad_tab[z - 1] : 0.0; }

static double atmcov_bohr(int z) {
  /* src/param.F90:289-303 atmcov0 (angstrom) / bohrtoa */
  static const double a[118] = {
      0.31, 0.28, 1.28, 0.96, 0.84, 0.73, 0.71, 0.66, 0.57, 0.58, 1.66, 1.41, 1.21, 1.11, 1.07, 1.05, 1.02, 1.06, 2.03, 1.76,
      1.70, 1.60, 1.53, 1.39, 1.50, 1.42, 1.38, 1.24, 1.32, 1.22, 1.22, 1.20, 1.19, 1.20, 1.20, 1.16, 2.20, 1.95, 1.90, 1.75,
      1.64, 1.54, 1.47, 1.46, 1.42, 1.39, 1.45, 1.44, 1.42, 1.39, 1.39, 1.38, 1.39, 1.40, 2.44, 2.15, 2.07, 2.04, 2.03, 2.01,
      1.99, 1.98, 1.98, 1.96, 1.94, 1.92, 1.92, 1.89, 1.90, 1.87, 1.87, 1.75, 1.70, 1.62, 1.51, 1.44, 1.41, 1.36, 1.36, 1.32,
      1.45, 1.46, 1.48, 1.40, 1.50, 1.50, 2.60, 2.21, 2.15, 2.06, 2.00, 1.96, 1.90, 1.87, 1.80, 1.69, 1.50, 1.50, 1.50, 1.50,
      1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50, 1.50};
  const double bohrtoa = 0.52917720859;
  if (z >= 1 && z <= 118) return a[z - 1] / bohrtoa;
  return 0.0;
}

static int p2i(const orc_environ *e, const int *xx) { /* :457-467 */
  int ix[3];
  for (int d = 0; d < 3; d++) {
    int v = xx[d];
    if (v < e->nmin[d]) v = e->nmin[d];
    if (v > e->nmax[d]) v = e->nmax[d];
    ix[d] = v - e->nmin[d];
  }
  return 1 + ix[0] + e->nreg[0] * (ix[1] + e->nreg[1] * ix[2]);
}
static void c2p(const orc_environ *e, const double *xx, int *res) { /* :447-454 */
  for (int d = 0; d < 3; d++) res[d] = (int)floor((xx[d] - e->x0[d]) / e->boxsize);
}
static int c2i(const orc_environ *e, const double *xx) { /* :470-480 */
  int p[3];
  c2p(e, xx, p);
  return p2i(e, p);
}

/* y2z_center from Cartesian to Cartesian (:426-444): fold to the main reduced cell for crystals */
void orc_center_cart(const orc_environ *e, double *xx, int *lvec) {
  if (e->ismolecule) {
    if (lvec) lvec[0] = lvec[1] = lvec[2] = 0;
    return;
  }
  double xr[3];
  orc_matvec(e->m_c2xr, xx, xr);
  for (int d = 0; d < 3; d++) {
    double fl = floor(xr[d]);
    if (lvec) lvec[d] = (int)fl;
    xr[d] = xr[d] - fl;
  }
  orc_matvec(e->m_xr2c, xr, xx);
}

typedef struct { double key; int idx; } dkey;
typedef struct { int key; int idx; } ikey;
static int cmp_dkey(const void *a, const void *b) {
  const dkey *x = a, *y = b;
  if (x->key < y->key) return -1;
  if (x->key > y->key) return 1;
  return x->idx - y->idx; /* stable; the reference's qcksort is not -- only traversal order differs */
}
static int cmp_ikey(const void *a, const void *b) {
  const ikey *x = a, *y = b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->idx - y->idx;
}

static void env_push(orc_environ *e, int *cap, const double *x, const double *r, int is, int cidx, const int *lvec) {
  if (e->n >= *cap) {
    *cap = 2 * (*cap) + 16;
    e->x = realloc(e->x, sizeof(double) * 3 * (*cap));
    e->r = realloc(e->r, sizeof(double) * 3 * (*cap));
    e->is = realloc(e->is, sizeof(int) * (*cap));
    e->cidx = realloc(e->cidx, sizeof(int) * (*cap));
    e->lvec = realloc(e->lvec, sizeof(int) * 3 * (*cap));
  }
  int i = e->n++;
  for (int d = 0; d < 3; d++) { e->x[3 * i + d] = x[d]; e->r[3 * i + d] = r[d]; e->lvec[3 * i + d] = lvec[d]; }
  e->is[i] = is;
  e->cidx[i] = cidx;
}

/* environ_build :65-331 */
orc_environ *orc_environ_build(int ismol, int nspc, const int *zspc, int ncell, const double *xcryst, const int *is,
                               const double *m_x2c, const double *m_x2xr, const double *m_xr2c, double dmax0_in) {
  orc_environ *e = calloc(1, sizeof(orc_environ));
  memcpy(e->m_xr2c, m_xr2c, 72);
  orc_matinv(m_xr2c, e->m_c2xr);
  memcpy(e->m_x2xr, m_x2xr, 72);
  orc_matinv(m_x2xr, e->m_xr2x);
  memcpy(e->m_x2c, m_x2c, 72);
  orc_matinv(m_x2c, e->m_c2x);
  e->ismolecule = ismol;
  e->nspc = nspc;
  e->zspc = malloc(sizeof(int) * (nspc > 0 ? nspc : 1));
  memcpy(e->zspc, zspc, sizeof(int) * nspc);

  /* boxsize :107-114 */
  double dmax = 0;
  for (int i = 0; i < nspc; i++)
    if (zspc[i] > 0) dmax = fmax(dmax, atmcov_bohr(zspc[i]));
  e->boxsize = fmax(4.0, 4.0 * dmax + 1e-10);

  /* sphmax :116-121 */
  static const double corners[4][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  double sphmax = 0;
  for (int c = 0; c < 4; c++) {
    double d[3] = {corners[c][0] - 0.5, corners[c][1] - 0.5, corners[c][2] - 0.5}, rr[3];
    orc_matvec(e->m_xr2c, d, rr);
    sphmax = fmax(sphmax, orc_norm2(rr, 3));
  }
  e->rsph_uc = sphmax;

  /* dmax0 :123-133 */
  if (dmax0_in >= 0) {
    e->dmax0 = dmax0_in + 1e-2;
  } else {
    e->dmax0 = 0;
    for (int i = 0; i < nspc; i++)
      if (zspc[i] > 0) e->dmax0 = fmax(e->dmax0, orc_cutrad(zspc[i]));
    e->dmax0 = fmax(e->dmax0, sphmax);
    e->dmax0 += 1e-2;
  }

  /* origin = centre of the unit cell :139-141 */
  double half[3] = {0.5, 0.5, 0.5};
  orc_matvec(e->m_xr2c, half, e->x0);

  int cap = 0;
  double rmin[3] = {1e40, 1e40, 1e40}, rmax[3] = {-1e40, -1e40, -1e40};
  e->n = 0;
  for (int i = 0; i < ncell; i++) {
    double xr[3], r[3];
    int lv[3] = {0, 0, 0};
    orc_matvec(e->m_x2xr, xcryst + 3 * i, xr);
    if (!ismol) { /* :178-184 */
      for (int d = 0; d < 3; d++) xr[d] = xr[d] - floor(xr[d]);
      double xb[3];
      orc_matvec(e->m_xr2x, xr, xb);
      for (int d = 0; d < 3; d++) lv[d] = (int)lround(xb[d] - xcryst[3 * i + d]);
    }
    orc_matvec(e->m_xr2c, xr, r);
    for (int d = 0; d < 3; d++) { rmin[d] = fmin(rmin[d], r[d]); rmax[d] = fmax(rmax[d], r[d]); }
    env_push(e, &cap, xr, r, is[i], i + 1, lv);
  }
  e->ncell = ncell;

  if (!ismol) { /* periodic images :191-212, addcell :301-330 */
    int dorepeat = 1, imax = 0;
    while (dorepeat) {
      dorepeat = 0;
      imax++;
      for (int i1 = -imax; i1 <= imax; i1++)
        for (int i2 = -imax; i2 <= imax; i2++) {
          int i3min = (abs(i1) == imax || abs(i2) == imax) ? 0 : imax;
          for (int i3 = i3min; i3 <= imax; i3++)
            for (int sgn = 0; sgn < 2; sgn++) {
              if (sgn == 1 && i3 == 0) continue;
              int p[3] = {i1, i2, sgn ? -i3 : i3};
              double pr[3] = {p[0], p[1], p[2]}, pxr[3];
              orc_matvec(e->m_xr2x, pr, pxr);
              int px[3] = {(int)lround(pxr[0]), (int)lround(pxr[1]), (int)lround(pxr[2])};
              for (int i = 0; i < ncell; i++) {
                double x[3], xc[3], dv[3];
                for (int d = 0; d < 3; d++) x[d] = e->x[3 * i + d] + p[d];
                orc_matvec(e->m_xr2c, x, xc);
                for (int d = 0; d < 3; d++) dv[d] = xc[d] - e->x0[d];
                if (orc_norm2(dv, 3) < sphmax + e->dmax0) {
                  dorepeat = 1;
                  int lv[3];
                  for (int d = 0; d < 3; d++) lv[d] = e->lvec[3 * i + d] + px[d];
                  for (int d = 0; d < 3; d++) { rmin[d] = fmin(rmin[d], xc[d]); rmax[d] = fmax(rmax[d], xc[d]); }
                  env_push(e, &cap, x, xc, e->is[i], e->cidx[i], lv);
                }
              }
            }
        }
    }
  }

  double dd[3] = {rmax[0] - rmin[0], rmax[1] - rmin[1], rmax[2] - rmin[2]};
  e->rsph_env = 0.5 * orc_norm2(dd, 3);                        /* :216 */
  e->boxsize = fmax(e->boxsize, 2.0 * e->rsph_env / 100.0);    /* :219 */
  e->nregion = 1;
  for (int d = 0; d < 3; d++) {                                /* :222-227 */
    e->nmin[d] = (int)floor((rmin[d] - e->x0[d]) / e->boxsize);
    e->nmax[d] = (int)floor((rmax[d] - e->x0[d]) / e->boxsize);
    e->nreg[d] = e->nmax[d] - e->nmin[d] + 1;
    e->nregion *= e->nreg[d];
  }

  /* ordered list of atoms by region :229-254 */
  ikey *ik = malloc(sizeof(ikey) * e->n);
  for (int i = 0; i < e->n; i++) { ik[i].key = c2i(e, e->r + 3 * i); ik[i].idx = i; }
  qsort(ik, e->n, sizeof(ikey), cmp_ikey);
  e->imap = malloc(sizeof(int) * e->n);
  e->nrlo = malloc(sizeof(int) * e->nregion);
  e->nrhi = malloc(sizeof(int) * e->nregion);
  for (int i = 0; i < e->nregion; i++) { e->nrlo[i] = 1; e->nrhi[i] = 0; }
  for (int i = 0; i < e->n; i++) {
    e->imap[i] = ik[i].idx;
    int reg = ik[i].key - 1;
    if (i == 0 || ik[i - 1].key != ik[i].key) e->nrlo[reg] = i + 1;
    e->nrhi[reg] = i + 1;
  }
  free(ik);

  /* search offsets sorted by distance :256-298 */
  e->rs_imax = (int)ceil(fmax(e->dmax0, e->rsph_env) / e->boxsize);
  e->rs_2imax1 = 2 * e->rs_imax + 1;
  e->rs_nreg = e->rs_2imax1 * e->rs_2imax1 * e->rs_2imax1;
  dkey *dk = malloc(sizeof(dkey) * e->rs_nreg);
  int *off = malloc(sizeof(int) * e->rs_nreg);
  const double ctsq32 = sqrt(3.0) / 2.0;
  int nreg = 0;
  for (int i1 = -e->rs_imax; i1 <= e->rs_imax; i1++)
    for (int i2 = -e->rs_imax; i2 <= e->rs_imax; i2++)
      for (int i3 = -e->rs_imax; i3 <= e->rs_imax; i3++) {
        int iv[3] = {i1, i2, i3};
        double s = 0;
        for (int d = 0; d < 3; d++) {
          double a = fmax(iv[d] - 0.5, 0.0), b = fmax(-(iv[d] + 0.5), 0.0);
          s += a * a + b * b;
        }
        /* the reference sums x0^2 terms first then x1^2 (:281); same value up to rounding of a 6-term sum */
        double dist = sqrt(s);
        double rcut0 = fmax(dist - ctsq32, 0.0) * e->boxsize;
        off[nreg] = i1 + e->rs_imax + e->rs_2imax1 * (i2 + e->rs_imax + e->rs_2imax1 * (i3 + e->rs_imax)); /* packoffset :1603 */
        dk[nreg].key = rcut0;
        dk[nreg].idx = nreg;
        nreg++;
      }
  qsort(dk, e->rs_nreg, sizeof(dkey), cmp_dkey);
  e->rs_ioffset = malloc(sizeof(int) * e->rs_nreg);
  e->rs_rcut = malloc(sizeof(double) * e->rs_nreg);
  for (int i = 0; i < e->rs_nreg; i++) { e->rs_ioffset[i] = off[dk[i].idx]; e->rs_rcut[i] = dk[i].key; }
  free(dk);
  free(off);
  return e;
}

void orc_environ_free(orc_environ *e) {
  if (!e) return;
  free(e->zspc); free(e->x); free(e->r); free(e->is); free(e->cidx); free(e->lvec);
  free(e->imap); free(e->nrlo); free(e->nrhi); free(e->rs_ioffset); free(e->rs_rcut);
  free(e);
}

/* list_near_atoms :895-1193, restricted to the two call shapes on the path:
   mode 0 = up2dcidx (rho2 :2070), mode 1 = up2dsp with lower limits 0 (promolecular :1268); sorted=.false. */
int orc_list_near_atoms(const orc_environ *e, const double *xp, int mode, const double *cut, int *nat_out, int *eid,
                        double *dist, int *lvec_out) {
  double x0[3] = {xp[0], xp[1], xp[2]};
  int lvec[3];
  orc_center_cart(e, x0, lvec);                                /* :938-939 */
  int ireg0[3];
  c2p(e, x0, ireg0);
  if (lvec_out) {                                              /* :941 */
    double lr[3] = {lvec[0], lvec[1], lvec[2]}, lx[3];
    orc_matvec(e->m_xr2x, lr, lx);
    for (int d = 0; d < 3; d++) lvec_out[d] = (int)lround(lx[d]);
  }
  double dv[3] = {x0[0] - e->x0[0], x0[1] - e->x0[1], x0[2] - e->x0[2]};
  double dmaxenv = fmax(e->rsph_env - orc_norm2(dv, 3), 0.0);  /* :945 */
  double up2rmax = 0;
  int ncut = (mode == 0) ? e->ncell : e->nspc;
  for (int i = 0; i < ncut; i++) up2rmax = (i == 0) ? cut[0] : fmax(up2rmax, cut[i]); /* maxval :958-959 */

  int nat = 0, nats = 0, enough = 0, i = 0;
  while (!enough && i < e->rs_nreg) {                          /* :966 */
    int ip = e->rs_ioffset[i];
    i++;
    int ioff[3];                                               /* unpackoffset :1613 */
    int idd = ((ip % e->rs_2imax1) + e->rs_2imax1) % e->rs_2imax1;
    ioff[0] = idd - e->rs_imax;
    ip = (ip - idd) / e->rs_2imax1;
    idd = ((ip % e->rs_2imax1) + e->rs_2imax1) % e->rs_2imax1;
    ioff[1] = idd - e->rs_imax;
    ip = (ip - idd) / e->rs_2imax1;
    ioff[2] = ip - e->rs_imax;
    int ireg[3] = {ireg0[0] + ioff[0], ireg0[1] + ioff[1], ireg0[2] + ioff[2]};
    int inside = 1;
    for (int d = 0; d < 3; d++)
      if (ireg[d] < e->nmin[d] || ireg[d] > e->nmax[d]) inside = 0;
    if (inside) {
      int idxreg = p2i(e, ireg) - 1;
      for (int j = e->nrlo[idxreg]; j <= e->nrhi[idxreg]; j++) {
        int k = e->imap[j - 1];
        nats++;
        double d3[3] = {e->r[3 * k] - x0[0], e->r[3 * k + 1] - x0[1], e->r[3 * k + 2] - x0[2]};
        double dist0 = orc_norm2(d3, 3);                       /* :988 */
        if (mode == 1) {
          int s = e->is[k] - 1;
          if (dist0 < 0.0 || dist0 > cut[s]) continue;         /* :998 */
        } else {
          if (dist0 > cut[e->cidx[k] - 1]) continue;           /* :1001 */
        }
        eid[nat] = k;
        dist[nat] = dist0;
        nat++;
      }
    }
    enough = (e->n == nats && e->ismolecule);                  /* :1014 */
    enough = enough || (e->rs_rcut[i - 1] > up2rmax && (dmaxenv > up2rmax || e->ismolecule)); /* :1025/:1029 */
  }
  *nat_out = nat;
  return enough ? 0 : 1;                                       /* :1134-1135 */
}

/* identify_atom :489-528 through nearest_atom_short :608-703 for distmax < boxsize/2 (the only use on the path is
   distmax = 1d-5, fieldmod@proc.f90:836).  Returns the 1-based cidx or 0. */
int orc_identify_atom(const orc_environ *e, const double *x0in, double distmax) {
  double d0 = fmin(distmax, 0.5 * e->boxsize - 1e-10);
  double eps2 = d0 * d0;
  double xp[3] = {x0in[0], x0in[1], x0in[2]};
  orc_center_cart(e, xp, NULL);
  int ireg0[3];
  c2p(e, xp, ireg0);
  int imin[3] = {0, 0, 0}, imax[3] = {0, 0, 0};
  for (int d = 0; d < 3; d++) {
    double x0r = e->x0[d] + e->boxsize * ireg0[d];
    double x1r = e->x0[d] + e->boxsize * (ireg0[d] + 1);
    if (xp[d] - x0r <= d0) imin[d] = -1;
    if (x1r - xp[d] <= d0) imax[d] = 1;
  }
  int kmin = -1;
  double d2min = 1e40;
  for (int i1 = imin[0]; i1 <= imax[0]; i1++)
    for (int i2 = imin[1]; i2 <= imax[1]; i2++)
      for (int i3 = imin[2]; i3 <= imax[2]; i3++) {
        int ireg[3] = {ireg0[0] + i1, ireg0[1] + i2, ireg0[2] + i3};
        int ok = 1;
        for (int d = 0; d < 3; d++)
          if (ireg[d] < e->nmin[d] || ireg[d] > e->nmax[d]) ok = 0;
        if (!ok) continue;
        int idx = p2i(e, ireg) - 1;
        for (int j = e->nrlo[idx]; j <= e->nrhi[idx]; j++) {
          int k = e->imap[j - 1];
          double x = e->r[3 * k] - xp[0], y = e->r[3 * k + 1] - xp[1], z = e->r[3 * k + 2] - xp[2];
          double d2 = x * x + y * y + z * z;
          if (d2 < d2min && d2 < eps2) { kmin = k; d2min = d2; } /* :677 */
        }
      }
  if (kmin >= 0 && d2min < eps2) return e->cidx[kmin];
  return 0;
}
