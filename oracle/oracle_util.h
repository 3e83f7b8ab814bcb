/* oracle_util.h -- small helpers shared by the oracle sources. TEST INFRASTRUCTURE ONLY (see oracle.h). */
#ifndef C2_ORACLE_UTIL_H
#define C2_ORACLE_UTIL_H
#include <math.h>

/* y = M x, M column-major 3x3 (Fortran matmul(M,x)) */
static inline void orc_matvec(const double *m, const double *x, double *y) {
  for (int i = 0; i < 3; i++) y[i] = m[i] * x[0] + m[i + 3] * x[1] + m[i + 6] * x[2];
}

/* gfortran's NORM2 intrinsic (libgfortran generated/norm2_r8.c): scaled sum of squares.  The reference uses
   norm2() for every atom distance (environmod@proc.f90:988) and for gfmod (fieldmod@proc.f90:839). */
static inline double orc_norm2(const double *x, int n) {
  double scale = 1.0, result = 0.0;
  for (int i = 0; i < n; i++) {
    if (x[i] != 0) {
      double absx = fabs(x[i]);
      if (scale < absx) {
        double val = scale / absx;
        result = 1 + result * val * val;
        scale = absx;
      } else {
        double val = absx / scale;
        result += val * val;
      }
    }
  }
  return scale * sqrt(result);
}

/* 3x3 inverse by cofactors (tools_math matinv is LU based; entries agree to rounding) */
static inline void orc_matinv(const double *m, double *inv) {
  double a = m[0], b = m[3], c = m[6], d = m[1], e = m[4], f = m[7], g = m[2], h = m[5], i = m[8];
  double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
  double r = 1.0 / det;
  inv[0] = (e * i - f * h) * r; inv[3] = (c * h - b * i) * r; inv[6] = (b * f - c * e) * r;
  inv[1] = (f * g - d * i) * r; inv[4] = (a * i - c * g) * r; inv[7] = (c * d - a * f) * r;
  inv[2] = (d * h - e * g) * r; inv[5] = (b * g - a * h) * r; inv[8] = (a * e - b * d) * r;
}

void orc_center_cart(const struct orc_environ *e, double *xx, int *lvec);
#endif
