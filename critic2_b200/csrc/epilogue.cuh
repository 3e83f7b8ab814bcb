// epilogue.cuh -- per-point tail of field%grd and the NCIPLOT formulas, fused into the evaluation kernels.
// Citations are relative to /root/reference/.
#pragma once
#include "common.cuh"

namespace c2 {

#ifdef __CUDACC__
// Eigenvalues (ascending) of a symmetric 3x3 matrix with the reference's own algorithm: Householder
// tridiagonalisation (src/oldlibs/tredig.f) + QL with implicit shifts, AMACH = 16.0E-13 single-precision literal,
// at most 30 iterations (src/oldlibs/tqlgrm.f), reached through eig (src/tools_math@proc.f90:822-841).
// Only the eigenvalue recurrences are kept (the eigenvector accumulation does not feed back into them).
// h is column-major hf(3,3); only its lower triangle is read, like tredig.  Returns ierr (0 ok).
__device__ inline int eig3_ref(const double *h, double *ev) {
  // lower triangle z(i,j), i>=j, 1-based names
  double z11 = h[0], z21 = h[1], z22 = h[4], z31 = h[2], z32 = h[5], z33 = h[8];
  double d[3], e[3];
  // ---- tredig, i = 3, l = 2
  {
    double scale = fabs(z31) + fabs(z32);
    if (scale == 0.0) {
      e[2] = z32;
      d[2] = 0.0;
    } else {
      double rscale = 1.0 / scale;
      z31 = z31 * rscale;
      z32 = z32 * rscale;
      double hh = z31 * z31;
      hh += z32 * z32;
      double f = z32;
      double g = -copysign(sqrt(hh), f);
      e[2] = scale * g;
      hh = hh - f * g;
      z32 = f - g;
      f = 0.0;
      double rh = 1.0 / hh;
      // j = 1: g = z(1,1)*z(3,1) + z(2,1)*z(3,2)
      double g1 = z11 * z31;
      g1 += z21 * z32;
      double e1 = g1 * rh;
      f += e1 * z31;
      // j = 2: g = z(2,1)*z(3,1) + z(2,2)*z(3,2)
      double g2 = z21 * z31;
      g2 += z22 * z32;
      double e2 = g2 * rh;
      f += e2 * z32;
      double hq = f / (hh + hh);
      // j = 1
      double f1 = z31;
      double gg1 = e1 - hq * f1;
      e1 = gg1;
      z11 = z11 - f1 * e1 - gg1 * z31;
      // j = 2
      double f2 = z32;
      double gg2 = e2 - hq * f2;
      e2 = gg2;
      z21 = z21 - f2 * e1 - gg2 * z31;
      z22 = z22 - f2 * e2 - gg2 * z32;
      d[2] = hh;
    }
  }
  // ---- tredig, i = 2, l = 1 (l < 2: simple branch)
  e[1] = z21;
  d[1] = 0.0;
  d[0] = 0.0;
  e[0] = 0.0;
  // final loop: d(i) = z(i,i)
  d[0] = z11;
  d[1] = z22;
  d[2] = z33;
  // ---- tqlgrm, n = 3
  const double amach = (double)16.0e-13f;
  e[0] = e[1];
  e[1] = e[2];
  e[2] = 0.0;
  double f = 0.0, b = 0.0;
  // every loop below has compile-time bounds (with predicates) so that d[] and e[] stay in registers
#pragma unroll
  for (int l = 0; l < 3; l++) {
    int j = 0;
    double hh = amach * (fabs(d[l]) + fabs(e[l]));
    if (b < hh) b = hh;
    int m = 3;
#pragma unroll
    for (int mm = 2; mm >= 0; mm--)
      if (mm >= l && fabs(e[mm]) <= b) m = mm;   // first m >= l with |e(m)| <= b (e[2] = 0 always qualifies)
    if (l < 2 && m != l) {   // l = 2: m = 2 always (e[2] = 0)
      do {
        if (j == 30) return l + 1;
        j++;
        double g = d[l];
        double p = (d[l + 1] - g) / (2.0 * e[l]);
        double r = (fabs(p * amach) > 1.0) ? p : sqrt(p * p + 1.0);
        d[l] = e[l] / (p + copysign(r, p));
        hh = g - d[l];
#pragma unroll
        for (int i = 1; i < 3; i++)
          if (i >= l + 1) d[i] = d[i] - hh;
        f = f + hh;
        p = (m == 2) ? d[2] : (m == 1 ? d[1] : d[0]);
        double c = 1.0, s = 0.0;
#pragma unroll
        for (int i = 1; i >= 0; i--) {
          if (i > m - 1 || i < l) continue;
          g = c * e[i];
          hh = c * p;
          if (fabs(p) >= fabs(e[i])) {
            c = e[i] / p;
            r = sqrt(c * c + 1.0);
            e[i + 1] = s * p * r;
            s = c / r;
            c = 1.0 / r;
          } else {
            c = p / e[i];
            r = sqrt(c * c + 1.0);
            e[i + 1] = s * e[i] * r;
            s = 1.0 / r;
            c = c * s;
          }
          p = c * d[i] - s * g;
          d[i + 1] = hh + s * (c * g + s * d[i]);
        }
        e[l] = s * p;
        d[l] = c * p;
      } while (fabs(e[l]) > b);
    }
    d[l] = d[l] + f;
  }
  // selection sort ascending (tqlgrm.f tail)
  for (int i = 0; i < 2; i++) {
    int k = i;
    double p = d[i];
    for (int jj = i + 1; jj < 3; jj++)
      if (d[jj] < p) { k = jj; p = d[jj]; }
    if (k != i) { d[k] = d[i]; d[i] = p; }
  }
  ev[0] = d[0]; ev[1] = d[1]; ev[2] = d[2];
  return 0;
}

// Tail of field%grd (src/fieldmod@proc.f90:834-840) + property selection (src/rhoplot@proc.f90:596-621) or the
// NCIPLOT per-point formulas (src/nci@proc.f90:523-525, 553-554).  hs = (xx,yy,zz,xy,xz,yz).
// nci_const = 2*(3*pi^2)^(1/3), evaluated on the host like the reference does at run time (src/nci@proc.f90:86-87).
__device__ __forceinline__ void emit_point(const OutDesc &o, size_t idx, double f, double gx, double gy, double gz,
                                           double h0, double h1, double h2, double h3, double h4, double h5,
                                           bool isnuc, double nci_const) {
  const size_t cs = o.core_stride;
  if (o.mode == 4) {                                                  // raw sums for the core pre-pass
    double *r = o.res + idx;
    r[0] = f;
    if (o.core_nd >= 4) { r[cs] = gx; r[2 * cs] = gy; r[3 * cs] = gz; }
    if (o.core_nd >= 10) { r[4 * cs] = h0; r[5 * cs] = h1; r[6 * cs] = h2; r[7 * cs] = h3; r[8 * cs] = h4; r[9 * cs] = h5; }
    return;
  }
  if (o.core) {                                                       // src/fieldmod@proc.f90:827-832
    const double *c = o.core + idx;
    f = add_rn(f, c[0]);
    if (o.core_nd >= 4) { gx = add_rn(gx, c[cs]); gy = add_rn(gy, c[2 * cs]); gz = add_rn(gz, c[3 * cs]); }
    if (o.core_nd >= 10) {
      h0 = add_rn(h0, c[4 * cs]); h1 = add_rn(h1, c[5 * cs]); h2 = add_rn(h2, c[6 * cs]);
      h3 = add_rn(h3, c[7 * cs]); h4 = add_rn(h4, c[8 * cs]); h5 = add_rn(h5, c[9 * cs]);
    }
  }
  if (isnuc) { gx = 0.0; gy = 0.0; gz = 0.0; }                        // :838
  if (o.mode == 5) {   // first pass of the two-pass NCIPLOT evaluation: f, gf, hs after the core and the nucleus test
    double *r = o.res + idx;
    r[0] = f; r[cs] = gx; r[2 * cs] = gy; r[3 * cs] = gz;
    r[4 * cs] = h0; r[5 * cs] = h1; r[6 * cs] = h2; r[7 * cs] = h3; r[8 * cs] = h4; r[9 * cs] = h5;
    return;
  }
  double gmod = norm2_ref(gx, gy, gz);                                // :839
  double lap = add_rn(add_rn(h0, h1), h2);                   // :840
  if (o.mode == 0) {
    const double v[12] = {f, gx, gy, gz, gmod, h0, h3, h4, h1, h5, h2, lap};
    int slot = 0;
#pragma unroll
    for (int nti = 0; nti < 12; nti++)
      if (o.outmask & (1u << nti)) o.prop[slot++][idx] = v[nti];
  } else if (o.mode == 1) {
    const double hf[9] = {h0, h3, h4, h3, h1, h5, h4, h5, h2};
    double ev[3];
    eig3_ref(hf, ev);
    const double vsmall = 1e-80;                                      // src/param.F90:27
    double dimgrad = gmod / (nci_const * pow(fmax(f, vsmall), 4.0 / 3.0));
    o.cgrad[idx] = dimgrad;
    o.crho[idx] = copysign(fabs(f), ev[1]) * 100.0;
  } else if (o.mode == 2) {
    double *r = o.res + idx * 16;
    r[0] = f; r[1] = gx; r[2] = gy; r[3] = gz;
    r[4] = h0; r[5] = h3; r[6] = h4;
    r[7] = h3; r[8] = h1; r[9] = h5;
    r[10] = h4; r[11] = h5; r[12] = h2;
    r[13] = gmod; r[14] = lap; r[15] = isnuc ? 1.0 : 0.0;
  } else {
    o.prop[0][idx] = f;
  }
}

// Out-of-line copy for the warp-specialised kernel: its point epilogue runs on 16 threads once per point, and keeping this
// code (eigenvalue iteration, pow, twelve stores) out of the kernel body keeps the register allocation of the DMMA
// loops independent of it (measured: inlining a larger emit_point cost 3 % of the whole pass through a different
// operand-register assignment in the consumer loop).
static __device__ __noinline__ void emit_point_call(const OutDesc &o, size_t idx, double f, double gx, double gy, double gz,
                                             double h0, double h1, double h2, double h3, double h4, double h5,
                                             bool isnuc, double nci_const) {
  emit_point(o, idx, f, gx, gy, gz, h0, h1, h2, h3, h4, h5, isnuc, nci_const);
}
#endif

}  // namespace c2
