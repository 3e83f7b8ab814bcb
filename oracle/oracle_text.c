/* oracle_text.c -- TEST INFRASTRUCTURE: C restatement of the formatted writes of the cube writers, for arrays too large
 * for the pure-Python restatement (oracle/cube_text.py, which it must agree with byte for byte) and as the timed CPU
 * baseline of the formatting step.  Formats and record structure: src/crystalmod@proc.f90:5026-5036 (writegrid_cube),
 * src/nci@proc.f90:867-883 (write_cube_body), :593-595 (.dat).  Digits come from glibc's correctly rounded "%.*E". */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "oracle.h"
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { int w, d, k, e, lead, per_line; } orc_fmt;
static orc_fmt orc_fmt_of(int fmt) {
  switch (fmt) {
    case 0: { orc_fmt f = {22, 14, 0, 2, 1, 6}; return f; }
    case 1: { orc_fmt f = {12, 5, 1, 2, 1, 6}; return f; }
    case 2: { orc_fmt f = {13, 5, 1, 3, 1, 6}; return f; }
    default: { orc_fmt f = {15, 7, 1, 2, 0, 2}; return f; }
  }
}

/* one field of f.lead + f.w characters (no terminator) */
static void orc_text_field(double x, orc_fmt f, char *dst) {
  const int W = f.w + f.lead;
  char out[64], tmp[64];
  int n = 0;
  memset(dst, ' ', W);
  if (isnan(x)) { memcpy(dst + W - 3, "NaN", 3); return; }
  if (isinf(x)) { const char *s = x < 0 ? "-Infinity" : "Infinity"; memcpy(dst + W - (int)strlen(s), s, strlen(s)); return; }
  const int nd = f.k == 0 ? f.d : f.d + 1;
  char digits[32];
  int e10;
  if (x == 0.0) { memset(digits, '0', nd); e10 = f.k == 0 ? -1 : 0; }
  else {
    snprintf(tmp, sizeof tmp, "%.*E", nd - 1, fabs(x));   /* d.dddE+xx */
    char *ep = strchr(tmp, 'E');
    e10 = atoi(ep + 1);
    int j = 0;
    for (char *c = tmp; c < ep; c++) if (*c != '.') digits[j++] = *c;
  }
  const int expo = f.k == 0 ? e10 + 1 : e10;
  if (signbit(x)) out[n++] = '-';
  if (f.k == 0) { out[n++] = '0'; out[n++] = '.'; memcpy(out + n, digits, nd); n += nd; }
  else { out[n++] = digits[0]; out[n++] = '.'; memcpy(out + n, digits + 1, nd - 1); n += nd - 1; }
  const int ae = expo < 0 ? -expo : expo;
  if (f.e == 3) n += snprintf(out + n, 8, "E%c%03d", expo < 0 ? '-' : '+', ae);
  else if (ae <= 99) n += snprintf(out + n, 8, "E%c%02d", expo < 0 ? '-' : '+', ae);
  else n += snprintf(out + n, 8, "%c%03d", expo < 0 ? '-' : '+', ae);
  memcpy(dst + W - n, out, n);
}

long orc_text_record_bytes(int nval, int fmt) {
  orc_fmt f = orc_fmt_of(fmt);
  return (long)nval * (f.w + f.lead) + (nval + f.per_line - 1) / f.per_line;
}

/* order 0: g(n1,n2,n3) x fastest, records over (ix,iy) with iz running; order 1: records are contiguous (n1*n2 records
   of n3 values).  text: n1*n2*orc_text_record_bytes(n3,fmt) bytes. */
void orc_cube_text(const double *g, const int *n, int fmt, int order, char *text, int nthreads) {
  const orc_fmt f = orc_fmt_of(fmt);
  const int W = f.w + f.lead;
  const long rb = orc_text_record_bytes(n[2], fmt);
  const long nrec = (long)n[0] * n[1];
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(static)
  for (long r = 0; r < nrec; r++) {
    char *row = text + r * rb;
    for (int k = 0; k < n[2]; k++) {
      double x;
      if (order == 0) { long ix = r / n[1], iy = r % n[1]; x = g[ix + (long)n[0] * (iy + (long)n[1] * k)]; }
      else x = g[r * n[2] + k];
      char *dst = row + (long)k * W + k / f.per_line;
      orc_text_field(x, f, dst);
      if ((k + 1) % f.per_line == 0 || k == n[2] - 1) dst[W] = '\n';
    }
  }
}
