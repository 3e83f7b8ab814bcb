// cube_format.cu -- text bodies of cube / .dat files, formatted on the device.
//
// Replaces the formatted-write loops that follow the grid evaluation in the reference:
//   writegrid_cube   src/crystalmod@proc.f90:5026-5036   '(6(1x,e22.14))' (PRECISECUBE) or '(1p,6(1x,e12.5))',
//                                                         g(ix,iy,iz): one record per (ix,iy), iz fastest
//   write_cube_body  src/nci@proc.f90:867-883             '(6(1x,1p,e13.5e3))', c(k,j,i): one record per (i,j), k fastest
//   .dat pairs       src/nci@proc.f90:593-595             '(1p,E15.7,E15.7)'
// Every value becomes a fixed-width field, so the byte offset of each field is known in advance: one thread per value,
// HBM-bound byte work (8 B read, 13-24 B written per value).  Decimal conversion is exact (round-half-even on the exact
// binary value, what gfortran's formatted output gives in the default rounding mode): the double is scaled by a
// 128-bit power of ten (pow10_table.inc) to a D-digit integer with ~80 fraction bits; the rare cases the fraction
// cannot decide (exact ties behind an inexact table entry) are settled by an exact big-integer comparison.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstring>

#include "c2gpu_internal.h"

namespace c2 {

#include "pow10_table.inc"

__device__ unsigned long long d_pow10_tab[POW10_MAX - POW10_MIN + 1][3];

struct FmtSpec {
  int D;        // significant digits
  int style;    // 0: 0.ddddE+ee (Ew.d), 1: d.dddE+ee (1P Ew.d)
  int w;        // field width (without the leading 1x)
  int lead;     // 1 if the field is preceded by 1x
  int expdig;   // exponent digits (2: E+ee / +eee, 3: E+eee)
  int per_line; // values per line
};

static __host__ __device__ inline FmtSpec fmt_spec(int fmt) {
  switch (fmt) {
    case 0: return FmtSpec{14, 0, 22, 1, 2, 6};   // (6(1x,e22.14))
    case 1: return FmtSpec{6, 1, 12, 1, 2, 6};    // (1p,6(1x,e12.5))
    case 2: return FmtSpec{6, 1, 13, 1, 3, 6};    // (6(1x,1p,e13.5e3))
    default: return FmtSpec{8, 1, 15, 0, 2, 2};   // (1p,E15.7,E15.7)
  }
}

// ---- exact comparison of m*2^q*10^p with (2N+1)/2 (slow path) -----------------------------------------------------
struct Big {
  static constexpr int NW = 64;
  uint32_t w[NW];
  int n;   // used words
};
__device__ static void big_set(Big &b, unsigned long long v) {
  for (int i = 0; i < Big::NW; i++) b.w[i] = 0;
  b.w[0] = (uint32_t)v; b.w[1] = (uint32_t)(v >> 32);
  b.n = b.w[1] ? 2 : (b.w[0] ? 1 : 0);
}
__device__ static void big_mul_small(Big &b, uint32_t f) {
  unsigned long long carry = 0;
  for (int i = 0; i < b.n; i++) {
    unsigned long long t = (unsigned long long)b.w[i] * f + carry;
    b.w[i] = (uint32_t)t;
    carry = t >> 32;
  }
  if (carry && b.n < Big::NW) b.w[b.n++] = (uint32_t)carry;
}
__device__ static void big_mul_pow5(Big &b, int k) {
  while (k >= 13) { big_mul_small(b, 1220703125u); k -= 13; }
  uint32_t f = 1;
  for (int i = 0; i < k; i++) f *= 5u;
  if (f > 1) big_mul_small(b, f);
}
__device__ static int big_bitlen(const Big &b) {
  if (b.n == 0) return 0;
  return 32 * (b.n - 1) + (32 - __clz(b.w[b.n - 1]));
}
__device__ static void big_shl(Big &b, int s) {
  const int ws = s >> 5, bs = s & 31;
  for (int i = Big::NW - 1; i >= 0; i--) {
    unsigned long long v = 0;
    const int j = i - ws;
    if (j >= 0) v = (unsigned long long)b.w[j] << bs;
    if (j - 1 >= 0 && bs) v |= (unsigned long long)b.w[j - 1] >> (32 - bs);
    b.w[i] = (uint32_t)v;
  }
  b.n = min(Big::NW, b.n + ws + 1);
  while (b.n > 0 && b.w[b.n - 1] == 0) b.n--;
}
// sign of m*2^q*10^p - (2N+1)/2
__device__ __noinline__ static int exact_cmp_half(unsigned long long m, int q, int p, unsigned long long N) {
  Big X, Y;
  int sx, sy;
  big_set(X, m);
  big_set(Y, 2 * N + 1);
  if (p >= 0) { big_mul_pow5(X, p); sx = q + 1 + p; sy = 0; }
  else { big_mul_pow5(Y, -p); sx = q + 1; sy = -p; }
  const int bx = big_bitlen(X) + sx, by = big_bitlen(Y) + sy;
  if (bx != by) return bx > by ? 1 : -1;
  const int d = sx - sy;
  if (d > 0) big_shl(X, d); else if (d < 0) big_shl(Y, -d);
  for (int i = Big::NW - 1; i >= 0; i--)
    if (X.w[i] != Y.w[i]) return X.w[i] > Y.w[i] ? 1 : -1;
  return 0;
}

__device__ __forceinline__ unsigned long long pow10_u64(int k) {
  unsigned long long v = 1;
  for (int i = 0; i < k; i++) v *= 10ull;
  return v;
}

// x finite, non-zero, positive magnitude given by its bits: N (D digits) and e10 with |x| ~ N * 10^(e10 - D + 1)
__device__ __forceinline__ static void to_decimal(unsigned long long bits, int D, int force_exact, unsigned long long &N, int &e10) {
  const int ex = (int)((bits >> 52) & 0x7ff);
  const unsigned long long frac = bits & 0x000fffffffffffffull;
  const unsigned long long m = ex ? (frac | 0x0010000000000000ull) : frac;
  const int q = ex ? ex - 1075 : -1074;              // |x| = m * 2^q
  const int lz = __clzll((long long)m);
  const unsigned long long m64 = m << lz;
  const int q64 = q - lz;                            // |x| = m64 * 2^q64, bit 63 of m64 set
  // floor(e2x log10 2) from a constant rounded towards the safe side: e10 is floor(log10|x|) or one less, never more
  const long long e2x = q64 + 63;
  e10 = (int)((e2x * (e2x >= 0 ? 1292913986ll : 1292913987ll)) >> 32);
  const unsigned long long lim_hi = pow10_u64(D), lim_lo = lim_hi / 10ull;
  unsigned long long P0 = 0, P1 = 0, rhi = 0;
  int k = 1, p = 0, exact = 0;
  for (int iter = 0; iter < 2; iter++) {
    p = D - 1 - e10;
    const unsigned long long *t = d_pow10_tab[p - POW10_MIN];
    const unsigned long long hi = t[0], lo = t[1];
    const int e2 = (int)(t[2] >> 1) - 2048;
    exact = (int)(t[2] & 1ull);
    const unsigned long long a_lo = m64 * lo, a_hi = __umul64hi(m64, lo);
    const unsigned long long b_lo = m64 * hi, b_hi = __umul64hi(m64, hi);
    P0 = a_lo;
    P1 = a_hi + b_lo;
    const unsigned long long P2 = b_hi + (P1 < a_hi ? 1ull : 0ull);
    k = -(q64 + e2) - 128;                           // fraction bits inside P2 (15 < k < 48 for D in 6..14)
    N = P2 >> k;
    rhi = P2 & ((1ull << k) - 1ull);
    // N == lim_lo - 1 can only come from the truncated table when |x| 10^p is (within 2^-79 of) lim_lo: the fraction
    // is then all ones and the rounding below carries N up to lim_lo, so it is not an exponent error.
    if (N < lim_hi) break;
    e10++;
  }
  const unsigned long long half = 1ull << (k - 1);
  int up;   // 1 round up, 0 round down, -1 tie
  bool ambiguous = force_exact != 0;
  if (!ambiguous) {
    if (exact) {
      if (rhi != half) up = rhi > half;
      else up = (P1 | P0) ? 1 : -1;
    } else {
      // table truncated: the true product lies in [P, P + 2^64) -> P1 may be one larger
      if (rhi > half || (rhi == half && P1 != 0)) up = 1;
      else if (rhi == half) { if (p > 0) up = 1; else { ambiguous = true; up = 0; } }   // p > 55: no exact ties exist
      else if (rhi == half - 1 && P1 == ~0ull) { ambiguous = true; up = 0; }
      else up = 0;
    }
  }
  if (ambiguous) {
    const int c = exact_cmp_half(m, q, p, N);
    up = c > 0 ? 1 : (c < 0 ? 0 : -1);
  }
  if (up < 0) up = (int)(N & 1ull);                  // tie: round half to even
  if (up) {
    N++;
    if (N == lim_hi) { N = lim_lo; e10++; }
  }
}

// Writes the field (lead + w characters) of value x at dst.  The layout of a finite value is static per format (the
// content always fills the same columns), so every character goes straight to its place with an immediate offset: no
// local buffer, no dynamic indexing.
template <int FMT>
__device__ __forceinline__ void format_field(double x, int force_exact, char *dst) {
  constexpr FmtSpec fs = FMT == 0 ? FmtSpec{14, 0, 22, 1, 2, 6}
                         : FMT == 1 ? FmtSpec{6, 1, 12, 1, 2, 6}
                         : FMT == 2 ? FmtSpec{6, 1, 13, 1, 3, 6} : FmtSpec{8, 1, 15, 0, 2, 2};
  constexpr int W = fs.w + fs.lead;
  constexpr int NEXP = fs.expdig == 3 ? 5 : 4;                 // E+eee | E+ee / +eee
  constexpr int NFRAC = fs.style == 0 ? fs.D : fs.D - 1;       // digits after the point
  constexpr int P_EXP = W - NEXP, P_FRAC = P_EXP - NFRAC, P_DOT = P_FRAC - 1, P_LEAD = P_DOT - 1, P_SIGN = P_LEAD - 1;
  const unsigned long long bits = (unsigned long long)__double_as_longlong(x);
  const bool neg = (bits >> 63) != 0;
  const unsigned long long mag = bits & 0x7fffffffffffffffull;
  if (mag >= 0x7ff0000000000000ull) {                          // NaN / Infinity, right-justified
    for (int i = 0; i < W; i++) dst[i] = ' ';
    if (mag > 0x7ff0000000000000ull) { dst[W - 3] = 'N'; dst[W - 2] = 'a'; dst[W - 1] = 'N'; }
    else {
      const char s[9] = "Infinity";
      for (int i = 0; i < 8; i++) dst[W - 8 + i] = s[i];
      if (neg) dst[W - 9] = '-';
    }
    return;
  }
  unsigned long long N = 0;
  int e10 = 0;
  if (mag != 0) to_decimal(mag, fs.D, force_exact, N, e10);
  const int E = mag == 0 ? 0 : (fs.style == 0 ? e10 + 1 : e10);
  const unsigned ae = (unsigned)(E < 0 ? -E : E);
#pragma unroll
  for (int i = 0; i < P_SIGN; i++) dst[i] = ' ';
  dst[P_SIGN] = neg ? '-' : ' ';
  // digits: N < 10^14 -> two 7-digit halves, least significant first
  uint32_t lo7 = (uint32_t)(N % 10000000ull), hi7 = (uint32_t)(N / 10000000ull);
#pragma unroll
  for (int i = 0; i < NFRAC; i++) {
    if (i < 7) { dst[P_EXP - 1 - i] = '0' + (char)(lo7 % 10u); lo7 /= 10u; }
    else { dst[P_EXP - 1 - i] = '0' + (char)(hi7 % 10u); hi7 /= 10u; }
  }
  dst[P_DOT] = '.';
  dst[P_LEAD] = fs.style == 0 ? '0' : (char)('0' + (NFRAC < 7 ? lo7 : hi7) % 10u);
  const char es = E < 0 ? '-' : '+';
  if (fs.expdig == 3) {
    dst[P_EXP] = 'E'; dst[P_EXP + 1] = es;
    dst[P_EXP + 2] = '0' + (char)(ae / 100u); dst[P_EXP + 3] = '0' + (char)((ae / 10u) % 10u); dst[P_EXP + 4] = '0' + (char)(ae % 10u);
  } else if (ae <= 99u) {
    dst[P_EXP] = 'E'; dst[P_EXP + 1] = es; dst[P_EXP + 2] = '0' + (char)(ae / 10u); dst[P_EXP + 3] = '0' + (char)(ae % 10u);
  } else {
    dst[P_EXP] = es; dst[P_EXP + 1] = '0' + (char)(ae / 100u); dst[P_EXP + 2] = '0' + (char)((ae / 10u) % 10u);
    dst[P_EXP + 3] = '0' + (char)(ae % 10u);
  }
}

// ---- kernel: groups of 4 records, segments of SEG values ----------------------------------------------------------
constexpr int FMT_NT = 256, FMT_ROWS = 4, FMT_SEG = 384;       // SEG is a multiple of 6 and of 2
constexpr int FMT_ROWBUF = FMT_SEG * 24 + FMT_SEG / 2 + 16;    // bytes per staged record segment (+16: alignment phase)

struct FmtArgs {
  const double *g;
  char *text;
  long nrec;        // records (rows of the file body)
  int nval;         // values per record
  int fmt, order;   // order 0: g(ix,iy,iz) x fastest, record (ix,iy) -> values over iz; 1: records contiguous
  int n1, n2;       // order 0 only
  int force_exact;
  long rec_bytes;
};

template <int FMT>
__global__ void __launch_bounds__(FMT_NT, 3) format_records_kernel(const FmtArgs a) {
  __shared__ __align__(16) char s_buf[FMT_ROWS][FMT_ROWBUF];
  const FmtSpec fs = fmt_spec(FMT);
  const int W = fs.w + fs.lead, L = fs.per_line;
  const int tid = threadIdx.x, dx = tid & (FMT_ROWS - 1), kk = tid >> 2;
  const long ngroup = a.order == 0 ? (long)((a.n1 + FMT_ROWS - 1) / FMT_ROWS) * a.n2 : (a.nrec + FMT_ROWS - 1) / FMT_ROWS;
  for (long grp = blockIdx.x; grp < ngroup; grp += gridDim.x) {
    long rec;          // this thread's record
    bool rvalid;
    size_t gbase, gstride;
    if (a.order == 0) {
      const int iy = (int)(grp % a.n2), ix = (int)(grp / a.n2) * FMT_ROWS + dx;
      rvalid = ix < a.n1;
      rec = (long)ix * a.n2 + iy;
      gbase = (size_t)ix + (size_t)a.n1 * iy;
      gstride = (size_t)a.n1 * a.n2;
    } else {
      rec = grp * FMT_ROWS + dx;
      rvalid = rec < a.nrec;
      gbase = (size_t)rec * a.nval;
      gstride = 1;
    }
    for (int k0 = 0; k0 < a.nval; k0 += FMT_SEG) {
      const int nk = min(FMT_SEG, a.nval - k0);
      const long seg_off = (long)k0 * W + k0 / L;                       // byte offset of the segment in its record
      const int seg_bytes = nk * W + (nk + L - 1) / L;                  // k0 + nk == nval or nk % L == 0
      __syncthreads();                                                  // previous copy-out finished
      if (rvalid) {
        const int phase = (int)((size_t)(rec * a.rec_bytes + seg_off) & 15);   // co-align staging and destination
        char *row = s_buf[dx] + phase;
        for (int k = kk; k < nk; k += FMT_NT / FMT_ROWS) {
          const double x = a.g[gbase + (size_t)(k0 + k) * gstride];
          char *dst = row + k * W + k / L;
          format_field<FMT>(x, a.force_exact, dst);
          if ((k + 1) % L == 0 || k0 + k == a.nval - 1) dst[W] = '\n';
        }
      }
      __syncthreads();
      // copy-out: 64 threads per record; staging and destination share the 16-byte phase
      {
        const int r = tid >> 6, j = tid & 63;
        long rrec;
        bool ok;
        if (a.order == 0) { const int ix = (int)(grp / a.n2) * FMT_ROWS + r; ok = ix < a.n1; rrec = (long)ix * a.n2 + (int)(grp % a.n2); }
        else { rrec = grp * FMT_ROWS + r; ok = rrec < a.nrec; }
        if (ok) {
          char *gdst = a.text + rrec * a.rec_bytes + seg_off;
          const int phase = (int)((size_t)(rrec * a.rec_bytes + seg_off) & 15);
          const char *src = s_buf[r] + phase;
          const int head = min(seg_bytes, (16 - phase) & 15);
          if (j < head) gdst[j] = src[j];
          const int nvec = (seg_bytes - head) >> 4;
          const uint4 *s4 = reinterpret_cast<const uint4 *>(src + head);
          uint4 *g4 = reinterpret_cast<uint4 *>(gdst + head);
          for (int v = j; v < nvec; v += 64) g4[v] = s4[v];
          const int done = head + (nvec << 4);
          if (j < seg_bytes - done) gdst[done + j] = src[done + j];
        }
      }
    }
  }
}

static bool g_tab_ready[16] = {false};

static cudaError_t ensure_table() {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 16 && g_tab_ready[dev]) return cudaSuccess;
  e = cudaMemcpyToSymbol(d_pow10_tab, POW10_TAB, sizeof(POW10_TAB));
  if (e == cudaSuccess && dev < 16) g_tab_ready[dev] = true;
  return e;
}

long format_record_bytes(int nval, int fmt) {
  const FmtSpec fs = fmt_spec(fmt);
  return (long)nval * (fs.w + fs.lead) + (nval + fs.per_line - 1) / fs.per_line;
}

// text must hold nrec * format_record_bytes(nval, fmt) bytes and be 16-byte aligned (cudaMalloc is).
cudaError_t format_records_launch(const double *g_dev, char *text_dev, long nrec, int nval, int fmt, int order, int n1,
                                  int n2, int force_exact, cudaStream_t st) {
  if (nrec <= 0 || nval <= 0) return cudaSuccess;
  cudaError_t e = ensure_table();
  if (e != cudaSuccess) return e;
  FmtArgs a{g_dev, text_dev, nrec, nval, fmt, order, n1, n2, force_exact, format_record_bytes(nval, fmt)};
  const long ngroup = order == 0 ? (long)((n1 + FMT_ROWS - 1) / FMT_ROWS) * n2 : (nrec + FMT_ROWS - 1) / FMT_ROWS;
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)std::min<long>(ngroup, (long)nsm * 6);
  switch (fmt) {
    case 0: format_records_kernel<0><<<grid, FMT_NT, 0, st>>>(a); break;
    case 1: format_records_kernel<1><<<grid, FMT_NT, 0, st>>>(a); break;
    case 2: format_records_kernel<2><<<grid, FMT_NT, 0, st>>>(a); break;
    default: format_records_kernel<3><<<grid, FMT_NT, 0, st>>>(a); break;
  }
  g_launches++;
  return cudaGetLastError();
}

}  // namespace c2
