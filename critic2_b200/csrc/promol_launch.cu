// promol_launch.cu -- instantiations of the promolecular kernel and the NCIPLOT cutoff pass (sm_100a).
#include "c2gpu_internal.h"

namespace c2 {

cudaError_t promol_launch(int nd, const PromolParams &p, int grid, cudaStream_t st) {
  if (grid <= 0) return cudaSuccess;
  const bool g = p.m.nspc > PROMOL_NSPC;   // species constants from global memory
  if (nd == 1) { if (g) promol_grid_kernel<1, true><<<grid, PROMOL_PT, 0, st>>>(p); else promol_grid_kernel<1, false><<<grid, PROMOL_PT, 0, st>>>(p); }
  else if (nd == 4) { if (g) promol_grid_kernel<4, true><<<grid, PROMOL_PT, 0, st>>>(p); else promol_grid_kernel<4, false><<<grid, PROMOL_PT, 0, st>>>(p); }
  else { if (g) promol_grid_kernel<10, true><<<grid, PROMOL_PT, 0, st>>>(p); else promol_grid_kernel<10, false><<<grid, PROMOL_PT, 0, st>>>(p); }
  g_launches++;
  return cudaGetLastError();
}

// ---- NCIPLOT cutoff pass (src/nci@proc.f90:569-604, nfrag = 0) -----------------------------------------------------
// HBM-bound streaming pass: 16-24 bytes read and up to 8 written per point.  Three kernels keep the reference's
// loop order for the (rho, s) pair list and make the sums reproducible: (1) per-block pair counts and partial sums
// (fixed-shape tree), (2) single-block exclusive scan of the counts + sequential sum of the partials,
// (3) ordered pair emission (ballot compaction) and the cgrad = 100 masking.
constexpr int CUT_NT = 256, CUT_PER = 8, CUT_BLK = CUT_NT * CUT_PER;

struct CutArgs {
  long npts;
  const double *crho;
  double *cgrad;
  const double *rhoat;
  double rhocut100, dimcut, rhoplot100, sr0, sr1, rho_void;
  int onlyneg;
  int nfrag;                 // fragments (:579-582): rhofrag is [nfrag][npts]
  const double *rhofrag;
  double rhoparam, rhoparam2;
};

__device__ __forceinline__ void cut_point(const CutArgs &a, long i, bool &isdat, bool &mask, double &q, double &qp,
                                          double &v, double &cr, double &cg) {
  cr = a.crho[i];
  cg = a.cgrad[i];
  bool inter = true;
  q = 0.0; qp = 0.0; v = 0.0;
  if (a.nfrag > 0) {                                                 // :579-582
    double s = 0.0;
    for (int f = 0; f < a.nfrag; f++) s += a.rhofrag[(size_t)f * a.npts + i];
    inter = s >= a.rhoparam2 * a.rhoat[i];
    for (int f = 0; f < a.nfrag; f++) inter = inter && (a.rhofrag[(size_t)f * a.npts + i] <= s * a.rhoparam);
  }
  if (a.rho_void > 0.0) {                                            // :583-590
    double ra = a.rhoat[i];
    inter = inter && ra < a.rho_void;
    if (ra < a.rho_void) { q = fabs(cr) / 100.0; qp = ra; v = 1.0; }
  }
  isdat = fabs(cr) < a.rhocut100 && cg < a.dimcut && inter;          // :593
  mask = fabs(cr) > a.rhoplot100 || !inter;                          // :596
  mask = mask || cr < a.sr0 || cr > a.sr1;                           // :597
  mask = mask || (cr > 0.0 && a.onlyneg);                            // :598
}

__global__ void __launch_bounds__(CUT_NT) cut_count_kernel(CutArgs a, unsigned *bcount, double *bsum) {
  __shared__ double s_sum[3][CUT_NT / 32];
  __shared__ unsigned s_c[CUT_NT / 32];
  const long base = (long)blockIdx.x * CUT_BLK;
  unsigned cnt = 0;
  double q = 0, qp = 0, v = 0;
  for (int k = 0; k < CUT_PER; k++) {
    long i = base + (long)k * CUT_NT + threadIdx.x;
    if (i < a.npts) {
      bool isdat, mask;
      double q1, qp1, v1, cr, cg;
      cut_point(a, i, isdat, mask, q1, qp1, v1, cr, cg);
      cnt += isdat;
      q += q1; qp += qp1; v += v1;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(~0u, cnt, o);
    q += __shfl_xor_sync(~0u, q, o);
    qp += __shfl_xor_sync(~0u, qp, o);
    v += __shfl_xor_sync(~0u, v, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_c[warp] = cnt; s_sum[0][warp] = q; s_sum[1][warp] = qp; s_sum[2][warp] = v; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned c = 0;
    double s0 = 0, s1 = 0, s2 = 0;
    for (int w = 0; w < CUT_NT / 32; w++) { c += s_c[w]; s0 += s_sum[0][w]; s1 += s_sum[1][w]; s2 += s_sum[2][w]; }
    bcount[blockIdx.x] = c;
    bsum[3 * (size_t)blockIdx.x] = s0; bsum[3 * (size_t)blockIdx.x + 1] = s1; bsum[3 * (size_t)blockIdx.x + 2] = s2;
  }
}

// one block: exclusive scan of bcount -> boff (64-bit), totals -> out[0..3] = ndat, sumq, sumqp, sumv
__global__ void __launch_bounds__(1024) cut_scan_kernel(int nblk, const unsigned *bcount, unsigned long long *boff,
                                                        const double *bsum, double *out) {
  __shared__ unsigned long long s_w[32];
  __shared__ unsigned long long s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    int i = base + threadIdx.x;
    unsigned long long c = i < nblk ? bcount[i] : 0, incl = c;
    for (int o = 1; o < 32; o <<= 1) {
      unsigned long long t = __shfl_up_sync(~0u, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    unsigned long long off = s_carry;
    for (int w = 0; w < warp; w++) off += s_w[w];
    if (i < nblk) boff[i] = off + incl - c;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = off + incl;
    __syncthreads();
  }
  // sums of the block partials in a fixed order (reproducible for a given grid): every thread adds a contiguous run of
  // blocks sequentially, lanes are folded by a butterfly, thread 0 adds the 32 warp sums in order
  __shared__ double s_p[3][32];
  const int chunk = (nblk + 1023) / 1024;
  double s0 = 0, s1 = 0, s2 = 0;
  for (int i = threadIdx.x * chunk; i < min(nblk, (threadIdx.x + 1) * chunk); i++) {
    s0 += bsum[3 * (size_t)i]; s1 += bsum[3 * (size_t)i + 1]; s2 += bsum[3 * (size_t)i + 2];
  }
  for (int o = 16; o > 0; o >>= 1) {
    s0 += __shfl_xor_sync(~0u, s0, o);
    s1 += __shfl_xor_sync(~0u, s1, o);
    s2 += __shfl_xor_sync(~0u, s2, o);
  }
  if (lane == 0) { s_p[0][warp] = s0; s_p[1][warp] = s1; s_p[2][warp] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    s0 = s1 = s2 = 0;
    for (int w = 0; w < 32; w++) { s0 += s_p[0][w]; s1 += s_p[1][w]; s2 += s_p[2][w]; }
    out[0] = (double)s_carry; out[1] = s0; out[2] = s1; out[3] = s2;
  }
}

__global__ void __launch_bounds__(CUT_NT) cut_emit_kernel(CutArgs a, const unsigned long long *boff, double *dat,
                                                          long maxdat) {
  __shared__ unsigned s_c[CUT_NT / 32];
  const long base = (long)blockIdx.x * CUT_BLK;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long off = boff[blockIdx.x];
  for (int k = 0; k < CUT_PER; k++) {
    long i = base + (long)k * CUT_NT + threadIdx.x;
    bool isdat = false, mask = false;
    double q1, qp1, v1, cr = 0, cg = 0;
    if (i < a.npts) cut_point(a, i, isdat, mask, q1, qp1, v1, cr, cg);
    unsigned m = __ballot_sync(~0u, isdat);
    if (lane == 0) s_c[warp] = __popc(m);
    __syncthreads();
    unsigned long long my = off;
    unsigned tot = 0;
    for (int w = 0; w < CUT_NT / 32; w++) {
      if (w < warp) my += s_c[w];
      tot += s_c[w];
    }
    my += __popc(m & ((1u << lane) - 1u));
    if (isdat && dat && (long)my < maxdat) { dat[2 * my] = cr / 100.0; dat[2 * my + 1] = cg; }   // :594
    if (mask && i < a.npts) a.cgrad[i] = 100.0;
    off += tot;
    __syncthreads();
  }
}

// per-device scratch of nci_cutoffs_launch, kept between calls (grow-only), released by c2gpu_finalize
static unsigned char *s_scr[64] = {nullptr};
static size_t s_cap[64] = {0};

void nci_cutoffs_release() {
  int prev = 0;
  cudaGetDevice(&prev);
  for (int d = 0; d < 64; d++)
    if (s_scr[d]) {
      cudaSetDevice(d);
      cudaFree(s_scr[d]);
      s_scr[d] = nullptr;
      s_cap[d] = 0;
    }
  cudaSetDevice(prev);
}

cudaError_t nci_cutoffs_launch(long npts, const double *crho, double *cgrad, const double *rhoat, int nfrag,
                               const double *rhofrag, double rhoparam, double rhoparam2, double rhocut,
                               double dimcut, double rhoplot, double sr0, double sr1, int onlyneg, double rho_void,
                               double *raw_sums, long *ndat, double *dat_dev, long maxdat, cudaStream_t st) {
  raw_sums[0] = raw_sums[1] = raw_sums[2] = 0.0;
  *ndat = 0;
  if (npts <= 0) return cudaSuccess;
  CutArgs a{npts, crho, cgrad, rhoat, rhocut * 100.0, dimcut, rhoplot * 100.0, sr0 * 100.0, sr1 * 100.0, rho_void, onlyneg,
            nfrag, rhofrag, rhoparam, rhoparam2};
  int nblk = (int)((npts + CUT_BLK - 1) / CUT_BLK);
  // one scratch allocation: out[4] | bsum[3*nblk] | boff[nblk] (64-bit) | bcount[nblk]
  // kept per device between calls (grow-only): cudaMalloc + cudaFree cost more than the pass itself on a 200^3 grid
  const size_t o_bsum = 4 * sizeof(double), o_boff = o_bsum + sizeof(double) * 3 * (size_t)nblk;
  const size_t o_bcount = o_boff + sizeof(unsigned long long) * (size_t)nblk;
  const size_t need = o_bcount + sizeof(unsigned) * (size_t)nblk;
  cudaError_t e;
  int dev = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (s_cap[dev] < need) {
    if (s_scr[dev]) cudaFree(s_scr[dev]);
    s_scr[dev] = nullptr; s_cap[dev] = 0;
    void *d = nullptr;
    if ((e = cudaMalloc(&d, need + need / 2)) != cudaSuccess) return e;
    s_scr[dev] = static_cast<unsigned char *>(d);
    s_cap[dev] = need + need / 2;
  }
  unsigned char *scr = s_scr[dev];
  double *out = reinterpret_cast<double *>(scr), *bsum = reinterpret_cast<double *>(scr + o_bsum);
  unsigned long long *boff = reinterpret_cast<unsigned long long *>(scr + o_boff);
  unsigned *bcount = reinterpret_cast<unsigned *>(scr + o_bcount);
  cut_count_kernel<<<nblk, CUT_NT, 0, st>>>(a, bcount, bsum);
  cut_scan_kernel<<<1, 1024, 0, st>>>(nblk, bcount, boff, bsum, out);
  cut_emit_kernel<<<nblk, CUT_NT, 0, st>>>(a, boff, dat_dev, maxdat);
  g_launches += 3;
  double h[4];
  e = cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return e;
  *ndat = (long)h[0];
  raw_sums[0] = h[1]; raw_sums[1] = h[2]; raw_sums[2] = h[3];
  return cudaGetLastError();
}

// ---- allocation-free variant of the cutoff pass (c2gpu_nci_cutoffs_dev_noalloc, include/c2gpu.h) ------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// The scratch of the three kernels above -- totals, per-block partial sums, per-block offsets, per-block counts -- lives in
// ONE file-scope device array sized for the largest grid the entry accepts; every GPU that loads the library holds these
// 2.25 MiB until exit.  In HOST code the array's name is only a stand-in for the symbol: its device address is asked for
// once per device (that device current) with cudaGetSymbolAddress and that pointer is what the kernels receive.
// Every call shares the array and finds the previous call's contents in it.  Nothing in it has to start at zero: a call
// with nblk blocks rewrites out[0..3] and entries [0, nblk) of the three per-block arrays before reading them, and reads
// nothing beyond nblk.  Two calls that overlap on one GPU (different streams, threads or processes sharing the context)
// race on it.
constexpr long NOALLOC_MAX_POINTS = 1L << 27;                                   // 134 217 728 = 512^3 (BASELINE config 5)
constexpr long NOALLOC_MAX_BLK = (NOALLOC_MAX_POINTS + CUT_BLK - 1) / CUT_BLK;   // 65 536 blocks of 2 048 points
constexpr size_t NOALLOC_O_BSUM = 256;                                           // out[4] in the first 256 bytes
constexpr size_t NOALLOC_O_BOFF = NOALLOC_O_BSUM + sizeof(double) * 3 * NOALLOC_MAX_BLK;
constexpr size_t NOALLOC_O_BCOUNT = NOALLOC_O_BOFF + sizeof(unsigned long long) * NOALLOC_MAX_BLK;
constexpr size_t NOALLOC_BYTES = NOALLOC_O_BCOUNT + sizeof(unsigned) * NOALLOC_MAX_BLK;   // 2 359 552 bytes
static __device__ __align__(256) unsigned char g_cut_scratch[NOALLOC_BYTES];

static cudaError_t noalloc_scratch_address(unsigned char **ptr) {
  static unsigned char *s_addr[64] = {nullptr};     // per device: the symbol has one instance (and address) per GPU
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!s_addr[dev]) {
    void *d = nullptr;
    if ((e = cudaGetSymbolAddress(&d, g_cut_scratch)) != cudaSuccess) return e;
    s_addr[dev] = static_cast<unsigned char *>(d);
  }
  *ptr = s_addr[dev];
  return cudaSuccess;
}

}  // namespace c2

extern "C" int c2gpu_nci_cutoffs_dev_noalloc(long npts, const double *crho_dev, double *cgrad_dev, const double *rhoat_dev,
                                             double rhocut, double dimcut, double rhoplot, const double srhorange[2],
                                             int onlyneg, double rho_void, double raw_sums[3], long *ndat, double *dat_dev,
                                             long maxdat, void *stream) {
  using namespace c2;
  if (!raw_sums || !ndat || !srhorange) return c2_fail(1, "c2gpu_nci_cutoffs_dev_noalloc: null argument");
  raw_sums[0] = raw_sums[1] = raw_sums[2] = 0.0;
  *ndat = 0;
  if (npts <= 0) return 0;
  if (npts > NOALLOC_MAX_POINTS)
    return c2_fail(1, "c2gpu_nci_cutoffs_dev_noalloc: " + std::to_string(npts) + " points exceed the static scratch limit of " +
                          std::to_string(NOALLOC_MAX_POINTS) + " (512^3); use c2gpu_nci_cutoffs_dev");
  if (!crho_dev || !cgrad_dev) return c2_fail(1, "c2gpu_nci_cutoffs_dev_noalloc: null device array");
  if (rho_void > 0.0 && !rhoat_dev) return c2_fail(1, "c2gpu_nci_cutoffs_dev_noalloc: VOID needs rhoat");
  unsigned char *scr = nullptr;
  cudaError_t e = noalloc_scratch_address(&scr);
  if (e != cudaSuccess) return c2_fail(2, std::string("cudaGetSymbolAddress(g_cut_scratch): ") + cudaGetErrorString(e));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CutArgs a{npts, crho_dev, cgrad_dev, rhoat_dev, rhocut * 100.0, dimcut, rhoplot * 100.0, srhorange[0] * 100.0,
            srhorange[1] * 100.0, rho_void, onlyneg, 0, nullptr, 0.95, 0.75};
  const int nblk = (int)((npts + CUT_BLK - 1) / CUT_BLK);
  double *out = reinterpret_cast<double *>(scr), *bsum = reinterpret_cast<double *>(scr + NOALLOC_O_BSUM);
  unsigned long long *boff = reinterpret_cast<unsigned long long *>(scr + NOALLOC_O_BOFF);
  unsigned *bcount = reinterpret_cast<unsigned *>(scr + NOALLOC_O_BCOUNT);
  cut_count_kernel<<<nblk, CUT_NT, 0, st>>>(a, bcount, bsum);
  cut_scan_kernel<<<1, 1024, 0, st>>>(nblk, bcount, boff, bsum, out);
  cut_emit_kernel<<<nblk, CUT_NT, 0, st>>>(a, boff, dat_dev, maxdat);
  g_launches += 3;
  double h[4];
  e = cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return c2_fail(2, std::string("c2gpu_nci_cutoffs_dev_noalloc: ") + cudaGetErrorString(e));
  *ndat = (long)h[0];
  raw_sums[0] = h[1]; raw_sums[1] = h[2]; raw_sums[2] = h[3];
  return 0;
}
