// wfn_launch.cu -- instantiations and launch dispatch of the wavefunction grid kernel (sm_100a).
#include "c2gpu_internal.h"
#include "wfn_ws_kernel.cuh"
#include "wfn_warp_kernel.cuh"

namespace c2 {

long g_launches = 0;

// name, WM, WN, NFRAG, KC, CTAS
#define C2_WFN_CONFIGS(X) \
  X(0, "S1", 8, 1, 1, 8, 2)   \
  X(1, "S3", 8, 1, 3, 8, 1)   \
  X(2, "M", 2, 4, 8, 16, 1)   \
  X(3, "L", 2, 7, 14, 28, 1)  \
  X(4, "L1", 1, 7, 14, 28, 2) \
  X(5, "L16", 2, 8, 14, 32, 1)
// M2 = M with two CTAs per SM (128 registers, a few spills): while one CTA sits in the latency-bound chi phase or at its
// chunk barrier the other one feeds the DMMA pipe.  60-atom molecule, 160^3: 57.3 ms per 64-column block against 72.7 ms
// for M (48 MOs: 37 -> 47 % of the DMMA peak; 64: 48 -> 62 %; 128: 49 -> 62 %).  Three CTAs (85 registers) or eight warp
// columns of one fragment (64 registers) spill heavily and are 1.6-3x slower.
// id after the warp-specialised kernel (id 6): 9..24 MOs with the 24 columns spread over three warp columns, so a thread
// holds 20 accumulators instead of 60 and two CTAs (24 warps) are resident per SM instead of one (8 warps); measured on
// benzene NCIPLOT 200^3: 20.6 ms against 22.0 ms for S3 (P = 64 and P = 16 variants of the same idea were slower: P = 16
// with four CTAs per SM 22.8 ms, P = 32 with three CTAs (56 registers, spilling) 33.6 ms; tools/probe_nci.py)
#define C2_WFN_CONFIGS2(X) \
  X(7, "S3w", 4, 3, 3, 12, 2) \
  X(8, "M2", 2, 4, 8, 16, 2)

// ids after the warp-specialised variants (12..14): M2 with 32-column blocks for 25..32 MOs (one fragment per warp column;
// M2 pads them to 64 columns); with two and with three CTAs per SM
#define C2_WFN_CONFIGS3(X)   \
  X(15, "M2h", 2, 4, 4, 16, 2) \
  X(16, "M2h3", 2, 4, 4, 16, 3) \
  X(17, "M2t", 2, 3, 6, 12, 2)

// warp-autonomous kernel (ids 9, 10, 11): NF fragments of 8 MO columns, CTAs per SM (register budget: 65536 / (128 x CTAS))
// Since the shell-wise chi code (fewer live registers) one more CTA per SM pays for two and three fragments in spite of ~140
// bytes of spills: 3 fragments (benzene NCIPLOT 200^3) 13.59 -> 12.59 ms with three CTAs at 168 registers, 2 fragments 20.56 ->
// 19.69 ms with four at 128; one fragment stays at four (five: 1.19 -> 1.32 ms on h2o 128^3)
#define C2_WARP_CONFIGS(X) \
  X(1, 4)                  \
  X(2, 4)                  \
  X(3, 3)

static void tile_shape(int P, int &ta, int &tb, int &tc) {
  switch (P) {
    case 8: ta = 2; tb = 2; tc = 2; break;
    case 16: ta = 4; tb = 2; tc = 2; break;
    case 32: ta = 4; tb = 4; tc = 2; break;
    default: ta = 4; tb = 4; tc = 4; break;
  }
}

static std::vector<WfnCfgInfo> build_infos() {
  std::vector<WfnCfgInfo> v;
#define X(ID, NAME, WM, WN, NF, KC, CTAS)                                                            \
  {                                                                                                  \
    WfnCfgInfo i;                                                                                    \
    i.name = NAME; i.wm = WM; i.wn = WN; i.ni = NF; i.kc = KC; i.ctas = CTAS;                        \
    i.P = 8 * WM; i.NB = 8 * NF; i.NT = 32 * WM * WN;                                           \
    tile_shape(i.P, i.ta, i.tb, i.tc);                                                               \
    i.smem[0] = TileCfg<1, WM, WN, NF, KC, CTAS>::SMEM;                                              \
    i.smem[1] = TileCfg<4, WM, WN, NF, KC, CTAS>::SMEM;                                              \
    i.smem[2] = TileCfg<10, WM, WN, NF, KC, CTAS>::SMEM;                                             \
    v.push_back(i);                                                                                  \
  }
  C2_WFN_CONFIGS(X)
#undef X
  {  // id 6: warp-specialised producer/consumer kernel (wfn_ws_kernel.cuh)
    WfnCfgInfo i;
    i.name = "WS"; i.wm = 2; i.wn = 4; i.ni = 14; i.kc = WsCfg<10>::KC; i.ctas = 1;
    i.P = 16; i.NB = 112; i.NT = WsCfg<10>::NT;
    tile_shape(i.P, i.ta, i.tb, i.tc);
    i.smem[0] = WsCfg<1>::SMEM; i.smem[1] = WsCfg<4>::SMEM; i.smem[2] = WsCfg<10>::SMEM;
    v.push_back(i);
  }
#define X(ID, NAME, WM, WN, NF, KC, CTAS)                                                            \
  {                                                                                                  \
    WfnCfgInfo i;                                                                                    \
    i.name = NAME; i.wm = WM; i.wn = WN; i.ni = NF; i.kc = KC; i.ctas = CTAS;                        \
    i.P = 8 * WM; i.NB = 8 * NF; i.NT = 32 * WM * WN;                                           \
    tile_shape(i.P, i.ta, i.tb, i.tc);                                                               \
    i.smem[0] = TileCfg<1, WM, WN, NF, KC, CTAS>::SMEM;                                              \
    i.smem[1] = TileCfg<4, WM, WN, NF, KC, CTAS>::SMEM;                                              \
    i.smem[2] = TileCfg<10, WM, WN, NF, KC, CTAS>::SMEM;                                             \
    v.push_back(i);                                                                                  \
  }
  C2_WFN_CONFIGS2(X)
#undef X
  // ids 9, 10, 11: warp-autonomous kernel (wfn_warp_kernel.cuh) with 1, 2, 3 fragments of 8 MO columns
#define W(NF, CT)                                                                                   \
  {                                                                                                 \
    WfnCfgInfo i;                                                                                   \
    i.name = "W" #NF; i.wm = 1; i.wn = 4; i.ni = NF; i.kc = 4; i.ctas = CT;                        \
    i.P = 8; i.NB = 8 * NF; i.NT = WarpCfg<10, NF, CT>::NT;                                         \
    tile_shape(i.P, i.ta, i.tb, i.tc);                                                              \
    i.smem[0] = i.smem[1] = i.smem[2] = 0;                                                          \
    v.push_back(i);                                                                                 \
  }
  C2_WARP_CONFIGS(W)
#undef W
  // id 12: the warp-specialised kernel with 96-column MO blocks: 3 + 3 fragments per SM sub-partition instead of the 4 + 3
  // of the 112-column blocks (a lone 4-fragment warp cannot keep the DMMA pipe full while its 3-fragment partner waits at
  // the stage barrier: tools/dmma_issue.cu).  128-column blocks (4 + 4) do not fit the register file (2 x 232 + 88 > 512).
  {
    WfnCfgInfo i = v[6];
    i.name = "WS96"; i.ni = 12; i.NB = 96;
    i.smem[0] = WsCfg<1, 12>::SMEM; i.smem[1] = WsCfg<4, 12>::SMEM; i.smem[2] = WsCfg<10, 12>::SMEM;
    v.push_back(i);
  }
  // id 13: the warp-specialised kernel for ONE TO FOUR MO blocks: 64-column blocks (two fragments per consumer warp), seven
  // generator warps instead of three (512 threads, 128 registers each)
  {
    WfnCfgInfo i = v[6];
    using W = WsCfg<10, 8, 32, 3, 8>;
    i.name = "WS64g"; i.ni = 8; i.NB = 64; i.NT = W::NT;
    i.smem[0] = WsCfg<1, 8, 32, 3, 8>::SMEM; i.smem[1] = WsCfg<4, 8, 32, 3, 8>::SMEM; i.smem[2] = W::SMEM;
    v.push_back(i);
  }
  // id 14: id 6 with the consumer loop of the balanced variants and paired chi rows (WsCfg::DENSE = 2) for systems with many MO
  // blocks per sub-tile
  {
    WfnCfgInfo i = v[6];
    i.name = "WSd";
    v.push_back(i);
  }
#define X(ID, NAME, WM, WN, NF, KC, CTAS)                                                            \
  {                                                                                                  \
    WfnCfgInfo i;                                                                                    \
    i.name = NAME; i.wm = WM; i.wn = WN; i.ni = NF; i.kc = KC; i.ctas = CTAS;                        \
    i.P = 8 * WM; i.NB = 8 * NF; i.NT = 32 * WM * WN;                                                \
    tile_shape(i.P, i.ta, i.tb, i.tc);                                                               \
    i.smem[0] = TileCfg<1, WM, WN, NF, KC, CTAS>::SMEM;                                              \
    i.smem[1] = TileCfg<4, WM, WN, NF, KC, CTAS>::SMEM;                                              \
    i.smem[2] = TileCfg<10, WM, WN, NF, KC, CTAS>::SMEM;                                             \
    v.push_back(i);                                                                                  \
  }
  C2_WFN_CONFIGS3(X)
#undef X
  return v;
}

static const std::vector<WfnCfgInfo> &infos() {
  static std::vector<WfnCfgInfo> v = build_infos();
  return v;
}
int wfn_num_configs() { return (int)infos().size(); }
const WfnCfgInfo &wfn_cfg_info(int cfg) { return infos()[cfg]; }

int wfn_choose_config(int nmo, int force) {
  if (force >= 0 && force < wfn_num_configs()) return force;
  int nfrag = (nmo + 7) / 8;
  // up to 24 MOs: the warp-autonomous kernel (wfn_warp_kernel.cuh).  Measured on BASELINE config 1 (h2o-aqz 128^3, 5 MOs):
  // 1.17 ms against 1.46 ms for S1; config 2 (benzene-adz NCIPLOT 200^3, 21 MOs): 13.4 ms against 17.3 ms for S3w
  if (nfrag <= 3) return 8 + nfrag;
  // 25..48 MOs: M2 with 32- / 48-column blocks instead of padding to 64 (60-atom molecule, 160^3: 36.4 ms for 25-32 MOs with
  // three CTAs per SM and 47.7 ms for 33-48 MOs against 55.7 ms for M2)
  if (nmo <= 32) return 16;
  if (nmo <= 48) return 17;
  // Measured cost model (60-atom molecule, 1770 primitives, 160^3; ms; tools/probe_mid.py, profiles/r2_probe_mid.log):
  //   blocks          1      2      3      4      5
  //   M2    (64)     57.2  110.5  163.7  216.9  269.9    53.5 per block + 3.5
  //   L1    (112)   102.6  198.6  294.3                  96 per block + 6
  //   WS    (112)   152.3  202.3  252.6                  102.3 + 50.1 per block (generators ahead, chi tiles re-used from L2)
  //   WS64g (64)     66.5   92.1  128.9  168.8  208.6    generator-bound for one block, then 9 + 40 per block
  // (first version of WS64g: 92.6 / 119.5 / 157.4 / 198.2 / 239.7.)  WS64g wins from 65 to 384 MOs: 128 MOs 74.2 % of the
  // DMMA peak, 192: 78.7 %, 256: 79.7 %, 300: 75.5 %; 500-atom config 4 (780 MOs): WS 89.8 % of the peak, L 62 %.
  const int bM = (nmo + 63) / 64, bL = (nmo + 111) / 112;
  const double tM = 53.5 * bM + 3.5, tL = 96 * bL + 6, tW = 102.3 + 50.1 * bL;
  const double tG = bM <= 1 ? 66.5 : (bM == 2 ? 92.1 : 9 + 40.0 * bM);
  // the model holds to 1 % up to nine 64-column blocks (7: 288.5 ms measured / 289 model, 9: 368.2 / 369; WS with 4 / 5 blocks:
  // 301.8 / 352.1), so WS64g also takes 385-512 and 561-576 MOs
  if (bM <= 9 && tG < tM && tG < tL && tG < tW) return 13;
  if (tM <= tL && tM <= tW) return 8;
  if (tL <= tW) return 4;
  // from nine 112-column blocks on the generators have slack to give and the denser consumer loop pays (WSd, id 14); bench
  // molecule, central slab, % of the DMMA peak for 7 / 8 / 9 / 10 / 14 blocks: WS 88.8 / 89.6 / 89.7 / 89.7 / 89.3, WSd 87.6 /
  // 89.6 / 90.8 / 91.0 / 90.6; the 28 blocks of BASELINE config 5: 89.6 -> 90.9
  return bL >= 9 ? 14 : 6;
}

template <class C>
static cudaError_t launch_one(const WfnParams &p, int grid, cudaStream_t st) {
  static bool attr_set[16] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 16 && !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(wfn_grid_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  wfn_grid_kernel<C><<<grid, C::NT, C::SMEM, st>>>(p);
  g_launches++;
  return cudaGetLastError();
}

template <class C>
static cudaError_t launch_ws(const WfnParams &p, int grid, cudaStream_t st) {
  static bool attr_set[16] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 16 && !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(wfn_ws_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  wfn_ws_kernel<C><<<grid, C::NT, C::SMEM, st>>>(p);
  g_launches++;
  return cudaGetLastError();
}

template <class C>
static cudaError_t launch_warp(const WfnParams &p, int grid, cudaStream_t st) {
  wfn_warp_kernel<C><<<grid, C::NT, 0, st>>>(p);
  g_launches++;
  return cudaGetLastError();
}

cudaError_t wfn_launch(int cfg, int nd, const WfnParams &p, int grid, cudaStream_t st) {
#define W(NF, CT)                                                        \
  if (cfg == 8 + NF) {                                                   \
    if (nd == 1) return launch_warp<WarpCfg<1, NF, CT>>(p, grid, st);    \
    if (nd == 4) return launch_warp<WarpCfg<4, NF, CT>>(p, grid, st);    \
    return launch_warp<WarpCfg<10, NF, CT>>(p, grid, st);                \
  }
  C2_WARP_CONFIGS(W)
#undef W
  if (cfg == 12) {
    if (nd == 1) return launch_ws<WsCfg<1, 12>>(p, grid, st);
    if (nd == 4) return launch_ws<WsCfg<4, 12>>(p, grid, st);
    return launch_ws<WsCfg<10, 12>>(p, grid, st);
  }
  if (cfg == 13) {
    if (nd == 1) return launch_ws<WsCfg<1, 8, 32, 3, 8>>(p, grid, st);
    if (nd == 4) return launch_ws<WsCfg<4, 8, 32, 3, 8>>(p, grid, st);
    return launch_ws<WsCfg<10, 8, 32, 3, 8>>(p, grid, st);
  }
  if (cfg == 14) {
    if (nd == 1) return launch_ws<WsCfg<1, 14, 32, 3, 4, 2>>(p, grid, st);
    if (nd == 4) return launch_ws<WsCfg<4, 14, 32, 3, 4, 2>>(p, grid, st);
    return launch_ws<WsCfg<10, 14, 32, 3, 4, 2>>(p, grid, st);
  }
  if (cfg == 6) {
    if (nd == 1) return launch_ws<WsCfg<1>>(p, grid, st);
    if (nd == 4) return launch_ws<WsCfg<4>>(p, grid, st);
    return launch_ws<WsCfg<10>>(p, grid, st);
  }
#define X(ID, NAME, WM, WN, NF, KC, CTAS)                                          \
  if (cfg == ID) {                                                                 \
    if (nd == 1) return launch_one<TileCfg<1, WM, WN, NF, KC, CTAS>>(p, grid, st); \
    if (nd == 4) return launch_one<TileCfg<4, WM, WN, NF, KC, CTAS>>(p, grid, st); \
    return launch_one<TileCfg<10, WM, WN, NF, KC, CTAS>>(p, grid, st);             \
  }
  C2_WFN_CONFIGS(X)
  C2_WFN_CONFIGS2(X)
  C2_WFN_CONFIGS3(X)
#undef X
  return cudaErrorInvalidValue;
}

// ---- diagnostic: sum over points of n_act(p), the reference's own per-point primitive screen (:2747) ------------
__global__ void __launch_bounds__(256) wfn_count_kernel(WfnDev w, GridDesc g, CellXf xf, unsigned long long *count) {
  __shared__ double4 s_xa[256];
  __shared__ double2 s_cut[256];
  __shared__ int s_np[256];
  const int na = grid_na(g), nb = grid_nb(g);
  const size_t npl = (size_t)na * nb;
  const size_t q = (size_t)blockIdx.x * 256 + threadIdx.x;
  const size_t ntot = npl * (size_t)(g.c1 - g.c0);
  const bool valid = q < ntot;
  double wcr[3] = {0, 0, 0};
  if (valid) {
    int c = g.c0 + (int)(q / npl), b = (int)((q % npl) / na), a = (int)(q % na);
    double x[3], wc[3];
    grid_point(g, a, b, c, x);
    grd_wrap(xf, x, g.periodic, wc, wcr);
  }
  unsigned long long n = 0;
  for (int base = 0; base < w.nshell; base += 256) {
    int s = base + threadIdx.x;
    __syncthreads();
    if (s < w.nshell) {
      s_xa[threadIdx.x] = w.shell_xa[s];
      s_cut[threadIdx.x] = w.shell_cut[s];
      s_np[threadIdx.x] = w.shell_pbeg[s + 1] - w.shell_pbeg[s];
    }
    __syncthreads();
    int m = min(256, w.nshell - base);
    if (valid)
      for (int i = 0; i < m; i++) {
        double dx = wcr[0] - s_xa[i].x, dy = wcr[1] - s_xa[i].y, dz = wcr[2] - s_xa[i].z;
        double d2 = dx * dx + dy * dy + dz * dz;
        if (within_cut(dx, dy, dz, d2, s_cut[i].x, s_cut[i].y)) n += s_np[i];
      }
  }
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(~0u, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(count, n);
}

// ---- second pass of the two-pass NCIPLOT evaluation: eigenvalues, RDG and sign(lambda2)*rho (src/nci@proc.f90:523-525,
// 553-554) of every point the first pass owned, one thread per point ------------------------------------------------
__global__ void __launch_bounds__(256) nci_epilogue_kernel(GridDesc g, OutDesc o, const double *raw, size_t stride,
                                                            double nci_const) {
  const int na = grid_na(g), nb = grid_nb(g);
  const size_t npl = (size_t)na * nb;
  const size_t q = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (q >= npl * (size_t)(g.c1 - g.c0)) return;
  const int c = g.c0 + (int)(q / npl), b = (int)((q % npl) / na);
  if (!tile_owned(g, b, c)) return;
  const size_t idx = q + (size_t)g.c0 * npl;
  const double *r = raw + idx;
  emit_point(o, idx, r[0], r[stride], r[2 * stride], r[3 * stride], r[4 * stride], r[5 * stride], r[6 * stride],
             r[7 * stride], r[8 * stride], r[9 * stride], false, nci_const);
}

cudaError_t nci_epilogue_launch(const GridDesc &g, const OutDesc &o, const double *raw, size_t stride, double nci_const,
                                cudaStream_t st) {
  const size_t ntot = (size_t)grid_na(g) * grid_nb(g) * (size_t)(g.c1 - g.c0);
  if (ntot == 0) return cudaSuccess;
  nci_epilogue_kernel<<<(unsigned)((ntot + 255) / 256), 256, 0, st>>>(g, o, raw, stride, nci_const);
  g_launches++;
  return cudaGetLastError();
}

// ---- diagnostic: FP64 peaks of this device, the roofline denominators of bench.py (the same chains as tools/fp64_peak.cu:
// 16 independent m8n8k4 DMMA accumulator pairs / 16 independent DFMA chains per thread, 2 CTAs of 256 threads per SM) -------
__global__ void __launch_bounds__(256) fp64_peak_dmma_kernel(double *out, int iters, double a, double b) {
  double c[16][2];
#pragma unroll
  for (int i = 0; i < 16; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) fp64_peak_dfma_kernel(double *out, int iters, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

cudaError_t fp64_peak_measure(double *dmma_tflops, double *dfma_tflops) {
  int dev = 0, nsm = 0;
  cudaGetDevice(&dev);
  cudaError_t e = cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return e;
  const int grid = nsm * 2, iters = 20000;
  double *out = nullptr;
  if ((e = cudaMalloc(&out, sizeof(double) * 256 * (size_t)grid)) != cudaSuccess) return e;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best[2] = {1e30f, 1e30f};
  for (int rep = 0; rep < 4; rep++)          // first repetition = warm-up
    for (int k = 0; k < 2; k++) {
      cudaEventRecord(e0);
      if (k == 0) fp64_peak_dmma_kernel<<<grid, 256>>>(out, iters, 1.0000001, 1e-9);
      else fp64_peak_dfma_kernel<<<grid, 256>>>(out, iters, 1.0000001, 1e-9);
      g_launches++;
      cudaEventRecord(e1);
      if ((e = cudaEventSynchronize(e1)) != cudaSuccess) break;
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep > 0 && ms < best[k]) best[k] = ms;
    }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  if (e != cudaSuccess) return e;
  // one warp-level m8n8k4 DMMA = 2*8*8*4 flops -> 2*256*... per thread: 16 per lane
  *dmma_tflops = 2.0 * 16 * 256 * (double)iters * 8.0 * grid / best[0] / 1e9;
  *dfma_tflops = 2.0 * 16 * (double)iters * 256.0 * grid / best[1] / 1e9;
  return cudaSuccess;
}

cudaError_t wfn_count_launch(const WfnDev &w, const GridDesc &g, const CellXf &xf, unsigned long long *count,
                             cudaStream_t st) {
  size_t ntot = (size_t)grid_na(g) * grid_nb(g) * (size_t)(g.c1 - g.c0);
  if (ntot == 0) return cudaSuccess;
  wfn_count_kernel<<<(unsigned)((ntot + 255) / 256), 256, 0, st>>>(w, g, xf, count);
  g_launches++;
  return cudaGetLastError();
}

}  // namespace c2

#if C2_WS_DIAG & 16
// development only (tools/ws_diag.sh): per (CTA, consumer warp) {cycles waiting for a full stage, cycles in the MMA loop,
// DMMA warp instructions, cycles alive, wait at the first stage of a sub-tile, cycles in the block / sub-tile epilogues,
// stages, sub-tiles} of the last wfn_ws_kernel launch on the current device
extern "C" int c2gpu_ws_gdiag_read(unsigned long long *out, int n) {   // generator warps 0/1: {wait free, S2, chi, alive}
  return cudaMemcpyFromSymbol(out, g_ws_gdiag, sizeof(unsigned long long) * (size_t)(n < 160 * 8 ? n : 160 * 8)) != cudaSuccess;
}
extern "C" int c2gpu_ws_diag_read(unsigned long long *out, int n) {
  return cudaMemcpyFromSymbol(out, g_ws_diag, sizeof(unsigned long long) * (size_t)(n < 160 * 64 ? n : 160 * 64)) != cudaSuccess;
}
#endif
