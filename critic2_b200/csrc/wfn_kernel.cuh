// wfn_kernel.cuh -- the wavefunction grid kernel: tile screening -> GTO primitives (FP64) -> chi x C contraction on
// the FP64 tensor cores (DMMA) -> fused rho / grad / Hessian / NCI epilogue.
//
// Replaces, for a whole lattice of points, the per-point chain field%grd -> molwfn%rho2 -> calculate_mo_gto
// (src/fieldmod@proc.f90:759-794,834-840; src/wfn_private@proc.F90:2032-2228, 2673-2781, 2783-2873).
//
// Work decomposition (persistent CTAs, dynamic scheduling over super-tiles of the lattice):
//   super-tile (e.g. 8x8x8 points)  : S1 ordered warp-ballot compaction of the shells whose screening sphere
//                                     (dran, :3102-3106) touches the super-tile's bounding box
//   sub-tile   (P = 8*WM points)    : S2 refine against the sub-tile box, expand to a primitive list
//   MO block   (NB = 8*NFRAG MOs)   : K loop over chunks of KC active primitives, double buffered:
//        A = chi[(d,point) x k]  evaluated in FP64 by all threads into shared memory (exact per-point dran mask)
//        B = C[k x mo]           gathered row by row with 1-D bulk TMA copies (mbarrier complete_tx)
//        acc(d,point; mo) += A*B mma.sync m8n8k4 f64; a thread owns all ND derivatives of one point for 2*NI MOs
//     block epilogue: occupation-weighted quadratic forms (:2091-2147) per thread, quad shuffle, fixed-order sum
//   point epilogue: EDF core, nucleus test, |grad|, Laplacian, 3x3 eigenvalues / RDG, coalesced property stores.
#pragma once
#include "epilogue.cuh"
#include "wfn_device.cuh"

namespace c2 {

struct WfnParams {
  WfnDev w;
  GridDesc g;
  CellXf xf;
  OutDesc o;
  int nblk;             // MO blocks (ldc = nblk*NB)
  int nsa, nsb, nsc;    // super-tile counts along (a,b) and layers of this part
  int nitems;           // work items (super-tiles) of this part
  int *work_counter;    // dynamic super-tile scheduler
  int *scratch;         // per CTA: shell list [shell_cap] then primitive list [prim_cap]
  int shell_cap, prim_cap;
  double nci_const;
  // warp-specialised kernel only: per-CTA cache of chi tiles in global memory (L2), re-used by MO blocks 1..nblk-1
  double *acache;       // [grid][2 tile parities][acache_chunks][A_STAGE]
  int *acache_nz;       // [grid][2][acache_chunks][8] k-quad masks
  double *ws_rec;       // [grid][shell_cap][12] records of the active shells of the current sub-tile
  int acache_chunks;
  // value-only (ND = 1) launches only.  bits 0-1 = k in 1..3: the basis value is replaced by its k-th Cartesian derivative,
  // so the pass returns sum_mo occ (d_k phi)^2 -- three such passes give the kinetic energy density G = res%gkin
  // (rho2, src/wfn_private@proc.F90:2104-2106,2209).  bit 2 (block-synchronous kernels): the MO sum is linear, sum_mo
  // occ phi -- with a one-MO handle that is the MO value of calculate_mo (:2447-2553).
  int chi_mode;
};

// NFRAG = 8-column MMA fragments per MO block, dealt to the WN warp columns: warp column wn owns
// NFRAG/WN (+1 for the first NFRAG%WN columns) consecutive fragments.
template <int ND_, int WM_, int WN_, int NFRAG_, int KC_, int CTAS_>
struct TileCfg {
  static constexpr int ND = ND_, WM = WM_, WN = WN_, NFRAG = NFRAG_, KC = KC_, CTAS = CTAS_;
  static constexpr int NI = (NFRAG + WN - 1) / WN;   // most fragments owned by one warp
  static constexpr int P = 8 * WM;            // points per sub-tile
  static constexpr int NB = 8 * NFRAG;        // MO columns per block
  static constexpr int NT = 32 * WM * WN;     // threads
  static constexpr int NW = WM * WN;
  static constexpr int SLOTS = NT / P;        // primitive slots evaluated concurrently (4*WN)
  static constexpr int R = KC / SLOTS;        // primitives per thread per chunk
  static constexpr int SB = ((NB + 11) / 16) * 16 + 4;  // B row stride (doubles), = 4 mod 16: conflict-free fragments
  static constexpr int A_STAGE = KC * ND * P;
  static constexpr int B_STAGE = KC * SB;
  static constexpr size_t SMEM = sizeof(double) * (2 * A_STAGE + 2 * B_STAGE + WN * P * ND + P * ND + 4 * P) +
                                 2 * (KC / 4) * WM * 32 + 16;
  static_assert(KC % SLOTS == 0 && KC % 4 == 0, "chunk must be a multiple of the slot count and of 4");
};

#ifdef __CUDACC__
// Bounding box of the valid points held by the block's threads -> bb[6] (lo xyz, hi xyz) in shared memory.
template <int NW>
__device__ __forceinline__ void block_bbox(double lo[3], double hi[3], double *s_red /*[NW*6]*/, double *bb) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int d = 0; d < 3; d++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[d] = fmin(lo[d], __shfl_xor_sync(~0u, lo[d], o));
      hi[d] = fmax(hi[d], __shfl_xor_sync(~0u, hi[d], o));
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < 3; d++) { s_red[warp * 6 + d] = lo[d]; s_red[warp * 6 + 3 + d] = hi[d]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    int d = threadIdx.x;
    double v = s_red[d];
    for (int w = 1; w < NW; w++) v = (d < 3) ? fmin(v, s_red[w * 6 + d]) : fmax(v, s_red[w * 6 + d]);
    bb[d] = v;
  }
  __syncthreads();
}

template <class C>
__global__ void __launch_bounds__(C::NT, C::CTAS) wfn_grid_kernel(const __grid_constant__ WfnParams p) {
  constexpr int ND = C::ND, P = C::P, NB = C::NB, NT = C::NT, NW = C::NW, WM = C::WM, WN = C::WN, NI = C::NI;
  constexpr int NFRAG = C::NFRAG;
  constexpr int KC = C::KC, SLOTS = C::SLOTS, R = C::R, SB = C::SB, A_STAGE = C::A_STAGE, B_STAGE = C::B_STAGE;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *As = reinterpret_cast<double *>(smem_raw);
  double *Bs = As + 2 * A_STAGE;
  double *s_part = Bs + 2 * B_STAGE;      // [WN][P][ND]
  double *s_tot = s_part + WN * P * ND;   // [P][ND]
  double *s_pt = s_tot + P * ND;          // [3][P] evaluation coordinates (wcr) + [P] validity
  unsigned char *s_nz = reinterpret_cast<unsigned char *>(s_pt + 4 * P);  // [2][KC/4][WM][32]
  __shared__ __align__(8) uint64_t mbar[2];
  __shared__ int s_cnt[NW + 1];
  __shared__ int s_int[8];                // 0 work item, 1 ns, 2 kact, 3 nnuc
  __shared__ int s_nuc[16];
  __shared__ double s_red[NW * 6];
  __shared__ double s_bb[12];             // super-tile box [0..5], sub-tile box [6..11]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp % WM, wn = warp / WM;
  const int pt = tid % P, slot = tid / P;
  // fragments of this warp column: [f0, f0 + nfw)
  const int nfw = NFRAG / WN + (wn < NFRAG % WN ? 1 : 0);
  const int f0 = wn * (NFRAG / WN) + min(wn, NFRAG % WN);
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  int *sh_list = p.scratch + (size_t)blockIdx.x * (p.shell_cap + p.prim_cap);
  int *pr_list = sh_list + p.shell_cap;
  const int na = grid_na(g), nb = grid_nb(g);

  if (tid == 0) {
    mbar_init(&mbar[0], 1);
    mbar_init(&mbar[1], 1);
    fence_mbar_init();
  }
  for (int i = tid; i < 2 * B_STAGE; i += NT) Bs[i] = 0.0;   // stale rows must stay finite (they meet chi = 0)
  fence_proxy_async();
  __syncthreads();
  uint32_t phase[2] = {0u, 0u};

  const int row = lane >> 2;
  const int aoff = row * 4 + ((lane & 3) ^ (((wm * 8 + row) >> 2) & 3));   // swizzled A fragment element
  const int npsuper = g.sa * g.sb * g.sc;
  const int nsub_a = g.sa / g.ta, nsub_b = g.sb / g.tb, nsub_c = g.sc / g.tc;

  for (;;) {
    __syncthreads();
    if (tid == 0) s_int[0] = atomicAdd(p.work_counter, 1);
    __syncthreads();
    const int st = s_int[0];
    if (st >= p.nitems) break;
    int ia, ib, ic;
    part_item(g, st, p.nsa, p.nsb, ia, ib, ic);
    const int sa0 = ia * g.sa, sb0 = ib * g.sb, sc0 = g.c0 + ic * g.sc;

    // ---- S1: super-tile bounding box of the (wrapped) evaluation points ---------------------------------------
    {
      double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
      for (int q = tid; q < npsuper; q += NT) {
        int a = sa0 + q % g.sa, b = sb0 + (q / g.sa) % g.sb, c = sc0 + q / (g.sa * g.sb);
        if (a < na && b < nb && c < g.c1) {
          double x[3], wc[3], wcr[3];
          grid_point(g, a, b, c, x);
          grd_wrap(p.xf, x, g.periodic, wc, wcr);
#pragma unroll
          for (int d = 0; d < 3; d++) { lo[d] = fmin(lo[d], wcr[d]); hi[d] = fmax(hi[d], wcr[d]); }
        }
      }
      block_bbox<NW>(lo, hi, s_red, s_bb);
    }
    // ordered compaction of the active shells (warp ballot + block prefix)
    if (tid == 0) s_int[3] = 0;
    __syncthreads();
    int ns = 0;
    for (int base = 0; base < w.nshell; base += NT) {
      int s = base + tid;
      bool act = false;
      if (s < w.nshell) {
        double4 xa = w.shell_xa[s];
        act = box_d2(s_bb, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12);
      }
      unsigned m = __ballot_sync(~0u, act);
      if (lane == 0) s_cnt[warp] = __popc(m);
      __syncthreads();
      int off = ns, tot = 0;
#pragma unroll
      for (int i = 0; i < NW; i++) {
        if (i < warp) off += s_cnt[i];
        tot += s_cnt[i];
      }
      if (act) sh_list[off + __popc(m & ((1u << lane) - 1u))] = s;
      ns += tot;
      __syncthreads();
    }
    // nucleus candidates (identify_atom with distmax = 1d-5, src/fieldmod@proc.f90:836)
    for (int i = tid; i < w.natom; i += NT) {
      if (box_d2(s_bb, w.atom_xyz[3 * i], w.atom_xyz[3 * i + 1], w.atom_xyz[3 * i + 2]) <= 1.0001e-10) {
        int k = atomicAdd(&s_int[3], 1);
        if (k < 16) s_nuc[k] = i;
      }
    }
    __syncthreads();
    const int nnuc = s_int[3];

    for (int sub = 0; sub < nsub_a * nsub_b * nsub_c; sub++) {
      const int ta0 = sa0 + (sub % nsub_a) * g.ta, tb0 = sb0 + ((sub / nsub_a) % nsub_b) * g.tb;
      const int tc0 = sc0 + (sub / (nsub_a * nsub_b)) * g.tc;
      if (ta0 >= na || tb0 >= nb || tc0 >= g.c1) continue;   // block-uniform
      // ---- points of the sub-tile ----------------------------------------------------------------------------
      {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        if (tid < P) {
          int a = ta0 + tid % g.ta, b = tb0 + (tid / g.ta) % g.tb, c = tc0 + tid / (g.ta * g.tb);
          bool valid = a < na && b < nb && c < g.c1;
          double x[3] = {0, 0, 0}, wc[3], wcr[3] = {0, 0, 0};
          if (valid) {
            grid_point(g, a, b, c, x);
            grd_wrap(p.xf, x, g.periodic, wc, wcr);
#pragma unroll
            for (int d = 0; d < 3; d++) { lo[d] = wcr[d]; hi[d] = wcr[d]; }
          }
          s_pt[tid] = wcr[0]; s_pt[P + tid] = wcr[1]; s_pt[2 * P + tid] = wcr[2];
          s_pt[3 * P + tid] = valid ? 1.0 : 0.0;
        }
        for (int i = tid; i < P * ND; i += NT) s_tot[i] = 0.0;
        block_bbox<NW>(lo, hi, s_red, s_bb + 6);
      }
      // ---- S2: refine the shell list against the sub-tile box, expand to primitives (ordered) ------------------
      int kact = 0;
      for (int base = 0; base < ns; base += NT) {
        int j = base + tid, cnt = 0, pb = 0;
        if (j < ns) {
          int s = sh_list[j];
          double4 xa = w.shell_xa[s];
          if (box_d2(s_bb + 6, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12)) {
            pb = w.shell_pbeg[s];
            cnt = w.shell_pbeg[s + 1] - pb;
          }
        }
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          int v = __shfl_up_sync(~0u, incl, o);
          if (lane >= o) incl += v;
        }
        if (lane == 31) s_cnt[warp] = incl;
        __syncthreads();
        int off = kact, tot = 0;
#pragma unroll
        for (int i = 0; i < NW; i++) {
          if (i < warp) off += s_cnt[i];
          tot += s_cnt[i];
        }
        off += incl - cnt;
        for (int i = 0; i < cnt; i++) pr_list[off + i] = pb + i;
        kact += tot;
        __syncthreads();
      }
      const int nch = (kact + KC - 1) / KC;
      const double px = s_pt[pt], py = s_pt[P + pt], pz = s_pt[2 * P + pt];
      const bool pvalid = s_pt[3 * P + pt] != 0.0;

      // ---- chi(c) -> A stage (all threads) -----------------------------------------------------------------------
      auto chi_phase = [&](int c, int stg) {
#pragma unroll
        for (int r = 0; r < R; r++) {
          const int k = slot + r * SLOTS, gk = c * KC + k;
          double chi[ND];
#pragma unroll
          for (int d = 0; d < ND; d++) chi[d] = 0.0;
          bool nz = false;
          const bool real = gk < kact && pvalid;
          const int prim = real ? pr_list[gk] : 0;
          const uchar4 l = real ? reinterpret_cast<const uchar4 *>(w.prim_l)[prim] : make_uchar4(0, 0, 0, 0);
          const bool big = __any_sync(~0u, (l.x | l.y | l.z) > 2);   // f shells and up: powers by multiplication
          if (real) {
            const int sh = w.prim_shell[prim];
            const double4 xa = w.shell_xa[sh];
            const double2 cut = w.shell_cut[sh];
            ShellAxis ax, ay;
            ax.x = px - xa.x; ay.x = py - xa.y;
            const double zr = pz - xa.z;
            ax.x2 = ax.x * ax.x; ay.x2 = ay.x * ay.x;
            const double z2 = zr * zr;
            const double d2 = __dadd_rn(__dadd_rn(ax.x2, ay.x2), z2);
            if (within_cut(ax.x, ay.x, zr, d2, cut.x, cut.y)) {      // :2747
              const double ta = xa.w + xa.w;
              ShellAxisE az;
              shell_axis_setup(ax, ta);
              shell_axis_setup(ay, ta);
              shell_axis_setup(az, zr, z2, ta, exp_neg(-xa.w * d2));  // :2749
              if (big) shell_prim_chi<ND, true>(ax, ay, az, l.x, l.y, l.z, ta, p.chi_mode, chi);   // :2755-2768
              else shell_prim_chi<ND, false>(ax, ay, az, l.x, l.y, l.z, ta, p.chi_mode, chi);
              nz = true;
            }
          }
          const int kq = k >> 2, k4 = k & 3;
          double *dst = As + stg * A_STAGE + (size_t)(kq * ND) * P * 4 + pt * 4 + (k4 ^ ((pt >> 2) & 3));
#pragma unroll
          for (int d = 0; d < ND; d++) dst[d * P * 4] = chi[d];
          s_nz[((stg * (KC / 4) + kq) * WM + (pt >> 3)) * 32 + (pt & 7) * 4 + k4] = nz ? 1 : 0;
        }
      };
      // ---- gather of C rows -> B stage (warp 0, one bulk TMA copy per primitive row) ------------------------------
      auto issue_b = [&](int c, int stg, int blk) {
        const int nrow = min(KC, kact - c * KC);
        if (lane == 0) mbar_expect_tx(&mbar[stg], (uint32_t)(nrow * NB * sizeof(double)));
        __syncwarp();
        for (int r = lane; r < nrow; r += 32) {
          const int prim = pr_list[c * KC + r];
          tma_bulk_g2s(Bs + stg * B_STAGE + r * SB, w.cmoT + (size_t)prim * w.ldc + (size_t)blk * NB,
                       (uint32_t)(NB * sizeof(double)), &mbar[stg]);
        }
      };

      if (kact > 0) {
        for (int blk = 0; blk < p.nblk; blk++) {
          double acc[ND][NI][2];
#pragma unroll
          for (int d = 0; d < ND; d++)
#pragma unroll
            for (int ni = 0; ni < NI; ni++) acc[d][ni][0] = acc[d][ni][1] = 0.0;
          if (warp == 0) issue_b(0, 0, blk);
          chi_phase(0, 0);
          for (int c = 0; c < nch; c++) {
            const int stg = c & 1;
            __syncthreads();   // chi(c) visible; every warp is done with stage stg^1
            if (c + 1 < nch && warp == 0) issue_b(c + 1, stg ^ 1, blk);
            mbar_wait(&mbar[stg], phase[stg]);
            phase[stg] ^= 1u;
            const double *Ast = As + stg * A_STAGE + (size_t)wm * 32 + aoff;
            const double *Bst = Bs + stg * B_STAGE + (lane & 3) * SB + f0 * 8 + (lane >> 2);
            const unsigned char *nzs = s_nz + ((stg * (KC / 4)) * WM + wm) * 32 + lane;
#pragma unroll 1
            for (int kq = 0; kq < KC / 4; kq++) {
              if (__ballot_sync(~0u, nzs[kq * WM * 32] != 0) == 0u) continue;   // all 8 points x 4 primitives masked
              double b[NI];
#pragma unroll
              for (int ni = 0; ni < NI; ni++) b[ni] = (NFRAG % WN == 0 || ni < nfw) ? Bst[kq * 4 * SB + ni * 8] : 0.0;
#pragma unroll
              for (int d = 0; d < ND; d++) {
                const double a = Ast[(size_t)(kq * ND + d) * P * 4];
#pragma unroll
                for (int ni = 0; ni < NI; ni++)
                  if (NFRAG % WN == 0 || ni < nfw) dmma884(acc[d][ni][0], acc[d][ni][1], a, b[ni]);
              }
            }
            if (c + 1 < nch) chi_phase(c + 1, stg ^ 1);
          }
          // ---- block epilogue: occupation-weighted sums over this thread's MOs (:2091-2147) -----------------------
          double q[ND];
#pragma unroll
          for (int d = 0; d < ND; d++) q[d] = 0.0;
#pragma unroll
          for (int ni = 0; ni < NI; ni++)
#pragma unroll
            for (int j = 0; j < 2; j++) {
              const double oc = (NFRAG % WN == 0 || ni < nfw) ? w.occ[blk * NB + (f0 + ni) * 8 + 2 * (lane & 3) + j] : 0.0;
              const double f0v = acc[0][ni][j];
              q[0] += (ND == 1 && (p.chi_mode & 4)) ? oc * f0v : oc * f0v * f0v;
              if (ND > 1) {
                const double t = 2.0 * oc;
                const double f1 = acc[ND > 1 ? 1 : 0][ni][j], f2 = acc[ND > 1 ? 2 : 0][ni][j], f3 = acc[ND > 1 ? 3 : 0][ni][j];
                q[ND > 1 ? 1 : 0] += t * f0v * f1;
                q[ND > 1 ? 2 : 0] += t * f0v * f2;
                q[ND > 1 ? 3 : 0] += t * f0v * f3;
                if (ND > 4) {
                  q[ND > 4 ? 4 : 0] += t * (f0v * acc[ND > 4 ? 4 : 0][ni][j] + f1 * f1);
                  q[ND > 4 ? 5 : 0] += t * (f0v * acc[ND > 4 ? 5 : 0][ni][j] + f2 * f2);
                  q[ND > 4 ? 6 : 0] += t * (f0v * acc[ND > 4 ? 6 : 0][ni][j] + f3 * f3);
                  q[ND > 4 ? 7 : 0] += t * (f0v * acc[ND > 4 ? 7 : 0][ni][j] + f1 * f2);
                  q[ND > 4 ? 8 : 0] += t * (f0v * acc[ND > 4 ? 8 : 0][ni][j] + f1 * f3);
                  q[ND > 4 ? 9 : 0] += t * (f0v * acc[ND > 4 ? 9 : 0][ni][j] + f2 * f3);
                }
              }
            }
#pragma unroll
          for (int d = 0; d < ND; d++) {
            q[d] += __shfl_xor_sync(~0u, q[d], 1);
            q[d] += __shfl_xor_sync(~0u, q[d], 2);
          }
          if ((lane & 3) == 0) {
#pragma unroll
            for (int d = 0; d < ND; d++) s_part[(wn * P + wm * 8 + row) * ND + d] = q[d];
          }
          __syncthreads();   // also: every warp is out of the K loop before the stages are reused
          for (int i = tid; i < P * ND; i += NT) {
            double s = s_tot[i];
            for (int v = 0; v < WN; v++) s += s_part[v * P * ND + i];
            s_tot[i] = s;
          }
        }
      }
      __syncthreads();
      // ---- EDF core density (calculate_edf, :2783-2873): fixed-order partial sums through the free A stage ---------
      if (w.nedf > 0) {
        double e[ND];
#pragma unroll
        for (int d = 0; d < ND; d++) e[d] = 0.0;
        if (pvalid) {
          for (int i = slot; i < w.nedf; i += SLOTS) {
            const double4 xa = w.edf_xa[i];
            const double4 cc = w.edf_cc[i];
            const double dx = px - xa.x, dy = py - xa.y, dz = pz - xa.z;
            const double d2 = dx * dx + dy * dy + dz * dz;
            if (!within_cut(dx, dy, dz, d2, cc.y, cc.z)) continue;
            const uchar4 l = reinterpret_cast<const uchar4 *>(w.edf_l)[i];
            double chi[ND];
            prim_chi<ND>(dx, dy, dz, l.x, l.y, l.z, xa.w, exp(-xa.w * d2), chi);
#pragma unroll
            for (int d = 0; d < ND; d++) e[d] += cc.x * chi[d];
          }
        }
#pragma unroll
        for (int d = 0; d < ND; d++) As[(slot * P + pt) * ND + d] = e[d];
        __syncthreads();
        for (int i = tid; i < P * ND; i += NT) {
          double s = s_tot[i];
          for (int v = 0; v < SLOTS; v++) s += As[v * P * ND + i];
          s_tot[i] = s;
        }
        __syncthreads();
      }
      // ---- point epilogue -------------------------------------------------------------------------------------------
      if (tid < P && s_pt[3 * P + tid] != 0.0) {
        const int a = ta0 + tid % g.ta, b = tb0 + (tid / g.ta) % g.tb, c = tc0 + tid / (g.ta * g.tb);
        const double *t = s_tot + tid * ND;
        double f = t[0], gx = 0, gy = 0, gz = 0, hs[6] = {0, 0, 0, 0, 0, 0};
        if (ND > 1) { gx = t[ND > 1 ? 1 : 0]; gy = t[ND > 1 ? 2 : 0]; gz = t[ND > 1 ? 3 : 0]; }
        if (ND > 4) {
#pragma unroll
          for (int d = 0; d < 6; d++) hs[d] = t[ND > 4 ? 4 + d : 0];
        }
        bool isnuc = false;
        const double x = s_pt[tid], y = s_pt[P + tid], z = s_pt[2 * P + tid];
        if (nnuc > 16) {
          for (int i = 0; i < w.natom && !isnuc; i++) {
            double ddx = w.atom_xyz[3 * i] - x, ddy = w.atom_xyz[3 * i + 1] - y, ddz = w.atom_xyz[3 * i + 2] - z;
            isnuc = (ddx * ddx + ddy * ddy + ddz * ddz) < 1e-10;      // src/environmod@proc.f90:677
          }
        } else {
          for (int k = 0; k < nnuc; k++) {
            int i = s_nuc[k];
            double ddx = w.atom_xyz[3 * i] - x, ddy = w.atom_xyz[3 * i + 1] - y, ddz = w.atom_xyz[3 * i + 2] - z;
            isnuc = isnuc || (ddx * ddx + ddy * ddy + ddz * ddz) < 1e-10;
          }
        }
        emit_point(p.o, out_index(g, a, b, c), f, gx, gy, gz, hs[0], hs[1], hs[2], hs[3], hs[4], hs[5], isnuc, p.nci_const);
      }
      __syncthreads();   // the epilogue threads are done with s_tot / s_pt before the next sub-tile overwrites them
    }
  }
}
#endif  // __CUDACC__

}  // namespace c2
