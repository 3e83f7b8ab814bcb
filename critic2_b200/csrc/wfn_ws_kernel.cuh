// wfn_ws_kernel.cuh -- warp-specialised version of the wavefunction grid kernel for large MO counts.
//
// Same mathematics and data layout as wfn_kernel.cuh (see there for the reference citations), different execution
// model.  ncu on the first version showed the FP64 tensor pipe (DMMA) idle ~38 % of the time because every warp
// alternated between latency-bound primitive evaluation and MMAs behind one block-wide barrier per chunk.  Here
//   * 4 PRODUCER warps (one per SM sub-partition) own everything that is not an MMA: one STREAMER (bulk-TMA copies of the
//     chi tiles and of the C rows into the stages) and three GENERATORS (super-tile claiming, shell screening S1/S2, the chi
//     tiles, evaluated shell by shell one sub-tile ahead into a per-CTA cache in L2);
//   * 8 CONSUMER warps (two per sub-partition, 4+3 MMA fragments each so that every sub-partition carries 7 of the
//     14 fragments of a 112-column MO block) only wait on mbarriers, issue DMMAs and run the fused epilogue;
//   * registers are re-balanced with setmaxnreg (consumers 232 / 184, producers 88): the consumers hold 10x4x2 accumulators;
//   * stages hand over through full/empty mbarriers (no block-wide barrier in steady state), and the stream of
//     (sub-tile, MO block, chunk) steps is continuous, so producers run ahead across block and tile boundaries.
// Variants (WsCfg): 64-column blocks with seven generator warps for a few MO blocks (WS64g), the dense consumer loop with
// paired chi rows for many MO blocks (WSd), 96-column blocks (WS96, kept for measurements).  What limits the kernel is the
// FP64 unit's arbitration between the consumers' DMMAs and the generators' ordinary FP64 work: DESIGN.md section 4.
#pragma once
#include "wfn_kernel.cuh"
#include <type_traits>

// Development diagnostics (tools/ws_diag.sh builds separate libraries with -DC2_WS_DIAG=bits; the shipped library has 0 and
// none of this code).  Results of a diagnostic build are NOT valid field values: the bits remove work to see what the
// tensor pipe is waiting for.  1: the streamer issues no TMA copies (stages complete at once; consumers multiply stale
// shared memory).  2: the generators skip the chi arithmetic (zeros stored, masks kept).  4: (retired).  16: consumer warps accumulate
// clock64() spent waiting for a full stage / inside the MMA loop and the DMMA count into g_ws_diag (c2gpu_ws_diag_read).  32: the generators
// do not store the chi tiles.  64: no exp in the generators.  128: the generators
// skip the arithmetic of every fourth shell pair (what a fourth generator warp would buy).  256: the consumers issue no
// DMMAs (the kernel time is then what the producers need when they have the FP64 unit to themselves).
#ifndef C2_WS_DIAG
#define C2_WS_DIAG 0
#endif
#if C2_WS_DIAG & 16
__device__ unsigned long long g_ws_diag[160 * 8 * 8];
__device__ unsigned long long g_ws_gdiag[160 * 8];
#endif

namespace c2 {

// NFRAG_ = 8-column fragments per MO block: 14 (NB = 112, large MO counts) or 8 (NB = 64, a few dozen to ~200 MOs, where
// 112-column blocks would pad too much).  Warp columns 0,1 (warpgroup 0) own NF0 fragments each, 2,3 (warpgroup 1) NF1.
// NPW_ = producer warps: 4 (one streamer + three generators; with 8 consumer warps 384 threads) or 8 (seven generators, 512
// threads: for one to four MO blocks, where the generators -- not the tensor pipe -- set the pace).
template <int ND_, int NFRAG_ = 14, int KC_ = 32, int NS_ = 3, int NPW_ = 4, int DENSE_ = 0>
struct WsCfg {
  // DENSE_: the 4 + 3 fragment split with the consumer loop of the balanced variants (ws_mma_stage; 2: and paired chi rows).
  // Denser consumers leave the generators less of the FP64 unit, which only pays when the generators have slack to give: many
  // MO blocks per sub-tile (wfn_choose_config)
  static constexpr bool DENSE = DENSE_ != 0;
  static constexpr int ND = ND_;
  static constexpr int P = 16, WM = 2;          // points per sub-tile (two 8-point MMA row groups)
  static constexpr int NFRAG = NFRAG_, NB = 8 * NFRAG_;    // MO columns per block
  static constexpr int NF0 = (NFRAG_ + 3) / 4, NF1 = (NFRAG_ - 2 * NF0) / 2;
  static_assert(2 * NF0 + 2 * NF1 == NFRAG_ && NF1 >= 1 && NB % 16 == 0, "fragment split");
  static constexpr int NCW = 8, NPW = NPW_;     // consumer / producer warps
  static constexpr int NT = 32 * (NCW + NPW);
  static constexpr int NCT = 32 * NCW, NPT = 32 * NPW;
  static constexpr int KC = KC_, NKQ = KC / 4;  // primitives per stage (<= 64: the streamer issues two rows per lane)
  static constexpr int NS = NS_;                // stages
  static constexpr int ACHUNK = 32 * ND * P;    // chi-cache doubles per 32 primitives (the unit of WfnParams::acache_chunks)
  static_assert(KC % 4 == 0 && KC <= 64 && NKQ <= 32, "stage size");
  static constexpr int NI = NF0;                // most fragments per consumer warp
  // chi tiles with derivative rows interleaved in pairs (one LDS.128 feeds two DMMA groups; see ws_a_index).  Only where
  // the two consumer warps of a sub-partition carry the same number of fragments: with 4 + 3 the lighter warp gets through
  // its stage even faster and the heavier one runs alone for longer (measured: 88.6 -> 87.4 % on the bench slab, while the
  // 64-column variant gains 4 %)
  static constexpr bool PAIRS = (ND_ % 2 == 0) && (NF0 == NF1 || DENSE_ == 2);
  static constexpr int SB = NB + 4;             // B row stride in doubles (= 4 mod 16)
  static constexpr int A_STAGE = KC * ND * P, B_STAGE = KC * SB;
  static constexpr int TSLOTS = 8;              // tile-info ring (>= NS + 3, see ws_generator)
  static constexpr int NPART = 2 * 4 * P * ND;  // s_part doubles: [2 sub-tile parities][4 warp columns][P][ND]
  static constexpr size_t SMEM = sizeof(double) * (NS * (A_STAGE + B_STAGE) + NPART + TSLOTS * 4 * P) + 256;
  // registers per warpgroup (setmaxnreg): 232 for four fragments (160 accumulator registers), 184 for three or fewer; the
  // three budgets add up to the 3 x 168 of the launch
  // three budgets add up to the 3 x 168 of the launch; with two producer warpgroups (512 threads) every warpgroup keeps the
  // 128 registers of the launch (two fragments per consumer warp need no more)
  static constexpr int REG_LAUNCH = 65536 / NT / 8 * 8;
  static constexpr int REG_CONS4 = NPW_ > 4 ? REG_LAUNCH : (NF0 >= 4 ? 232 : 184);
  static constexpr int REG_CONS3 = NPW_ > 4 ? REG_LAUNCH : (NF1 >= 4 ? 232 : 184);
  static constexpr int REG_PROD = NPW_ > 4 ? REG_LAUNCH : 504 - REG_CONS4 - REG_CONS3;
  static_assert(REG_PROD >= 88 && REG_PROD % 8 == 0, "register budget of the producer warpgroup");
  static_assert(NPW_ == 4 || (NPW_ == 8 && NF0 <= 2), "512-thread variant: at most two fragments per consumer warp");
};

// Per-stage header written by the producers, read by the consumers.
struct WsStageHdr {
  int flags;     // bit0 first chunk of an MO block, bit1 last chunk of an MO block, bit2 last block of the sub-tile,
                 // bit3 no data (sub-tile without active primitives), bit4 exit
  int blk;       // MO block
  int tslot;     // tile-info slot of the sub-tile
  int kqmask[2];   // per 8-point row group wm: bit kq set if k-quad kq has an unmasked (point, primitive) pair
};
struct WsTileInfo {
  int a0, b0, c0;   // sub-tile origin in output coordinates
  int nnuc;         // nucleus candidates of the super-tile (> 16: scan all atoms)
  int nuc[16];
};

// Split producer (generator warps -> streamer warp): one descriptor per chi-cache buffer.
struct WsTileDesc {
  int kact, nch, tslot, flags;   // flags bit0: exit
};

#ifdef __CUDACC__
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
template <int N>
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// ---------------------------------------------------------------------------------------------------------------
// Producer side: 128 threads, named barrier 2
// ---------------------------------------------------------------------------------------------------------------
template <class C>
struct WsShared {
  double *As, *Bs, *s_part, *s_pt;           // s_pt: [TSLOTS][4][P] (x,y,z,valid)
  uint64_t *full, *empty;                    // [NS]
  WsStageHdr *hdr;                           // [NS]
  WsTileInfo *tinfo;                         // [TSLOTS]
  double *p_red, *p_bb;                      // producer scratch: [NPW*6], [12]
  int *p_cnt, *p_int;                        // [NPW], [4]
  uint64_t *ready, *freeb;                   // [2] generator -> streamer / streamer -> generator (split producer)
  WsTileDesc *tdesc;                         // [2]
};

template <class C, int NW = C::NPW>
__device__ __forceinline__ void ws_prod_bbox(double lo[3], double hi[3], const WsShared<C> &S, double *bb, int ptid) {
  const int lane = ptid & 31, warp = ptid >> 5;
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
    for (int d = 0; d < 3; d++) { S.p_red[warp * 6 + d] = lo[d]; S.p_red[warp * 6 + 3 + d] = hi[d]; }
  }
  named_bar_sync(2, NW * 32);
  if (ptid < 6) {
    double v = S.p_red[ptid];
    for (int w = 1; w < NW; w++) v = (ptid < 3) ? fmin(v, S.p_red[w * 6 + ptid]) : fmax(v, S.p_red[w * 6 + ptid]);
    bb[ptid] = v;
  }
  named_bar_sync(2, NW * 32);
}

// ---------------------------------------------------------------------------------------------------------------
// Split producer.  An earlier version let all four producer warps compute the chi tile of MO block 0 straight into
// the stage and reload it from the L2 cache for blocks 1..6: ncu showed the consumers waiting ~17 % of the time on
// the computed stages while the producers idled ~60 % of the time during the cached blocks (74 % of the DMMA peak).
// Decoupling the two jobs took the same slab to 84 %:
//   * GENERATOR warps (3, named barrier 2 with 96 threads) run one sub-tile AHEAD: claim super-tiles, screen shells,
//     evaluate chi for every active primitive of the sub-tile and write the tiles (in the shared-memory layout of an
//     A stage) straight to the per-CTA L2 cache buffer `tcount & 1`;
//   * the STREAMER warp only moves data: for every (MO block, chunk) it acquires a stage and issues the bulk TMA
//     copies of the cached chi tile and of the C rows.  Every consumer stage is therefore a pure TMA stage.
// Hand-over: ready[b] (generator -> streamer, tile in buffer b complete) and freeb[b] (streamer -> generator, every
// TMA load of the tile in buffer b has landed), both single-arrival mbarriers used with the usual phase parity.
// ---------------------------------------------------------------------------------------------------------------
// Layout of the chi tile of one k-quad (4 primitives x 16 points x ND derivative rows).  Primitive-major slices, so the 16
// points a generator half-warp evaluates for one primitive are contiguous in the chi-tile cache (point-major slices made
// every store instruction touch 32 sectors, 8 bytes each: the generators were bound by the store path).  Derivative rows
// can be interleaved in PAIRS (WsCfg::PAIRS): element (d, k, pt) sits at ((d / 2) * 64 + ws_a_index(k, pt)) * 2 + d % 2, so a
// consumer lane fetches the operands of two derivative rows with one LDS.128 (half the operand loads of the MMA loop: a
// warp that has the tensor pipe to itself issues DMMAs at 79-87 % of the peak with one shared-memory load per group of
// 3-4, tools/dmma_issue.cu) and a generator lane stores them with one 16-byte store.  The point index is XOR-ed with 2k
// (pairs) or 4k (plain 8-byte slices): a shared-memory wavefront serves 128 bytes, i.e. the 8 (16) consecutive
// lanes = 2 (4) rows x 4 primitives of a consumer's A-fragment load, which this spreads over all 32 banks.
template <bool PAIRS>
__device__ __forceinline__ int ws_a_index(int k, int pt) { return k * 16 + (pt ^ ((PAIRS ? 2 : 4) * k)); }

template <int ND, bool PAIRS>
__device__ __forceinline__ void ws_store_chi(double *Aq, int k, int pt, const double *chi) {   // Aq: start of the k-quad's tile
  const int ai = ws_a_index<PAIRS>(k, pt);
  if (PAIRS) {
#pragma unroll
    for (int dp = 0; dp < ND / 2; dp++)
      *reinterpret_cast<double2 *>(Aq + (dp * 64 + ai) * 2) = make_double2(chi[2 * dp], chi[(ND > 1) ? 2 * dp + 1 : 0]);
  } else {
#pragma unroll
    for (int d = 0; d < ND; d++) Aq[d * 64 + ai] = chi[d];
  }
}

template <class C>
__device__ void ws_generator(const WfnParams &p, const WsShared<C> &S, int gtid) {
  constexpr int ND = C::ND, P = C::P, KC = C::KC, NGW = C::NPW - 1, NGT = 32 * NGW;
  const int lane = gtid & 31, warp = gtid >> 5;
  const int pt = gtid & 15;
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  // per-CTA scratch: shell list [shell_cap] then two (primitive, shell) lists [prim_cap] (one per cache buffer)
  int *sh_list = p.scratch + (size_t)blockIdx.x * (p.shell_cap + 4 * p.prim_cap);
  int2 *pr_list0 = reinterpret_cast<int2 *>(sh_list + p.shell_cap);
  constexpr int WS_REC = 12;   // doubles per active-shell record (96 bytes)
  double *rec = p.ws_rec + (size_t)blockIdx.x * (size_t)p.shell_cap * WS_REC;
  const int na = grid_na(g), nb = grid_nb(g);
  const int npsuper = g.sa * g.sb * g.sc;
  const int nsub_a = g.sa / g.ta, nsub_b = g.sb / g.tb, nsub_c = g.sc / g.tc;
  unsigned tcount = 0;
  __shared__ int s_nuc_tmp[16];
#if C2_WS_DIAG & 16
  long long gg_free = 0, gg_s2 = 0, gg_chi = 0, gg_t0 = 0, gg_t1 = 0;
  const long long gg_start = clock64();
#endif

  for (;;) {
    if (gtid == 0) S.p_int[0] = atomicAdd(p.work_counter, 1);
    named_bar_sync(2, NGT);
    const int st = S.p_int[0];
    if (st >= p.nitems) break;
    int ia, ib, ic;
    part_item(g, st, p.nsa, p.nsb, ia, ib, ic);
    const int sa0 = ia * g.sa, sb0 = ib * g.sb, sc0 = g.c0 + ic * g.sc;
    // ---- S1: super-tile box, ordered shell compaction, nucleus candidates ---------------------------------------
    {
      double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
      for (int q = gtid; q < npsuper; q += NGT) {
        int a = sa0 + q % g.sa, b = sb0 + (q / g.sa) % g.sb, c = sc0 + q / (g.sa * g.sb);
        if (a < na && b < nb && c < g.c1) {
          double x[3], wc[3], wcr[3];
          grid_point(g, a, b, c, x);
          grd_wrap(p.xf, x, g.periodic, wc, wcr);
#pragma unroll
          for (int d = 0; d < 3; d++) { lo[d] = fmin(lo[d], wcr[d]); hi[d] = fmax(hi[d], wcr[d]); }
        }
      }
      ws_prod_bbox<C, NGW>(lo, hi, S, S.p_bb, gtid);
    }
    if (gtid == 0) S.p_int[1] = 0;
    named_bar_sync(2, NGT);
    int ns = 0;
    for (int base = 0; base < w.nshell; base += NGT) {
      int s = base + gtid;
      bool act = false;
      if (s < w.nshell) {
        double4 xa = w.shell_xa[s];
        act = box_d2(S.p_bb, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12);
      }
      unsigned m = __ballot_sync(~0u, act);
      if (lane == 0) S.p_cnt[warp] = __popc(m);
      named_bar_sync(2, NGT);
      int off = ns, tot = 0;
#pragma unroll
      for (int i = 0; i < NGW; i++) {
        if (i < warp) off += S.p_cnt[i];
        tot += S.p_cnt[i];
      }
      if (act) sh_list[off + __popc(m & ((1u << lane) - 1u))] = s;
      ns += tot;
      named_bar_sync(2, NGT);
    }
    for (int i = gtid; i < w.natom; i += NGT) {
      if (box_d2(S.p_bb, w.atom_xyz[3 * i], w.atom_xyz[3 * i + 1], w.atom_xyz[3 * i + 2]) <= 1.0001e-10) {
        int k = atomicAdd(&S.p_int[1], 1);
        if (k < 16) s_nuc_tmp[k] = i;
      }
    }
    named_bar_sync(2, NGT);
    const int nnuc = S.p_int[1];

    for (int sub = 0; sub < nsub_a * nsub_b * nsub_c; sub++) {
      const int ta0 = sa0 + (sub % nsub_a) * g.ta, tb0 = sb0 + ((sub / nsub_a) % nsub_b) * g.tb;
      const int tc0 = sc0 + (sub / (nsub_a * nsub_b)) * g.tc;
      if (ta0 >= na || tb0 >= nb || tc0 >= g.c1) continue;
      const int tslot = tcount % C::TSLOTS;
      const int buf = tcount & 1;
      int2 *pr_list = pr_list0 + (size_t)buf * p.prim_cap;
      double *tp = S.s_pt + tslot * 4 * P;
      // ---- points of the sub-tile -> tile-info slot.  The generator is at most 2 sub-tiles ahead of the streamer and
      // the consumers at most NS stages (sub-tiles) behind it: TSLOTS >= NS + 3 keeps the slot free. -------------------
      {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        if (gtid < P) {
          int a = ta0 + gtid % g.ta, b = tb0 + (gtid / g.ta) % g.tb, c = tc0 + gtid / (g.ta * g.tb);
          bool valid = a < na && b < nb && c < g.c1;
          double x[3] = {0, 0, 0}, wc[3], wcr[3] = {0, 0, 0};
          if (valid) {
            grid_point(g, a, b, c, x);
            grd_wrap(p.xf, x, g.periodic, wc, wcr);
#pragma unroll
            for (int d = 0; d < 3; d++) { lo[d] = wcr[d]; hi[d] = wcr[d]; }
          }
          tp[gtid] = wcr[0]; tp[P + gtid] = wcr[1]; tp[2 * P + gtid] = wcr[2];
          tp[3 * P + gtid] = valid ? 1.0 : 0.0;
        }
        if (gtid == 0) {
          WsTileInfo &ti = S.tinfo[tslot];
          ti.a0 = ta0; ti.b0 = tb0; ti.c0 = tc0; ti.nnuc = nnuc;
          for (int k = 0; k < 16 && k < nnuc; k++) ti.nuc[k] = s_nuc_tmp[k];
        }
        ws_prod_bbox<C, NGW>(lo, hi, S, S.p_bb + 6, gtid);
      }
      // the streamer has finished with the tile that used this cache buffer (and its primitive list)
#if C2_WS_DIAG & 16
      gg_t0 = clock64();
#endif
      mbar_wait(&S.freeb[buf], ((tcount >> 1) & 1u) ^ 1u);
#if C2_WS_DIAG & 16
      gg_t1 = clock64(); gg_free += gg_t1 - gg_t0;
#endif
      // ---- S2: refine against the sub-tile box; ordered expansion into (primitive, shell) pairs (the streamer's C-row
      // gather) and one 96-byte record per ACTIVE SHELL: centre, exponent, dran, the two bounds of the rounding band of
      // the screen, (first slot in the primitive list, primitive count) and the angular momenta of its primitives, so the
      // evaluation below depends on ONE global load per shell ------------------------------------------------------------
      int kact = 0, nact_sh = 0;
      for (int base = 0; base < ns; base += NGT) {
        int j = base + gtid, cnt = 0, pb = 0, s = 0;
        if (j < ns) {
          s = sh_list[j];
          double4 xa = w.shell_xa[s];
          if (box_d2(S.p_bb + 6, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12)) {
            pb = w.shell_pbeg[s];
            cnt = w.shell_pbeg[s + 1] - pb;
          }
        }
        const int mine = cnt | (cnt > 0 ? 0x10000 : 0);   // packed: primitives (a batch has < 65536), shells
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          int v = __shfl_up_sync(~0u, incl, o);
          if (lane >= o) incl += v;
        }
        if (lane == 31) S.p_cnt[warp] = incl;
        named_bar_sync(2, NGT);
        int offp = incl - mine, tot = 0;
#pragma unroll
        for (int i = 0; i < NGW; i++) {
          if (i < warp) offp += S.p_cnt[i];
          tot += S.p_cnt[i];
        }
        if (cnt > 0) {
          const int off = kact + (offp & 0xffff);
          const double4 xa = w.shell_xa[s];
          const double2 cut = w.shell_cut[s];
          unsigned long long lc[4] = {0ull, 0ull, 0ull, 0ull};   // (lx, ly, lz) of the primitives: 10 bits each, 6 per word
          for (int i = 0; i < cnt; i++) {
            pr_list[off + i] = make_int2(pb + i, s);
            const uchar4 l = reinterpret_cast<const uchar4 *>(w.prim_l)[pb + i];
            lc[i / 6] |= (unsigned long long)(l.x | (l.y << 3) | (l.z << 6)) << (10 * (i % 6));
          }
          const long long bits = (long long)off | ((long long)cnt << 24);
          double4 *r = reinterpret_cast<double4 *>(rec + (size_t)(nact_sh + (offp >> 16)) * WS_REC);
          r[0] = xa;
          r[1] = make_double4(cut.x, cut.y * (1.0 - 1e-14), cut.y * (1.0 + 1e-14), __longlong_as_double(bits));
          r[2] = make_double4(__longlong_as_double((long long)lc[0]), __longlong_as_double((long long)lc[1]),
                              __longlong_as_double((long long)lc[2]), __longlong_as_double((long long)lc[3]));
        }
        kact += tot & 0xffff;
        nact_sh += tot >> 16;
        named_bar_sync(2, NGT);
      }
      const int nch = (kact + KC - 1) / KC;
      const int nquad = nch * (KC / 4);
      double *agbase = p.acache + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * C::ACHUNK;
      int *nzbase = p.acache_nz + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * 8;
      for (int q = gtid; q < nquad; q += NGT) nzbase[q] = 0;   // k-quad masks: OR-ed together below
      if (gtid == 0) S.p_int[2] = 0;                            // next shell pair to evaluate (claimed dynamically)
      __threadfence_block();
      named_bar_sync(2, NGT);   // records and cleared masks visible to every generator thread
      const double px = tp[pt], py = tp[P + pt], pz = tp[2 * P + pt];
      const bool pvalid = tp[3 * P + pt] != 0.0;
#if C2_WS_DIAG & 16
      gg_t0 = clock64(); gg_s2 += gg_t0 - gg_t1;
#endif
      // ---- chi tiles: a warp takes two active shells at a time (half-warp = shell, lane = point) and walks their
      // primitives together; slot k of the primitive list is column k % 4 of k-quad k / 4 (quads are contiguous in the
      // cache: chunk c starts at c * A_STAGE) ---------------------------------------------------------------------------
      const int hsh = lane & 16;
      double zero[ND];
#pragma unroll
      for (int d = 0; d < ND; d++) zero[d] = 0.0;
      for (;;) {   // shell pairs are claimed one at a time: shells differ in cost by the number of their primitives
        int j = 0;
        if (lane == 0) j = atomicAdd(&S.p_int[2], 1);
        j = __shfl_sync(~0u, j, 0);
        if (2 * j >= nact_sh) break;
        const int si = 2 * j + (hsh >> 4);
        const bool has = si < nact_sh;
        double4 xa = make_double4(0.0, 0.0, 0.0, 0.0), r1 = make_double4(0.0, 0.0, 0.0, 0.0), r2 = r1;
        if (has) {
          const double4 *r = reinterpret_cast<const double4 *>(rec + (size_t)si * WS_REC);
          xa = r[0];
          r1 = r[1];
          r2 = r[2];
        }
        if (si + 2 * NGW < nact_sh) {   // the records of this half-warp's next shell: into L1 while this one is evaluated
          const double *nx = rec + (size_t)(si + 2 * NGW) * WS_REC;
          asm volatile("prefetch.global.L1 [%0];" ::"l"(nx));
          asm volatile("prefetch.global.L1 [%0];" ::"l"(nx + WS_REC - 1));
        }
        const long long bits = __double_as_longlong(r1.w);
        const int off = (int)(bits & 0xffffff), cnt = (int)((bits >> 24) & 0xff);
        const int cmax = max(cnt, __shfl_xor_sync(~0u, cnt, 16));
        ShellAxis ax, ay;
        ax.x = px - xa.x; ay.x = py - xa.y;
        const double zr = pz - xa.z;
        ax.x2 = ax.x * ax.x; ay.x2 = ay.x * ay.x;
        const double z2 = zr * zr;
        const double d2 = __dadd_rn(__dadd_rn(ax.x2, ay.x2), z2);
        bool pass = has && pvalid && !(d2 > r1.z);                          // screen (:2747): outside the rounding band ...
        if (pass && !(d2 < r1.y)) pass = !(norm2_ref(ax.x, ay.x, zr) > r1.x);   // ... inside it, the reference's own norm2
        if (!__any_sync(~0u, pass)) {   // no point of the sub-tile within reach of either shell
          for (int i = 0; i < cnt; i++) {
            const int pos = off + i;
            ws_store_chi<ND, C::PAIRS>(agbase + (size_t)(pos >> 2) * ND * P * 4, pos & 3, pt, zero);
          }
          continue;
        }
        const double al = xa.w;
        const double ex = pass ? ((C2_WS_DIAG & 64) ? 1.0 - al * d2 : exp_neg(-al * d2)) : 0.0;   // :2749; 0 also zeroes every chi of the lane
        const double ta = al + al;
        ShellAxisE az;
        shell_axis_setup(ax, ta);
        shell_axis_setup(ay, ta);
        shell_axis_setup(az, zr, z2, ta, ex);
        unsigned long long lw = (unsigned long long)__double_as_longlong(r2.x);
        for (int i = 0; i < cmax; i++) {
          const bool on = i < cnt;
          const int lx = (int)(lw & 7u), ly = (int)((lw >> 3) & 7u), lz = (int)((lw >> 6) & 7u);
          lw >>= 10;
          if (i % 6 == 5) lw = (unsigned long long)__double_as_longlong(i < 6 ? r2.y : (i < 12 ? r2.z : r2.w));
          double chi[ND];
          if ((C2_WS_DIAG & 2) || ((C2_WS_DIAG & 128) && ((j / NGW) & 3) == 3)) {
#pragma unroll
            for (int d = 0; d < ND; d++) chi[d] = 0.0;
          } else if (__any_sync(~0u, (lx | ly | lz) > 2))
            shell_prim_chi<ND, true>(ax, ay, az, lx, ly, lz, ta, p.chi_mode, chi);
          else
            shell_prim_chi<ND, false>(ax, ay, az, lx, ly, lz, ta, p.chi_mode, chi);
          const int pos = off + i;
          if (on && (!(C2_WS_DIAG & 32) || chi[0] == 1.2345)) {
            ws_store_chi<ND, C::PAIRS>(agbase + (size_t)(pos >> 2) * ND * P * 4, pos & 3, pt, chi);
          }
          const unsigned bal = __ballot_sync(~0u, on && pass);
          const unsigned hm = (bal >> hsh) & 0xffffu;
          if (pt == 0 && hm) atomicOr(&nzbase[pos >> 2], ((hm & 0xffu) ? 1 : 0) | ((hm & 0xff00u) ? 2 : 0));
        }
      }
      // unused columns of the last k-quad: zeros (their C rows are whatever the stage held before)
      if (warp == 0)
        for (int pos = kact + (hsh >> 4); pos < ((kact + 3) & ~3); pos += 2) {
          ws_store_chi<ND, C::PAIRS>(agbase + (size_t)(pos >> 2) * ND * P * 4, pos & 3, pt, zero);
        }
#if C2_WS_DIAG & 16
      gg_chi += clock64() - gg_t0;
#endif
      // tile complete: make the generic-proxy global writes visible to the streamer's bulk (async-proxy) loads
      __threadfence();
      fence_proxy_async_global();
      named_bar_sync(2, NGT);
      if (gtid == 0) {
        WsTileDesc &td = S.tdesc[buf];
        td.kact = kact; td.nch = nch; td.tslot = tslot; td.flags = 0;
        mbar_arrive(&S.ready[buf]);
      }
      tcount++;
    }
  }
#if C2_WS_DIAG & 16
  if (lane == 0 && blockIdx.x < 160) {
    unsigned long long *o = g_ws_gdiag + (size_t)blockIdx.x * 8;
    if (warp == 0) { o[0] = gg_free; o[1] = gg_s2; o[2] = gg_chi; o[3] = clock64() - gg_start; }
    if (warp == 1) { o[4] = gg_free; o[5] = gg_s2; o[6] = gg_chi; o[7] = clock64() - gg_start; }
  }
#endif
  if (gtid == 0) {   // exit descriptor
    const int buf = tcount & 1;
    mbar_wait(&S.freeb[buf], ((tcount >> 1) & 1u) ^ 1u);
    WsTileDesc &td = S.tdesc[buf];
    td.kact = 0; td.nch = 0; td.tslot = 0; td.flags = 1;
    mbar_arrive(&S.ready[buf]);
  }
}

template <class C>
__device__ void ws_streamer(const WfnParams &p, const WsShared<C> &S, int lane) {
  constexpr int KC = C::KC, NS = C::NS, NB = C::NB, SB = C::SB;
  const WfnDev &w = p.w;
  const int2 *pr_list0 = reinterpret_cast<const int2 *>(p.scratch + (size_t)blockIdx.x * (p.shell_cap + 4 * p.prim_cap) + p.shell_cap);
  unsigned it = 0;                  // stage counter (stage = it % NS, phase = (it / NS) & 1)
  auto acquire = [&]() -> int {
    const int stg = it % NS;
    mbar_wait(&S.empty[stg], ((it / NS) & 1u) ^ 1u);
    return stg;
  };
  for (unsigned tcount = 0;; tcount++) {
    const int buf = tcount & 1;
    mbar_wait(&S.ready[buf], (tcount >> 1) & 1u);
    fence_proxy_async_global();
    const WsTileDesc td = S.tdesc[buf];
    if (td.flags & 1) break;
    const int2 *pr_list = pr_list0 + (size_t)buf * p.prim_cap;
    const double *agbase = p.acache + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * C::ACHUNK;
    const int *nzbase = p.acache_nz + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * 8;
    const unsigned it0 = it;
    if (td.kact == 0) {   // nothing active: a data-less stage tells the consumers to emit zeros (+ EDF) for this tile
      const int stg = acquire();
      if (lane == 0) {
        WsStageHdr &h = S.hdr[stg];
        h.flags = 1 | 2 | 4 | 8; h.blk = 0; h.tslot = td.tslot;
        mbar_expect_tx(&S.full[stg], 0u);
      }
      it++;
    } else {
      for (int blk = 0; blk < p.nblk; blk++) {
        for (int cc = 0; cc < td.nch; cc++) {
          // odd MO blocks walk the chunks backwards: the chi tiles of a sub-tile (~0.65 MB per CTA, x 148 CTAs x 2
          // buffers) exceed the L2, and a cyclic re-read would miss every time; the zigzag re-uses the most recent ones
          const int c = (blk & 1) ? td.nch - 1 - cc : cc;
          const int nrow = min(KC, td.kact - c * KC);
          const int prim = lane < nrow ? pr_list[c * KC + lane].x : 0;         // loads issued before the wait
          const int prim2 = (KC > 32 && lane + 32 < nrow) ? pr_list[c * KC + lane + 32].x : 0;
          const int nz = lane < C::NKQ ? nzbase[c * C::NKQ + lane] : 0;
          const unsigned m0 = __ballot_sync(~0u, (nz & 1) != 0), m1 = __ballot_sync(~0u, (nz & 2) != 0);
          const int stg = acquire();
          if (lane == 0) {
            WsStageHdr &h = S.hdr[stg];
            h.kqmask[0] = (int)m0; h.kqmask[1] = (int)m1;
            h.flags = (cc == 0 ? 1 : 0) | (cc == td.nch - 1 ? 2 : 0) | ((cc == td.nch - 1 && blk == p.nblk - 1) ? 4 : 0);
            h.blk = blk; h.tslot = td.tslot;
          }
          __syncwarp();
          if (lane == 0) {
#if C2_WS_DIAG & 1
            mbar_expect_tx(&S.full[stg], 0u);
#else
            mbar_expect_tx(&S.full[stg], (uint32_t)(nrow * NB * sizeof(double)) + (uint32_t)(C::A_STAGE * sizeof(double)));
            tma_bulk_g2s(S.As + stg * C::A_STAGE, agbase + (size_t)c * C::A_STAGE, (uint32_t)(C::A_STAGE * sizeof(double)),
                         &S.full[stg]);
#endif
          }
          __syncwarp();
          if (C2_WS_DIAG & 1) { it++; continue; }
          if (lane < nrow)
            tma_bulk_g2s(S.Bs + stg * C::B_STAGE + lane * SB, w.cmoT + (size_t)prim * w.ldc + (size_t)blk * NB,
                         (uint32_t)(NB * sizeof(double)), &S.full[stg]);
          if (KC > 32 && lane + 32 < nrow)
            tma_bulk_g2s(S.Bs + stg * C::B_STAGE + (lane + 32) * SB, w.cmoT + (size_t)prim2 * w.ldc + (size_t)blk * NB,
                         (uint32_t)(NB * sizeof(double)), &S.full[stg]);
          it++;
        }
      }
    }
    // every load of this tile has landed (wait on the full barriers of the stages still in flight), then hand the
    // cache buffer back to the generators
    if (lane == 0) {
      const unsigned first = it - it0 > (unsigned)NS ? it - NS : it0;
      for (unsigned i = first; i < it; i++) mbar_wait(&S.full[i % NS], (i / NS) & 1u);
      mbar_arrive(&S.freeb[buf]);
    }
    __syncwarp();
  }
  {   // exit stage
    const int stg = acquire();
    if (lane == 0) {
      S.hdr[stg].flags = 16; S.hdr[stg].blk = 0; S.hdr[stg].tslot = 0;
      mbar_expect_tx(&S.full[stg], 0u);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Consumer side: 256 threads, named barrier 1.  Warps 0..3 (warp columns 0,1) own 4 fragments each and run with 232
// registers, warps 4..7 (warp columns 2,3) own 3 fragments and run with 184: every SM sub-partition carries 4 + 3 = 7
// of the 14 fragments of an MO block.
// ---------------------------------------------------------------------------------------------------------------
// MMAs of one stage.  kqm: bit kq set if k-quad kq has an unmasked (point, primitive) pair in this warp's 8 points.
// Operands of the next k-quad / derivative row are fetched from shared memory while the current DMMAs issue, so the
// tensor pipe does not drain between k-quads (the first version reloaded everything at the top of each k-quad and
// spilled accumulators: ~8 % of the in-loop time was both warps of a sub-partition waiting on operands).
// MMAs of one stage, first form: operands of the next k-quad / derivative row fetched while the current DMMAs issue, in
// the order ptxas picks.  Kept for the 4 + 3 fragment split, where it measures 1 % better than ws_mma_stage below (whatever
// makes the consumers of that configuration stream more densely takes slack from the generators).
template <class C, int NF>
__device__ __forceinline__ void ws_mma_stage_v1(double (&acc)[C::ND][NF][2], const double *Ast, const double *Bst, unsigned kqm) {
  constexpr int ND = C::ND, P = C::P, SB = C::SB;
  if (kqm == 0) return;                                  // all 8 points x 32 primitives masked (warp-uniform)
  int kq = __ffs(kqm) - 1;
  kqm &= kqm - 1;
  double b[NF];
#pragma unroll
  for (int ni = 0; ni < NF; ni++) b[ni] = Bst[kq * 4 * SB + ni * 8];
  const double *Ak = Ast + (size_t)kq * ND * P * 4;
  bool more;
  if (C::PAIRS) {                                        // Ast points at this lane's pair of derivative rows 0, 1
    double2 a = *reinterpret_cast<const double2 *>(Ak);
#pragma unroll 1
    do {
      more = kqm != 0;
      const int kqn = more ? __ffs(kqm) - 1 : kq;
      kqm &= kqm - 1;
      double bn[NF];
#pragma unroll
      for (int ni = 0; ni < NF; ni++) bn[ni] = Bst[kqn * 4 * SB + ni * 8];
      const double *An = Ast + (size_t)kqn * ND * P * 4;
#pragma unroll
      for (int dp = 0; dp < ND / 2; dp++) {
        const double2 an = *reinterpret_cast<const double2 *>((dp + 1 < ND / 2) ? Ak + (dp + 1) * 128 : An);
#pragma unroll
        for (int ni = 0; ni < NF; ni++) dmma884(acc[2 * dp][ni][0], acc[2 * dp][ni][1], a.x, b[ni]);
#pragma unroll
        for (int ni = 0; ni < NF; ni++) dmma884(acc[(ND > 1) ? 2 * dp + 1 : 0][ni][0], acc[(ND > 1) ? 2 * dp + 1 : 0][ni][1], a.y, b[ni]);
        a = an;
      }
#pragma unroll
      for (int ni = 0; ni < NF; ni++) b[ni] = bn[ni];
      Ak = An;
      kq = kqn;
    } while (more);
  } else {
    double a = Ak[0];
#pragma unroll 1
    do {
      more = kqm != 0;
      const int kqn = more ? __ffs(kqm) - 1 : kq;
      kqm &= kqm - 1;
      double bn[NF];
#pragma unroll
      for (int ni = 0; ni < NF; ni++) bn[ni] = Bst[kqn * 4 * SB + ni * 8];
      const double *An = Ast + (size_t)kqn * ND * P * 4;
#pragma unroll
      for (int d = 0; d < ND; d++) {
        const double an = (d + 1 < ND) ? Ak[(d + 1) * P * 4] : An[0];
#pragma unroll
        for (int ni = 0; ni < NF; ni++) dmma884(acc[d][ni][0], acc[d][ni][1], a, b[ni]);
        a = an;
      }
#pragma unroll
      for (int ni = 0; ni < NF; ni++) b[ni] = bn[ni];
      Ak = An;
      kq = kqn;
    } while (more);
  }
}

// Shared-memory operand loads as volatile asm, so they keep their place between the (volatile asm) DMMAs.
__device__ __forceinline__ void ws_lds(double &v, uint32_t addr) { asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr)); }
__device__ __forceinline__ void ws_lds(double2 &v, uint32_t addr) {
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
}

// The DMMAs of one k-quad: derivative rows (or pairs of rows) in turn, NF fragments each; a_cur holds the operand of the
// first step on entry and the operand of the NEXT k-quad's first step on exit.  WHERE the loads sit matters (ncu, source
// page): a DMMA reads its source registers when it executes, up to four queue slots after it was issued, so a load that
// re-uses an operand register right after the group that read it waits for that group to drain (9.7 % of the 4-fragment
// warps' samples were this short-scoreboard wait).  Each step therefore issues its leading DMMA first -- that one waits for
// a queue slot, by which time the previous group has executed -- and only then the load of the next step's operand into
// the register the previous group used; the B operands of the next k-quad are loaded into the set the PREVIOUS k-quad
// used (two sets ping-pong, the caller alternates them: no register moves) during step 1.
template <class C, int NF, class AT>
__device__ __forceinline__ void ws_mma_kquad(double (&acc)[C::ND][NF][2], AT &a_cur, AT &a_oth, const double (&b)[NF], double (&bn)[NF],
                                             uint32_t ak, uint32_t an_first, uint32_t bn_addr) {
  constexpr int ND = C::ND;
  constexpr bool PAIRS = C::PAIRS;
  constexpr int NSTEP = PAIRS ? ND / 2 : ND;
  constexpr uint32_t STEP_BYTES = (PAIRS ? 128 : 64) * 8;   // one derivative row = 64 doubles, a pair = 128
#pragma unroll
  for (int st = 0; st < NSTEP; st++) {
    AT &cur = (st & 1) ? a_oth : a_cur;
    AT &nxt = (st & 1) ? a_cur : a_oth;
    const int d0 = PAIRS ? 2 * st : st;
    if (PAIRS) {
      const double2 &c2 = reinterpret_cast<const double2 &>(cur);
      dmma884(acc[d0][0][0], acc[d0][0][1], c2.x, b[0]);
      ws_lds(nxt, st + 1 < NSTEP ? ak + (st + 1) * STEP_BYTES : an_first);
#pragma unroll
      for (int ni = 1; ni < NF; ni++) dmma884(acc[d0][ni][0], acc[d0][ni][1], c2.x, b[ni]);
      if (st == 1) {
#pragma unroll
        for (int ni = 0; ni < NF; ni++) ws_lds(bn[ni], bn_addr + ni * 64);
      }
#pragma unroll
      for (int ni = 0; ni < NF; ni++) dmma884(acc[(ND > 1) ? d0 + 1 : 0][ni][0], acc[(ND > 1) ? d0 + 1 : 0][ni][1], c2.y, b[ni]);
    } else {
      const double &c1 = reinterpret_cast<const double &>(cur);
      dmma884(acc[d0][0][0], acc[d0][0][1], c1, b[0]);
      ws_lds(nxt, st + 1 < NSTEP ? ak + (st + 1) * STEP_BYTES : an_first);
#pragma unroll
      for (int ni = 1; ni < NF; ni++) dmma884(acc[d0][ni][0], acc[d0][ni][1], c1, b[ni]);
      if (st == (NSTEP > 1 ? 1 : 0)) {
#pragma unroll
        for (int ni = 0; ni < NF; ni++) ws_lds(bn[ni], bn_addr + ni * 64);
      }
    }
  }
}

template <class C, int NF>
__device__ __forceinline__ void ws_mma_stage(double (&acc)[C::ND][NF][2], const double *Ast, const double *Bst, unsigned kqm) {
  constexpr int ND = C::ND, P = C::P, SB = C::SB;
  constexpr int NSTEP = C::PAIRS ? ND / 2 : ND;
  using AT = typename std::conditional<C::PAIRS, double2, double>::type;
  if (kqm == 0) return;                                  // all 8 points x 32 primitives masked (warp-uniform)
  const uint32_t a0 = smem_u32(Ast), b0a = smem_u32(Bst);
  constexpr uint32_t KQ_A = ND * P * 4 * 8, KQ_B = 4 * SB * 8;   // bytes per k-quad of the chi tile / of the C rows
  int kq = __ffs(kqm) - 1;
  kqm &= kqm - 1;
  double b0[NF], b1[NF];
#pragma unroll
  for (int ni = 0; ni < NF; ni++) ws_lds(b0[ni], b0a + kq * KQ_B + ni * 64);
  AT ax, ay;
  ws_lds(ax, a0 + kq * KQ_A);
  // an odd number of steps per k-quad swaps the roles of ax / ay from one k-quad to the next
  bool more;
#pragma unroll 1
  for (;;) {
    more = kqm != 0;
    int kqn = more ? __ffs(kqm) - 1 : kq;
    kqm &= kqm - 1;
    ws_mma_kquad<C, NF, AT>(acc, ax, ay, b0, b1, a0 + kq * KQ_A, a0 + kqn * KQ_A, b0a + kqn * KQ_B);
    if (!more) break;
    kq = kqn;
    more = kqm != 0;
    kqn = more ? __ffs(kqm) - 1 : kq;
    kqm &= kqm - 1;
    if (NSTEP & 1) ws_mma_kquad<C, NF, AT>(acc, ay, ax, b1, b0, a0 + kq * KQ_A, a0 + kqn * KQ_A, b0a + kqn * KQ_B);
    else ws_mma_kquad<C, NF, AT>(acc, ax, ay, b1, b0, a0 + kq * KQ_A, a0 + kqn * KQ_A, b0a + kqn * KQ_B);
    if (!more) break;
    kq = kqn;
  }
}

template <class C, int NF>
__device__ void ws_consumer(const WfnParams &p, const WsShared<C> &S, int ctid) {
  constexpr int ND = C::ND, P = C::P, NS = C::NS, NB = C::NB, SB = C::SB, NCT = C::NCT;
  const int lane = ctid & 31, cw = ctid >> 5;
  const int wm = cw & 1, wn = cw >> 1;                 // row group (8 points) and warp column (0..3)
  const int f0 = wn < 2 ? wn * C::NF0 : 2 * C::NF0 + (wn - 2) * C::NF1;   // first fragment of this warp (4,4,3,3 for NB = 112)
  const int row = lane >> 2;
  const int aoff = ws_a_index<C::PAIRS>(lane & 3, wm * 8 + row) * (C::PAIRS ? 2 : 1);
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  double acc[ND][NF][2];
  double q[ND];                                         // this thread's share of the quadratic forms of the sub-tile
#pragma unroll
  for (int d = 0; d < ND; d++) {
    q[d] = 0.0;
#pragma unroll
    for (int ni = 0; ni < NF; ni++) acc[d][ni][0] = acc[d][ni][1] = 0.0;
  }
  unsigned tpar = 0;                                    // sub-tile parity: s_part is double-buffered
#if C2_WS_DIAG & 16
  long long dg_wait = 0, dg_mma = 0, dg_wait1 = 0, dg_epi = 0, dg_epi0 = 0;
  unsigned long long dg_cnt = 0, dg_stages = 0, dg_tiles = 0;
  const long long dg_start = clock64();
#endif

  for (unsigned it = 0;; it++) {
    const int stg = it % NS;
#if C2_WS_DIAG & 16
    const long long dg0 = clock64();
    if (dg_epi0) { dg_epi += dg0 - dg_epi0; dg_epi0 = 0; }
#endif
    mbar_wait(&S.full[stg], (it / NS) & 1u);
#if C2_WS_DIAG & 16
    const long long dg1 = clock64();
    dg_wait += dg1 - dg0;
#endif
    const WsStageHdr &h = S.hdr[stg];
    const int flags = h.flags, blk = h.blk, tslot = h.tslot;
    if (flags & 16) break;
#if C2_WS_DIAG & 16
    dg_stages++;
    if ((flags & 1) && blk == 0) { dg_wait1 += dg1 - dg0; dg_tiles++; }
#endif
    if (!(flags & 8)) {
      const double *Ast = S.As + stg * C::A_STAGE + aoff;
      const double *Bst = S.Bs + stg * C::B_STAGE + (lane & 3) * SB + f0 * 8 + (lane >> 2);
      if (!(C2_WS_DIAG & 256)) {
        if (C::NF0 == C::NF1 || C::DENSE) ws_mma_stage<C, NF>(acc, Ast, Bst, (unsigned)h.kqmask[wm]);
        else ws_mma_stage_v1<C, NF>(acc, Ast, Bst, (unsigned)h.kqmask[wm]);
      }
#if C2_WS_DIAG & 16
      dg_mma += clock64() - dg1;
      dg_cnt += (unsigned long long)__popc((unsigned)h.kqmask[wm]) * ND * NF;
#endif
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&S.empty[stg]);           // stage (and its header) no longer needed by this warp
    if (!(flags & 2)) continue;
#if C2_WS_DIAG & 16
    dg_epi0 = clock64();
#endif

    // ---- end of an MO block: occupation-weighted quadratic forms of this thread's MOs (:2091-2147).  No barrier:
    // the partial sums stay in registers until the end of the sub-tile, so the warps drift apart and one warp's
    // epilogue overlaps the other warps' MMAs. ------------------------------------------------------------------------
    if (!(flags & 8)) {
#pragma unroll
      for (int ni = 0; ni < NF; ni++)
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const double oc = w.occ[blk * NB + (f0 + ni) * 8 + 2 * (lane & 3) + j];
          const double v0 = acc[0][ni][j];
          q[0] += oc * v0 * v0;
          if (ND > 1) {
            const double t = 2.0 * oc;
            const double v1 = acc[ND > 1 ? 1 : 0][ni][j], v2 = acc[ND > 1 ? 2 : 0][ni][j], v3 = acc[ND > 1 ? 3 : 0][ni][j];
            q[ND > 1 ? 1 : 0] += t * v0 * v1;
            q[ND > 1 ? 2 : 0] += t * v0 * v2;
            q[ND > 1 ? 3 : 0] += t * v0 * v3;
            if (ND > 4) {
              q[ND > 4 ? 4 : 0] += t * (v0 * acc[ND > 4 ? 4 : 0][ni][j] + v1 * v1);
              q[ND > 4 ? 5 : 0] += t * (v0 * acc[ND > 4 ? 5 : 0][ni][j] + v2 * v2);
              q[ND > 4 ? 6 : 0] += t * (v0 * acc[ND > 4 ? 6 : 0][ni][j] + v3 * v3);
              q[ND > 4 ? 7 : 0] += t * (v0 * acc[ND > 4 ? 7 : 0][ni][j] + v1 * v2);
              q[ND > 4 ? 8 : 0] += t * (v0 * acc[ND > 4 ? 8 : 0][ni][j] + v1 * v3);
              q[ND > 4 ? 9 : 0] += t * (v0 * acc[ND > 4 ? 9 : 0][ni][j] + v2 * v3);
            }
          }
        }
#pragma unroll
      for (int d = 0; d < ND; d++)
#pragma unroll
        for (int ni = 0; ni < NF; ni++) acc[d][ni][0] = acc[d][ni][1] = 0.0;
    }
    if (!(flags & 4)) continue;

    // ---- end of the sub-tile: fold the partial sums (fixed order: lanes, then warp columns), EDF core, nucleus test,
    // derived quantities, stores ------------------------------------------------------------------------------------------
    double *part = S.s_part + tpar * 4 * P * ND;         // [4 warp columns][P][ND]
    double *other = S.s_part + (tpar ^ 1u) * 4 * P * ND; // free until the next sub-tile ends (EDF scratch)
    tpar ^= 1u;
#pragma unroll
    for (int d = 0; d < ND; d++) {
      q[d] += __shfl_xor_sync(~0u, q[d], 1);
      q[d] += __shfl_xor_sync(~0u, q[d], 2);
    }
    if ((lane & 3) == 0) {
#pragma unroll
      for (int d = 0; d < ND; d++) part[(wn * P + wm * 8 + row) * ND + d] = q[d];
    }
#pragma unroll
    for (int d = 0; d < ND; d++) q[d] = 0.0;
    named_bar_sync(1, NCT);
    const double *tp = S.s_pt + tslot * 4 * P;
    const WsTileInfo &ti = S.tinfo[tslot];
    if (w.nedf > 0) {                                    // core densities of ECP wavefunctions (calculate_edf, :2783)
      const int pt = ctid & 15, slot = ctid >> 4;        // 4 slots (threads 0..63)
      if (ctid < 64) {
        double e[ND];
#pragma unroll
        for (int d = 0; d < ND; d++) e[d] = 0.0;
        if (tp[3 * P + pt] != 0.0) {
          const double px = tp[pt], py = tp[P + pt], pz = tp[2 * P + pt];
          for (int i = slot; i < w.nedf; i += 4) {
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
        for (int d = 0; d < ND; d++) other[(slot * P + pt) * ND + d] = e[d];
      }
      named_bar_sync(1, NCT);
    }
    if (ctid < P) {
      if (tp[3 * P + ctid] != 0.0) {
        double t[ND];
#pragma unroll
        for (int d = 0; d < ND; d++) {
          double sum = 0.0;
#pragma unroll
          for (int v = 0; v < 4; v++) sum += part[(v * P + ctid) * ND + d];
          if (w.nedf > 0) {
#pragma unroll
            for (int v = 0; v < 4; v++) sum += other[(v * P + ctid) * ND + d];
          }
          t[d] = sum;
        }
        const int a = ti.a0 + ctid % g.ta, b = ti.b0 + (ctid / g.ta) % g.tb, c = ti.c0 + ctid / (g.ta * g.tb);
        double f = t[0], gx = 0, gy = 0, gz = 0, hs[6] = {0, 0, 0, 0, 0, 0};
        if (ND > 1) { gx = t[ND > 1 ? 1 : 0]; gy = t[ND > 1 ? 2 : 0]; gz = t[ND > 1 ? 3 : 0]; }
        if (ND > 4) {
#pragma unroll
          for (int d = 0; d < 6; d++) hs[d] = t[ND > 4 ? 4 + d : 0];
        }
        bool isnuc = false;
        const double x = tp[ctid], y = tp[P + ctid], z = tp[2 * P + ctid];
        if (ti.nnuc > 16) {
          for (int i = 0; i < w.natom && !isnuc; i++) {
            double ddx = w.atom_xyz[3 * i] - x, ddy = w.atom_xyz[3 * i + 1] - y, ddz = w.atom_xyz[3 * i + 2] - z;
            isnuc = (ddx * ddx + ddy * ddy + ddz * ddz) < 1e-10;      // src/environmod@proc.f90:677
          }
        } else {
          for (int k = 0; k < ti.nnuc; k++) {
            int i = ti.nuc[k];
            double ddx = w.atom_xyz[3 * i] - x, ddy = w.atom_xyz[3 * i + 1] - y, ddz = w.atom_xyz[3 * i + 2] - z;
            isnuc = isnuc || (ddx * ddx + ddy * ddy + ddz * ddz) < 1e-10;
          }
        }
        emit_point_call(p.o, out_index(g, a, b, c), f, gx, gy, gz, hs[0], hs[1], hs[2], hs[3], hs[4], hs[5], isnuc, p.nci_const);
      }
    }
    if (w.nedf > 0) named_bar_sync(1, NCT);              // EDF scratch (the other s_part buffer) free again
  }
#if C2_WS_DIAG & 16
  if (lane == 0 && blockIdx.x < 160) {
    unsigned long long *o = g_ws_diag + ((size_t)blockIdx.x * 8 + cw) * 8;
    o[0] = (unsigned long long)dg_wait; o[1] = (unsigned long long)dg_mma; o[2] = dg_cnt;
    o[3] = (unsigned long long)(clock64() - dg_start);
    o[4] = (unsigned long long)dg_wait1; o[5] = (unsigned long long)dg_epi; o[6] = dg_stages; o[7] = dg_tiles;
  }
#endif
}

template <class C>
__global__ void __launch_bounds__(C::NT, 1) wfn_ws_kernel(const __grid_constant__ WfnParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_full[C::NS], s_empty[C::NS], s_ready[2], s_freeb[2];
  __shared__ WsStageHdr s_hdr[C::NS];
  __shared__ WsTileInfo s_tinfo[C::TSLOTS];
  __shared__ WsTileDesc s_tdesc[2];
  __shared__ double s_pred[C::NPW * 6], s_pbb[12];
  __shared__ int s_pcnt[C::NPW], s_pint[4];
  WsShared<C> S;
  S.As = reinterpret_cast<double *>(smem_raw);
  S.Bs = S.As + C::NS * C::A_STAGE;
  S.s_part = S.Bs + C::NS * C::B_STAGE;
  S.s_pt = S.s_part + C::NPART;
  S.full = s_full; S.empty = s_empty; S.hdr = s_hdr; S.tinfo = s_tinfo;
  S.p_red = s_pred; S.p_bb = s_pbb; S.p_cnt = s_pcnt; S.p_int = s_pint;
  S.ready = s_ready; S.freeb = s_freeb; S.tdesc = s_tdesc;
#if C2_WS_DIAG & 8
  const int tid = (threadIdx.x + C::NCT) % C::NT;   // producer warps get the lowest warp ids
#else
  const int tid = threadIdx.x;
#endif
  if (tid == 0) {
    for (int s = 0; s < C::NS; s++) {
      mbar_init(&s_full[s], 1);            // the streamer's arrive.expect_tx (+ the TMA byte count)
      mbar_init(&s_empty[s], C::NCW);      // one arrival per consumer warp
    }
    for (int b = 0; b < 2; b++) { mbar_init(&s_ready[b], 1); mbar_init(&s_freeb[b], 1); }
    fence_mbar_init();
  }
  for (int i = tid; i < C::NS * C::B_STAGE; i += C::NT) S.Bs[i] = 0.0;   // stale rows stay finite (they meet chi = 0)
  fence_proxy_async();
  __syncthreads();
  // setmaxnreg works per warpgroup (4 warps): 128 x (232 + 184 + 88) = 64512 registers = 384 x 168, the launch figure
  constexpr bool REBALANCE = C::REG_CONS4 != C::REG_LAUNCH || C::REG_PROD != C::REG_LAUNCH;
  if (tid >= C::NCT) {
    if (REBALANCE) { if (C::REG_PROD <= C::REG_LAUNCH) reg_dec<C::REG_PROD>(); else reg_inc<C::REG_PROD>(); }
    if (tid < C::NCT + 32) ws_streamer<C>(p, S, tid - C::NCT);
    else ws_generator<C>(p, S, tid - C::NCT - 32);
  } else if (tid < 128) {
    if (REBALANCE) reg_inc<C::REG_CONS4>();
    ws_consumer<C, C::NF0>(p, S, tid);
  } else {
    if (REBALANCE) reg_inc<C::REG_CONS3>();
    ws_consumer<C, C::NF1>(p, S, tid);
  }
}
#endif

}  // namespace c2
