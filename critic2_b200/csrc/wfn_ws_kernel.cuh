// wfn_ws_kernel.cuh -- warp-specialised version of the wavefunction grid kernel for large MO counts.
//
// Same mathematics and data layout as wfn_kernel.cuh (see there for the reference citations), different execution
// model.  ncu on the first version showed the FP64 tensor pipe (DMMA) idle ~38 % of the time because every warp
// alternated between latency-bound primitive evaluation and MMAs behind one block-wide barrier per chunk.  Here
//   * 4 PRODUCER warps (one per SM sub-partition) own everything that is not an MMA: super-tile claiming, shell
//     screening (S1/S2), the chi tiles (A operand) and the bulk-TMA gather of C rows (B operand);
//   * 8 CONSUMER warps (two per sub-partition, 4+3 MMA fragments each so that every sub-partition carries 7 of the
//     14 fragments of a 112-column MO block) only wait on mbarriers, issue DMMAs and run the fused epilogue;
//   * registers are re-balanced with setmaxnreg (consumers 208, producers 88): the consumers hold 10x4x2 accumulators;
//   * stages hand over through full/empty mbarriers (no block-wide barrier in steady state), and the stream of
//     (sub-tile, MO block, chunk) steps is continuous, so producers run ahead across block and tile boundaries.
#pragma once
#include "wfn_kernel.cuh"

namespace c2 {

template <int ND_>
struct WsCfg {
  static constexpr int ND = ND_;
  static constexpr int P = 16, WM = 2;          // points per sub-tile (two 8-point MMA row groups)
  static constexpr int NFRAG = 14, NB = 112;    // MO columns per block
  static constexpr int NCW = 8, NPW = 4;        // consumer / producer warps
  static constexpr int NT = 32 * (NCW + NPW);
  static constexpr int NCT = 32 * NCW, NPT = 32 * NPW;
  static constexpr int KC = 32, NKQ = KC / 4;   // primitives per stage
  static constexpr int NS = 3;                  // stages
  static constexpr int NI = 4;                  // most fragments per consumer warp
  static constexpr int SB = 116;                // B row stride in doubles (= 4 mod 16)
  static constexpr int A_STAGE = KC * ND * P, B_STAGE = KC * SB;
  static constexpr int TSLOTS = 8;              // tile-info ring (>= NS + 3, see ws_generator)
  static constexpr int NPART = 2 * 4 * P * ND;  // s_part doubles: [2][4 warp columns][P][ND]
  static constexpr size_t SMEM = sizeof(double) * (NS * (A_STAGE + B_STAGE) + NPART + P * ND + TSLOTS * 4 * P) + 256;
  static constexpr int REG_CONS = 208, REG_PROD = 88;
};

// Per-stage header written by the producers, read by the consumers.
struct WsStageHdr {
  int flags;     // bit0 first chunk of an MO block, bit1 last chunk of an MO block, bit2 last block of the sub-tile,
                 // bit3 no data (sub-tile without active primitives), bit4 exit
  int blk;       // MO block
  int tslot;     // tile-info slot of the sub-tile
  int nzmask[WsCfg<1>::NKQ];   // per k-quad: bit wm set if any of the 8 points x 4 primitives is unmasked
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
  double *As, *Bs, *s_part, *s_tot, *s_pt;   // s_pt: [TSLOTS][4][P] (x,y,z,valid)
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

template <class C>
__device__ void ws_producer(const WfnParams &p, const WsShared<C> &S, int ptid) {
  constexpr int ND = C::ND, P = C::P, KC = C::KC, NS = C::NS, NB = C::NB, SB = C::SB, NPT = C::NPT, NPW = C::NPW;
  const int lane = ptid & 31, warp = ptid >> 5;
  const int grp = ptid >> 4;        // 8 groups of 16 lanes: group = k-quad of the stage
  const int pt = ptid & 15;         // point of the sub-tile
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  // per-CTA scratch: shell list [shell_cap, even] then (primitive, shell) pairs [prim_cap]
  int *sh_list = p.scratch + (size_t)blockIdx.x * (p.shell_cap + 2 * p.prim_cap);
  int2 *pr_list = reinterpret_cast<int2 *>(sh_list + p.shell_cap);
  // 64-byte records of the active primitives of the current sub-tile: x, y, z, alpha, dran, dran^2, bits, 0 with
  // bits = shell id | lx << 32 | ly << 40 | lz << 48 (everything the chi evaluation needs: one coalesced load per quad)
  double *rec = p.ws_rec + (size_t)blockIdx.x * (size_t)p.prim_cap * 8;
  const int na = grid_na(g), nb = grid_nb(g);
  const int npsuper = g.sa * g.sb * g.sc;
  const int nsub_a = g.sa / g.ta, nsub_b = g.sb / g.tb, nsub_c = g.sc / g.tc;
  unsigned it = 0;                  // stage counter (stage = it % NS, phase = (it / NS) & 1)
  unsigned tcount = 0;              // sub-tile counter (tile slot = tcount % TSLOTS)

  // acquire the next stage (wait until the consumers released it)
  auto acquire = [&]() -> int {
    const int stg = it % NS;
    mbar_wait(&S.empty[stg], ((it / NS) & 1u) ^ 1u);
    return stg;
  };
  auto publish = [&](int stg) {
    __syncwarp();
    if (lane == 0) mbar_arrive(&S.full[stg]);
    it++;
  };

  for (;;) {
    if (ptid == 0) S.p_int[0] = atomicAdd(p.work_counter, 1);
    named_bar_sync(2, NPT);
    const int st = S.p_int[0];
    if (st >= p.nsa * p.nsb * p.nsc) break;
    const int sa0 = (st % p.nsa) * g.sa, sb0 = ((st / p.nsa) % p.nsb) * g.sb;
    const int sc0 = g.c0 + (g.c_part + (st / (p.nsa * p.nsb)) * g.c_nparts) * g.sc;
    // ---- S1: super-tile box, ordered shell compaction, nucleus candidates ---------------------------------------
    {
      double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
      for (int q = ptid; q < npsuper; q += NPT) {
        int a = sa0 + q % g.sa, b = sb0 + (q / g.sa) % g.sb, c = sc0 + q / (g.sa * g.sb);
        if (a < na && b < nb && c < g.c1) {
          double x[3], wc[3], wcr[3];
          grid_point(g, a, b, c, x);
          grd_wrap(p.xf, x, g.periodic, wc, wcr);
#pragma unroll
          for (int d = 0; d < 3; d++) { lo[d] = fmin(lo[d], wcr[d]); hi[d] = fmax(hi[d], wcr[d]); }
        }
      }
      ws_prod_bbox<C>(lo, hi, S, S.p_bb, ptid);
    }
    if (ptid == 0) S.p_int[1] = 0;
    named_bar_sync(2, NPT);
    int ns = 0;
    for (int base = 0; base < w.nshell; base += NPT) {
      int s = base + ptid;
      bool act = false;
      if (s < w.nshell) {
        double4 xa = w.shell_xa[s];
        act = box_d2(S.p_bb, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12);
      }
      unsigned m = __ballot_sync(~0u, act);
      if (lane == 0) S.p_cnt[warp] = __popc(m);
      named_bar_sync(2, NPT);
      int off = ns, tot = 0;
#pragma unroll
      for (int i = 0; i < NPW; i++) {
        if (i < warp) off += S.p_cnt[i];
        tot += S.p_cnt[i];
      }
      if (act) sh_list[off + __popc(m & ((1u << lane) - 1u))] = s;
      ns += tot;
      named_bar_sync(2, NPT);
    }
    __shared__ int s_nuc_tmp[16];
    for (int i = ptid; i < w.natom; i += NPT) {
      if (box_d2(S.p_bb, w.atom_xyz[3 * i], w.atom_xyz[3 * i + 1], w.atom_xyz[3 * i + 2]) <= 1.0001e-10) {
        int k = atomicAdd(&S.p_int[1], 1);
        if (k < 16) s_nuc_tmp[k] = i;
      }
    }
    named_bar_sync(2, NPT);
    const int nnuc = S.p_int[1];

    for (int sub = 0; sub < nsub_a * nsub_b * nsub_c; sub++) {
      const int ta0 = sa0 + (sub % nsub_a) * g.ta, tb0 = sb0 + ((sub / nsub_a) % nsub_b) * g.tb;
      const int tc0 = sc0 + (sub / (nsub_a * nsub_b)) * g.tc;
      if (ta0 >= na || tb0 >= nb || tc0 >= g.c1) continue;
      const int tslot = tcount % C::TSLOTS;
      tcount++;
      double *tp = S.s_pt + tslot * 4 * P;
      // ---- points of the sub-tile -> tile-info slot (safe: consumers are at most NS sub-tiles behind) --------------
      {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        if (ptid < P) {
          int a = ta0 + ptid % g.ta, b = tb0 + (ptid / g.ta) % g.tb, c = tc0 + ptid / (g.ta * g.tb);
          bool valid = a < na && b < nb && c < g.c1;
          double x[3] = {0, 0, 0}, wc[3], wcr[3] = {0, 0, 0};
          if (valid) {
            grid_point(g, a, b, c, x);
            grd_wrap(p.xf, x, g.periodic, wc, wcr);
#pragma unroll
            for (int d = 0; d < 3; d++) { lo[d] = wcr[d]; hi[d] = wcr[d]; }
          }
          tp[ptid] = wcr[0]; tp[P + ptid] = wcr[1]; tp[2 * P + ptid] = wcr[2];
          tp[3 * P + ptid] = valid ? 1.0 : 0.0;
        }
        if (ptid == 0) {
          WsTileInfo &ti = S.tinfo[tslot];
          ti.a0 = ta0; ti.b0 = tb0; ti.c0 = tc0; ti.nnuc = nnuc;
          for (int k = 0; k < 16 && k < nnuc; k++) ti.nuc[k] = s_nuc_tmp[k];
        }
        ws_prod_bbox<C>(lo, hi, S, S.p_bb + 6, ptid);
      }
      // ---- S2: refine against the sub-tile box, expand to (primitive, shell) pairs (ordered) ------------------------
      int kact = 0;
      for (int base = 0; base < ns; base += NPT) {
        int j = base + ptid, cnt = 0, pb = 0, s = 0;
        if (j < ns) {
          s = sh_list[j];
          double4 xa = w.shell_xa[s];
          if (box_d2(S.p_bb + 6, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12)) {
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
        if (lane == 31) S.p_cnt[warp] = incl;
        named_bar_sync(2, NPT);
        int off = kact, tot = 0;
#pragma unroll
        for (int i = 0; i < NPW; i++) {
          if (i < warp) off += S.p_cnt[i];
          tot += S.p_cnt[i];
        }
        off += incl - cnt;
        if (cnt > 0) {
          const double4 xa = w.shell_xa[s];
          const double2 cut = w.shell_cut[s];
          for (int i = 0; i < cnt; i++) {
            pr_list[off + i] = make_int2(pb + i, s);
            const uchar4 l = reinterpret_cast<const uchar4 *>(w.prim_l)[pb + i];
            const long long bits = (long long)s | ((long long)l.x << 32) | ((long long)l.y << 40) | ((long long)l.z << 48);
            double4 *r = reinterpret_cast<double4 *>(rec + (size_t)(off + i) * 8);
            r[0] = xa;
            r[1] = make_double4(cut.x, cut.y, __longlong_as_double(bits), 0.0);
          }
        }
        kact += tot;
        named_bar_sync(2, NPT);
      }
      __threadfence_block();
      named_bar_sync(2, NPT);   // pr_list (global scratch) visible to every producer thread
      const int nch = (kact + KC - 1) / KC;
      const double px = tp[pt], py = tp[P + pt], pz = tp[2 * P + pt];
      const bool pvalid = tp[3 * P + pt] != 0.0;

      if (kact == 0) {   // nothing active: a data-less stage tells the consumers to emit zeros (+ EDF) for this tile
        const int stg = acquire();
        if (ptid == 0) {
          mbar_expect_tx(&S.full[stg], 0u);   // warp 0's second arrival (the one that carries the TMA byte count)
          WsStageHdr &h = S.hdr[stg];
          h.flags = 1 | 2 | 4 | 8; h.blk = 0; h.tslot = tslot;
        }
        publish(stg);
        continue;
      }
      auto load_rec = [&](int cc) -> double2 {
        return reinterpret_cast<const double2 *>(rec + (size_t)(cc * KC + grp * 4) * 8)[pt];
      };
      double2 rec_next = make_double2(0.0, 0.0);
      double *agbase = p.acache + ((size_t)blockIdx.x * 2 + (tcount & 1)) * (size_t)p.acache_chunks * C::A_STAGE;
      int *nzbase = p.acache_nz + ((size_t)blockIdx.x * 2 + (tcount & 1)) * (size_t)p.acache_chunks * 8;
      for (int blk = 0; blk < p.nblk; blk++) {
        if (blk == 1) {   // bulk stores of the chi tiles complete, k-quad masks visible -> safe to read them back
          if (ptid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
          __threadfence();
          named_bar_sync(2, NPT);
        }
        if (blk == 0 || nch > p.acache_chunks) rec_next = load_rec(blk == 0 ? 0 : p.acache_chunks);
        for (int c = 0; c < nch; c++) {
          const int stg = acquire();
          const bool cached = c < p.acache_chunks;           // this chunk's chi tile lives in the L2 cache
          const bool compute = blk == 0 || !cached;
          const bool save = blk == 0 && cached && p.nblk > 1;
          // B rows: one bulk TMA copy per active primitive of the chunk (warp 0: lane = row); A tile from the cache
          const int nrow = min(KC, kact - c * KC);
          if (warp == 0) {
            if (lane == 0)
              mbar_expect_tx(&S.full[stg], (uint32_t)(nrow * NB * sizeof(double)) +
                                               (compute ? 0u : (uint32_t)(C::A_STAGE * sizeof(double))));
            __syncwarp();
            if (lane < nrow) {
              const int prim = pr_list[c * KC + lane].x;
              tma_bulk_g2s(S.Bs + stg * C::B_STAGE + lane * SB, w.cmoT + (size_t)prim * w.ldc + (size_t)blk * NB,
                           (uint32_t)(NB * sizeof(double)), &S.full[stg]);
            }
            if (!compute) {
              if (lane == 0)
                tma_bulk_g2s(S.As + stg * C::A_STAGE, agbase + (size_t)c * C::A_STAGE,
                             (uint32_t)(C::A_STAGE * sizeof(double)), &S.full[stg]);
              if (lane < C::NKQ) S.hdr[stg].nzmask[lane] = nzbase[c * 8 + lane];
            }
            if (lane == 0) {
              WsStageHdr &h = S.hdr[stg];
              h.flags = (c == 0 ? 1 : 0) | (c == nch - 1 ? 2 : 0) | ((c == nch - 1 && blk == p.nblk - 1) ? 4 : 0);
              h.blk = blk; h.tslot = tslot;
            }
          }
          if (compute) {
            // A tile: this 16-lane group evaluates the 4 primitives of k-quad `grp` at its 16 points.  The quad's four
            // records (256 contiguous bytes) were loaded one chunk ahead, 16 bytes per lane, and are passed round
            // with shuffles: no dependent global load sits between acquiring the stage and the arithmetic.
            const double2 mine = rec_next;
            if (c + 1 < nch) rec_next = load_rec(c + 1);
            const size_t aoffq = (size_t)(grp * ND) * P * 4 + pt * 4;
            double *Aq = S.As + stg * C::A_STAGE + aoffq;
            unsigned nzbits = 0;
            int prev_sh = -1;
            double dx = 0, dy = 0, dz = 0, ex = 0, al = 0;
            bool pass = false;
            const int sbase = lane & 16;
#pragma unroll
            for (int k4 = 0; k4 < 4; k4++) {
              const int gk = c * KC + grp * 4 + k4;
              const double rx = __shfl_sync(~0u, mine.x, sbase + k4 * 4), ry = __shfl_sync(~0u, mine.y, sbase + k4 * 4);
              const double rz = __shfl_sync(~0u, mine.x, sbase + k4 * 4 + 1), ra = __shfl_sync(~0u, mine.y, sbase + k4 * 4 + 1);
              const double rc = __shfl_sync(~0u, mine.x, sbase + k4 * 4 + 2), rc2 = __shfl_sync(~0u, mine.y, sbase + k4 * 4 + 2);
              const long long bits = __double_as_longlong(__shfl_sync(~0u, mine.x, sbase + k4 * 4 + 3));
              double chi[ND];
#pragma unroll
              for (int d = 0; d < ND; d++) chi[d] = 0.0;
              if (gk < kact && pvalid) {
                const int sh = (int)(bits & 0xffffffffll);
                if (sh != prev_sh) {   // new shell: distance screen (:2747) and exponential (:2749) shared by its primitives
                  prev_sh = sh;
                  dx = px - rx; dy = py - ry; dz = pz - rz;
                  al = ra;
                  const double d2 = dx * dx + dy * dy + dz * dz;
                  pass = within_cut(dx, dy, dz, d2, rc, rc2);
                  ex = pass ? exp(-al * d2) : 0.0;
                }
                if (pass) {
                  prim_chi<ND>(dx, dy, dz, (int)((bits >> 32) & 0xff), (int)((bits >> 40) & 0xff), (int)((bits >> 48) & 0xff),
                               al, ex, chi);
                  nzbits |= 1u;
                }
              }
              const int sw = k4 ^ ((pt >> 2) & 3);
#pragma unroll
              for (int d = 0; d < ND; d++) Aq[d * P * 4 + sw] = chi[d];
            }
            // k-quad mask: bit wm set if any of the 8 points of row group wm has an unmasked primitive in this quad
            const unsigned bal = __ballot_sync(~0u, nzbits != 0);
            const unsigned half = (bal >> (lane & 16)) & 0xffffu;
            if (pt == 0) {
              const int m = ((half & 0xffu) ? 1 : 0) | ((half & 0xff00u) ? 2 : 0);
              S.hdr[stg].nzmask[grp] = m;
              if (save) nzbase[c * 8 + grp] = m;
            }
            if (save) {
              // copy the finished tile to the L2 cache with one bulk TMA store (shared -> global)
              fence_proxy_async();
              named_bar_sync(2, NPT);
              if (ptid == 0) {
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // previous store has read its stage
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(agbase + (size_t)c * C::A_STAGE),
                             "r"(smem_u32(S.As + stg * C::A_STAGE)), "r"((uint32_t)(C::A_STAGE * sizeof(double)))
                             : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              }
            }
          }
          publish(stg);
        }
      }
    }
  }
  // exit stage
  {
    const int stg = acquire();
    if (ptid == 0) {
      mbar_expect_tx(&S.full[stg], 0u);
      S.hdr[stg].flags = 16; S.hdr[stg].blk = 0; S.hdr[stg].tslot = 0;
    }
    publish(stg);
  }
}
// ---------------------------------------------------------------------------------------------------------------
// Split producer (config "WG").  ncu on the version above: the consumers still waited ~17 % of the time on the
// stages whose chi tile had to be computed (MO block 0), while the producer warps idled ~60 % of the time during the
// cached blocks.  Here the two jobs are decoupled:
//   * GENERATOR warps (3, named barrier 2 with 96 threads) run one sub-tile AHEAD: claim super-tiles, screen shells,
//     evaluate chi for every active primitive of the sub-tile and write the tiles (in the shared-memory layout of an
//     A stage) straight to the per-CTA L2 cache buffer `tcount & 1`;
//   * the STREAMER warp only moves data: for every (MO block, chunk) it acquires a stage and issues the bulk TMA
//     copies of the cached chi tile and of the C rows.  Every consumer stage is therefore a pure TMA stage.
// Hand-over: ready[b] (generator -> streamer, tile in buffer b complete) and freeb[b] (streamer -> generator, every
// TMA load of the tile in buffer b has landed), both single-arrival mbarriers used with the usual phase parity.
// ---------------------------------------------------------------------------------------------------------------
template <class C>
__device__ void ws_generator(const WfnParams &p, const WsShared<C> &S, int gtid) {
  constexpr int ND = C::ND, P = C::P, KC = C::KC, NGW = C::NPW - 1, NGT = 32 * NGW, NGRP = NGT / 16;
  const int lane = gtid & 31, warp = gtid >> 5;
  const int grp = gtid >> 4;        // 16-lane group: one k-quad (4 primitives x 16 points) at a time
  const int pt = gtid & 15;
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  // per-CTA scratch: shell list [shell_cap] then two (primitive, shell) lists [prim_cap] (one per cache buffer)
  int *sh_list = p.scratch + (size_t)blockIdx.x * (p.shell_cap + 4 * p.prim_cap);
  int2 *pr_list0 = reinterpret_cast<int2 *>(sh_list + p.shell_cap);
  double *rec = p.ws_rec + (size_t)blockIdx.x * (size_t)p.prim_cap * 8;
  const int na = grid_na(g), nb = grid_nb(g);
  const int npsuper = g.sa * g.sb * g.sc;
  const int nsub_a = g.sa / g.ta, nsub_b = g.sb / g.tb, nsub_c = g.sc / g.tc;
  unsigned tcount = 0;
  __shared__ int s_nuc_tmp[16];

  for (;;) {
    if (gtid == 0) S.p_int[0] = atomicAdd(p.work_counter, 1);
    named_bar_sync(2, NGT);
    const int st = S.p_int[0];
    if (st >= p.nsa * p.nsb * p.nsc) break;
    const int sa0 = (st % p.nsa) * g.sa, sb0 = ((st / p.nsa) % p.nsb) * g.sb;
    const int sc0 = g.c0 + (g.c_part + (st / (p.nsa * p.nsb)) * g.c_nparts) * g.sc;
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
      mbar_wait(&S.freeb[buf], ((tcount >> 1) & 1u) ^ 1u);
      // ---- S2: refine against the sub-tile box, expand to (primitive, shell) pairs + 64-byte records (ordered) ------
      int kact = 0;
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
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          int v = __shfl_up_sync(~0u, incl, o);
          if (lane >= o) incl += v;
        }
        if (lane == 31) S.p_cnt[warp] = incl;
        named_bar_sync(2, NGT);
        int off = kact, tot = 0;
#pragma unroll
        for (int i = 0; i < NGW; i++) {
          if (i < warp) off += S.p_cnt[i];
          tot += S.p_cnt[i];
        }
        off += incl - cnt;
        if (cnt > 0) {
          const double4 xa = w.shell_xa[s];
          const double2 cut = w.shell_cut[s];
          for (int i = 0; i < cnt; i++) {
            pr_list[off + i] = make_int2(pb + i, s);
            const uchar4 l = reinterpret_cast<const uchar4 *>(w.prim_l)[pb + i];
            const long long bits = (long long)s | ((long long)l.x << 32) | ((long long)l.y << 40) | ((long long)l.z << 48);
            double4 *r = reinterpret_cast<double4 *>(rec + (size_t)(off + i) * 8);
            r[0] = xa;
            r[1] = make_double4(cut.x, cut.y, __longlong_as_double(bits), 0.0);
          }
        }
        kact += tot;
        named_bar_sync(2, NGT);
      }
      __threadfence_block();
      named_bar_sync(2, NGT);   // records visible to every generator thread
      const int nch = (kact + KC - 1) / KC;
      const double px = tp[pt], py = tp[P + pt], pz = tp[2 * P + pt];
      const bool pvalid = tp[3 * P + pt] != 0.0;
      double *agbase = p.acache + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * C::A_STAGE;
      int *nzbase = p.acache_nz + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * 8;
      // ---- chi tiles: k-quad q = (chunk, quad within the chunk) -> 16-lane group q % NGRP ------------------------------
      const int nquad = nch * (KC / 4);
      const int sbase = lane & 16;
      int prev_sh = -1;
      double dx = 0, dy = 0, dz = 0, ex = 0, al = 0;
      bool pass = false;
      for (int q = grp; q < nquad; q += NGRP) {
        const double2 mine = reinterpret_cast<const double2 *>(rec + (size_t)q * 32)[pt];
        double *Aq = agbase + (size_t)(q >> 3) * C::A_STAGE + (size_t)((q & 7) * ND) * P * 4 + pt * 4;
        unsigned nzbits = 0;
#pragma unroll
        for (int k4 = 0; k4 < 4; k4++) {
          const int gk = q * 4 + k4;
          const double rx = __shfl_sync(~0u, mine.x, sbase + k4 * 4), ry = __shfl_sync(~0u, mine.y, sbase + k4 * 4);
          const double rz = __shfl_sync(~0u, mine.x, sbase + k4 * 4 + 1), ra = __shfl_sync(~0u, mine.y, sbase + k4 * 4 + 1);
          const double rc = __shfl_sync(~0u, mine.x, sbase + k4 * 4 + 2), rc2 = __shfl_sync(~0u, mine.y, sbase + k4 * 4 + 2);
          const long long bits = __double_as_longlong(__shfl_sync(~0u, mine.x, sbase + k4 * 4 + 3));
          double chi[ND];
#pragma unroll
          for (int d = 0; d < ND; d++) chi[d] = 0.0;
          if (gk < kact && pvalid) {
            const int sh = (int)(bits & 0xffffffffll);
            if (sh != prev_sh) {   // new shell: distance screen (:2747) and exponential (:2749) shared by its primitives
              prev_sh = sh;
              dx = px - rx; dy = py - ry; dz = pz - rz;
              al = ra;
              const double d2 = dx * dx + dy * dy + dz * dz;
              pass = within_cut(dx, dy, dz, d2, rc, rc2);
              ex = pass ? exp(-al * d2) : 0.0;
            }
            if (pass) {
              prim_chi<ND>(dx, dy, dz, (int)((bits >> 32) & 0xff), (int)((bits >> 40) & 0xff), (int)((bits >> 48) & 0xff),
                           al, ex, chi);
              nzbits |= 1u;
            }
          }
          const int sw = k4 ^ ((pt >> 2) & 3);
#pragma unroll
          for (int d = 0; d < ND; d++) Aq[d * P * 4 + sw] = chi[d];
        }
        const unsigned bal = __ballot_sync(~0u, nzbits != 0);
        const unsigned half = (bal >> (lane & 16)) & 0xffffu;
        if (pt == 0) nzbase[q] = ((half & 0xffu) ? 1 : 0) | ((half & 0xff00u) ? 2 : 0);
      }
      // tile complete: make the generic-proxy global writes visible to the streamer's bulk (async-proxy) loads
      __threadfence();
      fence_proxy_async();
      named_bar_sync(2, NGT);
      if (gtid == 0) {
        WsTileDesc &td = S.tdesc[buf];
        td.kact = kact; td.nch = nch; td.tslot = tslot; td.flags = 0;
        mbar_arrive(&S.ready[buf]);
      }
      tcount++;
    }
  }
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
    fence_proxy_async();
    const WsTileDesc td = S.tdesc[buf];
    if (td.flags & 1) break;
    const int2 *pr_list = pr_list0 + (size_t)buf * p.prim_cap;
    const double *agbase = p.acache + ((size_t)blockIdx.x * 2 + buf) * (size_t)p.acache_chunks * C::A_STAGE;
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
        for (int c = 0; c < td.nch; c++) {
          const int nrow = min(KC, td.kact - c * KC);
          const int prim = lane < nrow ? pr_list[c * KC + lane].x : 0;         // loads issued before the wait
          const int nz = lane < C::NKQ ? nzbase[c * 8 + lane] : 0;
          const int stg = acquire();
          if (lane < C::NKQ) S.hdr[stg].nzmask[lane] = nz;
          if (lane == 0) {
            WsStageHdr &h = S.hdr[stg];
            h.flags = (c == 0 ? 1 : 0) | (c == td.nch - 1 ? 2 : 0) | ((c == td.nch - 1 && blk == p.nblk - 1) ? 4 : 0);
            h.blk = blk; h.tslot = td.tslot;
          }
          __syncwarp();
          if (lane == 0) {
            mbar_expect_tx(&S.full[stg], (uint32_t)(nrow * NB * sizeof(double)) + (uint32_t)(C::A_STAGE * sizeof(double)));
            tma_bulk_g2s(S.As + stg * C::A_STAGE, agbase + (size_t)c * C::A_STAGE, (uint32_t)(C::A_STAGE * sizeof(double)),
                         &S.full[stg]);
          }
          __syncwarp();
          if (lane < nrow)
            tma_bulk_g2s(S.Bs + stg * C::B_STAGE + lane * SB, w.cmoT + (size_t)prim * w.ldc + (size_t)blk * NB,
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
// Consumer side: 256 threads, named barrier 1
// ---------------------------------------------------------------------------------------------------------------
// MMAs of one stage for a consumer warp that owns NF fragments (8-column groups) of the MO block.
template <class C, int NF>
__device__ __forceinline__ void ws_mma_stage(double (&acc)[C::ND][C::NI][2], const double *Ast, const double *Bst,
                                             const int *nzmask, int wm) {
  constexpr int ND = C::ND, P = C::P, SB = C::SB;
#pragma unroll 1
  for (int kq = 0; kq < C::NKQ; kq++) {
    if (!((nzmask[kq] >> wm) & 1)) continue;             // all 8 points x 4 primitives masked (warp-uniform)
    double b[NF];
#pragma unroll
    for (int ni = 0; ni < NF; ni++) b[ni] = Bst[kq * 4 * SB + ni * 8];
#pragma unroll
    for (int d = 0; d < ND; d++) {
      const double a = Ast[(size_t)(kq * ND + d) * P * 4];
#pragma unroll
      for (int ni = 0; ni < NF; ni++) dmma884(acc[d][ni][0], acc[d][ni][1], a, b[ni]);
    }
  }
}

template <class C>
__device__ void ws_consumer(const WfnParams &p, const WsShared<C> &S, int ctid) {
  constexpr int ND = C::ND, P = C::P, NS = C::NS, NB = C::NB, SB = C::SB, NI = C::NI, NCT = C::NCT, NKQ = C::NKQ;
  const int lane = ctid & 31, cw = ctid >> 5;
  const int wm = cw & 1, wn = cw >> 1;                 // row group (8 points) and warp column (0..3)
  const int nfw = wn < 2 ? 4 : 3;                      // fragments of this warp: 4,4,3,3 -> 7 per SM sub-partition
  const int f0 = wn < 2 ? wn * 4 : 8 + (wn - 2) * 3;
  const int row = lane >> 2;
  const int aoff = row * 4 + ((lane & 3) ^ (((wm * 8 + row) >> 2) & 3));
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  double acc[ND][NI][2];
#pragma unroll
  for (int d = 0; d < ND; d++)
#pragma unroll
    for (int ni = 0; ni < NI; ni++) acc[d][ni][0] = acc[d][ni][1] = 0.0;

  for (unsigned it = 0;; it++) {
    const int stg = it % NS;
    mbar_wait(&S.full[stg], (it / NS) & 1u);
    const WsStageHdr &h = S.hdr[stg];
    const int flags = h.flags, blk = h.blk, tslot = h.tslot;
    if (flags & 16) break;
    if (!(flags & 8)) {
      const double *Ast = S.As + stg * C::A_STAGE + (size_t)wm * 32 + aoff;
      const double *Bst = S.Bs + stg * C::B_STAGE + (lane & 3) * SB + f0 * 8 + (lane >> 2);
      // warp-uniform choice of the fragment count: a predicated-off DMMA still occupies the tensor pipe
      if (nfw == 4) ws_mma_stage<C, 4>(acc, Ast, Bst, h.nzmask, wm);
      else ws_mma_stage<C, 3>(acc, Ast, Bst, h.nzmask, wm);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&S.empty[stg]);           // stage (and its header) no longer needed by this warp
    if (!(flags & 2)) continue;

    // ---- end of an MO block: occupation-weighted quadratic forms of this thread's MOs (:2091-2147) ---------------
    if (!(flags & 8)) {
      double q[ND];
#pragma unroll
      for (int d = 0; d < ND; d++) q[d] = 0.0;
#pragma unroll
      for (int ni = 0; ni < NI; ni++)
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const double oc = (ni < 3 || nfw == 4) ? w.occ[blk * NB + (f0 + ni) * 8 + 2 * (lane & 3) + j] : 0.0;
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
      for (int d = 0; d < ND; d++) {
        q[d] += __shfl_xor_sync(~0u, q[d], 1);
        q[d] += __shfl_xor_sync(~0u, q[d], 2);
      }
      if ((lane & 3) == 0) {
#pragma unroll
        for (int d = 0; d < ND; d++) S.s_part[(((blk & 1) * 4 + wn) * P + wm * 8 + row) * ND + d] = q[d];
      }
#pragma unroll
      for (int d = 0; d < ND; d++)
#pragma unroll
        for (int ni = 0; ni < NI; ni++) acc[d][ni][0] = acc[d][ni][1] = 0.0;
    }
    named_bar_sync(1, NCT);
    if (!(flags & 8) && ctid < P * ND) {                 // fixed-order fold of the four warp columns
      double sum = S.s_tot[ctid];
#pragma unroll
      for (int v = 0; v < 4; v++) sum += S.s_part[((blk & 1) * 4 + v) * P * ND + ctid];
      S.s_tot[ctid] = sum;
    }
    if (!(flags & 4)) continue;

    // ---- end of the sub-tile: EDF core, nucleus test, derived quantities, stores ---------------------------------
    named_bar_sync(1, NCT);
    const double *tp = S.s_pt + tslot * 4 * P;
    const WsTileInfo &ti = S.tinfo[tslot];
    if (w.nedf > 0) {
      const int pt = ctid & 15, slot = ctid >> 4;        // 8 slots (threads 0..127) through the free s_part area
      if (ctid < 128) {
        double e[ND];
#pragma unroll
        for (int d = 0; d < ND; d++) e[d] = 0.0;
        if (tp[3 * P + pt] != 0.0) {
          const double px = tp[pt], py = tp[P + pt], pz = tp[2 * P + pt];
          for (int i = slot; i < w.nedf; i += 8) {
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
        for (int d = 0; d < ND; d++) S.s_part[(slot * P + pt) * ND + d] = e[d];
      }
      named_bar_sync(1, NCT);
      if (ctid < P * ND) {
        double sum = S.s_tot[ctid];
        for (int v = 0; v < 8; v++) sum += S.s_part[v * P * ND + ctid];
        S.s_tot[ctid] = sum;
      }
      named_bar_sync(1, NCT);
    }
    if (ctid < P) {
      double *t = S.s_tot + ctid * ND;
      if (tp[3 * P + ctid] != 0.0) {
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
        emit_point(p.o, out_index(g, a, b, c), f, gx, gy, gz, hs, isnuc, p.nci_const);
      }
#pragma unroll
      for (int d = 0; d < ND; d++) t[d] = 0.0;           // ready for the next sub-tile
    }
  }
}

template <class C, bool SPLIT>
__global__ void __launch_bounds__(C::NT, 1) wfn_ws_kernel(const __grid_constant__ WfnParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t s_full[C::NS], s_empty[C::NS];
  __shared__ WsStageHdr s_hdr[C::NS];
  __shared__ WsTileInfo s_tinfo[C::TSLOTS];
  __shared__ double s_pred[C::NPW * 6], s_pbb[12];
  __shared__ int s_pcnt[C::NPW], s_pint[4];
  __shared__ __align__(8) uint64_t s_ready[2], s_freeb[2];
  __shared__ WsTileDesc s_tdesc[2];
  WsShared<C> S;
  S.As = reinterpret_cast<double *>(smem_raw);
  S.Bs = S.As + C::NS * C::A_STAGE;
  S.s_part = S.Bs + C::NS * C::B_STAGE;
  S.s_tot = S.s_part + C::NPART;
  S.s_pt = S.s_tot + C::P * C::ND;
  S.full = s_full; S.empty = s_empty; S.hdr = s_hdr; S.tinfo = s_tinfo;
  S.p_red = s_pred; S.p_bb = s_pbb; S.p_cnt = s_pcnt; S.p_int = s_pint;
  S.ready = s_ready; S.freeb = s_freeb; S.tdesc = s_tdesc;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < C::NS; s++) {
      // unsplit: one arrival per producer warp + the expect_tx arrival of warp 0; split: the streamer's expect_tx
      mbar_init(&s_full[s], SPLIT ? 1 : C::NPW + 1);
      mbar_init(&s_empty[s], C::NCW);      // one arrival per consumer warp
    }
    for (int b = 0; b < 2; b++) { mbar_init(&s_ready[b], 1); mbar_init(&s_freeb[b], 1); }
    fence_mbar_init();
  }
  for (int i = tid; i < C::NS * C::B_STAGE; i += C::NT) S.Bs[i] = 0.0;   // stale rows stay finite (they meet chi = 0)
  for (int i = tid; i < C::P * C::ND; i += C::NT) S.s_tot[i] = 0.0;
  fence_proxy_async();
  __syncthreads();
  if (tid >= C::NCT) {
    reg_dec<C::REG_PROD>();
    if (!SPLIT) ws_producer<C>(p, S, tid - C::NCT);
    else if (tid < C::NCT + 32) ws_streamer<C>(p, S, tid - C::NCT);
    else ws_generator<C>(p, S, tid - C::NCT - 32);
  } else {
    reg_inc<C::REG_CONS>();
    ws_consumer<C>(p, S, tid);
  }
}
#endif

}  // namespace c2
