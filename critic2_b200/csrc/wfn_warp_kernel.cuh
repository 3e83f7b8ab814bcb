// wfn_warp_kernel.cuh -- warp-autonomous wavefunction kernel for small MO counts (one MO block of 8 / 16 / 24 columns).
//
// Same mathematics as wfn_kernel.cuh (see there for the reference citations: calculate_mo_gto
// src/wfn_private@proc.F90:2673-2781, rho2 :2032-2228, calculate_edf :2783-2873, the tail of grd
// src/fieldmod@proc.f90:834-840).  Different execution model, for the regime the block-synchronous S1 / S3w configurations
// handled badly (BASELINE configs 1 and 2: 12 % and 36 % of the DFMA peak; ncu: 25 % of the warp slots active, 2.2-2.6 warps
// per issue stalled at the per-chunk block barrier, instruction-cache misses from three unrolled phases run for a few
// hundred cycles each).  With 5-24 MOs the contraction is tiny and nothing is re-used across MO blocks, so the shared-memory
// staging, the TMA gather and the block barriers buy nothing:
//   * a WARP owns a brick of 8 points (2x2x2 on a lattice, 8 consecutive points of a list) from screening to stores;
//   * lane (r, k) = (lane >> 2, lane & 3) evaluates primitive k of the current k-quad at point r -- which IS the A fragment
//     of mma.sync m8n8k4 (row r, column k): chi goes from the FP64 pipe straight into the DMMA, no shared memory;
//   * the B fragment C[prim k][column r] is read from the L1-resident coefficient rows (62 KB for benzene);
//   * the brick-level screen (shell sphere against the brick's box) is a warp ballot over the CTA's shell list; passing
//     primitives are compacted into a per-warp list and consumed four at a time;
//   * k-quads without an unmasked (point, primitive) pair are skipped; the exact per-point screen of the reference
//     (dist > dran -> 0, :2747) masks the rest;
//   * the occupation-weighted sums (:2091-2147) stay in the accumulator layout (a lane holds two MOs of one point per
//     fragment), EDF terms are dealt over the four lanes of a point, one quad shuffle folds everything, lane k = 0 runs
//     the tail of grd.  No block barrier inside a super-tile; bricks are claimed dynamically from a shared-memory counter.
#pragma once
#include "wfn_kernel.cuh"

namespace c2 {

template <int ND_, int NF_, int CTAS_>
struct WarpCfg {
  static constexpr int ND = ND_, NF = NF_, NB = 8 * NF_, CTAS = CTAS_;
  static constexpr int NW = 4, NT = 32 * NW;   // warps / threads per CTA
  static constexpr int P = 8;                  // points per brick
  static constexpr int LCAP = 704;             // per-warp primitive list: 32 shells x 21 Cartesian h functions + a carried quad
};

#ifdef __CUDACC__
template <class C>
__global__ void __launch_bounds__(C::NT, C::CTAS) wfn_warp_kernel(const __grid_constant__ WfnParams p) {
  constexpr int ND = C::ND, NF = C::NF, NT = C::NT, NW = C::NW, LCAP = C::LCAP;
  __shared__ int2 s_list[NW][LCAP];
  __shared__ int s_cnt[NW + 1];
  __shared__ int s_int[8];                // 0 work item, 1 next brick, 3 nucleus candidates
  __shared__ int s_nuc[16];
  __shared__ double s_red[NW * 6];
  __shared__ double s_bb[6];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row = lane >> 2, kk = lane & 3;
  const WfnDev &w = p.w;
  const GridDesc &g = p.g;
  int *sh_list = p.scratch + (size_t)blockIdx.x * (p.shell_cap + p.prim_cap);
  int2 *wl = s_list[warp];
  const int na = grid_na(g), nb = grid_nb(g);
  const int npsuper = g.sa * g.sb * g.sc;
  const int nsub_a = g.sa / g.ta, nsub_b = g.sb / g.tb, nsub_c = g.sc / g.tc;
  const int nsub = nsub_a * nsub_b * nsub_c;

  for (;;) {
    __syncthreads();
    if (tid == 0) { s_int[0] = atomicAdd(p.work_counter, 1); s_int[1] = 0; s_int[3] = 0; }
    __syncthreads();
    const int st = s_int[0];
    if (st >= p.nitems) break;
    int ia, ib, ic;
    part_item(g, st, p.nsa, p.nsb, ia, ib, ic);
    const int sa0 = ia * g.sa, sb0 = ib * g.sb, sc0 = g.c0 + ic * g.sc;
    // ---- S1 (CTA): super-tile box, ordered compaction of the shells that reach it, nucleus candidates ------------------
    {
      double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
      for (int q = tid; q < npsuper; q += NT) {
        const int a = sa0 + q % g.sa, b = sb0 + (q / g.sa) % g.sb, c = sc0 + q / (g.sa * g.sb);
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
    int ns = 0;
    for (int base = 0; base < w.nshell; base += NT) {
      const int s = base + tid;
      bool act = false;
      if (s < w.nshell) {
        const double4 xa = w.shell_xa[s];
        act = box_d2(s_bb, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12);
      }
      const unsigned m = __ballot_sync(~0u, act);
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
    for (int i = tid; i < w.natom; i += NT) {
      if (box_d2(s_bb, w.atom_xyz[3 * i], w.atom_xyz[3 * i + 1], w.atom_xyz[3 * i + 2]) <= 1.0001e-10) {
        const int k = atomicAdd(&s_int[3], 1);
        if (k < 16) s_nuc[k] = i;
      }
    }
    __threadfence_block();
    __syncthreads();
    const int nnuc = s_int[3];

    // ---- bricks of 8 points, claimed dynamically by the warps -----------------------------------------------------------
    for (;;) {
      int sub = 0;
      if (lane == 0) sub = atomicAdd(&s_int[1], 1);
      sub = __shfl_sync(~0u, sub, 0);
      if (sub >= nsub) break;
      const int ta0 = sa0 + (sub % nsub_a) * g.ta, tb0 = sb0 + ((sub / nsub_a) % nsub_b) * g.tb;
      const int tc0 = sc0 + (sub / (nsub_a * nsub_b)) * g.tc;
      if (ta0 >= na || tb0 >= nb || tc0 >= g.c1) continue;
      const int a = ta0 + row % g.ta, b = tb0 + (row / g.ta) % g.tb, c = tc0 + row / (g.ta * g.tb);
      const bool valid = a < na && b < nb && c < g.c1;
      double px = 0, py = 0, pz = 0;
      double bb[6] = {1e300, 1e300, 1e300, -1e300, -1e300, -1e300};
      if (valid) {
        double x[3], wc[3], wcr[3];
        grid_point(g, a, b, c, x);
        grd_wrap(p.xf, x, g.periodic, wc, wcr);
        px = wcr[0]; py = wcr[1]; pz = wcr[2];
        bb[0] = bb[3] = px; bb[1] = bb[4] = py; bb[2] = bb[5] = pz;
      }
#pragma unroll
      for (int d = 0; d < 3; d++)
#pragma unroll
        for (int o = 16; o >= 4; o >>= 1) {       // the four lanes of a point hold the same coordinates
          bb[d] = fmin(bb[d], __shfl_xor_sync(~0u, bb[d], o));
          bb[3 + d] = fmax(bb[3 + d], __shfl_xor_sync(~0u, bb[3 + d], o));
        }

      double acc[ND][NF][2];
#pragma unroll
      for (int d = 0; d < ND; d++)
#pragma unroll
        for (int ni = 0; ni < NF; ni++) acc[d][ni][0] = acc[d][ni][1] = 0.0;

      // ---- brick-level screen: ballot over the CTA's shell list, ordered expansion into the warp's primitive list; the
      // list is consumed four primitives at a time (one call site of the MMA loop: the accumulators stay in registers) -----
      int nlist = 0, base = 0;
      bool more = bb[0] <= bb[3] && ns > 0;
      while (more) {
        for (; base < ns; base += 32) {
          const int j = base + lane;
          int cnt = 0, pb = 0, s = 0;
          if (j < ns) {
            s = sh_list[j];
            const double4 xa = w.shell_xa[s];
            if (box_d2(bb, xa.x, xa.y, xa.z) <= w.shell_cut[s].y * (1.0 + 1e-12)) {
              pb = w.shell_pbeg[s];
              cnt = w.shell_pbeg[s + 1] - pb;
            }
          }
          int incl = cnt;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(~0u, incl, o);
            if (lane >= o) incl += v;
          }
          const int tot = __shfl_sync(~0u, incl, 31);
          if (nlist + tot > LCAP) break;            // no room: consume what is there first (this step is redone)
          const int off = nlist + incl - cnt;
          for (int i = 0; i < cnt; i++) wl[off + i] = make_int2(pb + i, s);
          nlist += tot;
        }
        more = base < ns;
        if (!more && (nlist & 3)) {                 // end of the shell list: pad the last quad
          if (lane < 4 - (nlist & 3)) wl[nlist + lane] = make_int2(-1, 0);
          nlist = (nlist + 3) & ~3;
        }
        __syncwarp();
        const int nq = nlist >> 2;
        // k-quads [0, nq): chi in the A-fragment layout straight from the FP64 pipe, C rows from L1, DMMA.  (A software-
        // pipelined, branch-free version of this loop -- chi of quad q+1 scheduled between the DMMAs of quad q, a
        // branch-free exp -- measured the same: 13.6 against 13.4 ms on benzene 200^3; so did skipping fully masked quads
        // after the evaluation (14.3 ms).  Three CTAs per SM at 168 registers: 14.7 ms with that chi code, 12.6 ms with the
        // shell-wise one, which is what ships.  See DESIGN.md.)
#pragma unroll 1
        for (int q = 0; q < nq; q++) {
          const int2 e = wl[4 * q + kk];
          const bool real = e.x >= 0;
          const int prim = real ? e.x : 0;
          double chi[ND];
#pragma unroll
          for (int d = 0; d < ND; d++) chi[d] = 0.0;
          bool nz = false;
          const uchar4 l = reinterpret_cast<const uchar4 *>(w.prim_l)[prim];
          const bool big = __any_sync(~0u, real && (l.x | l.y | l.z) > 2);   // f shells and up: powers by multiplication
          if (real && valid) {
            const double4 xa = w.shell_xa[e.y];
            const double2 cut = w.shell_cut[e.y];
            ShellAxis ax, ay;
            ax.x = px - xa.x; ay.x = py - xa.y;
            const double zr = pz - xa.z;
            ax.x2 = ax.x * ax.x; ay.x2 = ay.x * ay.x;
            const double z2 = zr * zr;
            const double d2 = __dadd_rn(__dadd_rn(ax.x2, ay.x2), z2);
            if (within_cut(ax.x, ay.x, zr, d2, cut.x, cut.y)) {                   // :2747
              const double ta = xa.w + xa.w;
              ShellAxisE az;
              shell_axis_setup(ax, ta);
              shell_axis_setup(ay, ta);
              shell_axis_setup(az, zr, z2, ta, exp_neg(-xa.w * d2));              // :2749
              if (big) shell_prim_chi<ND, true>(ax, ay, az, l.x, l.y, l.z, ta, p.chi_mode, chi);   // :2755-2768
              else shell_prim_chi<ND, false>(ax, ay, az, l.x, l.y, l.z, ta, p.chi_mode, chi);
              nz = true;
            }
          }
          if (__ballot_sync(~0u, nz) == 0u) continue;
          const double *crow = w.cmoT + (size_t)prim * w.ldc + row;
          double bf[NF];
#pragma unroll
          for (int ni = 0; ni < NF; ni++) bf[ni] = __ldg(crow + ni * 8);
#pragma unroll
          for (int d = 0; d < ND; d++)
#pragma unroll
            for (int ni = 0; ni < NF; ni++) dmma884(acc[d][ni][0], acc[d][ni][1], chi[d], bf[ni]);
        }
        __syncwarp();
        if (more) {                                 // carry the incomplete quad to the front of the list
          const int rem = nlist & 3;
          int2 keep = make_int2(-1, 0);
          if (lane < rem) keep = wl[4 * nq + lane];
          __syncwarp();
          if (lane < rem) wl[lane] = keep;
          nlist = rem;
          __syncwarp();
        } else nlist = 0;
      }

      // ---- occupation-weighted sums of this lane's MOs (:2091-2147), EDF core (:2783-2873), quad fold, tail of grd --------
      double q[ND];
#pragma unroll
      for (int d = 0; d < ND; d++) q[d] = 0.0;
#pragma unroll
      for (int ni = 0; ni < NF; ni++)
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const double oc = w.occ[ni * 8 + 2 * kk + j];
          const double v0 = acc[0][ni][j];
          q[0] += (ND == 1 && (p.chi_mode & 4)) ? oc * v0 : oc * v0 * v0;
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
      if (w.nedf > 0 && valid) {
        for (int i = kk; i < w.nedf; i += 4) {
          const double4 xa = w.edf_xa[i];
          const double4 cc = w.edf_cc[i];
          const double dx = px - xa.x, dy = py - xa.y, dz = pz - xa.z;
          const double d2 = dx * dx + dy * dy + dz * dz;
          if (!within_cut(dx, dy, dz, d2, cc.y, cc.z)) continue;
          const uchar4 l = reinterpret_cast<const uchar4 *>(w.edf_l)[i];
          double chi[ND];
          prim_chi<ND>(dx, dy, dz, l.x, l.y, l.z, xa.w, exp(-xa.w * d2), chi);
#pragma unroll
          for (int d = 0; d < ND; d++) q[d] += cc.x * chi[d];
        }
      }
#pragma unroll
      for (int d = 0; d < ND; d++) {
        q[d] += __shfl_xor_sync(~0u, q[d], 1);
        q[d] += __shfl_xor_sync(~0u, q[d], 2);
      }
      if (valid && kk == 0) {
        double gx = 0, gy = 0, gz = 0, hs[6] = {0, 0, 0, 0, 0, 0};
        if (ND > 1) { gx = q[ND > 1 ? 1 : 0]; gy = q[ND > 1 ? 2 : 0]; gz = q[ND > 1 ? 3 : 0]; }
        if (ND > 4) {
#pragma unroll
          for (int d = 0; d < 6; d++) hs[d] = q[ND > 4 ? 4 + d : 0];
        }
        bool isnuc = false;
        if (nnuc > 16) {
          for (int i = 0; i < w.natom && !isnuc; i++) {
            const double ddx = w.atom_xyz[3 * i] - px, ddy = w.atom_xyz[3 * i + 1] - py, ddz = w.atom_xyz[3 * i + 2] - pz;
            isnuc = (ddx * ddx + ddy * ddy + ddz * ddz) < 1e-10;      // src/environmod@proc.f90:677
          }
        } else {
          for (int k = 0; k < nnuc; k++) {
            const int i = s_nuc[k];
            const double ddx = w.atom_xyz[3 * i] - px, ddy = w.atom_xyz[3 * i + 1] - py, ddz = w.atom_xyz[3 * i + 2] - pz;
            isnuc = isnuc || (ddx * ddx + ddy * ddy + ddz * ddz) < 1e-10;
          }
        }
        emit_point_call(p.o, out_index(g, a, b, c), q[0], gx, gy, gz, hs[0], hs[1], hs[2], hs[3], hs[4], hs[5], isnuc,
                        p.nci_const);
      }
      __syncwarp();
    }
  }
}
#endif

}  // namespace c2
