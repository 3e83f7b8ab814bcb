// promol_kernel.cuh -- promolecular density on a lattice: atom-binned radial-table sums.
//
// Replaces, per lattice, field%grd(type_promol) -> crystal%promolecular -> environ%promolecular
// (src/fieldmod@proc.f90:802-808,834-840; src/environmod@proc.f90:1202-1327) with grid1%interp
// (src/grid1mod@proc.f90:84-135) and the neighbour search list_near_atoms (src/environmod@proc.f90:895-1047).
//
// One CTA per tile of PT points (one thread per point).  The CTA walks the atom bins that can reach the tile's
// bounding box, keeps (ordered warp-ballot compaction) the atoms whose species cutoff sphere touches the box in a
// shared-memory list, and every thread sums its point's contributions from that list; the list is flushed in batches
// so any number of candidates works.  Each 4-node stencil of the radial tables is stored as its interpolating cubic
// (expanded on the host in extended precision about the stencil's second node), one 128-byte line per stencil.
#pragma once
#include "epilogue.cuh"

namespace c2 {

struct SpcDev {
  int off, n;         // first node in the packed tables, ngrid (0: not initialised)
  int z, pad;
  double a, b;        // r_i = a*exp(b*(i-1)) log grid parameters (src/grid1mod.f90:30-31)
  double rcut;        // min(cutrad(z), rmax) (:1246-1263); cutrad alone if the table is missing
  double cutrad;
  double rmax, r1;    // last and first node radius
  double ln2_b, lna_b;   // ln(2)/b and ln(a)/b: the pre-screen of the log-grid index (radial_interp); ln2_b = NaN
                         // when b < 1e-3 (pre-screen margins not valid: every pair takes the reference's own expression)
  double rmax_lo;        // rmax*(1 - 1e-14): below it r < rmax for certain
  double rfloor;         // max(r1, 1e-14): the clamp of :1302
};

struct PromolDev {
  int nenv, ncell, nspc;
  const double4 *atoms;     // nenv, sorted by bin: x, y, z, (species-1 | (cidx-1)<<8 ... packed as two ints)
  const int *bin_start;     // nbx*nby*nbz + 1
  double bin0[3], bin_inv;
  int nbin[3];
  const SpcDev *spc;
  // per stencil base the cubic through nodes base..base+3 of each table as sum_k c_k (r - xc)^k, xc = r(base+1):
  // xc, c0..c3 (f) [, d0..d3 (fp) [, e0..e3 (fpp)]] padded to 6 / 10 / 14 doubles; one table per derivative order
  const double *poly1, *poly4, *poly10;
  const unsigned char *fragmask;
  double rcutmax;
};

struct PromolParams {
  PromolDev m;
  GridDesc g;
  CellXf xf;
  OutDesc o;
  int nta, ntb, ntc;   // tile counts
  double nci_const;
  unsigned long long *count;   // optional: number of (point, atom) pairs inside the cutoffs
  int use_wc;                  // 1: evaluate at wc (the core call, src/fieldmod@proc.f90:828) instead of wcr (:803)
};

#ifdef __CUDACC__
__device__ __forceinline__ void unpack_atom(const double4 &a, int &is, int &cidx) {
  long long bits = __double_as_longlong(a.w);
  is = (int)(bits & 0xffffffffll);
  cidx = (int)(bits >> 32);
}

__device__ __forceinline__ float lg2_approx(float x) {   // MUFU.LG2 without the denormal fix-up (x is in [1,2))
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// grid1%interp (src/grid1mod@proc.f90:84-135) for r0 already clamped to >= r(1) by the caller (:1302) and < rmax (:98 is
// the caller's job).  r0 = r(1) needs no branch of its own: the index expression gives ir <= 1 there and the stencil base
// is 0 like in the reference's r0 <= r(1) branch (:103-105), with the same r.
template <int ND>
__device__ __forceinline__ void radial_interp(const PromolDev &m, const SpcDev &s, double r0, double &f, double &fp,
                                              double &fpp) {
  fp = 0.0; fpp = 0.0;
  int ir;
  const double r = r0;
  {                                                                    // ir = 1 + floor(log(r0/a)/b), :107
    // The FP64 log and the two divisions cost ~55 FP64 instructions per (point, atom) pair.  t = log(r0/a)/b is first
    // estimated as ((e + log2(m)) ln2 - ln a)/b with the mantissa's log2 from the SFU (absolute error <= 2^-22 on
    // [1,2] -> |dt| < 1.3e-4 for b = 0.002); unless t lies within 4e-4 of an integer its floor is already certain,
    // otherwise (0.08 % of the pairs) the reference's own expression decides.  The index is identical either way.
    const long long bits = __double_as_longlong(r0);
    const int e = (int)((bits >> 52) & 0x7ff) - 1023;
    const double mant = __longlong_as_double((bits & 0x000fffffffffffffll) | 0x3ff0000000000000ll);
    const double t = fma((double)e + (double)lg2_approx((float)mant), s.ln2_b, -s.lna_b);
    const int ti = __double2int_rd(t);
    const double fr = t - (double)ti;
    if (fr > 4e-4 && fr < 1.0 - 4e-4) ir = 1 + ti;                     // false for NaN (b < 1e-3)
    else ir = 1 + (int)floor(__ddiv_rn(log(__ddiv_rn(r0, s.a)), s.b));
  }
  const int base = min(max(ir, 2), s.n - 2) - 2;                       // :110, 0-based first node of the stencil
  // The 4-node Lagrange interpolant (:112-133) is one cubic per stencil: the host expands it once, in extended precision,
  // about the stencil's second node (PromolDev::poly*), so a pair costs 40 / 72 / 104 bytes of table reads (ND = 1 / 4 / 10)
  // and three FMAs per interpolated table instead of 192 bytes (four nodes + six reciprocal differences) and 24 multiplies.
  // The kernel is bound by the L1 data stage of these divergent gathers (ncu: l1tex wavefronts 91-93 % of peak, FP64 pipe
  // 20-26 %): one table per ND with record strides of 48 / 80 / 112 bytes, so that the 16-byte pieces fetched by the lanes
  // of a request fall on all eight 16-byte bank groups of a line instead of one (a 128-byte stride serialises them).
  constexpr int STR = ND == 1 ? 6 : (ND == 4 ? 10 : 14);
  const double *tab = ND == 1 ? m.poly1 : (ND == 4 ? m.poly4 : m.poly10);
  const double2 *pc = reinterpret_cast<const double2 *>(tab + STR * (size_t)(s.off + base));
  const double2 q0 = pc[0], q1 = pc[1], q2 = pc[2];                    // xc c0 | c1 c2 | c3 d0
  const double t = r - q0.x;
  f = fma(t, fma(t, fma(t, q2.x, q1.y), q1.x), q0.y);
  if (ND > 1) {
    const double2 q3 = pc[3], q4 = pc[4];                              // d1 d2 | d3 e0
    fp = fma(t, fma(t, fma(t, q4.x, q3.y), q3.x), q2.y);
    if (ND > 4) {
      const double2 q5 = pc[5], q6 = pc[6];                            // e1 e2 | e3 -
      fpp = fma(t, fma(t, fma(t, q6.x, q5.y), q5.x), q4.y);
    }
  }
}

constexpr int PROMOL_PT = 256;    // points per tile = threads per CTA
constexpr int PROMOL_CAP = 768;   // shared-memory atom list capacity (>= 2*PROMOL_PT)
constexpr int PROMOL_NSPC = 16;   // species whose constants are staged in shared memory (SPCG = false)

// SPCG: more than PROMOL_NSPC species -> their constants are read from global memory (same code otherwise)
template <int ND, bool SPCG>
__global__ void __launch_bounds__(PROMOL_PT, ND == 1 ? 4 : 3) promol_grid_kernel(const __grid_constant__ PromolParams p) {
  constexpr int NT = PROMOL_PT, NW = NT / 32;
  __shared__ double4 s_atom[PROMOL_CAP];
  __shared__ double s_rc2[PROMOL_CAP];       // squared cutoff of the candidate's species, upper edge of the rounding band
  __shared__ SpcDev s_spc[SPCG ? 1 : PROMOL_NSPC];   // species constants
  __shared__ int s_cnt[NW], s_rowoff[NT], s_rowi0[NT];
  __shared__ double s_red[NW * 6];
  __shared__ double s_bb[6];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const PromolDev &m = p.m;
  const GridDesc &g = p.g;
  const int na = grid_na(g), nb = grid_nb(g);
  const int t = blockIdx.x;
  int ia, ib, ic;
  part_item(g, t, p.nta, p.ntb, ia, ib, ic);
  const int ta0 = ia * g.ta, tb0 = ib * g.tb, tc0 = g.c0 + ic * g.tc;
  // Lattice tiles (8x8x4): every warp takes a compact 4x4x2 brick so that its 32 points see an atom at nearly the same
  // distance (lanes share table records: fewer distinct lines per gather, the L1 data stage is the busiest unit);
  // 4 consecutive points along a = one 32-byte sector per store.  Point lists (ta = 256) stay linear.
  int a, b, c;
  if (g.pts) { a = ta0 + tid; b = tb0; c = tc0; }
  else {
    a = ta0 + (lane & 3) + 4 * (warp & 1);
    b = tb0 + ((lane >> 2) & 3) + 4 * ((warp >> 1) & 1);
    c = tc0 + (lane >> 4) + 2 * (warp >> 2);
  }
  const bool valid = a < na && b < nb && c < g.c1;
  if (!tile_owned(g, tb0, tc0)) return;      // core pre-pass: another part's unit (block-uniform)

  if (!SPCG && tid < m.nspc) s_spc[tid] = m.spc[tid];   // visible after the barriers of the bounding-box reduction below
  auto species = [&](int is) -> const SpcDev & { return SPCG ? m.spc[is] : s_spc[is]; };
  double xc[3] = {0, 0, 0};
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  if (valid) {
    double x[3], wc[3];
    grid_point(g, a, b, c, x);
    grd_wrap(p.xf, x, g.periodic, wc, xc);   // xc = wcr (:802)
    if (p.use_wc) { xc[0] = wc[0]; xc[1] = wc[1]; xc[2] = wc[2]; }
    center_cart(p.xf, xc, nullptr);          // :1242-1243
#pragma unroll
    for (int d = 0; d < 3; d++) { lo[d] = xc[d]; hi[d] = xc[d]; }
  }
  {
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
    if (tid < 6) {
      double v = s_red[tid];
      for (int w = 1; w < NW; w++) v = (tid < 3) ? fmin(v, s_red[w * 6 + tid]) : fmax(v, s_red[w * 6 + tid]);
      s_bb[tid] = v;
    }
    __syncthreads();
  }
  if (s_bb[0] > s_bb[3]) return;   // no valid point in this tile (block-uniform)

  // bins that can reach the box
  int blo[3], bhi[3];
#pragma unroll
  for (int d = 0; d < 3; d++) {
    blo[d] = max(0, (int)floor((s_bb[d] - m.rcutmax - m.bin0[d]) * m.bin_inv));
    bhi[d] = min(m.nbin[d] - 1, (int)floor((s_bb[3 + d] + m.rcutmax - m.bin0[d]) * m.bin_inv));
  }

  double f = 0.0, gx = 0.0, gy = 0.0, gz = 0.0, hs[6] = {0, 0, 0, 0, 0, 0};
  bool isnuc = false;
  unsigned long long npair = 0;
  int nlist = 0;

  auto flush = [&]() {
    __syncthreads();
    if (valid) {
      for (int i = 0; i < nlist; i++) {
        const double4 at = s_atom[i];
        // the reference takes at%r - x0 for the distance (:988) and xc - at%r for the gradient (:1308)
        const double ex = at.x - xc[0], ey = at.y - xc[1], ez = at.z - xc[2];
        const double d2 = ex * ex + ey * ey + ez * ez;
        isnuc = isnuc || d2 < 1e-10;                                 // src/environmod@proc.f90:677
        const double rc2hi = s_rc2[i];                               // rcut^2 (1 + 1e-14)
        if (d2 > rc2hi) continue;                                    // :998, outside for certain
        int is, cidx;
        unpack_atom(at, is, cidx);
        const SpcDev &s = species(is);
        if (s.n == 0) continue;                                      // :1298-1299
        if (m.fragmask && !m.fragmask[cidx]) continue;               // :1287-1290
        // dist (:988, gfortran norm2) = sqrt(d2) to ~2 ulp, which moves rho by ~1e-13 relative.  Only inside the rounding
        // band of the cutoff (:998) and at the table end, where interp switches to exactly 0 (:98), is the reference's own
        // scaled arithmetic replayed.  The test r > cutrad (:1295) cannot fire: r <= rcut = min(cutrad, rmax).
        double r = __dsqrt_rn(d2);
        if (d2 >= rc2hi * (1.0 - 2.5e-14) || r > s.rmax_lo) {        // rare: decide with norm2 like the reference
          r = norm2_ref(ex, ey, ez);
          if (r > s.rcut) continue;                                  // :998
          if (r >= s.rmax) { npair++; continue; }                    // :98: inside the cutoff, contributes exactly 0
        }
        npair++;
        r = r < s.rfloor ? s.rfloor : r;                             // :1302
        double rho, rhop, rhopp;
        radial_interp<ND>(m, s, r, rho, rhop, rhopp);
        rho = rho < 0.0 ? 0.0 : rho;                                 // :1304
        f += rho;
        if (ND > 1) {
          const double xx = xc[0] - at.x, yy = xc[1] - at.y, zz = xc[2] - at.z;
          const double rinv1 = 1.0 / r;
          const double rinv1rp = rinv1 * rhop;
          gx += xx * rinv1rp; gy += yy * rinv1rp; gz += zz * rinv1rp;   // :1309-1311
          if (ND > 4) {
            const double rfac = (rhopp - rinv1rp) * rinv1 * rinv1;      // :1314
            hs[0] += rinv1rp + rfac * xx * xx;
            hs[1] += rinv1rp + rfac * yy * yy;
            hs[2] += rinv1rp + rfac * zz * zz;
            hs[3] += rfac * xx * yy;
            hs[4] += rfac * xx * zz;
            hs[5] += rfac * yy * zz;
          }
        }
      }
    }
    __syncthreads();
    nlist = 0;
  };

  // Candidate scan.  The atoms of a (bz, by) bin row with bx in [blo, bhi] are contiguous in `atoms`; the rows are
  // flattened into one virtual index space (prefix sums of the row lengths, NT rows at a time) so the block tests NT
  // candidates per pass instead of one short row per pass (a row holds ~10-30 atoms: two block barriers each).
  const int nby = bhi[1] - blo[1] + 1, nrow = nby * (bhi[2] - blo[2] + 1);
  for (int r0 = 0; r0 < nrow; r0 += NT) {
    const int r = r0 + tid;
    int i0 = 0, len = 0;
    if (r < nrow) {
      const int rowbase = ((blo[2] + r / nby) * m.nbin[1] + blo[1] + r % nby) * m.nbin[0];
      i0 = m.bin_start[rowbase + blo[0]];
      len = m.bin_start[rowbase + bhi[0] + 1] - i0;
    }
    int incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(~0u, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_cnt[warp] = incl;
    __syncthreads();
    int woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) {
      if (w < warp) woff += s_cnt[w];
      total += s_cnt[w];
    }
    s_rowoff[tid] = woff + incl - len;      // exclusive prefix: first virtual index of this row
    s_rowi0[tid] = i0;
    __syncthreads();
    const int nr = min(NT, nrow - r0);
    for (int base = 0; base < total; base += NT) {
      const int v = base + tid;
      bool act = false;
      double4 at;
      double rc2 = 0.0;
      if (v < total) {
        int lo_r = 0, hi_r = nr - 1;        // last row with rowoff <= v
        while (lo_r < hi_r) {
          const int mid = (lo_r + hi_r + 1) >> 1;
          if (s_rowoff[mid] <= v) lo_r = mid; else hi_r = mid - 1;
        }
        at = m.atoms[s_rowi0[lo_r] + (v - s_rowoff[lo_r])];
        int is, cidx;
        unpack_atom(at, is, cidx);
        rc2 = species(is).rcut;
        rc2 *= rc2;
        act = box_d2(s_bb, at.x, at.y, at.z) <= rc2 * (1.0 + 1e-12) + 1e-9;
        rc2 *= 1.0 + 1e-14;                 // upper edge of the rounding band of the per-point test
      }
      const unsigned mk = __ballot_sync(~0u, act);
      __syncthreads();                      // s_cnt of the previous pass / the prefix sums has been read by everyone
      if (lane == 0) s_cnt[warp] = __popc(mk);
      __syncthreads();
      int off = nlist, tot = 0;
#pragma unroll
      for (int w = 0; w < NW; w++) {
        if (w < warp) off += s_cnt[w];
        tot += s_cnt[w];
      }
      if (act) {
        const int slot = off + __popc(mk & ((1u << lane) - 1u));
        s_atom[slot] = at;
        s_rc2[slot] = rc2;
      }
      nlist += tot;
      if (nlist > PROMOL_CAP - NT) flush();
    }
    __syncthreads();
  }
  flush();
  if (p.count) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) npair += __shfl_xor_sync(~0u, npair, o);
    if (lane == 0 && npair) atomicAdd(p.count, npair);
  }
  if (valid) emit_point(p.o, out_index(g, a, b, c), f, gx, gy, gz, hs[0], hs[1], hs[2], hs[3], hs[4], hs[5], isnuc, p.nci_const);
}
#endif

}  // namespace c2
