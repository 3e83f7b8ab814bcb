// common.cuh -- shared device/host types of the critic2 grid field-evaluation module (sm_100a only).
// Citations are relative to /root/reference/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace c2 {

// ---------------------------------------------------------------------------------------------------------------
// Geometry carried to the device
// ---------------------------------------------------------------------------------------------------------------
// The matrices field%grd uses to wrap a point (src/fieldmod@proc.f90:698-721) and y2z_center uses to fold it into
// the main reduced cell (src/environmod@proc.f90:426-444).  Column-major like the reference.
struct CellXf {
  double c2x[9], x2c[9], c2xr[9], xr2c[9];
  int ismolecule;
};

// A lattice of points.  (i0,i1,i2) are the reference's loop indices in the textual order of its coordinate formula
// x = x0 + i0*v0 + i1*v1 + i2*v2 (src/rhoplot@proc.f90:591, src/nci@proc.f90:504).  The OUTPUT array's fastest
// axis is i0 for CUBE (lf(ix,iy,iz)) and i2 for NCIPLOT (crho(k,j,i)); (a,b,c) below are (fast,mid,slow) output axes.
struct GridDesc {
  double x0[3];
  double v[3][3];
  int n[3];        // extents of (i0,i1,i2)
  int nci_order;   // 0: a=i0,b=i1,c=i2   1: a=i2,b=i1,c=i0
  int c0, c1;      // evaluated range of the slowest output axis
  int c_part, c_nparts;  // of the layers of height sc (tc for the promolecular kernel) inside [c0,c1), evaluate those with index = c_part mod c_nparts
  int part_rows;   // 1: the interleaved unit is a b-row of (super-)tiles of a layer, unit = layer * ntb + ib (finer balance)
  int periodic;
  const double *pts;  // non-null: explicit point list (n[0] points, n[1]=n[2]=1), 3 doubles per point
  int ta, tb, tc;  // sub-tile shape (ta*tb*tc = points per tile)
  int sa, sb, sc;  // super-tile shape (multiples of the sub-tile shape)
  // Ownership filter of the core pre-pass (see OutDesc::core): this launch covers every tile of [c0,c1) but a tile is
  // evaluated only if the unit of ANOTHER decomposition (layers of height own_sc / rows of own_sb lines, own_ntb rows per
  // layer, dealt round-robin like c_part/c_nparts/part_rows) that contains it belongs to own_part.  own_nparts <= 1: off.
  int own_sb, own_sc, own_ntb, own_part, own_nparts, own_rows;
};

// true if the tile whose first point has output coordinates (b0, c0t) is evaluated under the ownership filter
__host__ __device__ inline bool tile_owned(const GridDesc &g, int b0, int c0t) {
  if (g.own_nparts <= 1) return true;
  const int layer = (c0t - g.c0) / g.own_sc;
  const int unit = g.own_rows ? layer * g.own_ntb + b0 / g.own_sb : layer;
  return unit % g.own_nparts == g.own_part;
}

// Work item `st` of this part -> (super-)tile indices along a, b and the layer index along c (nta, ntb tiles along a, b).
__host__ __device__ inline void part_item(const GridDesc &g, int st, int nta, int ntb, int &ia, int &ib, int &ic) {
  ia = st % nta;
  const int r = st / nta;
  if (g.part_rows) { const int u = g.c_part + r * g.c_nparts; ib = u % ntb; ic = u / ntb; }
  else { ib = r % ntb; ic = g.c_part + (r / ntb) * g.c_nparts; }
}
// Number of work items of this part for nlayers layers.
__host__ __device__ inline long part_items(const GridDesc &g, int nta, int ntb, int nlayers) {
  const long units = g.part_rows ? (long)nlayers * ntb : nlayers;
  const long mine = g.c_part < units ? (units - g.c_part + g.c_nparts - 1) / g.c_nparts : 0;
  return g.part_rows ? mine * nta : mine * nta * ntb;
}

__host__ __device__ inline int grid_na(const GridDesc &g) { return g.nci_order ? g.n[2] : g.n[0]; }
__host__ __device__ inline int grid_nb(const GridDesc &g) { return g.n[1]; }
__host__ __device__ inline int grid_nc(const GridDesc &g) { return g.nci_order ? g.n[0] : g.n[2]; }

// Output selection.  mode 0: CUBE properties (one array per set bit of outmask, src/rhoplot@proc.f90:596-621);
// mode 1: NCIPLOT crho/cgrad (src/nci@proc.f90:523-525,553-554); mode 2: point list, C2GPU_RES_LEN doubles per point;
// mode 3: value only into prop[0] (grd0, src/fieldmod@proc.f90:848-928);
// mode 4: the raw sums f, gf(3), hs(6) before the tail of grd, component k of point idx at res[k*core_stride + idx]
//         (k < core_nd) -- the core pre-pass below writes this;
// mode 5: like 4 (all ten components) but AFTER the core has been added and the gradient nullified on a nucleus: the first
//         pass of the two-pass NCIPLOT evaluation of the block-synchronous wavefunction kernels, whose eigenvalue / RDG
//         epilogue otherwise runs on P threads while the rest of the CTA waits (nci_epilogue_kernel is the second pass).
// core != null: f%usecore (src/fieldmod@proc.f90:827-832, :922-925): the core promolecular density and derivatives of
// the same lattice, written by a mode-4 pre-pass of the promolecular kernel over the core tables cgrid(iz,zpsp(iz)),
// are added to f, gf, hf BEFORE the nucleus test / gfmod / del2f and the NCIPLOT formulas.
struct OutDesc {
  int mode;
  unsigned outmask;
  double *prop[12];
  double *crho, *cgrad;
  double *res;
  const double *core;
  size_t core_stride;
  int core_nd;
};

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------------------------
// Arithmetic that must round like the reference's (no FMA contraction: gfortran on x86-64 emits mul + add)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

// y = M x, column-major, left-to-right like Fortran matmul on a 3x3 (oracle: orc_matvec)
__device__ __forceinline__ void matvec3(const double *m, const double *x, double *y) {
#pragma unroll
  for (int i = 0; i < 3; i++) y[i] = add_rn(add_rn(mul_rn(m[i], x[0]), mul_rn(m[i + 3], x[1])), mul_rn(m[i + 6], x[2]));
}

// gfortran's NORM2 (libgfortran norm2_r8): scaled sum of squares.  Used by the reference for every atom distance
// (src/environmod@proc.f90:988) and for gfmod (src/fieldmod@proc.f90:839).
__device__ __forceinline__ double norm2_ref(double x, double y, double z) {
  double scale = 1.0, result = 0.0;
  double v[3] = {x, y, z};
#pragma unroll
  for (int i = 0; i < 3; i++) {
    if (v[i] != 0.0) {
      double absx = fabs(v[i]);
      if (scale < absx) {
        double val = __ddiv_rn(scale, absx);
        result = add_rn(1.0, mul_rn(result, mul_rn(val, val)));
        scale = absx;
      } else {
        double val = __ddiv_rn(absx, scale);
        result = add_rn(result, mul_rn(val, val));
      }
    }
  }
  return mul_rn(scale, __dsqrt_rn(result));
}

// Reference screen `dist > cutoff -> skip` (src/wfn_private@proc.F90:2747, src/environmod@proc.f90:998-1001) decided
// from the plain squared distance, falling back to the reference's own norm2 arithmetic inside the rounding band.
__device__ __forceinline__ bool within_cut(double dx, double dy, double dz, double d2, double cut, double cut2) {
  if (d2 < cut2 * (1.0 - 1e-14)) return true;
  if (d2 > cut2 * (1.0 + 1e-14)) return false;
  return !(norm2_ref(dx, dy, dz) > cut);
}

// Lattice point of output coordinates (a,b,c): x = x0 + i0*v0 + i1*v1 + i2*v2, each operation rounded.
__device__ __forceinline__ void grid_point(const GridDesc &g, int a, int b, int c, double *x) {
  if (g.pts) {
    const double *p = g.pts + 3 * (size_t)a;
    x[0] = p[0]; x[1] = p[1]; x[2] = p[2];
    return;
  }
  int i0 = g.nci_order ? c : a, i1 = b, i2 = g.nci_order ? a : c;
#pragma unroll
  for (int d = 0; d < 3; d++)
    x[d] = add_rn(add_rn(add_rn(g.x0[d], mul_rn((double)i0, g.v[0][d])), mul_rn((double)i1, g.v[1][d])),
                  mul_rn((double)i2, g.v[2][d]));
}

// field%grd's wrap (src/fieldmod@proc.f90:698-721): wc via the cell matrices, wcr via the reduced-cell matrices.
__device__ __forceinline__ void grd_wrap(const CellXf &xf, const double *v, int periodic, double *wc, double *wcr) {
  const double flooreps = 1e-4;  // src/fieldmod@proc.f90:30
  double wx[3], wxr[3];
  matvec3(xf.c2x, v, wx);
  matvec3(xf.c2xr, v, wxr);
  if (periodic) {
#pragma unroll
    for (int i = 0; i < 3; i++) {
      if (wx[i] < -flooreps || wx[i] > 1.0 + flooreps) wx[i] = wx[i] - floor(wx[i]);
      if (wxr[i] < -flooreps || wxr[i] > 1.0 + flooreps) wxr[i] = wxr[i] - floor(wxr[i]);
    }
  }
  matvec3(xf.x2c, wx, wc);
  matvec3(xf.xr2c, wxr, wcr);
}

// y2z_center, Cartesian -> main reduced cell -> Cartesian (src/environmod@proc.f90:426-444); lvec = floor part.
__device__ __forceinline__ void center_cart(const CellXf &xf, double *xx, int *lvec) {
  if (xf.ismolecule) {
    if (lvec) lvec[0] = lvec[1] = lvec[2] = 0;
    return;
  }
  double xr[3];
  matvec3(xf.c2xr, xx, xr);
#pragma unroll
  for (int d = 0; d < 3; d++) {
    double fl = floor(xr[d]);
    if (lvec) lvec[d] = (int)fl;
    xr[d] = xr[d] - fl;
  }
  matvec3(xf.xr2c, xr, xx);
}

__device__ __forceinline__ double box_d2(const double *bb, double x, double y, double z) {
  double dx = fmax(fmax(bb[0] - x, 0.0), x - bb[3]);
  double dy = fmax(fmax(bb[1] - y, 0.0), y - bb[4]);
  double dz = fmax(fmax(bb[2] - z, 0.0), z - bb[5]);
  return dx * dx + dy * dy + dz * dz;
}

__device__ __forceinline__ size_t out_index(const GridDesc &g, int a, int b, int c) {
  return (size_t)a + (size_t)grid_na(g) * ((size_t)b + (size_t)grid_nb(g) * (size_t)c);
}
#endif  // __CUDACC__

}  // namespace c2
