// wfn_device.cuh -- device data of a wavefunction field and the primitive (GTO) evaluation.
// Restates calculate_mo_gto's per-primitive arithmetic (src/wfn_private@proc.F90:2713-2768) for the GPU.
#pragma once
#include "common.cuh"

namespace c2 {

// Device image of type molwfn (src/wfn_private.f90:83-131) after sorting the primitives by centre (icord order,
// :3112-3130) and grouping runs with equal centre+exponent into shells (they share exp(-a r^2) and dran).
struct WfnDev {
  int natom, npri, nshell, nmo, ldc;   // ldc = nblk*NB: padded MO count (row stride of cmoT)
  int nedf;
  const double *atom_xyz;              // 3*natom
  const double4 *shell_xa;             // nshell: centre x,y,z and exponent
  const double2 *shell_cut;            // nshell: dran, dran^2
  const int *shell_pbeg;               // nshell+1: first sorted primitive of each shell
  const int *prim_shell;               // npri: shell of each sorted primitive
  const unsigned char *prim_l;         // 4*npri: lx,ly,lz,0 of each sorted primitive (table :2695-2705)
  const double *cmoT;                  // npri x ldc row-major: cmoT[p][imo] = cmo(imo, icord(p)), zero padded
  const double *occ;                   // ldc, zero padded
  // EDF core functions (src/wfn_private@proc.F90:2783-2873), unsorted
  const double4 *edf_xa;               // nedf: centre + exponent
  const double4 *edf_cc;               // nedf: coefficient, dran, dran^2, 0
  const unsigned char *edf_l;          // 4*nedf
};

#ifdef __CUDACC__
// x^l for l = 0..5 (and the neighbours needed by the derivatives)
__device__ __forceinline__ double ipow_small(double x, int l) {
  double x2 = x * x;
  switch (l) {
    case 0: return 1.0;
    case 1: return x;
    case 2: return x2;
    case 3: return x2 * x;
    case 4: return x2 * x2;
    case 5: return x2 * x2 * x;
    case 6: return x2 * x2 * x2;
    default: return x2 * x2 * x2 * x;
  }
}

// One Cartesian axis of a primitive: f0 = x^l, f1 = d/dx, f2 = d2/dx2 of x^l exp(-a x^2) without the exponential
// (:2755-2766): f1 = l x^(l-1) - 2a x^(l+1); f2 = l(l-1) x^(l-2) - 2a(2l+1) x^l + 4a^2 x^(l+2).
template <int ND>
__device__ __forceinline__ void axis_terms(double x, int l, double al, double &f0, double &f1, double &f2) {
  f0 = ipow_small(x, l);
  if (ND > 1) {
    double xm1 = (l > 0) ? ipow_small(x, l - 1) : 0.0;
    double xp1 = f0 * x;
    double ta = 2.0 * al;
    f1 = (double)l * xm1 - ta * xp1;
    if (ND > 4) {
      double xm2 = (l > 1) ? ipow_small(x, l - 2) : 0.0;
      f2 = (double)(l * (l - 1)) * xm2 - ta * (double)(2 * l + 1) * f0 + ta * ta * (xp1 * x);
    }
  }
}

// chi(1:ND) of one primitive at displacement (x,y,z) from its centre, ex = exp(-al*d2):
// value; d/dx,d/dy,d/dz; dxx,dyy,dzz,dxy,dxz,dyz (:2755-2768).
template <int ND>
__device__ __forceinline__ void prim_chi(double x, double y, double z, int lx, int ly, int lz, double al, double ex,
                                         double *chi) {
  double x0, x1, x2, y0, y1, y2, z0, z1, z2;
  axis_terms<ND>(x, lx, al, x0, x1, x2);
  axis_terms<ND>(y, ly, al, y0, y1, y2);
  axis_terms<ND>(z, lz, al, z0, z1, z2);
  double xe = x0 * ex, ye = y0 * ex;
  chi[0] = xe * y0 * z0;
  if (ND > 1) {
    double yz = y0 * z0;
    chi[1] = x1 * yz * ex;
    chi[2] = y1 * xe * z0;
    chi[3] = z1 * xe * y0;
    if (ND > 4) {
      chi[4] = x2 * yz * ex;
      chi[5] = y2 * z0 * xe;
      chi[6] = z2 * xe * y0;
      chi[7] = x1 * y1 * z0 * ex;
      chi[8] = x1 * z1 * ye;
      chi[9] = z1 * y1 * xe;
    }
  }
}

// ---- chi of the primitives of ONE SHELL at one point -----------------------------------------------------------------
// The generator warps of wfn_ws_kernel.cuh share the FP64 unit with the consumers' DMMAs and lose almost every
// arbitration (tools/fp64_share.cu: 2500-9700 cycles per DFMA next to two DMMA-streaming warps), and the warp-autonomous
// kernel's FP64 ALU work shares the unit with its own DMMAs: what counts is the NUMBER of FP64 instructions a warp issues.
// Everything common to the primitives of a shell -- displacement, squares, 2a, 2a x, (2a x)^2 (and, in the generators,
// screen and exp) -- is computed once per (point, shell); what depends on (lx, ly, lz) is picked with selects (integer
// pipe), which also keeps lanes with different angular momenta on one instruction stream.  Same formulas as axis_terms /
// prim_chi above (src/wfn_private@proc.F90:2755-2768): f0 = x^l, f1 = l x^(l-1) - 2a x^(l+1),
// f2 = l(l-1) x^(l-2) - 2a(2l+1) x^l + (2a x)^2 x^l; 29 FP64 instructions per primitive for l <= 2 (a branch-free prim_chi: 65).
struct ShellAxis {
  double x, x2, twox, tx, tx2;   // displacement, its square, 2x, 2a x, (2a x)^2
};
static __constant__ double c_sh_lf[8] = {0, 1, 2, 3, 4, 5, 6, 7};          // l
static __constant__ double c_sh_ll1[8] = {0, 0, 2, 6, 12, 20, 30, 42};     // l (l - 1)
static __constant__ double c_sh_2l1[8] = {1, 3, 5, 7, 9, 11, 13, 15};      // 2 l + 1

// BIG = false: l <= 2 (s, p, d shells): no multiplication for the powers.  ta = 2a.
template <int ND, bool BIG>
__device__ __forceinline__ void shell_axis(const ShellAxis &a, int l, double ta, double &f0, double &f1, double &f2) {
  if (!BIG) {
    f0 = l == 2 ? a.x2 : (l == 1 ? a.x : 1.0);
    if (ND > 1) {
      const double lpm1 = l == 2 ? a.twox : (l == 1 ? 1.0 : 0.0);   // l x^(l-1)
      f1 = fma(-a.tx, f0, lpm1);
      if (ND > 4) f2 = fma(f0, fma(l == 2 ? -5.0 : (l == 1 ? -3.0 : -1.0), ta, a.tx2), l == 2 ? 2.0 : 0.0);
    }
  } else {
    const double x4 = a.x2 * a.x2;
    const int n = l - 2;
    const double pm2 = l >= 2 ? ((n & 1) ? a.x : 1.0) * ((n & 2) ? a.x2 : 1.0) * ((n & 4) ? x4 : 1.0) : 0.0;
    const double pm1 = l >= 2 ? pm2 * a.x : (l == 1 ? 1.0 : 0.0);
    f0 = l >= 1 ? pm1 * a.x : 1.0;
    if (ND > 1) {
      f1 = fma(-a.tx, f0, c_sh_lf[l & 7] * pm1);
      if (ND > 4) f2 = fma(f0, fma(-c_sh_2l1[l & 7], ta, a.tx2), c_sh_ll1[l & 7] * pm2);
    }
  }
}

template <int ND, bool BIG>
__device__ __forceinline__ void shell_prim_chi(const ShellAxis &ax, const ShellAxis &ay, const ShellAxis &az, int lx, int ly,
                                                  int lz, double ta, double ex, int mode, double *chi) {
  double x0, x1 = 0, x2 = 0, y0, y1 = 0, y2 = 0, z0, z1 = 0, z2 = 0;
  if (ND == 1 && (mode & 3)) {   // value-only launches: the k-th Cartesian derivative as the "value" (WfnParams::chi_mode)
    shell_axis<4, BIG>(ax, lx, ta, x0, x1, x2);
    shell_axis<4, BIG>(ay, ly, ta, y0, y1, y2);
    shell_axis<4, BIG>(az, lz, ta, z0, z1, z2);
    const int k = mode & 3;
    chi[0] = (k == 1 ? x1 * y0 * z0 : (k == 2 ? x0 * y1 * z0 : x0 * y0 * z1)) * ex;
    return;
  }
  shell_axis<ND, BIG>(ax, lx, ta, x0, x1, x2);
  shell_axis<ND, BIG>(ay, ly, ta, y0, y1, y2);
  shell_axis<ND, BIG>(az, lz, ta, z0, z1, z2);
  const double xe = x0 * ex, yz = y0 * z0;
  chi[0] = xe * yz;
  if (ND > 1) {
    const double yze = yz * ex, xze = xe * z0, xye = xe * y0;
    chi[ND > 1 ? 1 : 0] = x1 * yze;
    chi[ND > 1 ? 2 : 0] = y1 * xze;
    chi[ND > 1 ? 3 : 0] = z1 * xye;
    if (ND > 4) {
      const double ye = y0 * ex, ze = z0 * ex;
      chi[ND > 4 ? 4 : 0] = x2 * yze;
      chi[ND > 4 ? 5 : 0] = y2 * xze;
      chi[ND > 4 ? 6 : 0] = z2 * xye;
      chi[ND > 4 ? 7 : 0] = (x1 * y1) * ze;
      chi[ND > 4 ? 8 : 0] = (x1 * z1) * ye;
      chi[ND > 4 ? 9 : 0] = (z1 * y1) * xe;
    }
  }
}

// prim_chi with WfnParams::chi_mode: in value-only launches bits 0-1 select a Cartesian derivative as the "value"
template <int ND>
__device__ __forceinline__ void prim_chi_mode(double x, double y, double z, int lx, int ly, int lz, double al, double ex,
                                              int mode, double *chi) {
  if (ND == 1 && (mode & 3)) {
    double c[4];
    prim_chi<4>(x, y, z, lx, ly, lz, al, ex, c);
    const int k = mode & 3;
    chi[0] = k == 1 ? c[1] : (k == 2 ? c[2] : c[3]);
  } else
    prim_chi<ND>(x, y, z, lx, ly, lz, al, ex, chi);
}

// ---- PTX wrappers: mbarrier + 1-D bulk TMA copy + FP64 tensor-core MMA -------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t a = smem_u32(bar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(a), "r"(parity)
      : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// generic-proxy GLOBAL writes (the chi-tile cache) ordered against async-proxy bulk loads of the same addresses
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// D(8x8) += A(8x4, row) * B(4x8, col), FP64 tensor core (SASS: DMMA)
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
#endif

}  // namespace c2
