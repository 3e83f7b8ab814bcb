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

// exp(t) for t <= 0 with 12 FP64 instructions and a dependency chain of ~9 (CUDA's exp: ~21 in one chain): the generator
// warps pay for every FP64 instruction many times over (see below).  t = k ln2/32 + r, |r| <= ln2/64; exp(t) = 2^(k>>5)
// T[k&31] (1 + r q(r)) with the degree-6 Taylor polynomial (remainder 3.5e-18) and T = 2^(j/32) as hi + lo.  Maximum error
// 0.56 ulp against a long-double reference over 2e7 arguments in [-700, 0] (CPU emulation with fma()); below -700 (never
// inside a primitive's cutoff radius) the library function takes over.
static __device__ const double2 g_exp_tab[32] = {
    {1.0, 0.0}, {1.0218971486541166, 5.109225028973444e-17},
    {1.0442737824274138, 8.551889705537965e-17}, {1.0671404006768237, -7.899853966841582e-17},
    {1.0905077326652577, -3.046782079812471e-17}, {1.1143867425958924, 1.0410278456845571e-16},
    {1.1387886347566916, 8.912812676025408e-17}, {1.1637248587775775, 3.8292048369240935e-17},
    {1.189207115002721, 3.982015231465646e-17}, {1.215247359980469, -7.712630692681488e-17},
    {1.241857812073484, 4.658027591836937e-17}, {1.2690509571917332, 2.667932131342186e-18},
    {1.2968395546510096, 2.5382502794888315e-17}, {1.3252366431597413, -2.8587312100388614e-17},
    {1.3542555469368927, 7.70094837980299e-17}, {1.383909881963832, -6.770511658794786e-17},
    {1.4142135623730951, -9.667293313452913e-17}, {1.4451808069770467, -3.0237581349939873e-17},
    {1.4768261459394993, -3.483994556892796e-17}, {1.5091644275934228, -1.016455327754295e-16},
    {1.5422108254079407, 7.949834809697621e-17}, {1.5759808451078865, -1.0136916471278304e-17},
    {1.6104903319492543, 2.4707192569797888e-17}, {1.645755478153965, -1.0125679913674773e-16},
    {1.681792830507429, 8.199010020581497e-17}, {1.718619298122478, -1.851380418263111e-17},
    {1.7562521603732995, 2.960140695448873e-17}, {1.7947090750031072, 1.8227458427912087e-17},
    {1.8340080864093424, 3.283107224245627e-17}, {1.8741676341103, -6.122763413004143e-17},
    {1.9152065613971474, -1.0619946056195963e-16}, {1.9571441241754002, 8.960767791036668e-17},
};
__device__ __forceinline__ double exp_neg(double t) {
  if (t < -700.0) return exp(t);
  const double M = 6755399441055744.0;                       // 1.5 * 2^52: the low word of t*32/ln2 + M is round(t*32/ln2)
  const double kf = fma(t, 46.16624130844683, M);
  const int k = __double2loint(kf);
  const double kd = kf - M;
  double r = fma(kd, -0.02166084938653512, t);               // ln2/32: 32 significant bits, exact product
  r = fma(kd, -5.9631716539705866e-12, r);
  double q = fma(r, 1.0 / 720, 1.0 / 120);
  q = fma(q, r, 1.0 / 24);
  q = fma(q, r, 1.0 / 6);
  q = fma(q, r, 0.5);
  q = fma(q, r, 1.0);
  q = q * r;                                                 // exp(r) - 1
  const double2 tb = __ldg(&g_exp_tab[k & 31]);
  const double v = fma(tb.x, q, tb.y) + tb.x;
  return __hiloint2double(__double2hiint(v) + ((k >> 5) << 20), __double2loint(v));
}

// ---- chi of the primitives of ONE SHELL at one point -----------------------------------------------------------------
// The generator warps of wfn_ws_kernel.cuh share the FP64 unit with the consumers' DMMAs and lose almost every
// arbitration (tools/fp64_share.cu: 2500-9700 cycles per DFMA next to two DMMA-streaming warps), and the warp-autonomous
// kernel's FP64 ALU work shares the unit with its own DMMAs: what counts is the NUMBER of FP64 instructions a warp issues.
// Everything common to the primitives of a shell -- displacement, squares, 2a, 2a x, (2a x)^2, screen and exp -- is
// computed once per (point, shell); what depends on (lx, ly, lz) is picked with selects (integer pipe), which also keeps
// lanes with different angular momenta on one instruction stream; the z axis carries the exponential (its base values are
// multiplied by e = exp(-a r^2) once per shell), so no product of a primitive needs it.  Same formulas as axis_terms /
// prim_chi above (src/wfn_private@proc.F90:2755-2768): f0 = x^l, f1 = l x^(l-1) - 2a x^(l+1),
// f2 = l(l-1) x^(l-2) - 2a(2l+1) x^l + (2a x)^2 x^l; 25 FP64 instructions per primitive for l <= 2 (a branch-free
// prim_chi: 65).
struct ShellAxis {
  double x, x2, twox, tx, tx2;   // displacement, its square, 2x, 2a x, (2a x)^2
};
struct ShellAxisE {              // the axis that carries e = exp(-a r^2)
  double one, x, x2, twox, two;  // e, e x, e x^2, 2 e x, 2 e
  double tx, tx2, xr;            // 2a x, (2a x)^2, x (unscaled)
};
__device__ __forceinline__ void shell_axis_setup(ShellAxis &a, double ta) {   // a.x, a.x2 set by the caller
  a.twox = a.x + a.x;
  a.tx = ta * a.x;
  a.tx2 = a.tx * a.tx;
}
__device__ __forceinline__ void shell_axis_setup(ShellAxisE &a, double x, double x2, double ta, double e) {
  a.xr = x;
  a.tx = ta * x;
  a.tx2 = a.tx * a.tx;
  a.one = e;
  a.x = e * x;
  a.x2 = e * x2;
  a.twox = a.x + a.x;
  a.two = e + e;
}
static __constant__ double c_sh_lf[8] = {0, 1, 2, 3, 4, 5, 6, 7};          // l
static __constant__ double c_sh_ll1[8] = {0, 0, 2, 6, 12, 20, 30, 42};     // l (l - 1)
static __constant__ double c_sh_2l1[8] = {1, 3, 5, 7, 9, 11, 13, 15};      // 2 l + 1

// l <= 2 (s, p, d shells): no multiplication for the powers.  ta = 2a.
template <int ND>
__device__ __forceinline__ void shell_axis(const ShellAxis &a, int l, double ta, double &f0, double &f1, double &f2) {
  f0 = l == 2 ? a.x2 : (l == 1 ? a.x : 1.0);
  if (ND > 1) {
    const double lpm1 = l == 2 ? a.twox : (l == 1 ? 1.0 : 0.0);   // l x^(l-1)
    f1 = fma(-a.tx, f0, lpm1);
    if (ND > 4) f2 = fma(f0, fma(l == 2 ? -5.0 : (l == 1 ? -3.0 : -1.0), ta, a.tx2), l == 2 ? 2.0 : 0.0);
  }
}
template <int ND>
__device__ __forceinline__ void shell_axis(const ShellAxisE &a, int l, double ta, double &f0, double &f1, double &f2) {
  f0 = l == 2 ? a.x2 : (l == 1 ? a.x : a.one);
  if (ND > 1) {
    const double lpm1 = l == 2 ? a.twox : (l == 1 ? a.one : 0.0);
    f1 = fma(-a.tx, f0, lpm1);
    if (ND > 4) f2 = fma(f0, fma(l == 2 ? -5.0 : (l == 1 ? -3.0 : -1.0), ta, a.tx2), l == 2 ? a.two : 0.0);
  }
}
// any l up to 7 (f shells and up): powers by multiplication, e applied at the end
template <int ND>
__device__ __forceinline__ void shell_axis_big(double x, double x2, double tx, double tx2, int l, double ta, double e, double &f0,
                                               double &f1, double &f2) {
  const double x4 = x2 * x2;
  const int n = l - 2;
  const double pm2 = l >= 2 ? ((n & 1) ? x : 1.0) * ((n & 2) ? x2 : 1.0) * ((n & 4) ? x4 : 1.0) : 0.0;
  const double pm1 = l >= 2 ? pm2 * x : (l == 1 ? 1.0 : 0.0);
  f0 = l >= 1 ? pm1 * x : 1.0;
  if (ND > 1) {
    f1 = fma(-tx, f0, c_sh_lf[l & 7] * pm1) * e;
    if (ND > 4) f2 = fma(f0, fma(-c_sh_2l1[l & 7], ta, tx2), c_sh_ll1[l & 7] * pm2) * e;
  }
  f0 *= e;
}

// chi(1:ND) of one primitive of the shell; mode = WfnParams::chi_mode (value-only launches: bits 0-1 select the k-th
// Cartesian derivative as the "value")
template <int ND, bool BIG>
__device__ __forceinline__ void shell_prim_chi(const ShellAxis &ax, const ShellAxis &ay, const ShellAxisE &az, int lx, int ly,
                                               int lz, double ta, int mode, double *chi) {
  double x0, x1 = 0, x2 = 0, y0, y1 = 0, y2 = 0, z0, z1 = 0, z2 = 0;
  if (ND == 1 && (mode & 3)) {
    if (BIG) {
      shell_axis_big<4>(ax.x, ax.x2, ax.tx, ax.tx2, lx, ta, 1.0, x0, x1, x2);
      shell_axis_big<4>(ay.x, ay.x2, ay.tx, ay.tx2, ly, ta, 1.0, y0, y1, y2);
      shell_axis_big<4>(az.xr, az.xr * az.xr, az.tx, az.tx2, lz, ta, az.one, z0, z1, z2);
    } else {
      shell_axis<4>(ax, lx, ta, x0, x1, x2);
      shell_axis<4>(ay, ly, ta, y0, y1, y2);
      shell_axis<4>(az, lz, ta, z0, z1, z2);
    }
    const int k = mode & 3;
    chi[0] = k == 1 ? x1 * y0 * z0 : (k == 2 ? x0 * y1 * z0 : x0 * y0 * z1);
    return;
  }
  if (BIG) {
    shell_axis_big<ND>(ax.x, ax.x2, ax.tx, ax.tx2, lx, ta, 1.0, x0, x1, x2);
    shell_axis_big<ND>(ay.x, ay.x2, ay.tx, ay.tx2, ly, ta, 1.0, y0, y1, y2);
    shell_axis_big<ND>(az.xr, az.xr * az.xr, az.tx, az.tx2, lz, ta, az.one, z0, z1, z2);
  } else {
    shell_axis<ND>(ax, lx, ta, x0, x1, x2);
    shell_axis<ND>(ay, ly, ta, y0, y1, y2);
    shell_axis<ND>(az, lz, ta, z0, z1, z2);
  }
  if (ND == 1) {
    chi[0] = x0 * y0 * z0;
    return;
  }
  const double yz = y0 * z0, xz = x0 * z0, xy = x0 * y0;
  chi[0] = x0 * yz;
  chi[ND > 1 ? 1 : 0] = x1 * yz;
  chi[ND > 1 ? 2 : 0] = y1 * xz;
  chi[ND > 1 ? 3 : 0] = xy * z1;
  if (ND > 4) {
    chi[ND > 4 ? 4 : 0] = x2 * yz;
    chi[ND > 4 ? 5 : 0] = y2 * xz;
    chi[ND > 4 ? 6 : 0] = xy * z2;
    chi[ND > 4 ? 7 : 0] = (x1 * y1) * z0;
    chi[ND > 4 ? 8 : 0] = x1 * (y0 * z1);
    chi[ND > 4 ? 9 : 0] = x0 * (y1 * z1);
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
