// c2gpu.cu -- the C ABI of include/c2gpu.h: handle construction (host-side flattening of molwfn / environ data into
// the device layouts) and the grid drivers that replace rhoplot_cube / nciplot / promolecular_grid evaluation loops.
// No CPU evaluation path exists in this library: every entry either launches the sm_100a kernels or fails.
#include "../../include/c2gpu.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <numeric>
#include <thread>

#include "c2gpu_internal.h"

using namespace c2;

namespace {

thread_local std::string g_err;
std::vector<int> g_devs;
int g_force_cfg = -1;
int g_fmt_force_exact = 0;
int g_super = 8;
int g_nci_two_pass = 1;

int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}
#define CUDA_OK(call)                                                                               \
  do {                                                                                              \
    cudaError_t e__ = (call);                                                                       \
    if (e__ != cudaSuccess)                                                                         \
      return fail(2, std::string(#call) + ": " + cudaGetErrorString(e__));                          \
  } while (0)

struct DevGuard {
  int prev = 0;
  explicit DevGuard(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); }
  ~DevGuard() { cudaSetDevice(prev); }
};

int ensure_devices() {
  if (!g_devs.empty()) return 0;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) return fail(3, "no CUDA device visible (this library has no CPU path)");
  int cur = 0;
  cudaGetDevice(&cur);
  g_devs.push_back(cur);
  return 0;
}

// primitive type -> (lx,ly,lz), src/wfn_private@proc.F90:2695-2705
const unsigned char LI[56][3] = {
    {0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {2, 0, 0}, {0, 2, 0}, {0, 0, 2}, {1, 1, 0}, {1, 0, 1}, {0, 1, 1},
    {3, 0, 0}, {0, 3, 0}, {0, 0, 3}, {2, 1, 0}, {2, 0, 1}, {0, 2, 1}, {1, 2, 0}, {1, 0, 2}, {0, 1, 2}, {1, 1, 1},
    {4, 0, 0}, {0, 4, 0}, {0, 0, 4}, {3, 1, 0}, {3, 0, 1}, {1, 3, 0}, {0, 3, 1}, {1, 0, 3}, {0, 1, 3}, {2, 2, 0},
    {2, 0, 2}, {0, 2, 2}, {2, 1, 1}, {1, 2, 1}, {1, 1, 2}, {0, 0, 5}, {0, 1, 4}, {0, 2, 3}, {0, 3, 2}, {0, 4, 1},
    {0, 5, 0}, {1, 0, 4}, {1, 1, 3}, {1, 2, 2}, {1, 3, 1}, {1, 4, 0}, {2, 0, 3}, {2, 1, 2}, {2, 2, 1}, {2, 3, 0},
    {3, 0, 2}, {3, 1, 1}, {3, 2, 0}, {4, 0, 1}, {4, 1, 0}, {5, 0, 0}};

struct Replica {
  int dev = 0;
  int nsm = 0;
  std::vector<void *> allocs;
  cudaStream_t stream = nullptr;
  WfnDev w{};
  PromolDev m{};
  int *scratch = nullptr, *counter = nullptr;
  int grid = 0;
  double *acache = nullptr;   // warp-specialised kernel: chi-tile cache (see WfnParams)
  int *acache_nz = nullptr;
  double *ws_rec = nullptr;
  int acache_chunks = 0;
  // device output arrays of the host-array drivers, kept between calls (CUBE is typically issued once per property and
  // NCIPLOT evaluates crho/cgrad/rhoat on the same lattice): freed with the handle
  std::vector<double *> outbuf;
  size_t outbuf_n = 0;
  // scratch of the core pre-pass (f%usecore): nd raw sums per point of the evaluated range, grow-only
  double *corebuf = nullptr;
  size_t corebuf_n = 0;
  // scratch of the two-pass NCIPLOT evaluation: ten raw sums per point of the evaluated range, grow-only
  double *rawbuf = nullptr;
  size_t rawbuf_n = 0;
  // copy-back of the host-array drivers: own copy stream, two pinned bounce buffers and their events (host_eval)
  cudaStream_t cstream = nullptr;
  char *bounce[2] = {nullptr, nullptr};
  cudaEvent_t bev[2] = {nullptr, nullptr};
};

struct Handle {
  int kind = 0;   // 1 wavefunction, 2 promolecular
  CellXf xf{};
  int cfg = 0;    // wavefunction kernel configuration
  int natom = 0, npri = 0, nshell = 0, nmo = 0, nblk = 0;
  std::vector<Replica> reps;
  Handle *core = nullptr;   // f%usecore: promolecular handle over the core tables (c2gpu_field_set_core); not owned
};

template <class T>
int upload(Replica &r, const std::vector<T> &h, const T **dptr) {
  void *d = nullptr;
  size_t bytes = std::max<size_t>(sizeof(T) * h.size(), 16);
  CUDA_OK(cudaMalloc(&d, bytes));
  r.allocs.push_back(d);
  if (!h.empty()) CUDA_OK(cudaMemcpy(d, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice));
  *dptr = reinterpret_cast<const T *>(d);
  return 0;
}

void set_cell(CellXf &xf, int ismolecule, const double *x2c, const double *c2x, const double *xr2c, const double *c2xr) {
  memcpy(xf.x2c, x2c, 72);
  memcpy(xf.c2x, c2x, 72);
  memcpy(xf.xr2c, xr2c ? xr2c : x2c, 72);
  memcpy(xf.c2xr, c2xr ? c2xr : c2x, 72);
  xf.ismolecule = ismolecule;
}

double nci_const() {   // src/nci@proc.f90:86-87, evaluated at run time like the reference
  const double pi = 3.14159265358979323846264338328;
  return 2.0 * pow(3.0 * pi * pi, 1.0 / 3.0);
}

}  // namespace

namespace c2 {
int c2_fail(int code, const std::string &msg) { return fail(code, msg); }
}  // namespace c2

extern "C" {

int c2gpu_init(int ndev_req, int *ndev) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) return fail(3, "no CUDA device visible (this library has no CPU path)");
  if (ndev_req > 0) n = std::min(n, ndev_req);
  g_devs.clear();
  for (int i = 0; i < n; i++) g_devs.push_back(i);
  if (ndev) *ndev = n;
  return 0;
}

int c2gpu_init_devices(const int *devs, int n) {
  int nvis = 0;
  if (cudaGetDeviceCount(&nvis) != cudaSuccess || nvis <= 0) return fail(3, "no CUDA device visible (this library has no CPU path)");
  if (n <= 0 || !devs) return fail(1, "c2gpu_init_devices: empty device list");
  for (int i = 0; i < n; i++)
    if (devs[i] < 0 || devs[i] >= nvis) return fail(1, "c2gpu_init_devices: device index out of range");
  g_devs.assign(devs, devs + n);
  return 0;
}

int c2gpu_finalize(void) {
  nci_cutoffs_release();
  g_devs.clear();
  return 0;
}

int c2gpu_last_error(char *buf, int len) {
  if (!buf || len <= 0) return 1;
  snprintf(buf, (size_t)len, "%s", g_err.c_str());
  return 0;
}

long c2gpu_launch_count(void) { return g_launches; }

int c2gpu_set_option(const char *name, int value) {
  if (!name) return 1;
  if (!strcmp(name, "wfn_cfg")) { g_force_cfg = value; return 0; }
  if (!strcmp(name, "super")) { if (value < 4 || value % 4) return 1; g_super = value; return 0; }
  if (!strcmp(name, "nci_two_pass")) { g_nci_two_pass = value; return 0; }   // development: A/B of the NCIPLOT epilogue
  if (!strcmp(name, "fmt_force_exact")) { g_fmt_force_exact = value; return 0; }   // tests: big-integer rounding path
  return 1;
}

int c2gpu_wfn_choose_config(int nmoocc) { return wfn_choose_config(nmoocc, -1); }

int c2gpu_destroy(void *handle) {
  Handle *h = static_cast<Handle *>(handle);
  if (!h) return 0;
  for (auto &r : h->reps) {
    DevGuard gd(r.dev);
    for (void *p : r.allocs) cudaFree(p);
    for (double *p : r.outbuf) cudaFree(p);
    if (r.corebuf) cudaFree(r.corebuf);
    if (r.rawbuf) cudaFree(r.rawbuf);
    if (r.stream) cudaStreamDestroy(r.stream);
    if (r.cstream) cudaStreamDestroy(r.cstream);
    for (int b = 0; b < 2; b++) {
      if (r.bounce[b]) cudaFreeHost(r.bounce[b]);
      if (r.bev[b]) cudaEventDestroy(r.bev[b]);
    }
  }
  delete h;
  return 0;
}

int c2gpu_field_set_core(void *handle, void *hcore) {
  Handle *h = static_cast<Handle *>(handle), *c = static_cast<Handle *>(hcore);
  if (!h || (h->kind != 1 && h->kind != 2)) return fail(1, "c2gpu_field_set_core: invalid field handle");
  if (c && c->kind != 2) return fail(1, "c2gpu_field_set_core: the core handle must be a promolecular handle");
  if (c == h) return fail(1, "c2gpu_field_set_core: a field cannot be its own core");
  if (c)
    for (const Replica &r : h->reps) {
      bool found = false;
      for (const Replica &rc : c->reps) found = found || rc.dev == r.dev;
      if (!found) return fail(1, "c2gpu_field_set_core: the core handle is not resident on every device of the field");
    }
  h->core = c;
  return 0;
}

int c2gpu_wfn_create(int natom, const double *xyz, int npri, const int *icenter, const int *itype, const double *expo,
                     const double *dran, int nmoocc, int ldc_in, const double *cmo, const double *occ, int wfntyp,
                     int nalpha, int nedf, const int *icenter_edf, const int *itype_edf, const double *e_edf,
                     const double *c_edf, const double *dran_edf, int ismolecule, const double *m_x2c,
                     const double *m_c2x, void **handle) {
  (void)wfntyp; (void)nalpha;   // spin-resolved quantities are not exported by the CUBE / NCIPLOT loops
  if (!handle) return fail(1, "c2gpu_wfn_create: null handle pointer");
  *handle = nullptr;
  if (!ismolecule) return fail(-1, "wavefunction fields are supported for molecules only");
  if (natom <= 0 || npri <= 0 || nmoocc <= 0 || ldc_in < nmoocc) return fail(1, "c2gpu_wfn_create: bad sizes");
  for (int i = 0; i < npri; i++) {
    if (itype[i] < 1 || itype[i] > 56) return fail(-1, "primitive type outside 1..56");
    if (icenter[i] < 1 || icenter[i] > natom) return fail(1, "primitive centre out of range");
  }
  for (int i = 0; i < nedf; i++)
    if (itype_edf[i] < 1 || itype_edf[i] > 56 || icenter_edf[i] < 1 || icenter_edf[i] > natom)
      return fail(1, "EDF primitive out of range");
  if (int rc = ensure_devices()) return rc;

  Handle *h = new Handle;
  h->kind = 1;
  set_cell(h->xf, ismolecule, m_x2c, m_c2x, nullptr, nullptr);
  h->natom = natom; h->npri = npri; h->nmo = nmoocc;
  h->cfg = wfn_choose_config(nmoocc, g_force_cfg);
  // the split-producer kernel keeps every chi tile of a sub-tile in its cache: 80 KB per 32 primitives and CTA
  if ((h->cfg == 6 || h->cfg == 14) && g_force_cfg < 0 && (double)((npri + 31) / 32) * 40960.0 * 2.0 * 160.0 > 48e9) h->cfg = 3;
  if (h->cfg >= 9 && h->cfg <= 11 && nmoocc > wfn_cfg_info(h->cfg).NB) {   // the warp kernel holds ONE block of MO columns
    delete h;
    return fail(1, "c2gpu_wfn_create: the forced warp-kernel configuration holds fewer MO columns than nmoocc");
  }
  const WfnCfgInfo &ci = wfn_cfg_info(h->cfg);
  h->nblk = (nmoocc + ci.NB - 1) / ci.NB;
  const int ldc = h->nblk * ci.NB;

  // icord: primitives ordered by centre, file order kept inside an atom (complete_struct :3112-3130)
  std::vector<int> ord(npri);
  std::iota(ord.begin(), ord.end(), 0);
  std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return icenter[a] < icenter[b]; });
  // shells: runs with the same centre, exponent and screening radius
  std::vector<double4> shell_xa;
  std::vector<double2> shell_cut;
  std::vector<int> shell_pbeg, prim_shell(npri);
  std::vector<unsigned char> prim_l(4 * (size_t)npri);
  for (int p = 0; p < npri; p++) {
    int i = ord[p];
    bool newshell = (p == 0);
    if (!newshell) {
      int j = ord[p - 1];
      newshell = icenter[i] != icenter[j] || expo[i] != expo[j] || dran[i] != dran[j];
      // at most 21 primitives (a full Cartesian h shell) per device shell: the warp kernel's list holds 32 shells per step
      newshell = newshell || p - shell_pbeg.back() >= 21;
    }
    if (newshell) {
      const double *c = xyz + 3 * (size_t)(icenter[i] - 1);
      shell_xa.push_back(make_double4(c[0], c[1], c[2], expo[i]));
      shell_cut.push_back(make_double2(dran[i], dran[i] * dran[i]));
      shell_pbeg.push_back(p);
    }
    prim_shell[p] = (int)shell_xa.size() - 1;
    for (int d = 0; d < 3; d++) prim_l[4 * (size_t)p + d] = LI[itype[i] - 1][d];
    prim_l[4 * (size_t)p + 3] = 0;
  }
  shell_pbeg.push_back(npri);
  h->nshell = (int)shell_xa.size();
  // cmoT[p][imo] = cmo(imo, icord(p)), zero padded to ldc columns
  std::vector<double> cmoT((size_t)npri * ldc, 0.0), occp(ldc, 0.0);
  for (int p = 0; p < npri; p++)
    memcpy(&cmoT[(size_t)p * ldc], cmo + (size_t)ord[p] * ldc_in, sizeof(double) * nmoocc);
  memcpy(occp.data(), occ, sizeof(double) * nmoocc);
  std::vector<double> axyz(xyz, xyz + 3 * (size_t)natom);
  std::vector<double4> edf_xa(nedf), edf_cc(nedf);
  std::vector<unsigned char> edf_l(4 * (size_t)std::max(nedf, 1), 0);
  for (int i = 0; i < nedf; i++) {
    const double *c = xyz + 3 * (size_t)(icenter_edf[i] - 1);
    edf_xa[i] = make_double4(c[0], c[1], c[2], e_edf[i]);
    edf_cc[i] = make_double4(c_edf[i], dran_edf[i], dran_edf[i] * dran_edf[i], 0.0);
    for (int d = 0; d < 3; d++) edf_l[4 * (size_t)i + d] = LI[itype_edf[i] - 1][d];
  }

  for (int dev : g_devs) {
    h->reps.emplace_back();
    Replica &r = h->reps.back();
    r.dev = dev;
    DevGuard gd(dev);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) { c2gpu_destroy(h); return fail(2, "cudaGetDeviceProperties failed"); }
    if (prop.major != 10) { c2gpu_destroy(h); return fail(3, "this library is built for sm_100a (B200) only"); }
    r.nsm = prop.multiProcessorCount;
    int rc = 0;
    WfnDev &w = r.w;
    w.natom = natom; w.npri = npri; w.nshell = h->nshell; w.nmo = nmoocc; w.ldc = ldc; w.nedf = nedf;
    rc |= upload(r, axyz, &w.atom_xyz);
    rc |= upload(r, shell_xa, &w.shell_xa);
    rc |= upload(r, shell_cut, &w.shell_cut);
    rc |= upload(r, shell_pbeg, &w.shell_pbeg);
    rc |= upload(r, prim_shell, &w.prim_shell);
    rc |= upload(r, prim_l, &w.prim_l);
    rc |= upload(r, cmoT, &w.cmoT);
    rc |= upload(r, occp, &w.occ);
    rc |= upload(r, edf_xa, &w.edf_xa);
    rc |= upload(r, edf_cc, &w.edf_cc);
    rc |= upload(r, edf_l, &w.edf_l);
    r.grid = r.nsm * ci.ctas;
    size_t per_cta = (size_t)((h->nshell + 3) & ~3) + 4 * ((size_t)((npri + 1) & ~1) + 64);
    void *d = nullptr;
    if (!rc && cudaMalloc(&d, sizeof(int) * per_cta * r.grid) == cudaSuccess) { r.allocs.push_back(d); r.scratch = (int *)d; } else rc = 2;
    if (!rc && cudaMalloc(&d, 64) == cudaSuccess) { r.allocs.push_back(d); r.counter = (int *)d; } else rc = 2;
    if (!rc && wfn_is_ws(h->cfg)) {
      r.acache_chunks = (npri + 31) / 32 + 2;   // worst case: every primitive active in a sub-tile (+ rounding to stages of up to 64)
      size_t nt = (size_t)r.grid * 2 * r.acache_chunks;
      if (cudaMalloc(&d, nt * 5120 * sizeof(double)) == cudaSuccess) { r.allocs.push_back(d); r.acache = (double *)d; } else rc = 2;
      if (!rc && cudaMalloc(&d, nt * 8 * sizeof(int)) == cudaSuccess) { r.allocs.push_back(d); r.acache_nz = (int *)d; } else rc = 2;
      size_t shell_cap = (size_t)((h->nshell + 3) & ~3);   // one 96-byte record per active shell of a sub-tile
      if (!rc && cudaMalloc(&d, (size_t)r.grid * shell_cap * 96) == cudaSuccess) { r.allocs.push_back(d); r.ws_rec = (double *)d; } else rc = 2;
    }
    if (!rc && cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking) != cudaSuccess) rc = 2;
    if (rc) {
      std::string msg = g_err.empty() ? "device allocation failed" : g_err;
      c2gpu_destroy(h);
      return fail(2, msg);
    }
  }
  *handle = h;
  return 0;
}

int c2gpu_promol_create(int nenv, int ncell, const double *xyz_env, const int *ispc, const int *cidx, int nspc,
                        const int *zspc, const int *ngrid, const double *a, const double *b, const double *cutrad,
                        const double *tab_r, const double *tab_f, const double *tab_fp, const double *tab_fpp,
                        int ismolecule, const double *m_x2c, const double *m_c2x, const double *m_xr2c,
                        const double *m_c2xr, const unsigned char *fragmask, void **handle) {
  if (!handle) return fail(1, "c2gpu_promol_create: null handle pointer");
  *handle = nullptr;
  if (nenv <= 0 || nspc <= 0) return fail(1, "c2gpu_promol_create: bad sizes");
  if (int rc = ensure_devices()) return rc;
  Handle *h = new Handle;
  h->kind = 2;
  set_cell(h->xf, ismolecule, m_x2c, m_c2x, m_xr2c, m_c2xr);
  h->natom = nenv;

  // species tables: the interpolating cubic of every 4-node stencil (grid1%interp, src/grid1mod@proc.f90:108-133), as
  // coefficients of (r - xc)^k about the stencil's second node, from Newton divided differences in long double
  std::vector<SpcDev> spc(nspc);
  std::vector<double> poly1, poly4, poly10;
  size_t nnodes = 0;
  auto cubic = [](const double *x, const double *y, double *c) {
    const long double x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
    const long double f01 = ((long double)y[1] - y[0]) / (x1 - x0), f12 = ((long double)y[2] - y[1]) / (x2 - x1),
                      f23 = ((long double)y[3] - y[2]) / (x3 - x2);
    const long double f012 = (f12 - f01) / (x2 - x0), f123 = (f23 - f12) / (x3 - x1);
    const long double a0 = y[0], a1 = f01, a2 = f012, a3 = (f123 - f012) / (x3 - x0);
    const long double h0 = x1 - x0, h1 = x2 - x1;
    // p = a0 + a1 (t+h0) + a2 t (t+h0) + a3 t (t+h0) (t-h1),  t = r - x1
    c[0] = (double)(a0 + a1 * h0);
    c[1] = (double)(a1 + a2 * h0 - a3 * h0 * h1);
    c[2] = (double)(a2 + a3 * (h0 - h1));
    c[3] = (double)a3;
  };
  size_t off_in = 0;
  double rcutmax = 0.0;
  std::vector<char> used(nspc, 0);
  for (int i = 0; i < nenv; i++) {
    if (ispc[i] < 1 || ispc[i] > nspc) { delete h; return fail(1, "species index out of range"); }
    used[ispc[i] - 1] = 1;
  }
  for (int s = 0; s < nspc; s++) {
    SpcDev &d = spc[s];
    int n = ngrid[s];
    d.off = (int)nnodes; d.n = n; d.z = zspc[s]; d.pad = 0;
    d.a = a[s]; d.b = b[s]; d.cutrad = cutrad[s];
    d.ln2_b = b[s] >= 1e-3 ? std::log(2.0) / b[s] : std::nan("");   // NaN: the index pre-screen never accepts
    d.lna_b = (b[s] != 0.0 && a[s] > 0.0) ? std::log(a[s]) / b[s] : 0.0;
    if (n >= 4) {
      const double *r = tab_r + off_in;
      d.rmax = r[n - 1]; d.r1 = r[0];
      d.rcut = std::min(cutrad[s], d.rmax);                    // src/environmod@proc.f90:1257
      d.rmax_lo = d.rmax * (1.0 - 1e-14);
      d.rfloor = std::max(d.r1, 1e-14);
      for (int k = 0; k < n; k++) {
        double rec[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
        if (k + 3 < n) {
          rec[0] = r[k + 1];
          cubic(r + k, tab_f + off_in + k, rec + 1);
          cubic(r + k, tab_fp + off_in + k, rec + 5);
          cubic(r + k, tab_fpp + off_in + k, rec + 9);
        }
        poly1.insert(poly1.end(), rec, rec + 6);     // xc c0 c1 c2 c3 -
        poly4.insert(poly4.end(), rec, rec + 10);    // ... d0 d1 d2 d3 -
        poly10.insert(poly10.end(), rec, rec + 14);  // ... e0 e1 e2 e3 -
      }
      nnodes += (size_t)n;
    } else {
      d.n = 0; d.rmax = 0; d.r1 = 0; d.rmax_lo = 0; d.rfloor = 1e-14;
      d.rcut = cutrad[s];                                      // :1259
    }
    off_in += (size_t)std::max(n, 0);
    if (zspc[s] <= 0 || zspc[s] > 118) { d.n = 0; d.rcut = 0.0; }   // :1250, :1293
    if (used[s]) rcutmax = std::max(rcutmax, d.rcut);
  }
  // bins
  const double bs = 4.0;
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int i = 0; i < nenv; i++)
    for (int d = 0; d < 3; d++) { lo[d] = std::min(lo[d], xyz_env[3 * (size_t)i + d]); hi[d] = std::max(hi[d], xyz_env[3 * (size_t)i + d]); }
  int nbin[3];
  for (int d = 0; d < 3; d++) nbin[d] = (int)floor((hi[d] - lo[d]) / bs) + 1;
  size_t nbt = (size_t)nbin[0] * nbin[1] * nbin[2];
  std::vector<int> bin(nenv), ord(nenv), bin_start(nbt + 1, 0);
  for (int i = 0; i < nenv; i++) {
    int bi[3];
    for (int d = 0; d < 3; d++) bi[d] = std::min(nbin[d] - 1, std::max(0, (int)floor((xyz_env[3 * (size_t)i + d] - lo[d]) / bs)));
    bin[i] = bi[0] + nbin[0] * (bi[1] + nbin[1] * bi[2]);
    bin_start[bin[i] + 1]++;
  }
  for (size_t k = 0; k < nbt; k++) bin_start[k + 1] += bin_start[k];
  std::iota(ord.begin(), ord.end(), 0);
  std::stable_sort(ord.begin(), ord.end(), [&](int x, int y) { return bin[x] < bin[y]; });
  std::vector<double4> atoms(nenv);
  for (int k = 0; k < nenv; k++) {
    int i = ord[k];
    long long bits = ((long long)(cidx[i] - 1) << 32) | (long long)(unsigned)(ispc[i] - 1);
    double wv;
    memcpy(&wv, &bits, 8);
    atoms[k] = make_double4(xyz_env[3 * (size_t)i], xyz_env[3 * (size_t)i + 1], xyz_env[3 * (size_t)i + 2], wv);
  }
  std::vector<unsigned char> frag;
  if (fragmask) frag.assign(fragmask, fragmask + ncell);

  for (int dev : g_devs) {
    h->reps.emplace_back();
    Replica &r = h->reps.back();
    r.dev = dev;
    DevGuard gd(dev);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) { c2gpu_destroy(h); return fail(2, "cudaGetDeviceProperties failed"); }
    if (prop.major != 10) { c2gpu_destroy(h); return fail(3, "this library is built for sm_100a (B200) only"); }
    r.nsm = prop.multiProcessorCount;
    PromolDev &m = r.m;
    m.nenv = nenv; m.ncell = ncell; m.nspc = nspc; m.rcutmax = rcutmax;
    for (int d = 0; d < 3; d++) { m.bin0[d] = lo[d]; m.nbin[d] = nbin[d]; }
    m.bin_inv = 1.0 / bs;
    int rc = 0;
    rc |= upload(r, atoms, &m.atoms);
    rc |= upload(r, bin_start, &m.bin_start);
    rc |= upload(r, spc, &m.spc);
    rc |= upload(r, poly1, &m.poly1);
    rc |= upload(r, poly4, &m.poly4);
    rc |= upload(r, poly10, &m.poly10);
    m.fragmask = nullptr;
    if (fragmask) rc |= upload(r, frag, &m.fragmask);
    void *d = nullptr;
    if (!rc && cudaMalloc(&d, 64) == cudaSuccess) { r.allocs.push_back(d); r.counter = (int *)d; } else rc = 2;
    if (!rc && cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking) != cudaSuccess) rc = 2;
    if (rc) {
      std::string msg = g_err.empty() ? "device allocation failed" : g_err;
      c2gpu_destroy(h);
      return fail(2, msg);
    }
  }
  *handle = h;
  return 0;
}

int c2gpu_wfn_config(void *hh, int cfg[8]) {
  Handle *h = static_cast<Handle *>(hh);
  if (!h || h->kind != 1) return fail(1, "c2gpu_wfn_config: not a wavefunction handle");
  const WfnCfgInfo &ci = wfn_cfg_info(h->cfg);
  cfg[0] = h->cfg; cfg[1] = ci.P; cfg[2] = ci.NB; cfg[3] = h->nblk; cfg[4] = ci.NT; cfg[5] = ci.kc;
  cfg[6] = h->nshell; cfg[7] = h->reps.empty() ? 0 : h->reps[0].grid;
  return 0;
}

}  // extern "C"

#include "c2gpu_drivers.inc"
