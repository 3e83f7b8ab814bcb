// c2gpu_internal.h -- host-side declarations shared by the translation units of libc2gpu.so.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "promol_kernel.cuh"
#include "wfn_kernel.cuh"

namespace c2 {

// Kernel configurations of the wavefunction kernel (TileCfg instantiations in wfn_launch.cu)
struct WfnCfgInfo {
  const char *name;
  int wm, wn, ni, kc, ctas;   // template parameters
  int P, NB, NT;
  int ta, tb, tc;             // sub-tile shape of P points
  size_t smem[3];             // dynamic shared memory for ND = 1, 4, 10
};
int wfn_num_configs();
inline bool wfn_is_ws(int cfg) { return cfg == 6 || cfg == 12 || cfg == 13 || cfg == 14; }   // warp-specialised kernel and its variants
const WfnCfgInfo &wfn_cfg_info(int cfg);
// cfg minimising the padded MO count (then the larger tile); `force` >= 0 overrides
int wfn_choose_config(int nmo, int force);
cudaError_t wfn_launch(int cfg, int nd, const WfnParams &p, int grid, cudaStream_t st);
cudaError_t wfn_count_launch(const WfnDev &w, const GridDesc &g, const CellXf &xf, unsigned long long *count,
                             cudaStream_t st);
cudaError_t nci_epilogue_launch(const GridDesc &g, const OutDesc &o, const double *raw, size_t stride, double nci_const,
                                cudaStream_t st);
cudaError_t fp64_peak_measure(double *dmma_tflops, double *dfma_tflops);
cudaError_t promol_launch(int nd, const PromolParams &p, int grid, cudaStream_t st);
cudaError_t nci_cutoffs_launch(long npts, const double *crho, double *cgrad, const double *rhoat, int nfrag,
                               const double *rhofrag, double rhoparam, double rhoparam2, double rhocut,
                               double dimcut, double rhoplot, double sr0, double sr1, int onlyneg, double rho_void,
                               double *raw_sums, long *ndat, double *dat_dev, long maxdat, cudaStream_t st);

void nci_cutoffs_release();   // frees the per-device scratch of the cutoff pass

// cube_format.cu: text records of cube / .dat files (fmt and order as in include/c2gpu.h)
long format_record_bytes(int nval, int fmt);
cudaError_t format_records_launch(const double *g_dev, char *text_dev, long nrec, int nval, int fmt, int order, int n1,
                                  int n2, int force_exact, cudaStream_t st);

extern long g_launches;       // kernels launched by the library

// sets the message c2gpu_last_error returns and hands `code` back (for entry points defined outside c2gpu.cu)
int c2_fail(int code, const std::string &msg);

}  // namespace c2
