// FP64 peak microbenchmarks for B200 (sm_100a): the roofline denominator for the
// wavefunction path (DFMA chains, DMMA chains, mixed).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_peak fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int NCH>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
  double acc[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) acc[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NCH; i++) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NCH; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1684(double* c, const double* a, double b) {
  asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(b));
}
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma884(double* out, int iters, double a, double b) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NACC; i++) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC, int SHAPE>
__global__ void __launch_bounds__(256) k_dmma16(double* out, int iters, double a0, double b0) {
  double c[NACC][4];
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = a0 + i;
#pragma unroll
  for (int i = 0; i < 4; i++) b[i] = b0 + i;
#pragma unroll
  for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NACC; i++) {
      if (SHAPE == 4) dmma1684(c[i], a, b[0]);
      if (SHAPE == 8) dmma1688(c[i], a, b);
      if (SHAPE == 16) dmma16816(c[i], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mixed: NACC dmma884 + NF dfma per iteration (are the pipes separate?)
template <int NACC, int NF>
__global__ void __launch_bounds__(256) k_mixed(double* out, int iters, double a, double b) {
  double c[NACC][2];
  double f[NF];
#pragma unroll
  for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
#pragma unroll
  for (int i = 0; i < NF; i++) f[i] = i + threadIdx.x;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NACC; i++) dmma884(c[i][0], c[i][1], a, b);
#pragma unroll
    for (int i = 0; i < NF; i++) f[i] = fma(f[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
#pragma unroll
  for (int i = 0; i < NF; i++) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_exp(double* out, int iters, double a) {
  double x0 = -1e-3 * threadIdx.x, x1 = x0 - 0.5, x2 = x0 - 1.5, x3 = x0 - 2.5;
  double s = 0;
  for (int it = 0; it < iters; it++) {
    s += exp(x0) + exp(x1) + exp(x2) + exp(x3);
    x0 -= a; x1 -= a; x2 -= a; x3 -= a;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// shared-memory 64-bit load bandwidth (conflict-free)
__global__ void __launch_bounds__(256) k_lds(double* out, int iters) {
  __shared__ double sm[4096];
  for (int i = threadIdx.x; i < 4096; i += 256) sm[i] = i;
  __syncthreads();
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  int base = threadIdx.x;
  for (int it = 0; it < iters; it++) {
    s0 += sm[(base + 0) & 4095]; s1 += sm[(base + 256) & 4095];
    s2 += sm[(base + 512) & 4095]; s3 += sm[(base + 768) & 4095];
    base += 1024;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1 + s2 + s3;
}

template <typename F>
float timeit(F f, int reps = 3) {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount;
  printf("device %s sms %d clock %d kHz\n", p.name, nsm, p.clockRate);
  double* out; CK(cudaMalloc(&out, sizeof(double) * nsm * 16 * 1024));
  const int iters = 20000;
  for (int bps = 1; bps <= 8; bps *= 2) {
    int grid = nsm * bps;
    float ms;
    ms = timeit([&] { k_dfma<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dfma8\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 8 * iters * 256.0 * grid / ms / 1e9);
    ms = timeit([&] { k_dfma<16><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dfma16\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 16 * iters * 256.0 * grid / ms / 1e9);
    ms = timeit([&] { k_dmma884<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dmma884x8\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 8 * 256 * iters * 8.0 * grid / ms / 1e9);
    ms = timeit([&] { k_dmma884<16><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dmma884x16\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 16 * 256 * iters * 8.0 * grid / ms / 1e9);
    ms = timeit([&] { k_dmma16<8, 4><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dmma1684x8\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 8 * 512 * iters * 8.0 * grid / ms / 1e9);
    ms = timeit([&] { k_dmma16<8, 8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dmma1688x8\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 8 * 1024 * iters * 8.0 * grid / ms / 1e9);
    ms = timeit([&] { k_dmma16<8, 16><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"dmma16816x8\",\"ctas_per_sm\":%d,\"tflops\":%.2f}\n", bps, 2.0 * 8 * 2048 * iters * 8.0 * grid / ms / 1e9);
    ms = timeit([&] { k_mixed<8, 8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"mixed_dmma8_dfma8\",\"ctas_per_sm\":%d,\"tflops_dmma\":%.2f,\"tflops_dfma\":%.2f}\n", bps,
           2.0 * 8 * 256 * iters * 8.0 * grid / ms / 1e9, 2.0 * 8 * iters * 256.0 * grid / ms / 1e9);
    ms = timeit([&] { k_mixed<8, 2><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    printf("{\"bench\":\"mixed_dmma8_dfma2\",\"ctas_per_sm\":%d,\"tflops_dmma\":%.2f,\"tflops_dfma\":%.2f}\n", bps,
           2.0 * 8 * 256 * iters * 8.0 * grid / ms / 1e9, 2.0 * 2 * iters * 256.0 * grid / ms / 1e9);
  }
  {
    int grid = nsm * 8; int it = 2000;
    float ms = timeit([&] { k_exp<<<grid, 256>>>(out, it, 1e-4); });
    printf("{\"bench\":\"exp_f64\",\"gexp_per_s\":%.2f}\n", 4.0 * it * 256.0 * grid / ms / 1e6);
    ms = timeit([&] { k_lds<<<grid, 256>>>(out, 20000); });
    printf("{\"bench\":\"lds64\",\"tb_per_s\":%.2f,\"bytes_per_clk_per_sm_at_1.9GHz\":%.1f}\n",
           4.0 * 20000 * 256.0 * grid * 8 / ms / 1e9, 4.0 * 20000 * 256.0 * grid * 8 / ms / 1e3 / 1.9e9 / nsm * 1e3);
  }
  return 0;
}
