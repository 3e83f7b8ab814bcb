// How many warps per SM sub-partition does the FP64 tensor pipe need?  Chains of independent m8n8k4 DMMAs (NACC accumulator
// pairs per thread) from 1, 2, 3 warps per sub-partition (blocks of 128, 256, 384 threads, one block per SM), with and
// without a shared-memory operand load per DMMA group (the consumer loop of wfn_ws_kernel.cuh).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma_issue dmma_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int NACC, bool LDS>
__global__ void k(double* out, int iters, double a0, double b0) {
  __shared__ double sm[2048];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = a0 + i * 1e-9;
  __syncthreads();
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
  double a = a0;
  int idx = threadIdx.x & 31;
  for (int it = 0; it < iters; it++) {
    double an = LDS ? sm[(idx + it * 32) & 2047] : a;
#pragma unroll
    for (int i = 0; i < NACC; i++) dmma884(c[i][0], c[i][1], a, b0);
    a = an;
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// prefetch distance 2: the operand of group it+2 is loaded while group it issues
template <int NACC>
__global__ void k2(double* out, int iters, double a0, double b0) {
  __shared__ double sm[2048];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = a0 + i * 1e-9;
  __syncthreads();
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
  int idx = threadIdx.x & 31;
  double a = sm[idx], a1 = sm[idx + 32];
  for (int it = 0; it < iters; it++) {
    double a2 = sm[(idx + (it + 2) * 32) & 2047];
#pragma unroll
    for (int i = 0; i < NACC; i++) dmma884(c[i][0], c[i][1], a, b0);
    a = a1; a1 = a2;
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
void run2(double* out, int nsm, int threads) {
  const int iters = 40000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; r++) {
    cudaEventRecord(e0); k2<NACC><<<nsm, threads>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
  }
  printf("{\"warps_per_subpartition\":%d,\"nacc\":%d,\"lds\":2,\"tflops\":%.2f}\n", threads / 128, NACC,
         2.0 * NACC * 256 * (double)iters * (threads / 32) * nsm / best / 1e9);
}
template <int NACC, bool LDS>
void run(double* out, int nsm, int threads) {
  const int iters = 40000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; r++) {
    cudaEventRecord(e0); k<NACC, LDS><<<nsm, threads>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
  }
  printf("{\"warps_per_subpartition\":%d,\"nacc\":%d,\"lds\":%d,\"tflops\":%.2f}\n", threads / 128, NACC, (int)LDS,
         2.0 * NACC * 256 * (double)iters * (threads / 32) * nsm / best / 1e9);
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double* out; cudaMalloc(&out, sizeof(double) * p.multiProcessorCount * 512);
  for (int t = 128; t <= 384; t += 128) {
    run<2, false>(out, p.multiProcessorCount, t); run<3, false>(out, p.multiProcessorCount, t);
    run<4, false>(out, p.multiProcessorCount, t); run<3, true>(out, p.multiProcessorCount, t);
    run<4, true>(out, p.multiProcessorCount, t); run<8, false>(out, p.multiProcessorCount, t);
    run2<3>(out, p.multiProcessorCount, t); run2<4>(out, p.multiProcessorCount, t);
  }
  return 0;
}
