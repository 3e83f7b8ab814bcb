// What does a warp of ordinary FP64 arithmetic get from the FP64 unit of an SM sub-partition while consumer warps stream
// m8n8k4 DMMAs through it?  (The generator warps of wfn_ws_kernel.cuh.)  One block per SM: NC consumer warps and NW worker
// warps per sub-partition; the workers run DFMA chains (ILP independent chains) until every consumer has finished ITERS
// groups of 4 DMMAs; reported: cycles per worker DFMA warp instruction and the DMMA rate of the consumers.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_share fp64_share.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int ILP>
__global__ void k(int ncw, int nww, int workers_first, int iters, unsigned long long* res, double* out, double a0, double b0) {
  __shared__ double sm[2048];
  __shared__ int done;
  if (threadIdx.x == 0) done = 0;
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = a0 + i * 1e-9;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool worker = workers_first ? warp < nww : warp >= ncw;
  if (!worker) {
    double c[4][2];
#pragma unroll
    for (int i = 0; i < 4; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
    double a = a0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      double an = sm[(lane + it * 32) & 2047];
#pragma unroll
      for (int i = 0; i < 4; i++) dmma884(c[i][0], c[i][1], a, b0);
      a = an;
    }
    const long long t1 = clock64();
    __syncwarp();
    if (lane == 0) { atomicAdd(&done, 1); res[(blockIdx.x * 32 + warp) * 2] = 0; res[(blockIdx.x * 32 + warp) * 2 + 1] = t1 - t0; }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  } else {
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = a0 + i + lane;
    unsigned long long n = 0;
    const long long t0 = clock64();
    while (*(volatile int*)&done < ncw) {
#pragma unroll
      for (int r = 0; r < 64 / ILP; r++)
#pragma unroll
        for (int i = 0; i < ILP; i++) x[i] = fma(x[i], a0, b0);
      n += 64;
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (lane == 0) { res[(blockIdx.x * 32 + warp) * 2] = n; res[(blockIdx.x * 32 + warp) * 2 + 1] = t1 - t0; }
  }
}

// Second question: does it help the worker to issue its FP64 instructions in dense bursts?  BURST DFMAs (4 independent
// chains) followed by GAP dependent integer instructions, repeated.
template <int BURST, int GAP>
__global__ void kb(int ncw, int nww, int iters, unsigned long long* res, double* out, double a0, double b0, int seed) {
  __shared__ double sm[2048];
  __shared__ int done;
  if (threadIdx.x == 0) done = 0;
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = a0 + i * 1e-9;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool worker = warp >= ncw;
  if (!worker) {
    double c[4][2];
#pragma unroll
    for (int i = 0; i < 4; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
    double a = a0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      double an = sm[(lane + it * 32) & 2047];
#pragma unroll
      for (int i = 0; i < 4; i++) dmma884(c[i][0], c[i][1], a, b0);
      a = an;
    }
    const long long t1 = clock64();
    __syncwarp();
    if (lane == 0) { atomicAdd(&done, 1); res[(blockIdx.x * 32 + warp) * 2] = 0; res[(blockIdx.x * 32 + warp) * 2 + 1] = t1 - t0; }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  } else {
    double x[4];
#pragma unroll
    for (int i = 0; i < 4; i++) x[i] = a0 + i + lane;
    unsigned long long n = 0;
    int z = seed + lane;
    const long long t0 = clock64();
    while (*(volatile int*)&done < ncw) {
#pragma unroll
      for (int r = 0; r < BURST / 4; r++)
#pragma unroll
        for (int i = 0; i < 4; i++) x[i] = fma(x[i], a0, b0);
#pragma unroll
      for (int g = 0; g < GAP; g++) z = z * 1664525 + 1013904223;
      n += BURST;
    }
    const long long t1 = clock64();
    double s = z;
#pragma unroll
    for (int i = 0; i < 4; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (lane == 0) { res[(blockIdx.x * 32 + warp) * 2] = n; res[(blockIdx.x * 32 + warp) * 2 + 1] = t1 - t0; }
  }
}
template <int BURST, int GAP>
void runb(int nsm, unsigned long long* res, double* out) {
  const int iters = 20000, ncw = 8, nww = 4, threads = 32 * (ncw + nww);
  cudaMemset(res, 0, sizeof(unsigned long long) * nsm * 64);
  kb<BURST, GAP><<<nsm, threads>>>(ncw, nww, iters, res, out, 1.0000001, 1e-9, 3);
  kb<BURST, GAP><<<nsm, threads>>>(ncw, nww, iters, res, out, 1.0000001, 1e-9, 3);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return; }
  static unsigned long long h[256 * 64];
  cudaMemcpy(h, res, sizeof(unsigned long long) * nsm * 64, cudaMemcpyDeviceToHost);
  double wn = 0, wt = 0, ct = 0; int ncs = 0;
  for (int b = 0; b < nsm; b++)
    for (int w = 0; w < ncw + nww; w++) {
      if (w >= ncw) { wn += (double)h[(b * 32 + w) * 2]; wt += (double)h[(b * 32 + w) * 2 + 1]; }
      else { ct += (double)h[(b * 32 + w) * 2 + 1]; ncs++; }
    }
  printf("{\"consumers_per_sp\":2,\"workers_per_sp\":1,\"burst\":%d,\"gap_int_instr\":%d,\"cycles_per_worker_dfma\":%.1f,"
         "\"dmma_pipe_use\":%.3f}\n", BURST, GAP, wt / wn, 2 * iters * 4.0 * 16.0 / (ct / ncs));
}
template <int ILP>
void run(int nsm, int nc, int nw, int first, unsigned long long* res, double* out) {
  const int iters = 20000, ncw = 4 * nc, nww = 4 * nw, threads = 32 * (ncw + nww);
  cudaMemset(res, 0, sizeof(unsigned long long) * nsm * 64);
  k<ILP><<<nsm, threads>>>(ncw, nww, first, iters, res, out, 1.0000001, 1e-9);
  k<ILP><<<nsm, threads>>>(ncw, nww, first, iters, res, out, 1.0000001, 1e-9);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return; }
  static unsigned long long h[256 * 64];
  cudaMemcpy(h, res, sizeof(unsigned long long) * nsm * 64, cudaMemcpyDeviceToHost);
  double wn = 0, wt = 0, ct = 0; int nwk = 0, ncs = 0;
  for (int b = 0; b < nsm; b++)
    for (int w = 0; w < ncw + nww; w++) {
      const bool worker = first ? w < nww : w >= ncw;
      if (worker) { wn += (double)h[(b * 32 + w) * 2]; wt += (double)h[(b * 32 + w) * 2 + 1]; nwk++; }
      else { ct += (double)h[(b * 32 + w) * 2 + 1]; ncs++; }
    }
  // DMMA pipe use of a sub-partition: nc warps x iters x 4 DMMAs x 16 cycles / consumer cycles
  printf("{\"consumers_per_sp\":%d,\"workers_per_sp\":%d,\"workers_first\":%d,\"ilp\":%d,\"cycles_per_worker_dfma\":%.1f,"
         "\"dmma_pipe_use\":%.3f}\n", nc, nw, first, ILP, nwk ? wt / wn : 0.0, nc * iters * 4.0 * 16.0 / (ct / ncs));
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount;
  double* out; cudaMalloc(&out, sizeof(double) * nsm * 1024);
  unsigned long long* res; cudaMalloc(&res, sizeof(unsigned long long) * nsm * 64);
  for (int nc = 0; nc <= 2; nc++)
    for (int nw = 1; nw <= 2; nw++)
      for (int first = 0; first <= (nc ? 1 : 0); first++) {
        run<1>(nsm, nc, nw, first, res, out); run<4>(nsm, nc, nw, first, res, out); run<8>(nsm, nc, nw, first, res, out);
      }
  runb<4, 0>(nsm, res, out); runb<16, 0>(nsm, res, out); runb<64, 0>(nsm, res, out); runb<256, 0>(nsm, res, out);
  runb<16, 16>(nsm, res, out); runb<64, 16>(nsm, res, out); runb<64, 64>(nsm, res, out); runb<256, 64>(nsm, res, out); runb<256, 256>(nsm, res, out);
  return 0;
}
