// fp64_ilp.cu — how much instruction-level parallelism does the B200 FP64 pipe need at a given
// number of resident warps? Non-fused DMUL+DADD chains (the library is built --fmad=false).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o fp64_ilp fp64_ilp.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void k(double* out, int iters, double b, double c) {
  double a[CHAINS];
#pragma unroll
  for (int j = 0; j < CHAINS; j++) a[j] = threadIdx.x * 1e-3 + j;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < CHAINS; j++) a[j] = __dadd_rn(__dmul_rn(a[j], b), c);
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < CHAINS; j++) s += a[j];
  if (s == 123.456) out[0] = s;
}

template <int CHAINS>
void run(int warps_per_sm, int sms, double* d) {
  int threads = 128, blocks = sms * warps_per_sm / 4;
  int iters = 1 << 15;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<CHAINS><<<blocks, threads>>>(d, 64, 1.0000001, 1e-9);
  cudaEventRecord(e0);
  k<CHAINS><<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double flops = (double)blocks * threads * iters * CHAINS * 2.0;
  printf("warps/SM %2d chains %d : %7.2f TFLOP/s\n", warps_per_sm, CHAINS, flops / (ms * 1e-3) / 1e12);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double* d; cudaMalloc(&d, 64);
  for (int w : {4, 8, 12, 16, 20, 32, 64}) {
    run<1>(w, p.multiProcessorCount, d); run<2>(w, p.multiProcessorCount, d);
    run<4>(w, p.multiProcessorCount, d); run<8>(w, p.multiProcessorCount, d);
  }
  return 0;
}
