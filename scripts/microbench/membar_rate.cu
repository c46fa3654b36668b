// membar_rate.cu — how many device-scope release fences (MEMBAR.ALL.GPU) per second does a B200 sustain, chip-wide?
// Built and run on the GPU box: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o membar_rate membar_rate.cu
// Each thread (or lane 0 of each warp) loops: plain store to its own line, then an atom.release / red.relaxed on its
// own counter. Variants: per-lane vs per-warp, with and without the release.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>  // 4: release per lane, lanes diverged; 0: no fence (store + relaxed red), 1: release per lane, 2: release by lane 0 only, 3: __threadfence per lane
__global__ void k(double* data, unsigned* cnt, int iters) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool lead = (threadIdx.x & 31) == 0;
  for (int i = 0; i < iters; i++) {
    data[t * 8 + (i & 7)] = (double)i;
    if (MODE == 0) {
      asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(cnt + t), "r"(1u) : "memory");
    } else if (MODE == 1) {
      asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(cnt + t), "r"(1u) : "memory");
    } else if (MODE == 2) {
      __syncwarp();
      if (lead) asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(cnt + t), "r"(1u) : "memory");
      __syncwarp();
    } else if (MODE == 4) {  // divergent: every lane fences on its own, one after the other (as a per-lane delivery does)
      for (int l = 0; l < 32; l++)
        if ((int)(threadIdx.x & 31) == l) asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(cnt + t), "r"(1u) : "memory");
    } else {
      __threadfence();
      asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(cnt + t), "r"(1u) : "memory");
    }
  }
}

template <int MODE>
void run(const char* name, int blocks, int threads, int iters, double* d, unsigned* c) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<blocks, threads>>>(d, c, 8);
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  k<MODE><<<blocks, threads>>>(d, c, iters);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  const double fences = MODE == 2 ? (double)blocks * threads / 32 * iters : (double)blocks * threads * iters;
  printf("%-34s grid %4d x %3d  iters %5d  %8.3f ms  %8.1f M ops/s chip-wide  %7.1f ns per op per warp\n", name, blocks, threads, iters, ms,
         fences / ms / 1e3, ms * 1e6 / iters / (MODE == 2 ? 1 : 1));
}

int main() {
  const int maxb = 148 * 16, thr = 128;
  double* d; unsigned* c;
  cudaMalloc(&d, (size_t)maxb * thr * 8 * sizeof(double));
  cudaMalloc(&c, (size_t)maxb * thr * sizeof(unsigned));
  cudaMemset(c, 0, (size_t)maxb * thr * sizeof(unsigned));
  for (int blocks : {148 * 4, 148, 16, 1}) {
    run<0>("store + red.relaxed per lane", blocks, thr, 2000, d, c);
    run<1>("store + red.release per lane", blocks, thr, 200, d, c);
    run<2>("store + red.release lane 0 per warp", blocks, thr, 2000, d, c);
    run<3>("store + __threadfence per lane", blocks, thr, 200, d, c);
    run<4>("store + red.release, lanes diverged", blocks, thr, 50, d, c);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
