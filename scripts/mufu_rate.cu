// Micro-benchmark (development aid): per-SMSP issue cost of the softmax building blocks of the extend kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/mufu_rate.cu -o scripts/build/mufu_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(int reps, long long* out, float* sink, float seed) {
  float a[8];
  uint32_t h[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed * (i + 1 + threadIdx.x), h[i] = __float_as_uint(a[i]);
  __syncthreads();
  long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i]));
      if (MODE == 4) {
        unsigned long long v = ((unsigned long long)h[i] << 32) | h[i];
        asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(v));
        h[i] = (uint32_t)v ^ (uint32_t)(v >> 32);
      }
      if (MODE == 5) asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(a[(i + 1) & 7]));
      if (MODE == 6) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(a[i]), "f"(a[(i + 1) & 7]));
      if (MODE == 7) asm volatile("ex2.approx.f16 %0, %0;" : "+h"(*(unsigned short*)&h[i]));
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(h[i]);
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

int main() {
  long long* d; float* sink;
  cudaMalloc(&d, 8); cudaMalloc(&sink, 4096 * 4);
  const int reps = 4000;
  const char* names[] = {"ex2.f32", "ex2.f16x2", "ex2.bf16x2", "fma.f32", "fma.f32x2", "max.f32", "cvt.bf16x2", "ex2.f16"};
  for (int warps : {4, 8, 16}) {
    for (int m = 0; m < 8; ++m) {
      long long h = 0;
      switch (m) {
        case 0: k<0><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 1: k<1><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 2: k<2><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 3: k<3><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 4: k<4><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 5: k<5><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 6: k<6><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
        case 7: k<7><<<1, warps * 32>>>(reps, d, sink, 1e-3f); break;
      }
      cudaError_t e = cudaDeviceSynchronize();
      cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      printf("%2d warps/SM (%d per SMSP) %-10s: %s  %.2f clk per warp-instruction per SMSP\n", warps, warps / 4, names[m],
             cudaGetErrorString(e), (double)h / (reps * 8.0 * (warps / 4)));
    }
  }
  return 0;
}
