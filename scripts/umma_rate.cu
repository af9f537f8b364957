// Micro-benchmark (development aid): issue rate of tcgen05.mma.kind::f16 for the shapes the extend kernel uses.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I baby_sglang_b200/csrc scripts/umma_rate.cu -o scripts/build/umma_rate
#include <cstdio>
#include <vector>
#include "tcgen05.cuh"
namespace bsgl { void set_error(const char*, ...) {} int cuda_fail(cudaError_t, const char*) { return 1; } }
using namespace bsgl;

__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int ts, int mn_major_b, int reps, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (warp == 0) { tc::tmem_alloc(smem_u32(&tmem_slot), 512); tc::tmem_relinquish(); }
  tc::fence_before_sync(); __syncthreads(); tc::fence_after_sync();
  const uint32_t tm = tmem_slot;
  fence_proxy_async_smem();
  if (warp == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)mn_major_b << 16) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
    const uint32_t a = base, b = base + 32 * 1024;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (tc::elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          if (ts) tc::mma_ts(tm + 256, tm + kk * 8, mn_major_b ? tc::smem_desc_sw128(b + kk * 2048, 16384, 1024) : tc::smem_desc_sw128(b + off, 16, 1024), idesc, 1u);
          else tc::mma_ss(tm + (r & 1) * 128, tc::smem_desc_sw128(a + off, 16, 1024), tc::smem_desc_sw128(b + off, 16, 1024), idesc, kk > 0);
        }
      }
      __syncwarp();
    }
    if (tc::elect_one()) tc::commit(smem_u32(&bar));
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) { tc::fence_after_sync(); tc::tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const int reps = 2000;
  struct V { const char* name; int N, ts, mn; } vs[] = {
      {"SS K-major N=128", 128, 0, 0}, {"SS K-major N=64", 64, 0, 0}, {"SS K-major N=32", 32, 0, 0}, {"SS K-major N=256", 256, 0, 0},
      {"TS MN-major N=128", 128, 1, 1}, {"TS MN-major N=64", 64, 1, 1}, {"TS K-major N=128", 128, 1, 0}, {"TS K-major N=64", 64, 1, 0}};
  for (int grid : {1, 148})
    for (auto& v : vs) {
      rate_kernel<<<grid, 128, 100 * 1024>>>(v.N, v.ts, v.mn, reps, d);
      cudaError_t e = cudaDeviceSynchronize();
      std::vector<long long> h(148);
      cudaMemcpy(h.data(), d, grid * 8, cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("grid %3d %-20s: %s  %.1f clk / MMA (M128 K16)\n", grid, v.name, cudaGetErrorString(e), (double)mx / (reps * 8));
    }
  return 0;
}
