// Development aid: clock64 timeline of one CTA of the K3 extend kernel (softmax A/B, MMA issuer, producer 0).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -I baby_sglang_b200/csrc \
//        scripts/ext_trace.cu baby_sglang_b200/csrc/c_api.cu -o scripts/build/ext_trace
// Usage: ext_trace [prefix=4096] [batch=16] [traced_cta=0]
#define BSGL_EXT_TRACE 1
#include "../baby_sglang_b200/csrc/extend_attention.cu"

#include <cstdlib>
#include <vector>

int main(int argc, char** argv) {
  const int prefix = argc > 1 ? atoi(argv[1]) : 4096;
  const int B = argc > 2 ? atoi(argv[2]) : 16;
  const int cta = argc > 3 ? atoi(argv[3]) : 0;
  const int ext = 512, Hq = 28, Hkv = 4, D = 128;
  const int seq = prefix + ext, C = seq + 64;
  const int64_t S = (int64_t)B * seq + 1;
  const int64_t T = (int64_t)B * ext;
  std::vector<int32_t> r2t((size_t)B * C, 0);
  uint32_t lcg = 12345;
  for (int i = 0; i < B; ++i)
    for (int t = 0; t < seq; ++t) {   // scattered slots: multiplicative hash over the slab
      lcg = lcg * 1664525u + 1013904223u;
      r2t[(size_t)i * C + t] = 1 + (int32_t)(((uint64_t)i * seq + t) * 2654435761ull % (uint64_t)(S - 1));
    }
  std::vector<int64_t> rpi(B), pre(B, prefix);
  std::vector<int32_t> qo(B + 1);
  for (int i = 0; i < B; ++i) rpi[i] = i, qo[i] = i * ext;
  qo[B] = B * ext;
  const size_t slab = (size_t)S * Hkv * D, qn = (size_t)T * Hq * D;
  std::vector<__nv_bfloat16> hs(slab), hq(qn);
  for (size_t i = 0; i < slab; ++i) {
    lcg = lcg * 1664525u + 1013904223u;
    hs[i] = __float2bfloat16(((int)(lcg >> 16) % 2001 - 1000) * 1e-3f);
  }
  for (size_t i = 0; i < qn; ++i) {
    lcg = lcg * 1664525u + 1013904223u;
    hq[i] = __float2bfloat16(((int)(lcg >> 16) % 2001 - 1000) * 1e-3f);
  }
  __nv_bfloat16 *dk, *dv, *dq, *dout;
  int32_t *dr2t, *dqo;
  int64_t *drpi, *dpre;
  long long* dtrace;
  cudaMalloc(&dk, slab * 2); cudaMalloc(&dv, slab * 2); cudaMalloc(&dq, qn * 2); cudaMalloc(&dout, qn * 2);
  cudaMalloc(&dr2t, r2t.size() * 4); cudaMalloc(&dqo, (B + 1) * 4); cudaMalloc(&drpi, B * 8); cudaMalloc(&dpre, B * 8);
  cudaMalloc(&dtrace, 10 * 128 * 8 * 8);
  cudaMemcpy(dk, hs.data(), slab * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dv, hs.data(), slab * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dq, hq.data(), qn * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dr2t, r2t.data(), r2t.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dqo, qo.data(), (B + 1) * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(drpi, rpi.data(), B * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(dpre, pre.data(), B * 8, cudaMemcpyHostToDevice);
  cudaMemset(dtrace, 0, 10 * 128 * 8 * 8);
  auto run = [&]() {
    return bsgl_extend_attention(dq, dk, dv, dr2t, C, drpi, nullptr, dpre, dqo, nullptr, B, T, Hq, Hkv, D,
                                 0.08838834764f, BSGL_BF16, dout, nullptr, nullptr);
  };
  for (int i = 0; i < 3; ++i) run();
  cudaDeviceSynchronize();
  cudaMemcpyToSymbol(bsgl::g_ext_trace, &dtrace, sizeof(dtrace));
  cudaMemcpyToSymbol(bsgl::g_ext_trace_cta, &cta, sizeof(cta));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  int rc = run();
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("rc=%d err=%s  %.4f ms (traced run)\n", rc, cudaGetErrorString(err), ms);
  std::vector<long long> tr(10 * 128 * 8);
  cudaMemcpy(tr.data(), dtrace, tr.size() * 8, cudaMemcpyDeviceToHost);
  long long t0 = 0;
  for (auto v : tr) if (v && (!t0 || v < t0)) t0 = v;
  const char* names[10] = {"softmax w0 (A)", "softmax w1 (A)", "softmax w2 (A)", "softmax w3 (A)", "softmax w4 (B)",
                           "softmax w5 (B)", "softmax w6 (B)", "softmax w7 (B)", "mma: Vrdy pA0 pA1 S_A' pB0 pB1 S_B'",
                           "prod0: Kempty Kissued Vempty Vissued"};
  for (int role = 0; role < 10; ++role) {
    printf("== %s\n", names[role]);
    for (int j = 0; j < 128; ++j) {
      bool any = false;
      for (int e = 0; e < 8; ++e) any |= tr[(role * 128 + j) * 8 + e] != 0;
      if (!any) continue;
      printf("%3d:", j);
      for (int e = 0; e < 7; ++e) {
        long long v = tr[(role * 128 + j) * 8 + e];
        printf(" %8lld", v ? v - t0 : -1);
      }
      printf("\n");
    }
  }
  return 0;
}
