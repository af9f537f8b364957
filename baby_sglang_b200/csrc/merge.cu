// K2c  LSE merge of the split-KV partials written by the paged-attention kernel.
// No reference counterpart (the reference never splits a request).  One CTA per (sequence, q
// head): sequences that lived in a single CTA were already finalised and are skipped; rows with no
// KV tokens are zero-filled.  Traffic: a few KB per split sequence.
#include "common.cuh"

namespace bsgl {

constexpr int kMergeWarps = 4;

// One CTA per (sequence, q head): its warps stride over the partials with an online LSE merge
// (independent loads, unrolled for memory-level parallelism), then combine through shared memory.
template <typename T>
__global__ void __launch_bounds__(kMergeWarps * 32) merge_partials_kernel(
    const float* __restrict__ part_o, const float* __restrict__ part_lse, const int32_t* __restrict__ plan_hdr,
    const int32_t* __restrict__ tile_indptr, int n_q_heads, int n_kv_heads, int hpb, int D, T* __restrict__ out) {
  const int G = n_q_heads / n_kv_heads;
  const int heads = hpb * G;                     // q heads of one sequence
  const int j = blockIdx.x / heads;
  const int head = blockIdx.x - j * heads;
  const int n_tiles = plan_hdr[0], n_eff = plan_hdr[1];
  const int nhg = n_kv_heads / hpb;
  const int row = j / nhg, hg = j - row * nhg;
  const int tb = tile_indptr[j], te = tile_indptr[j + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = lane * 4;
  T* orow = out + ((int64_t)row * n_q_heads + (int64_t)hg * heads + head) * D;
  if (tb == te) {                                // no KV tokens: defined as zeros
    if (warp == 0 && d < D) *reinterpret_cast<uint2*>(orow + d) = make_uint2(0u, 0u);
    return;
  }
  const int c_first = cta_of_tile(tb, n_tiles, n_eff);
  const int c_last = cta_of_tile(te - 1, n_tiles, n_eff);
  if (c_first == c_last) return;                 // written in full by the attention kernel
  __shared__ float s_acc[kMergeWarps][128];
  __shared__ float s_m[kMergeWarps], s_den[kMergeWarps];
  float m = -INFINITY, den = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 4
  for (int c = c_first + warp; c <= c_last; c += kMergeWarps) {
    const int64_t slot = 2 * c + (c == c_first ? 1 : 0);
    const float lse = part_lse[slot * heads + head];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (d < D) v = *reinterpret_cast<const float4*>(part_o + (slot * heads + head) * D + d);
    const float mn = fmaxf(m, lse);
    const float sc = fast_exp2(m - mn), w = fast_exp2(lse - mn);   // m = -inf on the first part: sc = 0
    den = den * sc + w;
    a0 = a0 * sc + w * v.x; a1 = a1 * sc + w * v.y; a2 = a2 * sc + w * v.z; a3 = a3 * sc + w * v.w;
    m = mn;
  }
  if (d < D) {
    s_acc[warp][d] = a0; s_acc[warp][d + 1] = a1; s_acc[warp][d + 2] = a2; s_acc[warp][d + 3] = a3;
  }
  if (lane == 0) { s_m[warp] = m; s_den[warp] = den; }
  __syncthreads();
  if (warp == 0 && d < D) {
    float mm = -INFINITY;
#pragma unroll
    for (int w = 0; w < kMergeWarps; ++w) mm = fmaxf(mm, s_m[w]);
    float dsum = 0.f, x0 = 0.f, x1 = 0.f, x2 = 0.f, x3 = 0.f;
#pragma unroll
    for (int w = 0; w < kMergeWarps; ++w) {
      const float wm = s_m[w];
      const float wt = (wm == -INFINITY) ? 0.f : fast_exp2(wm - mm);
      dsum += wt * s_den[w];
      x0 += wt * s_acc[w][d]; x1 += wt * s_acc[w][d + 1]; x2 += wt * s_acc[w][d + 2]; x3 += wt * s_acc[w][d + 3];
    }
    const float inv = 1.f / dsum;
    uint2 pk;
    pk.x = pack2<T>(x0 * inv, x1 * inv);
    pk.y = pack2<T>(x2 * inv, x3 * inv);
    *reinterpret_cast<uint2*>(orow + d) = pk;
  }
}

}  // namespace bsgl

extern "C" int bsgl_merge_partials(const float* part_o, const float* part_lse, const int32_t* plan_hdr,
                                   const int32_t* tile_indptr, int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                                   int dtype, void* out, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_rows >= 0, "bsgl_merge_partials: negative row count");
  if (n_rows == 0) return BSGL_OK;
  BSGL_REQUIRE(part_o && part_lse && plan_hdr && tile_indptr && out, "bsgl_merge_partials: null pointer");
  BSGL_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0 && head_dim % 4 == 0 && head_dim <= 128,
               "bsgl_merge_partials: bad geometry");
  const int hpb = bsgl_heads_per_group(n_kv_heads);
  const int64_t n_blocks = (int64_t)n_rows * (n_kv_heads / hpb) * hpb * (n_q_heads / n_kv_heads);   // rows * q heads
  BSGL_REQUIRE(n_blocks < (1ll << 31), "bsgl_merge_partials: too many rows");
  const unsigned n_seqs = (unsigned)n_blocks;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == BSGL_BF16)
    merge_partials_kernel<__nv_bfloat16><<<n_seqs, kMergeWarps * 32, 0, s>>>(part_o, part_lse, plan_hdr, tile_indptr, n_q_heads,
                                                               n_kv_heads, hpb, head_dim, (__nv_bfloat16*)out);
  else if (dtype == BSGL_F16)
    merge_partials_kernel<__half><<<n_seqs, kMergeWarps * 32, 0, s>>>(part_o, part_lse, plan_hdr, tile_indptr, n_q_heads,
                                                        n_kv_heads, hpb, head_dim, (__half*)out);
  else {
    set_error("bsgl_merge_partials: dtype must be bf16 or fp16");
    return BSGL_ERR_INVALID;
  }
  return check_launch("merge_partials_kernel");
}
