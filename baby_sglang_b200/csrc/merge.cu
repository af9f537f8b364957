// K2c  LSE merge of the split-KV partials written by the paged-attention kernel.
// No reference counterpart (the reference never splits a request).  One CTA per sequence
// (q row, head group): sequences that lived in a single CTA were already finalised and are
// skipped; rows with no KV tokens are zero-filled.  Traffic: a few KB per split sequence.
#include "common.cuh"

namespace bsgl {

template <typename T>
__global__ void __launch_bounds__(128) merge_partials_kernel(const float* __restrict__ part_o,
                                                             const float* __restrict__ part_lse,
                                                             const int32_t* __restrict__ plan_hdr,
                                                             const int32_t* __restrict__ tile_indptr, int n_q_heads,
                                                             int n_kv_heads, int hpb, int D, T* __restrict__ out) {
  const int j = blockIdx.x;
  const int n_tiles = plan_hdr[0], n_eff = plan_hdr[1];
  const int nhg = n_kv_heads / hpb;
  const int G = n_q_heads / n_kv_heads;
  const int heads = hpb * G;                     // q heads of this sequence
  const int row = j / nhg, hg = j - row * nhg;
  const int tb = tile_indptr[j], te = tile_indptr[j + 1];
  T* orow = out + ((int64_t)row * n_q_heads + (int64_t)hg * heads) * D;
  const int n_vec = heads * D / 4;
  if (tb == te) {                                // no KV tokens: defined as zeros
    for (int i = threadIdx.x; i < n_vec; i += blockDim.x) *reinterpret_cast<uint2*>(orow + i * 4) = make_uint2(0u, 0u);
    return;
  }
  const int c_first = cta_of_tile(tb, n_tiles, n_eff);
  const int c_last = cta_of_tile(te - 1, n_tiles, n_eff);
  if (c_first == c_last) return;                 // written in full by the attention kernel
  for (int i = threadIdx.x; i < n_vec; i += blockDim.x) {
    const int head = (i * 4) / D;
    float mx = -INFINITY;
    for (int c = c_first; c <= c_last; ++c) {
      const int slot = 2 * c + (c == c_first ? 1 : 0);
      mx = fmaxf(mx, part_lse[(int64_t)slot * heads + head]);
    }
    float den = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int c = c_first; c <= c_last; ++c) {
      const int slot = 2 * c + (c == c_first ? 1 : 0);
      const float w = fast_exp2(part_lse[(int64_t)slot * heads + head] - mx);
      const float4 v = *reinterpret_cast<const float4*>(part_o + (int64_t)slot * heads * D + i * 4);
      den += w;
      a0 += w * v.x; a1 += w * v.y; a2 += w * v.z; a3 += w * v.w;
    }
    const float inv = 1.f / den;
    uint2 pk;
    pk.x = pack2<T>(a0 * inv, a1 * inv);
    pk.y = pack2<T>(a2 * inv, a3 * inv);
    *reinterpret_cast<uint2*>(orow + i * 4) = pk;
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
  BSGL_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0 && head_dim % 4 == 0, "bsgl_merge_partials: bad geometry");
  const int hpb = bsgl_heads_per_group(n_kv_heads);
  const int n_seqs = n_rows * (n_kv_heads / hpb);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == BSGL_BF16)
    merge_partials_kernel<__nv_bfloat16><<<n_seqs, 128, 0, s>>>(part_o, part_lse, plan_hdr, tile_indptr, n_q_heads,
                                                               n_kv_heads, hpb, head_dim, (__nv_bfloat16*)out);
  else if (dtype == BSGL_F16)
    merge_partials_kernel<__half><<<n_seqs, 128, 0, s>>>(part_o, part_lse, plan_hdr, tile_indptr, n_q_heads,
                                                        n_kv_heads, hpb, head_dim, (__half*)out);
  else {
    set_error("bsgl_merge_partials: dtype must be bf16 or fp16");
    return BSGL_ERR_INVALID;
  }
  return check_launch("merge_partials_kernel");
}
