// K2c  LSE merge of the split-KV partials written by the paged-attention kernel.
// No reference counterpart (the reference never splits a request).  A sequence is split only where a CTA
// boundary of the plan falls inside it, so there are at most n_ctas-1 split sequences: the grid is one CTA per
// (boundary, 4 q heads) -- the CTA of boundary b merges the sequence that crosses b|b+1 if b is the first
// boundary inside it -- plus a few CTAs that zero-fill rows with no KV tokens.  (A first version launched one
// CTA per (sequence, q head), almost all of which returned at once: 13-16 us per launch at B=256..1024.)
// Traffic: a few KB per split sequence.
#include "common.cuh"

namespace bsgl {

constexpr int kMergeWarps = 4;
constexpr int kZeroSeqsPerCta = 1024;

// Warp w of CTA (b, y) merges q head y*4+w of the sequence crossing boundary b: online LSE merge over its
// partials (independent loads, unrolled for memory-level parallelism).
template <typename T>
__global__ void __launch_bounds__(kMergeWarps * 32) merge_partials_kernel(
    const float* __restrict__ part_o, const float* __restrict__ part_lse, const int32_t* __restrict__ plan_hdr,
    const int32_t* __restrict__ tile_indptr, const int32_t* __restrict__ cta_first_seq, int n_bound, int n_q_heads,
    int n_kv_heads, int hpb, int D, T* __restrict__ out) {
  const int G = n_q_heads / n_kv_heads;
  const int heads = hpb * G;                     // q heads of one sequence
  const int nhg = n_kv_heads / hpb;
  const int n_tiles = plan_hdr[0], n_eff = plan_hdr[1], n_seqs = plan_hdr[2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Programmatic dependent launch: EVERY thread of this grid executes griddepcontrol.wait before it exits or
  // touches anything the attention kernel writes (PTX ISA: a dependent grid must wait when its prerequisite uses
  // launch_dependents; otherwise this grid could retire -- and release the next kernel on the stream, which is
  // ordered after THIS grid only -- while the attention kernel is still writing `out` and the partials).  The
  // plan look-ups below read only buffers written before the attention kernel started, so they overlap its tail.
  if ((int)blockIdx.x >= n_bound) {
    // ---- rows with no KV tokens are defined as zeros (nobody else writes them)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (blockIdx.y != 0) return;
    const int j0 = ((int)blockIdx.x - n_bound) * kZeroSeqsPerCta;
    for (int j = j0 + threadIdx.x; j < min(j0 + kZeroSeqsPerCta, n_seqs); j += blockDim.x) {
      if (tile_indptr[j] != tile_indptr[j + 1]) continue;
      const int row = j / nhg, hg = j - row * nhg;
      T* orow = out + ((int64_t)row * n_q_heads + (int64_t)hg * heads) * D;
      for (int e = 0; e < heads * D; e += 4) *reinterpret_cast<uint2*>(orow + e) = make_uint2(0u, 0u);
    }
    return;
  }
  const int b = blockIdx.x;
  const int head = blockIdx.y * kMergeWarps + warp;
  bool work = !(b + 1 >= n_eff || head >= heads || n_tiles == 0);
  int j = 0, tb = 0, te = 0, c_first = 0, c_last = 0;
  if (work) {
    const int t = cta_tile_begin(b + 1, n_tiles, n_eff);   // first tile of CTA b+1
    j = cta_first_seq[b + 1];                              // the sequence that holds it
    work = j < n_seqs;
    if (work) {
      tb = tile_indptr[j], te = tile_indptr[j + 1];
      work = tb < t;                                       // starts at the boundary: not split here
    }
    if (work) {
      c_first = cta_of_tile(tb, n_tiles, n_eff);
      work = c_first == b;                                 // an earlier boundary merges this sequence
      c_last = cta_of_tile(te - 1, n_tiles, n_eff);
    }
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (!work) return;
  const int row = j / nhg, hg = j - row * nhg;
  const int d = lane * 4;
  T* orow = out + ((int64_t)row * n_q_heads + (int64_t)hg * heads + head) * D;
  float m = -INFINITY, den = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 4
  for (int c = c_first; c <= c_last; ++c) {
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
    const float inv = 1.f / den;
    uint2 pk;
    pk.x = pack2<T>(a0 * inv, a1 * inv);
    pk.y = pack2<T>(a2 * inv, a3 * inv);
    *reinterpret_cast<uint2*>(orow + d) = pk;
  }
}

}  // namespace bsgl

extern "C" int bsgl_merge_partials(const float* part_o, const float* part_lse, const int32_t* plan_hdr,
                                   const int32_t* tile_indptr, const int32_t* cta_first_seq, int n_ctas, int n_rows,
                                   int n_q_heads, int n_kv_heads, int head_dim, int dtype, void* out, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_rows >= 0, "bsgl_merge_partials: negative row count");
  if (n_rows == 0) return BSGL_OK;
  BSGL_REQUIRE(part_o && part_lse && plan_hdr && tile_indptr && cta_first_seq && out, "bsgl_merge_partials: null pointer");
  BSGL_REQUIRE(n_ctas >= 1, "bsgl_merge_partials: n_ctas must be positive");
  BSGL_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0 && head_dim % 4 == 0 && head_dim <= 128,
               "bsgl_merge_partials: bad geometry");
  const int hpb = bsgl_heads_per_group(n_kv_heads);
  const int64_t n_seqs = (int64_t)n_rows * (n_kv_heads / hpb);
  BSGL_REQUIRE(n_seqs < (1ll << 31), "bsgl_merge_partials: too many rows");
  const int heads = hpb * (n_q_heads / n_kv_heads);
  const int n_bound = n_ctas - 1;                                  // CTA boundaries of the plan
  const int n_zero = (int)((n_seqs + kZeroSeqsPerCta - 1) / kZeroSeqsPerCta);
  const dim3 grid((unsigned)(n_bound + n_zero), (unsigned)((heads + kMergeWarps - 1) / kMergeWarps));
  // launched as a programmatic dependent of the attention kernel before it on the stream (which issues
  // griddepcontrol.launch_dependents at its start): launch latency and the plan look-ups hide behind its tail
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kMergeWarps * 32);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (dtype == BSGL_BF16)
    BSGL_CUDA_TRY(cudaLaunchKernelEx(&cfg, merge_partials_kernel<__nv_bfloat16>, part_o, part_lse, plan_hdr, tile_indptr,
                                     cta_first_seq, n_bound, n_q_heads, n_kv_heads, hpb, head_dim,
                                     (__nv_bfloat16*)out));
  else if (dtype == BSGL_F16)
    BSGL_CUDA_TRY(cudaLaunchKernelEx(&cfg, merge_partials_kernel<__half>, part_o, part_lse, plan_hdr, tile_indptr,
                                     cta_first_seq, n_bound, n_q_heads, n_kv_heads, hpb, head_dim, (__half*)out));
  else {
    set_error("bsgl_merge_partials: dtype must be bf16 or fp16");
    return BSGL_ERR_INVALID;
  }
  return check_launch("merge_partials_kernel");
}
